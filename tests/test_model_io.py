"""f3: the binary tree-model format (BinaryDTSerializer v4).  tests/golden/readablespec_model0.gbt is the reference's own
fixture src/test/resources/example/readablespec/model0.gbt (copied unchanged): our reader must consume it without
trailing bytes and our writer must reproduce its payload byte for byte."""
import gzip
import os

from shifu_b200 import model_io

HERE = os.path.dirname(os.path.abspath(__file__))
FIXTURE = os.path.join(HERE, "golden", "readablespec_model0.gbt")


def test_reference_fixture_round_trips_byte_exact(tmp_path):
    payload = gzip.open(FIXTURE, "rb").read()
    m = model_io.read_payload(payload)
    assert m["trailing"] == 0
    assert (m["version"], m["algorithm"], m["loss"], m["input_count"]) == (4, "GBT", "squared", 30)
    assert len(m["bags"]) == 1 and len(m["bags"][0]) == 100
    t0, t1 = m["bags"][0][0], m["bags"][0][1]
    assert t0["learning_rate"] == 1.0 and t1["learning_rate"] == 0.05 and t0["root_wgt_cnt"] == 348.0
    assert model_io.write_payload(m) == payload
    out = tmp_path / "m.gbt"
    model_io.write_model(m, out)
    assert model_io.read_model(out) == m


def test_scoring_walks_thresholds_like_independent_tree_model():
    m = model_io.read_model(FIXTURE)
    n_inputs = m["input_count"]
    lo = model_io.score_gbt(m, [-1e9] * n_inputs)   # everything left
    hi = model_io.score_gbt(m, [1e9] * n_inputs)    # everything right
    assert lo != hi and abs(lo) < 10 and abs(hi) < 10


def test_reader_agrees_with_the_reference_written_spec_of_the_same_model():
    """tests/golden/readablespec_model0.zip is the reference's own second rendering of model0.gbt (src/test/resources/
    example/readablespec/model0.zip, written by IndependentTreeModelUtils.convertBinaryToZipSpec, util/
    IndependentTreeModelUtils.java): `model.ini` = the header as JSON, `trees` = the tree section as raw bytes.  What our
    reader takes out of the .gbt must be what the reference itself took out of it."""
    import json
    import zipfile
    z = zipfile.ZipFile(os.path.join(HERE, "golden", "readablespec_model0.zip"))
    ini = json.loads(z.read("model.ini"))
    trees = z.read("trees")
    payload = gzip.open(FIXTURE, "rb").read()
    m = model_io.read_payload(payload)
    assert (ini["algorithm"], ini["lossStr"], ini["inputNode"], ini["classification"], ini["gbdt"]) == \
        (m["algorithm"], m["loss"], m["input_count"], m["is_classification"], True)
    assert {int(k): v for k, v in ini["numNameMapping"].items()} == dict(m["column_names"])
    assert {int(k): v for k, v in ini["columnNumIndexMapping"].items()} == dict(m["column_mapping"])
    assert {int(k): v for k, v in ini["numericalMeanMapping"].items()} == dict(m["numerical_mean"])
    assert {int(k): v for k, v in ini["categoricalColumnNameNames"].items()} == \
        {c: [s for _, s in vals] for c, vals in m["categories"]}
    assert ini["weights"] == [[t["learning_rate"] for t in bag] for bag in m["bags"]]
    # the tree section: the bytes the reference wrote are the tail of the payload, and what our writer emits for the
    # parsed trees
    assert payload.endswith(trees)
    assert model_io.write_payload(m).endswith(trees)
    only_trees = model_io.write_payload(dict(m, numerical_mean=[], column_names=[], categories=[], column_mapping=[]))
    assert only_trees.endswith(trees)
