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


def _walk(n):
    yield n
    for c in ("left", "right"):
        if n[c] is not None:
            yield from _walk(n[c])


def test_second_reference_fixture_round_trips_too():
    """src/test/resources/example/readablespec/model1.gbt (copied unchanged).  In the reference tree it is byte for byte
    the same file as model0.gbt, so it pins nothing new -- it is read and re-written here so that no reference
    fixture of the path is left unused."""
    other = os.path.join(HERE, "golden", "readablespec_model1.gbt")
    payload = gzip.open(other, "rb").read()
    m = model_io.read_payload(payload)
    assert m["trailing"] == 0 and model_io.write_payload(m) == payload
    assert open(other, "rb").read() == open(FIXTURE, "rb").read()


def test_invariants_the_reference_written_trees_allow():
    """What the 100 reference-grown trees of model0.gbt pin about DTMaster / Impurity, given what Node.write keeps
    (dt/Node.java:583-624: id, gain as float, wgtCnt, split; predict only for real leaves):
      * wgtCnt is what populateGainInfoToNode stored (dt/DTMaster.java:718): wherever both children of a split node
        were evaluated themselves, parent.wgtCnt == left.wgtCnt + right.wgtCnt exactly (unweighted rows: integers),
        and TreeNode.rootWgtCnt is the root's wgtCnt (:720-722);
      * a node was split only with gain > 0 (:350), heap ids follow Node.leftIndex / rightIndex, and no node sits below
        MaxDepth; children that were never evaluated (deepest level, or pure) carry wgtCnt 0 and a predict.
    What CANNOT be rebuilt from the file: the gain values themselves.  Variance gain needs the children's weighted
    sums, and the file stores neither the per-node sums nor the counts of leaves (every real leaf has wgtCnt 0), so
    orc_compute_impurity has no statistics to be run on; the float gains are only checked for their sign.  Parity of
    gains therefore stays pinned to the restated formulas only."""
    m = model_io.read_model(FIXTURE)
    trees = m["bags"][0]
    checked = 0
    max_level = 0
    for t in trees:
        root = t["root"]
        assert root["id"] == 1 and t["root_wgt_cnt"] == root["wgt_cnt"] == 348.0
        assert t["node_num"] == sum(1 for _ in _walk(root))
        for n in _walk(root):
            max_level = max(max_level, int(n["id"]).bit_length())
            if n["left"] is not None or n["right"] is not None:
                assert n["left"] is not None and n["right"] is not None and n["split"] is not None
                assert n["left"]["id"] == 2 * n["id"] and n["right"]["id"] == 2 * n["id"] + 1
                assert n["gain"] > 0.0 and n["wgt_cnt"] > 0.0 and n["predict"] is None
                l, r = n["left"], n["right"]
                if l["wgt_cnt"] > 0 and r["wgt_cnt"] > 0:
                    assert l["wgt_cnt"] + r["wgt_cnt"] == n["wgt_cnt"]
                    checked += 1
                else:
                    assert l["wgt_cnt"] + r["wgt_cnt"] <= n["wgt_cnt"]
            else:
                assert n["split"] is None and n["predict"] is not None and n["gain"] == 0.0 and n["wgt_cnt"] == 0.0
    assert checked == 709 and max_level <= 7
