"""f1, categorical half: the binCategory -> bin-index map at load (sb_category_map_*) against the oracle's restatement
of CommonUtils.flattenCatValGrp / DTWorker.init / DTWorker.load, on the reference's own labor-neg ColumnConfig and
data (tests/golden/labor_neg_categories.json) and on grouped categories.  Host code on both sides: no GPU needed."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from shifu_b200.capi import CategoryMap

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module", autouse=True)
def built():
    import __graft_entry__ as g
    g.build()


def oracle_index(lib, cats, value):
    arr = (C.c_char_p * max(len(cats), 1))(*[c.encode() for c in cats])
    return lib.orc_category_index(arr, len(cats), None if value is None else value.encode())


def test_group_delimiter_known_answers_of_the_reference(oracle_lib):
    """src/test/java/ml/shifu/shifu/core/dtrain/IndependentTreeModelTest.java:20-26 (testSplit, live): the "@^" split
    keeps empty pieces and does not treat '@' or '^' alone as a delimiter."""
    def pieces(group):
        buf = C.create_string_buffer(256)
        n = oracle_lib.orc_flatten_cat_val_grp(group.encode(), buf, 256)
        got = buf.value.decode().split("\n") if n > 0 else []
        assert len(got) == n or (n == 1 and got == [""])
        return got

    assert pieces("aa@^bb@cc")[0] == "aa" and pieces("aa@^bb@cc")[1] == "bb@cc"
    assert pieces("aa@bb@cc")[0] == "aa@bb@cc"
    assert pieces("aa@^bb@^")[2] == ""
    assert pieces("aa@^bb@^@")[2] == "@"
    assert pieces("@^bb@^@")[2] == "@"
    assert pieces("   ") == []  # StringUtils.isNotBlank


def test_labor_neg_columns_bin_like_the_reference(oracle_lib):
    cols = json.load(open(os.path.join(HERE, "golden", "labor_neg_categories.json")))
    assert len(cols) == 8
    for c in cols:
        m = CategoryMap(c["binCategory"])
        got = m.lookup(c["values"])
        m.close()
        want = [oracle_index(oracle_lib, c["binCategory"], v) for v in c["values"]]
        assert got.tolist() == want == c["expected"], c["columnName"]
        assert max(want) <= len(c["binCategory"])


def test_groups_duplicates_unknown_and_empty_values(oracle_lib):
    cats = ["us@^ca", "de@^fr@^", "jp", "us", "  ", "a^b", "@"]
    values = ["us", "ca", "de", "fr", "jp", "", None, "xx", "a^b", "@", "a", " ", "US"]
    m = CategoryMap(cats)
    got = m.lookup(values).tolist()
    m.close()
    want = [oracle_index(oracle_lib, cats, v) for v in values]
    assert got == want
    assert got[0] == 3          # "us" listed twice: the later put wins
    assert got[5] == got[6] == got[7] == len(cats)  # empty, null, unknown -> the missing-value bin
    # more than 255 categories -> uint16 output
    many = ["c%d" % i for i in range(1000)]
    m = CategoryMap(many)
    out = m.lookup(["c999", "c0", "nope"])
    assert out.dtype == np.uint16 and out.tolist() == [999, 0, 1000]
    m.close()
