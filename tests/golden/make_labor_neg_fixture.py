"""Builds tests/golden/labor_neg_categories.json from the reference's own fixture (run in the build container only):
  /root/reference/src/test/resources/example/labor-neg/DataStore/DataSet1/ColumnConfig.json  (8 categorical columns)
  .../data.dat                                                                               (57 rows, ','-delimited)
For every categorical column: ColumnConfig.binCategory as the reference's stats step wrote it, and the raw values of
that column.  The expected bin of a value is re-derived independently here (a dict built in binCategory order, the
missing bin for '' and for values not listed) so that the fixture does not depend on the oracle or on the product.
"""
import json
import os

REF = "/root/reference/src/test/resources/example/labor-neg/DataStore/DataSet1"
HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    cc = json.load(open(os.path.join(REF, "ColumnConfig.json")))
    lines = [l.rstrip("\n").split(",") for l in open(os.path.join(REF, "data.dat")) if l.strip()]
    rows = lines[1:]  # header first
    out = []
    for c in cc:
        cats = c["columnBinning"].get("binCategory")
        if not cats:
            continue
        values = [r[c["columnNum"]] for r in rows]
        table = {}
        for i, grp in enumerate(cats):
            for member in (grp.split("@^") if grp.strip() else []):
                table[member] = i
        expected = [len(cats) if v == "" else table.get(v, len(cats)) for v in values]
        out.append({"columnNum": c["columnNum"], "columnName": c["columnName"], "binCategory": cats, "values": values,
                    "expected": expected})
    json.dump(out, open(os.path.join(HERE, "labor_neg_categories.json"), "w"), indent=0)
    print("%d categorical columns, %d rows" % (len(out), len(rows)))


if __name__ == "__main__":
    main()
