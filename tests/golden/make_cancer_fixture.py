"""Builds tests/golden/cancer_judgement.npz from the reference's own fixture (run in the build container only):
  /root/reference/src/test/resources/example/cancer-judgement/DataStore/DataSet1/part-00  (429 rows, '|'-delimited)
  .../ModelStore/ModelSet1/ColumnConfig.json                                            (30 numeric columns x 10 bins)
Binning follows DTWorker.getBinIndex (dt/DTWorker.java:1001-1034), restated here with numpy.searchsorted so that it
is independent of the oracle's binary search (tests compare the two).  label: M -> 1, B -> 0 (ModelConfig posTags).
"""
import json
import os

import numpy as np

REF = "/root/reference/src/test/resources/example/cancer-judgement"
HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    cc = json.load(open(os.path.join(REF, "ModelStore/ModelSet1/ColumnConfig.json")))
    feats = [c for c in cc if c["finalSelect"] and c["columnFlag"] is None]
    rows = [l.rstrip("\n").split("|") for l in open(os.path.join(REF, "DataStore/DataSet1/part-00")) if l.strip()]
    label = np.array([1.0 if r[0] == "M" else 0.0 for r in rows], dtype=np.float32)
    bins, boundaries, raw = [], [], []
    for c in feats:
        bb = np.array([float(x) for x in c["columnBinning"]["binBoundary"]], dtype=np.float64)
        v = np.array([np.float32(r[c["columnNum"]]) for r in rows], dtype=np.float32)
        # value in [bb[i], bb[i+1]) -> i ; last bin is open-ended
        idx = np.searchsorted(bb, v.astype(np.float64), side="right") - 1
        idx = np.clip(idx, 0, len(bb) - 1)
        bins.append(idx.astype(np.uint8))
        boundaries.append(bb)
        raw.append(v)
    np.savez_compressed(os.path.join(HERE, "cancer_judgement.npz"), bins=np.stack(bins), label=label,
                        boundaries=np.stack(boundaries), raw=np.stack(raw),
                        column_num=np.array([c["columnNum"] for c in feats], dtype=np.int32),
                        weight=np.stack(raw)[0])  # column_3 is ModelConfig's weightColumnName
    print("rows", len(rows), "features", len(feats), "pos", int(label.sum()))


if __name__ == "__main__":
    main()
