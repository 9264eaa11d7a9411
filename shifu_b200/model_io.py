"""Reader / writer of Shifu's binary tree-model file (`models/model*.gbt|rf`), format version 4.

Mirrors dt/BinaryDTSerializer.java:60-159 (header + maps + bagging list), dt/TreeNode.java:204-234,
dt/Node.java:583-624, dt/Split.java:153-174, dt/Predict.java:70-73 and dt/SimpleBitSet.java (Bytable write), so that
trees grown by the device path land where the unmodified `shifu eval` / IndependentTreeModel reads them
(dt/IndependentTreeModel.java:955-1080).  Java DataOutput = big-endian; writeUTF = 2-byte length + modified UTF-8;
the stream is gzip-compressed.  Row SURVEY.md 8(f3).
"""
import gzip
import io
import struct

TREE_FORMAT_VERSION = 4  # core/dtrain/CommonConstants.java:141
CONTINUOUS, CATEGORICAL = 1, 2
UTF_BYTES_MARKER = -1


class _In:
    def __init__(self, data):
        self.b = io.BytesIO(data)

    def _u(self, fmt, n):
        return struct.unpack(fmt, self.b.read(n))[0]

    def int(self): return self._u(">i", 4)
    def short(self): return self._u(">h", 2)
    def byte(self): return self._u(">b", 1)
    def bool(self): return self._u(">b", 1) != 0
    def float(self): return self._u(">f", 4)
    def double(self): return self._u(">d", 8)

    def utf(self):
        n = self._u(">H", 2)
        return self.b.read(n).decode("utf-8", errors="surrogatepass")

    def rest(self): return self.b.read()


class _Out:
    def __init__(self):
        self.b = io.BytesIO()

    def int(self, v): self.b.write(struct.pack(">i", v))
    def short(self, v): self.b.write(struct.pack(">h", v))
    def byte(self, v): self.b.write(struct.pack(">b", v))
    def bool(self, v): self.b.write(b"\x01" if v else b"\x00")
    def float(self, v): self.b.write(struct.pack(">f", v))
    def double(self, v): self.b.write(struct.pack(">d", v))

    def utf(self, s):
        raw = s.encode("utf-8", errors="surrogatepass")
        self.b.write(struct.pack(">H", len(raw)))
        self.b.write(raw)


def _read_node(i):
    n = {"id": i.int(), "gain": i.float(), "wgt_cnt": i.double(), "split": None, "predict": None, "left": None, "right": None}
    if i.bool():
        s = {"column_num": i.int(), "feature_type": i.byte()}
        if s["feature_type"] == CATEGORICAL:
            s["is_left"] = i.bool()
            if i.bool():
                s["categories"] = None
            else:
                ln = i.int()
                s["categories"] = bytes(i.b.read(ln))  # SimpleBitSet words
        else:
            s["threshold"] = i.double()
        n["split"] = s
    if i.bool():  # isRealLeaf
        if i.bool():
            n["predict"] = (i.double(), i.byte())
    if i.bool():
        n["left"] = _read_node(i)
    if i.bool():
        n["right"] = _read_node(i)
    return n


def _write_node(o, n):
    o.int(n["id"])
    o.float(n["gain"])  # cast to float to save space (dt/Node.java:586-587)
    o.double(n["wgt_cnt"])
    s = n["split"]
    o.bool(s is not None)
    if s is not None:
        o.int(s["column_num"])
        o.byte(s["feature_type"])
        if s["feature_type"] == CATEGORICAL:
            o.bool(s["is_left"])
            o.bool(s["categories"] is None)
            if s["categories"] is not None:
                o.int(len(s["categories"]))
                o.b.write(s["categories"])
        else:
            o.double(s["threshold"])
    real_leaf = n["left"] is None and n["right"] is None
    o.bool(real_leaf)
    if real_leaf:
        o.bool(n["predict"] is not None)
        if n["predict"] is not None:
            o.double(n["predict"][0])
            o.byte(n["predict"][1])
    for side in ("left", "right"):
        o.bool(n[side] is not None)
        if n[side] is not None:
            _write_node(o, n[side])


def read_payload(data):
    """Parses the un-gzipped stream into a dict."""
    i = _In(data)
    m = {"version": i.int(), "algorithm": i.utf(), "loss": i.utf(), "is_classification": i.bool(), "is_one_vs_all": i.bool(),
         "input_count": i.int()}
    m["numerical_mean"] = [(i.int(), i.double()) for _ in range(i.int())]
    m["column_names"] = [(i.int(), i.utf()) for _ in range(i.int())]
    cats = []
    for _ in range(i.int()):
        col, cnt = i.int(), i.int()
        vals = []
        for _ in range(cnt):
            ln = i.short()
            if ln == UTF_BYTES_MARKER:
                blen = i.int()
                vals.append(("bytes", i.b.read(blen)))
            else:
                vals.append(("utf", i.b.read(ln & 0xFFFF).decode("utf-8", errors="surrogatepass")))
        cats.append((col, vals))
    m["categories"] = cats
    m["column_mapping"] = [(i.int(), i.int()) for _ in range(i.int())]
    bags = []
    for _ in range(i.int()):
        trees = []
        for _ in range(i.int()):
            t = {"tree_id": i.int(), "node_num": i.int()}
            t["root"] = _read_node(i)
            t["learning_rate"] = i.double()
            if t["root"]["id"] == 1:
                t["root_wgt_cnt"] = i.double()
            t["features"] = [i.int() for _ in range(i.int())]
            trees.append(t)
        bags.append(trees)
    m["bags"] = bags
    m["trailing"] = len(i.rest())
    return m


def write_payload(m):
    o = _Out()
    o.int(m["version"]); o.utf(m["algorithm"]); o.utf(m["loss"]); o.bool(m["is_classification"]); o.bool(m["is_one_vs_all"])
    o.int(m["input_count"])
    o.int(len(m["numerical_mean"]))
    for k, v in m["numerical_mean"]:
        o.int(k); o.double(v)
    o.int(len(m["column_names"]))
    for k, v in m["column_names"]:
        o.int(k); o.utf(v)
    o.int(len(m["categories"]))
    for col, vals in m["categories"]:
        o.int(col); o.int(len(vals))
        for kind, v in vals:
            if kind == "bytes":
                o.short(UTF_BYTES_MARKER); o.int(len(v)); o.b.write(v)
            else:
                o.utf(v)
    o.int(len(m["column_mapping"]))
    for k, v in m["column_mapping"]:
        o.int(k); o.int(v)
    o.int(len(m["bags"]))
    for trees in m["bags"]:
        o.int(len(trees))
        for t in trees:
            o.int(t["tree_id"]); o.int(t["node_num"])
            _write_node(o, t["root"])
            o.double(t["learning_rate"])
            if t["root"]["id"] == 1:
                o.double(t["root_wgt_cnt"])
            o.int(len(t["features"]))
            for f in t["features"]:
                o.int(f)
    return o.b.getvalue()


def read_model(path):
    with gzip.open(path, "rb") as f:
        return read_payload(f.read())


def write_model(m, path):
    with gzip.open(path, "wb") as f:
        f.write(write_payload(m))


def tree_from_export(nodes, pool, extras, column_num, tree_id, learning_rate, node_num=None):
    """Builds the dict form of one tree from NativeMaster.export_tree output (pre-order sb_node records)."""
    def build(i):
        s = nodes[i]
        n = {"id": s.id, "gain": float(extras[i][0]), "wgt_cnt": float(extras[i][6]), "split": None, "predict": None,
             "left": None, "right": None}
        if s.feature >= 0:
            sp = {"column_num": int(column_num[s.feature]), "feature_type": s.split_type}
            if s.split_type == CATEGORICAL:
                sp["is_left"] = bool(s.is_left)
                sp["categories"] = bytes(pool[s.cat_offset:s.cat_offset + s.cat_bytes])
            else:
                sp["threshold"] = s.threshold
            n["split"] = sp
        if s.has_predict:
            n["predict"] = (s.predict, 0)
        if s.left >= 0:
            n["left"] = build(s.left)
        if s.right >= 0:
            n["right"] = build(s.right)
        return n
    root = build(0)
    return {"tree_id": tree_id, "node_num": node_num if node_num is not None else len(nodes), "root": root,
            "learning_rate": learning_rate, "root_wgt_cnt": float(extras[0][6]), "features": []}


def score_gbt(m, raw_row):
    """IndependentTreeModel.computeRegressionScore for one GBT bag (dt/IndependentTreeModel.java:387-446, predictNode
    :516-560): sum of learningRate * leaf predict; numeric goes left iff value < threshold."""
    mapping = dict(m["column_mapping"])
    score = 0.0
    for t in m["bags"][0]:
        n = t["root"]
        while n["split"] is not None and (n["left"] is not None or n["right"] is not None):
            s = n["split"]
            v = float(raw_row[mapping[s["column_num"]]])  # the scorer compares doubles (dt/IndependentTreeModel.java:516-560)
            if s["feature_type"] == CONTINUOUS:
                n = n["left"] if v < s["threshold"] else n["right"]
            else:
                idx = int(v)
                has = (idx >> 3) < len(s["categories"]) and (s["categories"][idx >> 3] >> (idx % 8)) & 1
                n = n["left"] if bool(has) == s["is_left"] else n["right"]
        score += t["learning_rate"] * n["predict"][0]
    return score


def to_score_model(m, convert="RAW"):
    """The in-memory IndependentTreeModel of a parsed model file as a capi.ScoreModel (sb_score_model): split columns
    become positions in a data row through the file's columnNum -> index mapping (dt/IndependentTreeModel.java:566-568),
    categorical sizes come from the file's category lists (:570-573), tree weights are the trees' learning rates."""
    from .capi import ScoreModel
    mapping = dict(m["column_mapping"])
    n_columns = max(mapping.values()) + 1 if mapping else 1
    cat_size = [0] * n_columns
    for col, vals in m["categories"]:
        if col in mapping:
            cat_size[mapping[col]] = len(vals)

    def conv(n):
        sp = n["split"]
        out = {"split": None, "left": None, "right": None, "predict": n["predict"]}
        if sp is not None:
            o = {"column": mapping[sp["column_num"]], "feature_type": sp["feature_type"]}
            if sp["feature_type"] == CATEGORICAL:
                o["is_left"] = sp["is_left"]
                o["categories"] = sp["categories"] or b""
            else:
                o["threshold"] = sp["threshold"]
            out["split"] = o
        if n["left"] is not None:
            out["left"] = conv(n["left"])
        if n["right"] is not None:
            out["right"] = conv(n["right"])
        return out

    bags = [[(conv(t["root"]), t["learning_rate"]) for t in bag] for bag in m["bags"]]
    return ScoreModel(bags, n_columns, cat_size, is_gbdt=m["algorithm"].upper() == "GBT",
                      is_classification=m["is_classification"], convert=convert)
