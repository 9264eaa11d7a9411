// wire.cpp -- host-only helpers on the boundary (no CUDA in this file):
//   * sb_msg_pack / sb_msg_unpack: DTMasterParams as ONE position-independent byte buffer, the form in which a JVM
//     hands it over (a direct ByteBuffer written by jni/DTMessageFlattener.java, unpacked by jni/shifu_b200_jni.c).
//     The layout is documented in include/shifu_b200.h ("Wire format of sb_iter_in").
//   * sb_category_map_*: the categorical half of binning at load -- the value -> bin-index map DTWorker.init builds
//     from ColumnConfig.binCategory (dt/DTWorker.java:362-376, CommonUtils.flattenCatValGrp util/CommonUtils.java:1767)
//     and the lookup DTWorker.load does per value (dt/DTWorker.java:1152-1169).
#include <cstring>
#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/shifu_b200.h"

namespace sb { int fail(int code, const std::string& msg); }

namespace {

constexpr uint32_t WIRE_MAGIC = 0x314D4253u;  // "SBM1", little endian
constexpr int64_t WIRE_HEADER = 40;

inline int64_t pad8(int64_t n) { return (n + 7) & ~(int64_t)7; }

struct Reader {
    const uint8_t* p;
    int64_t len, off = 0;
    bool ok = true;
    const void* take(int64_t bytes) {
        if (!ok || bytes < 0 || off + bytes > len) { ok = false; return nullptr; }
        const void* r = p + off;
        off += bytes;
        return r;
    }
    int32_t i32() {
        int32_t v = 0;
        const void* q = take(4);
        if (q) std::memcpy(&v, q, 4);
        return v;
    }
};

}  // namespace

struct sb_msg {
    std::vector<sb_tree> trees;
    std::vector<sb_todo_node> todo;
    sb_iter_in in{};
};

struct sb_category_map {
    std::unordered_map<std::string, int32_t> index;  // Map<String, Integer> tmpMap of dt/DTWorker.java:366-372
    int32_t n_categories = 0;                        // binCategory.size(): the bin of missing / unknown values
};

extern "C" {

SB_API int64_t sb_msg_pack(const sb_iter_in* in, int64_t n_rows, void* buf, int64_t cap) {
    if (!in || in->n_trees < 0 || in->n_todo < 0 || n_rows < 0) return sb::fail(SB_ERR_INVALID, "sb_msg_pack: bad argument");
    int64_t need = WIRE_HEADER;
    for (int t = 0; t < in->n_trees; ++t)
        need += 16 + (int64_t)in->trees[t].n_nodes * (int64_t)sizeof(sb_node) + pad8(in->trees[t].mask_pool_bytes);
    for (int i = 0; i < in->n_todo; ++i) need += 16 + pad8((int64_t)in->todo[i].n_features * 4);
    if (in->new_weights) need += pad8(n_rows * 4);
    if (in->dropout) need += pad8(n_rows);
    if (!buf) return need;
    if (cap < need) return sb::fail(SB_ERR_INVALID, "sb_msg_pack: buffer too small");
    if (need > 0xffffffffLL) return sb::fail(SB_ERR_UNSUPPORTED, "sb_msg_pack: message beyond 4 GB");
    uint8_t* o = static_cast<uint8_t*>(buf);
    std::memset(o, 0, (size_t)need);
    auto put32 = [&](int64_t at, int32_t v) { std::memcpy(o + at, &v, 4); };
    const uint32_t magic = WIRE_MAGIC, total = (uint32_t)need;
    std::memcpy(o, &magic, 4);
    std::memcpy(o + 4, &total, 4);
    put32(8, in->n_trees);
    put32(12, in->n_todo);
    put32(16, in->switch_to_next_tree);
    put32(20, in->is_first_tree);
    put32(24, in->continuous_running_start);
    put32(28, (in->new_weights ? 1 : 0) | (in->dropout ? 2 : 0));
    std::memcpy(o + 32, &n_rows, 8);
    int64_t at = WIRE_HEADER;
    for (int t = 0; t < in->n_trees; ++t) {
        const sb_tree& tr = in->trees[t];
        put32(at, tr.tree_id);
        put32(at + 4, tr.n_nodes);
        put32(at + 8, tr.mask_pool_bytes);
        at += 16;
        if (tr.n_nodes > 0) std::memcpy(o + at, tr.nodes, (size_t)tr.n_nodes * sizeof(sb_node));
        at += (int64_t)tr.n_nodes * (int64_t)sizeof(sb_node);
        if (tr.mask_pool_bytes > 0) std::memcpy(o + at, tr.mask_pool, (size_t)tr.mask_pool_bytes);
        at += pad8(tr.mask_pool_bytes);
    }
    for (int i = 0; i < in->n_todo; ++i) {
        const sb_todo_node& td = in->todo[i];
        put32(at, td.key);
        put32(at + 4, td.tree_id);
        put32(at + 8, td.node_id);
        put32(at + 12, td.n_features);
        at += 16;
        if (td.n_features > 0) std::memcpy(o + at, td.features, (size_t)td.n_features * 4);
        at += pad8((int64_t)td.n_features * 4);
    }
    if (in->new_weights) {
        std::memcpy(o + at, in->new_weights, (size_t)n_rows * 4);
        at += pad8(n_rows * 4);
    }
    if (in->dropout) {
        std::memcpy(o + at, in->dropout, (size_t)n_rows);
        at += pad8(n_rows);
    }
    return at;
}

SB_API int sb_msg_unpack(const void* buf, int64_t len, sb_msg** out) {
    if (!buf || !out || len < WIRE_HEADER) return sb::fail(SB_ERR_INVALID, "sb_msg_unpack: null or truncated buffer");
    if (((uintptr_t)buf & 7) != 0) return sb::fail(SB_ERR_INVALID, "sb_msg_unpack: the buffer must be 8-byte aligned");
    *out = nullptr;
    Reader r{static_cast<const uint8_t*>(buf), len};
    uint32_t magic, total;
    std::memcpy(&magic, r.take(4), 4);
    std::memcpy(&total, r.take(4), 4);
    if (magic != WIRE_MAGIC) return sb::fail(SB_ERR_INVALID, "sb_msg_unpack: bad magic (expected \"SBM1\")");
    if ((int64_t)total > len) return sb::fail(SB_ERR_INVALID, "sb_msg_unpack: buffer shorter than the message says");
    r.len = total;
    std::unique_ptr<sb_msg> m(new sb_msg());
    const int32_t n_trees = r.i32(), n_todo = r.i32();
    m->in.switch_to_next_tree = r.i32();
    m->in.is_first_tree = r.i32();
    m->in.continuous_running_start = r.i32();
    const uint32_t has = (uint32_t)r.i32();
    int64_t n_rows = 0;
    std::memcpy(&n_rows, r.take(8), 8);
    if (n_trees < 0 || n_todo < 0 || n_rows < 0) return sb::fail(SB_ERR_INVALID, "sb_msg_unpack: negative count");
    m->trees.resize((size_t)n_trees);
    for (int t = 0; t < n_trees; ++t) {
        sb_tree& tr = m->trees[(size_t)t];
        tr.tree_id = r.i32();
        tr.n_nodes = r.i32();
        tr.mask_pool_bytes = r.i32();
        r.i32();
        if (tr.n_nodes < 0 || tr.mask_pool_bytes < 0) return sb::fail(SB_ERR_INVALID, "sb_msg_unpack: negative tree size");
        tr.nodes = static_cast<const sb_node*>(r.take((int64_t)tr.n_nodes * (int64_t)sizeof(sb_node)));
        tr.mask_pool = static_cast<const uint8_t*>(r.take(pad8(tr.mask_pool_bytes)));
        if (!r.ok) return sb::fail(SB_ERR_INVALID, "sb_msg_unpack: tree section runs past the end of the buffer");
        for (int i = 0; i < tr.n_nodes; ++i) {  // indices and offsets must stay inside the tree
            const sb_node& n = tr.nodes[i];
            if (n.left >= tr.n_nodes || n.right >= tr.n_nodes || n.left < -1 || n.right < -1)
                return sb::fail(SB_ERR_INVALID, "sb_msg_unpack: child index outside the tree");
            if (n.feature >= 0 && (n.left_mask_offset < 0 || n.left_mask_bytes < 0 || n.left_mask_offset + n.left_mask_bytes > tr.mask_pool_bytes ||
                                   n.cat_offset < 0 || n.cat_bytes < 0 || n.cat_offset + n.cat_bytes > tr.mask_pool_bytes))
                return sb::fail(SB_ERR_INVALID, "sb_msg_unpack: mask outside the tree's pool");
        }
    }
    m->todo.resize((size_t)n_todo);
    for (int i = 0; i < n_todo; ++i) {
        sb_todo_node& td = m->todo[(size_t)i];
        td.key = r.i32();
        td.tree_id = r.i32();
        td.node_id = r.i32();
        td.n_features = r.i32();
        if (td.n_features < 0) return sb::fail(SB_ERR_INVALID, "sb_msg_unpack: negative feature count");
        td.features = static_cast<const int32_t*>(r.take(pad8((int64_t)td.n_features * 4)));
        if (!r.ok) return sb::fail(SB_ERR_INVALID, "sb_msg_unpack: todo section runs past the end of the buffer");
    }
    if (has & 1u) m->in.new_weights = static_cast<const float*>(r.take(pad8(n_rows * 4)));
    if (has & 2u) m->in.dropout = static_cast<const uint8_t*>(r.take(pad8(n_rows)));
    if (!r.ok) return sb::fail(SB_ERR_INVALID, "sb_msg_unpack: per-row section runs past the end of the buffer");
    m->in.n_trees = n_trees;
    m->in.trees = m->trees.data();
    m->in.n_todo = n_todo;
    m->in.todo = m->todo.data();
    *out = m.release();
    return SB_OK;
}

SB_API const sb_iter_in* sb_msg_iter_in(const sb_msg* m) { return m ? &m->in : nullptr; }
SB_API void sb_msg_free(sb_msg* m) { delete m; }

SB_API int sb_category_map_create(const char* const* bin_category, int32_t n_categories, sb_category_map** out) {
    if (!out || n_categories < 0 || (n_categories > 0 && !bin_category)) return sb::fail(SB_ERR_INVALID, "sb_category_map_create: bad argument");
    if (n_categories > 32767) return sb::fail(SB_ERR_RANGE, "sb_category_map_create: more than Short.MAX_VALUE categories");
    std::unique_ptr<sb_category_map> m(new sb_category_map());
    m->n_categories = n_categories;
    static const std::string DELIM = "@^";  // Constants.CATEGORICAL_GROUP_VAL_DELIMITER (util/Constants.java:280)
    for (int32_t i = 0; i < n_categories; ++i) {
        if (!bin_category[i]) return sb::fail(SB_ERR_INVALID, "sb_category_map_create: null category string");
        const std::string grp(bin_category[i]);
        // StringUtils.isNotBlank: a group of whitespace only has no members
        bool blank = true;
        for (unsigned char c : grp)
            if (!(c == ' ' || (c >= 9 && c <= 13) || (c >= 28 && c <= 31))) { blank = false; break; }  // Character.isWhitespace, ASCII range
        if (blank) continue;
        // Guava Splitter.on("@^").split(grp): every piece counts, empty ones included; a later put of the same
        // value replaces the earlier index (HashMap.put)
        size_t from = 0;
        for (;;) {
            const size_t at = grp.find(DELIM, from);
            m->index[grp.substr(from, at == std::string::npos ? std::string::npos : at - from)] = i;
            if (at == std::string::npos) break;
            from = at + DELIM.size();
        }
    }
    *out = m.release();
    return SB_OK;
}

SB_API int sb_category_map_lookup(const sb_category_map* m, const char* const* values, int64_t n, void* bins_out, int32_t out_width) {
    if (!m || n < 0 || (n > 0 && (!values || !bins_out))) return sb::fail(SB_ERR_INVALID, "sb_category_map_lookup: bad argument");
    if (out_width != 1 && out_width != 2) return sb::fail(SB_ERR_INVALID, "sb_category_map_lookup: out_width must be 1 (uint8) or 2 (uint16)");
    if (out_width == 1 && m->n_categories + 1 > 256) return sb::fail(SB_ERR_INVALID, "sb_category_map_lookup: more than 256 bins need uint16 output");
    for (int64_t i = 0; i < n; ++i) {
        int32_t bin = m->n_categories;  // empty input, and a value that is in no group (dt/DTWorker.java:1153-1168)
        const char* v = values[i];
        if (v && v[0] != '\0') {
            auto it = m->index.find(v);
            if (it != m->index.end()) bin = it->second;
        }
        if (out_width == 2) static_cast<uint16_t*>(bins_out)[i] = (uint16_t)bin;
        else static_cast<uint8_t*>(bins_out)[i] = (uint8_t)bin;
    }
    return SB_OK;
}

SB_API void sb_category_map_destroy(sb_category_map* m) { delete m; }

}  // extern "C"
