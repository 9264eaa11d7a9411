// score.cu -- batch scoring of a tree ensemble: IndependentTreeModel.compute (dt/IndependentTreeModel.java:352-446)
// and predictNode (:516-560) over a block of rows of raw values (the step after training).
//
// One thread scores one row.  The rows of a block are staged in shared memory with coalesced loads (a data row is
// read by one thread, column after column, as the trees ask for them).  The model is re-laid out at creation: every
// tree in pre-order (the left child of node i is node i + 1), 16 bytes per node {threshold | leaf value | bitset
// reference, right child, column, flags}, so a step of the walk is one 16-byte load -- from shared memory when the
// nodes fit there beside the rows (100 trees of the reference's fixture = 69 KB), else through L1.  The walk and the
// accumulation follow the reference's loop order in fp64 without contraction (-fmad=false), so raw and cut-off
// scores are bit-identical to the Java loop; the sigmoid converts go through exp().
#include <algorithm>
#include <cmath>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "common.cuh"
#include "../../include/shifu_b200.h"

namespace sb {

namespace {

constexpr int SC_ROWS = 128;           // rows (= threads) per block
constexpr int SC_MAX_SMEM_COLS = 200;  // widest data row staged in shared memory (200 * 128 * 8 B = 200 KB)

// flags of a packed node
constexpr uint32_t PN_WALK = 1u;    // has a split and at least one child: predictNode goes on
constexpr uint32_t PN_CAT = 2u;     // categorical split: value = (bitset offset, bitset bytes)
constexpr uint32_t PN_ISLEFT = 4u;  // Split.isLeft
constexpr uint32_t PN_LEFT = 8u;    // left child exists (at index + 1)
constexpr uint32_t PN_RIGHT = 16u;  // right child exists (at .right)

struct __align__(16) PackedNode {
    double value;    // numeric split: threshold; leaf: Predict.predict (classification: the class value); categorical: bitset ref
    int32_t right;
    int16_t column;
    uint16_t flags;
};

struct ScoreView {
    const int32_t* bag_begin;
    const int32_t* tree_root;
    const double* weights;
    const PackedNode* nodes;
    const uint8_t* cat_pool;
    const int32_t* categorical_size;
    int n_bags, n_trees, n_nodes, cat_pool_bytes, is_gbdt, is_classification, convert, n_columns;
};

// Double.compare(a, b) < 0 with Java's total order (-0.0 < 0.0, NaN above everything)
__device__ __forceinline__ bool java_compare_lt(double a, double b) {
    if (a < b) return true;
    if (a > b) return false;
    if (a == b) return a == 0.0 && signbit(a) && !signbit(b);
    return !isnan(a) && isnan(b);
}

// predictNode (dt/IndependentTreeModel.java:516-560); returns the value of the node it stops at
template <bool NODES_IN_SMEM, class RowReader>
__device__ __forceinline__ double predict_node(const ScoreView& m, int node, const RowReader& row, uint32_t* error_flag) {
    while (true) {
        int4 raw;
        if (NODES_IN_SMEM) raw = *reinterpret_cast<const int4*>(m.nodes + node);  // m.nodes points into shared memory
        else raw = __ldg(reinterpret_cast<const int4*>(m.nodes + node));
        const double nvalue = __hiloint2double(raw.y, raw.x);
        const uint32_t flags = (uint32_t)raw.w >> 16;
        if (!(flags & PN_WALK)) return nvalue;  // split == null or isRealLeaf()
        const int column = (int)(short)(raw.w & 0xffff);
        const double value = row(column);
        bool go_left;
        if (!(flags & PN_CAT)) {
            go_left = value < nvalue;
        } else {
            const int size = m.categorical_size[column];
            int idx;
            // Double.compare(value, 0d) < 0 || Double.compare(value, categoricalSize) >= 0  -> the missing category
            if (java_compare_lt(value, 0.0) || !java_compare_lt(value, (double)size)) idx = (int)(short)size;
            else idx = (int)(short)(int)(value + 0.1);  // (short)(value + 0.1d): d2i then i2s
            bool has = false;
            if (idx >= 0 && (idx >> 3) < raw.y) has = (m.cat_pool[raw.x + (idx >> 3)] >> (idx & 7)) & 1;
            go_left = has == ((flags & PN_ISLEFT) != 0);
        }
        if (!(flags & (go_left ? PN_LEFT : PN_RIGHT))) {  // the reference would throw a NullPointerException here
            atomicOr(error_flag, 1u);
            return nvalue;
        }
        node = go_left ? node + 1 : raw.z;
    }
}

__device__ __forceinline__ double convert_score(int convert, double predict) {
    switch (convert) {
        case SB_CONVERT_OLD_SIGMOID: return 1 / (1 + fmin(1.0E19, exp(-predict)));
        case SB_CONVERT_SIGMOID: return 1 / (1 + fmin(1.0E19, exp(-20 * predict)));
        case SB_CONVERT_CUTOFF: return predict < 0.0 ? 0.0 : (predict > 1.0 ? 1.0 : predict);
    }
    return predict;
}

template <bool NODES_IN_SMEM, class RowReader>
__device__ __forceinline__ void score_row(const ScoreView& m, const RowReader& row, double* out, uint32_t* error_flag) {
    if (m.is_classification) {  // computeClassificationScore (:365-379): every tree's class value
        for (int t = 0; t < m.n_trees; ++t) out[t] = predict_node<NODES_IN_SMEM>(m, m.tree_root[t], row, error_flag);
        return;
    }
    double final_predict = 0.0;
    for (int b = 0; b < m.n_bags; ++b) {
        if (m.is_gbdt) {  // :390-420
            double predict = 0.0;
            for (int t = m.bag_begin[b]; t < m.bag_begin[b + 1]; ++t)
                predict += predict_node<NODES_IN_SMEM>(m, m.tree_root[t], row, error_flag) * m.weights[t];
            final_predict += convert_score(m.convert, predict);
        } else {          // :423-441
            double predict_sum = 0.0, weight_sum = 0.0;
            for (int t = m.bag_begin[b]; t < m.bag_begin[b + 1]; ++t) {
                const double score = predict_node<NODES_IN_SMEM>(m, m.tree_root[t], row, error_flag);
                const double weight = m.weights[t];
                weight_sum += weight;
                predict_sum += score * weight;
            }
            final_predict += predict_sum / weight_sum;
        }
    }
    out[0] = final_predict / m.n_bags;
}

struct SmemRow {
    const double* p;
    __device__ __forceinline__ double operator()(int c) const { return p[c]; }
};
struct GlobalRow {
    const double* p;
    __device__ __forceinline__ double operator()(int c) const { return __ldg(p + c); }
};

// STAGED: the block's rows x n_columns doubles are copied to shared memory first (contiguous in global memory:
// fully coalesced), each row at an odd stride so that the threads of a warp reading "their" column hit distinct banks.
// NODES_IN_SMEM: the packed nodes are copied to shared memory as well.  The 32 lanes of a warp sit in 32 different nodes,
// so through L1 a step of the walk is 32 tag lookups (measured 1.1 node visits per cycle and SM); in shared memory it
// is a few bank conflicts.  Used when nodes + rows fit one block's shared memory.
template <bool STAGED, bool NODES_IN_SMEM>
__global__ void score_kernel(ScoreView m, const double* __restrict__ data, int64_t n_rows, double* __restrict__ scores,
                             uint32_t* error_flag) {
    extern __shared__ __align__(16) unsigned char score_smem[];
    const int rows_per_block = blockDim.x;
    double* tile = reinterpret_cast<double*>(score_smem);
    if (NODES_IN_SMEM) {
        int4* dst = reinterpret_cast<int4*>(score_smem);
        const int4* src = reinterpret_cast<const int4*>(m.nodes);
        for (int i = threadIdx.x; i < m.n_nodes; i += blockDim.x) dst[i] = __ldg(src + i);
        m.nodes = reinterpret_cast<const PackedNode*>(score_smem);
        tile = reinterpret_cast<double*>(score_smem + (size_t)m.n_nodes * sizeof(PackedNode));
    }
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    const int64_t r = r0 + threadIdx.x;
    const int out_per_row = m.is_classification ? m.n_trees : 1;
    if (STAGED) {
        const int stride = m.n_columns | 1;
        const int64_t rows_here = min((int64_t)rows_per_block, n_rows - r0);
        const int64_t total = rows_here * m.n_columns;
        const double* src = data + r0 * m.n_columns;
        for (int64_t i = threadIdx.x; i < total; i += rows_per_block) {
            const int rr = (int)(i / m.n_columns), c = (int)(i - (int64_t)rr * m.n_columns);
            tile[rr * stride + c] = src[i];
        }
        __syncthreads();
        if (r < n_rows) score_row<NODES_IN_SMEM>(m, SmemRow{tile + threadIdx.x * stride}, scores + r * out_per_row, error_flag);
    } else {
        if (r < n_rows) score_row<false>(m, GlobalRow{data + r * m.n_columns}, scores + r * out_per_row, error_flag);
    }
}

}  // namespace

}  // namespace sb

using namespace sb;

struct sb_scorer {
    int device = 0;
    ScoreView view{};
    void* blob = nullptr;  // every model array in one device allocation
    uint32_t* d_error = nullptr;
    double* d_data = nullptr;
    double* d_scores = nullptr;
    size_t data_cap = 0, scores_cap = 0;
    float last_ms = 0.f;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
};

extern "C" {

SB_API int sb_scorer_create(const sb_score_model* m, int32_t device, sb_scorer** out) {
    if (!m || !out) return fail(SB_ERR_INVALID, "null argument");
    *out = nullptr;
    if (m->n_bags < 1 || m->n_trees < 1 || m->n_nodes < 1 || m->n_columns < 1 || !m->bag_begin || !m->tree_root || !m->weights ||
        !m->nodes || !m->categorical_size)
        return fail(SB_ERR_INVALID, "empty model");
    if (m->convert < SB_CONVERT_RAW || m->convert > SB_CONVERT_CUTOFF) return fail(SB_ERR_INVALID, "unknown convert");
    if (m->bag_begin[0] != 0 || m->bag_begin[m->n_bags] != m->n_trees) return fail(SB_ERR_INVALID, "bag_begin does not cover the trees");
    for (int t = 0; t < m->n_trees; ++t)
        if (m->tree_root[t] < 0 || m->tree_root[t] >= m->n_nodes) return fail(SB_ERR_INVALID, "tree root out of range");
    for (int i = 0; i < m->n_nodes; ++i) {
        const sb_score_node& nd = m->nodes[i];
        if (nd.left >= m->n_nodes || nd.right >= m->n_nodes || nd.column >= m->n_columns) return fail(SB_ERR_INVALID, "node index or column out of range");
        if (nd.column >= 0 && nd.split_type == SB_SPLIT_CATEGORICAL &&
            (nd.cat_bytes < 0 || nd.cat_offset < 0 || (int64_t)nd.cat_offset + nd.cat_bytes > m->cat_pool_bytes))
            return fail(SB_ERR_INVALID, "category bitset outside the pool");
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(SB_ERR_CUDA, "no CUDA device: shifu_b200 has no CPU fallback");
    }
    if (device < 0 || device >= ndev) return fail(SB_ERR_INVALID, "device ordinal out of range");
    SB_CUDA(cudaSetDevice(device));
    std::unique_ptr<sb_scorer> s(new sb_scorer());
    s->device = device;
    // pre-order re-layout of every tree: the left child of packed node i is packed node i + 1
    std::vector<PackedNode> packed;
    std::vector<int32_t> roots((size_t)m->n_trees);
    packed.reserve((size_t)m->n_nodes);
    for (int t = 0; t < m->n_trees; ++t) {
        roots[t] = (int32_t)packed.size();
        // explicit stack of (source node, packed index of the parent whose .right must be patched, or -1)
        std::vector<std::pair<int, int>> stack{{m->tree_root[t], -1}};
        while (!stack.empty()) {
            const auto [src, patch] = stack.back();
            stack.pop_back();
            if (packed.size() > (size_t)m->n_nodes * 2 + 16 || packed.size() >= (size_t)0x7fffffff)
                return fail(SB_ERR_INVALID, "the node links of a tree form a cycle (or a node is shared by trees more than once per node)");
            const sb_score_node& nd = m->nodes[src];
            const int me = (int)packed.size();
            if (patch >= 0) packed[patch].right = me;
            PackedNode pn;
            pn.right = -1;
            pn.column = (int16_t)-1;
            uint32_t flags = 0;
            const bool walk = nd.column >= 0 && (nd.left >= 0 || nd.right >= 0);
            if (walk) {
                if (nd.column > 32767) return fail(SB_ERR_UNSUPPORTED, "more than 32767 data columns");
                flags |= PN_WALK;
                pn.column = (int16_t)nd.column;
                if (nd.split_type == SB_SPLIT_CATEGORICAL) {
                    flags |= PN_CAT;
                    if (nd.is_left) flags |= PN_ISLEFT;
                    const int32_t ref[2] = {nd.cat_offset, nd.cat_bytes};
                    memcpy(&pn.value, ref, 8);
                } else {
                    pn.value = nd.threshold;
                }
                if (nd.left >= 0) flags |= PN_LEFT;
                if (nd.right >= 0) flags |= PN_RIGHT;
            } else {
                pn.value = m->is_classification ? (double)nd.class_value : nd.predict;
            }
            pn.flags = (uint16_t)flags;
            packed.push_back(pn);
            if (walk) {  // right is pushed first so that the left subtree is laid out right behind this node
                if (nd.right >= 0) stack.push_back({nd.right, me});
                if (nd.left >= 0) stack.push_back({nd.left, -1});
            }
        }
    }
    // one blob: [nodes][weights][bag_begin][tree_root][categorical_size][cat_pool], each 16-byte aligned
    auto up = [](size_t v) { return (v + 15) / 16 * 16; };
    const size_t o_nodes = 0;
    const size_t o_w = o_nodes + up(sizeof(PackedNode) * packed.size());
    const size_t o_bag = o_w + up(8 * (size_t)m->n_trees);
    const size_t o_root = o_bag + up(4 * (size_t)(m->n_bags + 1));
    const size_t o_cat = o_root + up(4 * (size_t)m->n_trees);
    const size_t o_pool = o_cat + up(4 * (size_t)m->n_columns);
    const size_t total = o_pool + up((size_t)std::max(m->cat_pool_bytes, 1));
    std::vector<uint8_t> host(total, 0);
    memcpy(host.data() + o_nodes, packed.data(), sizeof(PackedNode) * packed.size());
    memcpy(host.data() + o_w, m->weights, 8 * (size_t)m->n_trees);
    memcpy(host.data() + o_bag, m->bag_begin, 4 * (size_t)(m->n_bags + 1));
    memcpy(host.data() + o_root, roots.data(), 4 * (size_t)m->n_trees);
    memcpy(host.data() + o_cat, m->categorical_size, 4 * (size_t)m->n_columns);
    if (m->cat_pool_bytes > 0 && m->cat_pool) memcpy(host.data() + o_pool, m->cat_pool, (size_t)m->cat_pool_bytes);
    SB_CUDA(cudaMalloc(&s->blob, total));
    SB_CUDA(cudaMemcpy(s->blob, host.data(), total, cudaMemcpyHostToDevice));
    SB_CUDA(cudaMalloc(&s->d_error, 4));
    SB_CUDA(cudaEventCreate(&s->e0));
    SB_CUDA(cudaEventCreate(&s->e1));
    uint8_t* b = static_cast<uint8_t*>(s->blob);
    ScoreView& v = s->view;
    v.nodes = reinterpret_cast<const PackedNode*>(b + o_nodes);
    v.weights = reinterpret_cast<const double*>(b + o_w);
    v.bag_begin = reinterpret_cast<const int32_t*>(b + o_bag);
    v.tree_root = reinterpret_cast<const int32_t*>(b + o_root);
    v.categorical_size = reinterpret_cast<const int32_t*>(b + o_cat);
    v.cat_pool = b + o_pool;
    v.n_bags = m->n_bags; v.n_trees = m->n_trees; v.n_nodes = (int)packed.size(); v.cat_pool_bytes = m->cat_pool_bytes;
    v.is_gbdt = m->is_gbdt; v.is_classification = m->is_classification; v.convert = m->convert; v.n_columns = m->n_columns;
    *out = s.release();
    return SB_OK;
}

SB_API void sb_scorer_destroy(sb_scorer* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    if (s->blob) cudaFree(s->blob);
    if (s->d_error) cudaFree(s->d_error);
    if (s->d_data) cudaFree(s->d_data);
    if (s->d_scores) cudaFree(s->d_scores);
    if (s->e0) cudaEventDestroy(s->e0);
    if (s->e1) cudaEventDestroy(s->e1);
    delete s;
}

SB_API int sb_scorer_score(sb_scorer* s, int64_t n_rows, const double* data, int32_t data_on_device, double* scores) {
    if (!s || !scores || (!data && n_rows > 0) || n_rows < 0) return fail(SB_ERR_INVALID, "null argument");
    if (n_rows == 0) return SB_OK;
    SB_CUDA(cudaSetDevice(s->device));
    const ScoreView& v = s->view;
    const size_t n_in = (size_t)n_rows * v.n_columns;
    const size_t n_out = (size_t)n_rows * (v.is_classification ? v.n_trees : 1);
    const double* d_in = data;
    if (!data_on_device) {
        if (n_in > s->data_cap) {
            if (s->d_data) cudaFree(s->d_data);
            s->d_data = nullptr; s->data_cap = 0;
            SB_CUDA(cudaMalloc(&s->d_data, n_in * 8));
            s->data_cap = n_in;
        }
        SB_CUDA(cudaMemcpyAsync(s->d_data, data, n_in * 8, cudaMemcpyHostToDevice, 0));
        d_in = s->d_data;
    }
    if (n_out > s->scores_cap) {
        if (s->d_scores) cudaFree(s->d_scores);
        s->d_scores = nullptr; s->scores_cap = 0;
        SB_CUDA(cudaMalloc(&s->d_scores, n_out * 8));
        s->scores_cap = n_out;
    }
    SB_CUDA(cudaMemsetAsync(s->d_error, 0, 4, 0));
    // shared memory per block: rows x (n_columns | 1) doubles, plus the packed nodes when both fit
    const size_t budget = 220 * 1024;
    const size_t row_bytes = (size_t)(v.n_columns | 1) * 8;
    const size_t node_bytes = (size_t)v.n_nodes * sizeof(PackedNode);
    static bool attr_set = false;
    if (!attr_set) {
        SB_CUDA(cudaFuncSetAttribute(score_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        SB_CUDA(cudaFuncSetAttribute(score_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        // nodes read through L1: leave half of the L1/shared array to L1
        SB_CUDA(cudaFuncSetAttribute(score_kernel<true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, 50));
        attr_set = true;
    }
    SB_CUDA(cudaEventRecord(s->e0, 0));
    int rows_per_block = 0;
    for (int t : {512, 256, 128})
        if (node_bytes + (size_t)t * row_bytes <= budget) { rows_per_block = t; break; }
    if (rows_per_block > 0 && node_bytes <= 110 * 1024) {
        const unsigned grid = (unsigned)((n_rows + rows_per_block - 1) / rows_per_block);
        score_kernel<true, true><<<grid, rows_per_block, node_bytes + (size_t)rows_per_block * row_bytes, 0>>>(v, d_in, n_rows, s->d_scores, s->d_error);
    } else if (v.n_columns <= SC_MAX_SMEM_COLS) {
        const unsigned grid = (unsigned)((n_rows + SC_ROWS - 1) / SC_ROWS);
        score_kernel<true, false><<<grid, SC_ROWS, (size_t)SC_ROWS * row_bytes, 0>>>(v, d_in, n_rows, s->d_scores, s->d_error);
    } else {
        const unsigned grid = (unsigned)((n_rows + SC_ROWS - 1) / SC_ROWS);
        score_kernel<false, false><<<grid, SC_ROWS, 0, 0>>>(v, d_in, n_rows, s->d_scores, s->d_error);
    }
    SB_CUDA(cudaGetLastError());
    SB_CUDA(cudaEventRecord(s->e1, 0));
    uint32_t err = 0;
    SB_CUDA(cudaMemcpyAsync(scores, s->d_scores, n_out * 8, cudaMemcpyDeviceToHost, 0));
    SB_CUDA(cudaMemcpyAsync(&err, s->d_error, 4, cudaMemcpyDeviceToHost, 0));
    SB_CUDA(cudaStreamSynchronize(0));
    SB_CUDA(cudaEventElapsedTime(&s->last_ms, s->e0, s->e1));
    if (err & 1u) return fail(SB_ERR_STATE, "a split node is missing the child a row was sent to (NullPointerException in predictNode)");
    return SB_OK;
}

SB_API int sb_scorer_last_kernel_ms(sb_scorer* s, float* ms) {
    if (!s || !ms) return fail(SB_ERR_INVALID, "null argument");
    *ms = s->last_ms;
    return SB_OK;
}

}  // extern "C"
