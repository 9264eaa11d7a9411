// rowlist.cu -- per-pass row lists for levels whose todo nodes do not fit one shared-memory slot group.
//
// When a pass has more todo nodes than one block can hold histograms for (GS slots), the column-major tile kernel
// would have to stream every row tile once per slot group.  Instead the ACTIVE rows of the pass are grouped by slot
// group once per pass, their fixed-point operands written as a compact RowOps array in that order, and
// hist_rows_kernel (hist.cu) walks each group's list, reading the bins of a row and feature tile from a ROW-major copy
// of the bin matrix with one coalesced load.  Every active row is then touched once per feature tile, whatever the
// number of groups.
//
// The grouping is a stable counting sort with very few keys (the slot groups), done in two streaming passes over the
// rows and one prefix sum: (1) every warp counts, per group, the active rows of its own range of 1024 consecutive rows;
// (2) an exclusive scan over the (group, range) counts, group-major, gives every warp the position of its first row
// of every group; (3) every warp walks its range again and writes the RowOps of each active row to its position.
// Rows stay in ascending order inside a group, inactive rows (no todo node, zero weight) are neither sorted nor
// written, and no row index array or key array exists.
#include <cub/device/device_scan.cuh>

#include "common.cuh"
#include "kernels.cuh"

namespace sb {

namespace {

constexpr int TPB = 256;

// column-major [F][pitch] -> row-major [n][fpad]; block = 32 features x 128 rows through shared memory
__global__ void transpose_bins_kernel(const uint8_t* __restrict__ cm, int64_t pitch, int n_features,
                                      const int32_t* __restrict__ feat_of_pos, int64_t n, uint8_t* __restrict__ rm, int fpad) {
    __shared__ uint8_t tile[32][128 + 4];
    const int64_t r0 = (int64_t)blockIdx.x * 128;
    const int f0 = blockIdx.y * 32;
    // load: each warp takes 4 feature rows, a lane reads 4 consecutive bytes
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;  // 8 warps
    for (int k = warp; k < 32; k += 8) {
        const int f = feat_of_pos[f0 + k];  // byte f0 + k of a twin row holds this feature (-1: padding)
        uint32_t v = 0;
        if (f >= 0 && f < n_features) {
            const int64_t rb = r0 + lane * 4;
            const uint8_t* src = cm + (int64_t)f * pitch + rb;
            if (rb + 3 < pitch) v = *reinterpret_cast<const uint32_t*>(src);
            else
                for (int b = 0; b < 4; ++b)
                    if (rb + b < n) v |= (uint32_t)src[b] << (8 * b);
        }
        *reinterpret_cast<uint32_t*>(&tile[k][lane * 4]) = v;
    }
    __syncthreads();
    // store: a warp writes the 32 feature bytes of one row (one 32-byte segment)
    for (int rr = warp; rr < 128; rr += 8) {
        const int64_t r = r0 + rr;
        if (r < n) rm[r * fpad + f0 + lane] = tile[lane][rr];
    }
}

constexpr int RL_RANGE = 1024;  // consecutive rows counted / scattered by one warp
constexpr int RL_WARPS = 8;

// group id of a row: its slot group, or 255 when the row takes no part in this pass; *slot_out = the slot
__device__ __forceinline__ int row_group(int64_t r, int64_t n, const int32_t* __restrict__ node_id,
                                         const float* __restrict__ weights, const int16_t* __restrict__ slot_of_node,
                                         int slot_table_size, int gs, int* slot_out, float* w_out) {
    *slot_out = -1;
    *w_out = 0.0f;
    if (r >= n) return 255;
    const int nid = node_id[r];
    const int s = (nid >= 0 && nid < slot_table_size) ? (int)slot_of_node[nid] : -1;
    const float w = weights ? weights[r] : 1.0f;
    *slot_out = s;
    *w_out = w;
    return (s >= 0 && w != 0.0f) ? s / gs : 255;
}

// predictNodeIndex continued from node `id` (same walk as route_kernel in rows.cu)
__device__ __forceinline__ int walk_down(int id, int64_t r, const void* const* __restrict__ cols,
                                         const int32_t* __restrict__ feat_bins, const TreeView& t) {
    while (id > 0 && id < t.capacity) {
        const int f = t.feat[id];
        if (f < 0) break;
        const int bin = feat_bins[f] > 256 ? (int)static_cast<const uint16_t*>(cols[f])[r] : (int)static_cast<const uint8_t*>(cols[f])[r];
        const uint8_t m = t.masks[t.mask_off[id] + (bin >> 3)];
        id = 2 * id + (((m >> (bin & 7)) & 1) ? 0 : 1);
    }
    return id;
}

// counts[g * n_ranges + range] = active rows of group g in the range; one trailing zero so that the exclusive scan
// also yields the total.  ROUTE: rows are first moved from their node to the child the current tree sends them to.
template <bool ROUTE>
__global__ void __launch_bounds__(RL_WARPS * 32)
rowlist_count_kernel(int64_t n, int32_t* __restrict__ node_id, const float* __restrict__ weights,
                     const int16_t* __restrict__ slot_of_node, int slot_table_size, int gs, int n_groups, int64_t n_ranges,
                     uint32_t* __restrict__ counts, TreeView t, const void* const* __restrict__ cols,
                     const int32_t* __restrict__ feat_bins, uint32_t* error_flag) {
    __shared__ uint32_t cnt[RL_WARPS][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t range = (int64_t)blockIdx.x * RL_WARPS + warp;
    if (blockIdx.x == 0 && threadIdx.x == 0) counts[(int64_t)n_groups * n_ranges] = 0u;
    if (range >= n_ranges) return;
    for (int i = lane; i < 256; i += 32) cnt[warp][i] = 0u;
    __syncwarp();
    const int64_t r0 = range * RL_RANGE;
    constexpr int U = 8;
    for (int step = 0; step < RL_RANGE / 32; step += U) {
        int key[U];
        if (ROUTE) {
            // first step of the walk for U rows at once (loads of different rows in flight together); a row that has
            // further to go (a tree seen for the first time) finishes in walk_down
            int id[U], f[U], bin[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t r = r0 + (step + u) * 32 + lane;
                id[u] = r < n ? node_id[r] : 0;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) f[u] = (id[u] > 0 && id[u] < t.capacity) ? t.feat[id[u]] : -1;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t r = r0 + (step + u) * 32 + lane;
                bin[u] = 0;
                if (f[u] >= 0)
                    bin[u] = feat_bins[f[u]] > 256 ? (int)static_cast<const uint16_t*>(cols[f[u]])[r]
                                                   : (int)static_cast<const uint8_t*>(cols[f[u]])[r];
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t r = r0 + (step + u) * 32 + lane;
                int nid = id[u];
                if (f[u] >= 0) {
                    const uint8_t m = t.masks[t.mask_off[nid] + (bin[u] >> 3)];
                    nid = walk_down(2 * nid + (((m >> (bin[u] & 7)) & 1) ? 0 : 1), r, cols, feat_bins, t);
                    if (nid >= t.capacity) atomicOr(error_flag, 4u);  // child missing (dt/DTWorker.java:1089-1094)
                    node_id[r] = nid;
                }
                key[u] = 255;
                if (r < n) {
                    const int s = (nid >= 0 && nid < slot_table_size) ? (int)slot_of_node[nid] : -1;
                    const float w = weights ? weights[r] : 1.0f;
                    if (s >= 0 && w != 0.0f) key[u] = s / gs;
                }
            }
        } else {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                int sl;
                float w;
                key[u] = row_group(r0 + (step + u) * 32 + lane, n, node_id, weights, slot_of_node, slot_table_size, gs, &sl, &w);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t mask = __match_any_sync(0xffffffffu, key[u]);
            if (key[u] != 255 && lane == __ffs(mask) - 1) cnt[warp][key[u]] += __popc(mask);
            __syncwarp();
        }
    }
    for (int g = lane; g < n_groups; g += 32) counts[(int64_t)g * n_ranges + range] = cnt[warp][g];
}

// offsets[g] = first position of group g (offsets[n_groups] = number of active rows), from the scanned counts or, for
// the identity list (scanned == nullptr), a single group holding every row; then the chunk table: chunk c covers ops
// [begin, end) of one group.
__global__ void rowlist_chunks_kernel(const uint32_t* __restrict__ scanned, int64_t n_ranges, int64_t n, int n_groups,
                                      int64_t max_chunk_rows, int target_chunks, int n_tiles, int64_t* __restrict__ offsets,
                                      int32_t* __restrict__ chunk_group, int64_t* __restrict__ chunk_begin,
                                      int64_t* __restrict__ chunk_end, int32_t* __restrict__ n_chunks, int max_chunks,
                                      uint32_t* error_flag, unsigned long long* rows_streamed) {
    __shared__ int64_t off[257];
    for (int g = threadIdx.x; g <= n_groups; g += blockDim.x) {
        int64_t v;
        if (!scanned) v = (g == 0 ? 0 : n);  // identity list
        else v = (int64_t)scanned[(int64_t)g * n_ranges];
        off[g] = v;
        offsets[g] = v;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        // chunk size from the number of ACTIVE rows of this pass (with subtraction most passes touch a fraction of the
        // rows): about target_chunks chunks, so the grid fills the GPU several times over whatever the list length
        const int64_t total = off[n_groups];
        if (rows_streamed) atomicAdd(rows_streamed, (unsigned long long)total);
        // number of chunks: at least what the accumulator headroom asks for (max_chunk_rows), about target_chunks, and
        // such that chunks x tiles fills whole waves of 148 blocks (one block per SM) with room for the partial chunk
        // that ends every group -- a launch of 2.6 waves costs 3
        const int64_t k_min = (total + max_chunk_rows - 1) / max_chunk_rows;
        // target_chunks is about four waves; a short list takes two, so that zeroing and flushing a block's histogram
        // stays small next to its rows
        int64_t want = target_chunks;
        if (total / want < 16384 && want >= 2) want /= 2;
        int64_t k = k_min > want ? k_min : want;
        if (k_min > want) {
            const int64_t spare = n_groups < 74 ? n_groups : 74;
            int64_t waves = (k * n_tiles + 147) / 148;
            int64_t kw = (waves * 148) / n_tiles - spare;
            if (kw < k_min) kw = ((waves + 1) * 148) / n_tiles - spare;
            if (kw >= k_min) k = kw;
        }
        int64_t chunk_rows = ((total + k - 1) / k + 127) / 128 * 128;
        if (chunk_rows < 1024) chunk_rows = 1024;
        if (chunk_rows > max_chunk_rows) chunk_rows = max_chunk_rows;
        int c = 0;
        for (int g = 0; g < n_groups; ++g)
            for (int64_t b = off[g]; b < off[g + 1]; b += chunk_rows) {
                if (c >= max_chunks) {  // cannot happen with the host's bound; never drop rows silently
                    atomicOr(error_flag, 8u);
                    break;
                }
                chunk_group[c] = g;
                chunk_begin[c] = b;
                chunk_end[c] = min(off[g + 1], b + chunk_rows);
                ++c;
            }
        *n_chunks = c;
    }
}

// RowOps of one active row (see kernels.cuh): slot inside its group and the fixed-point operands
template <int KIND, int CNT_INT>
__device__ __forceinline__ RowOps make_row_ops(uint32_t r, int s, int gs, float w, float sig, float y, float sc0, float sc1,
                                               float sc2, int stats_size, uint32_t* error_flag) {
    RowOps o;
    o.row = r;
    o.slot = (uint32_t)(s % gs);
    const float cw = __fmul_rn(sig, w);
    long long qc;
    if (CNT_INT) qc = (long long)__float2int_rn(cw);
    else qc = __float2ll_rn(__fmul_rn(cw, sc0));
    o.cnt_lo = (uint32_t)qc; o.cnt_hi = (uint32_t)(qc >> 32);
    if (KIND == STAT_VARIANCE) {
        const float sw = __fmul_rn(__fmul_rn(y, sig), w);
        const float qw = __fmul_rn(__fmul_rn(__fmul_rn(y, y), sig), w);
        const long long qs = __float2ll_rn(__fmul_rn(sw, sc1));
        const long long qq = __float2ll_rn(__fmul_rn(qw, sc2));
        o.q[0] = (uint32_t)qs; o.q[1] = (uint32_t)(qs >> 32);
        o.q[2] = (uint32_t)qq; o.q[3] = (uint32_t)(qq >> 32);
        if (CNT_INT) {  // packed high words, see rowops_pack in kernels.cuh
            o.q[1] = rowops_pack(o.q[1], (uint32_t)qc);
            o.q[3] = rowops_pack(o.q[3], o.slot);
        }
    } else {
        const int cls = (int)(__fadd_rn(y, 0.000001f));
        if (cls < 0 || cls >= stats_size) { atomicOr(error_flag, 1u); o.slot = 0xffffffffu; o.q[0] = 0; }
        else o.q[0] = (uint32_t)(cls * (CNT_INT ? 128 : 256));  // byte offset of the class words
        o.q[1] = o.q[2] = o.q[3] = 0;
    }
    return o;
}

// pass (3): every warp writes the RowOps of the active rows of its range at the scanned positions
template <int KIND, int CNT_INT>
__global__ void __launch_bounds__(RL_WARPS * 32)
rowlist_scatter_kernel(int64_t n, const int32_t* __restrict__ node_id, const float* __restrict__ output,
                       const float* __restrict__ significance, const float* __restrict__ weights,
                       const int16_t* __restrict__ slot_of_node, int slot_table_size, int gs, int n_groups, int64_t n_ranges,
                       const uint32_t* __restrict__ scanned, int stats_size, const float* __restrict__ scales,
                       RowOps* __restrict__ ops, uint32_t* error_flag) {
    __shared__ uint32_t off[RL_WARPS][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t range = (int64_t)blockIdx.x * RL_WARPS + warp;
    if (range >= n_ranges) return;
    for (int g = lane; g < n_groups; g += 32) off[warp][g] = scanned[(int64_t)g * n_ranges + range];
    __syncwarp();
    const float sc0 = scales[0], sc1 = scales[1], sc2 = scales[2];
    const int64_t r0 = range * RL_RANGE;
    const uint32_t lt = (1u << lane) - 1u;
    constexpr int U = 4;
    for (int step = 0; step < RL_RANGE / 32; step += U) {
        int key[U], sl[U];
        float w[U], sig[U], y[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t r = r0 + (step + u) * 32 + lane;
            key[u] = row_group(r, n, node_id, weights, slot_of_node, slot_table_size, gs, &sl[u], &w[u]);
            sig[u] = y[u] = 0.0f;
            if (key[u] != 255) { sig[u] = significance[r]; y[u] = output[r]; }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t mask = __match_any_sync(0xffffffffu, key[u]);
            uint32_t pos = 0;
            if (key[u] != 255) pos = off[warp][key[u]] + __popc(mask & lt);
            __syncwarp();
            if (key[u] != 255 && lane == __ffs(mask) - 1) off[warp][key[u]] += __popc(mask);
            __syncwarp();
            if (key[u] != 255)
                ops[pos] = make_row_ops<KIND, CNT_INT>((uint32_t)(r0 + (step + u) * 32 + lane), sl[u], gs, w[u], sig[u], y[u], sc0,
                                                       sc1, sc2, stats_size, error_flag);
        }
    }
}

// identity list (a pass whose single todo node is a tree root): ops[i] for row i in storage order; rows that do not
// belong to the node are marked instead of being left out
template <int KIND, int CNT_INT>
__global__ void rowlist_ops_kernel(int64_t total, const int32_t* __restrict__ node_id, const float* __restrict__ output,
                                   const float* __restrict__ significance, const float* __restrict__ weights,
                                   const int16_t* __restrict__ slot_of_node, int slot_table_size, int gs, int stats_size,
                                   const float* __restrict__ scales, RowOps* __restrict__ ops, uint32_t* error_flag) {
    const float sc0 = scales[0], sc1 = scales[1], sc2 = scales[2];
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int nid = node_id[i];
        const int s = (nid >= 0 && nid < slot_table_size) ? (int)slot_of_node[nid] : -1;
        const float w = weights ? weights[i] : 1.0f;
        if (s < 0 || w == 0.0f) {
            RowOps skip;
            skip.row = (uint32_t)i;
            skip.slot = 0xffffffffu;
            skip.cnt_lo = skip.cnt_hi = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) skip.q[k] = 0;
            if (KIND == STAT_VARIANCE && CNT_INT) skip.q[3] = rowops_pack(0u, 0xffffu);
            ops[i] = skip;
            continue;
        }
        ops[i] = make_row_ops<KIND, CNT_INT>((uint32_t)i, s, gs, w, significance[i], output[i], sc0, sc1, sc2, stats_size, error_flag);
    }
}


// ---- forest passes: one list for the todo nodes of all trees of a pass, keyed by GLOBAL slot group ----------------------
// Same three steps as above -- count per (key, range), exclusive scan, scatter -- with two differences: a block (not a
// warp) owns a range of FOREST_RANGE rows and walks it once per tree, and the counters of all keys of the batch sit in
// shared memory (a pass of 50 trees at depth 10 has thousands of slot groups).  Positions inside a (key, range) cell are
// handed out by shared-memory atomics, so the order of the pairs inside a cell is not fixed; the integer sums
// the statistics kernels form do not depend on it.

// key of (row, tree): slot group inside the batch, or -1.  *slot_out = slot inside the batch.
__device__ __forceinline__ int forest_key(const ListTree& lt, int64_t r, int slot_base, int gs, int n_keys, int* slot_out,
                                          float* w_out) {
    const int nid = lt.node_id[r];
    const int s = (nid >= 0 && nid < lt.slot_table_size) ? lt.slot_of_node[nid] - slot_base : -1;
    const float w = lt.weights ? lt.weights[r] : 1.0f;
    *slot_out = s;
    *w_out = w;
    if (s < 0 || w == 0.0f) return -1;
    const int key = s / gs;
    return key < n_keys ? key : -1;  // a slot of a later batch
}

__global__ void __launch_bounds__(256) forest_count_kernel(const ForestListArgs a) {
    extern __shared__ uint32_t fcnt[];
    const int64_t range = blockIdx.x;
    if (blockIdx.x == 0 && threadIdx.x == 0) a.counts[(int64_t)a.n_keys * a.n_ranges] = 0u;
    for (int k = threadIdx.x; k < a.n_keys; k += 256) fcnt[k] = 0u;
    __syncthreads();
    const int64_t r0 = range * FOREST_RANGE;
    const int lane = threadIdx.x & 31;
    for (int t = 0; t < a.n_trees; ++t) {
        const ListTree lt = a.trees[t];
        constexpr int U = 4;
        for (int step = 0; step < FOREST_RANGE / 256; step += U) {
            int key[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t r = r0 + (int64_t)(step + u) * 256 + threadIdx.x;
                int sl;
                float w;
                key[u] = r < a.n ? forest_key(lt, r, a.slot_base, a.gs, a.n_keys, &sl, &w) : -1;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t mask = __match_any_sync(0xffffffffu, key[u]);
                if (key[u] >= 0 && lane == __ffs(mask) - 1) atomicAdd(&fcnt[key[u]], (uint32_t)__popc(mask));
            }
        }
    }
    __syncthreads();
    for (int k = threadIdx.x; k < a.n_keys; k += 256) a.counts[(int64_t)k * a.n_ranges + range] = fcnt[k];
}

// Chunk table of a forest list, slab-major.  The rows are cut into J slabs of whole ranges, the same for every key; cell
// (j, k) = the pairs of key k whose rows lie in slab j -- a contiguous piece of the key's list, read off the scanned
// counts -- and a cell becomes one chunk (more when it exceeds the accumulator headroom, chunk_rows).  Chunks are numbered
// slab by slab: blocks that run at the same time work on the same slab of rows for different keys (different trees), so
// a row's line of the twin is fetched from DRAM once and comes out of L2 for the other trees that hold the row (the
// twin of 10M rows is ten times the L2; a slab is a fraction of it).  Every chunk merges its block's histogram into the
// global accumulators, so J is bounded by what that costs: J sweeps over the statistics buffer must stay well below the
// bytes of bins the pass reads.  By default the slab size is unbounded and J only provides enough chunks to fill the GPU:
// slabs sized for L2 (SB_FOREST_SLAB_MB) were measured slower on C1 at every size, because the class kernel is not bound by
// HBM.  One block; the words behind n_chunks are the work counters of the persistent statistics kernels.
__global__ void __launch_bounds__(1024) forest_chunks_kernel(const ForestListArgs a) {
    __shared__ int s_J, s_R, s_base;
    __shared__ int s_warp[32];
    const int64_t nr = a.n_ranges;
    if (threadIdx.x == 0) {
        const int64_t total = (int64_t)a.scanned[(int64_t)a.n_keys * nr];
        if (a.rows_streamed) atomicAdd(a.rows_streamed, (unsigned long long)total);
        // the statistics kernel's bulk copies are 16 bytes wide and may take one 8-byte record past the last pair, and
        // it loads bins for whatever row that record names: make it row 0, count 0 (the scatter writes below `total` only)
        if (a.slim) {
            static_cast<ClassOps*>(a.ops)[total] = ClassOps{0u, 0u};
            static_cast<ClassOps*>(a.ops)[total + 1] = ClassOps{0u, 0u};
        }
        int64_t J = (a.n * a.row_bytes + a.slab_bytes - 1) / a.slab_bytes;            // a slab of the twin fits L2
        const int64_t j_merge = (total * a.row_bytes) / (4 * (a.stats_bytes > 0 ? a.stats_bytes : 1));  // merges stay a quarter of the bins
        if (J > j_merge) J = j_merge;
        const int64_t j_par = (a.target_chunks + a.n_keys - 1) / a.n_keys;            // enough chunks to fill the GPU
        if (J < j_par) J = j_par;
        if (J < 1) J = 1;
        if (J > nr) J = nr;
        const int64_t R = (nr + J - 1) / J;   // ranges per slab
        J = (nr + R - 1) / R;
        s_J = (int)J;
        s_R = (int)R;
        s_base = 0;
    }
    __syncthreads();
    const int J = s_J, R = s_R;
    const int64_t cells = (int64_t)J * a.n_keys;
    const int64_t cr = a.chunk_rows;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int64_t c0 = 0; c0 < cells; c0 += 1024) {
        const int64_t c = c0 + threadIdx.x;
        int64_t b = 0, e = 0;
        int k = 0;
        if (c < cells) {
            const int j = (int)(c / a.n_keys);
            k = (int)(c - (int64_t)j * a.n_keys);
            const int64_t r_lo = (int64_t)j * R, r_hi = min(nr, r_lo + R);
            b = (int64_t)a.scanned[(int64_t)k * nr + r_lo];
            e = (int64_t)a.scanned[(int64_t)k * nr + r_hi];  // r_hi == nr: the first position of key k + 1
        }
        const int nc = (int)((e - b + cr - 1) / cr);
        int incl = nc;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            int v = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, v, o);
                if (lane >= o) v += u;
            }
            s_warp[lane] = v;  // inclusive over warps
        }
        __syncthreads();
        const int base = s_base + (warp > 0 ? s_warp[warp - 1] : 0) + incl - nc;
        for (int q = 0; q < nc; ++q) {
            const int at = base + q;
            if (at >= a.max_chunks) {  // cannot happen with the host's bound; never drop pairs silently
                atomicOr(a.error_flag, 8u);
                break;
            }
            a.chunk_group[at] = k;
            a.chunk_begin[at] = b + (int64_t)q * cr;
            a.chunk_end[at] = min(e, b + (int64_t)(q + 1) * cr);
        }
        __syncthreads();
        if (threadIdx.x == 0) s_base += s_warp[31];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        a.n_chunks[0] = min(s_base, a.max_chunks);
        a.n_chunks[1] = 0;  // work counters
        a.n_chunks[2] = 0;
    }
}

template <int KIND, int CNT_INT, int SLIM>
__global__ void __launch_bounds__(256) forest_scatter_kernel(const ForestListArgs a) {
    extern __shared__ uint32_t foff[];
    const int64_t range = blockIdx.x;
    for (int k = threadIdx.x; k < a.n_keys; k += 256) foff[k] = a.scanned[(int64_t)k * a.n_ranges + range];
    __syncthreads();
    const float sc0 = a.scales[0], sc1 = a.scales[1], sc2 = a.scales[2];
    const int64_t r0 = range * FOREST_RANGE;
    const int lane = threadIdx.x & 31;
    const uint32_t lt_mask = (1u << lane) - 1u;
    for (int t = 0; t < a.n_trees; ++t) {
        const ListTree lt = a.trees[t];
        constexpr int U = 4;
        for (int step = 0; step < FOREST_RANGE / 256; step += U) {
            int key[U], sl[U];
            float w[U], sig[U], y[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t r = r0 + (int64_t)(step + u) * 256 + threadIdx.x;
                key[u] = r < a.n ? forest_key(lt, r, a.slot_base, a.gs, a.n_keys, &sl[u], &w[u]) : -1;
                sig[u] = y[u] = 0.0f;
                if (key[u] >= 0) { sig[u] = a.significance[r]; y[u] = a.output[r]; }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t mask = __match_any_sync(0xffffffffu, key[u]);
                const int leader = __ffs(mask) - 1;
                uint32_t first = 0;
                if (key[u] >= 0 && lane == leader) first = atomicAdd(&foff[key[u]], (uint32_t)__popc(mask));
                first = __shfl_sync(0xffffffffu, first, leader);
                if (key[u] < 0) continue;
                const uint32_t pos = first + (uint32_t)__popc(mask & lt_mask);
                const uint32_t row = (uint32_t)(r0 + (int64_t)(step + u) * 256 + threadIdx.x);
                if (SLIM) {
                    ClassOps o;
                    o.row = row;
                    const int cls = (int)(__fadd_rn(y[u], 0.000001f));  // dt/Impurity.java: (int) (label + 0.000001f)
                    const uint32_t cnt = (uint32_t)__float2int_rn(__fmul_rn(sig[u], w[u]));
                    if (cls < 0 || cls >= a.stats_size) { atomicOr(a.error_flag, 1u); o.where = 0u; }
                    else o.where = ((uint32_t)(sl[u] % a.gs) * a.slot_stride + (uint32_t)cls * a.class_stride) | (cnt << CLASSOPS_OFF_BITS);
                    static_cast<ClassOps*>(a.ops)[pos] = o;
                } else {
                    static_cast<RowOps*>(a.ops)[pos] = make_row_ops<KIND, CNT_INT>(row, sl[u], a.gs, w[u], sig[u], y[u], sc0, sc1, sc2,
                                                                                  a.stats_size, a.error_flag);
                }
            }
        }
    }
}

// pair_off[s][f] = offset (in statistics) of feature f's block of the todo node that slot s stands for, -1 if the node
// does not take f.  One warp per slot; the offsets inside a node block are a running sum over its list.
__global__ void __launch_bounds__(256) build_pair_off_kernel(const int32_t* __restrict__ slot_node, int n_slots,
                                                             const int32_t* __restrict__ feats, const int32_t* __restrict__ begin,
                                                             const int64_t* __restrict__ node_off,
                                                             const int32_t* __restrict__ feat_bins, int stats_size, int n_features,
                                                             int64_t* __restrict__ pair_off) {
    const int s = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (s >= n_slots) return;
    int64_t* row = pair_off + (int64_t)s * n_features;
    for (int f = lane; f < n_features; f += 32) row[f] = -1;
    __syncwarp();
    const int n = slot_node[s];
    const int b = begin[n], e = begin[n + 1];
    int64_t run = node_off[n];
    for (int k0 = b; k0 < e; k0 += 32) {
        const int k = k0 + lane;
        const int f = k < e ? feats[k] : -1;
        const int64_t len = f >= 0 ? (int64_t)feat_bins[f] * stats_size : 0;
        int64_t incl = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int64_t v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        if (f >= 0) row[f] = run + incl - len;
        run += __shfl_sync(0xffffffffu, incl, 31);
    }
}

}  // namespace

__global__ void transpose_bins16_kernel(const uint16_t* __restrict__ cm, int64_t pitch16, int n_wide, int64_t n,
                                        uint16_t* __restrict__ rm, int wpad) {
    // one-time layout change of the (few) wide columns: coalesced reads along rows, strided writes
    const int j = blockIdx.y;
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x)
        rm[r * wpad + j] = j < n_wide ? cm[(int64_t)j * pitch16 + r] : (uint16_t)0;
}
cudaError_t launch_transpose_bins16(const uint16_t* cm, int64_t pitch16, int n_wide, int64_t n, uint16_t* rm, int wpad,
                                    cudaStream_t s) {
    if (n <= 0 || wpad <= 0) return cudaSuccess;
    int64_t gx = (n + 255) / 256;
    if (gx > 148 * 8) gx = 148 * 8;
    dim3 grid((unsigned)gx, (unsigned)wpad);
    transpose_bins16_kernel<<<grid, 256, 0, s>>>(cm, pitch16, n_wide, n, rm, wpad);
    return cudaGetLastError();
}

cudaError_t launch_transpose_bins(const uint8_t* cm, int64_t pitch, int n_features, const int32_t* feat_of_pos, int64_t n,
                                  uint8_t* rm, int fpad, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    dim3 grid((unsigned)((n + 127) / 128), (unsigned)(fpad / 32));
    transpose_bins_kernel<<<grid, 256, 0, s>>>(cm, pitch, n_features, feat_of_pos, n, rm, fpad);
    return cudaGetLastError();
}

int64_t rowlist_n_ranges(int64_t n) { return (n + RL_RANGE - 1) / RL_RANGE; }

size_t rowlist_scan_temp_bytes(int64_t items) {
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, items);
    return bytes;
}

cudaError_t launch_rowlist_build(const RowListArgs& a, cudaStream_t s) {
    const int64_t n = a.n;
    if (n <= 0) return cudaSuccess;
    cudaError_t e;
    if (a.identity) {
        // no grouping: one group = all rows in storage order
        rowlist_chunks_kernel<<<1, 256, 0, s>>>(nullptr, 0, n, 1, a.chunk_rows, a.target_chunks, a.n_tiles, a.offsets, a.chunk_group,
                                                a.chunk_begin, a.chunk_end, a.n_chunks, a.max_chunks, a.error_flag, a.rows_streamed);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        int64_t g = (n + TPB - 1) / TPB;
        if (g > 148 * 16) g = 148 * 16;
#define OPS(K, C)                                                                                                      \
    rowlist_ops_kernel<K, C><<<(int)g, TPB, 0, s>>>(n, a.node_id, a.output, a.significance, a.weights, a.slot_of_node, \
                                                    a.slot_table_size, a.gs, a.stats_size, a.scales, a.ops, a.error_flag)
        if (a.kind == STAT_VARIANCE) { if (a.cnt_int) OPS(STAT_VARIANCE, 1); else OPS(STAT_VARIANCE, 0); }
        else { if (a.cnt_int) OPS(STAT_CLASS, 1); else OPS(STAT_CLASS, 0); }
#undef OPS
        return cudaGetLastError();
    }
    const int64_t n_ranges = rowlist_n_ranges(n);
    const int64_t items = (int64_t)a.n_groups * n_ranges + 1;
    const unsigned grid = (unsigned)((n_ranges + RL_WARPS - 1) / RL_WARPS);
    if (a.route)
        rowlist_count_kernel<true><<<grid, RL_WARPS * 32, 0, s>>>(n, a.node_id_rw, a.weights, a.slot_of_node, a.slot_table_size,
                                                                  a.gs, a.n_groups, n_ranges, a.counts, a.tree, a.cols,
                                                                  a.feat_bins, a.error_flag);
    else
        rowlist_count_kernel<false><<<grid, RL_WARPS * 32, 0, s>>>(n, const_cast<int32_t*>(a.node_id), a.weights, a.slot_of_node,
                                                                   a.slot_table_size, a.gs, a.n_groups, n_ranges, a.counts,
                                                                   TreeView{}, nullptr, nullptr, a.error_flag);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    size_t temp = a.temp_bytes;
    e = cub::DeviceScan::ExclusiveSum(a.temp, temp, (const uint32_t*)a.counts, a.scanned, items, s);
    if (e != cudaSuccess) return e;
    rowlist_chunks_kernel<<<1, 256, 0, s>>>(a.scanned, n_ranges, n, a.n_groups, a.chunk_rows, a.target_chunks, a.n_tiles, a.offsets,
                                            a.chunk_group, a.chunk_begin, a.chunk_end, a.n_chunks, a.max_chunks, a.error_flag, a.rows_streamed);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
#define SCATTER(K, C)                                                                                                     \
    rowlist_scatter_kernel<K, C><<<grid, RL_WARPS * 32, 0, s>>>(n, a.node_id, a.output, a.significance, a.weights,        \
                                                                a.slot_of_node, a.slot_table_size, a.gs, a.n_groups,      \
                                                                n_ranges, a.scanned, a.stats_size, a.scales, a.ops,       \
                                                                a.error_flag)
    if (a.kind == STAT_VARIANCE) { if (a.cnt_int) SCATTER(STAT_VARIANCE, 1); else SCATTER(STAT_VARIANCE, 0); }
    else { if (a.cnt_int) SCATTER(STAT_CLASS, 1); else SCATTER(STAT_CLASS, 0); }
#undef SCATTER
    return cudaGetLastError();
}

int64_t forest_n_ranges(int64_t n) { return (n + FOREST_RANGE - 1) / FOREST_RANGE; }

cudaError_t launch_forest_lists(const ForestListArgs& a, cudaStream_t s) {
    if (a.n <= 0 || a.n_keys <= 0) return cudaSuccess;
    static bool attr_set = false;
    const int max_smem = FOREST_MAX_KEYS * 4;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(forest_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(forest_scatter_kernel<STAT_VARIANCE, 0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(forest_scatter_kernel<STAT_VARIANCE, 1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(forest_scatter_kernel<STAT_CLASS, 0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(forest_scatter_kernel<STAT_CLASS, 1, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(forest_scatter_kernel<STAT_CLASS, 1, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    if (a.n_keys > FOREST_MAX_KEYS) return cudaErrorInvalidValue;
    const size_t smem = (size_t)a.n_keys * 4;
    const unsigned grid = (unsigned)a.n_ranges;
    forest_count_kernel<<<grid, 256, smem, s>>>(a);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    size_t temp = a.temp_bytes;
    e = cub::DeviceScan::ExclusiveSum(a.temp, temp, (const uint32_t*)a.counts, a.scanned, (int64_t)a.n_keys * a.n_ranges + 1, s);
    if (e != cudaSuccess) return e;
    forest_chunks_kernel<<<1, 1024, 0, s>>>(a);
    e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    if (a.slim) forest_scatter_kernel<STAT_CLASS, 1, 1><<<grid, 256, smem, s>>>(a);
    else if (a.kind == STAT_VARIANCE) {
        if (a.cnt_int) forest_scatter_kernel<STAT_VARIANCE, 1, 0><<<grid, 256, smem, s>>>(a);
        else forest_scatter_kernel<STAT_VARIANCE, 0, 0><<<grid, 256, smem, s>>>(a);
    } else {
        if (a.cnt_int) forest_scatter_kernel<STAT_CLASS, 1, 0><<<grid, 256, smem, s>>>(a);
        else forest_scatter_kernel<STAT_CLASS, 0, 0><<<grid, 256, smem, s>>>(a);
    }
    return cudaGetLastError();
}

__global__ void fill_pair_off_full_kernel(const int32_t* __restrict__ slot_node, int n_slots, int64_t node_stride,
                                          const int64_t* __restrict__ feat_off, int n_features, int64_t* __restrict__ pair_off) {
    const int64_t total = (int64_t)n_slots * n_features;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int s = (int)(i / n_features), f = (int)(i - (int64_t)s * n_features);
        const int64_t o = feat_off[f];
        pair_off[i] = o < 0 ? -1 : (int64_t)slot_node[s] * node_stride + o;
    }
}
cudaError_t launch_fill_pair_off_full(const int32_t* slot_node, int n_slots, int64_t node_stride, const int64_t* feat_off, int n_features,
                                      int64_t* pair_off, cudaStream_t s) {
    if (n_slots <= 0 || n_features <= 0) return cudaSuccess;
    const int64_t total = (int64_t)n_slots * n_features;
    fill_pair_off_full_kernel<<<(unsigned)std::min<int64_t>((total + 255) / 256, 148 * 8), 256, 0, s>>>(slot_node, n_slots, node_stride, feat_off,
                                                                                                    n_features, pair_off);
    return cudaGetLastError();
}

cudaError_t launch_build_pair_off(const int32_t* slot_node, int n_slots, const int32_t* feats, const int32_t* begin,
                                  const int64_t* node_off, const int32_t* feat_bins, int stats_size, int n_features,
                                  int64_t* pair_off, cudaStream_t s) {
    if (n_slots <= 0) return cudaSuccess;
    build_pair_off_kernel<<<(n_slots + 7) / 8, 256, 0, s>>>(slot_node, n_slots, feats, begin, node_off, feat_bins, stats_size,
                                                            n_features, pair_off);
    return cudaGetLastError();
}

}  // namespace sb
