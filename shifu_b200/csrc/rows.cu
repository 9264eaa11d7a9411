// rows.cu -- per-row kernels around the statistics build: node routing, GBT residual update, error sums,
// fixed-point scale selection, synthetic data.  All are one pass over per-row arrays (HBM-bound, coalesced).
// Built with -fmad=false: Java never contracts a*b+c, and these kernels restate Java float arithmetic.
#include "common.cuh"
#include "kernels.cuh"

namespace sb {

namespace {

constexpr int TPB = 256;

inline int grid_for(int64_t n, int tpb, int max_blocks) {
    int64_t g = (n + tpb - 1) / tpb;
    if (g < 1) g = 1;
    if (g > max_blocks) g = max_blocks;
    return (int)g;
}

__global__ void fill_i32_kernel(int32_t* p, int64_t n, int32_t v) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}
__global__ void fill_f32_kernel(float* p, int64_t n, float v) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}
__global__ void copy_f32_kernel(float* d, const float* s, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) d[i] = s[i];
}

// dt/DTWorker.java:1036-1096.  A node routes further iff it has a split and children (feat[id] >= 0);
// the left/right decision is the host-built left_mask bit of the row's bin (see sb_node in shifu_b200.h).
// A column holds uint8 bin indices when its feature has <= 256 bins, uint16 otherwise (the reference stores short).
__device__ __forceinline__ int load_bin(const void* const* cols, const int32_t* feat_bins, int f, int64_t r) {
    return feat_bins[f] > 256 ? (int)static_cast<const uint16_t*>(cols[f])[r] : (int)static_cast<const uint8_t*>(cols[f])[r];
}

__global__ void route_kernel(const void* const* __restrict__ cols, const int32_t* __restrict__ feat_bins, int64_t n,
                             int32_t* __restrict__ node_id, TreeView t, uint32_t* error_flag) {
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        int id = node_id[r];
        const int id0 = id;
        while (id > 0 && id < t.capacity) {
            const int f = t.feat[id];
            if (f < 0) break;
            const int bin = load_bin(cols, feat_bins, f, r);
            const uint8_t m = t.masks[t.mask_off[id] + (bin >> 3)];
            id = 2 * id + (((m >> (bin & 7)) & 1) ? 0 : 1);
        }
        if (id >= t.capacity) atomicOr(error_flag, 4u);  // child missing: the reference throws IllegalStateException (:1091)
        if (id != id0) node_id[r] = id;
    }
}

// dt/Loss.java:36-100, float arithmetic exactly as Java evaluates it.
__device__ __forceinline__ float loss_gradient(int kind, float predict, float label) {
    switch (kind) {
        case SB_LOSS_ABSOLUTE: {
            bool less;
            if (label < predict) less = true;
            else if (label > predict) less = false;
            else {  // Float.compare tie-break: -0.0f < 0.0f, NaN greatest
                int a = isnan(label) ? 0x7fc00000 : __float_as_int(label);
                int b = isnan(predict) ? 0x7fc00000 : __float_as_int(predict);
                less = a < b;
            }
            return less ? 1.0f : -1.0f;
        }
        case SB_LOSS_SQUARED: return 2.0f * (predict - label);
        case SB_LOSS_HALFGRADSQUARED: return predict - label;
        case SB_LOSS_LOG: {
            const float num = 2.0f - 4.0f * label;
            const float arg = (4.0f * label) * predict - 2.0f * predict;
            return num / (float)exp((double)arg);
        }
    }
    return 0.0f;
}
__device__ __forceinline__ float loss_error(int kind, float predict, float label) {
    switch (kind) {
        case SB_LOSS_ABSOLUTE: return fabsf(label - predict);
        case SB_LOSS_SQUARED:
        case SB_LOSS_HALFGRADSQUARED: {
            const float e = predict - label;
            return e * e;
        }
        case SB_LOSS_LOG: {
            const float arg = 2.0f * predict - (4.0f * predict) * label;
            return (float)log1p(1.0 + exp((double)arg));
        }
    }
    return 0.0f;
}

// dt/DTWorker.java:620-641 (training rows) and :721-735 (validation rows).
__global__ void gbt_update_kernel(int64_t n, const int32_t* __restrict__ node_id, TreeView prev, int is_first_tree,
                                  double learning_rate, int loss, const float* __restrict__ label,
                                  float* __restrict__ predict, float* __restrict__ output,
                                  const uint8_t* __restrict__ dropout) {
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const int id = node_id[r];
        if (id <= 0 || id >= prev.capacity || !prev.has_predict[id]) continue;
        const double p = prev.predict[id];
        float pr = predict[r];
        if (is_first_tree) pr = (float)p;
        else if (!(dropout && dropout[r])) pr = pr + (float)(learning_rate * p);
        predict[r] = pr;
        output[r] = -1.0f * loss_gradient(loss, pr, label[r]);
    }
}

// Per-block partial sums in a fixed order -> scratch; the last block to finish folds the partials, again in a fixed
// order, into acc.  Same input => same bits, whatever the scheduling.  Shared by errors_kernel and gbt_switch_kernel.
__device__ __forceinline__ void fold_error_sums(double tc, double vc, double te, double ve, double* scratch, double* acc) {
    __shared__ double sh[4][TPB / 32];
    double v[4] = {tc, vc, te, ve};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_down_sync(0xffffffffu, v[k], o);
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0)
        for (int k = 0; k < 4; ++k) sh[k][warp] = v[k];
    __syncthreads();
    __shared__ bool is_last;
    if (threadIdx.x == 0) {
        for (int k = 0; k < 4; ++k) {
            double s = 0;
            for (int w = 0; w < TPB / 32; ++w) s += sh[k][w];
            scratch[blockIdx.x * 4 + k] = s;
        }
        __threadfence();
        unsigned int* ticket_word = reinterpret_cast<unsigned int*>(scratch + ERR_BLOCKS * 4);
        const unsigned int ticket = atomicAdd(ticket_word, 1u);
        is_last = (ticket == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last && threadIdx.x < 4) {
        __threadfence();
        double s = 0;
        for (unsigned int b = 0; b < gridDim.x; ++b) s += __ldcg(&scratch[b * 4 + threadIdx.x]);
        acc[threadIdx.x] += s;
        if (threadIdx.x == 0) *reinterpret_cast<unsigned int*>(scratch + ERR_BLOCKS * 4) = 0u;
    }
}


// Error sums (dt/DTWorker.java:672-683, 737-748 GBT; 583-605, 691-705 RF).
__global__ void __launch_bounds__(TPB) errors_kernel(const ErrArgs a, double* scratch, double* acc) {
    double tc = 0, vc = 0, te = 0, ve = 0;
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < a.n; r += (int64_t)gridDim.x * blockDim.x) {
        float p;
        if (a.from_tree) {
            const int id = a.node_id[r];
            if (id <= 0 || id >= a.tree.capacity || !a.tree.has_predict[id]) continue;
            p = (float)a.tree.predict[id];
        } else {
            p = a.predict[r];
        }
        const float sig = a.significance[r];
        const float err = loss_error(a.loss, p, a.label[r]);
        if (a.is_validation) {
            ve += (double)(sig * err);
            vc += (double)sig;
        } else if (a.rf_oob) {
            const float w = a.weights ? a.weights[r] : 1.0f;  // rows adopted without a weight array: weight 1, never out-of-bag
            if (__float_as_int(w) == 0) {  // Float.compare(weight, 0f) == 0: out-of-bag row (dt/DTWorker.java:593-597)
                ve += (double)(sig * err);
                vc += (double)sig;
            } else {
                te += (double)((w * sig) * err);
                tc += (double)(w * sig);
            }
        } else {
            te += (double)(sig * err);
            tc += (double)sig;
        }
    }
    fold_error_sums(tc, vc, te, ve, scratch, acc);
}

// RF pass, rows side, in ONE launch for all trees (dt/DTWorker.java:583-605, 691-705): every row is walked from its node in
// every real tree to where the tree now sends it (predictNodeIndex continued), and the pass's error sums are taken
// over the predicts of those nodes.  A thread owns a row and loops over the trees: label and significance are read once
// per row, node id and bag weight once per (row, tree) -- coalesced, a tree's arrays are [n] -- and the bins the walks
// look at come from the row's line of the row-major twin, staged in shared memory by the block (one coalesced read
// of the 256 rows; row pitch + 4 bytes keeps the threads on different banks).  Without a twin (validation rows, the
// first pass) or for a uint16 column the column-major arrays are read.
// The kernel waits on dependent gathers (node id -> node -> bin -> mask -> child -> predict), so (1) a node is one
// 32-byte record with its left_mask inline (WalkNode): a row that moves one level costs two gathers, and (2) four trees
// are walked side by side, their loads issued together.
constexpr int WALK_MAX_FPAD = 256;  // twin rows up to this pitch are staged (256 rows x 260 B = 65 KB)
constexpr int WALK_TREES_SMEM = 128;
__device__ __forceinline__ WalkNode load_walk_node(const WalkNode* p) {
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(p));
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(p) + 1);
    WalkNode n;
    n.feat = (int32_t)a.x; n.pos = (int32_t)a.y; n.mask_off = a.z; n.predict = __uint_as_float(a.w);
    n.flags = b.x; n.pad = b.y; n.mask64 = ((uint64_t)b.w << 32) | b.z;
    return n;
}
__global__ void __launch_bounds__(TPB) forest_walk_kernel(const ForestWalkArgs a, double* scratch, double* acc) {
    extern __shared__ __align__(16) unsigned char walk_smem[];
    // the tree table is read by every thread for every tree: a copy in shared memory (the first WALK_TREES_SMEM trees;
    // a forest with more reads the rest from global memory)
    __shared__ WalkTree s_trees[WALK_TREES_SMEM];
    for (int i = threadIdx.x; i < min(a.n_trees, WALK_TREES_SMEM) * (int)(sizeof(WalkTree) / 8); i += TPB)
        reinterpret_cast<uint64_t*>(s_trees)[i] = reinterpret_cast<const uint64_t*>(a.trees)[i];
    __syncthreads();
    auto tree_of = [&](int t) -> const WalkTree& { return t < WALK_TREES_SMEM ? s_trees[t] : a.trees[t]; };
    double tc = 0, vc = 0, te = 0, ve = 0;
    const bool staged = a.rm_bins != nullptr && a.fpad <= (uint32_t)WALK_MAX_FPAD;
    const uint32_t pitch_s = a.fpad + 4u;
    constexpr int U = 4;
    for (int64_t r0 = (int64_t)blockIdx.x * TPB; r0 < a.n; r0 += (int64_t)gridDim.x * TPB) {
        if (staged) {
            __syncthreads();  // the previous step's readers are done
            const uint64_t first = (uint64_t)r0 * a.fpad, limit = (uint64_t)a.n * a.fpad;
            const uint32_t words = TPB * a.fpad / 4u;
            for (uint32_t w = threadIdx.x; w < words; w += TPB) {
                const uint64_t o = first + 4ull * w;
                const uint32_t v = o < limit ? __ldg(reinterpret_cast<const uint32_t*>(a.rm_bins + o)) : 0u;
                const uint32_t row = (4u * w) / a.fpad, col = (4u * w) - row * a.fpad;
                *reinterpret_cast<uint32_t*>(walk_smem + row * pitch_s + col) = v;
            }
            __syncthreads();
        }
        const int64_t r = r0 + threadIdx.x;
        if (r >= a.n) continue;
        const unsigned char* my_row = walk_smem + threadIdx.x * pitch_s;
        const float sig = a.significance[r], label = a.label[r];
        auto bin_of = [&](const WalkNode& nd) -> int {
            if (nd.pos >= 0 && a.rm_bins) return staged ? (int)my_row[nd.pos] : (int)a.rm_bins[(uint64_t)r * a.fpad + nd.pos];
            return load_bin(a.cols, a.feat_bins, nd.feat, r);
        };
        for (int t0 = 0; t0 < a.n_trees; t0 += U) {
            int id[U], id0[U];
            float w[U];
            WalkNode nd[U];
            bool live[U];
            // (requesting the node ids of the NEXT four trees here, ahead of the walk, made the kernel slower: 3.6 ms
            // against 3.0 per launch on C1 -- more registers, fewer resident blocks)
#pragma unroll
            for (int u = 0; u < U; ++u) {
                live[u] = t0 + u < a.n_trees;
                id[u] = live[u] ? tree_of(t0 + u).node_id[r] : 0;
                id0[u] = id[u];
                w[u] = 1.0f;
                if (live[u] && !a.is_validation && tree_of(t0 + u).weights) w[u] = tree_of(t0 + u).weights[r];
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                nd[u].feat = -1; nd[u].flags = 0; nd[u].predict = 0.f;
                if (live[u] && id[u] > 0 && id[u] < tree_of(t0 + u).capacity) nd[u] = load_walk_node(tree_of(t0 + u).nodes + id[u]);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {  // first step of every walk; a row that goes further (a tree seen for the first time) loops
                if (!live[u]) continue;
                const WalkTree& wt = tree_of(t0 + u);
                while (nd[u].feat >= 0) {
                    const int bin = bin_of(nd[u]);
                    bool left;
                    if (nd[u].flags & 2u) left = (nd[u].mask64 >> bin) & 1ull;
                    else left = (wt.masks[nd[u].mask_off + (bin >> 3)] >> (bin & 7)) & 1;
                    id[u] = 2 * id[u] + (left ? 0 : 1);
                    if (id[u] >= wt.capacity) {  // child missing: the reference throws IllegalStateException (:1091)
                        atomicOr(a.error_flag, 4u);
                        nd[u].feat = -1; nd[u].flags = 0;
                        break;
                    }
                    nd[u] = load_walk_node(wt.nodes + id[u]);
                }
                if (id[u] != id0[u]) wt.node_id[r] = id[u];
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (!live[u] || id[u] <= 0 || !(nd[u].flags & 1u)) continue;
                const float err = loss_error(a.loss, nd[u].predict, label);
                if (a.is_validation) {
                    ve += (double)(sig * err);
                    vc += (double)sig;
                } else if (__float_as_int(w[u]) == 0) {  // Float.compare(weight, 0f) == 0: out-of-bag row (dt/DTWorker.java:593-597)
                    ve += (double)(sig * err);
                    vc += (double)sig;
                } else {
                    te += (double)((w[u] * sig) * err);
                    tc += (double)(w[u] * sig);
                }
            }
        }
    }
    fold_error_sums(tc, vc, te, ve, scratch, acc);
}

// GBT tree switch in ONE pass over the rows (dt/DTWorker.java:620-641, 672-683 training; :721-748 validation): walk the
// finished tree from the row's node to its leaf (predictNodeIndex), update Data.predict / Data.output from the leaf,
// put the row back at the root for the next tree, add the row to the error sums over the new predict and -- training
// rows -- to the maxima behind the fixed-point scales of the next tree.  Replaces route + gbt_update + fill + errors +
// absmax (five passes over 20 bytes per row and one gathered bin per level walked).
__global__ void __launch_bounds__(TPB)
gbt_switch_kernel(const SwitchArgs a, double* scratch, double* acc) {
    double tc = 0, vc = 0, te = 0, ve = 0;
    float m0 = 0.f, m1 = 0.f, m2 = 0.f;
    // U rows per thread and step, their dependent loads (node id -> split feature -> bin -> mask) issued side by side:
    // one row at a time the kernel waits on four global-load latencies per row
    constexpr int U = 4;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t r0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r0 < a.n; r0 += U * stride) {
        int id[U], f[U], bin[U];
        float pr[U], out[U], lab[U], sig[U], w[U];
        uint8_t drop[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t r = r0 + u * stride;
            id[u] = r < a.n ? a.node_id[r] : 0;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t r = r0 + u * stride;
            f[u] = (id[u] > 0 && id[u] < a.prev.capacity) ? a.prev.feat[id[u]] : -1;
            if (r < a.n) {
                pr[u] = a.predict[r]; out[u] = a.output[r]; lab[u] = a.label[r]; sig[u] = a.significance[r];
                w[u] = (a.weights && !a.is_validation) ? a.weights[r] : 1.0f;
                drop[u] = a.dropout ? a.dropout[r] : (uint8_t)0;
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) bin[u] = f[u] >= 0 ? load_bin(a.cols, a.feat_bins, f[u], r0 + u * stride) : 0;
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t r = r0 + u * stride;
            if (r >= a.n) continue;
            int nid = id[u];
            if (f[u] >= 0) {
                const uint8_t m = a.prev.masks[a.prev.mask_off[nid] + (bin[u] >> 3)];
                nid = 2 * nid + (((m >> (bin[u] & 7)) & 1) ? 0 : 1);
                while (nid > 0 && nid < a.prev.capacity) {  // a tree seen for the first time: the rest of the way down
                    const int ff = a.prev.feat[nid];
                    if (ff < 0) break;
                    const int bb = load_bin(a.cols, a.feat_bins, ff, r);
                    const uint8_t mm = a.prev.masks[a.prev.mask_off[nid] + (bb >> 3)];
                    nid = 2 * nid + (((mm >> (bb & 7)) & 1) ? 0 : 1);
                }
            }
            if (nid >= a.prev.capacity) atomicOr(a.error_flag, 4u);
            if (nid > 0 && nid < a.prev.capacity && a.prev.has_predict[nid]) {
                const double p = a.prev.predict[nid];
                if (a.is_first_tree) pr[u] = (float)p;
                else if (!drop[u]) pr[u] = pr[u] + (float)(a.learning_rate * p);
                out[u] = -1.0f * loss_gradient(a.loss, pr[u], lab[u]);
                a.predict[r] = pr[u];
                a.output[r] = out[u];
            }
            a.node_id[r] = SB_ROOT_INDEX;
            const float err = loss_error(a.loss, pr[u], lab[u]);
            if (a.is_validation) {
                ve += (double)(sig[u] * err);
                vc += (double)sig[u];
            } else {
                te += (double)(sig[u] * err);
                tc += (double)sig[u];
                m0 = fmaxf(m0, fabsf(sig[u] * w[u]));
                m1 = fmaxf(m1, fabsf((out[u] * sig[u]) * w[u]));
                m2 = fmaxf(m2, fabsf(((out[u] * out[u]) * sig[u]) * w[u]));
                if (!isfinite(sig[u] * w[u]) || !isfinite((out[u] * sig[u]) * w[u]) || !isfinite(((out[u] * out[u]) * sig[u]) * w[u]))
                    m0 = __int_as_float(0x7f800000);
            }
        }
    }
    if (!a.is_validation) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            m0 = fmaxf(m0, __shfl_down_sync(0xffffffffu, m0, o));
            m1 = fmaxf(m1, __shfl_down_sync(0xffffffffu, m1, o));
            m2 = fmaxf(m2, __shfl_down_sync(0xffffffffu, m2, o));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMax(&a.maxbits[0], __float_as_uint(m0));
            atomicMax(&a.maxbits[1], __float_as_uint(m1));
            atomicMax(&a.maxbits[2], __float_as_uint(m2));
        }
    }
    fold_error_sums(tc, vc, te, ve, scratch, acc);
}

__global__ void __launch_bounds__(TPB) absmax_kernel(int64_t n, const float* __restrict__ output,
                                                    const float* __restrict__ significance,
                                                    const float* __restrict__ weights, uint32_t* out) {
    float m0 = 0.f, m1 = 0.f, m2 = 0.f;
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const float w = weights ? weights[r] : 1.0f;
        const float sig = significance[r];
        const float y = output[r];
        m0 = fmaxf(m0, fabsf(sig * w));
        m1 = fmaxf(m1, fabsf((y * sig) * w));
        m2 = fmaxf(m2, fabsf(((y * y) * sig) * w));
        // fmaxf drops NaN; make non-finite inputs visible as +inf so the host can refuse them
        if (!isfinite(sig * w) || !isfinite((y * sig) * w) || !isfinite(((y * y) * sig) * w)) m0 = __int_as_float(0x7f800000);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m0 = fmaxf(m0, __shfl_down_sync(0xffffffffu, m0, o));
        m1 = fmaxf(m1, __shfl_down_sync(0xffffffffu, m1, o));
        m2 = fmaxf(m2, __shfl_down_sync(0xffffffffu, m2, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMax(&out[0], __float_as_uint(m0));
        atomicMax(&out[1], __float_as_uint(m1));
        atomicMax(&out[2], __float_as_uint(m2));
    }
}

__global__ void check_int_weights_kernel(int64_t n, const float* __restrict__ significance,
                                         const float* __restrict__ weights, uint32_t* flag) {
    bool bad = false, not_one = false;
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const float w = weights ? weights[r] : 1.0f;
        const float c = significance[r] * w;
        bad |= !(c >= 0.0f && c <= 1024.0f && c == truncf(c));
        not_one |= w != 1.0f;
    }
    if (bad) atomicOr(flag, 1u);
    if (not_one) atomicOr(flag, 2u);  // bit 1: some sample weight differs from 1
}

__global__ void make_scales_kernel(const uint32_t* maxbits, int kind, float* scales, double* inv_scales, int* cnt_shift) {
    const int k = threadIdx.x;
    if (k >= 3) return;
    const float m = __uint_as_float(maxbits[k]);
    int e = 0;
    if (m > 0.0f && isfinite(m)) {
        e = QBITS - ilogbf(m) - 1;  // m * 2^e < 2^QBITS
        if (e > 100) e = 100;
        if (e < -100) e = -100;
    }
    if (kind == STAT_CLASS && k > 0) e = 0;
    scales[k] = ldexpf(1.0f, e);
    inv_scales[k] = ldexp(1.0, -e);
    if (k == 0) *cnt_shift = e > 0 ? e : 0;
}

// Counter-based generator of SURVEY.md 8(d): every (row, feature) value is a pure function of (seed, row, feature).
__host__ __device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

__global__ void synthetic_bins_kernel(uint8_t* bins, int64_t pitch, int n_features, const int32_t* feat_bins,
                                      int64_t row0, int64_t n, uint64_t seed) {
    // 8 rows per thread, one 8-byte store
    const int64_t groups = (n + 7) / 8;
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < groups * n_features;
         g += (int64_t)gridDim.x * blockDim.x) {
        const int f = (int)(g / groups);
        const int64_t rg = g - (int64_t)f * groups;
        const uint32_t nb = (uint32_t)feat_bins[f];
        uint64_t packed = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int64_t r = rg * 8 + k;
            const uint64_t h = splitmix64(seed ^ (uint64_t)((row0 + r) * (int64_t)n_features + f));
            const uint64_t b = (r < n && nb <= 256u) ? (uint64_t)((uint32_t)(h >> 33) % nb) : 0;
            packed |= b << (8 * k);
        }
        *reinterpret_cast<uint64_t*>(bins + (int64_t)f * pitch + rg * 8) = packed;
    }
}

__global__ void synthetic_labels_kernel(const uint8_t* bins, int64_t pitch, int n_features, const int32_t* feat_bins,
                                        float* label, float* significance, int64_t row0, int64_t n, uint64_t seed,
                                        int label_kind, int weighted) {
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t h = splitmix64(seed ^ 0xA5A5A5A5DEADBEEFULL ^ (uint64_t)(row0 + r));
        const float u = (float)(h >> 40) * (1.0f / 16777216.0f);          // [0,1)
        const float u2 = (float)((h >> 16) & 0xFFFFFF) * (1.0f / 16777216.0f);
        float y;
        if (label_kind == 0) {
            // 5-feature logistic with 10% flips (config C2)
            float z = 0.f;
            const int nf = n_features < 5 ? n_features : 5;
            for (int f = 0; f < nf; ++f) {
                const float x = (float)bins[(int64_t)f * pitch + r] / (float)feat_bins[f] - 0.5f;
                z += (f & 1 ? -2.0f : 3.0f) * x;
            }
            const float p = 1.0f / (1.0f + expf(-z));
            y = (u < p) ? 1.0f : 0.0f;
            if (u2 < 0.1f) y = 1.0f - y;
        } else {
            // threshold rule with noise (config C1)
            const float b0 = (float)bins[r], b1 = n_features > 1 ? (float)bins[pitch + r] : 0.f,
                        b2 = n_features > 2 ? (float)bins[2 * pitch + r] : 0.f;
            const float nz = 2.0f * u - 1.0f;
            y = ((b0 + b1 + 2.0f * b2) * 0.25f + 4.0f * nz > 0.5f * (float)feat_bins[0]) ? 1.0f : 0.0f;
        }
        label[r] = y;
        significance[r] = weighted ? (0.5f + 1.5f * u2) : 1.0f;
    }
}

__global__ void check_bins_kernel(const void* const* cols, const int32_t* feat_bins, int n_features, int64_t n,
                                  uint32_t* error_flag) {
    const int f = blockIdx.y;
    const int nb = feat_bins[f];
    bool bad = false;
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x)
        bad |= (load_bin(cols, feat_bins, f, r) >= nb);
    if (bad) atomicOr(error_flag, 2u);
    (void)n_features;
}

}  // namespace

cudaError_t launch_fill_i32(int32_t* p, int64_t n, int32_t v, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    fill_i32_kernel<<<grid_for(n, TPB, 148 * 8), TPB, 0, s>>>(p, n, v);
    return cudaGetLastError();
}
cudaError_t launch_fill_f32(float* p, int64_t n, float v, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    fill_f32_kernel<<<grid_for(n, TPB, 148 * 8), TPB, 0, s>>>(p, n, v);
    return cudaGetLastError();
}
cudaError_t launch_copy_f32(float* dst, const float* src, int64_t n, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    copy_f32_kernel<<<grid_for(n, TPB, 148 * 8), TPB, 0, s>>>(dst, src, n);
    return cudaGetLastError();
}
cudaError_t launch_route(const void* const* cols, const int32_t* feat_bins, int64_t n, int32_t* node_id, TreeView t,
                         uint32_t* error_flag, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    route_kernel<<<grid_for(n, TPB, 148 * 16), TPB, 0, s>>>(cols, feat_bins, n, node_id, t, error_flag);
    return cudaGetLastError();
}

__global__ void synthetic_bins16_kernel(uint16_t* col, int64_t n, int feature, int n_features, int n_bins, int64_t row0,
                                        uint64_t seed) {
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t h = splitmix64(seed ^ (uint64_t)((row0 + r) * (int64_t)n_features + feature));
        col[r] = (uint16_t)((uint32_t)(h >> 33) % (uint32_t)n_bins);
    }
}
cudaError_t launch_synthetic16(uint16_t* col, int64_t n, int feature, int n_features, int n_bins, int64_t row0,
                               uint64_t seed, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    synthetic_bins16_kernel<<<grid_for(n, TPB, 148 * 8), TPB, 0, s>>>(col, n, feature, n_features, n_bins, row0, seed);
    return cudaGetLastError();
}
cudaError_t launch_gbt_update(int64_t n, const int32_t* node_id, TreeView prev, int is_first_tree, double learning_rate,
                              int loss, const float* label, float* predict, float* output, const uint8_t* dropout,
                              cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    gbt_update_kernel<<<grid_for(n, TPB, 148 * 16), TPB, 0, s>>>(n, node_id, prev, is_first_tree, learning_rate, loss,
                                                                 label, predict, output, dropout);
    return cudaGetLastError();
}
cudaError_t launch_errors(const ErrArgs& a, double* scratch, double* acc, cudaStream_t s) {
    if (a.n <= 0) return cudaSuccess;
    errors_kernel<<<grid_for(a.n, TPB, ERR_BLOCKS), TPB, 0, s>>>(a, scratch, acc);
    return cudaGetLastError();
}
cudaError_t launch_forest_walk(const ForestWalkArgs& a, double* scratch, double* acc, cudaStream_t s) {
    if (a.n <= 0 || a.n_trees <= 0) return cudaSuccess;
    const bool staged = a.rm_bins != nullptr && a.fpad <= (uint32_t)WALK_MAX_FPAD;
    const size_t smem = staged ? (size_t)TPB * (a.fpad + 4u) : 0;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(forest_walk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TPB * (WALK_MAX_FPAD + 4));
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    forest_walk_kernel<<<grid_for(a.n, TPB, ERR_BLOCKS), TPB, smem, s>>>(a, scratch, acc);
    return cudaGetLastError();
}
cudaError_t launch_gbt_switch(const SwitchArgs& a, double* scratch, double* acc, cudaStream_t s) {
    if (a.n <= 0) return cudaSuccess;
    gbt_switch_kernel<<<grid_for(a.n, TPB, ERR_BLOCKS), TPB, 0, s>>>(a, scratch, acc);
    return cudaGetLastError();
}
cudaError_t launch_absmax(int64_t n, const float* output, const float* significance, const float* weights,
                          uint32_t* out, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    absmax_kernel<<<grid_for(n, TPB, 148 * 8), TPB, 0, s>>>(n, output, significance, weights, out);
    return cudaGetLastError();
}
cudaError_t launch_make_scales(const uint32_t* maxbits, int kind, float* scales, double* inv_scales, int* cnt_shift,
                               cudaStream_t s) {
    make_scales_kernel<<<1, 32, 0, s>>>(maxbits, kind, scales, inv_scales, cnt_shift);
    return cudaGetLastError();
}
cudaError_t launch_check_int_weights(int64_t n, const float* significance, const float* weights, uint32_t* flag,
                                     cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    check_int_weights_kernel<<<grid_for(n, TPB, 148 * 8), TPB, 0, s>>>(n, significance, weights, flag);
    return cudaGetLastError();
}
cudaError_t launch_synthetic(uint8_t* bins, int64_t pitch, int n_features, const int32_t* feat_bins, float* label,
                             float* significance, int64_t row0, int64_t n, uint64_t seed, int label_kind, int weighted,
                             cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    synthetic_bins_kernel<<<148 * 16, TPB, 0, s>>>(bins, pitch, n_features, feat_bins, row0, n, seed);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    synthetic_labels_kernel<<<grid_for(n, TPB, 148 * 16), TPB, 0, s>>>(bins, pitch, n_features, feat_bins, label,
                                                                       significance, row0, n, seed, label_kind, weighted);
    return cudaGetLastError();
}
cudaError_t launch_check_bins(const void* const* cols, const int32_t* feat_bins, int n_features, int64_t n,
                              uint32_t* error_flag, cudaStream_t s) {
    if (n <= 0 || n_features <= 0) return cudaSuccess;
    dim3 grid(grid_for(n, TPB, 148), n_features);
    check_bins_kernel<<<grid, TPB, 0, s>>>(cols, feat_bins, n_features, n, error_flag);
    return cudaGetLastError();
}

}  // namespace sb
