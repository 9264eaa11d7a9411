// hist.cu -- node x feature x bin statistics build (the hot kernel), sm_100a.
//
// Replaces the phase-B loop of DTWorker.doCompute (dt/DTWorker.java:802-862) and Impurity.featureUpdate
// (dt/Impurity.java:238-242 Variance/FriedmanMSE, :537-540 Entropy, :719-722 Gini).
//
// Shape of the kernel (DESIGN.md section 4 has the measurements behind each choice):
//   * one block = (row chunk) x (tile of 32 feature slots) x (group of GS todo-node slots);
//   * column tiles  [32 features][T_ROWS bins]  are staged into shared memory by the TMA engine
//     (cp.async.bulk, one 1-D bulk copy per feature column, completion on an mbarrier), NST stages deep;
//   * LANE == FEATURE: the 32 lanes of a warp work on the same row and on 32 different features, and the
//     privatised histogram is laid out [slot][bin][word][lane], so every shared-memory atomic of a warp
//     instruction hits 32 different banks and 32 different addresses -- no bank conflict and no
//     same-address serialisation can happen inside a warp (this is what warp-aggregation would buy,
//     without its __match_any cost);
//   * per-record contributions are the reference's fp32 products, converted to 64-bit fixed point
//     (q = rint(v * 2^e), |q| < 2^39) and added with native 32-bit ATOMS.ADD + carry: shared-memory
//     fp32/fp64/u64 atomicAdd compile to CAS loops on sm_100a and measured 6-10x slower (tools/ubench_smem.cu);
//   * at the end of its row chunk a block merges its privatised histogram into the global accumulators
//     (int64 lo/hi pairs, RED.ADD.64): integer addition is associative, so the merged result does not depend
//     on block order, launch geometry or the number of GPUs; the conversion to the reference's fp64
//     statistics happens once, in finalize_stats_kernel, with a single rounding per statistic.
#include "common.cuh"
#include "kernels.cuh"

namespace sb {

namespace {

constexpr int T_ROWS = HIST_T_ROWS;      // rows per tile
constexpr int NST = HIST_STAGES;         // TMA stages
constexpr int NTHREADS = HIST_THREADS;   // 16 warps
constexpr int NWARPS = NTHREADS / 32;
constexpr int ROWS_PER_WARP = T_ROWS / NWARPS;  // 8
constexpr int TILE_PITCH = T_ROWS + 16;  // bytes; pitch % 128 == 16 makes the per-lane 8/16-byte reads conflict-free
static_assert(ROWS_PER_WARP == 8, "the consumer loop reads 8 bins per lane (one LDS.64)");

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// TMA 1-D bulk copy global -> shared, completion signalled on the mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// 64-bit two's complement add into a (lo, hi) pair of shared words, lane-strided by 32 words.
__device__ __forceinline__ void smem_add64(uint32_t* lo_word, uint32_t lo, uint32_t hi) {
    uint32_t old = atomicAdd(lo_word, lo);
    uint32_t carry = (uint32_t)((old + lo) < old);
    atomicAdd(lo_word + 32, hi + carry);
}

struct __align__(16) RowVals {  // what every lane needs to know about one row (read as two broadcast LDS.128)
    int32_t slot;        // local slot in this block's group, or -1: skip the row
    uint32_t fmask;      // bit l set: feature l of the tile is in this node's feature list
    uint32_t q[6];       // VARIANCE: (cnt_lo, cnt_hi, sum_lo, sum_hi, sq_lo, sq_hi); CLASS: (lo, hi, class, -, -, -)
};

}  // namespace

template <int KIND>
__global__ void __launch_bounds__(NTHREADS, 1) hist_kernel(const HistArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr int NW = (KIND == STAT_VARIANCE) ? 6 : 0;  // words per bin for VARIANCE; CLASS uses 2*C at run time
    const int nw = (KIND == STAT_VARIANCE) ? NW : 2 * a.stats_size;

    // ---- carve shared memory
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);                         // NST barriers
    unsigned char* tiles = smem + 128;                                           // NST * 32 * TILE_PITCH
    RowVals* vals = reinterpret_cast<RowVals*>(tiles + NST * 32 * TILE_PITCH);   // 2 * T_ROWS
    uint32_t* hist = reinterpret_cast<uint32_t*>(vals + 2 * T_ROWS);             // gs * bins_pad * nw * 32 words

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    // ---- which work item
    const int items_per_chunk = a.n_tiles * a.n_groups;
    const int chunk = blockIdx.x / items_per_chunk;
    const int rem = blockIdx.x - chunk * items_per_chunk;
    const int tile_id = rem / a.n_groups;
    const int group = rem - tile_id * a.n_groups;
    const int g0 = group * a.gs;
    const int gcount = min(a.gs, a.n_slots - g0);
    const int64_t row_begin = (int64_t)chunk * a.rows_per_chunk;
    const int64_t row_end = min(a.n_rows, row_begin + a.rows_per_chunk);
    const int n_tile_iters = (int)((row_end - row_begin + T_ROWS - 1) / T_ROWS);

    const int my_feat = a.tile_feats[tile_id * 32 + lane];  // -1: padding lane
    const int my_bins = my_feat >= 0 ? a.feat_bins[my_feat] : 0;
    const unsigned char* my_col = my_feat >= 0 ? static_cast<const unsigned char*>(a.cols[my_feat]) : nullptr;
    const int bins_pad = a.bins_pad;
    const int hist_words = gcount * bins_pad * nw * 32;

    for (int i = threadIdx.x; i < hist_words; i += NTHREADS) hist[i] = 0u;
    if (threadIdx.x == 0) {
        for (int s = 0; s < NST; ++s) mbar_init(&bars[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const float sc0 = a.scales[0], sc1 = a.scales[1], sc2 = a.scales[2];

    auto issue_tile = [&](int it) {  // called by warp 0 only
        const int stage = it % NST;
        const int64_t r0 = row_begin + (int64_t)it * T_ROWS;
        const int rows = (int)min((int64_t)T_ROWS, row_end - r0);
        const uint32_t bytes = (uint32_t)((rows + 15) & ~15);
        const unsigned valid = __ballot_sync(0xffffffffu, my_feat >= 0);
        if (lane == 0) mbar_expect_tx(&bars[stage], bytes * __popc(valid));
        __syncwarp();
        if (my_feat >= 0) tma_load_1d(tiles + (stage * 32 + lane) * TILE_PITCH, my_col + r0, bytes, &bars[stage]);
    };

    if (warp == 0)
        for (int it = 0; it < NST - 1 && it < n_tile_iters; ++it) issue_tile(it);

    for (int it = 0; it < n_tile_iters; ++it) {
        const int stage = it % NST;
        const int64_t r0 = row_begin + (int64_t)it * T_ROWS;
        const int rows = (int)min((int64_t)T_ROWS, row_end - r0);
        RowVals* v = vals + (it & 1) * T_ROWS;

        // ---- per-row values: one thread per row of the tile (dt/Impurity.java:239-241: fp32 products, left to right)
        if (threadIdx.x < T_ROWS) {
            RowVals rv;
            rv.slot = -1;
            rv.fmask = 0;
#pragma unroll
            for (int k = 0; k < 6; ++k) rv.q[k] = 0;
            if (threadIdx.x < rows) {
                const int64_t r = r0 + threadIdx.x;
                const int nid = a.node_id[r];
                int s = (nid >= 0 && nid < a.slot_table_size) ? (int)a.slot_of_node[nid] : -1;
                const float w = a.weights ? a.weights[r] : 1.0f;
                if (s >= g0 && s < g0 + gcount && w != 0.0f) {
                    const float sig = a.significance[r];
                    const float y = a.output[r];
                    rv.slot = s - g0;
                    rv.fmask = a.slot_fmask[(size_t)s * a.n_tiles + tile_id];
                    const float cw = __fmul_rn(sig, w);
                    const long long qc = __float2ll_rn(__fmul_rn(cw, sc0));
                    if (KIND == STAT_VARIANCE) {
                        const float sw = __fmul_rn(__fmul_rn(y, sig), w);
                        const float qw = __fmul_rn(__fmul_rn(__fmul_rn(y, y), sig), w);
                        const long long qs = __float2ll_rn(__fmul_rn(sw, sc1));
                        const long long qq = __float2ll_rn(__fmul_rn(qw, sc2));
                        rv.q[0] = (uint32_t)qc; rv.q[1] = (uint32_t)(qc >> 32);
                        rv.q[2] = (uint32_t)qs; rv.q[3] = (uint32_t)(qs >> 32);
                        rv.q[4] = (uint32_t)qq; rv.q[5] = (uint32_t)(qq >> 32);
                    } else {
                        const int cls = (int)(__fadd_rn(y, 0.000001f));  // dt/Impurity.java:539,721
                        rv.q[0] = (uint32_t)qc; rv.q[1] = (uint32_t)(qc >> 32);
                        rv.q[2] = (uint32_t)cls;
                        if (cls < 0 || cls >= a.stats_size) { rv.slot = -1; atomicOr(a.error_flag, 1u); }
                    }
                }
            }
            v[threadIdx.x] = rv;
        }
        __syncthreads();  // vals visible; every warp is done with tile it-1, so its stage may be refilled

        if (warp == 0 && it + NST - 1 < n_tile_iters) issue_tile(it + NST - 1);

        mbar_wait(&bars[stage], (uint32_t)((it / NST) & 1));

        // ---- consume: warp w owns rows [w*8, w*8+8) of the tile; lane == feature
        const int wr0 = warp * ROWS_PER_WARP;
        if (wr0 < rows) {
            const uint2 packed = *reinterpret_cast<const uint2*>(tiles + (stage * 32 + lane) * TILE_PITCH + wr0);
#pragma unroll
            for (int j = 0; j < ROWS_PER_WARP; ++j) {
                const uint4 v0 = *reinterpret_cast<const uint4*>(&v[wr0 + j]);           // slot, fmask, q0, q1
                if ((int)v0.x < 0) continue;                                              // warp-uniform
                const uint4 v1 = *reinterpret_cast<const uint4*>(&v[wr0 + j].q[2]);       // q2..q5
                const uint32_t word = (j < 4) ? packed.x : packed.y;
                const int bin = (int)((word >> ((j & 3) * 8)) & 0xffu);
                if (!((v0.y >> lane) & 1u)) continue;
                if (bin >= my_bins) { atomicOr(a.error_flag, 2u); continue; }
                if (KIND == STAT_VARIANCE) {
                    uint32_t* p = hist + (((int)v0.x * bins_pad + bin) * NW) * 32 + lane;
                    smem_add64(p, v0.z, v0.w);
                    smem_add64(p + 64, v1.x, v1.y);
                    smem_add64(p + 128, v1.z, v1.w);
                } else {
                    uint32_t* p = hist + (((int)v0.x * bins_pad + bin) * nw + 2 * (int)v1.x) * 32 + lane;
                    smem_add64(p, v0.z, v0.w);
                }
            }
        }
    }
    __syncthreads();

    // ---- merge the privatised histogram into the global integer accumulators
    const int per_slot = bins_pad * (nw / 2) * 32;  // 64-bit statistics per slot
    for (int i = threadIdx.x; i < gcount * per_slot; i += NTHREADS) {
        const int l = i & 31;
        int t = i >> 5;
        const int st = t % (nw / 2);
        t /= (nw / 2);
        const int bin = t % bins_pad;
        const int sl = t / bins_pad;
        const uint32_t* p = hist + ((sl * bins_pad + bin) * nw + 2 * st) * 32 + l;
        const uint32_t lo = p[0], hi = p[32];
        if ((lo | hi) == 0u) continue;
        const int f = a.tile_feats[tile_id * 32 + l];
        if (f < 0) continue;
        const int64_t off = a.pair_off[(size_t)(g0 + sl) * a.n_features + f];
        if (off < 0) continue;
        const int64_t idx = off + (int64_t)bin * a.stats_size + st;
        atomicAdd(reinterpret_cast<unsigned long long*>(a.raw_lo + idx), (unsigned long long)lo);
        atomicAdd(reinterpret_cast<unsigned long long*>(a.raw_hi + idx), (unsigned long long)(long long)(int32_t)hi);
    }
}

size_t hist_smem_bytes(int gs, int bins_pad, int nw) {
    return 128 + (size_t)NST * 32 * TILE_PITCH + 2 * T_ROWS * sizeof(RowVals) + (size_t)gs * bins_pad * nw * 32 * 4;
}

int hist_max_slots_per_block(int bins_pad, int nw) {
    const size_t fixed = 128 + (size_t)NST * 32 * TILE_PITCH + 2 * T_ROWS * sizeof(RowVals);
    const size_t budget = 227 * 1024 - fixed;
    return (int)(budget / ((size_t)bins_pad * nw * 32 * 4));
}

cudaError_t hist_launch(int kind, const HistArgs& a, int grid, size_t smem, cudaStream_t stream) {
    static bool attr_set[2] = {false, false};
    if (kind == STAT_VARIANCE) {
        if (!attr_set[0]) {
            cudaError_t e = cudaFuncSetAttribute(hist_kernel<STAT_VARIANCE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
            if (e != cudaSuccess) return e;
            attr_set[0] = true;
        }
        hist_kernel<STAT_VARIANCE><<<grid, NTHREADS, smem, stream>>>(a);
    } else {
        if (!attr_set[1]) {
            cudaError_t e = cudaFuncSetAttribute(hist_kernel<STAT_CLASS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
            if (e != cudaSuccess) return e;
            attr_set[1] = true;
        }
        hist_kernel<STAT_CLASS><<<grid, NTHREADS, smem, stream>>>(a);
    }
    return cudaGetLastError();
}

}  // namespace sb
