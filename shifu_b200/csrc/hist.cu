// hist.cu -- node x feature x bin statistics build (the hot kernel), sm_100a.
//
// Replaces the phase-B loop of DTWorker.doCompute (dt/DTWorker.java:802-862) and Impurity.featureUpdate
// (dt/Impurity.java:238-242 Variance/FriedmanMSE, :537-540 Entropy, :719-722 Gini).
//
// Shape of the kernel (DESIGN.md section 4 has the measurements behind each choice):
//   * one block = (row chunk) x (tile of 32 feature slots) x (group of GS todo-node slots);
//   * warp-specialised: 4 PRODUCER warps turn per-row scalars into fixed-point operands and drive the TMA engine
//     (cp.async.bulk.tensor.2d: ONE box of [32 features][128 rows] bytes per stage from the column-major bin
//     matrix, 128-byte swizzled, completion on an mbarrier -- 32 separate 128-byte bulk copies per tile measured
//     ~3.5x slower end to end); 16 CONSUMER warps do the shared-memory atomics.  NST stages, full/empty
//     mbarriers, no block-wide barrier inside the row loop;
//   * LANE == FEATURE: the 32 lanes of a consumer warp work on the same row and on 32 different features, and
//     the privatised histogram is laid out [slot][bin][word][lane], so every shared-memory atomic of a warp
//     instruction hits 32 different banks and 32 different addresses -- no bank conflict and no same-address
//     serialisation can happen inside a warp (what warp-aggregation would buy, without its __match_any cost);
//   * per-record contributions are the reference's fp32 products, converted to 64-bit fixed point
//     (q = rint(v * 2^e), |q| < 2^44) and added with native 32-bit ATOMS.ADD + carry: shared-memory
//     fp32/fp64/u64 atomicAdd compile to CAS loops on sm_100a and measured 6-10x slower (tools/ubench_smem.cu).
//     When every sig*weight product of the pass is a small integer (the default: no weight column, 0/1 bagging
//     mask, Poisson bag counts) the count statistic is a single 32-bit add;
//   * at the end of its row chunk a block merges its privatised histogram into the global accumulators
//     (int64 lo/hi pairs, RED.ADD.64): integer addition is associative, so the merged result does not depend
//     on block order, launch geometry or the number of GPUs; the conversion to the reference's fp64
//     statistics happens once, in finalize_stats_kernel, with a single rounding per statistic.
#include <cuda.h>  // CUtensorMap (type only; the encoder is fetched through cudaGetDriverEntryPoint in api.cu)

#include <utility>

#include "common.cuh"
#include "kernels.cuh"

namespace sb {

namespace {

constexpr int T_ROWS = HIST_T_ROWS;             // rows per tile
constexpr int NST = HIST_STAGES;                // pipeline stages
constexpr int N_PROD_WARPS = T_ROWS / 32;       // one producer thread per tile row
constexpr int N_CONS_WARPS = 16;
constexpr int NTHREADS = (N_PROD_WARPS + N_CONS_WARPS) * 32;
constexpr int ROWS_PER_WARP = T_ROWS / N_CONS_WARPS;  // 8
constexpr int TILE_BYTES = 32 * T_ROWS;         // one TMA box: 32 feature rows of 128 bytes, SWIZZLE_128B
static_assert(T_ROWS == 128, "the TMA box is 128 bytes wide (the 128-byte swizzle span)");
constexpr int SLOT_TABLE_SMEM = 512;            // heap ids below this are mapped to slots from shared memory
static_assert(N_PROD_WARPS == 4 && N_CONS_WARPS == 16, "match[] packs 4 consumer warps per producer warp");
static_assert(ROWS_PER_WARP == 8, "the consumer loop reads 8 bins per lane (one LDS.64)");
static_assert(HIST_THREADS == NTHREADS, "kernels.cuh and hist.cu disagree on the block size");

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// 32-bit shared-window addressing for the hot loop: ATOMS with an immediate offset, no 64-bit pointer math.
template <int OFF>
__device__ __forceinline__ void sred_add(uint32_t addr, uint32_t v) {
    asm volatile("red.shared.add.u32 [%0+%1], %2;" ::"r"(addr), "n"(OFF), "r"(v) : "memory");
}
// hi + carry(old_lo + lo): kept as an opaque carry chain so the compiler cannot pull part of it next to the atomic
// that produced old_lo (that would stall the warp on the ATOMS return once per row).
__device__ __forceinline__ uint32_t hi_plus_carry(uint32_t old_lo, uint32_t lo, uint32_t hi) {
    uint32_t r;
    asm volatile("{\n.reg .u32 t;\nadd.cc.u32 t, %1, %2;\naddc.u32 %0, %3, 0;\n}" : "=r"(r) : "r"(old_lo), "r"(lo), "r"(hi));
    return r;
}
template <int OFF>
__device__ __forceinline__ uint32_t satom_add(uint32_t addr, uint32_t v) {
    uint32_t old;
    asm volatile("atom.shared.add.u32 %0, [%1+%2], %3;" : "=r"(old) : "r"(addr), "n"(OFF), "r"(v) : "memory");
    return old;
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
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
// TMA 2-D tiled copy global -> shared, completion signalled on the mbarrier (SASS: UTMALDG).
// c0 = first row of the box (inner, contiguous dimension), c1 = first feature slot.  Out-of-range parts of
// the box (tail rows, feature slots >= F) are zero-filled by the hardware.
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int32_t c0, int32_t c1, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
            smem_u32(smem_dst)),
        "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar))
        : "memory");
}

struct __align__(16) RowVals {  // what every consumer lane needs to know about one row (two broadcast LDS.128)
    uint32_t slot_off;   // byte offset of the row's slot inside the block's histogram (CLASS: + the class words)
    uint32_t fmask;      // bit l set: feature l of the tile takes this row; 0 = skip the row
    uint32_t q[6];       // VARIANCE: (cnt_lo, cnt_hi, sum_lo, sum_hi, sq_lo, sq_hi); CLASS: (lo, hi, -, -, -, -)
};

struct Stage {
    uint64_t tile_full;   // TMA bytes of the stage have landed
    uint64_t vals_full;   // the producers wrote every RowVals of the stage
    uint64_t empty;       // every consumer warp is done with the stage
    uint8_t match[N_PROD_WARPS];  // match[p] bit k: consumer warp 4p+k has at least one row to process
    uint8_t pad[8 - N_PROD_WARPS];
};

constexpr size_t CTRL_BYTES = 128 * ((sizeof(Stage) * NST + 127) / 128);
constexpr size_t SMEM_FIXED = (size_t)NST * TILE_BYTES                   // bin tiles (first: 1024-byte aligned)
                              + CTRL_BYTES                              // stage control
                              + SLOT_TABLE_SMEM * sizeof(int16_t)       // slot table
                              + (size_t)NST * T_ROWS * sizeof(RowVals); // row operands

}  // namespace

// words per bin of the privatised histogram
__host__ __device__ inline int hist_words(int kind, int cnt_int, int stats_size) {
    if (kind == STAT_VARIANCE) return cnt_int ? 5 : 6;
    return cnt_int ? stats_size : 2 * stats_size;
}

// Merge a block's privatised histogram into the global integer accumulators.  (lo, hi) split: the global sum over
// all blocks and GPUs may exceed 63 bits, each half cannot.
template <int KIND, int CNT_INT, int LW>
__device__ __forceinline__ void flush_hist(const uint32_t* hist, int gcount, int g0, int bins_pad, int nw, int n_stats,
                                           int cnt_shift, const int32_t* tile_feats, const int64_t* pair_off,
                                           int n_features, int64_t* raw_lo, int64_t* raw_hi, int nthreads) {
    // LW = features (words) per histogram word row: 32*FPL laid out [k][lane], or 8 for the wide-feature kernel.
    // A thread owns one column l (one feature) and every (nthreads / LW)-th (bin, statistic) of it, so the feature id
    // and the block offset are loaded once per slot instead of once per entry: the loop body is shared-memory reads
    // and fire-and-forget RED.ADD.64, no dependent global load.
    const int groups = nthreads / LW;
    const int l = threadIdx.x % LW, q = threadIdx.x / LW;
    if (q >= groups) return;
    const int f = tile_feats[l];
    if (f < 0) return;
    const int per_slot = bins_pad * n_stats;
    for (int sl = 0; sl < gcount; ++sl) {
        const int64_t off = pair_off[(size_t)(g0 + sl) * n_features + f];
        if (off < 0) continue;
        for (int e = q; e < per_slot; e += groups) {
            const int bin = e / n_stats, stt = e - bin * n_stats;
            const uint32_t* pb = hist + ((sl * bins_pad + bin) * nw) * LW + l;
            long long v;
            if (KIND == STAT_VARIANCE) {
                if (CNT_INT) {
                    if (stt == 0) v = (long long)((unsigned long long)pb[0] << cnt_shift);
                    else v = (long long)(((unsigned long long)pb[(2 * stt) * LW] << 32) | pb[(2 * stt - 1) * LW]);
                } else {
                    v = (long long)(((unsigned long long)pb[(2 * stt + 1) * LW] << 32) | pb[(2 * stt) * LW]);
                }
            } else {
                if (CNT_INT) v = (long long)((unsigned long long)pb[stt * LW] << cnt_shift);
                else v = (long long)(((unsigned long long)pb[(2 * stt + 1) * LW] << 32) | pb[(2 * stt) * LW]);
            }
            if (v == 0) continue;
            const int64_t idx = off + (int64_t)bin * n_stats + stt;
            atomicAdd(reinterpret_cast<unsigned long long*>(raw_lo + idx), (unsigned long long)(v & 0xffffffffLL));
            if ((v >> 32) != 0) atomicAdd(reinterpret_cast<unsigned long long*>(raw_hi + idx), (unsigned long long)(v >> 32));
        }
    }
}

template <int KIND, int CNT_INT>
__global__ void __launch_bounds__(NTHREADS, 1) hist_kernel(const HistArgs a, const __grid_constant__ CUtensorMap tmap) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const int nw = hist_words(KIND, CNT_INT, a.stats_size);

    // ---- carve shared memory
    unsigned char* tiles = smem;  // the 128-byte swizzle pattern repeats every 1024 bytes: keep tiles 1024-aligned
    Stage* stages = reinterpret_cast<Stage*>(smem + (size_t)NST * TILE_BYTES);
    int16_t* slot_tab = reinterpret_cast<int16_t*>(smem + (size_t)NST * TILE_BYTES + CTRL_BYTES);
    RowVals* vals = reinterpret_cast<RowVals*>(slot_tab + SLOT_TABLE_SMEM);
    uint32_t* hist = reinterpret_cast<uint32_t*>(vals + NST * T_ROWS);

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
    const int bins_pad = a.bins_pad;
    const int hist_words_total = gcount * bins_pad * nw * 32;

    for (int i = threadIdx.x; i < hist_words_total; i += NTHREADS) hist[i] = 0u;
    const int tab_n = min(a.slot_table_size, SLOT_TABLE_SMEM);
    for (int i = threadIdx.x; i < SLOT_TABLE_SMEM; i += NTHREADS) slot_tab[i] = i < tab_n ? a.slot_of_node[i] : (int16_t)-1;
    if (threadIdx.x == 0) {
        for (int s = 0; s < NST; ++s) {
            mbar_init(&stages[s].tile_full, 1);
            mbar_init(&stages[s].vals_full, N_PROD_WARPS * 32);
            mbar_init(&stages[s].empty, N_CONS_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp < N_PROD_WARPS) {
        // =========================================================== producers
        const int tid = threadIdx.x;  // row of the tile
        const float sc0 = a.scales[0], sc1 = a.scales[1], sc2 = a.scales[2];

        // software prefetch of the next tile's per-row scalars
        int nid = 0;
        float w = 0.f, sig = 0.f, y = 0.f;
        auto prefetch = [&](int it) {
            const int64_t r = row_begin + (int64_t)it * T_ROWS + tid;
            if (it < n_tile_iters && r < row_end) {
                nid = a.node_id[r];
                w = a.weights ? a.weights[r] : 1.0f;
                sig = a.significance[r];
                y = a.output[r];
            } else {
                nid = -1;
                w = 0.f;
            }
        };
        prefetch(0);
        for (int it = 0; it < n_tile_iters; ++it) {
            const int s = it % NST;
            Stage& st = stages[s];
            if (it >= NST) mbar_wait(&st.empty, (uint32_t)(((it / NST) - 1) & 1));
            const int64_t r0 = row_begin + (int64_t)it * T_ROWS;
            if (threadIdx.x == 0) {  // one TMA box per stage: 32 consecutive feature slots x rows [r0, r0+128)
                mbar_expect_tx(&st.tile_full, TILE_BYTES);
                tma_load_2d(tiles + s * TILE_BYTES, &tmap, (int32_t)r0, a.tile_slot0[tile_id], &st.tile_full);
            }
            // ---- this thread's row: the reference's fp32 products (dt/Impurity.java:239-241), then fixed point
            const int cur_nid = nid;
            const float cw_ = w, csig = sig, cy = y;
            prefetch(it + 1);
            RowVals rv;
            rv.slot_off = 0;
            rv.fmask = 0;
#pragma unroll
            for (int k = 0; k < 6; ++k) rv.q[k] = 0;
            int sl = -1;
            if (cur_nid >= 0) {
                if (cur_nid < SLOT_TABLE_SMEM) sl = slot_tab[cur_nid];
                else if (cur_nid < a.slot_table_size) sl = a.slot_of_node[cur_nid];
            }
            if (sl >= g0 && sl < g0 + gcount && cw_ != 0.0f) {
                rv.slot_off = (uint32_t)((sl - g0) * bins_pad * nw * 128);
                rv.fmask = a.slot_fmask[(size_t)sl * a.n_tiles + tile_id];
                const float cw = __fmul_rn(csig, cw_);
                long long qc;
                if (CNT_INT) qc = (long long)__float2int_rn(cw);  // exact: the host verified sig*w is a small integer
                else qc = __float2ll_rn(__fmul_rn(cw, sc0));
                if (KIND == STAT_VARIANCE) {
                    const float sw = __fmul_rn(__fmul_rn(cy, csig), cw_);
                    const float qw = __fmul_rn(__fmul_rn(__fmul_rn(cy, cy), csig), cw_);
                    const long long qs = __float2ll_rn(__fmul_rn(sw, sc1));
                    const long long qq = __float2ll_rn(__fmul_rn(qw, sc2));
                    rv.q[0] = (uint32_t)qc; rv.q[1] = (uint32_t)(qc >> 32);
                    rv.q[2] = (uint32_t)qs; rv.q[3] = (uint32_t)(qs >> 32);
                    rv.q[4] = (uint32_t)qq; rv.q[5] = (uint32_t)(qq >> 32);
                } else {
                    const int cls = (int)(__fadd_rn(cy, 0.000001f));  // dt/Impurity.java:539,721
                    rv.q[0] = (uint32_t)qc; rv.q[1] = (uint32_t)(qc >> 32);
                    if (cls < 0 || cls >= a.stats_size) { rv.fmask = 0; atomicOr(a.error_flag, 1u); }
                    else rv.slot_off += (uint32_t)(cls * (CNT_INT ? 128 : 256));
                }
            }
            vals[s * T_ROWS + tid] = rv;
            // which consumer warps have work: consumer warp c owns rows [8c, 8c+8)
            const unsigned hit = __ballot_sync(0xffffffffu, rv.fmask != 0u);
            if (lane == 0) {
                uint32_t m = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if ((hit >> (8 * k)) & 0xffu) m |= 1u << k;
                st.match[warp] = (uint8_t)m;
            }
            mbar_arrive(&st.vals_full);
        }
    } else {
        // =========================================================== consumers
        const int cw = warp - N_PROD_WARPS;
        const int wr0 = cw * ROWS_PER_WARP;
        const uint32_t lane_bit = 1u << lane;
        const uint32_t lane_base = smem_u32(hist) + (uint32_t)lane * 4u;
        const uint32_t bin_stride = (uint32_t)nw * 128u;
        for (int it = 0; it < n_tile_iters; ++it) {
            const int s = it % NST;
            Stage& st = stages[s];
            const uint32_t parity = (uint32_t)((it / NST) & 1);
            mbar_wait(&st.vals_full, parity);
            if ((st.match[cw >> 2] >> (cw & 3)) & 1u) {
                mbar_wait(&st.tile_full, parity);
                const RowVals* v = vals + s * T_ROWS + wr0;
                // row `lane` of the box; its 16-byte chunk c is stored at chunk (c ^ (lane & 7))  [SWIZZLE_128B]
                const uint2 packed = *reinterpret_cast<const uint2*>(
                    tiles + s * TILE_BYTES + lane * T_ROWS + ((((cw >> 1) ^ (lane & 7)) << 4) | ((cw & 1) << 3)));
                // Two phases so that the 8 rows' returning atomics are all in flight before any carry is needed
                // (one row at a time the loop is bound by the ATOMS return latency, not by its throughput).
                // Bin indices were validated against n_bins when the rows were loaded (check_bins_kernel).
                uint32_t ad[ROWS_PER_WARP];
                uint32_t lo0[ROWS_PER_WARP], lo1[ROWS_PER_WARP], lo2[ROWS_PER_WARP];  // low words added
                uint32_t hi0[ROWS_PER_WARP], hi1[ROWS_PER_WARP], hi2[ROWS_PER_WARP];  // high words to add
                uint32_t o0[ROWS_PER_WARP], o1[ROWS_PER_WARP], o2[ROWS_PER_WARP];     // returned old low words
                bool ok[ROWS_PER_WARP];
#pragma unroll
                for (int j = 0; j < ROWS_PER_WARP; ++j) {
                    const uint4 v0 = *reinterpret_cast<const uint4*>(&v[j]);  // slot_off, fmask, q0, q1
                    const uint32_t word = (j < 4) ? packed.x : packed.y;
                    const uint32_t bin = (word >> ((j & 3) * 8)) & 0xffu;
                    ok[j] = (v0.y & lane_bit) != 0u;
                    ad[j] = lane_base + v0.x + bin * bin_stride;
                    lo0[j] = v0.z; hi0[j] = v0.w;
                    lo1[j] = hi1[j] = lo2[j] = hi2[j] = 0;
                    o0[j] = o1[j] = o2[j] = 0;
                    if (KIND == STAT_VARIANCE) {
                        const uint4 v1 = *reinterpret_cast<const uint4*>(&v[j].q[2]);  // q2..q5
                        lo1[j] = v1.x; hi1[j] = v1.y; lo2[j] = v1.z; hi2[j] = v1.w;
                    }
                    if (ok[j]) {
                        if (KIND == STAT_VARIANCE) {
                            if (CNT_INT) {
                                sred_add<0>(ad[j], lo0[j]);
                                o1[j] = satom_add<128>(ad[j], lo1[j]);
                                o2[j] = satom_add<384>(ad[j], lo2[j]);
                            } else {
                                o0[j] = satom_add<0>(ad[j], lo0[j]);
                                o1[j] = satom_add<256>(ad[j], lo1[j]);
                                o2[j] = satom_add<512>(ad[j], lo2[j]);
                            }
                        } else {
                            if (CNT_INT) sred_add<0>(ad[j], lo0[j]);
                            else o0[j] = satom_add<0>(ad[j], lo0[j]);
                        }
                    }
                }
#pragma unroll
                for (int j = 0; j < ROWS_PER_WARP; ++j) {
                    if (!ok[j]) continue;
                    if (KIND == STAT_VARIANCE) {
                        if (CNT_INT) {
                            sred_add<256>(ad[j], hi_plus_carry(o1[j], lo1[j], hi1[j]));
                            sred_add<512>(ad[j], hi_plus_carry(o2[j], lo2[j], hi2[j]));
                        } else {
                            sred_add<128>(ad[j], hi_plus_carry(o0[j], lo0[j], hi0[j]));
                            sred_add<384>(ad[j], hi_plus_carry(o1[j], lo1[j], hi1[j]));
                            sred_add<640>(ad[j], hi_plus_carry(o2[j], lo2[j], hi2[j]));
                        }
                    } else if (!CNT_INT) {
                        sred_add<128>(ad[j], hi_plus_carry(o0[j], lo0[j], hi0[j]));
                    }
                }
            } else {
                // nothing to do for this warp, but the TMA of the stage must have landed before the stage is reused
                mbar_wait(&st.tile_full, parity);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&st.empty);
        }
    }
    __syncthreads();

    flush_hist<KIND, CNT_INT, 32>(hist, gcount, g0, bins_pad, nw, a.stats_size, CNT_INT ? *a.cnt_shift : 0, a.tile_feats + tile_id * 32,
                                 a.pair_off, a.n_features, a.raw_lo, a.raw_hi, NTHREADS);
}

size_t hist_smem_bytes(int gs, int bins_pad, int nw) { return SMEM_FIXED + (size_t)gs * bins_pad * nw * 32 * 4; }

int hist_max_slots_per_block(int bins_pad, int nw) {
    const size_t budget = 227 * 1024 - SMEM_FIXED;
    return (int)(budget / ((size_t)bins_pad * nw * 32 * 4));
}

int hist_words_per_bin(int kind, int cnt_int, int stats_size) { return hist_words(kind, cnt_int, stats_size); }

template <int KIND, int CNT_INT>
static cudaError_t launch_one(const HistArgs& a, const CUtensorMap& tmap, int grid, size_t smem, cudaStream_t stream) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(hist_kernel<KIND, CNT_INT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    hist_kernel<KIND, CNT_INT><<<grid, NTHREADS, smem, stream>>>(a, tmap);
    return cudaGetLastError();
}

cudaError_t hist_launch(int kind, int cnt_int, const HistArgs& a, const void* tensor_map, int grid, size_t smem,
                        cudaStream_t stream) {
    const CUtensorMap& tmap = *static_cast<const CUtensorMap*>(tensor_map);
    if (kind == STAT_VARIANCE)
        return cnt_int ? launch_one<STAT_VARIANCE, 1>(a, tmap, grid, smem, stream) : launch_one<STAT_VARIANCE, 0>(a, tmap, grid, smem, stream);
    return cnt_int ? launch_one<STAT_CLASS, 1>(a, tmap, grid, smem, stream) : launch_one<STAT_CLASS, 0>(a, tmap, grid, smem, stream);
}


// =====================================================================================================================
// hist_rows_kernel: the same accumulation driven by a per-group row list (rowlist.cu).  A block takes one chunk of one
// slot group's list and one tile of 32*FPL consecutive feature slots.  One producer warp streams the RowOps of the
// chunk into shared memory with TMA 1-D bulk copies (4 KB per stage); the consumer warps read, per row, the 32*FPL
// bins of the tile from the row-major bin copy with ONE coalesced load (lane l holds features FPL*l .. FPL*l+FPL-1),
// then issue the two-phase atomics.  The per-row operands are warp-uniform and cost as many LSU cycles to broadcast
// (8 words) as the atomics of 32 cells do; with FPL = 4 that cost is shared by 128 cells.  The histogram is laid out
// [slot][bin][word][k][lane] (feature = FPL*lane + k), so for every k a warp's atomics still hit 32 distinct banks.
namespace {
#ifndef SB_R_NST
#define SB_R_NST 4  // stages of the RowOps ring
#endif
#ifndef SB_ROWS_PF
#define SB_ROWS_PF 2  // tiles whose bins are in flight ahead of the one being accumulated
#endif
constexpr int R_NST = SB_R_NST;
constexpr int R_TILE = 128;  // rows per stage
constexpr size_t R_SMEM_FIXED = (size_t)R_NST * R_TILE * sizeof(RowOps) + 128 /* barriers */ + 256 /* fmask table */;
struct RStage { uint64_t full, empty; };
constexpr int R_MAX_GS = 16;  // fm_tab: 16 slots x FPL words <= 64 words
template <int FPL> struct RCfg {
    static constexpr int CONS_WARPS = 16;
    static constexpr int NTHREADS = (1 + CONS_WARPS) * 32;
    static constexpr int RPW = R_TILE / CONS_WARPS;                   // rows of a tile per consumer warp (8)
    static constexpr int BATCH = FPL == 4 ? 2 : (FPL == 2 ? 4 : 8);  // rows whose atomics are in flight together
};
}  // namespace

namespace {
template <int... Ks, class F>
__device__ __forceinline__ void static_for(std::integer_sequence<int, Ks...>, F&& f) {
    (f(std::integral_constant<int, Ks>{}), ...);
}
}  // namespace

// The consumer loop of hist_rows_kernel for a tile whose lanes hold KACT features each (KACT == FPL for a full tile).
// KACT is a template parameter so that the k loop has no branches and the k * 128 byte offsets fold into the
// immediates of the atomics.
template <int KIND, int CNT_INT, int FPL, int ALLF, int KACT>
__device__ __forceinline__ void rows_consume(const HistRowsArgs& a, const RowOps* ops_s, RStage* st, const uint32_t* fm_tab,
                                             uint32_t hist_base, int cw, int lane, int tile_id, int gcount, int64_t begin,
                                             int64_t end, int n_iters, int nw) {
    constexpr int RPW = RCfg<FPL>::RPW;
    constexpr int BATCH = RCfg<FPL>::BATCH;
    constexpr uint32_t WORD = 128u * FPL;                           // bytes between the words of one bin
    constexpr auto KS = std::make_integer_sequence<int, KACT>{};
    const uint32_t lane_bit = 1u << lane;
    const uint32_t lane_base = hist_base + (uint32_t)lane * 4u;
    const uint32_t bin_stride = (uint32_t)nw * 128u * FPL;         // bytes between bins
    const uint32_t slot_stride = (uint32_t)a.bins_pad * bin_stride;  // bytes between slots
    const uint8_t* my_bins = a.rm_bins + a.tile_slot0[tile_id] + lane * FPL;
    // ALLF (every todo node takes every feature): a (lane, k) without a feature reads a zero padding byte of the twin
    // and accumulates into its own, never flushed column, so the hot loop carries no per-cell feature test; a lane
    // without any feature does not load at all.  Otherwise the per-slot feature masks decide cell by cell.
    bool lane_loads = !ALLF;
#pragma unroll
    for (int k = 0; k < FPL; ++k)
        if (a.tile_feats[tile_id * 32 * FPL + k * 32 + lane] >= 0) lane_loads = true;
    const int wr0 = cw * RPW;
    // per-lane dummy bin right behind the block's histogram (one bin's worth of words)
    const uint32_t dummy_addr = lane_base + (uint32_t)gcount * slot_stride;

    // The bins of the next PF tiles are requested before the atomics of tile `it` are issued, so the global-load latency
    // hides behind the shared-memory work of the tiles in between (one tile ahead left the consumer warps waiting on
    // the loads for a quarter of their time: ncu, long_scoreboard at the first use of a tile's bins).
    constexpr int PF = SB_ROWS_PF;
    static_assert(PF >= 1 && PF < R_NST, "the RowOps ring must hold the tiles whose bins are in flight");
    uint32_t bw[PF + 1][RPW];
    auto fetch_bins = [&](int it, uint32_t (&b)[RPW]) {
        const int s = it % R_NST;
        mbar_wait(&st[s].full, (uint32_t)((it / R_NST) & 1));
        const int rows = (int)min((int64_t)R_TILE, end - (begin + (int64_t)it * R_TILE));
        const RowOps* v = ops_s + s * R_TILE + wr0;
#pragma unroll
        for (int j = 0; j < RPW; ++j) {
            b[j] = 0;
            if (wr0 + j < rows && lane_loads) {
                const uint8_t* src = my_bins + (size_t)v[j].row * a.fpad;
                if (FPL == 4) b[j] = __ldg(reinterpret_cast<const uint32_t*>(src));
                else if (FPL == 2) b[j] = __ldg(reinterpret_cast<const unsigned short*>(src));
                else b[j] = __ldg(src);
            }
        }
    };
#pragma unroll
    for (int d = 0; d < PF; ++d)
        if (d < n_iters) fetch_bins(d, bw[d]);
    for (int it = 0; it < n_iters; ++it) {
        const int s = it % R_NST;
        if (it + PF < n_iters) fetch_bins(it + PF, bw[PF]);
        const int rows = (int)min((int64_t)R_TILE, end - (begin + (int64_t)it * R_TILE));
        const RowOps* v = ops_s + s * R_TILE + wr0;
#pragma unroll
        for (int jb = 0; jb < RPW; jb += BATCH) {
            if (wr0 + jb >= rows) break;  // warp-uniform
            // Every atomic below is unconditional: a cell that must not count (row past the end of the list, skipped
            // row, feature not in the node's list) is redirected to the per-lane dummy bin behind the histogram.
            // No predicates, no divergence, no reconvergence bookkeeping in the hot loop.
            uint32_t ad[BATCH][KACT];
            uint32_t o0[BATCH][KACT], o1[BATCH][KACT], o2[BATCH][KACT];
            uint32_t lo0[BATCH], lo1[BATCH], lo2[BATCH], hi0[BATCH], hi1[BATCH], hi2[BATCH];
#pragma unroll
            for (int j = 0; j < BATCH; ++j) {  // phase 1: operands, addresses, low-word atomics
                uint4 v1;  // sum / sq words (VARIANCE) or the class byte offset (CLASS)
                if (KIND == STAT_VARIANCE) v1 = *reinterpret_cast<const uint4*>(&v[jb + j].q[0]);
                else v1 = make_uint4(v[jb + j].q[0], 0u, 0u, 0u);
                uint32_t slot;
                lo1[j] = hi1[j] = lo2[j] = hi2[j] = 0;
                if (KIND == STAT_VARIANCE && CNT_INT) {
                    // packed operands: one 128-bit broadcast load carries slot, count and both statistics
                    slot = rowops_upper(v1.w);
                    lo0[j] = rowops_upper(v1.y); hi0[j] = 0;
                    lo1[j] = v1.x; hi1[j] = rowops_hi_signed(v1.y); lo2[j] = v1.z; hi2[j] = rowops_hi_unsigned(v1.w);
                } else {
                    const uint4 v0 = *reinterpret_cast<const uint4*>(&v[jb + j]);  // row, cnt_hi, slot, cnt_lo
                    slot = v0.z;
                    lo0[j] = v0.w; hi0[j] = v0.y;
                    if (KIND == STAT_VARIANCE) { lo1[j] = v1.x; hi1[j] = v1.y; lo2[j] = v1.z; hi2[j] = v1.w; }
                }
                const bool rowok = (wr0 + jb + j < rows) && slot < (uint32_t)gcount;
                uint32_t base = lane_base + slot * slot_stride;
                if (KIND != STAT_VARIANCE) base += v1.x * FPL;  // class words
                // ALLF: one select per row; a bad row lands in the dummy bin whatever its bins are
                const uint32_t rbase = rowok ? base : dummy_addr;
                const uint32_t bmul = rowok ? bin_stride : 0u;
                const uint32_t bwj = bw[0][jb + j];
                // feature masks of the row's slot: the FPL words sit side by side, one vector load
                uint32_t fmw[FPL];
                if (!ALLF) {
                    const uint32_t* fmp = fm_tab + (slot & (R_MAX_GS - 1)) * FPL;
                    if (FPL == 4) {
                        const uint4 t = *reinterpret_cast<const uint4*>(fmp);
                        fmw[0] = t.x; fmw[1] = t.y; fmw[2] = t.z; fmw[3] = t.w;
                    } else if (FPL == 2) {
                        const uint2 t = *reinterpret_cast<const uint2*>(fmp);
                        fmw[0] = t.x; fmw[1] = t.y;
                    } else {
                        fmw[0] = fmp[0];
                    }
                }
                static_for(KS, [&](auto kc) {
                    constexpr int k = decltype(kc)::value;
                    constexpr uint32_t KO = 128u * k;
                    const uint32_t bin = (k == 3) ? (bwj >> 24) : ((bwj >> (8 * k)) & 0xffu);
                    uint32_t adr;
                    if (ALLF) adr = rbase + bin * bmul;
                    else adr = (rowok && (fmw[k] & lane_bit) != 0u) ? base + bin * bin_stride : dummy_addr;
                    ad[j][k] = adr;
                    if (KIND == STAT_VARIANCE) {
                        if (CNT_INT) {
                            if (CNT_INT != 2) sred_add<KO>(adr, lo0[j]);  // 2: the counts of this pass are known already
                            o1[j][k] = satom_add<KO + WORD>(adr, lo1[j]);
                            o2[j][k] = satom_add<KO + 3 * WORD>(adr, lo2[j]);
                        } else {
                            o0[j][k] = satom_add<KO>(adr, lo0[j]);
                            o1[j][k] = satom_add<KO + 2 * WORD>(adr, lo1[j]);
                            o2[j][k] = satom_add<KO + 4 * WORD>(adr, lo2[j]);
                        }
                    } else {
                        if (CNT_INT) sred_add<KO>(adr, lo0[j]);
                        else o0[j][k] = satom_add<KO>(adr, lo0[j]);
                    }
                });
            }
#pragma unroll
            for (int j = 0; j < BATCH; ++j) {  // phase 2: carries
                static_for(KS, [&](auto kc) {
                    constexpr int k = decltype(kc)::value;
                    constexpr uint32_t KO = 128u * k;
                    const uint32_t adr = ad[j][k];
                    if (KIND == STAT_VARIANCE) {
                        if (CNT_INT) {
                            sred_add<KO + 2 * WORD>(adr, hi_plus_carry(o1[j][k], lo1[j], hi1[j]));
                            sred_add<KO + 4 * WORD>(adr, hi_plus_carry(o2[j][k], lo2[j], hi2[j]));
                        } else {
                            sred_add<KO + WORD>(adr, hi_plus_carry(o0[j][k], lo0[j], hi0[j]));
                            sred_add<KO + 3 * WORD>(adr, hi_plus_carry(o1[j][k], lo1[j], hi1[j]));
                            sred_add<KO + 5 * WORD>(adr, hi_plus_carry(o2[j][k], lo2[j], hi2[j]));
                        }
                    } else if (!CNT_INT) {
                        sred_add<KO + WORD>(adr, hi_plus_carry(o0[j][k], lo0[j], hi0[j]));
                    }
                });
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&st[s].empty);
#pragma unroll
        for (int d = 0; d < PF; ++d)
#pragma unroll
            for (int j = 0; j < RPW; ++j) bw[d][j] = bw[d + 1][j];
    }
}

template <int KIND, int CNT_INT, int FPL, int ALLF>
__global__ void __launch_bounds__(RCfg<FPL>::NTHREADS, 1) hist_rows_kernel(const HistRowsArgs a) {
    constexpr int CONS_WARPS = RCfg<FPL>::CONS_WARPS;
    constexpr int NTHR = RCfg<FPL>::NTHREADS;
    extern __shared__ __align__(1024) unsigned char smem[];
    // Tile-major block order: the blocks of a full tile (fpl rounds of atomics per row) are dispatched before those of
    // the partly filled last tile (kact rounds), so the shorter blocks fill the tail of the launch.
    const int n_chunks = *a.n_chunks;
    if (n_chunks <= 0) return;  // no active row on this GPU in this pass (a shard whose todo nodes are all empty)
    const int tile_id = blockIdx.x / n_chunks;
    const int ci = blockIdx.x - tile_id * n_chunks;
    if (tile_id >= a.n_tiles) return;  // the grid is an upper bound; uniform for the block
    const int nw = hist_words(KIND, CNT_INT, a.stats_size);
    RowOps* ops_s = reinterpret_cast<RowOps*>(smem);
    RStage* st = reinterpret_cast<RStage*>(smem + (size_t)R_NST * R_TILE * sizeof(RowOps));
    uint32_t* fm_tab = reinterpret_cast<uint32_t*>(smem + (size_t)R_NST * R_TILE * sizeof(RowOps) + 128);
    uint32_t* hist = reinterpret_cast<uint32_t*>(smem + R_SMEM_FIXED);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int group = a.chunk_group[ci];
    const int64_t begin = a.chunk_begin[ci], end = a.chunk_end[ci];
    const int g0 = group * a.gs;
    const int gcount = min(a.gs, a.n_slots - g0);
    const int bins_pad = a.bins_pad;
    const int n_iters = (int)((end - begin + R_TILE - 1) / R_TILE);

    // the dummy bin behind the histogram is zeroed too: it only ever absorbs additions
    for (int i = threadIdx.x; i < (gcount * bins_pad + 1) * nw * 32 * FPL; i += NTHR) hist[i] = 0u;
    if (!ALLF && threadIdx.x < R_MAX_GS * FPL) {  // ALLF: no masks (slot_fmask may be null)
        const int sl = threadIdx.x / FPL, k = threadIdx.x % FPL;
        fm_tab[threadIdx.x] = sl < gcount ? a.slot_fmask[((size_t)(g0 + sl) * a.n_tiles + tile_id) * FPL + k] : 0u;
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < R_NST; ++s) {
            mbar_init(&st[s].full, 1);
            mbar_init(&st[s].empty, CONS_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == 0) {
        if (lane == 0) {
            for (int it = 0; it < n_iters; ++it) {
                const int s = it % R_NST;
                if (it >= R_NST) mbar_wait(&st[s].empty, (uint32_t)(((it / R_NST) - 1) & 1));
                const int64_t r0 = begin + (int64_t)it * R_TILE;
                const uint32_t bytes = (uint32_t)(min((int64_t)R_TILE, end - r0) * (int64_t)sizeof(RowOps));
                mbar_expect_tx(&st[s].full, bytes);
                tma_load_1d(ops_s + s * R_TILE, a.ops + r0, bytes, &st[s].full);
            }
        }
    } else {
        const int cw = warp - 1;
        const int kact = a.tile_kact[tile_id];  // a partly filled tile spreads its features over the lanes first
        const uint32_t hb = smem_u32(hist);
#define SB_CONSUME(K) rows_consume<KIND, CNT_INT, FPL, ALLF, K>(a, ops_s, st, fm_tab, hb, cw, lane, tile_id, gcount, begin, end, n_iters, nw)
        if (kact >= FPL) SB_CONSUME(FPL);
        else if constexpr (FPL == 4) {
            if (kact == 3) SB_CONSUME(3);
            else if (kact == 2) SB_CONSUME(2);
            else SB_CONSUME(1);
        } else if constexpr (FPL == 2) SB_CONSUME(1);
#undef SB_CONSUME
    }
    __syncthreads();
    flush_hist<KIND, CNT_INT, 32 * FPL>(hist, gcount, g0, bins_pad, nw, a.stats_size, CNT_INT ? *a.cnt_shift : 0,
                                   a.tile_feats + tile_id * 32 * FPL, a.pair_off, a.n_features, a.raw_lo, a.raw_hi, NTHR);
}

size_t hist_rows_smem_bytes(int gs, int bins_pad, int nw, int fpl) {
    return R_SMEM_FIXED + (size_t)gs * bins_pad * nw * 32 * fpl * 4 + (size_t)nw * 32 * fpl * 4 /* dummy bin */;
}
int hist_rows_max_slots_per_block(int bins_pad, int nw, int fpl) {
    const size_t budget = 227 * 1024 - R_SMEM_FIXED - (size_t)nw * 32 * fpl * 4;
    int gs = (int)(budget / ((size_t)bins_pad * nw * 32 * fpl * 4));
    return gs > R_MAX_GS ? R_MAX_GS : gs;
}

template <int KIND, int CNT_INT, int FPL, int ALLF>
static cudaError_t launch_rows_one(const HistRowsArgs& a, int grid, size_t smem, cudaStream_t stream) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(hist_rows_kernel<KIND, CNT_INT, FPL, ALLF>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    hist_rows_kernel<KIND, CNT_INT, FPL, ALLF><<<grid, RCfg<FPL>::NTHREADS, smem, stream>>>(a);
    return cudaGetLastError();
}

template <int KIND, int CNT_INT>
static cudaError_t launch_rows_kc(int fpl, int allf, const HistRowsArgs& a, int grid, size_t smem, cudaStream_t stream) {
    if (fpl == 4) return allf ? launch_rows_one<KIND, CNT_INT, 4, 1>(a, grid, smem, stream) : launch_rows_one<KIND, CNT_INT, 4, 0>(a, grid, smem, stream);
    if (fpl == 2) return allf ? launch_rows_one<KIND, CNT_INT, 2, 1>(a, grid, smem, stream) : launch_rows_one<KIND, CNT_INT, 2, 0>(a, grid, smem, stream);
    return allf ? launch_rows_one<KIND, CNT_INT, 1, 1>(a, grid, smem, stream) : launch_rows_one<KIND, CNT_INT, 1, 0>(a, grid, smem, stream);
}

cudaError_t hist_rows_launch(int kind, int cnt_int, int fpl, int allf, const HistRowsArgs& a, int grid, size_t smem,
                             cudaStream_t stream) {
    if (kind == STAT_VARIANCE && cnt_int == 2) {  // integer counts that are not accumulated (cached root counts)
        if (!allf) return cudaErrorInvalidValue;
        if (fpl == 4) return launch_rows_one<STAT_VARIANCE, 2, 4, 1>(a, grid, smem, stream);
        if (fpl == 2) return launch_rows_one<STAT_VARIANCE, 2, 2, 1>(a, grid, smem, stream);
        return launch_rows_one<STAT_VARIANCE, 2, 1, 1>(a, grid, smem, stream);
    }
    if (kind == STAT_VARIANCE)
        return cnt_int ? launch_rows_kc<STAT_VARIANCE, 1>(fpl, allf, a, grid, smem, stream) : launch_rows_kc<STAT_VARIANCE, 0>(fpl, allf, a, grid, smem, stream);
    return cnt_int ? launch_rows_kc<STAT_CLASS, 1>(fpl, allf, a, grid, smem, stream) : launch_rows_kc<STAT_CLASS, 0>(fpl, allf, a, grid, smem, stream);
}


// =====================================================================================================================
// hist_rows_class_kernel: class counts (Gini / Entropy) with integer weights over a forest list (rowlist.cu, forest_*).
// A cell costs ONE shared-memory word update here, so the instruction stream around the atomics decides the speed
// (the general kernel spent 61 instructions per row and tile on this case and ran at 61 % issue utilisation, ncu): the
// operands of a pair are 8 bytes (row; byte offset of its (slot, class) words and count in one word), read from the
// ring with broadcast loads, and a row and k cost one byte extract, one multiply-add and the atomic.
// Every slot takes every feature of the tile: features outside a node's list accumulate into columns that are never
// flushed (pair_off < 0), which costs nothing extra next to redirecting them.  The bins of two tiles are in flight
// while a third is accumulated.
// The kernel is persistent: one block per SM takes chunks off a work counter in table order, which is slab-major
// (forest_chunks_kernel), so the blocks sweep the rows together and share the twin's lines through L2.
namespace {
constexpr int C_NST = 8;                   // 1 KB stages: the operand ring runs well ahead of the bins
constexpr int C_WARPS = 15;                // consumer warps; with the producer warp 512 threads at 128 registers (30 warps of 4 rows at 64 registers: 15.4 ms against 14.3 on C1)
constexpr int C_RPW = 8;                   // rows of a tile per consumer warp
constexpr int C_TILE = C_WARPS * C_RPW;    // pairs per stage (120)
constexpr int C_NTHR = (1 + C_WARPS) * 32;
constexpr int C_STAGE = C_TILE + 2;               // a bulk copy starts and ends on 16 bytes: up to one record of the neighbours on either side
struct CStage { uint64_t full, empty; };
constexpr size_t C_SMEM_FIXED = (size_t)C_NST * C_STAGE * sizeof(ClassOps) + 128 + 16;  // ring, barriers, work item

template <int FPL, int KACT>
__device__ __noinline__ void class_consume(const HistClassArgs& a, const ClassOps* ops_s, CStage* st, uint32_t hist_base,
                                              int cw, int lane, int tile_id, int64_t begin, int64_t end, int n_iters,
                                              uint32_t it_base, uint32_t bin_stride) {
    constexpr int RPW = C_RPW;
    constexpr auto KS = std::make_integer_sequence<int, KACT>{};
    const uint32_t lane_base = hist_base + (uint32_t)lane * 4u;
    // Every lane loads, whatever its features: a byte that belongs to no used feature is a bin of an unused feature or
    // padding (zero) -- never more than bins_pad, which spans all uint8 features -- and lands in a column that is
    // not flushed.  The base is made opaque so that the address of a row is one 32 x 32 + 64 multiply-add.
    const uint8_t* my_bins = a.rm_bins + a.tile_slot0[tile_id] + lane * FPL;
    asm volatile("" : "+l"(my_bins));
    const int wr0 = cw * RPW;
    const uint32_t fpad = a.fpad;
    const int skew = (int)(begin & 1);  // records are 8 bytes, bulk copies start on 16: every stage begins one record early then

    // Only the bins travel through registers: three buffers of RPW words rotate statically (the loop is unrolled by
    // three; copying a buffer whose loads are still in flight would wait for them, and so would spilling it).  The
    // (offset, count) word of a row is read from the ring when its atomics are issued, so a ring slot is released
    // after the accumulation; the ring is deep enough (C_NST stages) for the producer to stay ahead.
    struct Buf { uint32_t bw[RPW]; };
    Buf b0, b1, b2;
    auto load_bins = [&](uint32_t row) -> uint32_t {
        const uint8_t* src = my_bins + (uint64_t)row * fpad;
        if (FPL == 4) return __ldg(reinterpret_cast<const uint32_t*>(src));
        if (FPL == 2) return __ldg(reinterpret_cast<const unsigned short*>(src));
        return __ldg(src);
    };
    // bins of tile `it` requested.  A record past the end of the chunk is whatever the ring slot held before (an older
    // pair, or the zeros the ring starts with): a valid row to load bins from; accumulate() gives it count zero.
    auto fetch = [&](int it, Buf& b) {
        const uint32_t g = it_base + (uint32_t)it;
        const int s = (int)(g % C_NST);
        mbar_wait(&st[s].full, (g / C_NST) & 1u);
        const ClassOps* v = ops_s + s * C_STAGE + skew + wr0;
#pragma unroll
        for (int j = 0; j < RPW; ++j) b.bw[j] = load_bins(v[j].row);
    };
    auto accumulate = [&](int it, const Buf& b) {
        const uint32_t g = it_base + (uint32_t)it;
        const int s = (int)(g % C_NST);
        const int mine = (int)min((int64_t)C_TILE, end - (begin + (int64_t)it * C_TILE)) - wr0;
        const ClassOps* v = ops_s + s * C_STAGE + skew + wr0;
        uint32_t where[RPW];
#pragma unroll
        for (int j = 0; j < RPW; ++j) where[j] = v[j].where;
        if (mine < RPW) {  // warp-uniform, the last tile of a chunk only
#pragma unroll
            for (int j = 0; j < RPW; ++j)
                if (j >= mine) where[j] = 0u;
        }
#pragma unroll
        for (int j = 0; j < RPW; ++j) {
            const uint32_t base = lane_base + (where[j] & ((1u << CLASSOPS_OFF_BITS) - 1u));
            const uint32_t cnt = where[j] >> CLASSOPS_OFF_BITS;
            static_for(KS, [&](auto kc) {
                constexpr int k = decltype(kc)::value;
                const uint32_t bin = __byte_perm(b.bw[j], 0u, 0x4440u + (uint32_t)k);
                sred_add<128 * k>(base + bin * bin_stride, cnt);
            });
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&st[s].empty);
    };
    if (n_iters > 0) fetch(0, b0);
    if (n_iters > 1) fetch(1, b1);
#pragma unroll 1
    for (int it = 0; it < n_iters; it += 3) {
        if (it + 2 < n_iters) fetch(it + 2, b2);
        accumulate(it, b0);
        if (it + 1 >= n_iters) break;
        if (it + 3 < n_iters) fetch(it + 3, b0);
        accumulate(it + 1, b1);
        if (it + 2 >= n_iters) break;
        if (it + 4 < n_iters) fetch(it + 4, b1);
        accumulate(it + 2, b2);
    }
}
}  // namespace

template <int FPL>
__global__ void __launch_bounds__(C_NTHR, 1) hist_rows_class_kernel(const HistClassArgs a) {
    constexpr int CONS_WARPS = C_WARPS;
    constexpr int NTHR = C_NTHR;
    extern __shared__ __align__(1024) unsigned char smem[];
    const int n_chunks = a.n_chunks[0];
    const int n_items = n_chunks * a.n_tiles;  // tile-major: item = tile * n_chunks + chunk
    const int nw = a.stats_size;               // one 32-bit word per class
    ClassOps* ops_s = reinterpret_cast<ClassOps*>(smem);
    CStage* st = reinterpret_cast<CStage*>(smem + (size_t)C_NST * C_STAGE * sizeof(ClassOps));
    int* s_item = reinterpret_cast<int*>(smem + (size_t)C_NST * C_STAGE * sizeof(ClassOps) + 128);
    uint32_t* hist = reinterpret_cast<uint32_t*>(smem + C_SMEM_FIXED);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int bins_pad = a.bins_pad;
    const uint32_t hb = smem_u32(hist);
    const uint32_t bin_stride = (uint32_t)nw * 128u * FPL;

    for (int i = threadIdx.x; i < C_NST * C_STAGE * 2; i += NTHR) reinterpret_cast<uint32_t*>(ops_s)[i] = 0u;  // see class_consume
    if (threadIdx.x == 0) {
        for (int s = 0; s < C_NST; ++s) {
            mbar_init(&st[s].full, 1);
            mbar_init(&st[s].empty, CONS_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    uint32_t it_base = 0;  // tiles that went through the ring so far: the barriers' phases run on across chunks
    for (;;) {
        __syncthreads();  // the previous item is flushed; s_item may be overwritten
        if (threadIdx.x == 0) *s_item = atomicAdd(&a.n_chunks[a.counter], 1);
        __syncthreads();
        const int item = *s_item;
        if (item >= n_items) break;
        const int tile_id = item / n_chunks;
        const int ci = item - tile_id * n_chunks;
        const int group = a.chunk_group[ci];
        const int64_t begin = a.chunk_begin[ci], end = a.chunk_end[ci];
        if (end <= begin) continue;
        const int g0 = group * a.gs;
        const int gcount = min(a.gs, a.n_slots - g0);
        const int n_iters = (int)((end - begin + C_TILE - 1) / C_TILE);

        for (int i = threadIdx.x; i < gcount * bins_pad * nw * 32 * FPL; i += NTHR) hist[i] = 0u;
        __syncthreads();

        if (warp == 0) {
            if (lane == 0) {
                for (int it = 0; it < n_iters; ++it) {
                    const uint32_t g = it_base + (uint32_t)it;
                    const int s = (int)(g % C_NST);
                    if (g >= (uint32_t)C_NST) mbar_wait(&st[s].empty, ((g / C_NST) - 1u) & 1u);
                    const int64_t r0 = begin + (int64_t)it * C_TILE;
                    const int64_t e0 = r0 & ~(int64_t)1;  // first record copied (16-byte aligned), e1: one past the last, even
                    const int64_t e1 = (min(r0 + C_TILE, end) + 1) & ~(int64_t)1;
                    const uint32_t bytes = (uint32_t)((e1 - e0) * (int64_t)sizeof(ClassOps));
                    mbar_expect_tx(&st[s].full, bytes);
                    tma_load_1d(ops_s + s * C_STAGE, a.ops + e0, bytes, &st[s].full);
                }
            }
        } else {
            const int cw = warp - 1;
            const int kact = a.tile_kact[tile_id];
#define SB_CCONSUME(K) class_consume<FPL, K>(a, ops_s, st, hb, cw, lane, tile_id, begin, end, n_iters, it_base, bin_stride)
            if (kact >= FPL) SB_CCONSUME(FPL);
            else if constexpr (FPL == 4) {
                if (kact == 3) SB_CCONSUME(3);
                else if (kact == 2) SB_CCONSUME(2);
                else SB_CCONSUME(1);
            } else if constexpr (FPL == 2) SB_CCONSUME(1);
#undef SB_CCONSUME
        }
        it_base += (uint32_t)n_iters;
        __syncthreads();
        flush_hist<STAT_CLASS, 1, 32 * FPL>(hist, gcount, g0, bins_pad, nw, a.stats_size, *a.cnt_shift, a.tile_feats + tile_id * 32 * FPL,
                                            a.pair_off, a.n_features, a.raw_lo, a.raw_hi, NTHR);
    }
}

// hist_rows_class_wide_kernel: the same pairs for the features with more than 256 bins (uint16 columns).  Tiles of 8
// features, a warp instruction covers 4 rows x 8 features (lane = (row q = lane >> 3, feature j = lane & 7)), histogram
// [slot][bin][class][8]; persistent blocks and work counter as above.  A pair's ClassOps word holds the byte offset of
// its (slot, class) in the NARROW tiles' histogram: slot and class are recovered from it (one division per row).
namespace {
constexpr int CW_RPW = 8;                     // rows of a tile per consumer warp = 2 steps of 4 rows
constexpr int CW_TILE = C_WARPS * CW_RPW;     // 120, as for the narrow tiles
}  // namespace

__global__ void __launch_bounds__(C_NTHR, 1) hist_rows_class_wide_kernel(const HistClassArgs a) {
    constexpr int NTHR = C_NTHR;
    extern __shared__ __align__(1024) unsigned char smem[];
    const int n_chunks = a.n_chunks[0];
    const int n_items = n_chunks * a.n_tiles;
    const int nw = a.stats_size;
    ClassOps* ops_s = reinterpret_cast<ClassOps*>(smem);
    CStage* st = reinterpret_cast<CStage*>(smem + (size_t)C_NST * C_STAGE * sizeof(ClassOps));
    int* s_item = reinterpret_cast<int*>(smem + (size_t)C_NST * C_STAGE * sizeof(ClassOps) + 128);
    uint32_t* hist = reinterpret_cast<uint32_t*>(smem + C_SMEM_FIXED);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int bins_pad = a.bins_pad;
    constexpr uint32_t WORD = 32u;  // bytes between the class words of one bin
    const uint32_t bin_stride = (uint32_t)nw * WORD;
    const uint32_t slot_stride = (uint32_t)bins_pad * bin_stride;
    // words / slot_words as a multiply: exact while words * slot_words < 2^32, and words < 2^18
    const uint32_t slot_magic = (uint32_t)((((uint64_t)1 << 32) + a.slot_words - 1) / a.slot_words);

    for (int i = threadIdx.x; i < C_NST * C_STAGE * 2; i += NTHR) reinterpret_cast<uint32_t*>(ops_s)[i] = 0u;
    if (threadIdx.x == 0) {
        for (int s = 0; s < C_NST; ++s) {
            mbar_init(&st[s].full, 1);
            mbar_init(&st[s].empty, C_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    uint32_t it_base = 0;
    for (;;) {
        __syncthreads();
        if (threadIdx.x == 0) *s_item = atomicAdd(&a.n_chunks[a.counter], 1);
        __syncthreads();
        const int item = *s_item;
        if (item >= n_items) break;
        const int tile_id = item / n_chunks;
        const int ci = item - tile_id * n_chunks;
        const int group = a.chunk_group[ci];
        const int64_t begin = a.chunk_begin[ci], end = a.chunk_end[ci];
        if (end <= begin) continue;
        const int g0 = group * a.gs;
        const int gcount = min(a.gs, a.n_slots - g0);
        const int n_iters = (int)((end - begin + CW_TILE - 1) / CW_TILE);
        // one spare bin behind the histogram takes what must not count (a row past the chunk, a bin out of range)
        for (int i = threadIdx.x; i < (gcount * bins_pad + 1) * nw * 8; i += NTHR) hist[i] = 0u;
        __syncthreads();

        if (warp == 0) {
            if (lane == 0) {
                for (int it = 0; it < n_iters; ++it) {
                    const uint32_t g = it_base + (uint32_t)it;
                    const int s = (int)(g % C_NST);
                    if (g >= (uint32_t)C_NST) mbar_wait(&st[s].empty, ((g / C_NST) - 1u) & 1u);
                    const int64_t r0 = begin + (int64_t)it * CW_TILE;
                    const int64_t e0 = r0 & ~(int64_t)1;
                    const int64_t e1 = (min(r0 + CW_TILE, end) + 1) & ~(int64_t)1;
                    const uint32_t bytes = (uint32_t)((e1 - e0) * (int64_t)sizeof(ClassOps));
                    mbar_expect_tx(&st[s].full, bytes);
                    tma_load_1d(ops_s + s * C_STAGE, a.ops + e0, bytes, &st[s].full);
                }
            }
        } else {
            const int cw = warp - 1;
            const int q = lane >> 3, j = lane & 7;
            const bool feat_ok = a.tile_feats[tile_id * 8 + j] >= 0;
            const uint32_t lane_base = smem_u32(hist) + (uint32_t)j * 4u;
            const uint32_t dummy_addr = lane_base + (uint32_t)gcount * slot_stride;
            const uint16_t* my_bins = reinterpret_cast<const uint16_t*>(a.rm_bins) + a.tile_slot0[tile_id] + j;
            const int wr0 = cw * CW_RPW;
            const int skew = (int)(begin & 1);
            uint32_t bw_cur[2], bw_next[2];
            auto fetch_bins = [&](int it, uint32_t (&bw)[2]) {
                const uint32_t g = it_base + (uint32_t)it;
                const int s = (int)(g % C_NST);
                mbar_wait(&st[s].full, (g / C_NST) & 1u);
                const int rows = (int)min((int64_t)CW_TILE, end - (begin + (int64_t)it * CW_TILE));
#pragma unroll
                for (int hs = 0; hs < 2; ++hs) {
                    const int r = wr0 + hs * 4 + q;
                    bw[hs] = 0;
                    if (r < rows) bw[hs] = __ldg(my_bins + (size_t)ops_s[s * C_STAGE + skew + r].row * a.fpad);
                }
            };
            // two register buffers, used alternately (the loop is unrolled by two): moving a buffer whose loads are still
            // in flight would wait for them and undo the prefetch
            auto accumulate = [&](int it, const uint32_t (&bw)[2]) {
                const uint32_t g = it_base + (uint32_t)it;
                const int s = (int)(g % C_NST);
                const int rows = (int)min((int64_t)CW_TILE, end - (begin + (int64_t)it * CW_TILE));
#pragma unroll
                for (int hs = 0; hs < 2; ++hs) {
                    const int r = wr0 + hs * 4 + q;
                    const uint32_t where = ops_s[s * C_STAGE + skew + min(r, CW_TILE - 1)].where;
                    const uint32_t words = (where & ((1u << CLASSOPS_OFF_BITS) - 1u)) >> a.word_shift;  // slot * slot_words + class
                    const uint32_t slot = a.slot_words > 1u ? __umulhi(words, slot_magic) : words, cls = words - slot * a.slot_words;
                    const bool ok = r < rows && feat_ok && bw[hs] < (uint32_t)bins_pad;
                    const uint32_t adr = ok ? lane_base + slot * slot_stride + bw[hs] * bin_stride + cls * WORD : dummy_addr;
                    sred_add<0>(adr, where >> CLASSOPS_OFF_BITS);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&st[s].empty);
            };
            if (n_iters > 0) fetch_bins(0, bw_cur);
#pragma unroll 1
            for (int it = 0; it < n_iters; it += 2) {
                if (it + 1 < n_iters) fetch_bins(it + 1, bw_next);
                accumulate(it, bw_cur);
                if (it + 1 >= n_iters) break;
                if (it + 2 < n_iters) fetch_bins(it + 2, bw_cur);
                accumulate(it + 1, bw_next);
            }
        }
        it_base += (uint32_t)n_iters;
        __syncthreads();
        flush_hist<STAT_CLASS, 1, 8>(hist, gcount, g0, bins_pad, nw, a.stats_size, *a.cnt_shift, a.tile_feats + tile_id * 8, a.pair_off,
                                     a.n_features, a.raw_lo, a.raw_hi, NTHR);
    }
}

size_t hist_rows_class_wide_smem_bytes(int gs, int bins_pad, int n_classes) {
    return C_SMEM_FIXED + ((size_t)gs * bins_pad + 1) * n_classes * 8 * 4;
}
int hist_rows_class_wide_max_slots_per_block(int bins_pad, int n_classes) {
    const size_t budget = 227 * 1024 - C_SMEM_FIXED - (size_t)n_classes * 8 * 4;
    const int gs = (int)(budget / ((size_t)bins_pad * n_classes * 8 * 4));
    return gs > 4096 ? 4096 : gs;
}
cudaError_t hist_rows_class_wide_launch(const HistClassArgs& a, int grid, size_t smem, cudaStream_t stream) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(hist_rows_class_wide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    hist_rows_class_wide_kernel<<<grid, C_NTHR, smem, stream>>>(a);
    return cudaGetLastError();
}

size_t hist_rows_class_smem_bytes(int gs, int bins_pad, int n_classes, int fpl) {
    return C_SMEM_FIXED + (size_t)gs * bins_pad * n_classes * 32 * fpl * 4;
}
int hist_rows_class_max_slots_per_block(int bins_pad, int n_classes, int fpl) {
    const size_t budget = 227 * 1024 - C_SMEM_FIXED;
    const int gs = (int)(budget / ((size_t)bins_pad * n_classes * 32 * fpl * 4));
    return gs > 4096 ? 4096 : gs;
}

template <int FPL>
static cudaError_t launch_class_one(const HistClassArgs& a, int grid, size_t smem, cudaStream_t stream) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(hist_rows_class_kernel<FPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    hist_rows_class_kernel<FPL><<<grid, C_NTHR, smem, stream>>>(a);
    return cudaGetLastError();
}
cudaError_t hist_rows_class_launch(int fpl, const HistClassArgs& a, int grid, size_t smem, cudaStream_t stream) {
    if (fpl == 4) return launch_class_one<4>(a, grid, smem, stream);
    if (fpl == 2) return launch_class_one<2>(a, grid, smem, stream);
    return launch_class_one<1>(a, grid, smem, stream);
}


// =====================================================================================================================
// hist_rows_wide_kernel: features with more than 256 bins (categorical columns, uint16 bin indices).  32 such
// features cannot have private histograms in one SM (1001 bins x 5 words x 32 lanes = 640 KB), so a tile is 8 wide
// features and a warp instruction covers 4 rows x 8 features: lane = (row slot q = lane >> 3, feature j = lane & 7),
// histogram [slot][bin][word][8].  Two lanes can now meet on one address (two of the 4 rows in the same bin of the
// same feature); the shared-memory atomic unit serialises exactly those.  Same row lists, same RowOps, same
// two-phase carry as hist_rows_kernel.
namespace {
constexpr size_t W_SMEM_FIXED = R_SMEM_FIXED;
constexpr int W_NTHREADS = (1 + 16) * 32;
}  // namespace

template <int KIND, int CNT_INT>
__global__ void __launch_bounds__(W_NTHREADS, 1) hist_rows_wide_kernel(const HistRowsArgs a) {
    constexpr int CONS_WARPS = 16;
    constexpr int RPW = R_TILE / CONS_WARPS;  // 8 rows of a stage per warp = 2 steps of 4 rows
    extern __shared__ __align__(1024) unsigned char smem[];
    const int ci = blockIdx.x / a.n_tiles;
    const int tile_id = blockIdx.x - ci * a.n_tiles;
    if (ci >= *a.n_chunks) return;
    const int nw = hist_words(KIND, CNT_INT, a.stats_size);
    RowOps* ops_s = reinterpret_cast<RowOps*>(smem);
    RStage* st = reinterpret_cast<RStage*>(smem + (size_t)R_NST * R_TILE * sizeof(RowOps));
    uint32_t* fm_tab = reinterpret_cast<uint32_t*>(smem + (size_t)R_NST * R_TILE * sizeof(RowOps) + 128);
    uint32_t* hist = reinterpret_cast<uint32_t*>(smem + W_SMEM_FIXED);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int group = a.chunk_group[ci];
    const int64_t begin = a.chunk_begin[ci], end = a.chunk_end[ci];
    const int g0 = group * a.gs;
    const int gcount = min(a.gs, a.n_slots - g0);
    const int bins_pad = a.bins_pad;
    const int n_iters = (int)((end - begin + R_TILE - 1) / R_TILE);

    for (int i = threadIdx.x; i < (gcount * bins_pad + 1) * nw * 8; i += W_NTHREADS) hist[i] = 0u;
    if (threadIdx.x < R_MAX_GS)
        fm_tab[threadIdx.x] = threadIdx.x < gcount ? a.slot_fmask[(size_t)(g0 + threadIdx.x) * a.n_tiles + tile_id] : 0u;
    if (threadIdx.x == 0) {
        for (int s = 0; s < R_NST; ++s) {
            mbar_init(&st[s].full, 1);
            mbar_init(&st[s].empty, CONS_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == 0) {
        if (lane == 0) {
            for (int it = 0; it < n_iters; ++it) {
                const int s = it % R_NST;
                if (it >= R_NST) mbar_wait(&st[s].empty, (uint32_t)(((it / R_NST) - 1) & 1));
                const int64_t r0 = begin + (int64_t)it * R_TILE;
                const uint32_t bytes = (uint32_t)(min((int64_t)R_TILE, end - r0) * (int64_t)sizeof(RowOps));
                mbar_expect_tx(&st[s].full, bytes);
                tma_load_1d(ops_s + s * R_TILE, a.ops + r0, bytes, &st[s].full);
            }
        }
    } else {
        const int cw = warp - 1;
        const int q = lane >> 3, j = lane & 7;
        const uint32_t feat_bit = 1u << j;
        const bool feat_ok = a.tile_feats[tile_id * 8 + j] >= 0;
        const uint32_t lane_base = smem_u32(hist) + (uint32_t)j * 4u;
        constexpr uint32_t WORD = 32u;                                  // bytes between the words of one bin
        const uint32_t bin_stride = (uint32_t)nw * WORD;
        const uint32_t slot_stride = (uint32_t)bins_pad * bin_stride;
        const uint32_t dummy_addr = lane_base + (uint32_t)gcount * slot_stride;
        const uint16_t* my_bins = reinterpret_cast<const uint16_t*>(a.rm_bins) + a.tile_slot0[tile_id] + j;
        const int wr0 = cw * RPW;
        uint32_t bw_cur[2], bw_next[2];
        auto fetch_bins = [&](int it, uint32_t (&bw)[2]) {
            const int s = it % R_NST;
            mbar_wait(&st[s].full, (uint32_t)((it / R_NST) & 1));
            const int rows = (int)min((int64_t)R_TILE, end - (begin + (int64_t)it * R_TILE));
#pragma unroll
            for (int hs = 0; hs < 2; ++hs) {
                const int r = wr0 + hs * 4 + q;
                bw[hs] = 0;
                if (r < rows) bw[hs] = __ldg(my_bins + (size_t)ops_s[s * R_TILE + r].row * a.fpad);
            }
        };
        // two register buffers, used alternately (the loop is unrolled by two): moving a buffer whose loads are still in
        // flight would wait for them and undo the prefetch
        auto accumulate = [&](int it, const uint32_t (&bw)[2]) {
            const int s = it % R_NST;
            const int rows = (int)min((int64_t)R_TILE, end - (begin + (int64_t)it * R_TILE));
            if (wr0 < rows) {
                uint32_t ad[2], o0[2], o1[2], o2[2], lo0[2], lo1[2], lo2[2], hi0[2], hi1[2], hi2[2];
#pragma unroll
                for (int hs = 0; hs < 2; ++hs) {
                    const int r = min(wr0 + hs * 4 + q, R_TILE - 1);
                    const RowOps* v = ops_s + s * R_TILE + r;
                    uint4 v0 = *reinterpret_cast<const uint4*>(v);                // row, cnt_hi, slot, cnt_lo
                    uint4 v1 = *reinterpret_cast<const uint4*>(&v->q[0]);         // sum / sq words
                    if (KIND == STAT_VARIANCE && CNT_INT) {  // packed high words (rowops_pack): same lists as the narrow tiles
                        v0.w = rowops_upper(v1.y); v0.y = 0u; v0.z = rowops_upper(v1.w);
                        if (v0.z == 0xffffu) v0.z = 0xffffffffu;
                        v1.y = rowops_hi_signed(v1.y); v1.w = rowops_hi_unsigned(v1.w);
                    }
                    const uint32_t slot = v0.z;
                    const bool ok = (wr0 + hs * 4 + q < rows) && slot < (uint32_t)gcount && feat_ok &&
                                    (fm_tab[slot & (R_MAX_GS - 1)] & feat_bit) != 0u && bw[hs] < (uint32_t)bins_pad;
                    uint32_t base = lane_base + slot * slot_stride + bw[hs] * bin_stride;
                    lo0[hs] = v0.w; hi0[hs] = v0.y;
                    lo1[hs] = hi1[hs] = lo2[hs] = hi2[hs] = 0;
                    if (KIND == STAT_VARIANCE) { lo1[hs] = v1.x; hi1[hs] = v1.y; lo2[hs] = v1.z; hi2[hs] = v1.w; }
                    else base += v1.x / 4u;  // class byte offset was computed for 128-byte words; here words are 32 bytes
                    ad[hs] = ok ? base : dummy_addr;
                    o0[hs] = o1[hs] = o2[hs] = 0;
                    if (KIND == STAT_VARIANCE) {
                        if (CNT_INT) {
                            sred_add<0>(ad[hs], lo0[hs]);
                            o1[hs] = satom_add<WORD>(ad[hs], lo1[hs]);
                            o2[hs] = satom_add<3 * WORD>(ad[hs], lo2[hs]);
                        } else {
                            o0[hs] = satom_add<0>(ad[hs], lo0[hs]);
                            o1[hs] = satom_add<2 * WORD>(ad[hs], lo1[hs]);
                            o2[hs] = satom_add<4 * WORD>(ad[hs], lo2[hs]);
                        }
                    } else {
                        if (CNT_INT) sred_add<0>(ad[hs], lo0[hs]);
                        else o0[hs] = satom_add<0>(ad[hs], lo0[hs]);
                    }
                }
#pragma unroll
                for (int hs = 0; hs < 2; ++hs) {
                    if (KIND == STAT_VARIANCE) {
                        if (CNT_INT) {
                            sred_add<2 * WORD>(ad[hs], hi_plus_carry(o1[hs], lo1[hs], hi1[hs]));
                            sred_add<4 * WORD>(ad[hs], hi_plus_carry(o2[hs], lo2[hs], hi2[hs]));
                        } else {
                            sred_add<WORD>(ad[hs], hi_plus_carry(o0[hs], lo0[hs], hi0[hs]));
                            sred_add<3 * WORD>(ad[hs], hi_plus_carry(o1[hs], lo1[hs], hi1[hs]));
                            sred_add<5 * WORD>(ad[hs], hi_plus_carry(o2[hs], lo2[hs], hi2[hs]));
                        }
                    } else if (!CNT_INT) {
                        sred_add<WORD>(ad[hs], hi_plus_carry(o0[hs], lo0[hs], hi0[hs]));
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&st[s].empty);
        };
        if (n_iters > 0) fetch_bins(0, bw_cur);
#pragma unroll 1
        for (int it = 0; it < n_iters; it += 2) {
            if (it + 1 < n_iters) fetch_bins(it + 1, bw_next);
            accumulate(it, bw_cur);
            if (it + 1 >= n_iters) break;
            if (it + 2 < n_iters) fetch_bins(it + 2, bw_cur);
            accumulate(it + 1, bw_next);
        }
    }
    __syncthreads();
    flush_hist<KIND, CNT_INT, 8>(hist, gcount, g0, bins_pad, nw, a.stats_size, CNT_INT ? *a.cnt_shift : 0,
                                 a.tile_feats + tile_id * 8, a.pair_off, a.n_features, a.raw_lo, a.raw_hi, W_NTHREADS);
}

size_t hist_rows_wide_smem_bytes(int gs, int bins_pad, int nw) {
    return W_SMEM_FIXED + ((size_t)gs * bins_pad + 1) * nw * 8 * 4;
}
int hist_rows_wide_max_slots_per_block(int bins_pad, int nw) {
    const size_t budget = 227 * 1024 - W_SMEM_FIXED - (size_t)nw * 8 * 4;
    int gs = (int)(budget / ((size_t)bins_pad * nw * 8 * 4));
    return gs > R_MAX_GS ? R_MAX_GS : gs;
}

template <int KIND, int CNT_INT>
static cudaError_t launch_wide_one(const HistRowsArgs& a, int grid, size_t smem, cudaStream_t stream) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(hist_rows_wide_kernel<KIND, CNT_INT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    hist_rows_wide_kernel<KIND, CNT_INT><<<grid, W_NTHREADS, smem, stream>>>(a);
    return cudaGetLastError();
}

cudaError_t hist_rows_wide_launch(int kind, int cnt_int, const HistRowsArgs& a, int grid, size_t smem, cudaStream_t stream) {
    if (kind == STAT_VARIANCE) return cnt_int ? launch_wide_one<STAT_VARIANCE, 1>(a, grid, smem, stream) : launch_wide_one<STAT_VARIANCE, 0>(a, grid, smem, stream);
    return cnt_int ? launch_wide_one<STAT_CLASS, 1>(a, grid, smem, stream) : launch_wide_one<STAT_CLASS, 0>(a, grid, smem, stream);
}

}  // namespace sb
