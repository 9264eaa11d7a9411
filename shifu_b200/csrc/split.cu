// split.cu -- (1) integer accumulators -> the reference's fp64 statistics; (2) on-device best split.
//
// (2) replaces the per-node body of DTMaster.doCompute (dt/DTMaster.java:316-340): Impurity.computeImpurity per
// feature (dt/Impurity.java:120-211 Variance, :269-360 FriedmanMSE, :378-477 Entropy, :563-661 Gini) and
// GainInfo.getGainInfoByMaxGain (dt/GainInfo.java:200-214).  One thread scans one (node, feature) block in the
// reference's own order (prefix sums bin by bin, strict '>' from Double.MIN_VALUE, first maximum wins), so given
// the same statistics it produces the same doubles as the Java code.  Built with -fmad=false.
#include <algorithm>

#include "common.cuh"
#include "kernels.cuh"

namespace sb {

namespace {

// value = hi * 2^32 + lo as a 128-bit integer, converted to double with ONE rounding (nearest-even).
__device__ __forceinline__ double int96_to_double(long long lo, long long hi) {
    __int128 v = ((__int128)hi << 32) + (__int128)lo;
    if (v == 0) return 0.0;
    const bool neg = v < 0;
    unsigned __int128 m = neg ? (unsigned __int128)(-v) : (unsigned __int128)v;
    const unsigned long long top = (unsigned long long)(m >> 64);
    int shift = 0;
    unsigned long long w;
    if (top != 0) {
        shift = 64 - __clzll((long long)top);  // bits above 64
        const unsigned long long lost = (unsigned long long)(m & (((unsigned __int128)1 << shift) - 1));
        w = (unsigned long long)(m >> shift) | (lost != 0 ? 1ULL : 0ULL);  // sticky bit keeps the rounding exact
    } else {
        w = (unsigned long long)m;
    }
    double d = __ull2double_rn(w);
    if (shift) d = ldexp(d, shift);
    return neg ? -d : d;
}

__global__ void finalize_stats_kernel(const int64_t* __restrict__ raw_lo, const int64_t* __restrict__ raw_hi,
                                      int64_t total, int stats_size, int kind, const double* __restrict__ inv_scales,
                                      double* __restrict__ stats) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int st = (kind == STAT_VARIANCE) ? (int)(i % 3) : 0;
        stats[i] = int96_to_double(raw_lo[i], raw_hi[i]) * inv_scales[st];
    }
    (void)stats_size;
}

__global__ void finalize_tasks_kernel(const int64_t* __restrict__ raw_lo, const int64_t* __restrict__ raw_hi,
                                      const SplitTask* __restrict__ tasks, int n_tasks, int64_t node_stride,
                                      const int64_t* __restrict__ feat_off, int stats_size, int kind,
                                      const double* __restrict__ inv_scales, double* __restrict__ stats) {
    const int t = (int)((blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (t >= n_tasks) return;
    const SplitTask tk = tasks[t];
    const int64_t src = (int64_t)tk.node * node_stride + feat_off[tk.feature];
    const int n = tk.n_bins * stats_size;
    for (int i = lane; i < n; i += 32) {
        const int st = (kind == STAT_VARIANCE) ? i % 3 : 0;
        stats[tk.offset + i] = int96_to_double(raw_lo[src + i], raw_hi[src + i]) * inv_scales[st];
    }
}

constexpr double JAVA_DOUBLE_MIN_VALUE = 4.9406564584124654e-324;

// Double.compareTo ordering: numeric, -0.0 < +0.0, NaN greatest.
__device__ __forceinline__ bool java_double_less(double a, double b) {
    if (a < b) return true;
    if (a > b) return false;
    if (a == b) {
        if (a == 0.0) return signbit(a) && !signbit(b);
        return false;
    }
    return !isnan(a) && isnan(b);
}

__device__ __forceinline__ double var_impurity(double count, double sum, double sum_square) {  // dt/Impurity.java:233-235
    return (count != 0.0) ? ((sum_square - (sum * sum) / count) / count) : 0.0;
}

struct ClsInfo { double sum_all; int largest; double impurity; };
constexpr int MAX_CLASSES = 16;

// Entropy.getEntropyInterInfo (dt/Impurity.java:499-522) / Gini.getGiniInfo (:683-704)
__device__ ClsInfo cls_info(const double* s, int c, bool entropy) {
    double sum_all = 0;
    for (int i = 0; i < c; ++i) sum_all += s[i];
    double impurity = 0.0;
    int idx = -1;
    if (sum_all != 0.0) {
        double max_element = JAVA_DOUBLE_MIN_VALUE;
        for (int i = 0; i < c; ++i) {
            const double rate = s[i] / sum_all;
            if (entropy) {
                if (rate != 0.0) {
                    impurity -= rate * (log(rate) / log(2.0));
                    if (s[i] > max_element) { max_element = s[i]; idx = i; }
                }
            } else {
                impurity -= rate * rate;
                if (s[i] > max_element) { max_element = s[i]; idx = i; }
            }
        }
    }
    return ClsInfo{sum_all, idx, impurity};
}

// Stable bottom-up merge sort of bin indices by key (Collections.sort is a stable merge sort; any stable sort gives the
// same permutation under Double.compareTo's total order), by the whole warp: in every round each element finds its
// place in the merged run with one binary search in the opposite run -- an element of the left run goes behind the
// elements of the right run that are strictly smaller, an element of the right run behind those of the left run that
// are smaller or equal (ties keep the left run first).  n log^2 n / 32 search steps per lane instead of n log n
// merge steps on one lane: a forest level with thousands of (node, 1001-category feature) tasks spent a quarter of its
// time here.
__device__ void warp_stable_sort_by_key(const double* key, int* idx, int* tmp, int n, int lane) {
    for (int i = lane; i < n; i += 32) idx[i] = i;
    __syncwarp();
    int* src = idx;
    int* dst = tmp;
    for (int width = 1; width < n; width *= 2) {
        for (int i = lane; i < n; i += 32) {
            const int lo = (i / (2 * width)) * (2 * width);
            const int mid = min(lo + width, n), hi = min(lo + 2 * width, n);
            const int me = src[i];
            const double k = key[me];
            int pos;
            if (i < mid) {
                int l = mid, r = hi;
                while (l < r) {
                    const int m = (l + r) >> 1;
                    if (java_double_less(key[src[m]], k)) l = m + 1;
                    else r = m;
                }
                pos = i + (l - mid);
            } else {
                int l = lo, r = mid;
                while (l < r) {
                    const int m = (l + r) >> 1;
                    if (!java_double_less(k, key[src[m]])) l = m + 1;
                    else r = m;
                }
                pos = l + (i - mid);
            }
            dst[pos] = me;
        }
        __syncwarp();
        int* t = src; src = dst; dst = t;
    }
    if (src != idx) {
        for (int i = lane; i < n; i += 32) idx[i] = src[i];
        __syncwarp();
    }
}

struct Cand {
    bool found;
    int split_i;       // loop index i of the winner
    double gain, impurity, predict, li, ri, lp, rp, wgt;
    int node_class, left_class, right_class;
};

// One (node, feature) block, scanned by one warp.  The running left-hand sums are a chain of double additions whose
// order fixes the result bits, so every lane walks the whole chain (the same additions in the same order -> the same
// bits); what is spread over the lanes is the expensive part, the divisions of the candidate evaluation: lane l keeps
// the prefix of candidates l, l + 32, ... and evaluates those.  Each lane keeps its first strictly-greater maximum; the
// warp then takes the largest gain, smallest candidate index among equals -- the candidate the sequential loop's
// "gain > maxGain" test selects (dt/Impurity.java:182-196).
// order == nullptr for numeric features.  pass_flags (categorical only) records which loop iterations produced a
// candidate: only those add their category to leftCategories (the `continue`s at dt/Impurity.java:157,166 skip the
// add), which is needed to rebuild the winner's bitset.
__device__ Cand warp_scan_feature(const double* __restrict__ s, int bins, int impurity_kind, int c, int min_inst,
                                  double min_gain, const int* order, uint8_t* pass_flags, int lane) {
    Cand best;
    best.found = false;
    best.split_i = 0x7fffffff;
    best.gain = 0.0;
    double max_gain = JAVA_DOUBLE_MIN_VALUE;
    if (impurity_kind == SB_IMPURITY_VARIANCE || impurity_kind == SB_IMPURITY_FRIEDMANMSE) {
        double count = 0, sum = 0, sum_sq = 0;
        for (int i = 0; i < bins; ++i) { count += s[i * 3]; sum += s[i * 3 + 1]; sum_sq += s[i * 3 + 2]; }
        const double impurity = var_impurity(count, sum, sum_sq);
        const double predict = count == 0.0 ? 0.0 : sum / count;
        double lc = 0, ls = 0, lq = 0;
        for (int base = 0; base < bins - 1; base += 32) {
            const int end = min(base + 32, bins - 1);
            double mc = 0, ms = 0, mq = 0;
            for (int i = base; i < end; ++i) {
                const int index = order ? order[i] : i;
                lc += s[index * 3]; ls += s[index * 3 + 1]; lq += s[index * 3 + 2];
                if (i - base == lane) { mc = lc; ms = ls; mq = lq; }
            }
            const int i = base + lane;
            if (i >= end) continue;
            const double rc = count - mc, rs = sum - ms, rq = sum_sq - mq;
            if (pass_flags) pass_flags[i] = 0;
            if (mc <= min_inst || rc <= min_inst) continue;
            const double lw = mc / count, rw = rc / count;
            const double li = var_impurity(mc, ms, mq), ri = var_impurity(rc, rs, rq);
            // "friedmanmse" is Variance on the master (dt/DTMaster.java:1026-1032)
            const double gain = impurity - lw * li - rw * ri;
            if (gain <= min_gain) continue;
            if (pass_flags) pass_flags[i] = 1;
            if (gain > max_gain) {
                max_gain = gain;
                best.found = true; best.split_i = i; best.gain = gain; best.impurity = impurity; best.predict = predict;
                best.li = li; best.ri = ri;
                best.lp = mc == 0.0 ? 0.0 : ms / mc;
                best.rp = rc == 0.0 ? 0.0 : rs / rc;
                best.wgt = count;
                best.node_class = best.left_class = best.right_class = 0;
            }
        }
    } else {
        const bool entropy = impurity_kind == SB_IMPURITY_ENTROPY;
        double tot[MAX_CLASSES], l[MAX_CLASSES], m[MAX_CLASSES], r[MAX_CLASSES];
        for (int j = 0; j < c; ++j) { tot[j] = 0; l[j] = 0; m[j] = 0; r[j] = 0; }
        for (int i = 0; i < bins; ++i)
            for (int j = 0; j < c; ++j) tot[j] += s[i * c + j];
        const ClsInfo info = cls_info(tot, c, entropy);
        const double predict = info.sum_all == 0.0 ? 0.0 : (tot[1] / info.sum_all);
        for (int base = 0; base < bins - 1; base += 32) {
            const int end = min(base + 32, bins - 1);
            for (int i = base; i < end; ++i) {
                const int index = order ? order[i] : i;
                for (int j = 0; j < c; ++j) l[j] += s[index * c + j];
                if (i - base == lane)
                    for (int j = 0; j < c; ++j) m[j] = l[j];
            }
            const int i = base + lane;
            if (i >= end) continue;
            const ClsInfo li = cls_info(m, c, entropy);
            const double lp = li.sum_all == 0.0 ? 0.0 : (m[1] / li.sum_all);
            for (int j = 0; j < c; ++j) r[j] = tot[j] - m[j];
            const ClsInfo ri = cls_info(r, c, entropy);
            if (pass_flags) pass_flags[i] = 0;
            if (li.sum_all <= min_inst || ri.sum_all <= min_inst) continue;
            const double rp = ri.sum_all == 0.0 ? 0.0 : (r[1] / ri.sum_all);
            const double lw = info.sum_all == 0.0 ? 0.0 : (li.sum_all / info.sum_all);
            const double rw = info.sum_all == 0.0 ? 0.0 : (ri.sum_all / info.sum_all);
            const double gain = info.impurity - lw * li.impurity - rw * ri.impurity;
            if (gain <= min_gain) continue;
            if (pass_flags) pass_flags[i] = 1;
            if (gain > max_gain) {
                max_gain = gain;
                best.found = true; best.split_i = i; best.gain = gain; best.impurity = info.impurity;
                best.predict = predict; best.li = li.impurity; best.ri = ri.impurity; best.lp = lp; best.rp = rp;
                best.wgt = info.sum_all;
                best.node_class = (int)(int8_t)info.largest;
                best.left_class = (int)(int8_t)li.largest;
                best.right_class = (int)(int8_t)ri.largest;
            }
        }
    }
    return best;
}

constexpr int SPLIT_WARPS = 4;       // tasks per block
constexpr int SPLIT_STAGE = 1024;    // doubles of one task's statistics staged in shared memory (64 bins x 16 classes)
constexpr int RANK_SORT_MAX = 128;   // categories up to which the warp sorts by counting ranks

__global__ void __launch_bounds__(SPLIT_WARPS * 32)
feature_split_kernel(const double* __restrict__ stats, const SplitTask* __restrict__ tasks, int n_tasks,
                     int impurity_kind, int c, int min_inst, double min_gain, sb_gain* cand, uint8_t* cand_bits,
                     double* sort_keys, int* sort_idx, int* sort_tmp, uint8_t* pass_flags, int max_bins, int* next_slot) {
    __shared__ double stage[SPLIT_WARPS][SPLIT_STAGE];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t = blockIdx.x * SPLIT_WARPS + warp;
    if (t >= n_tasks) return;  // warps are independent: no block-wide barrier below
    const SplitTask tk = tasks[t];
    const double* s = stats + tk.offset;
    const bool variance = impurity_kind == SB_IMPURITY_VARIANCE || impurity_kind == SB_IMPURITY_FRIEDMANMSE;
    const int ss = variance ? 3 : c;
    if (tk.n_bins * ss <= SPLIT_STAGE) {
        for (int i = lane; i < tk.n_bins * ss; i += 32) stage[warp][i] = s[i];
        __syncwarp();
        s = stage[warp];
    }
    int* order = nullptr;
    uint8_t* flags = nullptr;
    if (tk.categorical) {
        // scratch exists for the categorical tasks only (a forest level has 10^5 tasks, a fifth of them categorical, with
        // up to 1001 entries each): the task claims the next piece; which one it gets does not matter
        int slot = 0;
        if (lane == 0) slot = atomicAdd(next_slot, 1);
        slot = __shfl_sync(0xffffffffu, slot, 0);
        double* key = sort_keys + (size_t)slot * max_bins;
        order = sort_idx + (size_t)slot * max_bins;
        int* tmp = sort_tmp + (size_t)slot * max_bins;
        flags = pass_flags + (size_t)slot * max_bins;
        for (int i = lane; i < tk.n_bins; i += 32) {
            double k;
            if (variance) {  // dt/Impurity.java:213-231: empty bins sort first with key Double.MIN_VALUE
                k = JAVA_DOUBLE_MIN_VALUE;
                if (s[i * 3] != 0.0) k = s[i * 3 + 1] / s[i * 3];
            } else {         // :479-497, :663-681: positive rate over classes 0 and 1 only
                const double sum = s[i * ss] + s[i * ss + 1];
                k = 0.0;
                if (sum != 0.0) k = s[i * ss + 1] / sum;
            }
            key[i] = k;
        }
        __syncwarp();
        if (tk.n_bins <= RANK_SORT_MAX) {
            // stable sort by counting: position of i = #{j : key[j] < key[i], or equal and j < i}
            for (int i = lane; i < tk.n_bins; i += 32) {
                const double ki = key[i];
                int rank = 0;
                for (int j = 0; j < tk.n_bins; ++j) {
                    const double kj = key[j];
                    if (java_double_less(kj, ki) || (j < i && !java_double_less(ki, kj))) ++rank;
                }
                order[rank] = i;
            }
        } else {
            warp_stable_sort_by_key(key, order, tmp, tk.n_bins, lane);
        }
        __syncwarp();
    }
    const Cand b = warp_scan_feature(s, tk.n_bins, impurity_kind, c, min_inst, min_gain, order, flags, lane);
    // winner over the lanes: largest gain, smallest loop index among equal gains
    int win = b.found ? lane : -1;
    double wg = b.gain;
    int wi = b.split_i;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const int ow = __shfl_xor_sync(0xffffffffu, win, o);
        const double og = __shfl_xor_sync(0xffffffffu, wg, o);
        const int oi = __shfl_xor_sync(0xffffffffu, wi, o);
        if (ow >= 0 && (win < 0 || og > wg || (og == wg && oi < wi))) { win = ow; wg = og; wi = oi; }
    }
    __syncwarp();  // pass_flags of every lane are visible to the winner
    if (win < 0) {
        if (lane == 0) {
            sb_gain g;
            memset(&g, 0, sizeof(g));
            cand[t] = g;
        }
        return;
    }
    if (lane != win) return;
    sb_gain g;
    memset(&g, 0, sizeof(g));
    g.found = 1;
    g.feature = tk.feature;
    g.split_type = tk.categorical ? SB_SPLIT_CATEGORICAL : SB_SPLIT_CONTINUOUS;
    g.split_bin = b.split_i;
    g.gain = b.gain; g.impurity = b.impurity; g.predict = b.predict;
    g.left_impurity = b.li; g.right_impurity = b.ri; g.left_predict = b.lp; g.right_predict = b.rp;
    g.wgt_cnt = b.wgt;
    g.node_class = b.node_class; g.left_class = b.left_class; g.right_class = b.right_class;
    if (tk.categorical) {
        // rebuild leftCategories as of the winning iteration (dt/Impurity.java:169-198)
        const int n_cats = tk.n_bins - 1;
        const int bytes = (n_cats + 1) / 8 + 1;
        uint8_t* bits = cand_bits + (size_t)t * SB_MAX_CAT_BYTES;
        for (int i = 0; i < bytes; ++i) bits[i] = 0;
        int left_size = 0;
        for (int i = 0; i <= b.split_i; ++i) {
            if (!flags[i]) continue;
            const int index = order[i];
            const int v = index >= n_cats ? n_cats : index;
            bits[v >> 3] |= (uint8_t)(1 << (v % 8));
            left_size += 1;
        }
        int is_left = 1;
        if (n_cats + 1 <= left_size * 2) {  // too many on the left: store the right side instead
            is_left = 0;
            // complement within [0, n_cats]
            for (int i = 0; i < bytes; ++i) bits[i] = (uint8_t)~bits[i];
            for (int v = n_cats + 1; v < bytes * 8; ++v) bits[v >> 3] &= (uint8_t)~(1 << (v % 8));
        }
        g.is_left = is_left;
        g.cat_bytes = bytes;
    }
    cand[t] = g;
}

// GainInfo.getGainInfoByMaxGain over the features of one node, in the reference's visit order: one warp per node.
// "First strictly greater wins" == the maximum gain; among equal maxima the smallest visit rank.
__global__ void node_best_kernel(const sb_gain* __restrict__ cand, const uint8_t* __restrict__ cand_bits,
                                 const SplitTask* __restrict__ tasks, const int32_t* __restrict__ node_task_begin,
                                 int n_nodes, sb_gain* out, uint8_t* out_bits) {
    const int n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (n >= n_nodes) return;
    const int t0 = node_task_begin[n], t1 = node_task_begin[n + 1];
    double max_gain = JAVA_DOUBLE_MIN_VALUE;
    int best = -1, best_rank = 0x7fffffff;
    for (int t = t0 + lane; t < t1; t += 32) {
        if (!cand[t].found) continue;
        const double g = cand[t].gain;
        const int rk = tasks[t].visit_rank;
        if (g > max_gain || (g == max_gain && best >= 0 && rk < best_rank)) {
            max_gain = g;
            best = t;
            best_rank = rk;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double og = __shfl_down_sync(0xffffffffu, max_gain, o);
        const int ob = __shfl_down_sync(0xffffffffu, best, o);
        const int ork = __shfl_down_sync(0xffffffffu, best_rank, o);
        if (ob >= 0 && (best < 0 || og > max_gain || (og == max_gain && ork < best_rank))) {
            max_gain = og;
            best = ob;
            best_rank = ork;
        }
    }
    best = __shfl_sync(0xffffffffu, best, 0);
    if (best < 0) {
        if (lane == 0) {
            sb_gain z;
            memset(&z, 0, sizeof(z));
            out[n] = z;
        }
        return;
    }
    if (lane == 0) out[n] = cand[best];
    if (out_bits && cand[best].cat_bytes > 0) {
        const uint8_t* src = cand_bits + (size_t)best * SB_MAX_CAT_BYTES;
        uint8_t* dst = out_bits + (size_t)n * SB_MAX_CAT_BYTES;
        for (int i = lane; i < cand[best].cat_bytes; i += 32) dst[i] = src[i];
    }
}

// ---- histogram subtraction (sibling = parent - child), exact in the integer accumulators
// One warp per sibling pair: which child holds less (weighted) count under the parent's split?  Only that child is
// accumulated directly; the other one is derived after the merge.  The decision is made from the parent's merged
// (all-reduced) accumulators, so every GPU takes the same one.  Integer sums: the lane order does not matter.
template <typename SlotT>
__global__ void pair_decide_kernel(const PairDesc* __restrict__ pairs, int n_pairs, const int64_t* __restrict__ prev_lo,
                                   const int64_t* __restrict__ prev_hi, int stats_size, int kind, int n_features,
                                   SlotT* __restrict__ slot_tables, int64_t* __restrict__ pair_off, int32_t* __restrict__ small_is_b) {
    const int p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (p >= n_pairs) return;
    const PairDesc& d = pairs[p];
    const int per_bin = kind == STAT_VARIANCE ? 1 : stats_size;  // the count is stat 0 (variance) or the sum over classes
    __int128 left = 0, total = 0;
    for (int b = lane; b < d.n_bins; b += 32) {
        __int128 c = 0;
        for (int k = 0; k < per_bin; ++k) {
            const int64_t i = d.parent_split_block + (int64_t)b * stats_size + k;
            c += ((__int128)prev_hi[i] << 32) + (__int128)prev_lo[i];
        }
        total += c;
        if ((d.left_mask[b >> 3] >> (b & 7)) & 1) left += c;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long l0 = __shfl_xor_sync(0xffffffffu, (unsigned long long)left, o);
        const long long l1 = __shfl_xor_sync(0xffffffffu, (long long)(left >> 64), o);
        const unsigned long long t0 = __shfl_xor_sync(0xffffffffu, (unsigned long long)total, o);
        const long long t1 = __shfl_xor_sync(0xffffffffu, (long long)(total >> 64), o);
        left += ((__int128)l1 << 64) + (__int128)l0;
        total += ((__int128)t1 << 64) + (__int128)t0;
    }
    const int b_small = (2 * left > total) ? 1 : 0;  // left = child A (even id); ties -> A
    if (lane == 0) {
        small_is_b[p] = b_small;
        SlotT* table = slot_tables + d.table_off;
        table[b_small ? d.node_a : d.node_b] = (SlotT)-1;  // the larger child is not accumulated
    }
    if (b_small) {
        int64_t* row = pair_off + (size_t)d.slot * n_features;
        const int64_t delta = d.off_b - d.off_a;
        for (int f = lane; f < n_features; f += 32)
            if (row[f] >= 0) row[f] += delta;
    }
}

__global__ void pair_fixup_kernel(const PairDesc* __restrict__ pairs, const int32_t* __restrict__ small_is_b,
                                  const int64_t* __restrict__ prev_lo, const int64_t* __restrict__ prev_hi,
                                  int64_t* __restrict__ lo, int64_t* __restrict__ hi) {
    const PairDesc d = pairs[blockIdx.y];
    const int64_t small_off = small_is_b[blockIdx.y] ? d.off_b : d.off_a;
    const int64_t large_off = small_is_b[blockIdx.y] ? d.off_a : d.off_b;
    for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < d.len; e += (int64_t)gridDim.x * blockDim.x) {
        lo[large_off + e] = prev_lo[d.parent_off + e] - lo[small_off + e];
        hi[large_off + e] = prev_hi[d.parent_off + e] - hi[small_off + e];
    }
}

}  // namespace

// All todo nodes share one feature list (the usual case): the (node, feature) task table is the first node's row
// repeated with the node's block offset added, so only one row and the node offsets cross PCIe.
__global__ void expand_tasks_kernel(const SplitTask* __restrict__ tmpl, int fk, const int64_t* __restrict__ node_off,
                                    int n_nodes, SplitTask* __restrict__ tasks, int32_t* __restrict__ begin) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i <= n_nodes) begin[i] = i * fk;
    if (i >= n_nodes * fk) return;
    const int n = i / fk, k = i - n * fk;
    SplitTask t = tmpl[k];
    t.node = n;
    t.offset += node_off[n];
    tasks[i] = t;
}
cudaError_t launch_expand_tasks(const SplitTask* tmpl, int fk, const int64_t* node_off, int n_nodes, SplitTask* tasks,
                                int32_t* begin, cudaStream_t s) {
    const int total = std::max(n_nodes * fk, n_nodes + 1);
    expand_tasks_kernel<<<(total + 255) / 256, 256, 0, s>>>(tmpl, fk, node_off, n_nodes, tasks, begin);
    return cudaGetLastError();
}

// Todo nodes with their OWN feature lists (RF, FeatureSubsetStrategy != ALL): the task table of every node is built on
// the device from the concatenated lists -- block offsets (a running sum of bins * statsSize in list order) and the
// java.util.HashMap visit rank of every feature of the node (dt/DTMaster.java:324: buckets ascending, insertion order
// inside a bucket; capacity = the power of two >= the list length, as `new HashMap(n, 1f)` sizes it).  One warp per node;
// the bucket counters live in shared memory.  The host used to build and copy 32 bytes per (node, feature).
__global__ void build_tasks_kernel(const int32_t* __restrict__ feats, const int32_t* __restrict__ begin,
                                   const int64_t* __restrict__ node_off, int n_nodes, const int32_t* __restrict__ feat_bins,
                                   const uint8_t* __restrict__ is_cat, const int32_t* __restrict__ column_num, int stats_size,
                                   SplitTask* __restrict__ tasks) {
    extern __shared__ int bucket_first[];
    const int n = blockIdx.x, lane = threadIdx.x;
    if (n >= n_nodes) return;
    const int b0 = begin[n], k = begin[n + 1] - b0;
    int cap = 1;
    while (cap < k) cap <<= 1;
    for (int i = lane; i < cap; i += 32) bucket_first[i] = 0;
    __syncwarp();
    auto bucket = [cap](int key) {
        uint32_t x = (uint32_t)key;
        x ^= x >> 16;
        return (int)(x & (uint32_t)(cap - 1));
    };
    for (int i = lane; i < k; i += 32) atomicAdd(&bucket_first[bucket(column_num[feats[b0 + i]])], 1);
    __syncwarp();
    int running = 0;  // exclusive scan of the bucket sizes: first visit position of every bucket
    for (int c0 = 0; c0 < cap; c0 += 32) {
        const int v = (c0 + lane < cap) ? bucket_first[c0 + lane] : 0;
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (c0 + lane < cap) bucket_first[c0 + lane] = running + incl - v;
        running += __shfl_sync(0xffffffffu, incl, 31);
    }
    __syncwarp();
    int64_t carry = 0;  // block offset of the next feature inside the node block
    const uint32_t lt = (1u << lane) - 1u;
    for (int c0 = 0; c0 < k; c0 += 32) {
        const int i = c0 + lane;
        const bool valid = i < k;
        const int f = valid ? feats[b0 + i] : 0;
        const int b = valid ? bucket(column_num[f]) : -1 - lane;
        const uint32_t same = __match_any_sync(0xffffffffu, b);
        int rank = 0;
        if (valid) rank = bucket_first[b] + __popc(same & lt);
        __syncwarp();
        if (valid && lane == __ffs(same) - 1) bucket_first[b] += __popc(same);
        __syncwarp();
        const int64_t size = valid ? (int64_t)feat_bins[f] * stats_size : 0;
        int64_t incl = size;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int64_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (valid) {
            SplitTask t;
            t.node = n;
            t.feature = f;
            t.offset = node_off[n] + carry + incl - size;
            t.n_bins = feat_bins[f];
            t.categorical = is_cat[f];
            t.visit_rank = rank;
            t.pad = 0;
            tasks[b0 + i] = t;
        }
        carry += __shfl_sync(0xffffffffu, incl, 31);
    }
}
cudaError_t launch_build_tasks(const int32_t* feats, const int32_t* begin, const int64_t* node_off, int n_nodes, int max_list,
                               const int32_t* feat_bins, const uint8_t* is_cat, const int32_t* column_num, int stats_size,
                               SplitTask* tasks, cudaStream_t s) {
    if (n_nodes <= 0) return cudaSuccess;
    int cap = 1;
    while (cap < max_list) cap <<= 1;
    build_tasks_kernel<<<n_nodes, 32, (size_t)cap * sizeof(int), s>>>(feats, begin, node_off, n_nodes, feat_bins, is_cat, column_num,
                                                                       stats_size, tasks);
    return cudaGetLastError();
}

// Count statistic (stat 0) of every bin of a raw block, saved to / restored from a cache of (lo, hi) pairs.
__global__ void counts_cache_kernel(int64_t* __restrict__ raw_lo, int64_t* __restrict__ raw_hi, int64_t* __restrict__ cache,
                                    int64_t n_bins_total, int stats_size, int save) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n_bins_total) return;
    const int64_t idx = i * stats_size;
    if (save) { cache[2 * i] = raw_lo[idx]; cache[2 * i + 1] = raw_hi[idx]; }
    else { raw_lo[idx] = cache[2 * i]; raw_hi[idx] = cache[2 * i + 1]; }
}
cudaError_t launch_counts_cache(int64_t* raw_lo, int64_t* raw_hi, int64_t* cache, int64_t n_bins_total, int stats_size, int save,
                                cudaStream_t s) {
    if (n_bins_total <= 0) return cudaSuccess;
    counts_cache_kernel<<<(unsigned)((n_bins_total + 255) / 256), 256, 0, s>>>(raw_lo, raw_hi, cache, n_bins_total, stats_size, save);
    return cudaGetLastError();
}

cudaError_t launch_pair_decide(const PairDesc* pairs, int n_pairs, const int64_t* prev_lo, const int64_t* prev_hi,
                               int stats_size, int kind, int n_features, int16_t* slot_tables, int64_t* pair_off,
                               int32_t* small_is_b, cudaStream_t s) {
    if (n_pairs <= 0) return cudaSuccess;
    pair_decide_kernel<int16_t><<<(n_pairs * 32 + 127) / 128, 128, 0, s>>>(pairs, n_pairs, prev_lo, prev_hi, stats_size, kind, n_features,
                                                                   slot_tables, pair_off, small_is_b);
    return cudaGetLastError();
}

cudaError_t launch_pair_decide32(const PairDesc* pairs, int n_pairs, const int64_t* prev_lo, const int64_t* prev_hi,
                                 int stats_size, int kind, int n_features, int32_t* slot_tables, int64_t* pair_off,
                                 int32_t* small_is_b, cudaStream_t s) {
    if (n_pairs <= 0) return cudaSuccess;
    pair_decide_kernel<int32_t><<<(n_pairs * 32 + 127) / 128, 128, 0, s>>>(pairs, n_pairs, prev_lo, prev_hi, stats_size, kind, n_features,
                                                                   slot_tables, pair_off, small_is_b);
    return cudaGetLastError();
}

cudaError_t launch_pair_fixup(const PairDesc* pairs, int n_pairs, int64_t max_len, const int32_t* small_is_b,
                              const int64_t* prev_lo, const int64_t* prev_hi, int64_t* lo, int64_t* hi, cudaStream_t s) {
    if (n_pairs <= 0 || max_len <= 0) return cudaSuccess;
    int64_t gx = (max_len + 255) / 256;
    if (gx > 64) gx = 64;
    dim3 grid((unsigned)gx, (unsigned)n_pairs);
    pair_fixup_kernel<<<grid, 256, 0, s>>>(pairs, small_is_b, prev_lo, prev_hi, lo, hi);
    return cudaGetLastError();
}

cudaError_t launch_finalize(const int64_t* raw_lo, const int64_t* raw_hi, int64_t total, int stats_size, int kind,
                            const double* inv_scales, double* stats, cudaStream_t s) {
    if (total <= 0) return cudaSuccess;
    int64_t g = (total + 255) / 256;
    if (g > 148 * 8) g = 148 * 8;
    finalize_stats_kernel<<<(int)g, 256, 0, s>>>(raw_lo, raw_hi, total, stats_size, kind, inv_scales, stats);
    return cudaGetLastError();
}

cudaError_t launch_finalize_tasks(const int64_t* raw_lo, const int64_t* raw_hi, const SplitTask* tasks, int n_tasks, int64_t node_stride,
                                  const int64_t* feat_off, int stats_size, int kind, const double* inv_scales, double* stats,
                                  cudaStream_t s) {
    if (n_tasks <= 0) return cudaSuccess;
    finalize_tasks_kernel<<<(unsigned)(((int64_t)n_tasks * 32 + 255) / 256), 256, 0, s>>>(raw_lo, raw_hi, tasks, n_tasks, node_stride, feat_off,
                                                                                          stats_size, kind, inv_scales, stats);
    return cudaGetLastError();
}

cudaError_t launch_best_splits(const double* stats, const SplitTask* tasks, int n_tasks, int n_nodes, int impurity,
                               int stats_size, int min_instances, double min_info_gain, sb_gain* cand,
                               uint8_t* cand_bits, sb_gain* out, uint8_t* out_bits, const int32_t* node_task_begin,
                               const SplitScratch& sc, cudaStream_t s) {
    if (n_tasks <= 0 || n_nodes <= 0) return cudaSuccess;
    feature_split_kernel<<<(n_tasks + SPLIT_WARPS - 1) / SPLIT_WARPS, SPLIT_WARPS * 32, 0, s>>>(stats, tasks, n_tasks, impurity, stats_size, min_instances,
                                                            min_info_gain, cand, cand_bits, sc.keys, sc.idx, sc.tmp,
                                                            sc.flags, sc.max_bins, sc.next);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    node_best_kernel<<<(n_nodes * 32 + 127) / 128, 128, 0, s>>>(cand, cand_bits, tasks, node_task_begin, n_nodes, out, out_bits);
    return cudaGetLastError();
}

}  // namespace sb
