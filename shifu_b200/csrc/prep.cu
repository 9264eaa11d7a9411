// prep.cu -- the two steps on either side of the statistics build that SURVEY.md 8(f) ranks next:
//   f1  binning at load: DTWorker.getBinIndex (dt/DTWorker.java:1001-1034) for a whole column on the device;
//   f4  bagging masks from java.util.Random on the device (dt/DTWorker.java:1521-1544 at load, :643-668 per tree),
//       with jump-ahead so that row j gets exactly the j-th nextDouble() of `new Random(seed)`.
#include <cub/device/device_scan.cuh>

#include "common.cuh"
#include "kernels.cuh"

namespace {

constexpr int MAX_BOUNDARIES = 4096;

// getBinIndex: NaN is mapped to 0f first, as DTWorker.getFloatValue does for unparsable input (:1299-1305).
// The boundaries of the column are staged in shared memory once per block (<= 32 KB): every thread then bisects
// over on-chip data, and two columns binned at the same time on different streams share nothing.
__global__ void bin_values_kernel(const float* __restrict__ values, int64_t n, const double* __restrict__ boundaries,
                                  int n_bins, void* out, int out_width) {
    extern __shared__ double c_boundaries[];
    for (int i = threadIdx.x; i < n_bins; i += blockDim.x) c_boundaries[i] = boundaries[i];
    __syncthreads();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float v = values[i];
        if (isnan(v)) v = 0.0f;
        int bin = -1;
        if (n_bins <= 1) bin = 0;
        else if (v == __int_as_float(0x7f800000)) bin = n_bins - 1;
        else if (v == __int_as_float(0xff800000)) bin = 0;
        else {
            int low = 0, high = n_bins - 1;
            const double dv = (double)v;
            while (low <= high) {
                const int mid = (int)(((unsigned)low + (unsigned)high) >> 1);
                const double lo_t = c_boundaries[mid];
                const double hi_t = mid == n_bins - 1 ? 1.7976931348623157e308 : c_boundaries[mid + 1];
                if (dv >= lo_t && dv < hi_t) { bin = mid; break; }
                if (dv >= hi_t) low = mid + 1;
                else high = mid - 1;
            }
        }
        // -1 (value below the first boundary) is stored as the all-ones index, which sb_load_rows rejects
        if (out_width == 2) static_cast<uint16_t*>(out)[i] = (uint16_t)bin;
        else static_cast<uint8_t*>(out)[i] = (uint8_t)bin;
    }
}

constexpr unsigned long long LCG_A = 0x5DEECE66DULL, LCG_C = 0xBULL, LCG_MASK = (1ULL << 48) - 1;

// state after k steps from s: s * A^k + C * (A^(k-1) + ... + 1)  (mod 2^48), by binary decomposition of k
__device__ __forceinline__ unsigned long long lcg_jump(unsigned long long s, unsigned long long k) {
    unsigned long long a_acc = 1, c_acc = 0;        // identity map
    unsigned long long a_pow = LCG_A, c_pow = LCG_C;  // the map "one step", squared as we go
    while (k) {
        if (k & 1) {
            a_acc = (a_acc * a_pow) & LCG_MASK;
            c_acc = (c_acc * a_pow + c_pow) & LCG_MASK;
        }
        c_pow = (c_pow * (a_pow + 1)) & LCG_MASK;
        a_pow = (a_pow * a_pow) & LCG_MASK;
        k >>= 1;
    }
    return (s * a_acc + c_acc) & LCG_MASK;
}

constexpr int ROWS_PER_THREAD = 32;

// weights[j] = (j-th nextDouble() of new Random(seed), counted from first_draw) <= rate ? 1f : 0f
__global__ void java_bagging_mask_kernel(float* __restrict__ weights, int64_t n, long long seed, double rate, long long first_draw) {
    const int64_t chunks = (n + ROWS_PER_THREAD - 1) / ROWS_PER_THREAD;
    for (int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; c < chunks; c += (int64_t)gridDim.x * blockDim.x) {
        const int64_t j0 = c * ROWS_PER_THREAD;
        unsigned long long s = ((unsigned long long)seed ^ LCG_A) & LCG_MASK;  // Random(seed) scrambles the seed
        s = lcg_jump(s, 2ULL * (unsigned long long)(first_draw + j0));           // two next() calls per nextDouble()
        const int64_t j1 = min(n, j0 + ROWS_PER_THREAD);
        for (int64_t j = j0; j < j1; ++j) {
            s = (s * LCG_A + LCG_C) & LCG_MASK;
            const long long hi = (long long)(s >> (48 - 26));
            s = (s * LCG_A + LCG_C) & LCG_MASK;
            const long long lo = (long long)(s >> (48 - 27));
            const double d = (double)((hi << 27) + lo) * (1.0 / 9007199254740992.0);
            weights[j] = d <= rate ? 1.0f : 0.0f;
        }
    }
}

// Stratified variant (dt/DTWorker.java:1526-1532, 646-653): one Random per class, all created from the same seed, so a row
// gets the k-th nextDouble() of `new Random(seed)` where k counts the earlier rows of ITS class.
// class of a row = (int)(label + 0.01f) (:1520).
__global__ void class_indicator_kernel(const float* __restrict__ label, int64_t n, int cls, int n_classes, uint32_t* __restrict__ ind,
                                       uint32_t* error_flag) {
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(label[r] + 0.01f);
        if (cls == 0 && (c < 0 || c >= n_classes)) atomicOr(error_flag, 1u);
        ind[r] = c == cls ? 1u : 0u;
    }
}
__global__ void stratified_mask_kernel(float* __restrict__ weights, const float* __restrict__ label, const uint32_t* __restrict__ ordinal,
                                       int64_t n, int cls, long long seed, double rate) {
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        if ((int)(label[r] + 0.01f) != cls) continue;
        unsigned long long s = ((unsigned long long)seed ^ LCG_A) & LCG_MASK;
        s = lcg_jump(s, 2ULL * (unsigned long long)ordinal[r]);
        s = (s * LCG_A + LCG_C) & LCG_MASK;
        const long long hi = (long long)(s >> (48 - 26));
        s = (s * LCG_A + LCG_C) & LCG_MASK;
        const long long lo = (long long)(s >> (48 - 27));
        const double d = (double)((hi << 27) + lo) * (1.0 / 9007199254740992.0);
        weights[r] = d <= rate ? 1.0f : 0.0f;
    }
}

// RF bag counts (dt/DTWorker.java:1547-1571): the reference draws PoissonDistribution(rate).sample() from an unseeded
// commons-math3 generator, so only the distribution can be matched.  Here the count of (row, tree) is the inverse CDF of
// a counter-based uniform, splitmix64(seed ^ (global row * M1 + tree * M2)): any shard of the rows can be drawn anywhere
// (shifu_b200/synthetic.py:poisson_bags is the host twin).  exp(-rate) comes from the host so both sides share it.
__device__ __forceinline__ unsigned long long splitmix64_dev(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}
__global__ void poisson_bags_kernel(float* __restrict__ weights, int64_t n, unsigned long long seed, long long row0, int tree,
                                    double rate, double p0) {
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const unsigned long long key = (unsigned long long)(row0 + r) * 0xD1342543DE82EF95ULL + (unsigned long long)tree * 0x9E3779B97F4A7C15ULL;
        const double u = (double)(splitmix64_dev(seed ^ key) >> 11) * (1.0 / 9007199254740992.0);
        int k = 0;
        double pk = p0, cdf = p0;
        while (u > cdf && k < 1000) {
            ++k;
            pk = pk * rate / (double)k;
            cdf += pk;
        }
        weights[r] = (float)k;
    }
}

}  // namespace

namespace sb {
size_t stratified_temp_bytes(int64_t n) {
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, n);
    return bytes;
}
// ind / ordinal: n uint32 each; temp: stratified_temp_bytes(n)
cudaError_t launch_stratified_mask(float* weights, const float* label, int64_t n, int n_classes, long long seed, double rate,
                                   uint32_t* ind, uint32_t* ordinal, void* temp, size_t temp_bytes, uint32_t* error_flag,
                                   cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    int64_t g = (n + 255) / 256;
    if (g > 148 * 16) g = 148 * 16;
    for (int c = 0; c < n_classes; ++c) {
        class_indicator_kernel<<<(int)g, 256, 0, s>>>(label, n, c, n_classes, ind, error_flag);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        size_t tb = temp_bytes;
        e = cub::DeviceScan::ExclusiveSum(temp, tb, (const uint32_t*)ind, ordinal, n, s);
        if (e != cudaSuccess) return e;
        stratified_mask_kernel<<<(int)g, 256, 0, s>>>(weights, label, ordinal, n, c, seed, rate);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    return cudaSuccess;
}
cudaError_t launch_poisson_bags(float* weights, int64_t n, unsigned long long seed, long long row0, int tree, double rate, double p0,
                                cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    int64_t g = (n + 255) / 256;
    if (g > 148 * 16) g = 148 * 16;
    poisson_bags_kernel<<<(int)g, 256, 0, s>>>(weights, n, seed, row0, tree, rate, p0);
    return cudaGetLastError();
}
cudaError_t launch_java_bagging_mask(float* weights, int64_t n, long long seed, double rate, long long first_draw, cudaStream_t s) {
    if (n <= 0) return cudaSuccess;
    int64_t g = ((n + ROWS_PER_THREAD - 1) / ROWS_PER_THREAD + 255) / 256;
    if (g > 148 * 8) g = 148 * 8;
    if (g < 1) g = 1;
    java_bagging_mask_kernel<<<(int)g, 256, 0, s>>>(weights, n, seed, rate, first_draw);
    return cudaGetLastError();
}
}  // namespace sb

extern "C" SB_API int sb_bin_values(const float* values, int64_t n, const double* bin_boundary, int32_t n_bins, void* bins_out,
                                    int32_t out_width, void* cuda_stream) {
    if (!values || !bin_boundary || !bins_out || n < 0) return sb::fail(SB_ERR_INVALID, "sb_bin_values: null argument");
    if (n_bins < 1 || n_bins > MAX_BOUNDARIES) return sb::fail(SB_ERR_RANGE, "sb_bin_values: between 1 and 4096 boundaries are supported");
    if (out_width != 1 && out_width != 2) return sb::fail(SB_ERR_INVALID, "sb_bin_values: out_width must be 1 (uint8) or 2 (uint16)");
    if (out_width == 1 && n_bins > 256) return sb::fail(SB_ERR_INVALID, "sb_bin_values: more than 256 bins need uint16 output");
    // The reference's -1 (value below the first boundary) is stored as the all-ones index so that sb_load_rows can
    // reject it; with 256 bins in uint8 that index would be the valid bin 255.
    if (out_width == 1 && n_bins == 256) return sb::fail(SB_ERR_INVALID, "sb_bin_values: 256 bins need uint16 output (index 255 must stay free for values below the first boundary)");
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    if (n == 0) return SB_OK;
    // per-call boundary table, allocated and freed in stream order: calls on different streams or from different host
    // threads do not share it (a single __constant__ table would be overwritten under a running kernel)
    double* d_bounds = nullptr;
    SB_CUDA(cudaMallocAsync((void**)&d_bounds, sizeof(double) * n_bins, stream));
    cudaError_t e = cudaMemcpyAsync(d_bounds, bin_boundary, sizeof(double) * n_bins, cudaMemcpyHostToDevice, stream);
    if (e == cudaSuccess) {
        // 4096 doubles = 32 KB of dynamic shared memory: inside the default 48 KB limit
        int64_t g = (n + 255) / 256;
        if (g > 148 * 8) g = 148 * 8;
        bin_values_kernel<<<(int)g, 256, sizeof(double) * n_bins, stream>>>(values, n, d_bounds, n_bins, bins_out, out_width);
        e = cudaGetLastError();
    }
    cudaFreeAsync(d_bounds, stream);
    SB_CUDA(e);
    return SB_OK;
}
