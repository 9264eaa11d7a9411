// noalloc.cu -- sb_error_sums_static: weighted error sums of a row set with NO device allocation of any kind.
//
// Who needs this.  sb_compute gets its reduction scratch from the handle, and creating a handle allocates.  Two
// kinds of caller cannot do that.  (1) A monitor that re-scores a predict vector (early-stop check, validation
// curve) from a side thread while the trainer's histogram kernel is running: cudaMalloc / cudaFree synchronise the
// whole device and take the allocator lock, so a monitor that allocates stalls the training stream it was meant
// to watch.  (2) A deployment that has handed all of HBM to the bin matrix up front (the 1B-row shard is sized to
// fill the card): there a late cudaMalloc simply fails.  For them the scratch lives in a file-scope __device__
// array that is part of the loaded module: every GPU that loads libshifu_b200.so holds these 64 KB until the
// process exits, and nothing is ever allocated or freed by this entry point.
//
// Limits and rules, stated once:
//   * at most SB_STATIC_MAX_ROWS (67,108,864) rows per call; a larger n is rejected with SB_ERR_RANGE;
//   * all pointers are device pointers on the CURRENT device;
//   * the first call on each device looks up the scratch address (cudaGetSymbolAddress) -- make that first call
//     before any stream capture begins on that device;
//   * every call on a device shares the one scratch array, which keeps whatever the previous call left in it.
//     The only word that must start at zero (the finish ticket) is cleared on the stream before each launch; the
//     per-block partials are fully overwritten before they are read.  Two calls that overlap on the GPU (different
//     streams, different threads) race on the array and produce garbage: serialise them.
#include <atomic>

#include "common.cuh"

namespace {

constexpr int NA_TPB = 256;
constexpr int NA_ROWS_PER_BLOCK = 16384;
constexpr int NA_MAX_BLOCKS = 4096;
constexpr int64_t NA_MAX_ROWS = (int64_t)NA_MAX_BLOCKS * NA_ROWS_PER_BLOCK;  // 67,108,864
static_assert(NA_MAX_ROWS == SB_STATIC_MAX_ROWS, "header and kernel disagree on the row limit");

// [2*b], [2*b+1]: partial (count, error) of block b;  [2*NA_MAX_BLOCKS]: finish ticket (an unsigned in a double slot)
static __device__ __align__(256) double g_static_scratch[2 * NA_MAX_BLOCKS + 1];

// dt/Loss.java:36-100 computeError, evaluated in float exactly as Java does (this file is built with -fmad=false).
__device__ __forceinline__ float static_loss_error(int kind, float predict, float label) {
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

// `scratch` is the DEVICE address of g_static_scratch, looked up by the host with cudaGetSymbolAddress: in host code
// the symbol's name is only a stand-in and must never be passed to a kernel as if it were a pointer.
__global__ void __launch_bounds__(NA_TPB) error_sums_static_kernel(int64_t n, const float* __restrict__ label,
                                                                   const float* __restrict__ predict,
                                                                   const float* __restrict__ significance, int loss,
                                                                   double* scratch, double* out) {
    const int64_t begin = (int64_t)blockIdx.x * NA_ROWS_PER_BLOCK;
    const int64_t end = min(n, begin + NA_ROWS_PER_BLOCK);
    double cnt = 0.0, err = 0.0;
    for (int64_t r = begin + threadIdx.x; r < end; r += NA_TPB) {
        const float sig = significance ? significance[r] : 1.0f;
        err += (double)(sig * static_loss_error(loss, predict[r], label[r]));  // dt/DTWorker.java:681-682
        cnt += (double)sig;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        cnt += __shfl_down_sync(0xffffffffu, cnt, o);
        err += __shfl_down_sync(0xffffffffu, err, o);
    }
    __shared__ double sh[2][NA_TPB / 32];
    __shared__ bool is_last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
        sh[0][warp] = cnt;
        sh[1][warp] = err;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double c = 0.0, e = 0.0;
        for (int w = 0; w < NA_TPB / 32; ++w) {
            c += sh[0][w];
            e += sh[1][w];
        }
        scratch[2 * blockIdx.x] = c;
        scratch[2 * blockIdx.x + 1] = e;
        __threadfence();
        unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + 2 * NA_MAX_BLOCKS);
        is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (is_last && threadIdx.x < 2) {  // fixed block order: the same input always gives the same bits
        __threadfence();
        double s = 0.0;
        for (unsigned int b = 0; b < gridDim.x; ++b) s += __ldcg(&scratch[2 * b + threadIdx.x]);
        out[threadIdx.x] = s;
    }
}

constexpr int MAX_DEVICES = 64;
// device address of g_static_scratch, per device ordinal (null = not looked up yet).  Atomic because the first call on a
// device may come from two host threads at once (a monitor thread beside the trainer); both would store the same address.
std::atomic<double*> g_scratch_addr[MAX_DEVICES];

}  // namespace

extern "C" SB_API int sb_error_sums_static(const float* label, const float* predict, const float* significance,
                                           int64_t n_rows, int32_t loss, double* out, void* cuda_stream) {
    if (!label || !predict || !out) return sb::fail(SB_ERR_INVALID, "sb_error_sums_static: null label / predict / out");
    if (n_rows < 0) return sb::fail(SB_ERR_INVALID, "sb_error_sums_static: negative row count");
    if (n_rows > NA_MAX_ROWS)
        return sb::fail(SB_ERR_RANGE, "sb_error_sums_static: " + std::to_string(n_rows) + " rows exceed the static scratch limit of " +
                                          std::to_string(NA_MAX_ROWS) + " rows per call; split the rows or use sb_compute");
    if (loss < SB_LOSS_SQUARED || loss > SB_LOSS_LOG) return sb::fail(SB_ERR_INVALID, "sb_error_sums_static: unknown loss");
    cudaStream_t stream = (cudaStream_t)cuda_stream;
    int dev = -1;
    SB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES) return sb::fail(SB_ERR_UNSUPPORTED, "sb_error_sums_static: device ordinal beyond 63");
    if (!g_scratch_addr[dev].load(std::memory_order_relaxed)) {
        // one-time set-up for this device, with this device current (the caller made it current)
        double* p = nullptr;
        SB_CUDA(cudaGetSymbolAddress((void**)&p, g_static_scratch));
        if (!p) return sb::fail(SB_ERR_CUDA, "sb_error_sums_static: cudaGetSymbolAddress returned a null address");
        g_scratch_addr[dev].store(p, std::memory_order_relaxed);
    }
    double* scratch = g_scratch_addr[dev].load(std::memory_order_relaxed);
    if (n_rows == 0) {
        SB_CUDA(cudaMemsetAsync(out, 0, 2 * sizeof(double), stream));
        return SB_OK;
    }
    // the array still holds the previous call's contents: only the ticket has to start at zero
    SB_CUDA(cudaMemsetAsync(scratch + 2 * NA_MAX_BLOCKS, 0, sizeof(double), stream));
    const int blocks = (int)((n_rows + NA_ROWS_PER_BLOCK - 1) / NA_ROWS_PER_BLOCK);
    error_sums_static_kernel<<<blocks, NA_TPB, 0, stream>>>(n_rows, label, predict, significance, loss, scratch, out);
    SB_CUDA(cudaGetLastError());
    return SB_OK;
}
