// Micro-benchmark: shared-memory update throughput on sm_100a for the histogram inner step.
// Measures updates / cycle / SM for the candidate accumulation schemes (see DESIGN.md "why this layout").
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_smem tools/ubench_smem.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x;
}

// MODE 0: u32 ATOMS.ADD x NSTAT, lane-private bank (addr = bin*32 + lane)
// MODE 1: u32 ATOMS.ADD x NSTAT, random addresses (addr = hash % size): bank + address conflicts
// MODE 2: f64 CAS atomicAdd x NSTAT, lane-private column
// MODE 3: f64 plain RMW x NSTAT, warp-private histogram, lane-private column
// MODE 4: f32 plain RMW x NSTAT, warp-private
// MODE 5: f32 CAS atomicAdd x NSTAT lane-private column
// MODE 6: u32 atomic with return + carry into second word (64-bit fixed point emulation) x NSTAT
// MODE 7: u64 CAS atomicAdd x NSTAT lane-private column
// MODE 8: f64 plain RMW, 4 independent rows batched (loads first, then adds, then stores) -- unsafe aliasing ignored (upper bound)
template <int MODE, int NSTAT, int BINS>
__global__ void __launch_bounds__(1024) k(uint64_t* cyc_out, double* sink, int iters, int warp_private) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nwarps = blockDim.x >> 5;
    // zero
    uint32_t* s32 = reinterpret_cast<uint32_t*>(smem_raw);
    const int words_total = (MODE == 0 || MODE == 1 || MODE == 4 || MODE == 5) ? BINS * 32 * NSTAT * (warp_private ? nwarps : 1)
                          : BINS * 32 * NSTAT * 2 * (warp_private ? nwarps : 1);
    for (int i = threadIdx.x; i < words_total; i += blockDim.x) s32[i] = 0;
    __syncthreads();
    uint32_t st = hash32(blockIdx.x * 1024 + threadIdx.x + 1);
    uint64_t t0 = clock64();
    if (MODE == 0 || MODE == 1) {
        uint32_t* h = s32 + (warp_private ? warp * BINS * 32 * NSTAT : 0);
        for (int it = 0; it < iters; ++it) {
            st = st * 1664525u + 1013904223u;
            uint32_t bin = (st >> 16) % BINS;
            uint32_t a = (MODE == 0) ? (bin * 32 + lane) * NSTAT : ((st >> 8) % (BINS * 32)) * NSTAT;
#pragma unroll
            for (int s = 0; s < NSTAT; ++s) atomicAdd(&h[a + s], 1u + s);
        }
    } else if (MODE == 2) {
        double* h = reinterpret_cast<double*>(smem_raw) + (warp_private ? warp * BINS * 32 * NSTAT : 0);
        for (int it = 0; it < iters; ++it) {
            st = st * 1664525u + 1013904223u;
            uint32_t bin = (st >> 16) % BINS;
            uint32_t a = (bin * 32 + lane) * NSTAT;
#pragma unroll
            for (int s = 0; s < NSTAT; ++s) atomicAdd(&h[a + s], 1.0 + s);
        }
    } else if (MODE == 3) {
        double* h = reinterpret_cast<double*>(smem_raw) + warp * BINS * 32 * NSTAT;
        for (int it = 0; it < iters; ++it) {
            st = st * 1664525u + 1013904223u;
            uint32_t bin = (st >> 16) % BINS;
            // layout [bin][stat][lane] -> conflict-free 64-bit
            double* p = h + (bin * NSTAT) * 32 + lane;
#pragma unroll
            for (int s = 0; s < NSTAT; ++s) p[s * 32] += 1.0 + s;
        }
    } else if (MODE == 4) {
        float* h = reinterpret_cast<float*>(smem_raw) + warp * BINS * 32 * NSTAT;
        for (int it = 0; it < iters; ++it) {
            st = st * 1664525u + 1013904223u;
            uint32_t bin = (st >> 16) % BINS;
            float* p = h + (bin * NSTAT) * 32 + lane;
#pragma unroll
            for (int s = 0; s < NSTAT; ++s) p[s * 32] += 1.0f + s;
        }
    } else if (MODE == 5) {
        float* h = reinterpret_cast<float*>(smem_raw) + (warp_private ? warp * BINS * 32 * NSTAT : 0);
        for (int it = 0; it < iters; ++it) {
            st = st * 1664525u + 1013904223u;
            uint32_t bin = (st >> 16) % BINS;
            float* p = h + (bin * NSTAT) * 32 + lane;
#pragma unroll
            for (int s = 0; s < NSTAT; ++s) atomicAdd(&p[s * 32], 1.0f + s);
        }
    } else if (MODE == 6) {
        uint32_t* h = s32 + (warp_private ? warp * BINS * 32 * NSTAT * 2 : 0);
        for (int it = 0; it < iters; ++it) {
            st = st * 1664525u + 1013904223u;
            uint32_t bin = (st >> 16) % BINS;
            uint32_t* p = h + (bin * NSTAT * 2) * 32 + lane;
#pragma unroll
            for (int s = 0; s < NSTAT; ++s) {
                uint32_t lo = st ^ (s * 0x9e3779b9u);
                uint32_t old = atomicAdd(&p[(2 * s) * 32], lo);
                uint32_t carry = (old + lo < old) ? 1u : 0u;
                atomicAdd(&p[(2 * s + 1) * 32], 3u + carry);
            }
        }
    } else if (MODE == 7) {
        unsigned long long* h = reinterpret_cast<unsigned long long*>(smem_raw) + (warp_private ? warp * BINS * 32 * NSTAT : 0);
        for (int it = 0; it < iters; ++it) {
            st = st * 1664525u + 1013904223u;
            uint32_t bin = (st >> 16) % BINS;
            unsigned long long* p = h + (bin * NSTAT) * 32 + lane;
#pragma unroll
            for (int s = 0; s < NSTAT; ++s) atomicAdd(&p[s * 32], (unsigned long long)(1 + s));
        }
    } else if (MODE == 8) {
        double* h = reinterpret_cast<double*>(smem_raw) + warp * BINS * 32 * NSTAT;
        for (int it = 0; it < iters; it += 4) {
            double* p[4]; double v[4][NSTAT];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                st = st * 1664525u + 1013904223u;
                uint32_t bin = (st >> 16) % BINS;
                p[j] = h + (bin * NSTAT) * 32 + lane;
            }
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int s = 0; s < NSTAT; ++s) v[j][s] = p[j][s * 32];
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int s = 0; s < NSTAT; ++s) p[j][s * 32] = v[j][s] + (1.0 + s);
        }
    }
    __syncthreads();
    uint64_t t1 = clock64();
    if (threadIdx.x == 0) cyc_out[blockIdx.x] = t1 - t0;
    // sink so nothing is optimised away
    double acc = 0;
    for (int i = threadIdx.x; i < words_total; i += blockDim.x) acc += s32[i];
    if (acc == 1.2345) sink[0] = acc;
}

template <int MODE, int NSTAT, int BINS>
void run(const char* name, int threads, int warp_private, int iters, int bytes_per_slot) {
    int nwarps = threads / 32;
    size_t smem = (size_t)BINS * 32 * NSTAT * bytes_per_slot * (warp_private ? nwarps : 1);
    if (smem > 227 * 1024) { printf("%-44s thr=%4d priv=%d  SKIP smem %zu\n", name, threads, warp_private, smem); return; }
    CK(cudaFuncSetAttribute(k<MODE, NSTAT, BINS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    uint64_t* cyc; double* sink;
    int grid = 148;
    CK(cudaMalloc(&cyc, grid * sizeof(uint64_t))); CK(cudaMalloc(&sink, 8));
    k<MODE, NSTAT, BINS><<<grid, threads, smem>>>(cyc, sink, iters, warp_private);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MODE, NSTAT, BINS><<<grid, threads, smem>>>(cyc, sink, iters, warp_private);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    uint64_t h[148]; CK(cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost));
    double mean = 0; for (int i = 0; i < grid; ++i) mean += h[i]; mean /= grid;
    double cells = (double)threads * iters;
    printf("%-44s thr=%4d priv=%d smem=%6zuB  cyc=%9.0f  cells/cyc/SM=%7.3f  stat-upd/cyc/SM=%7.3f  ms=%.3f  Gcells/s(chip)=%8.1f\n",
           name, threads, warp_private, smem, mean, cells / mean, cells * NSTAT / mean, ms, cells * grid / (ms * 1e6));
    cudaFree(cyc); cudaFree(sink);
}

int main() {
    const int IT = 20000;
    int thr[] = {128, 256, 512, 1024};
    for (int t : thr) {
        run<0, 1, 64>("u32 ATOMS.ADD x1 lane-bank shared", t, 0, IT, 4);
        run<0, 3, 64>("u32 ATOMS.ADD x3 lane-bank shared", t, 0, IT, 4);
        run<0, 5, 64>("u32 ATOMS.ADD x5 lane-bank shared", t, 0, IT, 4);
        run<1, 1, 64>("u32 ATOMS.ADD x1 random-addr shared", t, 0, IT, 4);
        run<1, 3, 64>("u32 ATOMS.ADD x3 random-addr shared", t, 0, IT, 4);
        run<6, 2, 64>("u32 ATOMS ret+carry (2x64b fixed) shared", t, 0, IT, 8);
        run<2, 3, 64>("f64 CAS atomicAdd x3 lane-col shared", t, 0, IT, 8);
        run<7, 3, 64>("u64 CAS atomicAdd x3 lane-col shared", t, 0, IT, 8);
        run<5, 3, 64>("f32 CAS atomicAdd x3 lane-col shared", t, 0, IT, 4);
        run<3, 3, 64>("f64 plain RMW x3 warp-private", t, 1, IT, 8);
        run<8, 3, 64>("f64 plain RMW x3 warp-private batched4", t, 1, IT, 8);
        run<4, 3, 64>("f32 plain RMW x3 warp-private", t, 1, IT, 4);
        run<3, 3, 16>("f64 plain RMW x3 warp-private 16bins", t, 1, IT, 8);
        run<8, 3, 16>("f64 plain RMW x3 warp-priv batched4 16bins", t, 1, IT, 8);
        run<4, 3, 16>("f32 plain RMW x3 warp-private 16bins", t, 1, IT, 8);
    }
    return 0;
}
