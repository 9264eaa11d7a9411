// p2p.cu -- the DTMaster merge (dt/DTMaster.java:298-315: add the workers' statistics node by node) as an all-reduce of
// the integer accumulators over NVLink peer memory, for the GPUs of one box.
//
// Every rank keeps its accumulators in a buffer that all other ranks have mapped (CUDA IPC).  After the statistics
// kernels of a pass:
//   barrier   every rank has finished writing its own buffer
//   reduce    rank r owns the r-th slice: it reads that slice from all W buffers over NVLink (plain peer loads -- no TMA,
//             no multicast), adds the integers in rank order and writes the sum into its own buffer
//   barrier   every slice is final
//   gather    rank r copies the other slices from their owners' buffers
//   barrier   nobody still reads a buffer that the next pass clears (skipped when the two buffers alternate)
// Integer addition makes the result independent of the order of the reads, so it equals the NCCL sum bit for bit.  The
// barriers are flag words in peer memory written with system-scope release stores and polled with acquire loads; a
// poll gives up after about thirty seconds and raises the handle's error flag instead of hanging the GPU.
// NCCL still carries the setup (the IPC handles travel in an all-gather) and the two small floating-point
// reductions at a tree switch.  Measured at 8 GPUs: see DESIGN.md section 6.
#include "common.cuh"
#include "kernels.cuh"

namespace sb {

namespace {

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// peer data: bypass L1 (another GPU wrote it; the line must come from its home L2)
__device__ __forceinline__ longlong2 ld_peer(const longlong2* p) {
    longlong2 v;
    asm volatile("ld.relaxed.sys.global.v2.s64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
    return v;
}

// One block.  Thread t < world tells rank t "I reached barrier `slot` of round `epoch`" and waits for rank t's word.
__global__ void p2p_barrier_kernel(P2PView v, int slot, uint32_t epoch, uint32_t* error_flag) {
    const int t = threadIdx.x;
    if (t < v.world) {
        __threadfence_system();
        st_release_sys(v.peer_flags[t] + slot * P2P_MAX_WORLD + v.rank, epoch);
        const uint32_t* mine = v.peer_flags[v.rank] + slot * P2P_MAX_WORLD + t;
        const long long t0 = clock64();
        // epochs only grow; a peer that is already a round ahead has passed this barrier too
        while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
            if (clock64() - t0 > 60000000000LL) {  // ~30 s at 1.9 GHz: a peer is gone (ranks may start seconds apart)
                atomicOr(error_flag, 16u);
                break;
            }
            __nanosleep(64);
        }
    }
    __syncthreads();
}

// slice r of the element range, in 16-byte units
__device__ __forceinline__ void slice_of(int64_t n2, int world, int r, int64_t* b, int64_t* e) {
    const int64_t per = (n2 + world - 1) / world;
    *b = min(n2, per * r);
    *e = min(n2, per * (r + 1));
}

__global__ void __launch_bounds__(256) p2p_reduce_kernel(P2PView v, int which, int64_t n2) {
    int64_t b, e;
    slice_of(n2, v.world, v.rank, &b, &e);
    longlong2* mine = reinterpret_cast<longlong2*>(v.peer_buf[which][v.rank]);
    for (int64_t i = b + blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < e; i += (int64_t)gridDim.x * blockDim.x) {
        longlong2 s = make_longlong2(0, 0);
#pragma unroll
        for (int p = 0; p < P2P_MAX_WORLD; ++p) {
            if (p < v.world) {
                const longlong2 x = ld_peer(reinterpret_cast<const longlong2*>(v.peer_buf[which][p]) + i);
                s.x += x.x;
                s.y += x.y;
            }
        }
        mine[i] = s;
    }
}

__global__ void __launch_bounds__(256) p2p_gather_kernel(P2PView v, int which, int64_t n2) {
    longlong2* mine = reinterpret_cast<longlong2*>(v.peer_buf[which][v.rank]);
    const int64_t per = (n2 + v.world - 1) / v.world;
    // one flat index over the other ranks' slices: (slice, offset)
    const int64_t total = per * v.world;
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < total; j += (int64_t)gridDim.x * blockDim.x) {
        const int p = (int)(j / per);
        const int64_t i = j;  // slices are contiguous: element j belongs to rank j / per
        if (p == v.rank || i >= n2) continue;
        mine[i] = ld_peer(reinterpret_cast<const longlong2*>(v.peer_buf[which][p]) + i);
    }
}

}  // namespace

cudaError_t launch_p2p_allreduce(const P2PView& v, int which, int64_t n_elems, uint32_t epoch, int tail_barrier, uint32_t* error_flag,
                                 cudaStream_t s) {
    // n_elems int64 values, an even number (lo and hi planes of equal length): 16-byte units
    const int64_t n2 = (n_elems + 1) / 2;
    p2p_barrier_kernel<<<1, 32, 0, s>>>(v, 0, epoch, error_flag);
    if (n2 > 0) {
        const int64_t per = (n2 + v.world - 1) / v.world;
        int grid = (int)std::min<int64_t>((per + 255) / 256, 148 * 4);
        if (grid < 1) grid = 1;
        p2p_reduce_kernel<<<grid, 256, 0, s>>>(v, which, n2);
        p2p_barrier_kernel<<<1, 32, 0, s>>>(v, 1, epoch, error_flag);
        grid = (int)std::min<int64_t>((n2 + 255) / 256, 148 * 8);
        if (grid < 1) grid = 1;
        p2p_gather_kernel<<<grid, 256, 0, s>>>(v, which, n2);
    }
    // When the two buffers alternate from pass to pass (histogram subtraction keeps the previous pass), a buffer is
    // cleared again only after its owner passed the first two barriers of the NEXT pass, which every rank reaches after
    // its gather of this one: the third barrier is then implied.
    if (tail_barrier) p2p_barrier_kernel<<<1, 32, 0, s>>>(v, 2, epoch, error_flag);
    return cudaGetLastError();
}

}  // namespace sb
