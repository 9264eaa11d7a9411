// rowlist.cu -- per-pass row lists for levels whose todo nodes do not fit one shared-memory slot group.
//
// When a pass has more todo nodes than one block can hold histograms for (GS slots), the column-major tile kernel
// would have to stream every row tile once per slot group.  Instead the rows are grouped by slot group once per
// pass (a one-digit radix sort of (group id, row index) -- row indices stay ascending inside a group), their
// fixed-point operands are gathered into a compact array in that order, and hist_rows_kernel (hist.cu) walks each
// group's list, reading the 32 bins of a row and feature tile from a ROW-major copy of the bin matrix with one
// coalesced 32-byte load.  Every active row is then touched once per feature tile, whatever the number of groups.
#include <cub/device/device_radix_sort.cuh>

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

__global__ void rowlist_keys_kernel(int64_t n, const int32_t* __restrict__ node_id, const float* __restrict__ weights,
                                    const int16_t* __restrict__ slot_of_node, int slot_table_size, int gs,
                                    uint8_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < n; r += (int64_t)gridDim.x * blockDim.x) {
        const int nid = node_id[r];
        int s = (nid >= 0 && nid < slot_table_size) ? (int)slot_of_node[nid] : -1;
        const float w = weights ? weights[r] : 1.0f;
        keys[r] = (s >= 0 && w != 0.0f) ? (uint8_t)(s / gs) : (uint8_t)255;
        vals[r] = (uint32_t)r;
    }
}

// offsets[g] = first position of key >= g in the sorted keys, g = 0..n_groups (offsets[n_groups] = number of active
// rows); then the chunk table: chunk c covers ops [begin, end) of one group.
__global__ void rowlist_chunks_kernel(const uint8_t* __restrict__ sorted_keys, int64_t n, int n_groups, int64_t max_chunk_rows,
                                      int target_chunks, int64_t* __restrict__ offsets, int32_t* __restrict__ chunk_group,
                                      int64_t* __restrict__ chunk_begin, int64_t* __restrict__ chunk_end,
                                      int32_t* __restrict__ n_chunks, int max_chunks) {
    __shared__ int64_t off[257];
    for (int g = threadIdx.x; g <= n_groups; g += blockDim.x) {
        int64_t lo = 0, hi = n;
        if (!sorted_keys) lo = hi = (g == 0 ? 0 : n);  // identity list: a single group holding every row
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if ((int)sorted_keys[mid] < g) lo = mid + 1;
            else hi = mid;
        }
        off[g] = lo;
        offsets[g] = lo;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        // chunk size from the number of ACTIVE rows of this pass (with subtraction most passes touch a fraction of the
        // rows): about target_chunks chunks, so the grid fills the GPU several times over whatever the list length
        const int64_t total = off[n_groups];
        int64_t chunk_rows = ((total + target_chunks - 1) / target_chunks + 127) / 128 * 128;
        if (chunk_rows < 1024) chunk_rows = 1024;
        if (chunk_rows > max_chunk_rows) chunk_rows = max_chunk_rows;
        int c = 0;
        for (int g = 0; g < n_groups; ++g)
            for (int64_t b = off[g]; b < off[g + 1] && c < max_chunks; b += chunk_rows) {
                chunk_group[c] = g;
                chunk_begin[c] = b;
                chunk_end[c] = min(off[g + 1], b + chunk_rows);
                ++c;
            }
        *n_chunks = c;
    }
}

// ops[i] for the i-th row of the sorted order: {row index, local slot, fixed-point operands} (see RowOps in kernels.cuh)
template <int KIND, int CNT_INT>
__global__ void rowlist_ops_kernel(const uint32_t* __restrict__ sorted_rows, const int64_t* __restrict__ offsets, int n_groups,
                                   const int32_t* __restrict__ node_id, const float* __restrict__ output,
                                   const float* __restrict__ significance, const float* __restrict__ weights,
                                   const int16_t* __restrict__ slot_of_node, int slot_table_size, int gs, int stats_size,
                                   const float* __restrict__ scales, RowOps* __restrict__ ops, uint32_t* error_flag) {
    const int64_t total = offsets[n_groups];
    const float sc0 = scales[0], sc1 = scales[1], sc2 = scales[2];
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        // sorted_rows == nullptr: identity list (a pass whose single todo node is a tree root takes the rows in
        // storage order; rows that do not belong to it are marked instead of being sorted out)
        const uint32_t r = sorted_rows ? sorted_rows[i] : (uint32_t)i;
        int s;
        if (sorted_rows) s = slot_of_node[node_id[r]];
        else {
            const int nid = node_id[r];
            s = (nid >= 0 && nid < slot_table_size) ? (int)slot_of_node[nid] : -1;
        }
        const float w = weights ? weights[r] : 1.0f;
        if (s < 0 || w == 0.0f) {
            RowOps skip;
            skip.row = r;
            skip.slot = 0xffffffffu;
#pragma unroll
            skip.cnt_lo = skip.cnt_hi = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) skip.q[k] = 0;
            ops[i] = skip;
            continue;
        }
        const float sig = significance[r];
        const float y = output[r];
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
        } else {
            const int cls = (int)(__fadd_rn(y, 0.000001f));
            if (cls < 0 || cls >= stats_size) { atomicOr(error_flag, 1u); o.slot = 0xffffffffu; o.q[0] = 0; }
            else o.q[0] = (uint32_t)(cls * (CNT_INT ? 128 : 256));  // byte offset of the class words
            o.q[1] = o.q[2] = o.q[3] = 0;
        }
        ops[i] = o;
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

size_t rowlist_sort_temp_bytes(int64_t n) {
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint8_t*)nullptr, (uint8_t*)nullptr, (const uint32_t*)nullptr,
                                    (uint32_t*)nullptr, n, 0, 8);
    return bytes;
}

cudaError_t launch_rowlist_build(const RowListArgs& a, cudaStream_t s) {
    const int64_t n = a.n;
    if (n <= 0) return cudaSuccess;
    int64_t g = (n + TPB - 1) / TPB;
    if (g > 148 * 16) g = 148 * 16;
    cudaError_t e;
    if (a.identity) {
        // no keys, no sort: one group = all rows in storage order
        rowlist_chunks_kernel<<<1, 256, 0, s>>>(nullptr, n, 1, a.chunk_rows, a.target_chunks, a.offsets, a.chunk_group,
                                                a.chunk_begin, a.chunk_end, a.n_chunks, a.max_chunks);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    } else {
        rowlist_keys_kernel<<<(int)g, TPB, 0, s>>>(n, a.node_id, a.weights, a.slot_of_node, a.slot_table_size, a.gs, a.keys_in, a.vals_in);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        size_t temp = a.temp_bytes;
        e = cub::DeviceRadixSort::SortPairs(a.temp, temp, (const uint8_t*)a.keys_in, a.keys_out, (const uint32_t*)a.vals_in, a.vals_out, n, 0, 8, s);
        if (e != cudaSuccess) return e;
        rowlist_chunks_kernel<<<1, 256, 0, s>>>(a.keys_out, n, a.n_groups, a.chunk_rows, a.target_chunks, a.offsets, a.chunk_group,
                                                a.chunk_begin, a.chunk_end, a.n_chunks, a.max_chunks);
        e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
#define OPS(K, C)                                                                                                          \
    rowlist_ops_kernel<K, C><<<(int)g, TPB, 0, s>>>(a.identity ? nullptr : a.vals_out, a.offsets, a.identity ? 1 : a.n_groups,  \
                                                    a.node_id, a.output, a.significance, a.weights, a.slot_of_node,            \
                                                    a.slot_table_size, a.gs, a.stats_size, a.scales, a.ops, a.error_flag)
    if (a.kind == STAT_VARIANCE) { if (a.cnt_int) OPS(STAT_VARIANCE, 1); else OPS(STAT_VARIANCE, 0); }
    else { if (a.cnt_int) OPS(STAT_CLASS, 1); else OPS(STAT_CLASS, 0); }
#undef OPS
    return cudaGetLastError();
}

}  // namespace sb
