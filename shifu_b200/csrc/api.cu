// api.cu -- the C ABI of include/shifu_b200.h: handle, data load, the DTWorker.doCompute replacement and the
// on-device DTMaster best-split step.  Host logic only; kernels live in hist.cu / rows.cu / split.cu.
#include <cuda.h>
#include <dlfcn.h>
#include <nccl.h>  // types and enums only: the library is opened at run time (sb_dist_init)

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>

#include "common.cuh"
#include "kernels.cuh"

namespace sb {

static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }
int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

template <typename T>
struct DevBuf {  // grow-only device buffer; or an alias of a piece of the per-pass staging arena (not owned)
    T* p = nullptr;
    size_t cap = 0;
    bool owned = true;
    size_t fixed = 0;  // capacity of an aliased buffer that may be kept as long as it is large enough
    // Growth is stream-ordered (cudaMallocAsync / cudaFreeAsync): cudaFree would wait for everything queued on the
    // device, and the buffers of a pass grow while the kernels of its first half are still running.
    cudaError_t reserve(cudaStream_t stream, size_t n) {
        if (owned && n <= cap) return cudaSuccess;
        if (!owned && n <= fixed) return cudaSuccess;
        fixed = 0;
        // 2x: the statistics of a level-wise tree double with every level (a buffer then serves two levels), and a
        // pool that has to map new memory costs about 10 us per MB (prewarm_pool reserves the usual working set)
        const size_t want = std::max<size_t>((owned && cap) ? 2 * n : n, 16);
        if (p && owned) cudaFreeAsync(p, stream);
        p = nullptr;
        cap = 0;
        owned = true;
        cudaError_t e = cudaMallocAsync((void**)&p, want * sizeof(T), stream);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void alias(T* q, size_t fixed_capacity = 0) {
        if (p && owned) cudaFreeAsync(p, 0);
        p = q;
        cap = 0;
        owned = false;
        fixed = fixed_capacity;
    }
    void release() {
        if (p && owned) cudaFreeAsync(p, 0);
        p = nullptr;
        cap = 0;
        owned = true;
    }
};

// The small tables of a pass (tree arrays, tile / slot / offset tables, sibling pairs, split tasks) are written into ONE
// pinned host arena and reach the device with one asynchronous copy per flush, instead of a dozen separate copies
// from pageable memory -- those are staged by the driver and, above 64 KB, wait for the work already queued on the
// stream.  The arena is recycled at the start of every pass: a pass ends with a stream synchronisation, so by then
// both the copies and the kernels that read the device side are done.
struct Stager {
    uint8_t* host = nullptr;  // pinned
    uint8_t* dev = nullptr;
    size_t cap = 0, off = 0, flushed = 0;
    std::vector<std::pair<uint8_t*, uint8_t*>> retired;  // arenas that were too small; alive until the next reset
    cudaError_t grow(size_t need, cudaStream_t stream) {
        cudaError_t e = flush(stream);  // what was handed out from the old arena must still arrive
        if (e != cudaSuccess) return e;
        if (host) retired.push_back({host, dev});
        host = dev = nullptr;
        const size_t want = std::max<size_t>(std::max<size_t>(2 * cap, need + need / 2), 1 << 20);
        e = cudaMallocHost((void**)&host, want);
        if (e != cudaSuccess) return e;
        e = cudaMalloc((void**)&dev, want);
        if (e != cudaSuccess) return e;
        cap = want;
        off = flushed = 0;
        return cudaSuccess;
    }
    // room for n elements: *hp = where to write them now, *d = where they will be on the device after the next flush
    template <typename T>
    cudaError_t take(size_t n, T** hp, T** d, cudaStream_t stream) {
        const size_t bytes = n * sizeof(T);
        size_t at = (off + 255) & ~(size_t)255;
        if (at + bytes > cap) {
            cudaError_t e = grow(bytes + 256, stream);
            if (e != cudaSuccess) return e;
            at = 0;
        }
        *hp = reinterpret_cast<T*>(host + at);
        *d = reinterpret_cast<T*>(dev + at);
        off = at + bytes;
        return cudaSuccess;
    }
    template <typename T>
    cudaError_t put(const T* src, size_t n, T** d, cudaStream_t stream) {
        T* hp = nullptr;
        cudaError_t e = take(n, &hp, d, stream);
        if (e == cudaSuccess && n) memcpy(hp, src, n * sizeof(T));
        return e;
    }
    cudaError_t flush(cudaStream_t stream) {
        if (off == flushed) return cudaSuccess;
        cudaError_t e = cudaMemcpyAsync(dev + flushed, host + flushed, off - flushed, cudaMemcpyHostToDevice, stream);
        flushed = off;
        return e;
    }
    void reset() {  // only after a stream synchronisation
        for (auto& r : retired) {
            cudaFreeHost(r.first);
            cudaFree(r.second);
        }
        retired.clear();
        off = flushed = 0;
    }
    void release() {
        reset();
        if (host) cudaFreeHost(host);
        if (dev) cudaFree(dev);
        host = dev = nullptr;
        cap = 0;
    }
};

struct HistLaunch {
    int kind;
    int cnt_int;
    HistArgs args;
    int grid;
    size_t smem;
    // row-list path (more todo nodes than one slot group): lists are rebuilt before the kernel, every pass
    bool use_rows = false, has_narrow = false, has_wide = false;
    int fpl = 1, allf = 0;
    HistRowsArgs wide{};
    int wide_grid = 0;
    size_t wide_smem = 0;
    RowListArgs list{};
    HistRowsArgs rows{};
    int root_counts = 0;   // 1: save the bin counts of this root pass after the kernel; 2: the kernel skips them, restore
    int root_wcol = 0;     // sample-weight column the counts belong to
    int64_t root_bins = 0;
};

}  // namespace sb

using namespace sb;

struct sb_handle {
    sb_config cfg{};
    std::vector<int32_t> n_bins, column_num;
    std::vector<uint8_t> is_cat;
    std::vector<double> bin_boundary;
    std::vector<int> all_features;  // getAllValidFeatures (dt/DTWorker.java:965-990)
    std::vector<int> all_feature_pos;   // feature slot -> its position in all_features, or -1
    std::vector<int64_t> all_feat_off;  // offset of each of them inside a node block that takes all of them
    int64_t all_feat_total = 0;
    std::vector<int> wide_index;    // per feature slot: index among the wide (> 256 bins, uint16) features, or -1
    int n_wide = 0;
    int stats_size = 3;
    int stat_kind = STAT_VARIANCE;
    int max_bins = 0;
    cudaStream_t stream = nullptr;

    RowSet train, valid;
    Stager stage;
    int32_t* d_feat_bins = nullptr;
    uint8_t* d_is_cat = nullptr;       // [n_features]
    int32_t* d_column_num = nullptr;   // [n_features]
    bool any_categorical = false;
    int max_categorical_bins = 0;
    std::map<int, DeviceTree> trees;  // by tree id
    int tracked_tree = -1;            // GBT: tree id whose node ids train.node_id currently holds

    // iteration state
    TodoLayout layout;
    bool begun = false, stats_ready = false;
    std::vector<HistLaunch> launches;
    DevBuf<int64_t> raw;      // [2 * total]: lo then hi
    // histogram subtraction: the previous pass's merged accumulators and layout stay around for one pass
    DevBuf<int64_t> raw_prev;
    TodoLayout prev_layout;
    bool prev_valid = false;   // raw_prev / prev_layout hold the merged accumulators of the last FINISHED pass
    bool use_prev = false;     // this pass may derive children from them (prev_valid as seen when the pass began)
    bool enable_subtraction = true;
    std::vector<PairDesc> pairs;
    DevBuf<PairDesc> d_pairs;
    DevBuf<int32_t> d_small_is_b;
    int64_t pair_max_len = 0;
    DevBuf<double> stats;     // [total]
    DevBuf<int32_t> tile_feats;
    DevBuf<uint32_t> slot_fmask;
    DevBuf<int64_t> pair_off;
    DevBuf<int16_t> slot_of_node;
    DevBuf<SplitTask> tasks;
    DevBuf<int64_t> task_node_off;
    DevBuf<int32_t> node_task_begin;
    DevBuf<sb_gain> cand, gain_out;
    DevBuf<uint8_t> cand_bits, out_bits, new_dropout;
    DevBuf<float> new_weights;
    DevBuf<double> sort_keys;
    DevBuf<int> sort_idx, sort_tmp;
    DevBuf<uint8_t> sort_flags;
    int n_tasks = 0;
    bool has_cat_tasks = false;
    double* d_acc = nullptr;          // 4 doubles
    ncclComm_t comm = nullptr;        // set by sb_dist_init: the pass merges itself over NCCL
    struct P2P {  // accumulators in peer-mapped memory: the merge is launch_p2p_allreduce (p2p.cu) instead of ncclAllReduce
        bool ok = false;
        P2PView view{};
        int64_t* buf[2] = {nullptr, nullptr};
        uint32_t* flags = nullptr;
        size_t cap_elems = 0;
        uint32_t epoch = 0;
        std::vector<void*> opened;  // peer mappings to close
    } p2p;
    int world = 1, rank = 0;
    void* pin = nullptr;              // pinned host block: PinnedOut, then the sb_gain records (+ category bits)
    size_t pin_bytes = 0;
    double* d_err_scratch = nullptr;  // ERR_SCRATCH_DOUBLES
    uint32_t* d_maxbits = nullptr;    // 3 words (+1 pad)
    float* d_scales = nullptr;        // 3
    double* d_inv_scales = nullptr;   // 3
    uint32_t* d_error_flag = nullptr;
    unsigned long long* d_rows_streamed = nullptr;  // active rows the statistics kernels read (row-list passes)
    int64_t rows_streamed_host = 0;                 // the same for passes without row lists (every row, once)
    int* d_cnt_shift = nullptr;
    // row-list path buffers
    DevBuf<uint32_t> rl_counts;   // [2][n_groups * n_ranges + 1]: counts, then their exclusive scan
    DevBuf<uint8_t> rl_temp;
    DevBuf<int64_t> rl_tables;    // offsets[257] + chunk_begin[max] + chunk_end[max]
    DevBuf<int32_t> rl_chunk_group;  // [max] + n_chunks scalar at the end
    DevBuf<RowOps> rl_ops;        // [n]
    DevBuf<int32_t> twin_map;     // feature held at each byte position of a twin row
    DevBuf<int32_t> twin_pos;     // feature slot -> byte position in a twin row (-1: a uint16 column)
    std::vector<int32_t> twin_pos_host;  // the same on the host (empty: no twin yet)
    DevBuf<int32_t> rt_tiles;     // row-list path: tile_feats [n_rt][32*fpl] then tile_slot0 [n_rt]
    DevBuf<uint32_t> rt_fmask;    // [n_slots][n_rt][fpl]
    DevBuf<int32_t> wt_tiles;     // wide tiles: tile_feats [n_wt][8] then first wide index [n_wt]
    DevBuf<uint32_t> wt_fmask;    // [n_slots][n_wt]
    bool enable_row_lists = true;
    // forest passes (RF): all trees of a pass in one set of launches (plan_forest / run_forest)
    bool enable_forest = true;
    int sm_count = 148;
    bool forest_pass = false;       // this pass's statistics are planned as a forest pass (layout.flat)
    struct ForestBatch {            // trees whose pair list fits the operand buffer and the key limit
        ForestListArgs list{};
        int slim = 0, fpl = 1;
        HistClassArgs cls{};
        HistClassArgs cls_wide{};    // tiles of 8 uint16 features over the same list
        HistRowsArgs rows{};
        bool has_narrow = false, has_wide = false;
        int grid = 0;
        size_t smem = 0, smem_wide = 0;
    };
    std::vector<ForestBatch> forest_batches;
    DevBuf<uint8_t> forest_ops;     // ClassOps or RowOps of the active (row, tree) pairs of one batch
    DevBuf<uint32_t> forest_counts; // [2][n_keys * n_ranges + 1]
    DevBuf<int64_t> forest_pair_off;
    size_t forest_scan_items = 0, forest_scan_temp = 0;
    size_t forest_ops_budget = 0;   // bytes the operand buffer may take (half of the free memory at first use)
    // the flat feature lists of the pass on the device (shared by the pair_off and the split task kernels)
    const int32_t *d_flat_feats = nullptr, *d_flat_begin = nullptr;
    const int64_t* d_flat_node_off = nullptr;
    int flat_max_list = 0, flat_max_cat_bins = 0;
    bool flat_has_cat = false;
    std::vector<uint8_t> flat_used;  // feature slot appears in some caller list of the pass
    size_t flat_n_cat = 0;           // categorical entries over all caller lists
    int64_t* d_int_feat_off = nullptr;  // [F] offset of a feature's block inside a node block that holds every valid feature, -1
    std::vector<uint32_t> col_flags;  // per sample-weight column: bit0 non-integer products, bit1 a weight != 1
    size_t scan_temp_bytes = 0;
    bool split_planned = false;
    bool acc_valid = false;     // d_acc holds the error sums of the current Data.predict (GBT, inside a tree)
    bool acc_fresh = false, scales_fresh = false;  // recomputed by the last sb_compute_begin (=> need an all-reduce)
    bool scales_valid = false;  // d_maxbits holds the maxima of the current tree's contributions
    // Bin counts of a tree root: they depend on the bins and the sample weights only, so consecutive trees share them.
    DevBuf<int64_t> root_cnt;
    bool root_cnt_valid = false, enable_root_cache = true;
    uint64_t weights_version = 0, root_cnt_version = 0;
    int root_cnt_wcol = -1;
    int64_t root_cnt_total = 0;
    std::vector<int> root_cnt_feats;
    bool weights_all_one = false;
    bool pending_route = false;  // compute_begin left the walk of the training rows to the row-list counting pass
    TreeView pending_tree{};
    bool cnt_int = false;  // every sig*weight product of the training rows is an integer in [0, 1024]

    int64_t launches_count = 0, h2d_bytes = 0, d2h_bytes = 0;

    // SB_TRACE=1 in the environment: host / device timeline of every pass, printed by sb_destroy (developer aid)
    struct PassTrace { double t0, t_begin, t_stats, t_enq, t_sync; cudaEvent_t e0, e1, c0, c1; int n_todo; double sub[6]; };
    bool trace = false;
    std::vector<PassTrace> passes_trace;
    // optional per-pass timing of the histogram kernel (CUDA events on the launching stream)
    bool profiling = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events;
    std::vector<int> prof_launches;
};

extern "C" {
static void p2p_teardown(sb_handle* h);
}

namespace {

double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

void dump_trace(sb_handle* h) {
    double last_sync = 0;
    fprintf(stderr, "SB_TRACE rank %d: pass todo | host gap before, begin, stats enqueue, finish enqueue, sync wait | wall, gpu busy (ms)\n", h->rank);
    for (size_t i = 0; i < h->passes_trace.size(); ++i) {
        auto& p = h->passes_trace[i];
        float gpu = 0, coll = 0;
        cudaEventElapsedTime(&gpu, p.e0, p.e1);
        if (cudaEventElapsedTime(&coll, p.c0, p.c1) != cudaSuccess) { coll = -1; cudaGetLastError(); }
        fprintf(stderr, "SB_TRACE %4zu %4d | %7.3f %7.3f %7.3f %7.3f %7.3f | %7.3f %7.3f\n", i, p.n_todo,
                last_sync > 0 ? p.t0 - last_sync : 0.0, p.t_begin - p.t0, p.t_stats - p.t_begin, p.t_enq - p.t_stats,
                p.t_sync - p.t_enq, p.t_sync - p.t0, (double)gpu);
        fprintf(stderr, "SB_TRACE      all-reduce (wait for the slowest rank + transfer) %.3f ms\n", (double)coll);
        fprintf(stderr, "SB_TRACE      begin parts: layout+memset %.3f  trees+switch kernels %.3f  scales %.3f  plan_hist %.3f  plan_split %.3f\n",
                p.sub[0], p.sub[1], p.sub[2], p.sub[3], p.sub[4]);
        last_sync = p.t_sync;
        cudaEventDestroy(p.e0);
        cudaEventDestroy(p.e1);
    }
    h->passes_trace.clear();
}

// stage `n` elements and let `buf` point at their place in the device arena
template <typename T>
cudaError_t stage_into(sb_handle* h, DevBuf<T>& buf, const T* src, size_t n) {
    T* d = nullptr;
    cudaError_t e = h->stage.put(src, n, &d, h->stream);
    if (e == cudaSuccess) {
        buf.alias(d);
        h->h2d_bytes += (int64_t)(n * sizeof(T));
    }
    return e;
}

bool is_variance(int imp) { return imp == SB_IMPURITY_VARIANCE || imp == SB_IMPURITY_FRIEDMANMSE; }

void free_rowset(RowSet& r) {
    if (r.d_col) cudaFree((void*)r.d_col);
    if (r.owns_inputs) {
        if (r.owned_bins) cudaFree(r.owned_bins);
        if (r.label) cudaFree(r.label);
        if (r.significance) cudaFree(r.significance);
        if (r.weights) cudaFree(r.weights);
    }
    if (r.output) cudaFree(r.output);
    if (r.predict) cudaFree(r.predict);
    if (r.node_id) cudaFree(r.node_id);
    if (r.bins_rm) cudaFree(r.bins_rm);
    if (r.bins16) cudaFree(r.bins16);
    if (r.bins16_rm) cudaFree(r.bins16_rm);
    r = RowSet();
}

void free_tree(DeviceTree& t) { t = DeviceTree(); }  // the arrays live in the per-pass staging arena

TreeView view_of(const DeviceTree& t) { return TreeView{t.feat, t.mask_off, t.predict, t.has_predict, t.masks, t.capacity}; }

int n_node_arrays(const sb_handle* h) { return h->cfg.is_gbdt ? 1 : std::max(1, h->cfg.tree_num); }

// Decides whether the count statistic can be accumulated as plain 32-bit integers (see hist.cu).  Synchronises;
// called at load time and whenever the caller injects new sample weights.
int refresh_cnt_int(sb_handle* h, int only_col = -1) {
    RowSet& r = h->train;
    uint32_t* flag_dev = reinterpret_cast<uint32_t*>(h->d_cnt_shift) + 1;  // second word of the same allocation
    const int W = r.weights ? h->cfg.weights_per_row : 1;
    // a column's verdict depends on that column and the significances alone: after a change of one column only that
    // one is looked at again (an RF draws its 50 bags one call at a time)
    if ((int)h->col_flags.size() != W || only_col < 0 || only_col >= W) {
        h->col_flags.assign((size_t)W, 0u);
        only_col = -1;
    }
    uint32_t flag = 0;
    for (int c = (only_col >= 0 ? only_col : 0); c < (only_col >= 0 ? only_col + 1 : W); ++c) {
        SB_CUDA(cudaMemsetAsync(flag_dev, 0, 4, h->stream));
        SB_CUDA(launch_check_int_weights(r.n, r.significance, r.weights ? r.weights + (int64_t)c * r.n : nullptr, flag_dev,
                                         h->stream));
        h->launches_count += 1;
        SB_CUDA(cudaMemcpyAsync(&flag, flag_dev, 4, cudaMemcpyDeviceToHost, h->stream));
        SB_CUDA(cudaStreamSynchronize(h->stream));
        h->col_flags[(size_t)c] = flag;
    }
    flag = 0;
    for (uint32_t f : h->col_flags) flag |= f;
    h->cnt_int = (flag & 1u) == 0;
    h->weights_all_one = (flag & 2u) == 0;  // then the weight arrays need not be read at all (x * 1.0f == x)
    h->weights_version += 1;  // every caller changed (or loaded) sample weights
    // the previous pass's accumulators were built with the old weights: a child must not be derived from them
    h->prev_valid = false;
    h->use_prev = false;
    return SB_OK;
}

// CUtensorMap over the uint8 bin matrix [F][pitch]: box = {128 rows, 32 feature slots}, 128-byte swizzle, zero fill
// outside [0, n) x [0, F).  cuTensorMapEncodeTiled is resolved through the runtime, so libcuda is not linked.
int make_tensor_map(sb_handle* h, RowSet& r) {
    typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        SB_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
        if (qres != cudaDriverEntryPointSuccess || !fn) return fail(SB_ERR_CUDA, "cuTensorMapEncodeTiled not available in this driver");
        encode = (encode_fn)fn;
    }
    (void)h;
    const cuuint64_t gdim[2] = {(cuuint64_t)std::max<int64_t>(r.n, 1), (cuuint64_t)r.n_features};
    const cuuint64_t gstride[1] = {(cuuint64_t)r.pitch};
    const cuuint32_t box[2] = {(cuuint32_t)HIST_T_ROWS, 32u};
    const cuuint32_t estride[2] = {1u, 1u};
    CUresult res = encode(reinterpret_cast<CUtensorMap*>(r.tensor_map), CU_TENSOR_MAP_DATA_TYPE_UINT8, 2,
                          const_cast<void*>(r.bins_base), gdim, gstride, box, estride, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (res != CUDA_SUCCESS) return fail(SB_ERR_CUDA, "cuTensorMapEncodeTiled failed with code " + std::to_string((int)res));
    return SB_OK;
}

// The work buffers of a pass come from the stream-ordered pool and grow level after level; a pool that has to map new
// memory does so at about 10 us per MB, in the middle of a pass.  Reserve the working set once, when the rows are
// loaded: allocate and free it, the pool keeps what it is given back (release threshold, sb_create).
int prewarm_pool(sb_handle* h) {
    // GBT: the statistics of one tree level.  RF: 10^4 node histograms at depth 10 (4 GB) + the operand list of a forest
    // pass, 8 bytes per (row, tree) pair
    const long long def_mb = h->cfg.is_gbdt ? 512 : 8192 + (long long)(((double)h->train.n * std::max(h->cfg.tree_num, 1) * 8.0) / (1 << 20));
    const long long want_mb = getenv("SB_POOL_RESERVE_MB") ? atoll(getenv("SB_POOL_RESERVE_MB")) : def_mb;
    if (want_mb <= 0) return SB_OK;
    // One block of the wanted size, allocated and freed: if the pool already holds such a block (an earlier handle of
    // this process reserved it) this costs nothing; a pool that is large but fragmented by what earlier handles
    // left behind grows here, once, and not level by level inside the passes.
    size_t free_b = 0, total_b = 0;
    SB_CUDA(cudaMemGetInfo(&free_b, &total_b));
    cudaMemPool_t pool = nullptr;
    uint64_t reserved = 0, used = 0;
    if (cudaDeviceGetDefaultMemPool(&pool, h->cfg.device) == cudaSuccess && pool) {
        if (cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &reserved) != cudaSuccess) reserved = 0;
        if (cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &used) != cudaSuccess) used = 0;
    }
    cudaGetLastError();
    const uint64_t idle = reserved > used ? reserved - used : 0;  // what the pool holds without anyone using it
    const size_t want = (size_t)std::min<uint64_t>((uint64_t)want_mb << 20, (uint64_t)free_b / 2 + idle);
    if (want < ((size_t)1 << 20)) return SB_OK;
    void* p = nullptr;
    if (cudaMallocAsync(&p, want, h->stream) == cudaSuccess) cudaFreeAsync(p, h->stream);
    else cudaGetLastError();  // not enough room: the buffers grow on demand as before
    return SB_OK;
}

// Allocates the derived per-row state and finishes a load (shared by the host and device loaders).
int finish_load(sb_handle* h, RowSet& r) {
    const int64_t n = r.n;
    r.h_col.resize((size_t)r.n_features);
    for (int f = 0; f < r.n_features; ++f) {
        if (h->wide_index[f] >= 0) r.h_col[f] = r.bins16 + (size_t)h->wide_index[f] * r.pitch16;
        else r.h_col[f] = static_cast<const uint8_t*>(r.bins_base) + (size_t)f * r.pitch;
    }
    int trc = make_tensor_map(h, r);
    if (trc != SB_OK) return trc;
    SB_CUDA(cudaMalloc(&r.d_col, sizeof(void*) * std::max(1, r.n_features)));
    SB_CUDA(cudaMemcpyAsync((void*)r.d_col, r.h_col.data(), sizeof(void*) * r.n_features, cudaMemcpyHostToDevice, h->stream));
    SB_CUDA(cudaMalloc(&r.output, sizeof(float) * std::max<int64_t>(n, 1)));
    SB_CUDA(cudaMalloc(&r.predict, sizeof(float) * std::max<int64_t>(n, 1)));
    r.n_node_arrays = n_node_arrays(h);
    SB_CUDA(cudaMalloc(&r.node_id, sizeof(int32_t) * std::max<int64_t>(n * r.n_node_arrays, 1)));
    // output = predict = label (dt/DTWorker.java:1264-1273); every row starts at the root
    SB_CUDA(launch_copy_f32(r.output, r.label, n, h->stream));
    SB_CUDA(launch_copy_f32(r.predict, r.label, n, h->stream));
    SB_CUDA(launch_fill_i32(r.node_id, n * r.n_node_arrays, SB_ROOT_INDEX, h->stream));
    // bin indices must address the statistics arrays (the reference would throw ArrayIndexOutOfBounds)
    SB_CUDA(cudaMemsetAsync(h->d_error_flag, 0, 4, h->stream));
    SB_CUDA(launch_check_bins(r.d_col, h->d_feat_bins, r.n_features, n, h->d_error_flag, h->stream));
    h->launches_count += 4;
    uint32_t flag = 0;
    SB_CUDA(cudaMemcpyAsync(&flag, h->d_error_flag, 4, cudaMemcpyDeviceToHost, h->stream));
    SB_CUDA(cudaStreamSynchronize(h->stream));
    if (flag & 2u) return fail(SB_ERR_RANGE, "a bin index is outside [0, n_bins) of its feature");
    r.loaded = true;
    if (&r == &h->train) {
        const int prc = prewarm_pool(h);
        if (prc != SB_OK) return prc;
        return refresh_cnt_int(h);
    }
    return SB_OK;
}

int upload_tree(sb_handle* h, const sb_tree& t, DeviceTree& d) {
    int max_id = 1;
    for (int i = 0; i < t.n_nodes; ++i) {
        if (t.nodes[i].id <= 0) return fail(SB_ERR_INVALID, "tree node with non-positive id");
        max_id = std::max(max_id, t.nodes[i].id);
    }
    if (max_id >= (1 << 24)) return fail(SB_ERR_UNSUPPORTED, "node id beyond 2^24 (MaxDepth <= 20 in the reference)");
    const int cap = max_id + 1;
    // the arrays of a tree are rebuilt from the message every pass: they live in the staging arena of the pass
    int32_t *feat = nullptr, *moff = nullptr;
    double* pred = nullptr;
    uint8_t *hasp = nullptr, *masks = nullptr;
    SB_CUDA(h->stage.take((size_t)cap, &feat, &d.feat, h->stream));
    SB_CUDA(h->stage.take((size_t)cap, &moff, &d.mask_off, h->stream));
    SB_CUDA(h->stage.take((size_t)cap, &pred, &d.predict, h->stream));
    SB_CUDA(h->stage.take((size_t)cap, &hasp, &d.has_predict, h->stream));
    SB_CUDA(h->stage.take((size_t)std::max(t.mask_pool_bytes, 1), &masks, &d.masks, h->stream));
    d.capacity = cap;
    d.mask_capacity = t.mask_pool_bytes;
    for (int i = 0; i < cap; ++i) { feat[i] = -1; moff[i] = 0; pred[i] = 0.0; hasp[i] = 0; }
    for (int i = 0; i < t.n_nodes; ++i) {
        const sb_node& nd = t.nodes[i];
        const bool routes = nd.feature >= 0 && (nd.left >= 0 || nd.right >= 0);  // dt/DTWorker.java:1041
        if (routes) {
            if (nd.feature >= h->cfg.n_features) return fail(SB_ERR_INVALID, "split feature slot out of range");
            const int need = h->n_bins[nd.feature] / 8 + 1;
            if (nd.left_mask_bytes < need || nd.left_mask_offset < 0 || nd.left_mask_offset + nd.left_mask_bytes > t.mask_pool_bytes)
                return fail(SB_ERR_INVALID, "left_mask of a split node is missing or too short");
            feat[nd.id] = nd.feature;
            moff[nd.id] = nd.left_mask_offset;
        }
        pred[nd.id] = nd.predict;
        hasp[nd.id] = nd.has_predict ? 1 : 0;
    }
    if (t.mask_pool_bytes > 0) memcpy(masks, t.mask_pool, (size_t)t.mask_pool_bytes);
    h->h2d_bytes += (int64_t)cap * 17 + t.mask_pool_bytes;
    d.walk = nullptr;
    if (!h->cfg.is_gbdt && h->enable_forest) {  // the same nodes as 32-byte records for forest_walk_kernel
        WalkNode *wn = nullptr, *d_wn = nullptr;
        SB_CUDA(h->stage.take((size_t)cap, &wn, &d_wn, h->stream));
        memset(wn, 0, (size_t)cap * sizeof(WalkNode));
        for (int i = 0; i < cap; ++i) { wn[i].feat = -1; wn[i].pos = -1; }
        for (int i = 0; i < t.n_nodes; ++i) {
            const sb_node& nd = t.nodes[i];
            WalkNode& w = wn[nd.id];
            if (feat[nd.id] >= 0) {
                w.feat = nd.feature;
                w.pos = h->twin_pos_host.empty() ? -1 : h->twin_pos_host[(size_t)nd.feature];
                w.mask_off = (uint32_t)nd.left_mask_offset;
                if (h->n_bins[nd.feature] <= 64) {
                    uint64_t m = 0;
                    memcpy(&m, t.mask_pool + nd.left_mask_offset, (size_t)std::min(8, nd.left_mask_bytes));
                    w.mask64 = m;
                    w.flags |= 2u;
                }
            }
            w.predict = (float)nd.predict;
            if (nd.has_predict) w.flags |= 1u;
        }
        d.walk = d_wn;
        h->h2d_bytes += (int64_t)cap * (int64_t)sizeof(WalkNode);
    }
    return SB_OK;
}

// dt/DTWorker.java:903-929 initTodoNodeStats: sizes and order of the statistics buffer.
int build_layout(const sb_handle* h, const sb_iter_in* in, TodoLayout& lay, bool flat = false, bool full = false, bool keep_lists = true) {
    // `lay` is recycled from an earlier pass: resize + assign keep the capacity of the inner vectors
    const int N = std::max(in->n_todo, 0);
    lay.nodes.resize((size_t)N);
    lay.flat = flat;
    lay.full = full;
    lay.flat_begin.clear();
    lay.flat_feats.clear();
    lay.caller_off.resize((size_t)N);
    int64_t off = 0;
    if (flat) {
        // the caller's lists as one flat array: a pass of an RF at depth 10 has 10^4 todo nodes with their own subsets
        const int F = h->cfg.n_features;
        const int n_all = (int)h->all_features.size();
        size_t total_feats = 0;
        for (int i = 0; i < N; ++i) total_feats += (size_t)(in->todo[i].n_features == 0 ? n_all : in->todo[i].n_features);
        if (total_feats > (size_t)0x7fffffff) return fail(SB_ERR_UNSUPPORTED, "more than 2^31 (todo node, feature) pairs in one pass");
        lay.flat_feats.resize(total_feats);
        lay.flat_begin.resize((size_t)N + 1);
        std::vector<int64_t> len_of((size_t)F);
        for (int f = 0; f < F; ++f) len_of[(size_t)f] = (int64_t)h->n_bins[(size_t)f] * h->stats_size;
        size_t at = 0;
        for (int i = 0; i < N; ++i) {
            const sb_todo_node& t = in->todo[i];
            TodoLayout::Node& nd = lay.nodes[(size_t)i];
            nd.key = t.key; nd.tree_id = t.tree_id; nd.node_id = t.node_id;
            lay.caller_off[(size_t)i] = off;
            lay.flat_begin[(size_t)i] = (int32_t)at;
            if (full) {  // internal layout: every valid feature, the same block size for every node
                nd.offset = (int64_t)i * h->all_feat_total;
                if (keep_lists) { nd.features = h->all_features; nd.feat_off = h->all_feat_off; }
                else { nd.features.clear(); nd.feat_off.clear(); }
            } else {
                nd.offset = off;
                nd.features.clear();
                nd.feat_off.clear();
            }
            if (t.n_features == 0) {
                if (n_all) memcpy(&lay.flat_feats[at], h->all_features.data(), (size_t)n_all * 4);
                at += (size_t)n_all;
                off += h->all_feat_total;
                continue;
            }
            int64_t l = 0;
            for (int k = 0; k < t.n_features; ++k) {
                const int f = t.features[k];
                if (f < 0 || f >= F) return fail(SB_ERR_INVALID, "todo node feature slot out of range");
                if (full && h->all_feature_pos[(size_t)f] < 0) return fail(SB_ERR_INVALID, "todo node lists a feature without bins to split");
                lay.flat_feats[at + (size_t)k] = f;
                l += len_of[(size_t)f];
            }
            at += (size_t)t.n_features;
            off += l;
        }
        lay.flat_begin[(size_t)N] = (int32_t)at;
        lay.caller_total = off;
        lay.total = full ? (int64_t)N * h->all_feat_total : off;
        return SB_OK;
    }
    for (int i = 0; i < N; ++i) {
        const sb_todo_node& t = in->todo[i];
        TodoLayout::Node& nd = lay.nodes[(size_t)i];
        nd.key = t.key; nd.tree_id = t.tree_id; nd.node_id = t.node_id; nd.offset = off;
        lay.caller_off[(size_t)i] = off;
        if (t.n_features == 0) {  // every valid feature: sizes and offsets are the same for all such nodes
            nd.features = h->all_features;
            nd.feat_off = h->all_feat_off;
            off += h->all_feat_total;
            continue;
        }
        nd.features.assign(t.features, t.features + t.n_features);
        int64_t l = 0;
        nd.feat_off.clear();
        for (int f : nd.features) {
            if (f < 0 || f >= h->cfg.n_features) return fail(SB_ERR_INVALID, "todo node feature slot out of range");
            nd.feat_off.push_back(l);
            l += (int64_t)h->n_bins[f] * h->stats_size;
        }
        off += l;
    }
    lay.total = off;
    lay.caller_total = off;
    return SB_OK;
}

// The caller's flat feature lists of the pass on the device (shared by the (slot, feature) offset kernel of a forest
// pass, the split task kernel and the gather of a full pass), and what the split step wants to know about them.
int stage_flat_lists(sb_handle* h) {
    const TodoLayout& lay = h->layout;
    const int N = (int)lay.nodes.size();
    const int F = h->cfg.n_features;
    h->d_flat_feats = h->d_flat_begin = nullptr;
    h->d_flat_node_off = nullptr;
    h->flat_max_list = 0;
    h->flat_has_cat = false;
    h->flat_max_cat_bins = 0;
    if (N == 0) return SB_OK;
    h->flat_used.assign((size_t)F, 0);
    h->flat_n_cat = 0;
    for (int32_t f : lay.flat_feats) {
        h->flat_used[(size_t)f] = 1;
        h->flat_n_cat += h->is_cat[(size_t)f] ? 1 : 0;
    }
    for (int i = 0; i < N; ++i) h->flat_max_list = std::max(h->flat_max_list, lay.flat_begin[(size_t)i + 1] - lay.flat_begin[(size_t)i]);
    for (int f = 0; f < F; ++f)
        if (h->flat_used[(size_t)f] && h->is_cat[(size_t)f]) { h->flat_has_cat = true; h->flat_max_cat_bins = std::max(h->flat_max_cat_bins, h->n_bins[(size_t)f]); }
    int32_t *hf = nullptr, *df = nullptr, *hb = nullptr, *db = nullptr;
    int64_t *ho = nullptr, *dof = nullptr;
    SB_CUDA(h->stage.take(std::max<size_t>(lay.flat_feats.size(), 1), &hf, &df, h->stream));
    SB_CUDA(h->stage.take((size_t)N + 1, &hb, &db, h->stream));
    SB_CUDA(h->stage.take((size_t)N, &ho, &dof, h->stream));
    if (!lay.flat_feats.empty()) memcpy(hf, lay.flat_feats.data(), lay.flat_feats.size() * 4);
    memcpy(hb, lay.flat_begin.data(), ((size_t)N + 1) * 4);
    memcpy(ho, lay.caller_off.data(), (size_t)N * 8);
    h->d_flat_feats = df; h->d_flat_begin = db; h->d_flat_node_off = dof;
    h->h2d_bytes += (int64_t)lay.flat_feats.size() * 4 + ((int64_t)N + 1) * 4 + (int64_t)N * 8;
    return SB_OK;
}

// RF passes whose statistics go through one list for all trees (plan_forest)
bool forest_lists(const sb_handle* h) {
    // uint16 columns: only the class kernels over 8-byte operands have a wide-tile variant
    const bool wide_ok = h->n_wide == 0 || (h->stat_kind == STAT_CLASS && h->cnt_int);
    return !h->cfg.is_gbdt && h->enable_forest && h->enable_row_lists && wide_ok && h->train.n < (int64_t)0xffffffff;
}

// Iteration order of a java.util.HashMap<Integer, double[]> created with (n, 1f) and filled in list order
// (dt/DTWorker.java:910, dt/DTWorkerParams.java:235), as walked at dt/DTMaster.java:324: buckets ascending,
// insertion order inside a bucket.  Returns rank[i] = visit position of list entry i.
std::vector<int> hashmap_visit_rank(const std::vector<int>& column_nums) {
    const int n = (int)column_nums.size();
    int cap = 1;
    while (cap < n) cap <<= 1;
    auto bucket = [cap](int k) {
        uint32_t x = (uint32_t)k;
        x ^= x >> 16;
        return (int)(x & (uint32_t)(cap - 1));
    };
    // a stable counting sort by bucket: first position of every bucket, then list order inside it
    static thread_local std::vector<int> first;
    first.assign((size_t)cap + 1, 0);
    for (int i = 0; i < n; ++i) first[(size_t)bucket(column_nums[i]) + 1] += 1;
    for (int b = 0; b < cap; ++b) first[(size_t)b + 1] += first[(size_t)b];
    std::vector<int> rank((size_t)n);
    for (int i = 0; i < n; ++i) rank[(size_t)i] = first[(size_t)bucket(column_nums[i])]++;
    return rank;
}

const sb_tree* find_tree(const sb_iter_in* in, int tree_id) {
    for (int i = 0; i < in->n_trees; ++i)
        if (in->trees[i].tree_id == tree_id && in->trees[i].n_nodes > 0) return &in->trees[i];
    return nullptr;
}

// Byte layout of a row of the row-major twin for `fpl` features per lane: full tiles of 32*fpl feature slots hold feature
// fpl*lane + k in natural order; the last, partly filled tile spreads its r features over the lanes first (kact =
// ceil(r / 32) features per lane), so the kernels issue kact instead of fpl rounds of atomics per row for it.
void twin_positions(const sb_handle* h, int fpl, std::vector<int32_t>& pos_of, int& twin_pad) {
    const int F = h->cfg.n_features;
    pos_of.assign((size_t)F, -1);
    const int tw = 32 * fpl;
    for (int f0 = 0; f0 < F; f0 += tw) {
        const int r = std::min(tw, F - f0), kact = (r + 31) / 32;
        for (int j = 0; j < r; ++j) pos_of[f0 + j] = f0 + fpl * (j / kact) + (j % kact);
    }
    int max_pos = 0;
    for (int f = 0; f < F; ++f) max_pos = std::max(max_pos, pos_of[f]);
    // Row pitch: DRAM is fetched in whole 128-byte lines, so a tile of a row must not straddle lines -- rows
    // wider than one line are padded to a multiple of 128 bytes (a 224-byte pitch read 450 bytes per row).
    twin_pad = (max_pos / 32 + 1) * 32;
    if (twin_pad > 128) twin_pad = (twin_pad + 127) / 128 * 128;
}

// The row-major twin of the training rows' bin matrix for this layout, built on first use (and again when the
// statistics layout -- integer <-> real counts -- changes the features per lane).
int ensure_twin(sb_handle* h, int fpl, const std::vector<int32_t>& pos_of, int twin_pad) {
    RowSet& tr = h->train;
    const int F = h->cfg.n_features;
    const int64_t n = tr.n;
    if (tr.bins_rm && tr.rm_fpl != fpl) {
        SB_CUDA(cudaStreamSynchronize(h->stream));
        cudaFree(tr.bins_rm);
        tr.bins_rm = nullptr;
        h->twin_pos_host.clear();
    }
    if (tr.bins_rm) return SB_OK;
    // 32-byte row pitch: a lane group that reaches past F reads into the next row (masked out in
    // the kernel); the allocation is padded so the last row's over-read stays inside it
    tr.fpad = twin_pad;
    tr.rm_fpl = fpl;
    std::vector<int32_t> feat_of_pos((size_t)twin_pad, -1);
    for (int f = 0; f < F; ++f)
        if (h->wide_index[f] < 0) feat_of_pos[pos_of[f]] = f;  // a uint16 column has no bytes to show: its position stays zero
    SB_CUDA(h->twin_map.reserve(h->stream, (size_t)twin_pad));
    SB_CUDA(cudaMemcpyAsync(h->twin_map.p, feat_of_pos.data(), (size_t)twin_pad * 4, cudaMemcpyHostToDevice, h->stream));
    // feature slot -> byte position, for kernels that look single bins up in the twin (forest_walk_kernel); uint16 columns are not in it
    std::vector<int32_t> pos_dev(pos_of);
    for (int f = 0; f < F; ++f)
        if (h->wide_index[f] >= 0) pos_dev[f] = -1;
    SB_CUDA(h->twin_pos.reserve(h->stream, (size_t)F));
    SB_CUDA(cudaMemcpyAsync(h->twin_pos.p, pos_dev.data(), (size_t)F * 4, cudaMemcpyHostToDevice, h->stream));
    h->twin_pos_host = pos_dev;
    SB_CUDA(cudaStreamSynchronize(h->stream));  // the tables are locals
    SB_CUDA(cudaMalloc(&tr.bins_rm, (size_t)std::max<int64_t>(n, 1) * tr.fpad + 256));
    SB_CUDA(cudaMemsetAsync(tr.bins_rm + (size_t)std::max<int64_t>(n, 1) * tr.fpad, 0, 256, h->stream));  // what a lane group reads past the last row
    SB_CUDA(launch_transpose_bins(static_cast<const uint8_t*>(tr.bins_base), tr.pitch, F, h->twin_map.p, n, tr.bins_rm, tr.fpad, h->stream));
    h->launches_count += 1;
    return SB_OK;
}

// row-major twin of the uint16 columns [n][wpad], built on first use
int ensure_twin16(sb_handle* h) {
    RowSet& tr = h->train;
    if (tr.bins16_rm || h->n_wide == 0) return SB_OK;
    tr.wpad = ((h->n_wide + 7) / 8) * 8;
    SB_CUDA(cudaMalloc(&tr.bins16_rm, ((size_t)std::max<int64_t>(tr.n, 1) * tr.wpad + 64) * 2));
    SB_CUDA(launch_transpose_bins16(tr.bins16, tr.pitch16, h->n_wide, tr.n, tr.bins16_rm, tr.wpad, h->stream));
    h->launches_count += 1;
    return SB_OK;
}

// Tiles of 32*fpl consecutive feature slots of the twin that hold a used feature: rt_feats [n_rt][32*fpl] (entry k*32+lane =
// feature at byte slot0 + fpl*lane + k, or -1), rt_slot0, rt_kact (features per lane actually populated); rt_of[f] = tile
void build_row_tiles(const sb_handle* h, const std::vector<uint8_t>& used, int fpl, const std::vector<int32_t>& pos_of,
                     std::vector<int32_t>& rt_feats, std::vector<int32_t>& rt_slot0, std::vector<int32_t>& rt_kact,
                     std::vector<int>& rt_of) {
    const int F = h->cfg.n_features, tw = 32 * fpl;
    rt_feats.clear(); rt_slot0.clear(); rt_kact.clear();
    rt_of.assign((size_t)F, -1);
    for (int f0 = 0; f0 < F; f0 += tw) {
        bool any = false;
        for (int f = f0; f < std::min(F, f0 + tw); ++f) any |= used[f] != 0;
        if (!any) continue;
        const int t = (int)rt_slot0.size();
        rt_slot0.push_back(f0);
        const size_t base = rt_feats.size();
        rt_feats.resize(base + tw, -1);
        int kact = 1;
        for (int f = f0; f < std::min(F, f0 + tw); ++f) {
            if (!used[f]) continue;
            const int rel = pos_of[f] - f0;
            rt_feats[base + (size_t)(rel % fpl) * 32 + rel / fpl] = f;
            rt_of[f] = t;
            kact = std::max(kact, rel % fpl + 1);
        }
        rt_kact.push_back(kact);
    }
}

int plan_histograms(sb_handle* h, const sb_iter_in* in) {
    h->launches.clear();
    const TodoLayout& lay = h->layout;
    const int F = h->cfg.n_features;
    if (lay.nodes.empty() || h->train.n == 0) return SB_OK;

    // Slots: one per todo node -- or one per SIBLING PAIR whose parent's merged accumulators are still on the device
    // (the previous pass): then only the smaller child is accumulated and the other is parent - child, which is
    // exact in the integer accumulators (same scales within a tree).  `order[s]` is the todo index a slot stands for
    // (child A of a pair); slots are grouped by tree (one launch per tree: a row has one node id per tree).
    const int S_all = (int)lay.nodes.size();
    std::vector<int> partner((size_t)S_all, -1);  // todo index of child B for child A of a pair; -2 for child B
    h->pairs.clear();
    h->pair_max_len = 0;
    std::vector<int> pair_parent_prev;            // per pair: index into prev_layout
    std::map<std::pair<int, int>, int> prev_by_key;  // (tree id, heap id) -> index into prev_layout
    std::vector<int> msg_index;                      // heap id -> index into the message's node array, for one tree
    int msg_index_tree = -1;
    const sb_tree* tr_msg = nullptr;
    if (h->enable_subtraction && h->use_prev) {
        std::map<std::pair<int, int>, int> cur_idx;
        std::map<std::pair<int, int>, int>& prev_idx = prev_by_key;
        for (int i = 0; i < S_all; ++i) cur_idx[{lay.nodes[i].tree_id, lay.nodes[i].node_id}] = i;
        for (size_t i = 0; i < h->prev_layout.nodes.size(); ++i)
            prev_idx[{h->prev_layout.nodes[i].tree_id, h->prev_layout.nodes[i].node_id}] = (int)i;
        for (int i = 0; i < S_all; ++i) {
            const auto& a = lay.nodes[i];
            if (a.node_id < 2 || (a.node_id & 1)) continue;
            auto ib = cur_idx.find({a.tree_id, a.node_id + 1});
            auto ip = prev_idx.find({a.tree_id, a.node_id / 2});
            if (ib == cur_idx.end() || ip == prev_idx.end()) continue;
            const auto& b = lay.nodes[ib->second];
            const auto& par = h->prev_layout.nodes[ip->second];
            if (a.features != b.features || a.features != par.features) continue;
            if (find_tree(in, a.tree_id) == nullptr) continue;
            partner[i] = ib->second;
            partner[ib->second] = -2;
            pair_parent_prev.push_back(ip->second);
        }
    }
    std::vector<int> order;
    for (int i = 0; i < S_all; ++i)
        if (partner[i] != -2) order.push_back(i);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return lay.nodes[a].tree_id < lay.nodes[b].tree_id; });
    const int S = (int)order.size();  // shadows the node count from here on: number of slots

    // tiles of 32 CONSECUTIVE feature slots (one TMA box each); only tiles with a used feature are launched
    // `used`: narrow (uint8) features some todo node asks for; wide (uint16) ones go to their own 8-feature tiles
    std::vector<uint8_t> used((size_t)F, 0), used_wide((size_t)F, 0);
    bool any_wide = false;
    // common case (GBT with FeatureSubsetStrategy ALL): every todo node takes every valid feature -- one list to look at
    bool all_feats_nodes = true;  // (a full pass lays every node out with every valid feature, whatever its own list)
    for (int i = 0; i < in->n_todo && all_feats_nodes && !lay.full; ++i) all_feats_nodes = in->todo[i].n_features == 0;
    auto mark_used = [&](const std::vector<int>& feats) {
        for (int f : feats) {
            if (h->wide_index[f] >= 0) { used_wide[f] = 1; any_wide = true; }
            else used[f] = 1;
        }
    };
    if (all_feats_nodes) mark_used(h->all_features);
    else
        for (const auto& nd : lay.nodes) mark_used(nd.features);
    std::vector<int32_t> tile_feats, tile_slot0;
    std::vector<int> tile_of((size_t)F, -1), lane_of((size_t)F, -1);
    int bins_pad = 1;
    for (int f0 = 0; f0 < F; f0 += 32) {
        bool any = false;
        for (int f = f0; f < std::min(F, f0 + 32); ++f) any |= used[f] != 0;
        if (!any) continue;
        const int t = (int)tile_slot0.size();
        tile_slot0.push_back(f0);
        for (int l = 0; l < 32; ++l) {
            const int f = f0 + l;
            if (f < F && used[f]) {
                tile_feats.push_back(f);
                tile_of[f] = t;
                lane_of[f] = l;
                bins_pad = std::max(bins_pad, h->n_bins[f]);
            } else {
                tile_feats.push_back(-1);
            }
        }
    }
    const int n_tiles = (int)tile_slot0.size();
    if (n_tiles == 0 && !any_wide) return SB_OK;
    if (any_wide && !(h->enable_row_lists && h->train.n < (int64_t)0xffffffff))
        return fail(SB_ERR_UNSUPPORTED, "features wider than 256 bins need the row-list path (SB_FLAG_NO_ROW_LISTS is set)");

    std::vector<uint32_t> fmask((size_t)S * n_tiles, 0u);
    std::vector<int64_t> pair_off((size_t)S * F, -1);
    for (int s = 0; s < S; ++s) {
        const auto& nd = lay.nodes[order[s]];
        if (all_feats_nodes && s > 0) {  // same features as slot 0: its mask row, its offsets shifted to this node's block
            const auto& n0 = lay.nodes[order[0]];
            for (int t = 0; t < n_tiles; ++t) fmask[(size_t)s * n_tiles + t] = fmask[(size_t)t];
            const int64_t shift = nd.offset - n0.offset;
            for (int f = 0; f < F; ++f) {
                const int64_t o = pair_off[(size_t)f];
                pair_off[(size_t)s * F + f] = o < 0 ? -1 : o + shift;
            }
            continue;
        }
        for (size_t k = 0; k < nd.features.size(); ++k) {
            const int f = nd.features[k];
            if (tile_of[f] >= 0) fmask[(size_t)s * n_tiles + tile_of[f]] |= 1u << lane_of[f];
            pair_off[(size_t)s * F + f] = nd.offset + nd.feat_off[k];
        }
    }
    {
        const size_t n_tf = tile_feats.size();
        tile_feats.insert(tile_feats.end(), tile_slot0.begin(), tile_slot0.end());  // one table: feats, then slot0
        SB_CUDA(stage_into(h, h->tile_feats, tile_feats.data(), tile_feats.size()));
        tile_feats.resize(n_tf);
    }
    SB_CUDA(stage_into(h, h->slot_fmask, fmask.data(), fmask.size()));
    SB_CUDA(stage_into(h, h->pair_off, pair_off.data(), pair_off.size()));

    const int cnt_int = h->cnt_int ? 1 : 0;
    const int nw = hist_words_per_bin(h->stat_kind, cnt_int, h->stats_size);
    const int max_gs = hist_max_slots_per_block(bins_pad, nw);
    if (max_gs < 1) return fail(SB_ERR_UNSUPPORTED, "one node's histogram tile does not fit in shared memory");

    // ---- wide features (uint16 bins): tiles of 8, histogram [slot][bin][word][8] (hist_rows_wide_kernel)
    int n_wt = 0, bins_pad_w = 1, wide_max_gs = 0;
    if (any_wide) {
        std::vector<int> feat_of_wide((size_t)h->n_wide, -1);
        for (int f = 0; f < F; ++f)
            if (h->wide_index[f] >= 0) feat_of_wide[(size_t)h->wide_index[f]] = f;
        std::vector<int32_t> wt_feats, wt_slot0;
        std::vector<int> wt_of((size_t)F, -1), wj_of((size_t)F, -1);
        for (int w0 = 0; w0 < h->n_wide; w0 += 8) {
            bool any = false;
            for (int w = w0; w < std::min(h->n_wide, w0 + 8); ++w) any |= used_wide[feat_of_wide[(size_t)w]] != 0;
            if (!any) continue;
            const int t = (int)wt_slot0.size();
            wt_slot0.push_back(w0);
            for (int j = 0; j < 8; ++j) {
                const int w = w0 + j;
                const int f = w < h->n_wide ? feat_of_wide[(size_t)w] : -1;
                if (f >= 0 && used_wide[f]) {
                    wt_feats.push_back(f);
                    wt_of[f] = t;
                    wj_of[f] = j;
                    bins_pad_w = std::max(bins_pad_w, h->n_bins[f]);
                } else {
                    wt_feats.push_back(-1);
                }
            }
        }
        n_wt = (int)wt_slot0.size();
        std::vector<uint32_t> wfm((size_t)S * n_wt, 0u);
        for (int s = 0; s < S; ++s)
            for (int f : lay.nodes[order[s]].features)
                if (wt_of[f] >= 0) wfm[(size_t)s * n_wt + wt_of[f]] |= 1u << wj_of[f];
        wide_max_gs = hist_rows_wide_max_slots_per_block(bins_pad_w, nw);
        if (wide_max_gs < 1) return fail(SB_ERR_UNSUPPORTED, "a tile of 8 wide features does not fit in shared memory");
        {
            std::vector<int32_t> both(wt_feats);
            both.insert(both.end(), wt_slot0.begin(), wt_slot0.end());
            SB_CUDA(stage_into(h, h->wt_tiles, both.data(), both.size()));
        }
        SB_CUDA(stage_into(h, h->wt_fmask, wfm.data(), wfm.size()));
    }

    // ---- row-list path tables: tiles of 32*fpl consecutive feature slots, fpl = features per lane
    int fpl = 0, rows_max_gs = 0, n_rt = 0, allf = 1, allf_kernel = 1;
    const bool rows_enabled = h->enable_row_lists && h->train.n < (int64_t)0xffffffff;
    if (rows_enabled && n_tiles > 0) {
        for (int f : {4, 2, 1}) {
            rows_max_gs = hist_rows_max_slots_per_block(bins_pad, nw, f);
            if (rows_max_gs >= 1) { fpl = f; break; }
        }
    }
    std::vector<int32_t> pos_of((size_t)F, -1);
    int twin_pad = 32;
    if (fpl > 0) twin_positions(h, fpl, pos_of, twin_pad);
    if (fpl > 0) {
        const int tw = 32 * fpl;
        std::vector<int32_t> rt_feats, rt_slot0, rt_kact;
        std::vector<int> rt_of;
        build_row_tiles(h, used, fpl, pos_of, rt_feats, rt_slot0, rt_kact, rt_of);
        n_rt = (int)rt_slot0.size();
        std::vector<uint32_t> rfm((size_t)S * n_rt * fpl, 0u);
        for (int s = 0; s < S; ++s) {
            const auto& nd = lay.nodes[order[s]];
            if (all_feats_nodes && s > 0) {
                memcpy(&rfm[(size_t)s * n_rt * fpl], rfm.data(), (size_t)n_rt * fpl * 4);
                continue;
            }
            for (int f : nd.features) {
                const int t = rt_of[f];
                if (t < 0) continue;  // wide feature
                const int rel = pos_of[f] - rt_slot0[t];
                rfm[((size_t)s * n_rt + t) * fpl + rel % fpl] |= 1u << (rel / fpl);
            }
        }
        // allf: every slot takes every feature that lives in a launched tile
        std::vector<uint32_t> full((size_t)n_rt * fpl, 0u);
        for (int t = 0; t < n_rt; ++t)
            for (int f = rt_slot0[t]; f < std::min(F, rt_slot0[t] + tw); ++f) {
                const int rel = pos_of[f] - rt_slot0[t];
                full[(size_t)t * fpl + rel % fpl] |= 1u << (rel / fpl);
            }
        for (int s = 0; s < S && allf; ++s)
            if (memcmp(&rfm[(size_t)s * n_rt * fpl], full.data(), full.size() * 4) != 0) allf = 0;
        // Nodes with feature lists of their own: the kernel's masked mode issues the atomics of an unlisted cell all the
        // same (redirected to a spare bin) and pays a mask lookup per row on top, so the all-features mode is the
        // cheaper one whenever it is safe: unlisted cells then land in columns of the block's histogram that are never
        // flushed (pair_off < 0).  Safe = no feature of a launched tile has more bins than the histogram was sized for.
        allf_kernel = allf;
        if (!allf) {
            static const bool force = getenv("SB_NO_ALLF_FORCE") == nullptr;
            int bins_in_tiles = 1;
            for (int t = 0; t < n_rt; ++t)
                for (int f = rt_slot0[t]; f < std::min(F, rt_slot0[t] + tw); ++f)
                    if (h->wide_index[(size_t)f] < 0) bins_in_tiles = std::max(bins_in_tiles, h->n_bins[(size_t)f]);
            if (force && bins_in_tiles <= bins_pad) allf_kernel = 1;
        }
        {
            std::vector<int32_t> all3(rt_feats);  // one table: feats, then slot0, then kact
            all3.insert(all3.end(), rt_slot0.begin(), rt_slot0.end());
            all3.insert(all3.end(), rt_kact.begin(), rt_kact.end());
            SB_CUDA(stage_into(h, h->rt_tiles, all3.data(), all3.size()));
        }
        SB_CUDA(stage_into(h, h->rt_fmask, rfm.data(), rfm.size()));
    }

    // per tree: slot table + launch
    size_t table_total = 0;
    std::vector<std::pair<int, int>> tree_ranges;  // [begin, end) in `order`
    for (int s = 0; s < S;) {
        int e = s;
        while (e < S && lay.nodes[order[e]].tree_id == lay.nodes[order[s]].tree_id) ++e;
        tree_ranges.push_back({s, e});
        s = e;
    }
    std::vector<int> table_size(tree_ranges.size());
    for (size_t t = 0; t < tree_ranges.size(); ++t) {
        int max_id = 1;
        for (int s = tree_ranges[t].first; s < tree_ranges[t].second; ++s)
            max_id = std::max(max_id, lay.nodes[order[s]].node_id + (partner[order[s]] >= 0 ? 1 : 0));
        table_size[t] = max_id + 1;
        table_total += (size_t)table_size[t];
    }
    int16_t* tables = nullptr;  // written in place in the staging arena
    {
        int16_t* d_tables = nullptr;
        SB_CUDA(h->stage.take(std::max<size_t>(table_total, 1), &tables, &d_tables, h->stream));
        h->slot_of_node.alias(d_tables);
        for (size_t i = 0; i < table_total; ++i) tables[i] = (int16_t)-1;
        h->h2d_bytes += (int64_t)table_total * 2;
    }
    size_t toff = 0;
    const int64_t n = h->train.n;
    const int W = h->cfg.weights_per_row;
    for (size_t t = 0; t < tree_ranges.size(); ++t) {
        const int s0 = tree_ranges[t].first, s1 = tree_ranges[t].second;
        const int tree_id = lay.nodes[order[s0]].tree_id;
        const int ns = s1 - s0;
        if (ns > 32767) return fail(SB_ERR_UNSUPPORTED, "more than 32767 todo nodes in one tree");
        for (int s = s0; s < s1; ++s) {
            const auto& na = lay.nodes[order[s]];
            const int nid = na.node_id;
            if (nid <= 0) return fail(SB_ERR_INVALID, "todo node with non-positive id");
            tables[toff + nid] = (int16_t)(s - s0);
            if (partner[order[s]] >= 0) {
                // both children map to the pair's slot for now; pair_decide_kernel removes the larger one
                const auto& nb = lay.nodes[partner[order[s]]];
                tables[toff + nb.node_id] = (int16_t)(s - s0);
                // the parent: its split feature, left mask and node block in the previous pass
                // the parent in the message (by heap id, indexed once per tree) and in the previous pass's layout
                if (msg_index_tree != na.tree_id) {
                    msg_index_tree = na.tree_id;
                    tr_msg = find_tree(in, na.tree_id);
                    msg_index.clear();
                    for (int q = 0; q < tr_msg->n_nodes; ++q) {
                        const int id = tr_msg->nodes[q].id;
                        if (id >= (int)msg_index.size()) msg_index.resize((size_t)id + 1, -1);
                        msg_index[(size_t)id] = q;
                    }
                }
                const sb_node* pn = (nid / 2 < (int)msg_index.size() && msg_index[(size_t)(nid / 2)] >= 0)
                                        ? &tr_msg->nodes[msg_index[(size_t)(nid / 2)]] : nullptr;
                const TodoLayout::Node* par = nullptr;
                {
                    auto ip = prev_by_key.find({na.tree_id, nid / 2});
                    if (ip != prev_by_key.end()) par = &h->prev_layout.nodes[(size_t)ip->second];
                }
                int kf = -1;
                if (pn && par) {
                    if (par->features.size() == h->all_features.size() && pn->feature >= 0 && pn->feature < F &&
                        h->all_feature_pos[(size_t)pn->feature] >= 0 && par->features[(size_t)h->all_feature_pos[(size_t)pn->feature]] == pn->feature)
                        kf = h->all_feature_pos[(size_t)pn->feature];  // the list is "every valid feature": position known
                    else
                        for (size_t k = 0; k < par->features.size(); ++k)
                            if (par->features[k] == pn->feature) kf = (int)k;
                }
                if (!pn || !par || kf < 0 || pn->feature < 0) return fail(SB_ERR_STATE, "histogram subtraction: parent split not found");
                PairDesc d{};
                d.parent_split_block = par->offset + par->feat_off[(size_t)kf];
                d.parent_off = par->offset;
                d.off_a = na.offset;
                d.off_b = nb.offset;
                int64_t len = 0;
                for (int f : na.features) len += (int64_t)h->n_bins[f] * h->stats_size;
                d.len = len;
                d.left_mask = h->trees[na.tree_id].masks + pn->left_mask_offset;
                d.n_bins = h->n_bins[pn->feature];
                d.node_a = nid;
                d.node_b = nb.node_id;
                d.slot = s;
                d.table_off = (int32_t)toff;
                h->pairs.push_back(d);
                h->pair_max_len = std::max(h->pair_max_len, len);
            }
        }
        if (table_size[t] < 0) return fail(SB_ERR_STATE, "slot table");
        HistLaunch L{};
        L.kind = h->stat_kind;
        L.cnt_int = cnt_int;
        HistArgs& a = L.args;
        a.feat_bins = h->d_feat_bins;
        a.output = h->train.output;
        a.significance = h->train.significance;
        a.weights = (h->train.weights && !h->weights_all_one) ? h->train.weights + (int64_t)(tree_id % W) * n : nullptr;  // [treeId % len], :847
        const int arr = h->cfg.is_gbdt ? 0 : tree_id;
        if (arr < 0 || arr >= h->train.n_node_arrays) return fail(SB_ERR_INVALID, "todo node tree id out of range");
        a.node_id = h->train.node_id + (int64_t)arr * n;
        a.slot_of_node = h->slot_of_node.p + toff;
        a.slot_table_size = table_size[t];
        a.n_rows = n;
        a.n_features = F;
        a.tile_feats = h->tile_feats.p;
        a.tile_slot0 = h->tile_feats.p + (size_t)n_tiles * 32;
        a.slot_fmask = h->slot_fmask.p + (size_t)s0 * n_tiles;
        a.pair_off = h->pair_off.p + (size_t)s0 * F;
        a.n_tiles = n_tiles;
        a.gs = std::min(ns, max_gs);
        a.n_groups = (ns + a.gs - 1) / a.gs;
        a.n_slots = ns;
        a.bins_pad = bins_pad;
        if (fpl > 0 || any_wide) {
            // group the rows by slot group once per pass and walk the lists (rowlist.cu, hist_rows_kernel);
            // narrow and wide tiles share the lists, so they share the group size
            int rgs = ns;
            if (fpl > 0) rgs = std::min(rgs, rows_max_gs);
            if (any_wide) rgs = std::min(rgs, wide_max_gs);
            const int rgroups = (ns + rgs - 1) / rgs;
            if (rgroups > 254 && any_wide) return fail(SB_ERR_UNSUPPORTED, "more than 254 slot groups with wide features");
            if (rgroups <= 254) {
                RowSet& tr = h->train;
                if (any_wide) {
                    const int wrc = ensure_twin16(h);
                    if (wrc != SB_OK) return wrc;
                }
                if (fpl > 0) {
                    const int trc = ensure_twin(h, fpl, pos_of, twin_pad);
                    if (trc != SB_OK) return trc;
                }
                const int tiles_total = n_rt + n_wt;
                static const int waves = getenv("SB_WAVES") ? atoi(getenv("SB_WAVES")) : 4;  // measured on C2 / C3: 54.4 ms per tree against 56.0 (1), 56.6 (2), 55.1 (3), 54.7 (8)
                // every group ends with a partial chunk: leave room for those, or the launch spills into one more wave
                const int64_t blocks_target = std::max<int64_t>(148 * waves - (int64_t)tiles_total * std::min(rgroups, 74), 74);
                const int target_chunks = (int)std::max<int64_t>(1, blocks_target / tiles_total);
                const int64_t chunk_rows = MAX_ROWS_PER_FLUSH;  // upper bound; integer counts: 2^18 rows * 1024 < 2^32 as well
                const int max_chunks = (int)(n / chunk_rows + target_chunks + rgroups + 2 + 2 * 148);  // + up to two waves of rounding
                const size_t scan_items = (size_t)rgroups * (size_t)rowlist_n_ranges(n) + 1;
                SB_CUDA(h->rl_counts.reserve(h->stream, 2 * scan_items));
                // cub's size query asks the driver about the device every time: one query for the largest scan of this handle
                if (h->scan_temp_bytes == 0) h->scan_temp_bytes = rowlist_scan_temp_bytes((int64_t)254 * rowlist_n_ranges(n) + 1);
                const size_t temp_bytes = h->scan_temp_bytes;
                SB_CUDA(h->rl_temp.reserve(h->stream, temp_bytes + 16));
                SB_CUDA(h->rl_tables.reserve(h->stream, (size_t)257 + 2 * (size_t)max_chunks));
                SB_CUDA(h->rl_chunk_group.reserve(h->stream, (size_t)max_chunks + 1));
                SB_CUDA(h->rl_ops.reserve(h->stream, (size_t)n));
                L.use_rows = true;
                L.fpl = fpl;
                L.allf = allf_kernel;
                RowListArgs& la = L.list;
                la.n = n; la.node_id = a.node_id; la.output = a.output; la.significance = a.significance; la.weights = a.weights;
                la.slot_of_node = a.slot_of_node; la.slot_table_size = a.slot_table_size; la.gs = rgs; la.n_groups = rgroups;
                la.stats_size = h->stats_size; la.kind = h->stat_kind; la.cnt_int = cnt_int; la.scales = h->d_scales;
                la.counts = h->rl_counts.p; la.scanned = h->rl_counts.p + scan_items;
                la.temp = h->rl_temp.p; la.temp_bytes = temp_bytes; la.chunk_rows = chunk_rows; la.target_chunks = target_chunks; la.n_tiles = std::max(tiles_total, 1);
                la.identity = (ns == 1 && lay.nodes[order[s0]].node_id == SB_ROOT_INDEX && partner[order[s0]] < 0) ? 1 : 0;
                la.route = 0;
                if (h->pending_route && !la.identity && tree_ranges.size() == 1 && h->cfg.is_gbdt) {
                    la.route = 1;
                    la.tree = h->pending_tree;
                    la.cols = tr.d_col;
                    la.feat_bins = h->d_feat_bins;
                    la.node_id_rw = tr.node_id;
                    h->pending_route = false;
                }
                la.offsets = h->rl_tables.p; la.chunk_begin = h->rl_tables.p + 257; la.chunk_end = h->rl_tables.p + 257 + max_chunks;
                la.chunk_group = h->rl_chunk_group.p; la.n_chunks = h->rl_chunk_group.p + max_chunks; la.max_chunks = max_chunks;
                la.ops = h->rl_ops.p; la.error_flag = h->d_error_flag; la.rows_streamed = h->d_rows_streamed;
                HistRowsArgs& ra = L.rows;
                ra.rm_bins = tr.bins_rm; ra.fpad = tr.fpad; ra.ops = la.ops; ra.chunk_group = la.chunk_group;
                ra.chunk_begin = la.chunk_begin; ra.chunk_end = la.chunk_end; ra.n_chunks = la.n_chunks;
                ra.tile_feats = h->rt_tiles.p; ra.tile_slot0 = h->rt_tiles.p + (size_t)n_rt * 32 * fpl;
                ra.tile_kact = ra.tile_slot0 + n_rt;
                ra.slot_fmask = h->rt_fmask.p + (size_t)s0 * n_rt * fpl; ra.pair_off = a.pair_off; ra.n_tiles = n_rt; ra.gs = rgs;
                ra.n_slots = ns; ra.bins_pad = bins_pad; ra.n_features = F; ra.stats_size = h->stats_size;
                ra.cnt_shift = h->d_cnt_shift; ra.raw_lo = h->raw.p; ra.raw_hi = h->raw.p + lay.total;
                L.has_narrow = fpl > 0 && n_rt > 0;
                // a pass that is one tree root over every feature: its bin counts repeat from tree to tree
                if (h->enable_root_cache && la.identity && S == 1 && cnt_int == 1 && h->stat_kind == STAT_VARIANCE && allf &&
                    !any_wide && L.has_narrow) {
                    const int wcol = tree_id % W;
                    const auto& feats = lay.nodes[order[s0]].features;
                    L.root_bins = lay.total / h->stats_size;
                    if (h->root_cnt_valid && h->root_cnt_version == h->weights_version && h->root_cnt_wcol == wcol &&
                        h->root_cnt_total == lay.total && h->root_cnt_feats == feats) {
                        L.root_counts = 2;
                    } else {
                        SB_CUDA(h->root_cnt.reserve(h->stream, (size_t)(2 * L.root_bins)));
                        L.root_counts = 1;
                        L.root_wcol = wcol;
                        h->root_cnt_valid = false;  // valid again once the saving kernel of THIS pass is enqueued (run_histograms)
                    }
                }
                L.grid = max_chunks * std::max(n_rt, 1);
                L.smem = L.has_narrow ? hist_rows_smem_bytes(rgs, bins_pad, nw, fpl) : 0;
                if (any_wide) {
                    L.has_wide = true;
                    L.wide = ra;
                    L.wide.rm_bins = reinterpret_cast<const uint8_t*>(tr.bins16_rm);
                    L.wide.fpad = tr.wpad;
                    L.wide.tile_feats = h->wt_tiles.p;
                    L.wide.tile_slot0 = h->wt_tiles.p + (size_t)n_wt * 8;
                    L.wide.slot_fmask = h->wt_fmask.p + (size_t)s0 * n_wt;
                    L.wide.n_tiles = n_wt;
                    L.wide.bins_pad = bins_pad_w;
                    L.wide_grid = max_chunks * n_wt;
                    L.wide_smem = hist_rows_wide_smem_bytes(rgs, bins_pad_w, nw);
                }
            }
        }
        const int64_t tiles_rows = (n + HIST_T_ROWS - 1) / HIST_T_ROWS;
        const int64_t tile_items = std::max<int64_t>(1, (int64_t)n_tiles * a.n_groups);  // n_tiles == 0: wide features only
        int64_t n_chunks = (148 * 3 + tile_items - 1) / tile_items;
        n_chunks = std::max<int64_t>(1, std::min(n_chunks, tiles_rows));
        int64_t rpc = ((tiles_rows + n_chunks - 1) / n_chunks) * HIST_T_ROWS;
        rpc = std::min<int64_t>(rpc, MAX_ROWS_PER_FLUSH);  // integer counts: 2^18 rows * 1024 < 2^32 as well
        n_chunks = (n + rpc - 1) / rpc;
        a.rows_per_chunk = rpc;
        a.stats_size = h->stats_size;
        a.scales = h->d_scales;
        a.cnt_shift = h->d_cnt_shift;
        a.raw_lo = h->raw.p;
        a.raw_hi = h->raw.p + lay.total;
        a.error_flag = h->d_error_flag;
        if (!L.use_rows) {
            const int64_t grid = n_chunks * n_tiles * a.n_groups;
            if (grid > 0x7fffffff) return fail(SB_ERR_UNSUPPORTED, "histogram grid too large");
            L.grid = (int)grid;
            L.smem = hist_smem_bytes(a.gs, bins_pad, nw);
        }
        h->launches.push_back(L);
        toff += (size_t)table_size[t];
    }
    if (!h->pairs.empty()) {
        SB_CUDA(stage_into(h, h->d_pairs, h->pairs.data(), h->pairs.size()));
        SB_CUDA(h->d_small_is_b.reserve(h->stream, h->pairs.size()));
    }
    SB_CUDA(h->stage.flush(h->stream));  // every table of the pass so far, in one copy
    if (!h->pairs.empty()) {
        SB_CUDA(launch_pair_decide(h->d_pairs.p, (int)h->pairs.size(), h->raw_prev.p, h->raw_prev.p + h->prev_layout.total,
                                   h->stats_size, h->stat_kind, F, h->slot_of_node.p, h->pair_off.p, h->d_small_is_b.p, h->stream));
        h->launches_count += 1;
    }
    return SB_OK;
}

// RF: the statistics of ALL trees of the pass from one list (dt/DTWorker.java:810-826 visits every tree of a row in one
// pass over the rows).  Slots are numbered over the trees of the pass (todo nodes ordered by tree), slot groups are
// ranges of that numbering -- a group may span trees -- and every active (row, tree) pair becomes one operand record
// in its group's part of the list (rowlist.cu, forest_*).  One counting pass, one scan, one scatter and ONE statistics
// launch per batch of trees; a batch is as many trees as the operand buffer and the key limit take (all 50 trees of a
// 10M-row forest).  Host work per pass is linear in the flat feature lists: no per-node vectors, no per-slot masks
// (every slot accumulates every feature of a launched tile; columns outside a node's list are never flushed), and the
// (slot, feature) offset table is filled on the device from the lists.
int plan_forest(sb_handle* h, const sb_iter_in* in) {
    h->forest_batches.clear();
    h->launches.clear();
    h->pairs.clear();
    h->pair_max_len = 0;
    const TodoLayout& lay = h->layout;
    const int F = h->cfg.n_features;
    const int N = (int)lay.nodes.size();
    RowSet& tr = h->train;
    const int64_t n = tr.n;
    if (N == 0) return SB_OK;

    // ---- features in use (the caller's lists are on the device already: stage_flat_lists)
    // (a full pass accumulates and keeps every valid feature, whatever the lists say)
    std::vector<uint8_t> used_any(h->flat_used);
    if (lay.full)
        for (int f : h->all_features) used_any[(size_t)f] = 1;
    std::vector<uint8_t> used((size_t)F, 0), used_wide((size_t)F, 0);
    bool any_wide = false;
    for (int f = 0; f < F; ++f) {
        if (!used_any[(size_t)f]) continue;
        if (h->wide_index[(size_t)f] >= 0) { used_wide[(size_t)f] = 1; any_wide = true; }
        else used[(size_t)f] = 1;
    }
    if (n == 0) return SB_OK;  // a shard without rows still takes part in the merge

    // ---- slots: todo nodes ordered by tree (stable counting sort), one slot each -- or, on a full pass, one slot per
    // SIBLING PAIR whose parent's merged accumulators are still on the device (the previous pass was full as well): only
    // the smaller child is accumulated, the other one is parent - child (pair_decide / pair_fixup, exact in the integers)
    const int T_arr = tr.n_node_arrays;
    std::vector<int> first((size_t)T_arr + 1, 0);
    for (int i = 0; i < N; ++i) {
        const int t = lay.nodes[(size_t)i].tree_id;
        if (t < 0 || t >= T_arr) return fail(SB_ERR_INVALID, "todo node tree id out of range");
        if (lay.nodes[(size_t)i].node_id <= 0) return fail(SB_ERR_INVALID, "todo node with non-positive id");
        first[(size_t)t + 1] += 1;
    }
    for (int t = 0; t < T_arr; ++t) first[(size_t)t + 1] += first[(size_t)t];
    std::vector<int> by_tree((size_t)N);  // todo indices ordered by tree, todo order inside a tree
    {
        std::vector<int> at(first.begin(), first.end() - 1);
        for (int i = 0; i < N; ++i) by_tree[(size_t)at[(size_t)lay.nodes[(size_t)i].tree_id]++] = i;
    }
    h->pairs.clear();
    h->pair_max_len = 0;
    std::vector<int> partner((size_t)N, -1);     // todo index of child B for child A of a pair; -2 for child B
    std::vector<int> pair_parent_prev;           // per pair (in the order of their child A in by_tree): index into prev_layout
    if (lay.full && h->enable_subtraction && h->use_prev && h->prev_layout.full) {
        // (tree, heap id) -> index, for this pass's nodes and the previous pass's: flat tables, one row per tree
        int stride = 2;
        for (int i = 0; i < N; ++i) stride = std::max(stride, lay.nodes[(size_t)i].node_id + 2);
        for (const auto& pn : h->prev_layout.nodes) stride = std::max(stride, pn.node_id + 2);
        if ((int64_t)stride * T_arr <= ((int64_t)1 << 26)) {
            std::vector<int> cur_idx((size_t)stride * T_arr, -1), prev_idx((size_t)stride * T_arr, -1);
            for (int i = 0; i < N; ++i) cur_idx[(size_t)lay.nodes[(size_t)i].tree_id * stride + lay.nodes[(size_t)i].node_id] = i;
            for (size_t i = 0; i < h->prev_layout.nodes.size(); ++i) {
                const auto& pn = h->prev_layout.nodes[i];
                if (pn.tree_id >= 0 && pn.tree_id < T_arr && pn.node_id >= 0) prev_idx[(size_t)pn.tree_id * stride + pn.node_id] = (int)i;
            }
            int have_tree = -1;
            bool tree_ok = false;
            for (int q = 0; q < N; ++q) {
                const int i = by_tree[(size_t)q];
                const auto& a = lay.nodes[(size_t)i];
                if (a.node_id < 2 || (a.node_id & 1)) continue;
                const int ib = cur_idx[(size_t)a.tree_id * stride + a.node_id + 1];
                const int ip = prev_idx[(size_t)a.tree_id * stride + a.node_id / 2];
                if (ib < 0 || ip < 0) continue;
                if (have_tree != a.tree_id) { have_tree = a.tree_id; tree_ok = find_tree(in, a.tree_id) != nullptr; }
                if (!tree_ok) continue;
                partner[(size_t)i] = ib;
                partner[(size_t)ib] = -2;
                pair_parent_prev.push_back(ip);
            }
        }
    }
    std::vector<int> slot_of_todo((size_t)N, -1);
    int S = 0;
    for (int q = 0; q < N; ++q)
        if (partner[(size_t)by_tree[(size_t)q]] != -2) slot_of_todo[(size_t)by_tree[(size_t)q]] = S++;
    for (int i = 0; i < N; ++i)
        if (partner[(size_t)i] >= 0) slot_of_todo[(size_t)partner[(size_t)i]] = slot_of_todo[(size_t)i];
    int32_t *slot_node = nullptr, *d_slot_node = nullptr;   // the todo node a slot stands for (child A of a pair)
    SB_CUDA(h->stage.take((size_t)S, &slot_node, &d_slot_node, h->stream));
    for (int i = 0; i < N; ++i)
        if (partner[(size_t)i] != -2) slot_node[slot_of_todo[(size_t)i]] = i;
    // slot tables: heap id -> global slot, one table per tree with todo nodes
    struct TreeSlots { int tree_id, s0, s1, table_off, table_size; };
    std::vector<TreeSlots> tl;
    size_t table_total = 0;
    for (int t = 0; t < T_arr; ++t) {
        if (first[(size_t)t + 1] == first[(size_t)t]) continue;
        int max_id = 1, s0 = S, s1 = 0;
        for (int q = first[(size_t)t]; q < first[(size_t)t + 1]; ++q) {
            const int i = by_tree[(size_t)q];
            max_id = std::max(max_id, lay.nodes[(size_t)i].node_id);
            s0 = std::min(s0, slot_of_todo[(size_t)i]);
            s1 = std::max(s1, slot_of_todo[(size_t)i] + 1);
        }
        tl.push_back({t, s0, s1, (int)table_total, max_id + 1});
        table_total += (size_t)max_id + 1;
    }
    int32_t *tables = nullptr, *d_tables = nullptr;
    SB_CUDA(h->stage.take(table_total, &tables, &d_tables, h->stream));
    for (size_t i = 0; i < table_total; ++i) tables[i] = -1;
    // both children of a pair map to the pair's slot for now; pair_decide_kernel removes the larger one
    for (const TreeSlots& ts : tl)
        for (int q = first[(size_t)ts.tree_id]; q < first[(size_t)ts.tree_id + 1]; ++q) {
            const int i = by_tree[(size_t)q];
            tables[(size_t)ts.table_off + (size_t)lay.nodes[(size_t)i].node_id] = slot_of_todo[(size_t)i];
        }
    // the pairs: parent's split block and node block in the previous pass, the children's blocks in this one
    if (!pair_parent_prev.empty()) {
        std::vector<int> msg_index;
        int msg_index_tree = -1, li = -1;
        const sb_tree* tr_msg = nullptr;
        size_t pi = 0;
        for (int q = 0; q < N; ++q) {
            const int i = by_tree[(size_t)q];
            if (partner[(size_t)i] < 0) continue;
            const auto& na = lay.nodes[(size_t)i];
            const auto& nb = lay.nodes[(size_t)partner[(size_t)i]];
            const auto& par = h->prev_layout.nodes[(size_t)pair_parent_prev[pi++]];
            if (msg_index_tree != na.tree_id) {
                msg_index_tree = na.tree_id;
                for (li = 0; li < (int)tl.size() && tl[(size_t)li].tree_id != na.tree_id; ++li) {}
                tr_msg = find_tree(in, na.tree_id);
                msg_index.clear();
                for (int k = 0; k < tr_msg->n_nodes; ++k) {
                    const int id = tr_msg->nodes[k].id;
                    if (id >= (int)msg_index.size()) msg_index.resize((size_t)id + 1, -1);
                    msg_index[(size_t)id] = k;
                }
            }
            const int pid = na.node_id / 2;
            const sb_node* pn = (pid < (int)msg_index.size() && msg_index[(size_t)pid] >= 0) ? &tr_msg->nodes[msg_index[(size_t)pid]] : nullptr;
            if (!pn || pn->feature < 0 || pn->feature >= F || h->all_feature_pos[(size_t)pn->feature] < 0)
                return fail(SB_ERR_STATE, "histogram subtraction: parent split not found");
            PairDesc d{};
            d.parent_split_block = par.offset + h->all_feat_off[(size_t)h->all_feature_pos[(size_t)pn->feature]];
            d.parent_off = par.offset;
            d.off_a = na.offset;
            d.off_b = nb.offset;
            d.len = h->all_feat_total;
            d.left_mask = h->trees[na.tree_id].masks + pn->left_mask_offset;
            d.n_bins = h->n_bins[(size_t)pn->feature];
            d.node_a = na.node_id;
            d.node_b = nb.node_id;
            d.slot = slot_of_todo[(size_t)i];
            d.table_off = tl[(size_t)li].table_off;
            h->pairs.push_back(d);
        }
        h->pair_max_len = h->all_feat_total;
    }
    ListTree *lts = nullptr, *d_lts = nullptr;
    SB_CUDA(h->stage.take(tl.size(), &lts, &d_lts, h->stream));
    const int W = h->cfg.weights_per_row;
    for (size_t i = 0; i < tl.size(); ++i) {
        ListTree& lt = lts[i];
        lt.node_id = tr.node_id + (int64_t)tl[i].tree_id * n;
        lt.weights = (tr.weights && !h->weights_all_one) ? tr.weights + (int64_t)(tl[i].tree_id % W) * n : nullptr;  // [treeId % len], :847
        lt.slot_of_node = d_tables + tl[i].table_off;
        lt.slot_table_size = tl[i].table_size;
        lt.pad = 0;
    }
    h->h2d_bytes += (int64_t)S * 4 + (int64_t)table_total * 4 + (int64_t)tl.size() * (int64_t)sizeof(ListTree);

    // ---- feature tiles of the twin.  The tile geometry depends on the widest histogram, taken over ALL uint8 features so
    // that it (and with it the twin's byte layout) does not change from pass to pass with the sampled subsets.
    int bins_pad = 1, bins_pad_w = 1;
    for (int f = 0; f < F; ++f) {
        if (h->wide_index[(size_t)f] < 0) bins_pad = std::max(bins_pad, h->n_bins[(size_t)f]);
        else bins_pad_w = std::max(bins_pad_w, h->n_bins[(size_t)f]);
    }
    const int cnt_int = h->cnt_int ? 1 : 0;
    const int slim = (h->stat_kind == STAT_CLASS && cnt_int) ? 1 : 0;
    if (any_wide && !slim) return fail(SB_ERR_STATE, "internal: forest pass with uint16 columns needs integer class counts");
    const int nw = hist_words_per_bin(h->stat_kind, cnt_int, h->stats_size);
    int fpl = 0, gs = 0;
    for (int f : {4, 2, 1}) {
        gs = slim ? hist_rows_class_max_slots_per_block(bins_pad, h->stats_size, f) : hist_rows_max_slots_per_block(bins_pad, nw, f);
        if (gs >= 1) { fpl = f; break; }
    }
    if (fpl == 0) return fail(SB_ERR_UNSUPPORTED, "one node's histogram tile does not fit in shared memory");
    std::vector<int32_t> pos_of;
    int twin_pad = 32;
    twin_positions(h, fpl, pos_of, twin_pad);
    std::vector<int32_t> rt_feats, rt_slot0, rt_kact;
    std::vector<int> rt_of;
    build_row_tiles(h, used, fpl, pos_of, rt_feats, rt_slot0, rt_kact, rt_of);
    const int n_rt = (int)rt_slot0.size();
    // tiles of 8 uint16 features (hist_rows_class_wide_kernel)
    int n_wt = 0;
    if (any_wide) {
        std::vector<int> feat_of_wide((size_t)h->n_wide, -1);
        for (int f = 0; f < F; ++f)
            if (h->wide_index[(size_t)f] >= 0) feat_of_wide[(size_t)h->wide_index[(size_t)f]] = f;
        std::vector<int32_t> wt;  // [n_wt][8] features, then [n_wt] first wide index
        std::vector<int32_t> wt_slot0;
        for (int w0 = 0; w0 < h->n_wide; w0 += 8) {
            bool any = false;
            for (int w = w0; w < std::min(h->n_wide, w0 + 8); ++w) any |= used_wide[(size_t)feat_of_wide[(size_t)w]] != 0;
            if (!any) continue;
            wt_slot0.push_back(w0);
            for (int j = 0; j < 8; ++j) {
                const int w = w0 + j;
                const int f = w < h->n_wide ? feat_of_wide[(size_t)w] : -1;
                wt.push_back((f >= 0 && used_wide[(size_t)f]) ? f : -1);
            }
        }
        n_wt = (int)wt_slot0.size();
        wt.insert(wt.end(), wt_slot0.begin(), wt_slot0.end());
        SB_CUDA(stage_into(h, h->wt_tiles, wt.data(), wt.size()));
        const int gs_w = hist_rows_class_wide_max_slots_per_block(bins_pad_w, h->stats_size);
        if (gs_w < 1) return fail(SB_ERR_UNSUPPORTED, "a tile of 8 wide features does not fit in shared memory");
        gs = n_rt > 0 ? std::min(gs, gs_w) : gs_w;
        const int wrc = ensure_twin16(h);
        if (wrc != SB_OK) return wrc;
    }
    if (n_rt == 0 && n_wt == 0) return SB_OK;
    if (n_rt > 0) {
        std::vector<int32_t> all3(rt_feats);  // one table: feats, then slot0, then kact
        all3.insert(all3.end(), rt_slot0.begin(), rt_slot0.end());
        all3.insert(all3.end(), rt_kact.begin(), rt_kact.end());
        SB_CUDA(stage_into(h, h->rt_tiles, all3.data(), all3.size()));
        const int rc2 = ensure_twin(h, fpl, pos_of, twin_pad);
        if (rc2 != SB_OK) return rc2;
    }
    int rc = SB_OK;
    // more slots per block than the pass has: smaller histograms to clear and merge
    gs = std::min(gs, S);
    // A pair's operand word carries the byte offset of its (slot, class) words in the narrow tiles' histogram,
    // (slot * enc_slot_words + class) << enc_shift; a pass without narrow tiles encodes with the smallest geometry
    const int enc_fpl = n_rt > 0 ? fpl : 1;
    const uint32_t enc_shift = 7u + (enc_fpl == 4 ? 2u : (enc_fpl == 2 ? 1u : 0u));
    const uint32_t enc_slot_words = (uint32_t)(n_rt > 0 ? bins_pad : 1) * (uint32_t)h->stats_size;
    if (slim && ((uint64_t)gs * enc_slot_words << enc_shift) > ((uint64_t)1 << CLASSOPS_OFF_BITS))
        return fail(SB_ERR_STATE, "internal: forest pass operand offset out of range");
    const int n_tiles_all = n_rt + n_wt;

    // ---- (slot, feature) -> offset in the statistics, on the device
    SB_CUDA(h->forest_pair_off.reserve(h->stream, (size_t)S * F));
    SB_CUDA(h->stage.flush(h->stream));
    if (lay.full) {
        SB_CUDA(launch_fill_pair_off_full(d_slot_node, S, h->all_feat_total, h->d_int_feat_off, F, h->forest_pair_off.p, h->stream));
        h->launches_count += 1;
        if (!h->pairs.empty()) {
            SB_CUDA(stage_into(h, h->d_pairs, h->pairs.data(), h->pairs.size()));
            SB_CUDA(h->d_small_is_b.reserve(h->stream, h->pairs.size()));
            SB_CUDA(h->stage.flush(h->stream));
            SB_CUDA(launch_pair_decide32(h->d_pairs.p, (int)h->pairs.size(), h->raw_prev.p, h->raw_prev.p + h->prev_layout.total,
                                         h->stats_size, h->stat_kind, F, d_tables, h->forest_pair_off.p, h->d_small_is_b.p, h->stream));
            h->launches_count += 1;
        }
    } else {
        SB_CUDA(launch_build_pair_off(d_slot_node, S, h->d_flat_feats, h->d_flat_begin, h->d_flat_node_off, h->d_feat_bins,
                                      h->stats_size, F, h->forest_pair_off.p, h->stream));
        h->launches_count += 1;
    }

    // ---- batches of trees
    const size_t ops_bytes = slim ? sizeof(ClassOps) : sizeof(RowOps);
    if (h->forest_ops_budget == 0) {
        size_t free_b = 0, total_b = 0;
        SB_CUDA(cudaMemGetInfo(&free_b, &total_b));
        h->forest_ops_budget = std::max<size_t>(free_b / 2, (size_t)std::max<int64_t>(n, 1) * ops_bytes);
        if (const char* e = getenv("SB_FOREST_OPS_MB")) h->forest_ops_budget = std::max<size_t>((size_t)atoll(e) << 20, (size_t)n * ops_bytes);
    }
    const int64_t n_ranges = forest_n_ranges(n);
    const int64_t max_pairs = std::min<int64_t>((int64_t)0xfffffff0, (int64_t)(h->forest_ops_budget / ops_bytes));
    const int trees_cap = (int)std::min<int64_t>(std::max<int64_t>(1, max_pairs / n), 1 << 20);
    const int64_t keys_cap = std::min<int64_t>(FOREST_MAX_KEYS, std::max<int64_t>(1, ((int64_t)1 << 28) / n_ranges));
    const int64_t chunk_rows = MAX_ROWS_PER_FLUSH;  // integer counts: 2^18 pairs * 1024 < 2^32 as well
    const int target_chunks = std::max(1, 148 * 8 / std::max(n_rt, 1));
    static const int64_t slab_bytes = (getenv("SB_FOREST_SLAB_MB") ? std::max<int64_t>(1, atoll(getenv("SB_FOREST_SLAB_MB"))) : (int64_t)1 << 20) << 20;  // default: no slabs beyond what fills the GPU
    const int64_t j_l2 = ((int64_t)n * tr.fpad + slab_bytes - 1) / slab_bytes;
    size_t need_ops = 0, need_counts = 0, need_scan_items = 0;
    int need_chunks = 0;
    for (size_t i0 = 0; i0 < tl.size();) {
        size_t i1 = i0;
        int slots = 0;
        while (i1 < tl.size() && (int)(i1 - i0) < trees_cap) {
            const int more = tl[i1].s1 - tl[i1].s0;
            if (i1 > i0 && ((int64_t)slots + more + gs - 1) / gs > keys_cap) break;
            slots += more;
            ++i1;
        }
        const int n_keys = (slots + gs - 1) / gs;
        if (n_keys > keys_cap) return fail(SB_ERR_UNSUPPORTED, "one tree has more todo nodes than a forest pass can group");
        sb_handle::ForestBatch b{};
        b.slim = slim;
        b.fpl = fpl;
        const int batch_trees = (int)(i1 - i0);
        const int64_t pairs_max = n * batch_trees;
        const int64_t j_max = std::max<int64_t>(j_l2, (target_chunks + n_keys - 1) / n_keys) + 1;  // slabs: see forest_chunks_kernel
        const int64_t max_chunks64 = j_max * n_keys + pairs_max / chunk_rows + 2;
        if (max_chunks64 * std::max(n_tiles_all, 1) > 0x7fffffff) return fail(SB_ERR_UNSUPPORTED, "forest pass: chunk table too large");
        const int max_chunks = (int)max_chunks64;
        ForestListArgs& la = b.list;
        la.n = n; la.trees = d_lts + i0; la.n_trees = batch_trees; la.slot_base = tl[i0].s0; la.gs = gs; la.n_keys = n_keys;
        la.n_ranges = n_ranges; la.output = tr.output; la.significance = tr.significance; la.scales = h->d_scales;
        la.stats_size = h->stats_size; la.kind = h->stat_kind; la.cnt_int = cnt_int; la.slim = slim;
        la.class_stride = 1u << enc_shift;  // one 32-bit word per class and (k, lane)
        la.slot_stride = enc_slot_words << enc_shift;
        la.chunk_rows = chunk_rows; la.target_chunks = target_chunks; la.max_chunks = max_chunks;
        la.row_bytes = tr.fpad; la.slab_bytes = slab_bytes; la.stats_bytes = lay.total * 16;
        la.error_flag = h->d_error_flag; la.rows_streamed = h->d_rows_streamed;
        need_ops = std::max(need_ops, (size_t)pairs_max * ops_bytes + 16);  // a bulk copy may take one record past the last pair
        need_scan_items = std::max(need_scan_items, (size_t)n_keys * (size_t)n_ranges + 1);
        need_counts = std::max(need_counts, 2 * ((size_t)n_keys * (size_t)n_ranges + 1));
        need_chunks = std::max(need_chunks, max_chunks);
        b.has_narrow = n_rt > 0;
        b.has_wide = n_wt > 0;
        if (slim) {
            HistClassArgs& ca = b.cls;
            ca.rm_bins = tr.bins_rm; ca.fpad = (uint32_t)tr.fpad;
            ca.tile_feats = h->rt_tiles.p; ca.tile_slot0 = h->rt_tiles.p + (size_t)n_rt * 32 * fpl; ca.tile_kact = ca.tile_slot0 + n_rt;
            ca.pair_off = h->forest_pair_off.p + (size_t)la.slot_base * F;
            ca.n_tiles = n_rt; ca.gs = gs; ca.n_slots = slots; ca.bins_pad = bins_pad; ca.n_features = F; ca.stats_size = h->stats_size;
            ca.cnt_shift = h->d_cnt_shift; ca.raw_lo = h->raw.p; ca.raw_hi = h->raw.p + lay.total;
            ca.word_shift = enc_shift; ca.slot_words = enc_slot_words; ca.counter = 1;
            b.smem = hist_rows_class_smem_bytes(gs, bins_pad, h->stats_size, fpl);
            if (b.has_wide) {
                HistClassArgs& cw = b.cls_wide;
                cw = ca;
                cw.rm_bins = reinterpret_cast<const uint8_t*>(tr.bins16_rm); cw.fpad = (uint32_t)tr.wpad;
                cw.tile_feats = h->wt_tiles.p; cw.tile_slot0 = h->wt_tiles.p + (size_t)n_wt * 8; cw.tile_kact = nullptr;
                cw.n_tiles = n_wt; cw.bins_pad = bins_pad_w; cw.counter = 2;
                b.smem_wide = hist_rows_class_wide_smem_bytes(gs, bins_pad_w, h->stats_size);
            }
        } else {
            HistRowsArgs& ra = b.rows;
            ra.rm_bins = tr.bins_rm; ra.fpad = tr.fpad;
            ra.tile_feats = h->rt_tiles.p; ra.tile_slot0 = h->rt_tiles.p + (size_t)n_rt * 32 * fpl; ra.tile_kact = ra.tile_slot0 + n_rt;
            ra.slot_fmask = nullptr;  // every slot takes every feature of a launched tile
            ra.pair_off = h->forest_pair_off.p + (size_t)la.slot_base * F;
            ra.n_tiles = n_rt; ra.gs = gs; ra.n_slots = slots; ra.bins_pad = bins_pad; ra.n_features = F; ra.stats_size = h->stats_size;
            ra.cnt_shift = h->d_cnt_shift; ra.raw_lo = h->raw.p; ra.raw_hi = h->raw.p + lay.total;
            b.smem = hist_rows_smem_bytes(gs, bins_pad, nw, fpl);
        }
        b.grid = slim ? h->sm_count : max_chunks * n_rt;  // the class kernel is persistent: one block per SM
        h->forest_batches.push_back(b);
        i0 = i1;
    }
    // buffers shared by the batches (they run one after the other on the stream)
    SB_CUDA(h->forest_ops.reserve(h->stream, need_ops));
    SB_CUDA(h->forest_counts.reserve(h->stream, need_counts));
    if (need_scan_items > h->forest_scan_items) {
        h->forest_scan_items = need_scan_items * 2;
        h->forest_scan_temp = rowlist_scan_temp_bytes((int64_t)h->forest_scan_items);
    }
    SB_CUDA(h->rl_temp.reserve(h->stream, h->forest_scan_temp + 16));
    SB_CUDA(h->rl_tables.reserve(h->stream, 2 * (size_t)need_chunks));
    SB_CUDA(h->rl_chunk_group.reserve(h->stream, (size_t)need_chunks + 3));
    for (auto& b : h->forest_batches) {
        ForestListArgs& la = b.list;
        const size_t items = (size_t)la.n_keys * (size_t)n_ranges + 1;
        la.counts = h->forest_counts.p; la.scanned = h->forest_counts.p + items;
        la.temp = h->rl_temp.p; la.temp_bytes = h->forest_scan_temp;
        la.ops = h->forest_ops.p;
        la.chunk_begin = h->rl_tables.p; la.chunk_end = h->rl_tables.p + need_chunks;
        la.chunk_group = h->rl_chunk_group.p; la.n_chunks = h->rl_chunk_group.p + need_chunks;
        if (b.slim) {
            for (HistClassArgs* ca : {&b.cls, &b.cls_wide}) {
                ca->ops = reinterpret_cast<const ClassOps*>(la.ops);
                ca->chunk_group = la.chunk_group; ca->chunk_begin = la.chunk_begin; ca->chunk_end = la.chunk_end; ca->n_chunks = la.n_chunks;
            }
        } else {
            b.rows.ops = reinterpret_cast<const RowOps*>(la.ops);
            b.rows.chunk_group = la.chunk_group; b.rows.chunk_begin = la.chunk_begin; b.rows.chunk_end = la.chunk_end; b.rows.n_chunks = la.n_chunks;
        }
    }
    return SB_OK;
}

int run_forest(sb_handle* h) {
    for (const auto& b : h->forest_batches) {
        SB_CUDA(launch_forest_lists(b.list, h->stream));
        h->launches_count += 5;  // count, scan (2 kernels), chunk table, scatter
        if (b.slim) {
            if (b.has_narrow) {
                SB_CUDA(hist_rows_class_launch(b.fpl, b.cls, b.grid, b.smem, h->stream));
                h->launches_count += 1;
            }
            if (b.has_wide) {
                SB_CUDA(hist_rows_class_wide_launch(b.cls_wide, b.grid, b.smem_wide, h->stream));
                h->launches_count += 1;
            }
        } else {
            SB_CUDA(hist_rows_launch(b.list.kind, b.list.cnt_int, b.fpl, 1, b.rows, b.grid, b.smem, h->stream));
            h->launches_count += 1;
        }
    }
    return SB_OK;
}

int plan_split_tasks(sb_handle* h) {
    const TodoLayout& lay = h->layout;
    const size_t N = lay.nodes.size();
    h->has_cat_tasks = false;
    int max_cat_bins = 0;
    size_t n_cat = 0, n_cat_tmpl = 0;  // categorical tasks of the pass / of the one list all nodes share
    bool same_lists = N > 0 && !lay.flat;
    for (size_t n = 1; n < N && same_lists; ++n) same_lists = lay.nodes[n].features == lay.nodes[0].features;
    // tasks of one node: (feature, block offset relative to the node, bins, kind, HashMap visit rank), written at `out`
    std::vector<int> cols;
    auto node_tasks = [&](size_t n, int64_t base, SplitTask* out) {
        const auto& nd = lay.nodes[n];
        cols.clear();
        for (int f : nd.features) cols.push_back(h->column_num[f]);
        const std::vector<int> rank = hashmap_visit_rank(cols);
        for (size_t k = 0; k < nd.features.size(); ++k) {
            const int f = nd.features[k];
            SplitTask t;
            t.node = (int32_t)n;
            t.feature = f;
            t.offset = base + nd.feat_off[k];
            t.n_bins = h->n_bins[f];
            t.categorical = h->is_cat[f];
            t.visit_rank = rank[k];
            t.pad = 0;
            if (t.categorical) { h->has_cat_tasks = true; max_cat_bins = std::max(max_cat_bins, t.n_bins); n_cat_tmpl += 1; }
            out[k] = t;
        }
    };
    size_t n_tasks = 0;
    if (lay.flat) n_tasks = lay.flat_feats.size();
    else
        for (const auto& nd : lay.nodes) n_tasks += nd.features.size();
    h->n_tasks = (int)n_tasks;
    SB_CUDA(h->cand.reserve(h->stream, n_tasks));
    SB_CUDA(h->gain_out.reserve(h->stream, N));
    if (lay.flat && N > 0) {
        // forest pass: the lists are on the device already (plan_forest); build_tasks_kernel turns them into tasks
        if (h->flat_max_list > 8192) return fail(SB_ERR_UNSUPPORTED, "a todo node with more than 8192 features of its own");
        h->has_cat_tasks = h->flat_has_cat;
        max_cat_bins = h->flat_max_cat_bins;
        n_cat = h->flat_n_cat;
        h->node_task_begin.alias(const_cast<int32_t*>(h->d_flat_begin));
        SB_CUDA(h->tasks.reserve(h->stream, n_tasks + 1));
        SB_CUDA(launch_build_tasks(h->d_flat_feats, h->d_flat_begin, h->d_flat_node_off, (int)N, std::max(h->flat_max_list, 1), h->d_feat_bins,
                                   h->d_is_cat, h->d_column_num, h->stats_size, h->tasks.p, h->stream));
        h->launches_count += 1;
    } else if (same_lists && N > 1) {
        // one template row + the node offsets; expanded on the device
        const size_t fk = lay.nodes[0].features.size();
        SplitTask *tmpl = nullptr, *d_tmpl = nullptr;
        int64_t *node_off = nullptr, *d_node_off = nullptr;
        SB_CUDA(h->stage.take(std::max<size_t>(fk, 1), &tmpl, &d_tmpl, h->stream));
        SB_CUDA(h->stage.take(N, &node_off, &d_node_off, h->stream));
        node_tasks(0, 0, tmpl);
        n_cat = n_cat_tmpl * N;
        for (size_t n = 0; n < N; ++n) node_off[n] = lay.nodes[n].offset;
        SB_CUDA(h->stage.flush(h->stream));
        SB_CUDA(h->tasks.reserve(h->stream, n_tasks + 1));
        SB_CUDA(h->node_task_begin.reserve(h->stream, N + 1));
        SB_CUDA(launch_expand_tasks(d_tmpl, (int)fk, d_node_off, (int)N, h->tasks.p, h->node_task_begin.p, h->stream));
        h->launches_count += 1;
        h->h2d_bytes += fk * sizeof(SplitTask) + N * 8;
    } else {
        // every node has its own list: the lists cross PCIe (4 bytes per entry) and build_tasks_kernel turns them into
        // tasks -- block offsets and HashMap visit ranks -- on the device
        int32_t *feats = nullptr, *d_feats = nullptr, *begin = nullptr, *d_begin = nullptr;
        int64_t *node_off = nullptr, *d_node_off = nullptr;
        SB_CUDA(h->stage.take(std::max<size_t>(n_tasks, 1), &feats, &d_feats, h->stream));
        SB_CUDA(h->stage.take(N + 1, &begin, &d_begin, h->stream));
        SB_CUDA(h->stage.take(std::max<size_t>(N, 1), &node_off, &d_node_off, h->stream));
        size_t at = 0, max_list = 1;
        for (size_t n = 0; n < N; ++n) {
            const auto& nd = lay.nodes[n];
            begin[n] = (int32_t)at;
            node_off[n] = nd.offset;
            if (!nd.features.empty()) memcpy(feats + at, nd.features.data(), nd.features.size() * 4);
            at += nd.features.size();
            max_list = std::max(max_list, nd.features.size());
            if (h->any_categorical)
                for (int f : nd.features)
                    if (h->is_cat[(size_t)f]) { h->has_cat_tasks = true; max_cat_bins = std::max(max_cat_bins, h->n_bins[(size_t)f]); n_cat += 1; }
        }
        begin[N] = (int32_t)at;
        if (max_list > 8192) return fail(SB_ERR_UNSUPPORTED, "a todo node with more than 8192 features of its own");
        h->node_task_begin.alias(d_begin);
        SB_CUDA(h->stage.flush(h->stream));
        SB_CUDA(h->tasks.reserve(h->stream, n_tasks + 1));
        SB_CUDA(launch_build_tasks(d_feats, d_begin, d_node_off, (int)N, (int)max_list, h->d_feat_bins, h->d_is_cat, h->d_column_num,
                                   h->stats_size, h->tasks.p, h->stream));
        h->launches_count += 1;
        h->h2d_bytes += n_tasks * 4 + (N + 1) * 4 + N * 8;
    }
    if (h->has_cat_tasks) {
        SB_CUDA(h->cand_bits.reserve(h->stream, n_tasks * SB_MAX_CAT_BYTES));
        SB_CUDA(h->out_bits.reserve(h->stream, N * SB_MAX_CAT_BYTES));
        const size_t scratch = std::max<size_t>(n_cat, 1) * (size_t)max_cat_bins;  // per categorical task
        SB_CUDA(h->sort_keys.reserve(h->stream, scratch));
        SB_CUDA(h->sort_idx.reserve(h->stream, scratch));
        SB_CUDA(h->sort_tmp.reserve(h->stream, scratch));
        SB_CUDA(h->sort_flags.reserve(h->stream, scratch));
    }
    h->max_bins = max_cat_bins;
    return SB_OK;
}

// ---- NCCL, opened at run time so that single-GPU users need no NCCL at all.  In a process that already holds a
// libnccl.so.2 (PyTorch's), dlopen returns that copy.
struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};
NcclApi& nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api;
    tried = true;
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) return api;
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(dlsym(lib, "ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(dlsym(lib, "ncclCommInitRank"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(dlsym(lib, "ncclCommDestroy"));
    api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(dlsym(lib, "ncclAllReduce"));
    api.AllGather = reinterpret_cast<decltype(api.AllGather)>(dlsym(lib, "ncclAllGather"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(dlsym(lib, "ncclGetErrorString"));
    api.ok = api.GetUniqueId && api.CommInitRank && api.CommDestroy && api.AllReduce && api.GetErrorString;
    return api;
}
#define SB_NCCL(call)                                                                              \
    do {                                                                                           \
        ncclResult_t r_ = (call);                                                                  \
        if (r_ != ncclSuccess) return fail(SB_ERR_CUDA, std::string("NCCL: ") + nccl_api().GetErrorString(r_)); \
    } while (0)

int run_histograms(sb_handle* h) {
    if (h->forest_pass) return run_forest(h);
    for (const HistLaunch& L : h->launches) {
        if (L.use_rows) {
            SB_CUDA(launch_rowlist_build(L.list, h->stream));
            h->launches_count += L.list.identity ? 2 : 5;  // count, scan (2 kernels), chunk table, scatter | chunk table, operands
            if (L.has_narrow) {
                SB_CUDA(hist_rows_launch(L.kind, L.root_counts == 2 ? 2 : L.cnt_int, L.fpl, L.allf, L.rows, L.grid, L.smem, h->stream));
                h->launches_count += 1;
            }
            if (L.root_counts) {
                SB_CUDA(launch_counts_cache(L.rows.raw_lo, L.rows.raw_hi, h->root_cnt.p, L.root_bins, h->stats_size,
                                            L.root_counts == 1, h->stream));
                h->launches_count += 1;
                if (L.root_counts == 1) {  // only now does the cache describe something: a pass that stopped earlier leaves it invalid
                    h->root_cnt_valid = true;
                    h->root_cnt_version = h->weights_version;
                    h->root_cnt_wcol = L.root_wcol;
                    h->root_cnt_total = h->layout.total;
                    h->root_cnt_feats = h->layout.nodes.empty() ? std::vector<int>() : h->layout.nodes[0].features;
                }
            }
            if (L.has_wide) {
                SB_CUDA(hist_rows_wide_launch(L.kind, L.cnt_int, L.wide, L.wide_grid, L.wide_smem, h->stream));
                h->launches_count += 1;
            }
        } else {
            SB_CUDA(hist_launch(L.kind, L.cnt_int, L.args, h->train.tensor_map, L.grid, L.smem, h->stream));
            h->launches_count += 1;
            h->rows_streamed_host += h->train.n * L.args.n_groups;
        }
    }
    return SB_OK;
}

}  // namespace

extern "C" {

SB_API const char* sb_last_error(void) { return g_last_error.c_str(); }
SB_API const char* sb_version(void) { return "shifu_b200 0.1 (sm_100a)"; }

SB_API int sb_create(const sb_config* cfg, sb_handle** out) {
    if (!cfg || !out) return fail(SB_ERR_INVALID, "null argument");
    *out = nullptr;
    if (cfg->n_features <= 0 || !cfg->n_bins) return fail(SB_ERR_INVALID, "n_features / n_bins missing");
    if (cfg->weights_per_row < 1) return fail(SB_ERR_INVALID, "weights_per_row must be >= 1");
    if (cfg->impurity < SB_IMPURITY_VARIANCE || cfg->impurity > SB_IMPURITY_GINI) return fail(SB_ERR_INVALID, "unknown impurity");
    if (cfg->loss < SB_LOSS_SQUARED || cfg->loss > SB_LOSS_LOG) return fail(SB_ERR_INVALID, "unknown loss");
    if (!is_variance(cfg->impurity) && (cfg->num_classes < 2 || cfg->num_classes > 16))
        return fail(SB_ERR_INVALID, "num_classes must be in [2, 16] for entropy / gini");
    if (cfg->is_gbdt && !is_variance(cfg->impurity))  // core/validator/ModelInspector.java:898-904
        return fail(SB_ERR_INVALID, "GBT supports only variance / friedmanmse impurity");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(SB_ERR_CUDA, "no CUDA device: shifu_b200 has no CPU fallback");
    }
    if (cfg->device < 0 || cfg->device >= ndev) return fail(SB_ERR_INVALID, "device ordinal out of range");
    SB_CUDA(cudaSetDevice(cfg->device));
    {   // keep what cudaFreeAsync returns inside the pool: the work buffers of a pass are re-sized level after level
        cudaMemPool_t pool = nullptr;
        if (cudaDeviceGetDefaultMemPool(&pool, cfg->device) == cudaSuccess && pool) {
            uint64_t keep = UINT64_MAX;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        } else {
            cudaGetLastError();
        }
    }
    std::unique_ptr<sb_handle> h(new sb_handle());
    h->cfg = *cfg;
    h->n_bins.assign(cfg->n_bins, cfg->n_bins + cfg->n_features);
    h->is_cat.assign((size_t)cfg->n_features, 0);
    if (cfg->is_categorical) h->is_cat.assign(cfg->is_categorical, cfg->is_categorical + cfg->n_features);
    h->column_num.resize((size_t)cfg->n_features);
    for (int f = 0; f < cfg->n_features; ++f) h->column_num[f] = cfg->column_num ? cfg->column_num[f] : f;
    size_t nb = 0;
    for (int f = 0; f < cfg->n_features; ++f) {
        if (h->n_bins[f] < 1 || h->n_bins[f] > 32768) return fail(SB_ERR_INVALID, "n_bins out of range");
        if (!h->is_cat[f]) nb += (size_t)h->n_bins[f];
        // getAllValidFeatures: numeric with > 1 boundary, categorical with > 0 categories
        if ((!h->is_cat[f] && h->n_bins[f] > 1) || (h->is_cat[f] && h->n_bins[f] - 1 > 0)) h->all_features.push_back(f);
    }
    h->wide_index.assign((size_t)cfg->n_features, -1);
    for (int f = 0; f < cfg->n_features; ++f)
        if (h->n_bins[f] > 256) h->wide_index[f] = h->n_wide++;
    if (cfg->bin_boundary) h->bin_boundary.assign(cfg->bin_boundary, cfg->bin_boundary + nb);
    h->cfg.n_bins = nullptr; h->cfg.is_categorical = nullptr; h->cfg.column_num = nullptr; h->cfg.bin_boundary = nullptr;
    h->trace = getenv("SB_TRACE") != nullptr;
    h->enable_row_lists = (cfg->flags & SB_FLAG_NO_ROW_LISTS) == 0;
    h->enable_subtraction = (cfg->flags & SB_FLAG_NO_SUBTRACTION) == 0;
    h->enable_root_cache = (cfg->flags & SB_FLAG_NO_ROOT_COUNT_CACHE) == 0;
    h->enable_forest = (cfg->flags & SB_FLAG_NO_FOREST_PASS) == 0;
    // an RF pass stages the feature lists of thousands of todo nodes: start with an arena that does not have to grow level by level
    if (!cfg->is_gbdt) SB_CUDA(h->stage.grow((size_t)32 << 20, h->stream));
    if (cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, cfg->device) != cudaSuccess || h->sm_count < 1) h->sm_count = 148;
    h->stats_size = is_variance(cfg->impurity) ? 3 : cfg->num_classes;
    h->stat_kind = is_variance(cfg->impurity) ? STAT_VARIANCE : STAT_CLASS;
    h->all_feature_pos.assign((size_t)cfg->n_features, -1);
    for (size_t k = 0; k < h->all_features.size(); ++k) h->all_feature_pos[(size_t)h->all_features[k]] = (int)k;
    for (int f : h->all_features) {
        h->all_feat_off.push_back(h->all_feat_total);
        h->all_feat_total += (int64_t)h->n_bins[f] * h->stats_size;
    }
    SB_CUDA(cudaMalloc(&h->d_feat_bins, sizeof(int32_t) * cfg->n_features));
    {
        std::vector<int64_t> int_off((size_t)cfg->n_features, -1);
        for (size_t k = 0; k < h->all_features.size(); ++k) int_off[(size_t)h->all_features[k]] = h->all_feat_off[k];
        SB_CUDA(cudaMalloc(&h->d_int_feat_off, sizeof(int64_t) * cfg->n_features));
        SB_CUDA(cudaMemcpy(h->d_int_feat_off, int_off.data(), sizeof(int64_t) * cfg->n_features, cudaMemcpyHostToDevice));
    }
    SB_CUDA(cudaMalloc(&h->d_is_cat, (size_t)cfg->n_features));
    SB_CUDA(cudaMalloc(&h->d_column_num, sizeof(int32_t) * cfg->n_features));
    SB_CUDA(cudaMemcpy(h->d_is_cat, h->is_cat.data(), (size_t)cfg->n_features, cudaMemcpyHostToDevice));
    SB_CUDA(cudaMemcpy(h->d_column_num, h->column_num.data(), sizeof(int32_t) * cfg->n_features, cudaMemcpyHostToDevice));
    for (int f : h->all_features)
        if (h->is_cat[(size_t)f]) { h->any_categorical = true; h->max_categorical_bins = std::max(h->max_categorical_bins, h->n_bins[(size_t)f]); }
    SB_CUDA(cudaMemcpy(h->d_feat_bins, h->n_bins.data(), sizeof(int32_t) * cfg->n_features, cudaMemcpyHostToDevice));
    if (!cfg->is_gbdt) {
        // RF: a pinned result block for 10^4 todo nodes up front (growing it synchronises the device and takes tens of
        // milliseconds); with categorical features a node also returns up to 4 KB of category bits
        const size_t pin0 = h->any_categorical ? (size_t)64 << 20 : (size_t)4 << 20;
        SB_CUDA(cudaMallocHost(&h->pin, pin0));
        h->pin_bytes = pin0;
    }
    SB_CUDA(cudaMalloc(&h->d_acc, sizeof(double) * 4));
    SB_CUDA(cudaMalloc(&h->d_err_scratch, sizeof(double) * ERR_SCRATCH_DOUBLES));
    SB_CUDA(cudaMemset(h->d_err_scratch, 0, sizeof(double) * ERR_SCRATCH_DOUBLES));
    SB_CUDA(cudaMalloc(&h->d_maxbits, 16));
    SB_CUDA(cudaMalloc(&h->d_scales, 16));
    SB_CUDA(cudaMalloc(&h->d_inv_scales, 32));
    SB_CUDA(cudaMalloc(&h->d_error_flag, 4));
    SB_CUDA(cudaMalloc(&h->d_rows_streamed, 8));
    SB_CUDA(cudaMemset(h->d_rows_streamed, 0, 8));
    SB_CUDA(cudaMemset(h->d_error_flag, 0, 4));
    SB_CUDA(cudaMalloc(&h->d_cnt_shift, 16));  // count shift, integer-weights flag, scratch slot counter of the best-split kernel
    SB_CUDA(cudaMemset(h->d_cnt_shift, 0, 16));
    *out = h.release();
    return SB_OK;
}

SB_API void sb_destroy(sb_handle* h) {
    if (!h) return;
    cudaSetDevice(h->cfg.device);
    cudaDeviceSynchronize();
    if (h->trace) dump_trace(h);
    if (h->p2p.buf[0] || h->p2p.flags) p2p_teardown(h);
    h->stage.release();
    free_rowset(h->train);
    free_rowset(h->valid);
    for (auto& kv : h->trees) free_tree(kv.second);
    h->raw.release(); h->raw_prev.release(); h->d_pairs.release(); h->d_small_is_b.release(); h->stats.release(); h->tile_feats.release(); h->twin_map.release(); h->root_cnt.release(); h->slot_fmask.release(); h->pair_off.release();
    h->slot_of_node.release(); h->tasks.release(); h->task_node_off.release(); h->node_task_begin.release(); h->cand.release(); h->gain_out.release();
    h->cand_bits.release(); h->out_bits.release(); h->new_dropout.release(); h->new_weights.release();
    h->sort_keys.release(); h->sort_idx.release(); h->sort_tmp.release(); h->sort_flags.release();
    h->rl_counts.release(); h->rl_temp.release(); h->rl_tables.release(); h->rl_chunk_group.release(); h->rl_ops.release(); h->rt_tiles.release(); h->rt_fmask.release(); h->wt_tiles.release(); h->wt_fmask.release();
    if (h->d_feat_bins) cudaFree(h->d_feat_bins);
    if (h->d_is_cat) cudaFree(h->d_is_cat);
    if (h->d_column_num) cudaFree(h->d_column_num);
    if (h->d_acc) cudaFree(h->d_acc);
    if (h->pin) cudaFreeHost(h->pin);
    if (h->comm && nccl_api().ok) nccl_api().CommDestroy(h->comm);
    if (h->d_err_scratch) cudaFree(h->d_err_scratch);
    if (h->d_maxbits) cudaFree(h->d_maxbits);
    if (h->d_scales) cudaFree(h->d_scales);
    if (h->d_inv_scales) cudaFree(h->d_inv_scales);
    if (h->d_error_flag) cudaFree(h->d_error_flag);
    if (h->d_int_feat_off) cudaFree(h->d_int_feat_off);
    if (h->d_rows_streamed) cudaFree(h->d_rows_streamed);
    if (h->d_cnt_shift) cudaFree(h->d_cnt_shift);
    delete h;
}

SB_API int sb_set_stream(sb_handle* h, void* cuda_stream) {
    if (!h) return fail(SB_ERR_INVALID, "null handle");
    h->stream = (cudaStream_t)cuda_stream;
    return SB_OK;
}

SB_API int sb_load_rows(sb_handle* h, int32_t is_validation, int64_t n_rows, const void* const* bins, const float* label,
                        const float* significance, const float* weights) {
    if (!h || !bins || !label || n_rows < 0) return fail(SB_ERR_INVALID, "null argument");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    RowSet& r = is_validation ? h->valid : h->train;
    if (r.loaded) return fail(SB_ERR_UNSUPPORTED, "rows of this set were already loaded (one load call per set)");
    if (n_rows >= (int64_t)0xffffffff) return fail(SB_ERR_UNSUPPORTED, "2^32 - 1 rows or more in one handle (row indices are 32-bit on the device): shard the rows over several GPUs");
    const int F = h->cfg.n_features, W = h->cfg.weights_per_row;
    r.n = n_rows;
    r.n_features = F;
    r.owns_inputs = true;
    const int64_t pitch = ((n_rows + 255) / 256) * 256 + 256;  // tail tile reads stay inside the allocation
    SB_CUDA(cudaMalloc(&r.owned_bins, (size_t)pitch * F));
    SB_CUDA(cudaMemsetAsync(r.owned_bins, 0, (size_t)pitch * F, h->stream));
    r.bins_base = r.owned_bins;
    r.pitch = pitch;
    r.n_wide = h->n_wide;
    r.pitch16 = pitch;
    if (h->n_wide > 0) SB_CUDA(cudaMalloc(&r.bins16, (size_t)pitch * h->n_wide * 2));
    for (int f = 0; f < F; ++f) {
        if (!bins[f]) return fail(SB_ERR_INVALID, "null feature column");
        if (n_rows <= 0) continue;
        if (h->wide_index[f] >= 0)  // uint16 column (the reference stores short, dt/DTWorker.java:1593)
            SB_CUDA(cudaMemcpyAsync(r.bins16 + (size_t)h->wide_index[f] * pitch, bins[f], (size_t)n_rows * 2, cudaMemcpyHostToDevice, h->stream));
        else
            SB_CUDA(cudaMemcpyAsync(static_cast<uint8_t*>(r.owned_bins) + (size_t)f * pitch, bins[f], (size_t)n_rows, cudaMemcpyHostToDevice, h->stream));
    }
    const size_t nf = (size_t)std::max<int64_t>(n_rows, 1);
    SB_CUDA(cudaMalloc(&r.label, nf * 4));
    SB_CUDA(cudaMalloc(&r.significance, nf * 4));
    SB_CUDA(cudaMemcpyAsync(r.label, label, (size_t)n_rows * 4, cudaMemcpyHostToDevice, h->stream));
    if (significance) SB_CUDA(cudaMemcpyAsync(r.significance, significance, (size_t)n_rows * 4, cudaMemcpyHostToDevice, h->stream));
    else SB_CUDA(launch_fill_f32(r.significance, n_rows, 1.0f, h->stream));
    if (!is_validation) {
        SB_CUDA(cudaMalloc(&r.weights, nf * 4 * W));
        if (weights) SB_CUDA(cudaMemcpyAsync(r.weights, weights, (size_t)n_rows * 4 * W, cudaMemcpyHostToDevice, h->stream));
        else SB_CUDA(launch_fill_f32(r.weights, n_rows * W, 1.0f, h->stream));
    }
    h->h2d_bytes += n_rows * (F + 8 + (weights ? 4 * W : 0));
    return finish_load(h, r);
}

SB_API int sb_load_rows_device(sb_handle* h, int32_t is_validation, int64_t n_rows, const void* bins, int64_t pitch,
                               float* label, float* significance, float* weights) {
    if (!h || !bins || !label || !significance || n_rows < 0) return fail(SB_ERR_INVALID, "null argument");
    if (((uintptr_t)bins & 15) || (pitch & 15) || pitch < n_rows) return fail(SB_ERR_INVALID, "bin matrix must be 16-byte aligned with a pitch that is a multiple of 16 and >= n_rows");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    RowSet& r = is_validation ? h->valid : h->train;
    if (r.loaded) return fail(SB_ERR_UNSUPPORTED, "rows of this set were already loaded (one load call per set)");
    if (n_rows >= (int64_t)0xffffffff) return fail(SB_ERR_UNSUPPORTED, "2^32 - 1 rows or more in one handle (row indices are 32-bit on the device): shard the rows over several GPUs");
    const int F = h->cfg.n_features;
    r.n = n_rows;
    r.n_features = F;
    r.owns_inputs = false;
    if (h->n_wide > 0) return fail(SB_ERR_UNSUPPORTED, "sb_load_rows_device takes the uint8 matrix only: features wider than 256 bins need sb_load_rows");
    r.bins_base = bins;
    r.pitch = pitch;
    r.label = label;
    r.significance = significance;
    r.weights = is_validation ? nullptr : weights;
    return finish_load(h, r);
}

SB_API int sb_load_synthetic(sb_handle* h, int32_t is_validation, int64_t row0, int64_t n_rows, uint64_t seed,
                             int32_t label_kind, int32_t weighted) {
    if (!h || n_rows < 0) return fail(SB_ERR_INVALID, "null argument");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    RowSet& r = is_validation ? h->valid : h->train;
    if (r.loaded) return fail(SB_ERR_UNSUPPORTED, "rows of this set were already loaded (one load call per set)");
    if (n_rows >= (int64_t)0xffffffff) return fail(SB_ERR_UNSUPPORTED, "2^32 - 1 rows or more in one handle (row indices are 32-bit on the device): shard the rows over several GPUs");
    const int F = h->cfg.n_features, W = h->cfg.weights_per_row;
    r.n = n_rows;
    r.n_features = F;
    r.owns_inputs = true;
    const int64_t pitch = ((n_rows + 255) / 256) * 256 + 256;
    SB_CUDA(cudaMalloc(&r.owned_bins, (size_t)pitch * F));
    r.bins_base = r.owned_bins;
    r.pitch = pitch;
    r.n_wide = h->n_wide;
    r.pitch16 = pitch;
    if (h->n_wide > 0) {
        SB_CUDA(cudaMalloc(&r.bins16, (size_t)pitch * h->n_wide * 2));
        for (int f = 0; f < F; ++f)
            if (h->wide_index[f] >= 0) {
                SB_CUDA(launch_synthetic16(r.bins16 + (size_t)h->wide_index[f] * pitch, n_rows, f, F, h->n_bins[f], row0, seed, h->stream));
                h->launches_count += 1;
            }
    }
    const size_t nf = (size_t)std::max<int64_t>(n_rows, 1);
    SB_CUDA(cudaMalloc(&r.label, nf * 4));
    SB_CUDA(cudaMalloc(&r.significance, nf * 4));
    if (!is_validation) {
        SB_CUDA(cudaMalloc(&r.weights, nf * 4 * W));
        SB_CUDA(launch_fill_f32(r.weights, n_rows * W, 1.0f, h->stream));
    }
    SB_CUDA(launch_synthetic(static_cast<uint8_t*>(r.owned_bins), pitch, F, h->d_feat_bins, r.label, r.significance, row0,
                             n_rows, seed, label_kind, weighted, h->stream));
    h->launches_count += 3;
    return finish_load(h, r);
}

SB_API int sb_read_bins(sb_handle* h, int32_t is_validation, int32_t feature, int64_t row0, int64_t n_rows, void* out) {
    if (!h || !out) return fail(SB_ERR_INVALID, "null argument");
    RowSet& r = is_validation ? h->valid : h->train;
    if (!r.loaded || feature < 0 || feature >= r.n_features || row0 < 0 || row0 + n_rows > r.n)
        return fail(SB_ERR_INVALID, "range outside the loaded rows");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    const size_t width = h->wide_index[feature] >= 0 ? 2 : 1;  // `out` receives uint16 for features wider than 256 bins
    SB_CUDA(cudaMemcpyAsync(out, static_cast<const uint8_t*>(r.h_col[feature]) + row0 * width, (size_t)n_rows * width, cudaMemcpyDeviceToHost, h->stream));
    SB_CUDA(cudaStreamSynchronize(h->stream));
    h->d2h_bytes += n_rows;
    return SB_OK;
}

SB_API int sb_read_labels(sb_handle* h, int32_t is_validation, int64_t row0, int64_t n_rows, float* label, float* significance) {
    if (!h) return fail(SB_ERR_INVALID, "null argument");
    RowSet& r = is_validation ? h->valid : h->train;
    if (!r.loaded || row0 < 0 || row0 + n_rows > r.n) return fail(SB_ERR_INVALID, "range outside the loaded rows");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    if (label) SB_CUDA(cudaMemcpyAsync(label, r.label + row0, (size_t)n_rows * 4, cudaMemcpyDeviceToHost, h->stream));
    if (significance) SB_CUDA(cudaMemcpyAsync(significance, r.significance + row0, (size_t)n_rows * 4, cudaMemcpyDeviceToHost, h->stream));
    SB_CUDA(cudaStreamSynchronize(h->stream));
    return SB_OK;
}

SB_API int sb_read_rows(sb_handle* h, int32_t is_validation, int32_t tree_id, int32_t w_index, float* predict, float* output,
                        int32_t* node_id, float* weight) {
    if (!h) return fail(SB_ERR_INVALID, "null argument");
    RowSet& r = is_validation ? h->valid : h->train;
    if (!r.loaded) return fail(SB_ERR_STATE, "rows not loaded");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    const size_t nb = (size_t)r.n * 4;
    if (predict) SB_CUDA(cudaMemcpyAsync(predict, r.predict, nb, cudaMemcpyDeviceToHost, h->stream));
    if (output) SB_CUDA(cudaMemcpyAsync(output, r.output, nb, cudaMemcpyDeviceToHost, h->stream));
    if (node_id) {
        const int arr = h->cfg.is_gbdt ? 0 : tree_id;
        if (arr < 0 || arr >= r.n_node_arrays) return fail(SB_ERR_INVALID, "tree id out of range");
        SB_CUDA(cudaMemcpyAsync(node_id, r.node_id + (int64_t)arr * r.n, nb, cudaMemcpyDeviceToHost, h->stream));
    }
    if (weight) {
        if (!r.weights || w_index < 0 || w_index >= h->cfg.weights_per_row) return fail(SB_ERR_INVALID, "weight index out of range");
        SB_CUDA(cudaMemcpyAsync(weight, r.weights + (int64_t)w_index * r.n, nb, cudaMemcpyDeviceToHost, h->stream));
    }
    SB_CUDA(cudaStreamSynchronize(h->stream));
    return SB_OK;
}

SB_API int sb_set_weights(sb_handle* h, int32_t w_index, const float* weights) {
    if (!h || !weights) return fail(SB_ERR_INVALID, "null argument");
    RowSet& r = h->train;
    if (!r.loaded || !r.weights) return fail(SB_ERR_STATE, "training rows not loaded (or adopted without a weight array)");
    if (w_index < 0 || w_index >= h->cfg.weights_per_row) return fail(SB_ERR_INVALID, "weight index out of range");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    SB_CUDA(cudaMemcpyAsync(r.weights + (int64_t)w_index * r.n, weights, (size_t)r.n * 4, cudaMemcpyHostToDevice, h->stream));
    h->h2d_bytes += r.n * 4;
    h->scales_valid = false;
    return refresh_cnt_int(h, w_index);
}

SB_API int sb_sample_weights(sb_handle* h, int32_t w_index, int64_t seed, double rate, int64_t first_draw) {
    if (!h) return fail(SB_ERR_INVALID, "null handle");
    RowSet& r = h->train;
    if (!r.loaded || !r.weights) return fail(SB_ERR_STATE, "training rows not loaded (or adopted without a weight array)");
    if (w_index < 0 || w_index >= h->cfg.weights_per_row) return fail(SB_ERR_INVALID, "weight index out of range");
    if (first_draw < 0) return fail(SB_ERR_INVALID, "first_draw must be >= 0");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    SB_CUDA(launch_java_bagging_mask(r.weights + (int64_t)w_index * r.n, r.n, seed, rate, first_draw, h->stream));
    h->launches_count += 1;
    h->scales_valid = false;
    return refresh_cnt_int(h, w_index);
}

SB_API int sb_sample_weights_stratified(sb_handle* h, int32_t w_index, int64_t seed, double rate, int32_t n_classes) {
    if (!h) return fail(SB_ERR_INVALID, "null handle");
    RowSet& r = h->train;
    if (!r.loaded || !r.weights) return fail(SB_ERR_STATE, "training rows not loaded (or adopted without a weight array)");
    if (w_index < 0 || w_index >= h->cfg.weights_per_row) return fail(SB_ERR_INVALID, "weight index out of range");
    if (n_classes < 1 || n_classes > 64) return fail(SB_ERR_INVALID, "n_classes must be in [1, 64]");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    const size_t n = (size_t)std::max<int64_t>(r.n, 1);
    const size_t temp_bytes = stratified_temp_bytes(r.n);
    SB_CUDA(h->rl_counts.reserve(h->stream, 2 * n));
    SB_CUDA(h->rl_temp.reserve(h->stream, temp_bytes + 16));
    SB_CUDA(cudaMemsetAsync(h->d_error_flag, 0, 4, h->stream));
    SB_CUDA(launch_stratified_mask(r.weights + (int64_t)w_index * r.n, r.label, r.n, n_classes, seed, rate, h->rl_counts.p,
                                   h->rl_counts.p + n, h->rl_temp.p, temp_bytes, h->d_error_flag, h->stream));
    h->launches_count += 4 * n_classes;
    uint32_t flag = 0;
    SB_CUDA(cudaMemcpyAsync(&flag, h->d_error_flag, 4, cudaMemcpyDeviceToHost, h->stream));
    SB_CUDA(cudaStreamSynchronize(h->stream));
    if (flag & 1u) return fail(SB_ERR_RANGE, "class (int)(label + 0.01f) outside [0, n_classes)");
    h->scales_valid = false;
    return refresh_cnt_int(h, w_index);
}

SB_API int sb_sample_poisson(sb_handle* h, int32_t w_index, uint64_t seed, double rate, int64_t row0) {
    if (!h) return fail(SB_ERR_INVALID, "null handle");
    RowSet& r = h->train;
    if (!r.loaded || !r.weights) return fail(SB_ERR_STATE, "training rows not loaded (or adopted without a weight array)");
    if (w_index < 0 || w_index >= h->cfg.weights_per_row) return fail(SB_ERR_INVALID, "weight index out of range");
    if (!(rate > 0.0) || rate > 64.0 || row0 < 0) return fail(SB_ERR_INVALID, "rate must be in (0, 64], row0 >= 0");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    SB_CUDA(launch_poisson_bags(r.weights + (int64_t)w_index * r.n, r.n, seed, row0, w_index, rate, std::exp(-rate), h->stream));
    h->launches_count += 1;
    h->scales_valid = false;
    return refresh_cnt_int(h, w_index);
}

SB_API int64_t sb_stats_len(sb_handle* h, const sb_iter_in* in) {
    if (!h || !in) return fail(SB_ERR_INVALID, "null argument");
    // the sum build_layout forms, without building the layout (callers ask before every pass)
    int64_t total = 0;
    for (int i = 0; i < in->n_todo; ++i) {
        const sb_todo_node& t = in->todo[i];
        if (t.n_features == 0) { total += h->all_feat_total; continue; }
        for (int k = 0; k < t.n_features; ++k) {
            const int f = t.features[k];
            if (f < 0 || f >= h->cfg.n_features) return fail(SB_ERR_INVALID, "todo node feature slot out of range");
            total += (int64_t)h->n_bins[(size_t)f] * h->stats_size;
        }
    }
    return total;
}

static int pinned_reserve(sb_handle* h, size_t bytes);
static size_t pinned_need(const sb_handle* h);

SB_API int sb_compute_begin(sb_handle* h, const sb_iter_in* in) {
    if (!h || !in) return fail(SB_ERR_INVALID, "null argument");
    if (!h->train.loaded) return fail(SB_ERR_STATE, "sb_compute before sb_load_rows");
    if (in->continuous_running_start) return fail(SB_ERR_UNSUPPORTED, "continuous training is handled on the Java side");
    if (in->n_trees < 1 || !in->trees) return fail(SB_ERR_INVALID, "no trees");
    // A rank without a single training row plans no sibling pairs, so after the merge it would not derive the larger
    // children the other ranks derive: its statistics, and the trees its master grows from them, would differ.
    if (h->comm && h->train.n == 0) return fail(SB_ERR_UNSUPPORTED, "multi-GPU merge: every rank must hold at least one training row");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    if (h->trace) {
        sb_handle::PassTrace p{};
        p.t0 = now_ms();
        p.n_todo = in->n_todo;
        cudaEventCreate(&p.e0);
        cudaEventCreate(&p.e1);
        cudaEventCreate(&p.c0);
        cudaEventCreate(&p.c1);
        cudaEventRecord(p.e0, h->stream);
        h->passes_trace.push_back(p);
    }
    h->begun = false;
    h->stats_ready = false;
    // A pass that does not reach finish_check (an error, or a second begin) must not leave the next one pairing its
    // children with accumulators of an older pass: the claim is renewed only by a finished pass.
    h->use_prev = h->prev_valid;
    h->prev_valid = false;
    h->stage.reset();  // the previous pass ended with a stream synchronisation: its tables are no longer in use
    std::swap(h->layout, h->prev_layout);  // the last pass's layout goes with raw_prev; its storage before that is recycled
    h->forest_pass = forest_lists(h);
    // Nodes with feature lists of their own (FeatureSubsetStrategy): when the lists cover a fair share of the features,
    // accumulate EVERY valid feature for every node (the kernels do that in shared memory anyway, see plan_histograms) and
    // keep it, so that histogram subtraction and the root count cache work as they do without lists; the caller's
    // statistics are gathered from that internal layout (finish_enqueue).
    bool full = false;
    if (h->enable_subtraction && in->n_todo > 0 && !h->all_features.empty()) {
        int64_t listed = 0;
        bool own = false;
        bool listable = true;  // every listed feature has a block in the all-features layout
        for (int i = 0; i < in->n_todo; ++i) {
            const sb_todo_node& t = in->todo[i];
            own |= t.n_features != 0;
            listed += t.n_features == 0 ? (int64_t)h->all_features.size() : t.n_features;
            for (int k = 0; k < t.n_features && listable; ++k)
                listable = t.features[k] >= 0 && t.features[k] < h->cfg.n_features && h->all_feature_pos[(size_t)t.features[k]] >= 0;
        }
        static const bool off = getenv("SB_NO_FULL_PASS") != nullptr;
        full = own && listable && !off && 4 * listed >= (int64_t)in->n_todo * (int64_t)h->all_features.size();
    }
    int rc = build_layout(h, in, h->layout, h->forest_pass || full, full, !h->forest_pass);
    if (rc != SB_OK) return rc;
    if (h->layout.flat) {
        rc = stage_flat_lists(h);
        if (rc != SB_OK) return rc;
    }
    rc = pinned_reserve(h, pinned_need(h));  // now, while the GPU is idle: growing the pinned block waits for the device
    if (rc != SB_OK) return rc;
    const int64_t total = h->layout.total;
    SB_CUDA(h->raw.reserve(h->stream, (size_t)std::max<int64_t>(2 * total, 1)));
    SB_CUDA(h->stats.reserve(h->stream, (size_t)std::max<int64_t>(h->layout.caller_total, 1)));
    SB_CUDA(cudaMemsetAsync(h->raw.p, 0, (size_t)(2 * total) * 8, h->stream));
    // GBT error sums are sums over Data.predict, which only changes at a tree switch (dt/DTWorker.java:627-640, 681):
    // inside a tree (after the first one) the four scalars of the previous pass are still the answer.
    const bool errors_from_tree = h->cfg.is_gbdt && in->is_first_tree && !in->switch_to_next_tree;
    const bool need_errors = !h->cfg.is_gbdt || in->switch_to_next_tree || errors_from_tree || !h->acc_valid;
    h->acc_fresh = need_errors;
    if (need_errors) SB_CUDA(cudaMemsetAsync(h->d_acc, 0, 32, h->stream));
    SB_CUDA(cudaMemsetAsync(h->d_error_flag, 0, 4, h->stream));
    RowSet& tr = h->train;
    RowSet& va = h->valid;
    const int loss = h->cfg.loss;
    double tt0 = h->trace ? now_ms() : 0;
    if (h->trace) h->passes_trace.back().sub[0] = tt0 - h->passes_trace.back().t0;

    bool fused_switch = false;  // the GBT tree switch already produced the error sums and the scale maxima of this pass
    if (h->cfg.is_gbdt) {
        const int curr = in->n_trees - 1;  // currTreeIndex (dt/DTWorker.java:620)
        const sb_tree& cur_tree = in->trees[curr];
        if (cur_tree.n_nodes < 1) return fail(SB_ERR_INVALID, "current tree has no nodes");
        if (in->switch_to_next_tree) {
            if (curr >= 1) {
                const sb_tree& prev = in->trees[curr - 1];
                DeviceTree& dprev = h->trees[prev.tree_id];
                rc = upload_tree(h, prev, dprev);
                if (rc != SB_OK) return rc;
                SB_CUDA(h->stage.flush(h->stream));
                if (h->tracked_tree != prev.tree_id) {  // first time we see it: walk from the root
                    SB_CUDA(launch_fill_i32(tr.node_id, tr.n, SB_ROOT_INDEX, h->stream));
                    if (va.loaded) SB_CUDA(launch_fill_i32(va.node_id, va.n, SB_ROOT_INDEX, h->stream));
                }
                const TreeView pv = view_of(dprev);
                if (in->new_weights) {  // redraw of subsampleWeights[currTreeIndex % len] (dt/DTWorker.java:643-668), injected
                    const int W = h->cfg.weights_per_row;
                    SB_CUDA(cudaMemcpyAsync(tr.weights + (int64_t)(curr % W) * tr.n, in->new_weights, (size_t)tr.n * 4,
                                            cudaMemcpyHostToDevice, h->stream));
                    h->h2d_bytes += tr.n * 4;
                    rc = refresh_cnt_int(h, curr % W);
                    if (rc != SB_OK) return rc;
                }
                const uint8_t* d_drop = nullptr;
                if (in->dropout && !in->is_first_tree) {
                    SB_CUDA(h->new_dropout.reserve(h->stream, (size_t)tr.n));
                    SB_CUDA(cudaMemcpyAsync(h->new_dropout.p, in->dropout, (size_t)tr.n, cudaMemcpyHostToDevice, h->stream));
                    h->h2d_bytes += tr.n;
                    d_drop = h->new_dropout.p;
                }
                // one pass over the rows: walk the finished tree, update predict / output, back to the root, error
                // sums over the new predict, maxima for the next tree's fixed-point scales (gbt_switch_kernel)
                SB_CUDA(cudaMemsetAsync(h->d_maxbits, 0, 16, h->stream));
                SwitchArgs sa{};
                sa.n = tr.n; sa.cols = tr.d_col; sa.feat_bins = h->d_feat_bins; sa.node_id = tr.node_id; sa.prev = pv;
                sa.is_first_tree = in->is_first_tree; sa.learning_rate = h->cfg.learning_rate; sa.loss = loss;
                sa.label = tr.label; sa.predict = tr.predict; sa.output = tr.output; sa.significance = tr.significance;
                const int wcol = cur_tree.tree_id % (tr.weights ? h->cfg.weights_per_row : 1);
                sa.weights = (tr.weights && !h->weights_all_one) ? tr.weights + (int64_t)wcol * tr.n : nullptr;
                sa.dropout = d_drop; sa.is_validation = 0; sa.maxbits = h->d_maxbits; sa.error_flag = h->d_error_flag;
                SB_CUDA(launch_gbt_switch(sa, h->d_err_scratch, h->d_acc, h->stream));
                h->launches_count += 1;
                if (va.loaded) {
                    sa.n = va.n; sa.cols = va.d_col; sa.node_id = va.node_id; sa.label = va.label; sa.predict = va.predict;
                    sa.output = va.output; sa.significance = va.significance; sa.weights = nullptr; sa.dropout = nullptr;
                    sa.is_validation = 1; sa.maxbits = nullptr;
                    SB_CUDA(launch_gbt_switch(sa, h->d_err_scratch, h->d_acc, h->stream));
                    h->launches_count += 1;
                }
                fused_switch = true;
            }
            if (!fused_switch) {  // the very first tree: nothing to walk, every row starts at the root
                SB_CUDA(launch_fill_i32(tr.node_id, tr.n, SB_ROOT_INDEX, h->stream));
                if (va.loaded) SB_CUDA(launch_fill_i32(va.node_id, va.n, SB_ROOT_INDEX, h->stream));
                h->launches_count += 1 + (va.loaded ? 1 : 0);
            }
        } else if (h->tracked_tree != cur_tree.tree_id) {
            SB_CUDA(launch_fill_i32(tr.node_id, tr.n, SB_ROOT_INDEX, h->stream));
            if (va.loaded) SB_CUDA(launch_fill_i32(va.node_id, va.n, SB_ROOT_INDEX, h->stream));
            h->launches_count += 1 + (va.loaded ? 1 : 0);
        }
        DeviceTree& dcur = h->trees[cur_tree.tree_id];
        rc = upload_tree(h, cur_tree, dcur);
        if (rc != SB_OK) return rc;
        SB_CUDA(h->stage.flush(h->stream));
        h->tracked_tree = cur_tree.tree_id;
        // keep the device copies of the current and the previous tree only
        const int keep_prev = curr >= 1 ? in->trees[curr - 1].tree_id : cur_tree.tree_id;
        for (auto it = h->trees.begin(); it != h->trees.end();) {
            if (it->first != cur_tree.tree_id && it->first != keep_prev) { free_tree(it->second); it = h->trees.erase(it); }
            else ++it;
        }
        const TreeView cv = view_of(dcur);
        const bool from_tree = errors_from_tree;  // dt/DTWorker.java:672-683
        // predictNodeIndex for the training rows.  A tree that is only its root has nowhere to send them; otherwise
        // the walk is folded into the counting pass of the row lists (plan_histograms) unless the error sums of this
        // pass need the node ids first.
        h->pending_route = false;
        if (cur_tree.n_nodes > 1) {
            if (h->enable_row_lists && !(need_errors && from_tree)) {
                h->pending_route = true;
                h->pending_tree = cv;
            } else {
                SB_CUDA(launch_route(tr.d_col, h->d_feat_bins, tr.n, tr.node_id, cv, h->d_error_flag, h->stream));
                h->launches_count += 1;
            }
        }
        if (va.loaded && from_tree) {
            SB_CUDA(launch_route(va.d_col, h->d_feat_bins, va.n, va.node_id, cv, h->d_error_flag, h->stream));
            h->launches_count += 1;
        }
        ErrArgs e{};
        e.n = tr.n; e.label = tr.label; e.predict = tr.predict; e.significance = tr.significance; e.weights = nullptr;
        e.node_id = tr.node_id; e.tree = cv; e.from_tree = from_tree ? 1 : 0; e.loss = loss; e.is_validation = 0; e.rf_oob = 0;
        if (need_errors && fused_switch) {
            h->acc_valid = true;  // sums over Data.predict: they hold until the next tree switch
        } else if (need_errors) {
            SB_CUDA(launch_errors(e, h->d_err_scratch, h->d_acc, h->stream));
            h->launches_count += 1;
            if (va.loaded) {
                e.n = va.n; e.label = va.label; e.predict = va.predict; e.significance = va.significance; e.node_id = va.node_id;
                e.is_validation = 1;
                SB_CUDA(launch_errors(e, h->d_err_scratch, h->d_acc, h->stream));
                h->launches_count += 1;
            }
            h->acc_valid = !from_tree;  // sums over the current tree's leaves change with every level
        }
    } else {
        // RF (dt/DTWorker.java:583-605, 691-705): every real tree is walked for the error sums
        for (int t = 0; t < in->n_trees; ++t) {  // all trees reach the device with one copy
            const sb_tree& tree = in->trees[t];
            if (tree.n_nodes < 1) continue;  // INVALID_INDEX placeholder
            if (tree.tree_id < 0 || tree.tree_id >= tr.n_node_arrays) return fail(SB_ERR_INVALID, "tree id out of range");
            if (tree.tree_id >= h->cfg.weights_per_row) return fail(SB_ERR_INVALID, "RF needs one weight per tree");
            rc = upload_tree(h, tree, h->trees[tree.tree_id]);
            if (rc != SB_OK) return rc;
        }
        if (h->enable_forest) {
            // one launch walks every tree and takes the error sums of the pass (forest_walk_kernel)
            int n_real = 0;
            for (int t = 0; t < in->n_trees; ++t) n_real += in->trees[t].n_nodes >= 1 ? 1 : 0;
            WalkTree *wt = nullptr, *d_wt = nullptr, *wv = nullptr, *d_wv = nullptr;
            SB_CUDA(h->stage.take((size_t)std::max(n_real, 1), &wt, &d_wt, h->stream));
            if (va.loaded) SB_CUDA(h->stage.take((size_t)std::max(n_real, 1), &wv, &d_wv, h->stream));
            int k = 0;
            for (int t = 0; t < in->n_trees; ++t) {
                const sb_tree& tree = in->trees[t];
                if (tree.n_nodes < 1) continue;
                const DeviceTree& dt = h->trees[tree.tree_id];
                wt[k].nodes = static_cast<const WalkNode*>(dt.walk);
                wt[k].masks = dt.masks;
                wt[k].capacity = dt.capacity;
                wt[k].pad = 0;
                wt[k].node_id = tr.node_id + (int64_t)tree.tree_id * tr.n;
                wt[k].weights = tr.weights ? tr.weights + (int64_t)tree.tree_id * tr.n : nullptr;  // [treeId], no modulo (:592); adopted rows without weights: all 1
                if (va.loaded) {
                    wv[k] = wt[k];
                    wv[k].node_id = va.node_id + (int64_t)tree.tree_id * va.n;
                    wv[k].weights = nullptr;
                }
                ++k;
            }
            h->h2d_bytes += (int64_t)n_real * (int64_t)sizeof(WalkTree) * (va.loaded ? 2 : 1);
            SB_CUDA(h->stage.flush(h->stream));
            ForestWalkArgs wa{};
            wa.n = tr.n; wa.trees = d_wt; wa.n_trees = n_real; wa.cols = tr.d_col; wa.feat_bins = h->d_feat_bins;
            wa.rm_bins = (tr.bins_rm && !h->twin_pos_host.empty()) ? tr.bins_rm : nullptr; wa.fpad = (uint32_t)tr.fpad;
            wa.label = tr.label; wa.significance = tr.significance; wa.loss = loss; wa.is_validation = 0; wa.error_flag = h->d_error_flag;
            SB_CUDA(launch_forest_walk(wa, h->d_err_scratch, h->d_acc, h->stream));
            h->launches_count += 1;
            if (va.loaded) {
                wa.n = va.n; wa.trees = d_wv; wa.cols = va.d_col; wa.rm_bins = nullptr; wa.fpad = 0;
                wa.label = va.label; wa.significance = va.significance; wa.is_validation = 1;
                SB_CUDA(launch_forest_walk(wa, h->d_err_scratch, h->d_acc, h->stream));
                h->launches_count += 1;
            }
        } else {
        SB_CUDA(h->stage.flush(h->stream));
        for (int t = 0; t < in->n_trees; ++t) {
            const sb_tree& tree = in->trees[t];
            if (tree.n_nodes < 1) continue;
            const TreeView v = view_of(h->trees[tree.tree_id]);
            int32_t* nid = tr.node_id + (int64_t)tree.tree_id * tr.n;
            SB_CUDA(launch_route(tr.d_col, h->d_feat_bins, tr.n, nid, v, h->d_error_flag, h->stream));
            ErrArgs e{};
            e.n = tr.n; e.label = tr.label; e.predict = tr.predict; e.significance = tr.significance;
            e.weights = tr.weights ? tr.weights + (int64_t)tree.tree_id * tr.n : nullptr;  // [treeId], no modulo (:592); adopted rows without weights: all 1
            e.node_id = nid; e.tree = v; e.from_tree = 1; e.loss = loss; e.is_validation = 0; e.rf_oob = 1;
            SB_CUDA(launch_errors(e, h->d_err_scratch, h->d_acc, h->stream));
            h->launches_count += 2;
            if (va.loaded) {
                int32_t* vnid = va.node_id + (int64_t)tree.tree_id * va.n;
                SB_CUDA(launch_route(va.d_col, h->d_feat_bins, va.n, vnid, v, h->d_error_flag, h->stream));
                e.n = va.n; e.label = va.label; e.predict = va.predict; e.significance = va.significance; e.weights = nullptr;
                e.node_id = vnid; e.is_validation = 1; e.rf_oob = 0;
                SB_CUDA(launch_errors(e, h->d_err_scratch, h->d_acc, h->stream));
                h->launches_count += 2;
            }
        }
        }
    }

    if (h->trace) { const double t = now_ms(); h->passes_trace.back().sub[1] = t - tt0; tt0 = t; }
    // Fixed-point scales: maxima of the per-row contributions.  They change only when Data.output or the sample
    // weights change, i.e. at a GBT tree switch (RF: never after load), so they are computed once per tree and
    // the accumulators of all levels of a tree share one scale.
    h->scales_fresh = false;
    if (fused_switch) {
        h->scales_fresh = true;
        h->scales_valid = true;
    } else if (!h->scales_valid || (h->cfg.is_gbdt && in->switch_to_next_tree)) {
        h->scales_fresh = true;
        SB_CUDA(cudaMemsetAsync(h->d_maxbits, 0, 16, h->stream));
        const int W = tr.weights ? h->cfg.weights_per_row : 1;
        std::vector<int> wcols;
        if (h->cfg.is_gbdt) {
            for (const auto& nd : h->layout.nodes) {
                const int c = nd.tree_id % W;
                if (std::find(wcols.begin(), wcols.end(), c) == wcols.end()) wcols.push_back(c);
            }
            if (wcols.empty()) wcols.push_back((in->n_trees - 1) % W);
        } else {
            for (int c = 0; c < W; ++c) wcols.push_back(c);
        }
        for (int c : wcols) {
            SB_CUDA(launch_absmax(tr.n, tr.output, tr.significance, (tr.weights && !h->weights_all_one) ? tr.weights + (int64_t)c * tr.n : nullptr,
                                  h->d_maxbits, h->stream));
            h->launches_count += 1;
        }
        h->scales_valid = true;
    }
    if (h->trace) { const double t = now_ms(); h->passes_trace.back().sub[2] = t - tt0; tt0 = t; }
    rc = h->forest_pass ? plan_forest(h, in) : plan_histograms(h, in);
    if (rc != SB_OK) return rc;
    if (h->trace) { const double t = now_ms(); h->passes_trace.back().sub[3] = t - tt0; tt0 = t; }
    if (h->pending_route) {  // no row list took the walk along: do it on its own
        SB_CUDA(launch_route(tr.d_col, h->d_feat_bins, tr.n, tr.node_id, h->pending_tree, h->d_error_flag, h->stream));
        h->launches_count += 1;
        h->pending_route = false;
    }
    h->split_planned = false;  // the best-split task table is built while the statistics kernels run (sb_compute_stats)
    SB_CUDA(h->stage.flush(h->stream));
    h->begun = true;
    if (h->trace) { h->passes_trace.back().t_begin = now_ms(); h->passes_trace.back().sub[4] = h->passes_trace.back().t_begin - tt0; }
    return SB_OK;
}

SB_API int sb_scale_buffer(sb_handle* h, void** dev_ptr, int64_t* bytes) {
    if (!h || !dev_ptr) return fail(SB_ERR_INVALID, "null argument");
    *dev_ptr = h->d_maxbits;
    if (bytes) *bytes = 12;
    return SB_OK;
}

SB_API int sb_pass_info(sb_handle* h, int32_t* scales_fresh, int32_t* acc_fresh) {
    if (!h) return fail(SB_ERR_INVALID, "null handle");
    // after sb_dist_init the library merges by itself: nothing is left for the caller to reduce
    if (scales_fresh) *scales_fresh = (h->scales_fresh && !h->comm) ? 1 : 0;
    if (acc_fresh) *acc_fresh = (h->acc_fresh && !h->comm) ? 1 : 0;
    return SB_OK;
}

SB_API int sb_compute_stats(sb_handle* h) {
    if (!h) return fail(SB_ERR_INVALID, "null argument");
    if (!h->begun) return fail(SB_ERR_STATE, "sb_compute_stats before sb_compute_begin");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    // one process per GPU: the maxima behind the fixed-point scales are non-negative floats, so their bit patterns
    // order like unsigned integers
    if (h->comm && h->scales_fresh)
        SB_NCCL(nccl_api().AllReduce(h->d_maxbits, h->d_maxbits, 3, ncclUint32, ncclMax, h->comm, h->stream));
    SB_CUDA(launch_make_scales(h->d_maxbits, h->stat_kind, h->d_scales, h->d_inv_scales, h->d_cnt_shift, h->stream));
    h->launches_count += 1;
    int rc;
    if (!h->profiling) rc = run_histograms(h);
    else {
        cudaEvent_t e0, e1;
        SB_CUDA(cudaEventCreate(&e0));
        SB_CUDA(cudaEventCreate(&e1));
        SB_CUDA(cudaEventRecord(e0, h->stream));
        rc = run_histograms(h);
        SB_CUDA(cudaEventRecord(e1, h->stream));
        h->prof_events.push_back({e0, e1});
        h->prof_launches.push_back((int)h->launches.size());
    }
    if (rc != SB_OK) return rc;
    if (!h->split_planned) {  // host work that only the best-split step needs: done while the GPU is busy
        rc = plan_split_tasks(h);
        if (rc != SB_OK) return rc;
        SB_CUDA(h->stage.flush(h->stream));
        h->split_planned = true;
    }
    if (h->trace && !h->passes_trace.empty()) cudaEventRecord(h->passes_trace.back().c0, h->stream);
    if (h->comm) {
        // the DTMaster merge (dt/DTMaster.java:316-331 adds the workers' statistics node by node): one integer sum
        // over all GPUs, exact and order-independent; plus the four error scalars when this pass recomputed them
        if (h->layout.total > 0) {
            const int which = h->raw.p == h->p2p.buf[0] ? 0 : (h->raw.p == h->p2p.buf[1] ? 1 : -1);
            if (h->p2p.ok && which >= 0) {  // the accumulators sit in peer-mapped memory: merge over NVLink loads (p2p.cu)
                // the tail barrier is implied only when the two buffers alternate from pass to pass (finish_check)
                const int alternate = (h->enable_subtraction && (!h->forest_pass || h->layout.full)) ? 1 : 0;
                SB_CUDA(launch_p2p_allreduce(h->p2p.view, which, 2 * h->layout.total, ++h->p2p.epoch, alternate ? 0 : 1,
                                             h->d_error_flag, h->stream));
                h->launches_count += 5;
            } else {
                SB_NCCL(nccl_api().AllReduce(h->raw.p, h->raw.p, (size_t)(2 * h->layout.total), ncclInt64, ncclSum, h->comm, h->stream));
            }
        }
        if (h->acc_fresh) SB_NCCL(nccl_api().AllReduce(h->d_acc, h->d_acc, 4, ncclFloat64, ncclSum, h->comm, h->stream));
    }
    if (h->trace && !h->passes_trace.empty()) {
        cudaEventRecord(h->passes_trace.back().c1, h->stream);
        h->passes_trace.back().t_stats = now_ms();
    }
    return SB_OK;
}

SB_API int sb_dist_unique_id(uint8_t* id128) {
    if (!id128) return fail(SB_ERR_INVALID, "null argument");
    NcclApi& api = nccl_api();
    if (!api.ok) return fail(SB_ERR_UNSUPPORTED, "libnccl.so.2 not found");
    ncclUniqueId id;
    SB_NCCL(api.GetUniqueId(&id));
    static_assert(sizeof(id) == 128, "ncclUniqueId is 128 bytes");
    memcpy(id128, &id, 128);
    return SB_OK;
}

// Accumulators in peer-mapped memory (p2p.cu).  Any failure leaves the NCCL merge in place; every rank must reach the
// same answer, so the ranks vote.  Two buffers: the accumulators of this pass and of the previous one swap every pass.
static void p2p_teardown(sb_handle* h) {
    for (void* q : h->p2p.opened) cudaIpcCloseMemHandle(q);
    h->p2p.opened.clear();
    for (int k = 0; k < 2; ++k) {
        if (h->raw.p == h->p2p.buf[k]) h->raw.alias(nullptr);
        if (h->raw_prev.p == h->p2p.buf[k]) h->raw_prev.alias(nullptr);
        if (h->p2p.buf[k]) cudaFree(h->p2p.buf[k]);
        h->p2p.buf[k] = nullptr;
    }
    if (h->p2p.flags) cudaFree(h->p2p.flags);
    h->p2p.flags = nullptr;
    h->p2p.ok = false;
    cudaGetLastError();
}

static int p2p_setup(sb_handle* h) {
    NcclApi& api = nccl_api();
    if (h->world < 2 || h->world > P2P_MAX_WORLD || !api.AllGather || getenv("SB_NO_P2P")) return SB_OK;
    struct Wire { cudaIpcMemHandle_t mem[3]; int32_t ok; int32_t pad; };
    static_assert(sizeof(Wire) % 8 == 0, "all-gathered as bytes");
    const size_t mb = getenv("SB_P2P_MB") ? (size_t)atol(getenv("SB_P2P_MB")) : 96;  // per buffer; C2 needs 39 MB
    const size_t cap_bytes = mb << 20;
    Wire mine{};
    mine.ok = 1;
    bool local_ok = cudaMalloc((void**)&h->p2p.buf[0], cap_bytes) == cudaSuccess && cudaMalloc((void**)&h->p2p.buf[1], cap_bytes) == cudaSuccess &&
                    cudaMalloc((void**)&h->p2p.flags, 3 * P2P_MAX_WORLD * sizeof(uint32_t)) == cudaSuccess;
    if (local_ok) {
        cudaMemset(h->p2p.flags, 0, 3 * P2P_MAX_WORLD * sizeof(uint32_t));
        local_ok = cudaIpcGetMemHandle(&mine.mem[0], h->p2p.buf[0]) == cudaSuccess && cudaIpcGetMemHandle(&mine.mem[1], h->p2p.buf[1]) == cudaSuccess &&
                   cudaIpcGetMemHandle(&mine.mem[2], h->p2p.flags) == cudaSuccess;
    }
    if (!local_ok) { mine.ok = 0; cudaGetLastError(); }
    // the handles travel in one all-gather
    Wire *d_send = nullptr, *d_recv = nullptr;
    SB_CUDA(cudaMalloc((void**)&d_send, sizeof(Wire)));
    SB_CUDA(cudaMalloc((void**)&d_recv, sizeof(Wire) * (size_t)h->world));
    SB_CUDA(cudaMemcpyAsync(d_send, &mine, sizeof(Wire), cudaMemcpyHostToDevice, h->stream));
    SB_NCCL(api.AllGather(d_send, d_recv, sizeof(Wire), ncclChar, h->comm, h->stream));
    std::vector<Wire> all((size_t)h->world);
    SB_CUDA(cudaMemcpyAsync(all.data(), d_recv, sizeof(Wire) * (size_t)h->world, cudaMemcpyDeviceToHost, h->stream));
    SB_CUDA(cudaStreamSynchronize(h->stream));
    int32_t ok = 1;
    for (auto& w : all) ok &= w.ok;
    P2PView v{};
    v.world = h->world;
    v.rank = h->rank;
    for (int p = 0; p < h->world && ok; ++p) {
        if (p == h->rank) {
            v.peer_buf[0][p] = h->p2p.buf[0];
            v.peer_buf[1][p] = h->p2p.buf[1];
            v.peer_flags[p] = h->p2p.flags;
            continue;
        }
        void* q[3] = {nullptr, nullptr, nullptr};
        for (int k = 0; k < 3 && ok; ++k) {
            if (cudaIpcOpenMemHandle(&q[k], all[(size_t)p].mem[k], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { ok = 0; cudaGetLastError(); }
            else h->p2p.opened.push_back(q[k]);
        }
        v.peer_buf[0][p] = static_cast<int64_t*>(q[0]);
        v.peer_buf[1][p] = static_cast<int64_t*>(q[1]);
        v.peer_flags[p] = static_cast<uint32_t*>(q[2]);
    }
    // vote: a rank that could not map a peer takes everybody back to NCCL
    int32_t* d_vote = reinterpret_cast<int32_t*>(d_send);
    SB_CUDA(cudaMemcpyAsync(d_vote, &ok, 4, cudaMemcpyHostToDevice, h->stream));
    SB_NCCL(api.AllReduce(d_vote, d_vote, 1, ncclInt32, ncclMin, h->comm, h->stream));
    SB_CUDA(cudaMemcpyAsync(&ok, d_vote, 4, cudaMemcpyDeviceToHost, h->stream));
    SB_CUDA(cudaStreamSynchronize(h->stream));
    cudaFree(d_send);
    cudaFree(d_recv);
    if (!ok) {
        if (h->trace) fprintf(stderr, "SB_TRACE rank %d: peer mapping not available, merging with ncclAllReduce\n", h->rank);
        p2p_teardown(h);
        return SB_OK;
    }
    h->p2p.view = v;
    h->p2p.cap_elems = cap_bytes / sizeof(int64_t);
    h->p2p.ok = true;
    if (h->trace) fprintf(stderr, "SB_TRACE rank %d: accumulators in peer-mapped memory (%zu MB x 2): merge over NVLink loads\n", h->rank, mb);
    h->raw.alias(h->p2p.buf[0], h->p2p.cap_elems);
    h->raw_prev.alias(h->p2p.buf[1], h->p2p.cap_elems);
    h->prev_valid = false;
    return SB_OK;
}

SB_API int sb_dist_init(sb_handle* h, const uint8_t* id128, int32_t rank, int32_t world_size) {
    if (!h || !id128) return fail(SB_ERR_INVALID, "null argument");
    if (world_size < 1 || rank < 0 || rank >= world_size) return fail(SB_ERR_INVALID, "rank / world_size out of range");
    if (h->comm) return fail(SB_ERR_STATE, "sb_dist_init was already called on this handle");
    NcclApi& api = nccl_api();
    if (!api.ok) return fail(SB_ERR_UNSUPPORTED, "libnccl.so.2 not found");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    SB_NCCL(api.CommInitRank(&h->comm, world_size, id, rank));
    h->world = world_size;
    h->rank = rank;
    return p2p_setup(h);
}

SB_API int sb_set_profiling(sb_handle* h, int32_t on) {
    if (!h) return fail(SB_ERR_INVALID, "null handle");
    h->profiling = on != 0;
    return SB_OK;
}

SB_API int sb_hist_time(sb_handle* h, float* ms_total, int32_t* n_passes, int32_t* n_launches) {
    if (!h) return fail(SB_ERR_INVALID, "null handle");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    double ms = 0;
    int launches = 0;
    for (size_t i = 0; i < h->prof_events.size(); ++i) {
        SB_CUDA(cudaEventSynchronize(h->prof_events[i].second));
        float t = 0;
        SB_CUDA(cudaEventElapsedTime(&t, h->prof_events[i].first, h->prof_events[i].second));
        ms += t;
        launches += h->prof_launches[i];
        cudaEventDestroy(h->prof_events[i].first);
        cudaEventDestroy(h->prof_events[i].second);
    }
    if (ms_total) *ms_total = (float)ms;
    if (n_passes) *n_passes = (int32_t)h->prof_events.size();
    if (n_launches) *n_launches = launches;
    h->prof_events.clear();
    h->prof_launches.clear();
    return SB_OK;
}

SB_API int sb_raw_buffer(sb_handle* h, void** dev_ptr, int64_t* bytes) {
    if (!h || !dev_ptr) return fail(SB_ERR_INVALID, "null argument");
    *dev_ptr = h->raw.p;
    if (bytes) *bytes = 2 * h->layout.total * 8;
    return SB_OK;
}

SB_API int sb_acc_buffer(sb_handle* h, void** dev_ptr, int64_t* bytes) {
    if (!h || !dev_ptr) return fail(SB_ERR_INVALID, "null argument");
    *dev_ptr = h->d_acc;
    if (bytes) *bytes = 32;
    return SB_OK;
}

// Small results travel through one pinned block so that a pass ends with a single stream synchronisation.
struct PinnedOut {
    double acc[4];
    uint32_t flags[2];
    uint32_t maxbits[4];
};

static int pinned_reserve(sb_handle* h, size_t bytes) {
    if (bytes <= h->pin_bytes) return SB_OK;
    if (h->pin) cudaFreeHost(h->pin);
    h->pin = nullptr;
    h->pin_bytes = 0;
    const size_t cap = std::max<size_t>(bytes * 4, 1 << 16);  // pinned allocations synchronise the device: rarely
    SB_CUDA(cudaMallocHost(&h->pin, cap));
    h->pin_bytes = cap;
    return SB_OK;
}

static int finish_enqueue(sb_handle* h, sb_iter_out* out) {
    const int64_t total = h->layout.total, caller_total = h->layout.caller_total;
    if (out->stats && out->stats_len < caller_total) return fail(SB_ERR_INVALID, "stats buffer too small");
    if (!h->pairs.empty()) {  // the larger child of every sibling pair = parent - smaller child (after the merge)
        SB_CUDA(launch_pair_fixup(h->d_pairs.p, (int)h->pairs.size(), h->pair_max_len, h->d_small_is_b.p, h->raw_prev.p,
                                  h->raw_prev.p + h->prev_layout.total, h->raw.p, h->raw.p + total, h->stream));
        h->launches_count += 1;
        h->pairs.clear();
    }
    if (h->layout.full) {
        // the caller's blocks out of the internal all-features layout, one (node, listed feature) task at a time
        if (!h->split_planned) return fail(SB_ERR_STATE, "sb_compute_finish before sb_compute_stats");
        SB_CUDA(launch_finalize_tasks(h->raw.p, h->raw.p + total, h->tasks.p, h->n_tasks, h->all_feat_total, h->d_int_feat_off, h->stats_size,
                                      h->stat_kind, h->d_inv_scales, h->stats.p, h->stream));
    } else {
        SB_CUDA(launch_finalize(h->raw.p, h->raw.p + total, total, h->stats_size, h->stat_kind, h->d_inv_scales, h->stats.p, h->stream));
    }
    h->launches_count += 1;
    PinnedOut* po = static_cast<PinnedOut*>(h->pin);
    SB_CUDA(cudaMemcpyAsync(po->acc, h->d_acc, 32, cudaMemcpyDeviceToHost, h->stream));
    SB_CUDA(cudaMemcpyAsync(po->flags, h->d_error_flag, 4, cudaMemcpyDeviceToHost, h->stream));
    SB_CUDA(cudaMemcpyAsync(po->maxbits, h->d_maxbits, 12, cudaMemcpyDeviceToHost, h->stream));
    if (out->stats && caller_total > 0) {
        SB_CUDA(cudaMemcpyAsync(out->stats, h->stats.p, (size_t)caller_total * 8, cudaMemcpyDeviceToHost, h->stream));
        h->d2h_bytes += caller_total * 8;
    }
    h->d2h_bytes += 48;
    return SB_OK;
}

static int finish_check(sb_handle* h, sb_iter_out* out) {
    const PinnedOut* po = static_cast<const PinnedOut*>(h->pin);
    const uint32_t flags = po->flags[0];
    h->begun = false;
    if (flags & 1u) return fail(SB_ERR_RANGE, "class index (int)(label + 1e-6f) outside [0, num_classes)");
    if (flags & 2u) return fail(SB_ERR_RANGE, "a bin index is outside [0, n_bins) of its feature");
    if (flags & 4u) return fail(SB_ERR_STATE, "NextNode is null: a split node is missing a child (dt/DTWorker.java:1089-1094)");
    if (flags & 8u) return fail(SB_ERR_STATE, "internal: the chunk table of a row list overflowed");
    if (flags & 16u) return fail(SB_ERR_CUDA, "multi-GPU merge: a peer did not reach the barrier within thirty seconds");
    for (int k = 0; k < 3; ++k)
        if ((po->maxbits[k] & 0x7f800000u) == 0x7f800000u) return fail(SB_ERR_RANGE, "non-finite per-row contribution (label, significance or weight)");
    if ((po->maxbits[0] >> 23) >= 127 + 38) return fail(SB_ERR_RANGE, "significance * weight beyond 2^38");
    out->train_count = po->acc[0];
    out->validation_count = po->acc[1];
    out->train_error = po->acc[2];
    out->validation_error = po->acc[3];
    h->stats_ready = true;
    if (h->enable_subtraction && (!h->forest_pass || h->layout.full)) {  // keep this pass's merged accumulators for the children of its nodes (a forest pass pairs only when it is full)
        std::swap(h->raw, h->raw_prev);  // (the layouts are swapped by the next sb_compute_begin)
        h->prev_valid = true;
    }
    return SB_OK;
}

// device side of sb_best_splits; the records land in the pinned block behind PinnedOut
static int best_enqueue(sb_handle* h, bool want_bits) {
    const int n_nodes = (int)h->layout.nodes.size();
    if (n_nodes == 0) return SB_OK;
    if (!h->split_planned) {  // a caller that skipped sb_compute_stats
        int rc = plan_split_tasks(h);
        if (rc != SB_OK) return rc;
        SB_CUDA(h->stage.flush(h->stream));
        h->split_planned = true;
    }
    // stream-ordered allocations are not zero-filled: bytes the kernels leave alone (category bits past cat_bytes, the
    // record of a node without any candidate) must not carry an earlier pass's data to the caller
    SB_CUDA(cudaMemsetAsync(h->gain_out.p, 0, sizeof(sb_gain) * (size_t)n_nodes, h->stream));
    if (h->has_cat_tasks) SB_CUDA(cudaMemsetAsync(h->out_bits.p, 0, (size_t)n_nodes * SB_MAX_CAT_BYTES, h->stream));
    SplitScratch sc;
    sc.keys = h->sort_keys.p; sc.idx = h->sort_idx.p; sc.tmp = h->sort_tmp.p; sc.flags = h->sort_flags.p; sc.max_bins = h->max_bins;
    sc.next = h->d_cnt_shift + 2;  // third word of the same small allocation
    if (h->has_cat_tasks) SB_CUDA(cudaMemsetAsync(sc.next, 0, 4, h->stream));
    SB_CUDA(launch_best_splits(h->stats.p, h->tasks.p, h->n_tasks, n_nodes, h->cfg.impurity, h->stats_size,
                               h->cfg.min_instances_per_node, h->cfg.min_info_gain, h->cand.p,
                               h->has_cat_tasks ? h->cand_bits.p : nullptr, h->gain_out.p,
                               h->has_cat_tasks ? h->out_bits.p : nullptr, h->node_task_begin.p, sc, h->stream));
    h->launches_count += 2;
    uint8_t* dst = static_cast<uint8_t*>(h->pin) + 256;
    SB_CUDA(cudaMemcpyAsync(dst, h->gain_out.p, sizeof(sb_gain) * n_nodes, cudaMemcpyDeviceToHost, h->stream));
    h->d2h_bytes += sizeof(sb_gain) * n_nodes;
    if (want_bits && h->has_cat_tasks) {
        SB_CUDA(cudaMemcpyAsync(dst + sizeof(sb_gain) * n_nodes, h->out_bits.p, (size_t)n_nodes * SB_MAX_CAT_BYTES, cudaMemcpyDeviceToHost, h->stream));
        h->d2h_bytes += (int64_t)n_nodes * SB_MAX_CAT_BYTES;
    }
    return SB_OK;
}

static void best_deliver(sb_handle* h, sb_gain* gains, uint8_t* cat_bits) {
    const size_t n_nodes = h->layout.nodes.size();
    const uint8_t* src = static_cast<const uint8_t*>(h->pin) + 256;
    memcpy(gains, src, sizeof(sb_gain) * n_nodes);
    if (cat_bits && h->has_cat_tasks) memcpy(cat_bits, src + sizeof(sb_gain) * n_nodes, n_nodes * SB_MAX_CAT_BYTES);
}

static size_t pinned_need(const sb_handle* h) {
    // category bits only travel when the handle has a categorical feature at all
    return 256 + h->layout.nodes.size() * (sizeof(sb_gain) + (h->any_categorical ? SB_MAX_CAT_BYTES : 0));
}

SB_API int sb_compute_finish(sb_handle* h, sb_iter_out* out) {
    if (!h || !out) return fail(SB_ERR_INVALID, "null argument");
    if (!h->begun) return fail(SB_ERR_STATE, "sb_compute_finish before sb_compute_begin");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    int rc = pinned_reserve(h, pinned_need(h));
    if (rc != SB_OK) return rc;
    rc = finish_enqueue(h, out);
    if (rc != SB_OK) return rc;
    if (h->trace && !h->passes_trace.empty()) { cudaEventRecord(h->passes_trace.back().e1, h->stream); h->passes_trace.back().t_enq = now_ms(); }
    SB_CUDA(cudaStreamSynchronize(h->stream));
    if (h->trace && !h->passes_trace.empty()) h->passes_trace.back().t_sync = now_ms();
    return finish_check(h, out);
}

SB_API int sb_compute_finish_split(sb_handle* h, sb_iter_out* out, sb_gain* gains, uint8_t* cat_bits) {
    if (!h || !out || !gains) return fail(SB_ERR_INVALID, "null argument");
    if (!h->begun) return fail(SB_ERR_STATE, "sb_compute_finish_split before sb_compute_begin");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    int rc = pinned_reserve(h, pinned_need(h));
    if (rc != SB_OK) return rc;
    rc = finish_enqueue(h, out);
    if (rc != SB_OK) return rc;
    rc = best_enqueue(h, cat_bits != nullptr);
    if (rc != SB_OK) return rc;
    if (h->trace && !h->passes_trace.empty()) { cudaEventRecord(h->passes_trace.back().e1, h->stream); h->passes_trace.back().t_enq = now_ms(); }
    SB_CUDA(cudaStreamSynchronize(h->stream));
    if (h->trace && !h->passes_trace.empty()) h->passes_trace.back().t_sync = now_ms();
    rc = finish_check(h, out);
    if (rc != SB_OK) return rc;
    best_deliver(h, gains, cat_bits);
    return SB_OK;
}

SB_API int sb_compute(sb_handle* h, const sb_iter_in* in, sb_iter_out* out) {
    int rc = sb_compute_begin(h, in);
    if (rc != SB_OK) return rc;
    rc = sb_compute_stats(h);
    if (rc != SB_OK) return rc;
    return sb_compute_finish(h, out);
}

SB_API int sb_best_splits(sb_handle* h, sb_gain* gains, uint8_t* cat_bits) {
    if (!h || !gains) return fail(SB_ERR_INVALID, "null argument");
    if (!h->stats_ready) return fail(SB_ERR_STATE, "sb_best_splits before sb_compute_finish");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    if (h->layout.nodes.empty()) return SB_OK;
    int rc = pinned_reserve(h, pinned_need(h));
    if (rc != SB_OK) return rc;
    rc = best_enqueue(h, cat_bits != nullptr);
    if (rc != SB_OK) return rc;
    SB_CUDA(cudaStreamSynchronize(h->stream));
    best_deliver(h, gains, cat_bits);
    return SB_OK;
}

SB_API int sb_time_stats_pass(sb_handle* h, int32_t n_repeat, float* ms_hist_kernel, float* ms_total) {
    if (!h || n_repeat < 1) return fail(SB_ERR_INVALID, "bad argument");
    if (!h->begun) return fail(SB_ERR_STATE, "sb_time_stats_pass before sb_compute_begin");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    cudaEvent_t e0, e1;
    SB_CUDA(cudaEventCreate(&e0));
    SB_CUDA(cudaEventCreate(&e1));
    const int64_t total = h->layout.total;
    double sum_ms = 0;
    for (int r = 0; r < n_repeat; ++r) {
        SB_CUDA(cudaMemsetAsync(h->raw.p, 0, (size_t)(2 * total) * 8, h->stream));
        SB_CUDA(cudaEventRecord(e0, h->stream));
        int rc = run_histograms(h);
        if (rc != SB_OK) return rc;
        SB_CUDA(cudaEventRecord(e1, h->stream));
        SB_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        SB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        sum_ms += ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (ms_hist_kernel) *ms_hist_kernel = (float)(sum_ms / n_repeat);
    if (ms_total) *ms_total = (float)(sum_ms / n_repeat);
    return SB_OK;
}

SB_API int sb_rows_streamed(sb_handle* h, int64_t* rows, int32_t reset) {
    if (!h || !rows) return fail(SB_ERR_INVALID, "null argument");
    SB_CUDA(cudaSetDevice(h->cfg.device));
    unsigned long long dev = 0;
    SB_CUDA(cudaMemcpyAsync(&dev, h->d_rows_streamed, 8, cudaMemcpyDeviceToHost, h->stream));
    SB_CUDA(cudaStreamSynchronize(h->stream));
    *rows = (int64_t)dev + h->rows_streamed_host;
    if (reset) {
        SB_CUDA(cudaMemsetAsync(h->d_rows_streamed, 0, 8, h->stream));
        h->rows_streamed_host = 0;
    }
    return SB_OK;
}

SB_API int sb_counters(sb_handle* h, int64_t* launches, int64_t* h2d_bytes, int64_t* d2h_bytes) {
    if (!h) return fail(SB_ERR_INVALID, "null handle");
    if (launches) *launches = h->launches_count;
    if (h2d_bytes) *h2d_bytes = h->h2d_bytes;
    if (d2h_bytes) *d2h_bytes = h->d2h_bytes;
    return SB_OK;
}

}  // extern "C"
