// kernels.cuh -- launch wrappers of the device kernels (definitions in hist.cu, rows.cu, split.cu).
#pragma once
#include "common.cuh"

namespace sb {

constexpr int HIST_T_ROWS = 128;
constexpr int HIST_STAGES = 3;
constexpr int HIST_THREADS = 640;  // 4 producer + 16 consumer warps

struct HistArgs {
    const int32_t* feat_bins;    // [F] stats bins per feature slot
    const float* output;         // Data.output (the `label` argument of featureUpdate, dt/DTWorker.java:851)
    const float* significance;
    const float* weights;        // subsampleWeights[treeId % len] as one column, or nullptr (all 1f)
    const int32_t* node_id;      // current heap id per row in the tree being accumulated
    const int16_t* slot_of_node; // heap id -> todo slot (or -1)
    int slot_table_size;
    int64_t n_rows;
    int n_features;
    const int32_t* tile_feats;   // [n_tiles][32] feature slot of each lane (tile_slot0 + lane), or -1 if unused
    const int32_t* tile_slot0;   // [n_tiles] first feature slot of the tile (multiple of 32)
    const uint32_t* slot_fmask;  // [n_slots][n_tiles] bit l: tile feature l is in the slot's feature list
    const int64_t* pair_off;     // [n_slots][F] offset (in statistics) of the (slot, feature) block, -1 if absent
    int n_tiles, n_groups, gs, n_slots, bins_pad;
    int64_t rows_per_chunk;
    int stats_size;              // 3 (variance) or numClasses
    const float* scales;         // [3] 2^e per statistic (count, sum, sum of squares)
    const int* cnt_shift;        // device: e of the count statistic (integer-count mode shifts by it when merging)
    int64_t* raw_lo;             // global accumulators, one (lo, hi) pair per statistic
    int64_t* raw_hi;
    uint32_t* error_flag;        // bit0: class index out of range, bit1: bin index out of range
};

// ---- row-list path (rowlist.cu + hist_rows_kernel in hist.cu): passes with more todo nodes than one slot group
struct __align__(16) RowOps {   // one active row of the pass, in slot-group order
    // (slot, cnt_lo) share an aligned 8-byte word: with integer counts a consumer needs those two and q[0..3],
    // one 64-bit and one 128-bit broadcast load per row
    uint32_t row;               // row index (ascending inside a group)
    uint32_t cnt_hi;
    uint32_t slot;              // slot inside its group, 0..gs-1
    uint32_t cnt_lo;
    uint32_t q[4];              // VARIANCE: (sum_lo, sum_hi, sq_lo, sq_hi); CLASS: (class byte offset, -, -, -)
};
// Variance statistics with integer counts (the GBT default): the high words of the two fixed-point operands need 13
// bits (|q| < 2^44), so their upper halves carry what a consumer would otherwise fetch with a second shared-memory
// load -- q[1] = (count << 16) | (sum_hi & 0xffff), q[3] = (slot << 16) | (sq_hi & 0xffff): ONE 128-bit broadcast load per
// row and tile gives a consumer warp everything (the shared-memory wavefronts of these loads are a fifth of the
// kernel's LSU work).  A row that must be skipped has slot 0xffff.
__host__ __device__ inline uint32_t rowops_pack(uint32_t hi_word, uint32_t upper) { return (hi_word & 0xffffu) | (upper << 16); }
__host__ __device__ inline uint32_t rowops_hi_signed(uint32_t packed) { return (uint32_t)((int32_t)(packed << 16) >> 16); }
__host__ __device__ inline uint32_t rowops_hi_unsigned(uint32_t packed) { return packed & 0xffffu; }
__host__ __device__ inline uint32_t rowops_upper(uint32_t packed) { return packed >> 16; }

struct TreeView {  // DeviceTree as kernels see it
    const int32_t* feat;
    const int32_t* mask_off;
    const double* predict;
    const uint8_t* has_predict;
    const uint8_t* masks;
    int capacity;
};

struct RowListArgs {
    int64_t n;
    const int32_t* node_id;
    const float* output;
    const float* significance;
    const float* weights;
    const int16_t* slot_of_node;
    int slot_table_size, gs, n_groups, stats_size, kind, cnt_int;
    const float* scales;
    // route == 1: the counting pass first walks every row down the current tree from its node (predictNodeIndex,
    // dt/DTWorker.java:1036-1096) and stores the new node id, instead of a separate routing pass over the rows
    int route;
    TreeView tree;
    const void* const* cols;
    const int32_t* feat_bins;
    int32_t* node_id_rw;
    uint32_t* counts;          // [n_groups * n_ranges + 1] active rows per (group, range of 1024 rows), group-major
    uint32_t* scanned;         // exclusive prefix sum of counts: first position of each (group, range)
    void* temp;
    size_t temp_bytes;
    int64_t chunk_rows;        // upper bound of rows per chunk (accumulator headroom)
    int identity;              // 1: no sort -- the list is all rows in storage order (single todo node that is a tree root)
    int target_chunks;         // the chunk size is chosen on the device so that about this many chunks result
    int n_tiles;               // feature tiles per chunk (blocks = chunks x tiles)
    int64_t* offsets;          // [n_groups + 1]
    int32_t* chunk_group;      // [max_chunks]
    int64_t* chunk_begin;
    int64_t* chunk_end;
    int32_t* n_chunks;         // device scalar
    int max_chunks;
    RowOps* ops;               // [n]
    uint32_t* error_flag;
    unsigned long long* rows_streamed;  // device counter: += active rows of this list (what the statistics kernels read)
};
cudaError_t launch_transpose_bins(const uint8_t* cm, int64_t pitch, int n_features, const int32_t* feat_of_pos, int64_t n, uint8_t* rm, int fpad,
                                  cudaStream_t s);
// wide (uint16) columns [n_wide][pitch16] -> row-major [n][wpad]
cudaError_t launch_transpose_bins16(const uint16_t* cm, int64_t pitch16, int n_wide, int64_t n, uint16_t* rm, int wpad,
                                    cudaStream_t s);

int64_t rowlist_n_ranges(int64_t n);
size_t rowlist_scan_temp_bytes(int64_t items);
cudaError_t launch_rowlist_build(const RowListArgs& a, cudaStream_t s);
struct HistRowsArgs {
    const uint8_t* rm_bins;      // row-major bins [n][fpad]
    int fpad;
    const RowOps* ops;
    const int32_t* chunk_group;
    const int64_t* chunk_begin;
    const int64_t* chunk_end;
    const int32_t* n_chunks;
    const int32_t* tile_feats;   // [n_tiles][32*fpl]: entry k*32+lane = feature held at byte tile_slot0 + fpl*lane + k of a twin row, or -1
    const int32_t* tile_slot0;   // [n_tiles] byte position of the tile in a row of the twin (multiple of 32*fpl)
    const int32_t* tile_kact;    // [n_tiles] features per lane actually populated (<= fpl): the k loop stops there
    const uint32_t* slot_fmask;  // [n_slots][n_tiles][fpl]: word k, bit lane: that feature is in the slot's list
    const int64_t* pair_off;     // [n_slots][F]
    int n_tiles, gs, n_slots, bins_pad, n_features, stats_size;
    const int* cnt_shift;
    int64_t* raw_lo;
    int64_t* raw_hi;
};
// fpl = features per lane (1, 2 or 4); allf = every slot takes every feature of the tile (skips the mask lookups)
cudaError_t hist_rows_launch(int kind, int cnt_int, int fpl, int allf, const HistRowsArgs& a, int grid, size_t smem,
                             cudaStream_t stream);
// hist_rows_wide_kernel: tiles of 8 wide features, lane = (row slot, feature); args as hist_rows (rm_bins = uint16 twin,
// fpad = wpad in elements, tile tables with 8 entries per tile)
cudaError_t hist_rows_wide_launch(int kind, int cnt_int, const HistRowsArgs& a, int grid, size_t smem, cudaStream_t stream);
size_t hist_rows_wide_smem_bytes(int gs, int bins_pad, int nw);
int hist_rows_wide_max_slots_per_block(int bins_pad, int nw);
size_t hist_rows_smem_bytes(int gs, int bins_pad, int nw, int fpl);
int hist_rows_max_slots_per_block(int bins_pad, int nw, int fpl);

size_t hist_smem_bytes(int gs, int bins_pad, int nw);
int hist_max_slots_per_block(int bins_pad, int nw);
int hist_words_per_bin(int kind, int cnt_int, int stats_size);
// tensor_map: a CUtensorMap (128 bytes) over the column-major uint8 bin matrix [F][pitch], box {128 rows, 32 features}
cudaError_t hist_launch(int kind, int cnt_int, const HistArgs& a, const void* tensor_map, int grid, size_t smem,
                        cudaStream_t stream);

// ---- forest passes (RF): the todo nodes of ALL trees of a pass in one set of launches -- one list keyed by global slot
// group (slots numbered over the trees of the pass), one statistics launch (dt/DTWorker.java:810-826 walks every tree per
// row in one pass over the rows).  rowlist.cu: forest_* kernels; hist.cu: hist_rows_class_kernel.
struct __align__(32) WalkNode {   // everything the walk wants to know about a node, in one 32-byte sector
    int32_t feat;                 // split feature slot if the node routes further, else -1
    int32_t pos;                  // byte position of that feature in a row of the twin, -1: read the column
    uint32_t mask_off;            // left_mask in the tree's mask pool (used when the feature has more than 64 bins)
    float predict;                // (float) Node.predict
    uint32_t flags;               // bit 0: the node has a predict; bit 1: mask64 holds the left_mask
    uint32_t pad;
    uint64_t mask64;              // bit b: bin b goes left
};
struct WalkTree {                 // a real tree of the message, for the walk + error sums (rows.cu: forest_walk_kernel)
    const WalkNode* nodes;        // [capacity], by heap id
    const uint8_t* masks;
    int32_t* node_id;             // [n] node of every row in this tree
    const float* weights;         // [n] subsampleWeights[treeId] of every row, nullptr: all 1f (adopted rows, validation rows)
    int32_t capacity;
    int32_t pad;
};
struct ListTree {                 // a tree with todo nodes in this pass
    const int32_t* node_id;
    const float* weights;
    const int32_t* slot_of_node;  // [slot_table_size] heap id -> GLOBAL slot of the pass, -1: not a todo node
    int32_t slot_table_size;
    int32_t pad;
};
struct __align__(8) ClassOps {    // an active (row, tree) pair of an integer-count class pass (Gini / Entropy)
    uint32_t row;
    uint32_t where;               // low 18 bits: byte offset of the (slot inside its group, class) words in the block's
                                  // histogram; upper 14 bits: significance * weight, an integer in [0, 1024]
};
constexpr uint32_t CLASSOPS_OFF_BITS = 18;  // 227 KB of shared memory < 2^18 bytes
constexpr int FOREST_RANGE = 4096;   // consecutive rows counted / scattered by one block
constexpr int FOREST_MAX_KEYS = 24576;  // slot groups of one batch of trees (counters in shared memory, 96 KB)
struct ForestListArgs {
    int64_t n;
    const ListTree* trees;
    int n_trees;
    int slot_base;                // first global slot of this batch of trees
    int gs, n_keys;               // slots per group, groups of the batch: key = (slot - slot_base) / gs
    int64_t n_ranges;
    uint32_t* counts;             // [n_keys * n_ranges + 1], key-major
    uint32_t* scanned;
    void* temp;
    size_t temp_bytes;
    const float* output;
    const float* significance;
    const float* scales;
    int stats_size, kind, cnt_int;
    int slim;                     // 1: ops are ClassOps (kind == STAT_CLASS with integer counts), 0: RowOps
    uint32_t slot_stride, class_stride;  // slim: bytes between slots / classes in the block's histogram
    void* ops;
    int64_t chunk_rows;           // upper bound (accumulator headroom)
    int target_chunks;
    int64_t row_bytes;            // bytes of a twin row (what a pair reads)
    int64_t slab_bytes;           // bytes of the twin a slab of rows may span (a fraction of L2)
    int64_t stats_bytes;          // bytes of the pass's global accumulators (what the chunks of one slab merge into)
    int32_t* chunk_group;
    int64_t* chunk_begin;
    int64_t* chunk_end;
    int32_t* n_chunks;            // [3]: number of chunks; work counters of the persistent statistics kernels
    int max_chunks;
    uint32_t* error_flag;
    unsigned long long* rows_streamed;
};
int64_t forest_n_ranges(int64_t n);
cudaError_t launch_forest_lists(const ForestListArgs& a, cudaStream_t s);
// pair_off[s][f] for every slot from the per-node feature lists (CSR in todo order): one warp per slot
cudaError_t launch_build_pair_off(const int32_t* slot_node, int n_slots, const int32_t* feats, const int32_t* begin,
                                  const int64_t* node_off, const int32_t* feat_bins, int stats_size, int n_features,
                                  int64_t* pair_off, cudaStream_t s);
// full passes: pair_off[s][f] = slot_node[s] * node_stride + feat_off[f] (every valid feature of every node), -1 where feat_off is
cudaError_t launch_fill_pair_off_full(const int32_t* slot_node, int n_slots, int64_t node_stride, const int64_t* feat_off, int n_features,
                                      int64_t* pair_off, cudaStream_t s);
struct HistClassArgs {
    const uint8_t* rm_bins;
    uint32_t fpad;
    const ClassOps* ops;
    const int32_t* chunk_group;
    const int64_t* chunk_begin;
    const int64_t* chunk_end;
    int32_t* n_chunks;            // [3]: number of chunks, work counters of the narrow and the wide launch (zeroed by forest_chunks_kernel)
    const int32_t* tile_feats;
    const int32_t* tile_slot0;
    const int32_t* tile_kact;
    const int64_t* pair_off;      // [n_slots][F], first row = the batch's first slot
    int n_tiles, gs, n_slots, bins_pad, n_features, stats_size;
    const int* cnt_shift;
    int64_t* raw_lo;
    int64_t* raw_hi;
    // wide tiles only (hist_rows_class_wide_kernel): how to take a ClassOps offset apart -- it was built for the narrow
    // tiles' histogram, (slot * slot_words + class) << word_shift bytes
    uint32_t word_shift, slot_words;
    int counter;                  // which word of n_chunks is this launch's work counter (1: narrow tiles, 2: wide tiles)
};
cudaError_t hist_rows_class_launch(int fpl, const HistClassArgs& a, int grid, size_t smem, cudaStream_t stream);
// tiles of 8 uint16 features over the same ClassOps list; rm_bins = the uint16 twin, fpad = its pitch in elements
cudaError_t hist_rows_class_wide_launch(const HistClassArgs& a, int grid, size_t smem, cudaStream_t stream);
size_t hist_rows_class_wide_smem_bytes(int gs, int bins_pad, int n_classes);
int hist_rows_class_wide_max_slots_per_block(int bins_pad, int n_classes);
size_t hist_rows_class_smem_bytes(int gs, int bins_pad, int n_classes, int fpl);
int hist_rows_class_max_slots_per_block(int bins_pad, int n_classes, int fpl);

// ---- rows.cu

cudaError_t launch_fill_i32(int32_t* p, int64_t n, int32_t v, cudaStream_t s);
cudaError_t launch_fill_f32(float* p, int64_t n, float v, cudaStream_t s);
cudaError_t launch_copy_f32(float* dst, const float* src, int64_t n, cudaStream_t s);
// predictNodeIndex (dt/DTWorker.java:1036-1096) continued from each row's current node id.
cudaError_t launch_route(const void* const* cols, const int32_t* feat_bins, int64_t n, int32_t* node_id, TreeView t,
                         uint32_t* error_flag, cudaStream_t s);
cudaError_t launch_synthetic16(uint16_t* col, int64_t n, int feature, int n_features, int n_bins, int64_t row0,
                               uint64_t seed, cudaStream_t s);
// GBT tree switch (dt/DTWorker.java:620-669, 721-735): predict / output update from the finished tree's leaves.
cudaError_t launch_gbt_update(int64_t n, const int32_t* node_id, TreeView prev, int is_first_tree, double learning_rate,
                              int loss, const float* label, float* predict, float* output, const uint8_t* dropout,
                              cudaStream_t s);
// Error sums (dt/DTWorker.java:672-683, 737-748 GBT; 583-605, 691-705 RF).  acc: 4 doubles
// (train_count, validation_count, train_error, validation_error), accumulated via per-block partials in `scratch`.
struct ErrArgs {
    int64_t n;
    const float* label;
    const float* predict;      // Data.predict (used when from_tree == 0)
    const float* significance;
    const float* weights;      // RF training rows: subsampleWeights[treeId] column; nullptr otherwise
    const int32_t* node_id;    // used when from_tree == 1
    TreeView tree;
    int from_tree;             // 1: error of the current node's predict (skipped when the node has none)
    int loss;
    int is_validation;         // validation rows: everything goes to (validation_count, validation_error)
    int rf_oob;                // RF training rows: weight == 0 -> out-of-bag "validation" sums
};
// GBT tree switch, fused (gbt_switch_kernel): route through the finished tree, update predict / output, reset the node id,
// error sums over the new predict, and (training rows) the maxima behind the next tree's fixed-point scales.
struct SwitchArgs {
    int64_t n;
    const void* const* cols;
    const int32_t* feat_bins;
    int32_t* node_id;
    TreeView prev;             // the finished tree
    int is_first_tree;
    double learning_rate;
    int loss;
    const float* label;
    float* predict;
    float* output;
    const float* significance;
    const float* weights;      // training rows: the next tree's sample-weight column, nullptr = all 1f
    const uint8_t* dropout;    // training rows, may be nullptr
    int is_validation;
    uint32_t* maxbits;         // [3], training rows only
    uint32_t* error_flag;
};
cudaError_t launch_gbt_switch(const SwitchArgs& a, double* scratch /*[ERR_SCRATCH_DOUBLES]*/, double* acc /*[4], += */, cudaStream_t s);
constexpr int ERR_BLOCKS = 1184;  // 8 per SM
constexpr int ERR_SCRATCH_DOUBLES = ERR_BLOCKS * 4 + 1;  // per-block partials + the finish ticket (zeroed once at allocation)
cudaError_t launch_errors(const ErrArgs& a, double* scratch /*[ERR_SCRATCH_DOUBLES]*/, double* acc /*[4], += */, cudaStream_t s);
// RF: predictNodeIndex continued from every row's node in every tree + the error sums of the pass, one launch
// (dt/DTWorker.java:583-605 training rows, 691-705 validation rows).  rm_bins / pos_of: the row-major twin of the bin
// matrix and feature -> byte position in a twin row, or nullptr (then the column-major columns are read).
struct ForestWalkArgs {
    int64_t n;
    const WalkTree* trees;
    int n_trees;
    const void* const* cols;
    const int32_t* feat_bins;
    const uint8_t* rm_bins;       // row-major twin (the positions are in the nodes), or nullptr
    uint32_t fpad;
    const float* label;
    const float* significance;
    int loss;
    int is_validation;
    uint32_t* error_flag;
};
cudaError_t launch_forest_walk(const ForestWalkArgs& a, double* scratch, double* acc, cudaStream_t s);
// max |sig*w|, |y*sig*w|, |y*y*sig*w| over rows (as fp32 bit patterns, atomicMax) -> out[3]
cudaError_t launch_absmax(int64_t n, const float* output, const float* significance, const float* weights,
                          uint32_t* out, cudaStream_t s);
// out[3] (bit patterns of maxima) -> scales[3] = 2^e, inv_scales[3] = 2^-e (doubles), cnt_shift = max(e[0], 0)
cudaError_t launch_make_scales(const uint32_t* maxbits, int kind, float* scales, double* inv_scales, int* cnt_shift,
                               cudaStream_t s);
// flag |= 1 unless every sig*w product is an integer in [0, 1024]
cudaError_t launch_check_int_weights(int64_t n, const float* significance, const float* weights, uint32_t* flag,
                                     cudaStream_t s);
cudaError_t launch_synthetic(uint8_t* bins, int64_t pitch, int n_features, const int32_t* feat_bins, float* label,
                             float* significance, int64_t row0, int64_t n, uint64_t seed, int label_kind, int weighted,
                             cudaStream_t s);
cudaError_t launch_check_bins(const void* const* cols, const int32_t* feat_bins, int n_features, int64_t n,
                              uint32_t* error_flag, cudaStream_t s);

// ---- prep.cu
cudaError_t launch_java_bagging_mask(float* weights, int64_t n, long long seed, double rate, long long first_draw, cudaStream_t s);
size_t stratified_temp_bytes(int64_t n);
cudaError_t launch_stratified_mask(float* weights, const float* label, int64_t n, int n_classes, long long seed, double rate,
                                   uint32_t* ind, uint32_t* ordinal, void* temp, size_t temp_bytes, uint32_t* error_flag,
                                   cudaStream_t s);
cudaError_t launch_poisson_bags(float* weights, int64_t n, unsigned long long seed, long long row0, int tree, double rate, double p0,
                                cudaStream_t s);

// ---- p2p.cu: all-reduce of the integer accumulators over NVLink peer memory (GPUs of one box)
constexpr int P2P_MAX_WORLD = 8;
struct P2PView {
    int world, rank;
    int64_t* peer_buf[2][P2P_MAX_WORLD];   // the two accumulator buffers (this pass / previous pass) of every rank, as mapped here
    uint32_t* peer_flags[P2P_MAX_WORLD];   // every rank's barrier words: [3 barriers][P2P_MAX_WORLD arrivals]
};
cudaError_t launch_p2p_allreduce(const P2PView& v, int which, int64_t n_elems, uint32_t epoch, int tail_barrier, uint32_t* error_flag,
                                 cudaStream_t s);

// ---- split.cu
// raw (lo, hi) integer accumulators -> the reference's double statistics (one rounding each).
cudaError_t launch_finalize(const int64_t* raw_lo, const int64_t* raw_hi, int64_t total, int stats_size, int kind,
                            const double* inv_scales, double* stats, cudaStream_t s);
// Sibling pair whose parent's merged accumulators are still on the device (histogram subtraction).
struct PairDesc {
    int64_t parent_split_block;  // offset (in statistics) of the parent's block of its own split feature, in the previous pass
    int64_t parent_off;          // offset of the parent's node block in the previous pass
    int64_t off_a, off_b;        // offsets of the two children's node blocks in this pass (same feature lists => same layout)
    int64_t len;                 // statistics per node block
    const uint8_t* left_mask;    // device: the parent's left_mask (bit b: bin b goes to child A = 2*id)
    int32_t n_bins;              // bins of the parent's split feature
    int32_t node_a, node_b;      // heap ids 2p, 2p+1
    int32_t slot;                // global slot index of the pair (row of pair_off)
    int32_t table_off;           // offset of the tree's slot table inside the slot-table buffer
    int32_t pad;
};
// the same for a full pass: task t's block is read at tasks[t].node * node_stride + feat_off[tasks[t].feature] of the
// internal (every valid feature) layout and written at tasks[t].offset of the caller's
struct SplitTask;
cudaError_t launch_finalize_tasks(const int64_t* raw_lo, const int64_t* raw_hi, const SplitTask* tasks, int n_tasks, int64_t node_stride,
                                  const int64_t* feat_off, int stats_size, int kind, const double* inv_scales, double* stats,
                                  cudaStream_t s);
cudaError_t launch_pair_decide(const PairDesc* pairs, int n_pairs, const int64_t* prev_lo, const int64_t* prev_hi,
                               int stats_size, int kind, int n_features, int16_t* slot_tables, int64_t* pair_off,
                               int32_t* small_is_b, cudaStream_t s);
// the same over 32-bit slot tables (forest passes)
cudaError_t launch_pair_decide32(const PairDesc* pairs, int n_pairs, const int64_t* prev_lo, const int64_t* prev_hi,
                                 int stats_size, int kind, int n_features, int32_t* slot_tables, int64_t* pair_off,
                                 int32_t* small_is_b, cudaStream_t s);
cudaError_t launch_pair_fixup(const PairDesc* pairs, int n_pairs, int64_t max_len, const int32_t* small_is_b,
                              const int64_t* prev_lo, const int64_t* prev_hi, int64_t* lo, int64_t* hi, cudaStream_t s);
struct SplitTask {       // one (todo node, feature) pair
    int32_t node;        // todo index
    int32_t feature;     // feature slot
    int64_t offset;      // offset of the feature block in stats
    int32_t n_bins;
    int32_t categorical;
    int32_t visit_rank;  // position in the reference's HashMap iteration order (dt/DTMaster.java:324)
    int32_t pad;
};
cudaError_t launch_expand_tasks(const SplitTask* tmpl, int fk, const int64_t* node_off, int n_nodes, SplitTask* tasks,
                                int32_t* begin, cudaStream_t s);
// per-node feature lists -> (node, feature) tasks on the device; max_list = the longest list (<= 8192 features)
cudaError_t launch_build_tasks(const int32_t* feats, const int32_t* begin, const int64_t* node_off, int n_nodes, int max_list,
                               const int32_t* feat_bins, const uint8_t* is_cat, const int32_t* column_num, int stats_size,
                               SplitTask* tasks, cudaStream_t s);
cudaError_t launch_counts_cache(int64_t* raw_lo, int64_t* raw_hi, int64_t* cache, int64_t n_bins_total, int stats_size, int save,
                                cudaStream_t s);
struct SplitScratch {    // categorical features only: max_bins entries per CATEGORICAL task; a task claims its piece from `next`
    double* keys = nullptr;
    int* idx = nullptr;
    int* tmp = nullptr;
    uint8_t* flags = nullptr;
    int max_bins = 0;
    int* next = nullptr;     // device counter, zero before the launch
};
cudaError_t launch_best_splits(const double* stats, const SplitTask* tasks, int n_tasks, int n_nodes, int impurity,
                               int stats_size, int min_instances, double min_info_gain, sb_gain* cand /*[n_tasks]*/,
                               uint8_t* cand_bits /*[n_tasks*SB_MAX_CAT_BYTES] or null*/, sb_gain* out /*[n_nodes]*/,
                               uint8_t* out_bits, const int32_t* node_task_begin /*[n_nodes+1]*/,
                               const SplitScratch& sc, cudaStream_t s);

}  // namespace sb
