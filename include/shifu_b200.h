/*
 * shifu_b200.h — C ABI of the B200-native tree-ensemble training hot path.
 *
 * This is the drop-in boundary for Shifu's GBT / RF trainer (reference: DevinWu/shifu, package
 * ml.shifu.shifu.core.dtrain.dt).  The reference has no FFI for this path: the boundary is the Java
 * class surface DTWorker / DTMaster / Impurity / TreeNode.  Every entry point below cites the Java
 * method whose body it replaces; INTEGRATION.md shows the JNI stub a Shifu maintainer adds.
 *
 * Conventions
 *   - plain C, plain pointers and sizes; no C++ or torch types cross this boundary;
 *   - every function returns 0 on success and a negative sb_status on failure; sb_last_error()
 *     returns the thread-local message (the JNI layer turns it into RuntimeException(msg), which is
 *     what the reference throws: dt/DTWorker.java:359,878,1229);
 *   - a handle drives ONE GPU (one Guagua worker == one GPU == one process); it is not thread-safe,
 *     like a DTWorker instance (dt/DTWorker.java:338-349 is called from one Guagua thread);
 *   - out-buffers are caller-allocated; the library keeps no caller pointer after a call returns,
 *     except device pointers registered through sb_load_rows_device;
 *   - there is no CPU fallback: without a CUDA device sb_create fails with SB_ERR_CUDA.
 *
 * Encodings that leak through from the reference (kept identical so BinaryDTSerializer output is
 * unchanged): Node.ROOT_INDEX = 1, Node.INVALID_INDEX = -1 (dt/Node.java:249,254); heap numbering
 * left = 2*id, right = 2*id+1 (dt/Node.java:534-558); Split.CONTINUOUS = 1, Split.CATEGORICAL = 2
 * (dt/Split.java:63-64); SimpleBitSet = byte array, value v is bit (v % 8) of byte (v >> 3)
 * (dt/SimpleBitSet.java:57-68,89-101).
 */
#ifndef SHIFU_B200_H
#define SHIFU_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SB_API __attribute__((visibility("default")))

typedef enum sb_status {
    SB_OK = 0,
    SB_ERR_INVALID = -1,   /* bad argument / inconsistent config (reference: RuntimeException / IllegalStateException) */
    SB_ERR_CUDA = -2,      /* CUDA runtime error, or no device */
    SB_ERR_STATE = -3,     /* call order violated (e.g. compute before load) */
    SB_ERR_RANGE = -4,     /* bin index outside [0, n_bins) (reference: ArrayIndexOutOfBounds in featureUpdate) */
    SB_ERR_NOMEM = -5,
    SB_ERR_UNSUPPORTED = -6
} sb_status;

/* Impurity kinds: dt/Impurity.java (Variance :103, FriedmanMSE :255, Entropy :368, Gini :553). */
enum { SB_IMPURITY_VARIANCE = 0, SB_IMPURITY_FRIEDMANMSE = 1, SB_IMPURITY_ENTROPY = 2, SB_IMPURITY_GINI = 3 };
/* Loss kinds: dt/Loss.java (Absolute :36, Squared :55, HalfGradSquared :75, Log :90). */
enum { SB_LOSS_SQUARED = 0, SB_LOSS_HALFGRADSQUARED = 1, SB_LOSS_ABSOLUTE = 2, SB_LOSS_LOG = 3 };
/* Split.featureType: dt/Split.java:63-64. */
enum { SB_SPLIT_CONTINUOUS = 1, SB_SPLIT_CATEGORICAL = 2 };

#define SB_ROOT_INDEX 1
#define SB_INVALID_INDEX (-1)

/* sb_config.flags */
#define SB_FLAG_DETERMINISTIC 1u /* reserved: the histogram path is order-independent by construction */
#define SB_FLAG_NO_SUBTRACTION 4u /* accumulate every todo node directly instead of deriving the larger child of a
                                     sibling pair as parent - smaller child (same numbers either way); nodes with
                                     feature lists of their own are then accumulated for the listed features only */
#define SB_FLAG_NO_ROOT_COUNT_CACHE 8u /* accumulate the bin counts of a tree root every time instead of reusing those of the
                                          previous root pass while the sample weights are unchanged */
#define SB_FLAG_NO_FOREST_PASS 16u /* RF: one set of launches per tree and pass instead of one for all trees of the pass
                                      (same numbers either way; the forest pass needs room for one operand record per
                                      active (row, tree) pair) */
#define SB_FLAG_NO_ROW_LISTS 2u  /* never build per-pass row lists / the row-major bin copy: passes with more todo
                                    nodes than one shared-memory slot group then stream the column tiles once per group
                                    (slower, but needs no extra memory) */

typedef struct sb_handle sb_handle;

/*
 * What DTWorker.init / DTMaster.init read from ModelConfig.json, ColumnConfig.json and train.params
 * (dt/DTWorker.java:349-550, dt/DTMaster.java:896-1116), flattened.  A "feature slot" is the position
 * of a selected column in Data.inputs (DTWorker.inputIndexMap, dt/DTWorker.java:224,1148).
 */
typedef struct sb_config {
    int32_t n_features;            /* DTWorker.inputCount */
    const int32_t* n_bins;         /* [n_features] stats bins: numeric binBoundary.size(), categorical
                                      binCategory.size()+1 (dt/DTWorker.java:915-919) */
    const uint8_t* is_categorical; /* [n_features] */
    const int32_t* column_num;     /* [n_features] ColumnConfig.columnNum of each slot */
    const double* bin_boundary;    /* concatenated ColumnConfig.binBoundary of the numeric slots (n_bins[f] doubles
                                      each, in slot order); may be NULL if the caller builds sb_node.left_mask itself */
    int32_t impurity;              /* SB_IMPURITY_*  (train.params "Impurity") */
    int32_t num_classes;           /* Entropy/Gini statsSize (dt/DTWorker.java:462); ignored for variance */
    int32_t loss;                  /* SB_LOSS_*      (train.params "Loss") */
    int32_t is_gbdt;               /* 1 = GBT, 0 = RF (dt/DTWorker.java:376-377) */
    int32_t tree_num;              /* "TreeNum" */
    int32_t weights_per_row;       /* Data.subsampleWeights.length: 1 or tree_num (dt/DTWorker.java:1521-1547) */
    int32_t max_depth;             /* "MaxDepth" */
    int32_t min_instances_per_node;/* "MinInstancesPerNode" */
    double min_info_gain;          /* "MinInfoGain" */
    double learning_rate;          /* "LearningRate" */
    int32_t device;                /* CUDA device ordinal this handle drives */
    uint32_t flags;
} sb_config;

/*
 * One node of a flattened tree == dt/Node.java fields the worker needs.  Children are addressed by
 * index into the same array (-1 = none).  A node is a leaf for routing when it has no split or no
 * children, exactly as dt/DTWorker.java:1041.
 *
 * Routing (dt/DTWorker.java:1052-1088) is carried as `left_mask`: a SimpleBitSet-encoded byte string
 * over this feature's stats bins, bit b set <=> a row whose bin index is b goes LEFT.  The host builds
 * it from the reference's own rule: numeric  bit b = (binBoundary[b] < threshold)   (:1057-1062);
 * categorical  idx = (0 <= bin < nCats) ? bin : nCats,  bit = (bitset.contains(idx) == isLeft)
 * (:1064-1087).  Carrying the mask instead of the threshold keeps "compare boundary VALUES, not bin
 * indices" exact even with duplicate boundaries.
 */
typedef struct sb_node {
    int32_t id;              /* heap id, root = 1 */
    int32_t left;            /* index of left child in the array, -1 if none */
    int32_t right;           /* index of right child, -1 if none */
    int32_t feature;         /* feature slot of the split, -1 if the node has no split */
    int32_t split_type;      /* SB_SPLIT_* (informational on the device side) */
    int32_t left_mask_offset;/* byte offset of this node's left_mask in sb_tree.mask_pool */
    int32_t left_mask_bytes; /* n_bins/8 + 1 */
    int32_t has_predict;     /* Node.predict != null */
    double predict;          /* Node.predict.predict */
    /* The reference's raw Split fields (dt/Split.java:39-58).  The device routes by left_mask only; these are
       carried so the tree can be handed back unchanged (and so the CPU oracle can route the reference's way). */
    double threshold;        /* Split.threshold (numeric) */
    int32_t is_left;         /* Split.isLeft (categorical) */
    int32_t cat_offset;      /* byte offset of Split.leftOrRightCategories (SimpleBitSet bytes) in mask_pool */
    int32_t cat_bytes;
    int32_t reserved;
} sb_node;

typedef struct sb_tree {
    int32_t tree_id;         /* TreeNode.treeId */
    int32_t n_nodes;         /* 0 => placeholder tree with root id INVALID_INDEX (dt/DTMaster.java:500-504) */
    const sb_node* nodes;    /* nodes[0] is the root */
    const uint8_t* mask_pool;
    int32_t mask_pool_bytes;
} sb_tree;

/* One entry of DTMasterParams.todoNodes (dt/DTMasterParams.java:64, dt/TreeNode.java:41-104). */
typedef struct sb_todo_node {
    int32_t key;             /* map key (nodeIndexInGroup, dt/DTMaster.java:467-468) */
    int32_t tree_id;
    int32_t node_id;         /* heap id */
    int32_t n_features;      /* TreeNode.features.size(); 0 => all valid features (dt/DTWorker.java:841-843) */
    const int32_t* features; /* feature SLOTS, in list order */
} sb_todo_node;

/* DTMasterParams as the worker consumes it (dt/DTWorker.java:563-571, dt/DTMasterParams.java:44-203). */
typedef struct sb_iter_in {
    int32_t n_trees;
    const sb_tree* trees;    /* GBT: the last one or two trees (dt/DTMaster.java:479-487); RF: all trees */
    int32_t n_todo;
    const sb_todo_node* todo;
    int32_t switch_to_next_tree;
    int32_t is_first_tree;
    int32_t continuous_running_start; /* must be 0: continuous training stays on the Java side */
    /* Injected randomness (the reference draws it from unseeded java.util.Random, dt/DTWorker.java:643-668,
       634-638).  NULL => keep the current weights / drop nothing. */
    const float* new_weights;   /* [n_train_rows] replacement for subsampleWeights[currTreeIndex % len] on tree switch */
    const uint8_t* dropout;     /* [n_train_rows] 1 = drop this tree's contribution for the row */
} sb_iter_in;

/* DTWorkerParams as the master consumes it (dt/DTWorkerParams.java:44-117). */
typedef struct sb_iter_out {
    double train_count;       /* weightedTrainCount */
    double validation_count;  /* weightedValidationCount */
    double train_error;
    double validation_error;
    double* stats;            /* caller-allocated, sb_stats_len(...) doubles, or NULL to leave stats on the device.
                                 Layout: todo order; within a node its feature list order; within a feature
                                 bin-major, stat-minor -- i.e. each feature block is the reference's
                                 double[bins*statsSize] (dt/DTWorkerParams.java:160-245) */
    int64_t stats_len;
} sb_iter_out;

/* GainInfo + Split of the winning candidate of one todo node (dt/GainInfo.java:27-43, dt/Split.java). */
typedef struct sb_gain {
    int32_t found;            /* 0 => maxGainInfo == null: node becomes a leaf (dt/DTMaster.java:335-339) */
    int32_t feature;          /* feature slot */
    int32_t split_type;       /* SB_SPLIT_* */
    int32_t split_bin;        /* numeric: i of the winning prefix; threshold = binBoundary[i+1] (dt/Impurity.java:200).
                                 categorical: number of sorted bins scanned (i+1) */
    int32_t is_left;          /* categorical: bitset holds the left (1) or right (0) categories (dt/Impurity.java:184-193) */
    int32_t cat_bytes;        /* bytes used in cat_bits */
    int32_t left_class;       /* Predict.classValue of the left child (Entropy/Gini), else 0 */
    int32_t right_class;
    int32_t node_class;
    int32_t reserved;
    double gain, impurity, predict;
    double left_impurity, right_impurity, left_predict, right_predict;
    double wgt_cnt;
} sb_gain;

#define SB_MAX_CAT_BYTES 4096 /* Short.MAX_VALUE categories / 8 + 1 (dt/DTWorker.java:100-101) */

/* ------------------------------------------------------------------ lifecycle */

/* Replaces DTWorker.init (dt/DTWorker.java:349-550): fixes the column layout and train params. */
SB_API int sb_create(const sb_config* cfg, sb_handle** out);
SB_API void sb_destroy(sb_handle* h);
/* Thread-local message of the last failing call on this thread. */
SB_API const char* sb_last_error(void);
SB_API const char* sb_version(void);

/* ------------------------------------------------------------------ data load */

/*
 * Replaces the per-line DTWorker.load (dt/DTWorker.java:1099-1297) for already-binned rows: appends
 * n_rows records to the training (is_validation = 0) or validation list.  bins[f] points at n_rows
 * bin indices of feature slot f, uint8 when n_bins[f] <= 256 else uint16 (the reference stores
 * short, dt/DTWorker.java:1593).  weights is tree-major [weights_per_row][n_rows]
 * (Data.subsampleWeights transposed); NULL => all 1f.  significance NULL => all 1f.
 * output = predict = label initially (dt/DTWorker.java:1264-1273).
 * Host pointers; the call copies to the device and returns when the copy is complete.
 * Limits: ONE load call per set and handle (a second sb_load_rows* / sb_load_synthetic of the same set returns
 * SB_ERR_UNSUPPORTED -- the Java side batches DTWorker.load lines into direct buffers and loads in postLoad);
 * fewer than 2^32 - 1 rows per handle (row indices are 32-bit on the device: SB_ERR_UNSUPPORTED beyond; shard
 * larger sets over several GPUs); at most 16 classes for Entropy / Gini (sb_create rejects more).
 */
SB_API int sb_load_rows(sb_handle* h, int32_t is_validation, int64_t n_rows, const void* const* bins,
                        const float* label, const float* significance, const float* weights);
/* Same, but every pointer is a device pointer on the handle's GPU and is ADOPTED, not copied (the
   caller keeps it alive until sb_destroy).  bins is the column-major uint8 bin matrix: feature slot f
   starts at bins + f*pitch; bins 16-byte aligned, pitch a multiple of 16 and >= n_rows. */
SB_API int sb_load_rows_device(sb_handle* h, int32_t is_validation, int64_t n_rows, const void* bins, int64_t pitch,
                               float* label, float* significance, float* weights);
/* Replaces Data.subsampleWeights[w_index] of every training row (host array of n floats): the bag counts / bagging
   mask DTWorker.sampleWeights draws at load time (dt/DTWorker.java:1515-1575), injected by the caller. */
SB_API int sb_set_weights(sb_handle* h, int32_t w_index, const float* weights);
/* The GBT bagging mask of DTWorker.sampleWeights / the per-tree redraw (dt/DTWorker.java:1521-1544, 643-668), drawn
   on the device: subsampleWeights[w_index] of training row j becomes 1f iff the (first_draw + j)-th nextDouble() of
   `new java.util.Random(seed)` is <= rate, else 0f -- the same numbers a seeded (baggingSampleSeed) reference worker
   draws for its rows in load order.  Not stratified (the reference keeps one Random per class for that). */
SB_API int sb_sample_weights(sb_handle* h, int32_t w_index, int64_t seed, double rate, int64_t first_draw);
/* Stratified variant (isStratifiedSampling, dt/DTWorker.java:1526-1532 and :646-653): the reference keeps one Random
   per class value (int)(label + 0.01f), every one created from the same seed, so training row j gets the k-th
   nextDouble() of `new java.util.Random(seed)`, k = number of earlier rows of its class in load order. */
SB_API int sb_sample_weights_stratified(sb_handle* h, int32_t w_index, int64_t seed, double rate, int32_t n_classes);
/* RF bag counts (dt/DTWorker.java:1547-1571): subsampleWeights[w_index] of training row j becomes a Poisson(rate) count.
   The reference samples commons-math3's PoissonDistribution from an unseeded generator, so this matches it in
   distribution only: the count is the inverse Poisson CDF of a counter-based uniform of (seed, row0 + j, w_index) --
   row0 = global index of the handle's first row, so any sharding of the rows draws the same bags. */
SB_API int sb_sample_poisson(sb_handle* h, int32_t w_index, uint64_t seed, double rate, int64_t row0);
/* DTWorker.getBinIndex (dt/DTWorker.java:1001-1034) for a whole numeric column: values (device, n floats; NaN counts as
   0f like DTWorker.getFloatValue :1299-1305) -> bin indices (device, uint8 if out_width == 1 else uint16).
   bin_boundary is a HOST array of n_bins (<= 4096) doubles; it is copied into a per-call device table allocated and
   freed in stream order (cudaMallocAsync), so calls on different streams or host threads are independent; pinned host
   memory must stay valid until the stream reaches the copy.  A value below the first boundary gets the all-ones index
   (the reference returns -1), which sb_load_rows rejects -- for that reason 256 bins need out_width == 2.
   Handle-free; runs on cuda_stream (NULL = default). */
SB_API int sb_bin_values(const float* values, int64_t n, const double* bin_boundary, int32_t n_bins, void* bins_out,
                         int32_t out_width, void* cuda_stream);
/* Reads back per-row state (any pointer may be NULL): Data.predict, Data.output, current node id of
   tree `tree_id`, subsampleWeights[w_index]. */
SB_API int sb_read_rows(sb_handle* h, int32_t is_validation, int32_t tree_id, int32_t w_index, float* predict,
                        float* output, int32_t* node_id, float* weight);

/* ------------------------------------------------------------------ iteration */

/* Number of doubles sb_compute writes for this todo list (dt/DTWorker.java:903-929 initTodoNodeStats). */
SB_API int64_t sb_stats_len(sb_handle* h, const sb_iter_in* in);

/*
 * Replaces DTWorker.doCompute (dt/DTWorker.java:558-891): phase A / A' (GBT predict + residual update,
 * train / validation error, :582-751), phase B (node x feature x bin statistics via
 * Impurity.featureUpdate, :762-863) and the thread merge (:865-883).  Synchronous: on return `out`
 * holds the four scalars and (if out->stats != NULL) the statistics on the host.
 */
SB_API int sb_compute(sb_handle* h, const sb_iter_in* in, sb_iter_out* out);

/*
 * The same iteration in pieces, so that the DTMaster merge of worker results (dt/DTMaster.java:298-315) can run
 * as collectives between them when rows are sharded over several GPUs (one handle / process per GPU):
 *   sb_compute_begin   phase A / A' (routing, GBT update, error sums) + per-statistic maxima of the pass
 *   sb_scale_buffer    device pointer to 3 x int32 maxima (fp32 bit patterns)      -> all-reduce MAX
 *   sb_compute_stats   phase B: fixed-point scales from the maxima, then the histogram kernel, which leaves
 *                      order-independent integer accumulators on the device
 *   sb_raw_buffer      device pointer + byte size of those accumulators (int64)    -> all-reduce SUM
 *   sb_acc_buffer      device pointer to the 4 fp64 scalars (counts, errors)       -> all-reduce SUM
 *   sb_compute_finish  converts accumulators to the reference's double[] statistics (device-resident),
 *                      fills the scalars of `out`, copies stats to out->stats when non-NULL.
 * With one GPU the collectives are simply skipped (that is what sb_compute does).
 * All work is enqueued on the stream given to sb_set_stream (default: the legacy default stream);
 * sb_compute_begin returns without synchronising.
 */
SB_API int sb_compute_begin(sb_handle* h, const sb_iter_in* in);
/* After sb_compute_begin: were the scale maxima / the 4 scalars recomputed by this pass?  Inside a GBT tree neither
   changes (output, weights and predict move only at a tree switch), so a multi-GPU driver all-reduces them only when
   the flag is set; a buffer that was not recomputed still holds the already-reduced values. */
SB_API int sb_pass_info(sb_handle* h, int32_t* scales_fresh, int32_t* acc_fresh);
SB_API int sb_scale_buffer(sb_handle* h, void** dev_ptr, int64_t* bytes);
SB_API int sb_compute_stats(sb_handle* h);
SB_API int sb_raw_buffer(sb_handle* h, void** dev_ptr, int64_t* bytes);
SB_API int sb_acc_buffer(sb_handle* h, void** dev_ptr, int64_t* bytes);
SB_API int sb_compute_finish(sb_handle* h, sb_iter_out* out);
SB_API int sb_set_stream(sb_handle* h, void* cuda_stream);

/*
 * Batch scoring: IndependentTreeModel.compute(double[] data) (dt/IndependentTreeModel.java:352-446) over a block of
 * rows.  The model is the in-memory form of a BinaryDTSerializer file: bags of trees with a weight each (the tree's
 * learning rate for GBT, :397-403; bag weights for RF, :426-435).  predictNode (:516-560) is followed literally: a
 * node is walked through while it has a split and at least one child; numeric: left iff value < threshold;
 * categorical: index = categoricalSize when Double.compare(value, 0) < 0 or Double.compare(value, size) >= 0, else
 * (short)(value + 0.1); left iff SimpleBitSet.contains(index) == isLeft.
 * Output: regression -> one double per row (GBT: sum of score*weight per bag, converted, averaged over the bags;
 * RF: weighted mean per bag, averaged); classification -> n_trees doubles per row (class value of every tree).
 */
#define SB_CONVERT_RAW 0
#define SB_CONVERT_OLD_SIGMOID 1 /* 1 / (1 + min(1e19, exp(-score)))       :505-507 */
#define SB_CONVERT_SIGMOID 2     /* 1 / (1 + min(1e19, exp(-20 * score)))  :491-493 */
#define SB_CONVERT_CUTOFF 3      /* clamp to [0, 1]                        :471-479 */

typedef struct {
    int32_t left, right;      /* index in the model's node array, -1 = null (Node.left / Node.right) */
    int32_t column;           /* getColumnIndex(split.columnNum): position in a data row; -1 = Node.split == null */
    int32_t split_type;       /* SB_SPLIT_CONTINUOUS / SB_SPLIT_CATEGORICAL */
    double threshold;         /* Split.threshold */
    int32_t is_left;          /* Split.isLeft */
    int32_t cat_offset;       /* Split.leftOrRightCategories as SimpleBitSet bytes inside cat_pool */
    int32_t cat_bytes;
    int32_t class_value;      /* Predict.classValue */
    double predict;           /* Predict.predict */
} sb_score_node;

typedef struct {
    int32_t n_bags;
    const int32_t* bag_begin;         /* [n_bags + 1] range of trees of each bag */
    int32_t n_trees;
    const int32_t* tree_root;         /* [n_trees] node index of each tree's root */
    const double* weights;            /* [n_trees] */
    int32_t n_nodes;
    const sb_score_node* nodes;
    const uint8_t* cat_pool;
    int32_t cat_pool_bytes;
    int32_t is_gbdt, is_classification, convert;
    int32_t n_columns;                /* doubles per data row */
    const int32_t* categorical_size;  /* [n_columns] getCategoricalSize of the column at that position, 0 if numeric */
} sb_score_model;

typedef struct sb_scorer sb_scorer;
/* copies the model to `device`; the sb_score_model and what it points at may be released afterwards */
SB_API int sb_scorer_create(const sb_score_model* model, int32_t device, sb_scorer** out);
SB_API void sb_scorer_destroy(sb_scorer* s);
/* data: row-major [n_rows][n_columns] doubles, host memory (or device memory when data_on_device != 0);
   scores: host buffer of n_rows (regression) or n_rows * n_trees (classification) doubles */
SB_API int sb_scorer_score(sb_scorer* s, int64_t n_rows, const double* data, int32_t data_on_device, double* scores);
/* kernel time of the last sb_scorer_score call in milliseconds (CUDA events around the kernel) */
SB_API int sb_scorer_last_kernel_ms(sb_scorer* s, float* ms);

/*
 * One process per GPU, rows sharded over the processes: the merge of the workers' statistics that Guagua hands to
 * DTMaster.doCompute (dt/DTMaster.java:316-331 sums the NodeStats of every worker; :342-352 the error sums)
 * becomes an NCCL all-reduce inside sb_compute / sb_compute_stats (integer sum of the accumulators, MAX of the
 * fixed-point scales at a tree switch, sum of the four error scalars).  Rank 0 calls sb_dist_unique_id and hands
 * the 128 bytes to every process (any side channel); each process then calls sb_dist_init once on its handle.
 * Afterwards every rank's sb_compute returns the merged statistics and sb_best_splits the same splits.
 * NCCL is opened with dlopen("libnccl.so.2") at the first call; single-GPU use needs no NCCL.
 * Every rank must hold at least one training row (a rank without rows is refused: it would not take part in the
 * sibling subtraction the other ranks plan).
 * No counterpart call in the reference: there the merge is the Guagua master's message loop.
 */
SB_API int sb_dist_unique_id(uint8_t* id128);
SB_API int sb_dist_init(sb_handle* h, const uint8_t* id128, int32_t rank, int32_t world_size);

/*
 * Replaces the per-node body of DTMaster.doCompute (dt/DTMaster.java:316-340): for every todo node of the
 * last sb_compute*, Impurity.computeImpurity per feature (dt/Impurity.java:120-211,269-360,378-477,563-661)
 * and GainInfo.getGainInfoByMaxGain across features (dt/GainInfo.java:200-214), on the device, from the
 * (all-reduced) statistics.  visit_order: feature slots in the order the reference's HashMap iterates
 * them (dt/DTMaster.java:324), or NULL for each node's own feature list order.
 * gains: [n_todo]; cat_bits: [n_todo * SB_MAX_CAT_BYTES] or NULL when no categorical feature exists.
 */
SB_API int sb_best_splits(sb_handle* h, sb_gain* gains, uint8_t* cat_bits);
/* sb_compute_finish followed by sb_best_splits with one stream synchronisation between them (the per-pass
   latency matters once a level's statistics take under a millisecond). */
SB_API int sb_compute_finish_split(sb_handle* h, sb_iter_out* out, sb_gain* gains, uint8_t* cat_bits);

/* ------------------------------------------------------------------ host mirror of DTMaster (no CUDA) */

/* core/dtrain/FeatureSubsetStrategy.java:23-31; SBH_FSS_RATE = "featureSubsetStrategy is a number" (dt/DTMaster.java:823-825) */
enum { SBH_FSS_ALL = 0, SBH_FSS_HALF = 1, SBH_FSS_ONETHIRD = 2, SBH_FSS_TWOTHIRDS = 3, SBH_FSS_AUTO = 4, SBH_FSS_SQRT = 5,
       SBH_FSS_LOG2 = 6, SBH_FSS_RATE = 7 };

/* The train.params DTMaster.init reads (dt/DTMaster.java:932-1051). */
typedef struct sbh_params {
    int32_t tree_num;                /* "TreeNum" */
    int32_t max_depth;               /* "MaxDepth" */
    int32_t is_gbdt;
    int32_t feature_subset_strategy; /* SBH_FSS_* */
    double feature_subset_rate;      /* used with SBH_FSS_RATE */
    double learning_rate;            /* "LearningRate": TreeNode.learningRate of every tree but the first (:421, :1108) */
    int64_t feature_sampling_seed;   /* the reference's featureSamplingRandom is unseeded (:246); injected here */
    int32_t max_leaves;              /* "MaxLeaves": > 0 switches to leaf-wise growth (:980-982) */
    int32_t max_batch_split_size;    /* "MaxBatchSplitSize" (unused by the reference's loop, :359-363) */
    int64_t max_stats_memory;        /* "MaxStatsMemoryMB" in bytes (:1001-1013) */
    int32_t worker_number;           /* workers the statistics memory estimate is scaled by (:745) */
    int32_t reserved;
} sbh_params;

typedef struct sbh_master sbh_master;

/* DTMaster.init (dt/DTMaster.java:896-1116): cfg needs bin_boundary (split thresholds are boundary values). */
SB_API int sbh_master_create(const sb_config* cfg, const sbh_params* params, sbh_master** out);
SB_API void sbh_master_destroy(sbh_master* m);
/* buildInitialMasterParams (dt/DTMaster.java:759-820): the first DTMasterParams, flattened.  The returned message and
   everything it points at stay valid until the next sbh_master_* call on this object. */
SB_API const sb_iter_in* sbh_master_initial(sbh_master* m);
/* DTMaster.doCompute (dt/DTMaster.java:316-537) fed with the sb_gain records of sb_best_splits for the todo nodes of
   the previous message (same order).  Returns the next message, or NULL when training halts (:396-407, :431-435). */
SB_API const sb_iter_in* sbh_master_do_compute(sbh_master* m, const sb_gain* gains, const uint8_t* cat_bits);
/* The iteration loop of one tree in native code: starting from message `first`, run sb_compute_begin /
   sb_compute_stats / sb_compute_finish_split on `worker` and feed sbh_master_do_compute until the master switches to
   the next tree (or halts).  *next: the first message of the next tree (NULL when training halted); *n_passes: worker
   iterations run; *last: the scalars of the last iteration.  What the Guagua iteration loop does between
   DTWorker.doCompute and DTMaster.doCompute (guagua MasterComputable / WorkerComputable contract). */
SB_API int sbh_grow_tree(sbh_master* m, sb_handle* worker, const sb_iter_in* first, const sb_iter_in** next,
                         int32_t* n_passes, sb_iter_out* last);
SB_API int sbh_master_tree_count(sbh_master* m);
SB_API int sbh_master_halted(sbh_master* m);
/* Copies tree `index` out: sb_node records in pre-order + mask pool, and per node 7 doubles
   (gain, impurity, leftImpurity, rightImpurity, leftPredict, rightPredict, wgtCnt).  Call with nodes == NULL to size. */
SB_API int sbh_master_export_tree(sbh_master* m, int index, sb_node* nodes, int node_cap, uint8_t* pool, int pool_cap,
                                  double* extras, int* n_nodes, int* pool_bytes);

/* ------------------------------------------------------------------ categorical binning at load (host side)
 *
 * The value -> bin-index map DTWorker.init builds from ColumnConfig.binCategory (dt/DTWorker.java:362-376): entry i
 * of bin_category may be a group "a@^b@^c" (CommonUtils.flattenCatValGrp, util/CommonUtils.java:1767; delimiter
 * Constants.CATEGORICAL_GROUP_VAL_DELIMITER) whose members all map to i; a value listed twice keeps the LAST index
 * (HashMap.put).  sb_category_map_lookup is the per-value rule of DTWorker.load (dt/DTWorker.java:1152-1169): the
 * empty string (or NULL) and any value that is in no group get bin n_categories, the missing-value bin.  Host strings
 * in, host bin indices out (uint8 if out_width == 1 else uint16) -- ready for sb_load_rows.
 */
typedef struct sb_category_map sb_category_map;
SB_API int sb_category_map_create(const char* const* bin_category, int32_t n_categories, sb_category_map** out);
SB_API int sb_category_map_lookup(const sb_category_map* m, const char* const* values, int64_t n, void* bins_out,
                                  int32_t out_width);
SB_API void sb_category_map_destroy(sb_category_map* m);

/* ------------------------------------------------------------------ wire format of sb_iter_in
 *
 * DTMasterParams (dt/DTMasterParams.java:148-173 write / readFields) as ONE position-independent buffer: what a JVM
 * hands over in a direct ByteBuffer (jni/DTMessageFlattener.java writes it, jni/shifu_b200_jni.c unpacks it).
 * Little endian, every section starts on a multiple of 8 bytes:
 *
 *   offset  0  u32 magic 0x314D4253 ("SBM1")      4  u32 total bytes of the message
 *           8  i32 n_trees                        12  i32 n_todo
 *          16  i32 switch_to_next_tree            20  i32 is_first_tree
 *          24  i32 continuous_running_start       28  u32 bit 0: new_weights present, bit 1: dropout present
 *          32  i64 n_rows (length of new_weights / dropout)
 *          40  per tree:  i32 tree_id, i32 n_nodes, i32 mask_pool_bytes, i32 0,
 *                         n_nodes x sb_node (64 bytes each, fields in declaration order),
 *                         mask_pool_bytes bytes, zero-padded to a multiple of 8
 *              per todo:  i32 key, i32 tree_id, i32 node_id, i32 n_features,
 *                         n_features x i32 feature slots, zero-padded to a multiple of 8
 *              new_weights: n_rows x f32, padded to 8 (if present);  dropout: n_rows x u8, padded to 8 (if present)
 *
 * sb_msg_unpack validates sizes, child indices and mask offsets and returns a view whose sb_iter_in points INTO the
 * buffer (nodes, pools, feature lists and per-row arrays are not copied): the buffer must be 8-byte aligned and
 * outlive the view.  sb_msg_pack writes the same layout from an sb_iter_in (buf == NULL: returns the size needed).
 */
typedef struct sb_msg sb_msg;
SB_API int64_t sb_msg_pack(const sb_iter_in* in, int64_t n_rows, void* buf, int64_t cap);
SB_API int sb_msg_unpack(const void* buf, int64_t len, sb_msg** out);
SB_API const sb_iter_in* sb_msg_iter_in(const sb_msg* m);
SB_API void sb_msg_free(sb_msg* m);

/*
 * No counterpart in the reference.  Weighted error sums of a row set -- out[0] = sum(significance),
 * out[1] = sum(significance * Loss.computeError(predict, label)), the two numbers a GBT iteration reports as
 * trainCount / trainError -- for callers that must not allocate device memory: a monitoring thread next to a
 * running trainer (cudaMalloc / cudaFree would stall the training stream), or a process whose HBM is already
 * fully committed.  The reduction scratch is a static 64 KB device array inside the library, so no handle is
 * needed and nothing is allocated or freed, ever.
 *   - at most SB_STATIC_MAX_ROWS rows per call (more => SB_ERR_RANGE);
 *   - label / predict / significance (NULL = all 1f) / out are device pointers on the CURRENT device; out holds
 *     2 doubles and is valid once `cuda_stream` (NULL = default stream) has reached this call;
 *   - the first call on a device looks up the scratch address; make it before any stream capture on that device;
 *   - all calls on a device share the scratch and must not overlap on the GPU (they would race).
 */
#define SB_STATIC_MAX_ROWS 67108864LL
SB_API int sb_error_sums_static(const float* label, const float* predict, const float* significance, int64_t n_rows,
                                int32_t loss, double* out, void* cuda_stream);

/* ------------------------------------------------------------------ pieces of the path (bench / tests) */

/* Re-runs only the statistics pass (phase B) of the last sb_compute_begin, n_repeat times, and reports the
   mean device time of the histogram kernel in milliseconds (CUDA events on the handle's stream). */
SB_API int sb_time_stats_pass(sb_handle* h, int32_t n_repeat, float* ms_hist_kernel, float* ms_total);
/* Per-pass timing of the histogram kernel: with profiling on, sb_compute_stats brackets the kernel with CUDA
   events on the handle's stream; sb_hist_time returns (and clears) the accumulated device time, the number of
   passes and of histogram launches since the last call. */
SB_API int sb_set_profiling(sb_handle* h, int32_t on);
SB_API int sb_hist_time(sb_handle* h, float* ms_total, int32_t* n_passes, int32_t* n_launches);
/* Counters since creation: kernels launched by this library, bytes copied H2D / D2H. */
SB_API int sb_counters(sb_handle* h, int64_t* launches, int64_t* h2d_bytes, int64_t* d2h_bytes);
/* Rows the statistics kernels actually read since creation (or since the last call with reset != 0): with histogram
   subtraction a pass below the root reads only the rows of the smaller child of every sibling pair.  For reporting a
   bandwidth that counts streamed bytes, not the bytes a pass over every row would move.  Synchronises the stream. */
SB_API int sb_rows_streamed(sb_handle* h, int64_t* rows, int32_t reset);
/* Fills device memory with the synthetic data set of SURVEY.md 8(d) (counter-based splitmix64) and loads it:
   n_rows rows starting at global row `row0`, all features uint8 with n_bins[f] uniform bins.
   label_kind 0: binary label from a 5-feature logistic with 10% flips (C2); 1: threshold rule (C1). */
SB_API int sb_load_synthetic(sb_handle* h, int32_t is_validation, int64_t row0, int64_t n_rows, uint64_t seed,
                             int32_t label_kind, int32_t weighted);
/* Copies one synthetic column / label range to the host so the oracle can be fed identical rows. */
SB_API int sb_read_bins(sb_handle* h, int32_t is_validation, int32_t feature, int64_t row0, int64_t n_rows, void* out);
SB_API int sb_read_labels(sb_handle* h, int32_t is_validation, int64_t row0, int64_t n_rows, float* label,
                          float* significance);

#ifdef __cplusplus
}
#endif
#endif /* SHIFU_B200_H */
