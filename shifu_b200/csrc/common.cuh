// common.cuh -- shared declarations of the shifu_b200 device library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/shifu_b200.h"

namespace sb {

void set_error(const std::string& msg);
int fail(int code, const std::string& msg);

#define SB_CUDA(expr)                                                                                   \
    do {                                                                                                \
        cudaError_t _e = (expr);                                                                        \
        if (_e != cudaSuccess)                                                                          \
            return ::sb::fail(SB_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" +  \
                                               __FILE__ + ":" + std::to_string(__LINE__) + ")");        \
    } while (0)

// ------------------------------------------------------------------------------------------------
// Fixed-point accumulation (DESIGN.md "Why integer accumulators").
// A per-record fp32 contribution v is turned into q = rint(v * 2^e) with |q| < 2^QBITS and summed in
// 64-bit two's complement (shared memory, per block), then in a (lo, hi) pair of int64 words (global,
// across blocks and GPUs).  e is chosen per statistic from max|v| over the rows of the pass.
constexpr int QBITS = 44;                   // |q| < 2^44: fp32 values down to max * 2^-20 are represented exactly
constexpr int64_t MAX_ROWS_PER_FLUSH = 1 << 18;  // rows per block between merges: 2^18 * 2^44 < 2^63

enum StatKind { STAT_VARIANCE = 0, STAT_CLASS = 1 };

// One set of rows (training or validation) resident in HBM, column-major.
struct RowSet {
    int64_t n = 0;
    int n_features = 0;
    std::vector<const void*> h_col;  // per feature slot device pointer to n bin indices (uint8 or uint16)
    const void** d_col = nullptr;    // the same table on the device
    float* label = nullptr;
    float* significance = nullptr;
    float* output = nullptr;   // Data.output
    float* predict = nullptr;  // Data.predict
    float* weights = nullptr;  // [weights_per_row][n]
    int32_t* node_id = nullptr;  // [n_node_arrays][n] current heap id per tracked tree
    int n_node_arrays = 0;
    bool owns_inputs = false;
    void* owned_bins = nullptr;
    uint8_t* bins_rm = nullptr;       // row-major twin [n][fpad] for the row-list path (built on first use)
    int fpad = 0;
    int rm_fpl = 0;                   // features per lane the twin's byte layout was built for
    // features with more than 256 bins: uint16 columns [n_wide][pitch16] (+ row-major twin [n][wpad])
    uint16_t* bins16 = nullptr;
    int64_t pitch16 = 0;  // elements
    int n_wide = 0;
    uint16_t* bins16_rm = nullptr;
    int wpad = 0;
    const void* bins_base = nullptr;  // column-major uint8 bin matrix: feature f starts at bins_base + f * pitch
    int64_t pitch = 0;
    alignas(64) unsigned char tensor_map[128];  // CUtensorMap over the bin matrix (see hist.cu)
    bool loaded = false;
};

// A flattened tree on the device, indexed by heap id.
struct DeviceTree {
    int capacity = 0;        // ids [0, capacity)
    int32_t* feat = nullptr; // split feature slot if the node routes further (has split AND children), else -1
    int32_t* mask_off = nullptr;
    double* predict = nullptr;
    uint8_t* has_predict = nullptr;
    uint8_t* masks = nullptr;
    int mask_capacity = 0;
    void* walk = nullptr;    // WalkNode[capacity] (RF passes), or nullptr
};

struct TodoLayout {  // host description of the statistics buffer of one iteration
    struct Node {
        int key, tree_id, node_id;
        std::vector<int> features;
        std::vector<int64_t> feat_off;  // offset (doubles) of each feature block inside the node block
        int64_t offset;                 // offset (doubles) of the node block
    };
    std::vector<Node> nodes;
    int64_t total = 0;  // doubles
    // flat: the CALLER's feature lists are kept as one flat array (forest passes of an RF with thousands of todo nodes,
    // where nodes[i].features / feat_off stay empty; and full passes).
    bool flat = false;
    std::vector<int32_t> flat_begin;  // [n + 1]
    std::vector<int32_t> flat_feats;  // lists in todo order; a node that takes every valid feature has them spelled out
    // full: todo nodes have feature lists of their own, but the accumulators are laid out (and filled) for EVERY valid
    // feature of every node -- nodes[i] describes that internal layout, `total` is its size -- so that a child can be
    // derived from its parent and sibling whatever the lists are.  The caller still gets its own layout: caller_off[i] =
    // offset of node i's block there, caller_total its size (== total and nodes[i].offset when not full).
    bool full = false;
    std::vector<int64_t> caller_off;
    int64_t caller_total = 0;
};

}  // namespace sb
