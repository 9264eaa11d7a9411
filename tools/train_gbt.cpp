// train_gbt.cpp -- a host program that uses nothing but the C ABI of include/shifu_b200.h (no Python, no PyTorch):
// synthetic rows generated on the device, a GBT grown tree by tree with the native iteration loop
// (sbh_grow_tree = DTWorker.doCompute + DTMaster.doCompute per level), the training error printed per tree.
// It is what a C++ (or, through JNI, Java) caller of the library looks like; build.sh compiles it to tools/train_gbt.
//
//   tools/train_gbt [rows=2000000] [features=50] [bins=32] [trees=5] [depth=6]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../include/shifu_b200.h"

#define CHECK(call)                                                          \
    do {                                                                     \
        const int rc_ = (call);                                              \
        if (rc_ != SB_OK) {                                                  \
            std::fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, sb_last_error()); \
            return 1;                                                        \
        }                                                                    \
    } while (0)

int main(int argc, char** argv) {
    const long long rows = argc > 1 ? std::atoll(argv[1]) : 2000000;
    const int features = argc > 2 ? std::atoi(argv[2]) : 50;
    const int bins = argc > 3 ? std::atoi(argv[3]) : 32;
    const int trees = argc > 4 ? std::atoi(argv[4]) : 5;
    const int depth = argc > 5 ? std::atoi(argv[5]) : 6;

    std::vector<int32_t> n_bins((size_t)features, bins), column_num((size_t)features);
    std::vector<double> boundary;  // ColumnConfig.binBoundary of every feature: the split thresholds come from here
    for (int f = 0; f < features; ++f) {
        column_num[f] = f + 1;
        for (int b = 0; b < bins; ++b) boundary.push_back((double)b);
    }
    sb_config cfg{};
    cfg.n_features = features;
    cfg.n_bins = n_bins.data();
    cfg.column_num = column_num.data();
    cfg.bin_boundary = boundary.data();
    cfg.impurity = SB_IMPURITY_VARIANCE;
    cfg.loss = SB_LOSS_SQUARED;
    cfg.is_gbdt = 1;
    cfg.tree_num = trees;
    cfg.weights_per_row = 1;
    cfg.max_depth = depth;
    cfg.min_instances_per_node = 5;
    cfg.learning_rate = 0.05;
    cfg.device = 0;

    sb_handle* worker = nullptr;
    CHECK(sb_create(&cfg, &worker));
    CHECK(sb_load_synthetic(worker, /*is_validation=*/0, /*row0=*/0, rows, /*seed=*/1002, /*label_kind=*/0, /*weighted=*/0));

    sbh_params mp{};
    mp.tree_num = trees;
    mp.max_depth = depth;
    mp.is_gbdt = 1;
    mp.feature_subset_strategy = SBH_FSS_ALL;
    mp.learning_rate = cfg.learning_rate;
    mp.max_leaves = -1;
    mp.max_batch_split_size = 32;
    mp.max_stats_memory = 1LL << 30;
    mp.worker_number = 1;
    sbh_master* master = nullptr;
    CHECK(sbh_master_create(&cfg, &mp, &master));

    const sb_iter_in* msg = sbh_master_initial(master);
    int tree = 0;
    while (msg) {
        const auto t0 = std::chrono::steady_clock::now();
        const sb_iter_in* next = nullptr;
        int32_t passes = 0;
        sb_iter_out last{};
        CHECK(sbh_grow_tree(master, worker, msg, &next, &passes, &last));
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        std::printf("tree %d: %d passes, %.2f ms, train error %.6f over %.0f rows (error of the model before this tree)\n", tree,
                    passes, ms, last.train_count > 0 ? last.train_error / last.train_count : 0.0, last.train_count);
        msg = next;
        ++tree;
    }
    std::printf("%d trees grown, %s\n", sbh_master_tree_count(master), sb_version());
    sbh_master_destroy(master);
    sb_destroy(worker);
    return 0;
}
