// host_master.cpp -- host-side mirror of dt/DTMaster.java: the tree-growth state machine around the device path.
//
// What stays on the host is what the reference keeps in the master's scalar logic: the todo queue, level-wise /
// leaf-wise growth, per-node feature sub-sampling, the GBT tree switch and the halt rules (dt/DTMaster.java:274-537,
// 543-605, 710-727, 759-893).  What moved to the GPU is the merge of worker statistics and Impurity.computeImpurity
// + max-gain per node: do_compute consumes sb_gain records (sb_best_splits) instead of NodeStats maps, and hands back
// the next DTMasterParams already flattened as an sb_iter_in (the message sb_compute takes).
//
// Method names follow the Java class.  No CUDA in this file.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <deque>
#include <memory>
#include <string>
#include <vector>

#include "../../include/shifu_b200.h"

namespace sb { int fail(int code, const std::string& msg); }

namespace {

// java.util.Random (48-bit LCG): seeded runs of the reference draw exactly these numbers.
struct JavaRandom {
    uint64_t seed;
    explicit JavaRandom(int64_t s) : seed(((uint64_t)s ^ 0x5DEECE66DULL) & ((1ULL << 48) - 1)) {}
    int32_t next(int bits) {
        seed = (seed * 0x5DEECE66DULL + 0xBULL) & ((1ULL << 48) - 1);
        return (int32_t)((int64_t)seed >> (48 - bits));
    }
    double nextDouble() { return (double)(((int64_t)next(26) << 27) + next(27)) * (1.0 / (double)(1LL << 53)); }
};

struct Node {  // dt/Node.java
    int id = 1;
    int feature = -1;
    int split_type = SB_SPLIT_CONTINUOUS;
    double threshold = 0;
    bool is_left = false;
    std::vector<uint8_t> cat_bits;   // Split.leftOrRightCategories (SimpleBitSet bytes)
    std::vector<uint8_t> left_mask;  // derived once the split is final
    std::unique_ptr<Node> left, right;
    bool has_predict = false;
    double predict = 0;
    double gain = 0, impurity = 0, left_impurity = 0, right_impurity = 0, left_predict = 0, right_predict = 0;
    double wgt_cnt = 0, wgt_cnt_ratio = 0;
    int8_t predict_class = 0;
    bool is_leaf = true;
};

struct TreeNode {  // dt/TreeNode.java:41-104
    int tree_id;
    Node* node;               // the tree root (owned by `root`) or a todo node inside some tree
    std::unique_ptr<Node> root;
    double learning_rate = 1.0;
    std::vector<int> features;
    int node_num = 1;
    double root_wgt_cnt = 0;
};

inline int index_to_level(int id) {  // dt/Node.java:459-461
    int l = 0;
    while (id) { ++l; id >>= 1; }
    return l;
}
inline bool is_pos_zero(double x) { return x == 0.0 && !std::signbit(x); }  // Double.compare(x, 0d) == 0

// Double.compare as a total order: -0.0 below 0.0, NaN above everything
inline int double_compare(double a, double b) {
    if (a < b) return -1;
    if (a > b) return 1;
    const bool an = std::isnan(a), bn = std::isnan(b);
    if (an || bn) return an == bn ? 0 : (an ? 1 : -1);
    const bool as = std::signbit(a), bs = std::signbit(b);
    return as == bs ? 0 : (as ? -1 : 1);
}

// The leaf-wise split queue is a java.util.PriorityQueue (dt/DTMaster.java:1064-1070): a binary heap, so the order in
// which entries of EQUAL priority leave it is a property of the heap operations, not of the values -- and every tree
// root has priority 0 (Node.wgtCntRatio is never set for a root, :721-726).  Same sift-up / sift-down here.
struct SplitQueue {
    struct Item { double prio; int tree_id; Node* node; };
    std::vector<Item> heap;
    static bool before(const Item& a, const Item& b) { return double_compare(b.prio, a.prio) < 0; }  // larger first
    bool empty() const { return heap.empty(); }
    void offer(Item e) {
        size_t k = heap.size();
        heap.push_back(e);
        while (k > 0) {
            const size_t parent = (k - 1) / 2;
            if (!before(e, heap[parent])) break;
            heap[k] = heap[parent];
            k = parent;
        }
        heap[k] = e;
    }
    Item poll() {
        const Item top = heap.front();
        const Item last = heap.back();
        heap.pop_back();
        const size_t n = heap.size();
        size_t k = 0;
        while (2 * k + 1 < n) {
            size_t c = 2 * k + 1;
            if (c + 1 < n && before(heap[c + 1], heap[c])) c = c + 1;
            if (!before(heap[c], last)) break;
            heap[k] = heap[c];
            k = c;
        }
        if (n > 0) heap[k] = last;
        return top;
    }
};

}  // namespace

struct sbh_master {
    // config
    int F = 0;
    std::vector<int32_t> n_bins;
    std::vector<uint8_t> is_cat;
    std::vector<double> bb;            // concatenated boundaries of numeric slots
    std::vector<int64_t> bb_off;
    int stats_size = 3;
    sbh_params p{};
    bool is_gbdt = true, is_leaf_wise = false;
    std::vector<int> all_features;
    JavaRandom feature_sampling_random{0};
    // state
    std::vector<std::unique_ptr<TreeNode>> trees;
    struct QItem { int tree_id; Node* node; };
    std::deque<QItem> to_do_queue;
    struct Todo { int key; int tree_id; Node* node; std::vector<int> features; };
    std::vector<Todo> todo_nodes;
    bool halt = false, switch_to_next_tree = false, is_first_tree = false;
    // the flattened message handed out (valid until the next call)
    std::vector<std::vector<sb_node>> msg_nodes;
    std::vector<std::vector<uint8_t>> msg_pools;
    std::vector<sb_tree> msg_trees;
    std::vector<sb_todo_node> msg_todo;
    std::vector<int32_t> msg_feats;
    sb_iter_in msg{};

    const double* boundaries_of(int f) const { return bb.data() + bb_off[(size_t)f]; }

    // ---- feature sub-sampling, dt/DTMaster.java:822-850, 880-893
    std::vector<int> sample(int k) {
        std::vector<int> feats(all_features.begin(), all_features.begin() + std::min<size_t>((size_t)std::max(k, 0), all_features.size()));
        for (int i = k; i < (int)all_features.size(); ++i) {
            const int j = (int)(feature_sampling_random.nextDouble() * i);
            if (j >= 0 && j < k) feats[(size_t)j] = all_features[(size_t)i];
        }
        return feats;
    }
    std::vector<int> get_subsampling_features() {
        const int n = (int)all_features.size();
        switch (p.feature_subset_strategy) {
            case SBH_FSS_RATE: return sample((int)(n * p.feature_subset_rate));
            case SBH_FSS_HALF: return sample(n / 2);
            case SBH_FSS_ONETHIRD: return sample(n / 3);
            case SBH_FSS_TWOTHIRDS: return sample(n * 2 / 3);
            case SBH_FSS_SQRT: return sample((int)(n * std::sqrt((double)F) / F));
            case SBH_FSS_LOG2: return sample((int)(n * std::log((double)F) / std::log(2.0) / F));
            case SBH_FSS_AUTO: return p.tree_num > 1 ? sample(n / 2) : std::vector<int>();
            default: return {};  // ALL: empty list == every valid feature
        }
    }
    int64_t get_stats_mem(const std::vector<int>& subset) const {  // :729-747
        const std::vector<int>& feats = subset.empty() ? all_features : subset;
        int64_t mem = 0;
        for (int f : feats) mem += (int64_t)n_bins[(size_t)f] * stats_size * 8 * 2;
        return mem * p.worker_number / 2;
    }

    void build_left_mask(Node& n) const {  // dt/DTWorker.java:1052-1088
        const int f = n.feature;
        const int nb = n_bins[(size_t)f];
        n.left_mask.assign((size_t)(nb / 8 + 1), 0);
        if (!is_cat[(size_t)f]) {
            const double* b = boundaries_of(f);
            for (int i = 0; i < nb; ++i)
                if (b[i] < n.threshold) n.left_mask[(size_t)(i >> 3)] |= (uint8_t)(1 << (i % 8));
        } else {
            const int n_cats = nb - 1;
            for (int i = 0; i < nb; ++i) {
                const int idx = (i >= 0 && i < n_cats) ? i : n_cats;
                const size_t wi = (size_t)(idx >> 3);
                const bool has = wi < n.cat_bits.size() && ((n.cat_bits[wi] >> (idx % 8)) & 1);
                if (has == n.is_left) n.left_mask[(size_t)(i >> 3)] |= (uint8_t)(1 << (i % 8));
            }
        }
    }

    void flatten(const Node* n, std::vector<sb_node>& out, std::vector<uint8_t>& pool) const {
        const size_t me = out.size();
        out.emplace_back();
        {
            sb_node& s = out[me];
            std::memset(&s, 0, sizeof(s));
            s.id = n->id; s.left = -1; s.right = -1; s.feature = n->feature; s.split_type = n->split_type;
            s.has_predict = n->has_predict ? 1 : 0; s.predict = n->predict; s.threshold = n->threshold;
            s.is_left = n->is_left ? 1 : 0;
            if (n->feature >= 0) {
                s.left_mask_offset = (int32_t)pool.size();
                s.left_mask_bytes = (int32_t)n->left_mask.size();
                pool.insert(pool.end(), n->left_mask.begin(), n->left_mask.end());
                s.cat_offset = (int32_t)pool.size();
                s.cat_bytes = (int32_t)n->cat_bits.size();
                pool.insert(pool.end(), n->cat_bits.begin(), n->cat_bits.end());
            }
        }
        if (n->left) { const int32_t li = (int32_t)out.size(); flatten(n->left.get(), out, pool); out[me].left = li; }
        if (n->right) { const int32_t ri = (int32_t)out.size(); flatten(n->right.get(), out, pool); out[me].right = ri; }
    }

    const sb_iter_in* message() {
        std::vector<TreeNode*> sent;
        if (is_gbdt) {  // dt/DTMaster.java:479-487
            if (switch_to_next_tree && trees.size() >= 2) sent = {trees[trees.size() - 2].get(), trees.back().get()};
            else sent = {trees.back().get()};
        } else {
            for (auto& t : trees) sent.push_back(t.get());
        }
        msg_nodes.assign(sent.size(), {});
        msg_pools.assign(sent.size(), {});
        msg_trees.assign(sent.size(), sb_tree{});
        for (size_t i = 0; i < sent.size(); ++i) {
            flatten(sent[i]->root.get(), msg_nodes[i], msg_pools[i]);
            msg_pools[i].push_back(0);
            msg_trees[i].tree_id = sent[i]->tree_id;
            msg_trees[i].n_nodes = (int32_t)msg_nodes[i].size();
            msg_trees[i].nodes = msg_nodes[i].data();
            msg_trees[i].mask_pool = msg_pools[i].data();
            msg_trees[i].mask_pool_bytes = (int32_t)msg_pools[i].size() - 1;
        }
        size_t nf = 0;
        for (auto& t : todo_nodes) nf += t.features.size();
        msg_feats.assign(std::max<size_t>(nf, 1), 0);
        msg_todo.assign(todo_nodes.size(), sb_todo_node{});
        size_t off = 0;
        for (size_t i = 0; i < todo_nodes.size(); ++i) {
            msg_todo[i].key = todo_nodes[i].key;
            msg_todo[i].tree_id = todo_nodes[i].tree_id;
            msg_todo[i].node_id = todo_nodes[i].node->id;
            msg_todo[i].n_features = (int32_t)todo_nodes[i].features.size();
            msg_todo[i].features = msg_feats.data() + off;
            std::copy(todo_nodes[i].features.begin(), todo_nodes[i].features.end(), msg_feats.begin() + (long)off);
            off += todo_nodes[i].features.size();
        }
        std::memset(&msg, 0, sizeof(msg));
        msg.n_trees = (int32_t)msg_trees.size();
        msg.trees = msg_trees.data();
        msg.n_todo = (int32_t)msg_todo.size();
        msg.todo = msg_todo.data();
        msg.switch_to_next_tree = switch_to_next_tree ? 1 : 0;
        msg.is_first_tree = is_first_tree ? 1 : 0;
        return &msg;
    }

    // ---- iteration 0, dt/DTMaster.java:759-820
    const sb_iter_in* build_initial_master_params() {
        todo_nodes.clear();
        if (!is_gbdt) {
            int64_t curr_mem = 0;
            int key = 0;
            for (auto& tn : trees) {
                tn->features = get_subsampling_features();
                curr_mem += get_stats_mem(tn->features);
                if (curr_mem < p.max_stats_memory) todo_nodes.push_back({key++, tn->tree_id, tn->node, tn->features});
                else to_do_queue.push_back({tn->tree_id, tn->node});
            }
            switch_to_next_tree = false;
            is_first_tree = false;
        } else {
            TreeNode* tn = trees.back().get();
            tn->features = get_subsampling_features();
            todo_nodes.push_back({0, tn->tree_id, tn->node, tn->features});
            switch_to_next_tree = true;
            is_first_tree = trees.size() == 1;
        }
        return message();
    }

    void populate_gain_info_to_node(int tree_id, Node* n, const sb_gain& g, const uint8_t* cat_bits, size_t idx) {  // :710-727
        n->has_predict = true;
        n->predict = g.predict;
        n->predict_class = (int8_t)g.node_class;
        n->feature = g.feature;
        n->split_type = g.split_type;
        if (g.split_type == SB_SPLIT_CONTINUOUS) {
            n->threshold = boundaries_of(g.feature)[g.split_bin + 1];  // dt/Impurity.java:200
        } else {
            n->is_left = g.is_left != 0;
            if (cat_bits) n->cat_bits.assign(cat_bits + idx * SB_MAX_CAT_BYTES, cat_bits + idx * SB_MAX_CAT_BYTES + g.cat_bytes);
        }
        n->gain = g.gain; n->impurity = g.impurity;
        n->left_impurity = g.left_impurity; n->right_impurity = g.right_impurity;
        n->left_predict = g.left_predict; n->right_predict = g.right_predict;
        n->wgt_cnt = g.wgt_cnt;
        build_left_mask(*n);
        if (n->id == 1) trees[(size_t)tree_id]->root_wgt_cnt = g.wgt_cnt;
        else n->wgt_cnt_ratio = g.wgt_cnt / trees[(size_t)tree_id]->root_wgt_cnt;
    }

    Node* add_child(Node* n, int side) {
        auto c = std::make_unique<Node>();
        c->id = 2 * n->id + side;
        c->has_predict = true;
        c->predict = side == 0 ? n->left_predict : n->right_predict;
        c->impurity = side == 0 ? n->left_impurity : n->right_impurity;
        Node* raw = c.get();
        if (side == 0) n->left = std::move(c);
        else n->right = std::move(c);
        return raw;
    }

    void split_node_for_level_wised_tree(bool is_leaf, int tree_id, Node* n) {  // :568-605
        if (is_leaf) return;
        const int level = index_to_level(n->id);
        for (int side = 0; side < 2; ++side) {
            const double imp = side == 0 ? n->left_impurity : n->right_impurity;
            const bool child_is_leaf = level + 1 == p.max_depth || is_pos_zero(imp);
            Node* c = add_child(n, side);
            trees[(size_t)tree_id]->node_num += 1;
            if (!child_is_leaf) to_do_queue.push_back({tree_id, c});
        }
    }
    void split_node_for_leaf_wised_tree(int tree_id, Node* n) {  // :543-562
        const bool over = trees[(size_t)tree_id]->node_num + 1 > p.max_leaves;
        if (!over && !is_pos_zero(n->left_impurity)) {
            for (int side = 0; side < 2; ++side) {
                Node* c = add_child(n, side);
                trees[(size_t)tree_id]->node_num += 1;
                to_do_queue.push_back({tree_id, c});
            }
        }
    }

    // ---- dt/DTMaster.java:316-537 with the merged statistics already reduced to gain records on the device
    const sb_iter_in* do_compute(const sb_gain* gains, const uint8_t* cat_bits) {
        SplitQueue to_split;
        for (size_t idx = 0; idx < todo_nodes.size(); ++idx) {
            const sb_gain& g = gains[idx];
            const int tree_id = todo_nodes[idx].tree_id;
            Node* done = todo_nodes[idx].node;
            if (!g.found) { done->is_leaf = true; continue; }  // :335-339
            populate_gain_info_to_node(tree_id, done, g, cat_bits, idx);
            if (is_leaf_wise) {
                if (!(g.gain <= 0.0)) to_split.offer({done->wgt_cnt_ratio * g.gain, tree_id, done});  // priority :1064-1070
            } else {
                const bool is_leaf = g.gain <= 0.0 || index_to_level(done->id) == p.max_depth;
                done->is_leaf = is_leaf;
                split_node_for_level_wised_tree(is_leaf, tree_id, done);
            }
        }
        if (is_leaf_wise) {  // the reference never increments currSplitIndex (:359-363): the whole queue is split
            while (!to_split.empty()) {
                const SplitQueue::Item it = to_split.poll();
                split_node_for_leaf_wised_tree(it.tree_id, it.node);
            }
        }
        todo_nodes.clear();
        if (to_do_queue.empty()) {
            if (is_gbdt) {
                TreeNode* tree = trees.back().get();
                if ((int)trees.size() >= p.tree_num) halt = true;                                   // :396-399
                else if (!tree->root->left && !tree->root->right) halt = true;                       // :400-407
                else {
                    is_first_tree = trees.size() == 1;                                               // :418
                    tree->features.clear();
                    auto nt = std::make_unique<TreeNode>();
                    nt->tree_id = (int)trees.size();
                    nt->root = std::make_unique<Node>();
                    nt->node = nt->root.get();
                    nt->learning_rate = p.learning_rate;
                    nt->features = get_subsampling_features();
                    todo_nodes.push_back({0, nt->tree_id, nt->node, nt->features});
                    trees.push_back(std::move(nt));
                    switch_to_next_tree = true;
                }
            } else {
                halt = true;                                                                         // :431-435
            }
        } else {
            int64_t curr_mem = 0;
            int key = 0;
            while (!to_do_queue.empty() && curr_mem <= p.max_stats_memory) {                          // :450-470
                QItem q = to_do_queue.front();
                to_do_queue.pop_front();
                std::vector<int> feats = get_subsampling_features();
                curr_mem += get_stats_mem(feats);
                todo_nodes.push_back({key++, q.tree_id, q.node, std::move(feats)});
            }
            switch_to_next_tree = false;
            is_first_tree = trees.size() == 1;                                                       // :475
        }
        if (halt) return nullptr;
        return message();
    }
};

extern "C" {

SB_API int sbh_master_create(const sb_config* cfg, const sbh_params* params, sbh_master** out) {
    if (!cfg || !params || !out) return sb::fail(SB_ERR_INVALID, "null argument");
    auto m = std::make_unique<sbh_master>();
    m->F = cfg->n_features;
    m->n_bins.assign(cfg->n_bins, cfg->n_bins + cfg->n_features);
    m->is_cat.assign((size_t)cfg->n_features, 0);
    if (cfg->is_categorical) m->is_cat.assign(cfg->is_categorical, cfg->is_categorical + cfg->n_features);
    m->bb_off.assign((size_t)cfg->n_features + 1, 0);
    for (int f = 0; f < cfg->n_features; ++f) m->bb_off[(size_t)f + 1] = m->bb_off[(size_t)f] + (m->is_cat[(size_t)f] ? 0 : m->n_bins[(size_t)f]);
    if (!cfg->bin_boundary) return sb::fail(SB_ERR_INVALID, "sbh_master_create needs sb_config.bin_boundary (split thresholds)");
    m->bb.assign(cfg->bin_boundary, cfg->bin_boundary + m->bb_off[(size_t)cfg->n_features]);
    const bool variance = cfg->impurity == SB_IMPURITY_VARIANCE || cfg->impurity == SB_IMPURITY_FRIEDMANMSE;
    m->stats_size = variance ? 3 : cfg->num_classes;
    m->p = *params;
    m->is_gbdt = params->is_gbdt != 0;
    m->is_leaf_wise = params->max_leaves > 0;  // dt/DTMaster.java:980-982
    for (int f = 0; f < cfg->n_features; ++f)  // getAllFeatureList (dt/DTMaster.java:852-878), as feature slots
        if ((!m->is_cat[(size_t)f] && m->n_bins[(size_t)f] > 1) || (m->is_cat[(size_t)f] && m->n_bins[(size_t)f] - 1 > 0)) m->all_features.push_back(f);
    // the reference uses an unseeded Random (dt/DTMaster.java:246); a seed is injected for reproducibility
    m->feature_sampling_random = JavaRandom(params->feature_sampling_seed);
    const int n_init = m->is_gbdt ? 1 : params->tree_num;  // :1078-1080, :1108
    for (int i = 0; i < n_init; ++i) {
        auto t = std::make_unique<TreeNode>();
        t->tree_id = i;
        t->root = std::make_unique<Node>();
        t->node = t->root.get();
        t->learning_rate = 1.0;
        m->trees.push_back(std::move(t));
    }
    *out = m.release();
    return SB_OK;
}
SB_API void sbh_master_destroy(sbh_master* m) { delete m; }
SB_API const sb_iter_in* sbh_master_initial(sbh_master* m) { return m ? m->build_initial_master_params() : nullptr; }
SB_API const sb_iter_in* sbh_master_do_compute(sbh_master* m, const sb_gain* gains, const uint8_t* cat_bits) {
    return (m && gains) ? m->do_compute(gains, cat_bits) : nullptr;
}
SB_API int sbh_grow_tree(sbh_master* m, sb_handle* worker, const sb_iter_in* first, const sb_iter_in** next,
                         int32_t* n_passes, sb_iter_out* last) {
    if (!m || !worker || !first || !next) return SB_ERR_INVALID;
    const size_t tree0 = m->trees.size();
    const sb_iter_in* it = first;
    std::vector<sb_gain> gains;
    std::vector<uint8_t> bits;
    bool any_cat = false;
    for (uint8_t c : m->is_cat) any_cat |= c != 0;
    int passes = 0;
    sb_iter_out out{};
    while (it) {
        int rc = sb_compute_begin(worker, it);
        if (rc != SB_OK) return rc;
        rc = sb_compute_stats(worker);
        if (rc != SB_OK) return rc;
        const size_t n = (size_t)std::max(it->n_todo, 1);
        gains.resize(n);
        // category bits travel only when a categorical feature exists (4 KB per todo node: 52 MB for a forest level of
        // 12800 nodes, zero-filled by resize)
        if (any_cat) bits.resize(n * SB_MAX_CAT_BYTES);
        uint8_t* bits_p = any_cat ? bits.data() : nullptr;
        out = sb_iter_out{};
        rc = sb_compute_finish_split(worker, &out, gains.data(), bits_p);
        if (rc != SB_OK) return rc;
        ++passes;
        it = m->do_compute(gains.data(), bits_p);
        if (m->trees.size() != tree0) break;
    }
    *next = it;
    if (n_passes) *n_passes = passes;
    if (last) *last = out;
    return SB_OK;
}

SB_API int sbh_master_tree_count(sbh_master* m) { return m ? (int)m->trees.size() : 0; }
SB_API int sbh_master_halted(sbh_master* m) { return m && m->halt ? 1 : 0; }

// Flattened copy of tree `index` (sb_node records + mask pool) and, per node, the GainInfo fields Node keeps
// (gain, impurity, left/right impurity, left/right predict, wgtCnt -- 7 doubles per node), for model export / tests.
SB_API int sbh_master_export_tree(sbh_master* m, int index, sb_node* nodes, int node_cap, uint8_t* pool, int pool_cap,
                                  double* extras, int* n_nodes, int* pool_bytes) {
    if (!m || index < 0 || index >= (int)m->trees.size() || !n_nodes || !pool_bytes) return sb::fail(SB_ERR_INVALID, "bad argument");
    std::vector<sb_node> out;
    std::vector<uint8_t> pl;
    m->flatten(m->trees[(size_t)index]->root.get(), out, pl);
    *n_nodes = (int)out.size();
    *pool_bytes = (int)pl.size();
    if (!nodes) return SB_OK;
    if ((int)out.size() > node_cap || (int)pl.size() > pool_cap) return sb::fail(SB_ERR_INVALID, "export buffers too small");
    std::memcpy(nodes, out.data(), out.size() * sizeof(sb_node));
    if (pool && !pl.empty()) std::memcpy(pool, pl.data(), pl.size());
    if (extras) {
        std::vector<const Node*> stack{m->trees[(size_t)index]->root.get()};
        size_t i = 0;
        // same pre-order as flatten()
        std::vector<const Node*> order;
        while (!stack.empty()) {
            const Node* n = stack.back();
            stack.pop_back();
            order.push_back(n);
            if (n->right) stack.push_back(n->right.get());
            if (n->left) stack.push_back(n->left.get());
        }
        for (const Node* n : order) {
            double* e = extras + 7 * i++;
            e[0] = n->gain; e[1] = n->impurity; e[2] = n->left_impurity; e[3] = n->right_impurity;
            e[4] = n->left_predict; e[5] = n->right_predict; e[6] = n->wgt_cnt;
        }
    }
    return SB_OK;
}

}  // extern "C"
