// oracle_master.cpp -- TEST INFRASTRUCTURE ONLY (part of liboracle.so; see the header of shifu_oracle.cpp).
//
// CPU restatement of the reference's master-side tree growth, method by method:
//   DTMaster.doCompute                 dt/DTMaster.java:274-537
//   splitNodeForLeafWisedTree          dt/DTMaster.java:543-562
//   splitNodeForLevelWisedTree         dt/DTMaster.java:568-605
//   populateGainInfoToNode             dt/DTMaster.java:710-727
//   getStatsMem                        dt/DTMaster.java:729-747
//   buildInitialMasterParams           dt/DTMaster.java:759-820
//   getSubsamplingFeatures             dt/DTMaster.java:822-850
//   getAllFeatureList                  dt/DTMaster.java:852-878
//   sampleFeaturesForNodeStats         dt/DTMaster.java:880-893
//   init (queues, trees)               dt/DTMaster.java:1062-1112
// with the Java containers restated too, because their iteration order decides the order of todo nodes:
// java.util.LinkedList as a FIFO, java.util.PriorityQueue with its binary-heap sift-up / sift-down (ties leave the
// heap in an order a stable sort would not give), HashMap<Integer, ...> with small non-negative keys (ascending).
//
// What it consumes: where the reference calls Impurity.computeImpurity + GainInfo.getGainInfoByMaxGain per todo node
// (dt/DTMaster.java:322-334), this restatement takes the resulting GainInfo as an sb_gain record (the per-node body is
// restated separately: orc_best_splits in shifu_oracle.cpp).  What it produces: the next DTMasterParams in the
// flattened form of include/shifu_b200.h, so that tests can compare it field by field with the product's master
// (shifu_b200/csrc/host_master.cpp), and the grown trees.
//
// Not restated (master-side features outside SURVEY.md 8): checkpoint / fail-over recovery (:281-294, :1113-1160),
// continuous training (:1077-1100), DTEarlyStopDecider and validationTolerance (:366-382, :408-415) -- a run that
// enables them is not comparable.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <deque>
#include <memory>
#include <vector>

#include "../include/shifu_b200.h"

namespace {

// java.util.Random: next(bits) and nextDouble(), as published in the class documentation.
struct JRandom {
    int64_t seed;
    explicit JRandom(int64_t s) : seed((s ^ 0x5DEECE66DLL) & ((1LL << 48) - 1)) {}
    int next(int bits) {
        seed = (seed * 0x5DEECE66DLL + 0xBLL) & ((1LL << 48) - 1);
        return (int)(seed >> (48 - bits));
    }
    double nextDouble() { return (double)(((int64_t)next(26) << 27) + next(27)) * 0x1.0p-53; }
};

// java.lang.Double.compare: -0.0 < 0.0, NaN above everything (and equal to itself)
int doubleCompare(double a, double b) {
    if (a < b) return -1;
    if (a > b) return 1;
    int64_t x, y;
    std::memcpy(&x, &a, 8);
    std::memcpy(&y, &b, 8);
    if (a != a) x = 0x7ff8000000000000LL;  // doubleToLongBits canonicalises NaN
    if (b != b) y = 0x7ff8000000000000LL;
    return x == y ? 0 : (x < y ? -1 : 1);
}

struct MNode {  // dt/Node.java: the fields the master reads and writes
    int id = 1;
    bool hasSplit = false;       // Node.split != null
    int splitFeature = -1;       // Split.columnNum, as a feature slot
    int splitType = SB_SPLIT_CONTINUOUS;
    double threshold = 0.0;
    bool isLeftCategories = false;
    std::vector<uint8_t> categories;  // SimpleBitSet words of Split.leftOrRightCategories
    std::unique_ptr<MNode> left, right;
    bool hasPredict = false;     // Node.predict != null
    double predict = 0.0;
    double gain = 0.0, impurity = 0.0, leftImpurity = 0.0, rightImpurity = 0.0, leftPredict = 0.0, rightPredict = 0.0;
    double wgtCnt = 0.0;
    double wgtCntRatio = 0.0;    // NodeStats field default: a root never gets one (dt/DTMaster.java:721-726)
    bool isLeaf = true;
};

struct MTreeNode {  // dt/TreeNode.java; for a queue entry `node` points into some tree
    int treeId = 0;
    MNode* node = nullptr;
    std::vector<int> features;
    bool featuresNull = true;
};

struct MTree {  // the entry of DTMaster.trees
    int treeId = 0;
    std::unique_ptr<MNode> root;
    double learningRate = 1.0;
    int nodeNum = 1;            // TreeNode.nodeNum starts at 1 (dt/TreeNode.java:62)
    double rootWgtCnt = 0.0;
    std::vector<int> features;
};

int indexToLevel(int index) {  // dt/Node.java:459-461: Integer.SIZE - Integer.numberOfLeadingZeros(index)
    int n = 0;
    for (unsigned v = (unsigned)index; v; v >>= 1) ++n;
    return n;
}

// java.util.PriorityQueue<TreeNode> with the comparator of dt/DTMaster.java:1064-1070 (larger wgtCntRatio * gain first)
struct SplitHeap {
    std::vector<MTreeNode> q;
    static int cmp(const MTreeNode& o1, const MTreeNode& o2) {
        return doubleCompare(o2.node->wgtCntRatio * o2.node->gain, o1.node->wgtCntRatio * o1.node->gain);
    }
    bool isEmpty() const { return q.empty(); }
    void offer(const MTreeNode& e) {  // PriorityQueue.offer -> siftUpUsingComparator
        size_t k = q.size();
        q.push_back(e);
        while (k > 0) {
            const size_t parent = (k - 1) >> 1;
            if (cmp(e, q[parent]) >= 0) break;
            q[k] = q[parent];
            k = parent;
        }
        q[k] = e;
    }
    MTreeNode poll() {  // PriorityQueue.poll -> siftDownUsingComparator
        const MTreeNode result = q[0];
        const MTreeNode x = q.back();
        q.pop_back();
        const size_t n = q.size();
        if (n > 0) {
            size_t k = 0;
            const size_t half = n >> 1;
            while (k < half) {
                size_t child = (k << 1) + 1;
                const size_t right = child + 1;
                if (right < n && cmp(q[child], q[right]) > 0) child = right;
                if (cmp(x, q[child]) <= 0) break;
                q[k] = q[child];
                k = child;
            }
            q[k] = x;
        }
        return result;
    }
};

}  // namespace

struct orc_master {
    // ---- what init() reads (dt/DTMaster.java:932-1051)
    int inputNum = 0;  // ColumnConfig count of the feature slots (all slots are inputs here)
    std::vector<int> binCount;      // numeric: binBoundary.size(); categorical: binCategory.size() + 1
    std::vector<uint8_t> categorical;
    std::vector<std::vector<double>> boundaries;
    int statsSize = 3;
    sbh_params prm{};
    bool isGBDT = true, isRF = false, isLeafWise = false;
    std::vector<int> allFeatures;
    JRandom featureSamplingRandom{0};
    // ---- state
    std::vector<std::unique_ptr<MTree>> trees;
    std::deque<MTreeNode> toDoQueue;  // LinkedList used as a Queue: offer = addLast, poll = removeFirst
    SplitHeap toSplitQueue;
    std::vector<MTreeNode> todoNodes;  // the HashMap<Integer, TreeNode> of the last message, key = position
    bool halt = false, switchToNextTree = false, firstTree = false;
    // ---- flattened message
    std::vector<std::vector<sb_node>> outNodes;
    std::vector<std::vector<uint8_t>> outPools;
    std::vector<sb_tree> outTrees;
    std::vector<sb_todo_node> outTodo;
    std::vector<int32_t> outFeatures;
    sb_iter_in out{};

    std::vector<int> sampleFeaturesForNodeStats(int sample) {  // :880-893
        std::vector<int> features;
        for (int i = 0; i < sample; i++) features.push_back(allFeatures.at((size_t)i));
        for (int i = sample; i < (int)allFeatures.size(); i++) {
            const int replacementIndex = (int)(featureSamplingRandom.nextDouble() * i);
            if (replacementIndex >= 0 && replacementIndex < sample) features[(size_t)replacementIndex] = allFeatures[(size_t)i];
        }
        return features;
    }
    std::vector<int> getSubsamplingFeatures() {  // :822-850 (strategy null -> rate)
        const int size = (int)allFeatures.size();
        switch (prm.feature_subset_strategy) {
            case SBH_FSS_RATE: return sampleFeaturesForNodeStats((int)(size * prm.feature_subset_rate));
            case SBH_FSS_HALF: return sampleFeaturesForNodeStats(size / 2);
            case SBH_FSS_ONETHIRD: return sampleFeaturesForNodeStats(size / 3);
            case SBH_FSS_TWOTHIRDS: return sampleFeaturesForNodeStats(size * 2 / 3);
            case SBH_FSS_SQRT: return sampleFeaturesForNodeStats((int)(size * std::sqrt((double)inputNum) / inputNum));
            case SBH_FSS_LOG2: return sampleFeaturesForNodeStats((int)(size * std::log((double)inputNum) / std::log(2.0) / inputNum));
            case SBH_FSS_AUTO:
                if (prm.tree_num > 1) return sampleFeaturesForNodeStats(size / 2);
                return {};
            default: return {};  // ALL
        }
    }
    int64_t getStatsMem(const std::vector<int>& subsetFeatures) {  // :729-747
        int64_t statsMem = 0;
        const std::vector<int>& tempFeatures = subsetFeatures.empty() ? allFeatures : subsetFeatures;
        for (int columnNum : tempFeatures) statsMem += (int64_t)binCount[(size_t)columnNum] * statsSize * 8LL * 2;
        statsMem = statsMem * prm.worker_number / 2;
        return statsMem;
    }

    void populateGainInfoToNode(int treeId, MNode* doneNode, const sb_gain& g, const uint8_t* bits) {  // :710-727
        doneNode->hasPredict = true;
        doneNode->predict = g.predict;
        doneNode->hasSplit = true;
        doneNode->splitFeature = g.feature;
        doneNode->splitType = g.split_type;
        if (g.split_type == SB_SPLIT_CONTINUOUS) {
            doneNode->threshold = boundaries[(size_t)g.feature].at((size_t)g.split_bin + 1);  // dt/Impurity.java:200
        } else {
            doneNode->isLeftCategories = g.is_left != 0;
            doneNode->categories.assign(bits, bits + g.cat_bytes);
        }
        doneNode->gain = g.gain;
        doneNode->impurity = g.impurity;
        doneNode->leftImpurity = g.left_impurity;
        doneNode->rightImpurity = g.right_impurity;
        doneNode->leftPredict = g.left_predict;
        doneNode->rightPredict = g.right_predict;
        doneNode->wgtCnt = g.wgt_cnt;
        if (doneNode->id == SB_ROOT_INDEX) trees[(size_t)treeId]->rootWgtCnt = g.wgt_cnt;
        else doneNode->wgtCntRatio = g.wgt_cnt / trees[(size_t)treeId]->rootWgtCnt;
    }

    MNode* newChild(MNode* parent, bool leftSide) {  // new Node(index, predict, impurity, true)
        std::unique_ptr<MNode> n(new MNode());
        n->id = leftSide ? parent->id * 2 : parent->id * 2 + 1;  // Node.leftIndex / rightIndex
        n->hasPredict = true;
        n->predict = leftSide ? parent->leftPredict : parent->rightPredict;
        n->impurity = leftSide ? parent->leftImpurity : parent->rightImpurity;
        n->isLeaf = true;
        MNode* raw = n.get();
        (leftSide ? parent->left : parent->right) = std::move(n);
        return raw;
    }

    void splitNodeForLeafWisedTree(int treeId, MNode* doneNode) {  // :543-562
        const bool isOverMaxLeaves = trees[(size_t)treeId]->nodeNum + 1 > prm.max_leaves;
        const bool canSplit = !isOverMaxLeaves && doubleCompare(doneNode->leftImpurity, 0.0) != 0;
        if (canSplit) {
            MNode* left = newChild(doneNode, true);
            trees[(size_t)treeId]->nodeNum += 1;
            toDoQueue.push_back(MTreeNode{treeId, left, {}, true});
            MNode* right = newChild(doneNode, false);
            trees[(size_t)treeId]->nodeNum += 1;
            toDoQueue.push_back(MTreeNode{treeId, right, {}, true});
        }
    }

    void splitNodeForLevelWisedTree(bool isLeaf, int treeId, MNode* doneNode) {  // :568-605
        if (!isLeaf) {
            const bool leftChildIsLeaf =
                indexToLevel(doneNode->id) + 1 == prm.max_depth || doubleCompare(doneNode->leftImpurity, 0.0) == 0;
            MNode* left = newChild(doneNode, true);
            trees[(size_t)treeId]->nodeNum += 1;
            if (!leftChildIsLeaf) toDoQueue.push_back(MTreeNode{treeId, left, {}, true});
            const bool rightChildIsLeaf =
                indexToLevel(doneNode->id) + 1 == prm.max_depth || doubleCompare(doneNode->rightImpurity, 0.0) == 0;
            MNode* right = newChild(doneNode, false);
            trees[(size_t)treeId]->nodeNum += 1;
            if (!rightChildIsLeaf) toDoQueue.push_back(MTreeNode{treeId, right, {}, true});
        }
    }

    // ---- the message (our flattened DTMasterParams).  Which trees travel: dt/DTMaster.java:479-513, :797, :809-811.
    // left_mask of a split node: bit b set <=> DTWorker.predictNodeIndex sends bin b to the left child
    // (dt/DTWorker.java:1052-1088: continuous -> binBoundary[b] < threshold ... see below; categorical -> membership)
    std::vector<uint8_t> leftMaskOf(const MNode& n) const {
        const int f = n.splitFeature, nb = binCount[(size_t)f];
        std::vector<uint8_t> m((size_t)(nb / 8 + 1), 0);
        for (int b = 0; b < nb; ++b) {
            bool goesLeft;
            if (n.splitType == SB_SPLIT_CONTINUOUS) {
                // the worker compares the bin's lower boundary with the threshold (a threshold IS a boundary value)
                goesLeft = boundaries[(size_t)f][(size_t)b] < n.threshold;
            } else {
                // short indexValue = binCategory.size() when the bin index is out of [0, binCategory.size())
                const int nCats = nb - 1;
                const int indexValue = (b >= 0 && b < nCats) ? b : nCats;
                const size_t word = (size_t)indexValue / 8;
                const bool contains = word < n.categories.size() && ((n.categories[word] >> (indexValue % 8)) & 1) != 0;
                goesLeft = n.isLeftCategories ? contains : !contains;
            }
            if (goesLeft) m[(size_t)b / 8] |= (uint8_t)(1u << (b % 8));
        }
        return m;
    }
    int32_t emit(const MNode* n, std::vector<sb_node>& nodes, std::vector<uint8_t>& pool) const {
        const int32_t me = (int32_t)nodes.size();
        sb_node s;
        std::memset(&s, 0, sizeof(s));
        s.id = n->id;
        s.left = s.right = -1;
        s.feature = n->hasSplit ? n->splitFeature : -1;
        s.split_type = n->splitType;
        s.has_predict = n->hasPredict ? 1 : 0;
        s.predict = n->predict;
        s.threshold = n->threshold;
        s.is_left = n->isLeftCategories ? 1 : 0;
        if (n->hasSplit) {
            const std::vector<uint8_t> mask = leftMaskOf(*n);
            s.left_mask_offset = (int32_t)pool.size();
            s.left_mask_bytes = (int32_t)mask.size();
            pool.insert(pool.end(), mask.begin(), mask.end());
            s.cat_offset = (int32_t)pool.size();
            s.cat_bytes = (int32_t)n->categories.size();
            pool.insert(pool.end(), n->categories.begin(), n->categories.end());
        }
        nodes.push_back(s);
        if (n->left) { const int32_t c = emit(n->left.get(), nodes, pool); nodes[(size_t)me].left = c; }
        if (n->right) { const int32_t c = emit(n->right.get(), nodes, pool); nodes[(size_t)me].right = c; }
        return me;
    }
    const sb_iter_in* flatten() {
        std::vector<const MTree*> sent;
        if (isGBDT) {
            if (switchToNextTree && trees.size() >= 2) {  // subList(size - 2, size), :481-483
                sent.push_back(trees[trees.size() - 2].get());
                sent.push_back(trees[trees.size() - 1].get());
            } else {
                sent.push_back(trees.back().get());       // subList(size - 1, size), :484-486, :809-811
            }
        } else {
            // RF: every tree travels.  The reference replaces trees without todo nodes by an INVALID_INDEX mock to save
            // message space (:489-507); the worker reads the same trees either way, and the flattened message keeps them.
            for (auto& t : trees) sent.push_back(t.get());
        }
        outNodes.assign(sent.size(), {});
        outPools.assign(sent.size(), {});
        outTrees.assign(sent.size(), sb_tree{});
        for (size_t i = 0; i < sent.size(); ++i) {
            emit(sent[i]->root.get(), outNodes[i], outPools[i]);
            outPools[i].push_back(0);
            outTrees[i].tree_id = sent[i]->treeId;
            outTrees[i].n_nodes = (int32_t)outNodes[i].size();
            outTrees[i].nodes = outNodes[i].data();
            outTrees[i].mask_pool = outPools[i].data();
            outTrees[i].mask_pool_bytes = (int32_t)outPools[i].size() - 1;
        }
        size_t nf = 0;
        for (auto& t : todoNodes) nf += t.features.size();
        outFeatures.assign(nf ? nf : 1, 0);
        outTodo.assign(todoNodes.size(), sb_todo_node{});
        size_t off = 0;
        for (size_t i = 0; i < todoNodes.size(); ++i) {
            outTodo[i].key = (int32_t)i;
            outTodo[i].tree_id = todoNodes[i].treeId;
            outTodo[i].node_id = todoNodes[i].node->id;
            outTodo[i].n_features = (int32_t)todoNodes[i].features.size();
            outTodo[i].features = outFeatures.data() + off;
            for (int f : todoNodes[i].features) outFeatures[off++] = f;
        }
        std::memset(&out, 0, sizeof(out));
        out.n_trees = (int32_t)outTrees.size();
        out.trees = outTrees.data();
        out.n_todo = (int32_t)outTodo.size();
        out.todo = outTodo.data();
        out.switch_to_next_tree = switchToNextTree ? 1 : 0;
        out.is_first_tree = firstTree ? 1 : 0;
        return &out;
    }

    const sb_iter_in* buildInitialMasterParams() {  // :759-820
        todoNodes.clear();
        if (isRF) {
            int64_t currMem = 0;
            for (auto& tree : trees) {
                std::vector<int> features = getSubsamplingFeatures();
                tree->features = features;
                currMem += getStatsMem(features);
                MTreeNode tn{tree->treeId, tree->root.get(), features, false};
                if (currMem < prm.max_stats_memory) todoNodes.push_back(tn);
                else toDoQueue.push_back(tn);
            }
            switchToNextTree = false;  // DTMasterParams default
            firstTree = false;
        } else if (isGBDT) {
            std::vector<int> features = getSubsamplingFeatures();
            MTree* tree = trees.back().get();
            tree->features = features;
            todoNodes.push_back(MTreeNode{tree->treeId, tree->root.get(), features, false});
            switchToNextTree = true;
            firstTree = trees.size() == 1;
        }
        return flatten();
    }

    const sb_iter_in* doCompute(const sb_gain* gains, const uint8_t* catBits) {  // :316-537
        // nodeStatsMap.entrySet(): Integer keys 0..n-1 in a HashMap iterate in ascending order
        for (size_t key = 0; key < todoNodes.size(); ++key) {
            const int treeId = todoNodes[key].treeId;
            MNode* doneNode = todoNodes[key].node;
            const sb_gain& maxGainInfo = gains[key];
            if (!maxGainInfo.found) {  // maxGainInfo == null, :335-339
                doneNode->isLeaf = true;
                continue;
            }
            populateGainInfoToNode(treeId, doneNode, maxGainInfo, catBits ? catBits + key * SB_MAX_CAT_BYTES : nullptr);
            if (isLeafWise) {
                const bool isNotSplit = maxGainInfo.gain <= 0.0;
                if (!isNotSplit) toSplitQueue.offer(MTreeNode{treeId, doneNode, {}, true});
            } else {
                const bool isLeaf = maxGainInfo.gain <= 0.0 || indexToLevel(doneNode->id) == prm.max_depth;
                doneNode->isLeaf = isLeaf;
                splitNodeForLevelWisedTree(isLeaf, treeId, doneNode);
            }
        }
        if (isLeafWise) {
            const int currSplitIndex = 0;  // never incremented in the reference (:359-363)
            while (!toSplitQueue.isEmpty() && currSplitIndex < prm.max_batch_split_size) {
                const MTreeNode treeNode = toSplitQueue.poll();
                splitNodeForLeafWisedTree(treeNode.treeId, treeNode.node);
            }
        }
        todoNodes.clear();
        if (toDoQueue.empty()) {
            if (isGBDT) {
                MTree* treeNode = trees.back().get();
                const MNode* node = treeNode->root.get();
                if ((int)trees.size() >= prm.tree_num) {
                    halt = true;
                } else if (!node->left && !node->right) {
                    halt = true;
                } else {
                    firstTree = trees.size() == 1;
                    treeNode->features.clear();
                    std::unique_ptr<MTree> newRoot(new MTree());
                    newRoot->treeId = (int)trees.size();
                    newRoot->root.reset(new MNode());
                    newRoot->learningRate = prm.learning_rate;
                    MTree* raw = newRoot.get();
                    trees.push_back(std::move(newRoot));
                    raw->features = getSubsamplingFeatures();
                    todoNodes.push_back(MTreeNode{raw->treeId, raw->root.get(), raw->features, false});
                    switchToNextTree = true;
                }
            } else {
                halt = true;
            }
        } else {
            int64_t currMem = 0;
            while (!toDoQueue.empty() && currMem <= prm.max_stats_memory) {
                MTreeNode node = toDoQueue.front();
                toDoQueue.pop_front();
                std::vector<int> subsetFeatures = getSubsamplingFeatures();
                node.features = subsetFeatures;
                node.featuresNull = false;
                currMem += getStatsMem(subsetFeatures);
                todoNodes.push_back(node);
            }
            switchToNextTree = false;
            firstTree = trees.size() == 1;
        }
        if (halt) return nullptr;
        return flatten();
    }
};

extern "C" {

SB_API int orc_master_create(const sb_config* cfg, const sbh_params* params, orc_master** outp) {
    if (!cfg || !params || !outp || !cfg->bin_boundary) return SB_ERR_INVALID;
    std::unique_ptr<orc_master> m(new orc_master());
    m->inputNum = cfg->n_features;
    m->binCount.assign(cfg->n_bins, cfg->n_bins + cfg->n_features);
    m->categorical.assign((size_t)cfg->n_features, 0);
    if (cfg->is_categorical) m->categorical.assign(cfg->is_categorical, cfg->is_categorical + cfg->n_features);
    const double* bb = cfg->bin_boundary;
    m->boundaries.resize((size_t)cfg->n_features);
    for (int f = 0; f < cfg->n_features; ++f)
        if (!m->categorical[(size_t)f]) {
            m->boundaries[(size_t)f].assign(bb, bb + m->binCount[(size_t)f]);
            bb += m->binCount[(size_t)f];
        }
    m->statsSize = (cfg->impurity == SB_IMPURITY_VARIANCE || cfg->impurity == SB_IMPURITY_FRIEDMANMSE) ? 3 : cfg->num_classes;
    m->prm = *params;
    m->isGBDT = params->is_gbdt != 0;
    m->isRF = !m->isGBDT;
    m->isLeafWise = params->max_leaves > 0;  // :980-982: "MaxLeaves" present and positive
    for (int f = 0; f < cfg->n_features; ++f) {  // getAllFeatureList, :852-878
        const bool numerical = !m->categorical[(size_t)f];
        if ((numerical && m->binCount[(size_t)f] > 1) || (!numerical && m->binCount[(size_t)f] - 1 > 0)) m->allFeatures.push_back(f);
    }
    m->featureSamplingRandom = JRandom(params->feature_sampling_seed);
    const int n = m->isRF ? params->tree_num : 1;  // :1075-1080, :1108
    for (int i = 0; i < n; ++i) {
        std::unique_ptr<MTree> t(new MTree());
        t->treeId = i;
        t->root.reset(new MNode());
        t->learningRate = 1.0;
        m->trees.push_back(std::move(t));
    }
    *outp = m.release();
    return SB_OK;
}
SB_API void orc_master_destroy(orc_master* m) { delete m; }
SB_API const sb_iter_in* orc_master_initial(orc_master* m) { return m->buildInitialMasterParams(); }
SB_API const sb_iter_in* orc_master_do_compute(orc_master* m, const sb_gain* gains, const uint8_t* cat_bits) {
    return m->doCompute(gains, cat_bits);
}
SB_API int orc_master_tree_count(orc_master* m) { return (int)m->trees.size(); }
// pre-order dump of one tree: per node (id, isLeaf, hasSplit, feature, gain, impurity, leftImpurity, rightImpurity,
// leftPredict, rightPredict, wgtCnt, wgtCntRatio, predict) as 13 doubles; returns the node count
SB_API int orc_master_dump_tree(orc_master* m, int index, double* out, int cap_nodes) {
    if (index < 0 || index >= (int)m->trees.size()) return -1;
    std::vector<const MNode*> stack{m->trees[(size_t)index]->root.get()};
    int n = 0;
    while (!stack.empty()) {
        const MNode* nd = stack.back();
        stack.pop_back();
        if (out && n < cap_nodes) {
            double* o = out + 13 * (size_t)n;
            o[0] = nd->id; o[1] = nd->isLeaf; o[2] = nd->hasSplit; o[3] = nd->splitFeature; o[4] = nd->gain;
            o[5] = nd->impurity; o[6] = nd->leftImpurity; o[7] = nd->rightImpurity; o[8] = nd->leftPredict;
            o[9] = nd->rightPredict; o[10] = nd->wgtCnt; o[11] = nd->wgtCntRatio; o[12] = nd->predict;
        }
        ++n;
        if (nd->right) stack.push_back(nd->right.get());
        if (nd->left) stack.push_back(nd->left.get());
    }
    return n;
}

}  // extern "C"
