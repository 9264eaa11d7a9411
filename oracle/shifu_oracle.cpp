/*
 * shifu_oracle.cpp -- CPU restatement of Shifu's GBT/RF training hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it.  The product library (libshifu_b200.so) never links,
 * loads or calls anything in this directory.
 *
 * PARITY UNPINNED: the reference (DevinWu/shifu 0.12.0) is Java + un-vendored Guagua/Hadoop jars and cannot be
 * compiled or run in this image (no JDK), and none of its live tests asserts a histogram value, gain, split
 * or tree.  The restatement below therefore follows the Java source line by line, with
 * each function citing the file:line it restates (paths relative to
 * /root/reference/src/main/java/ml/shifu/shifu/core/dtrain/), and is validated by self-consistency and by
 * independent re-derivations in tests/test_oracle_*.py -- not by reference-produced golden vectors.
 *
 * Java semantics reproduced: IEEE fp32/fp64 with no FMA contraction (build with -ffp-contract=off), float
 * expressions evaluated in float and widened on `double += float`, (int) casts truncating toward zero,
 * Double.compareTo ordering, stable Collections.sort, java.util.Random's 48-bit LCG, HashMap iteration
 * order for Integer keys.
 */
#include "../include/shifu_b200.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <deque>
#include <limits>
#include <memory>
#include <string>
#include <thread>
#include <vector>

namespace {

thread_local std::string g_err;
const double JAVA_DOUBLE_MIN_VALUE = 4.9e-324;  // Double.MIN_VALUE: smallest positive subnormal

// ---------------------------------------------------------------- java.util.Random (48-bit LCG)
struct JavaRandom {
    uint64_t seed;
    explicit JavaRandom(int64_t s) { seed = ((uint64_t)s ^ 0x5DEECE66DULL) & ((1ULL << 48) - 1); }
    int32_t next(int bits) {
        seed = (seed * 0x5DEECE66DULL + 0xBULL) & ((1ULL << 48) - 1);
        return (int32_t)((int64_t)seed >> (48 - bits));
    }
    double nextDouble() {
        int64_t a = (int64_t)next(26);
        int64_t b = (int64_t)next(27);
        return (double)((a << 27) + b) * (1.0 / (double)(1LL << 53));
    }
};

// ---------------------------------------------------------------- dt/SimpleBitSet.java:29-183
struct SimpleBitSet {
    std::vector<uint8_t> words;
    SimpleBitSet() {}
    explicit SimpleBitSet(int bitLen) : words((size_t)(bitLen / 8 + 1), 0) {}  // :38-40
    bool contains(int v) const {                                             // :57-68
        int wi = v >> 3;
        if (wi < 0 || wi >= (int)words.size()) return false;
        return (words[wi] & (1 << (v % 8))) != 0;
    }
    void add(int v) {  // :89-101
        int bi = v >> 3;
        while (bi >= (int)words.size()) words.resize(words.empty() ? 8 : 2 * words.size(), 0);
        words[bi] |= (uint8_t)(1 << (v % 8));
    }
};

// ---------------------------------------------------------------- dt/Predict.java, dt/Split.java, dt/Node.java
struct Predict {
    double predict = 0;
    int8_t classValue = 0;
};
struct Split {
    int feature = -1;  // feature slot (the reference keys by columnNum; slots are 1:1 with columnNum)
    int featureType = SB_SPLIT_CONTINUOUS;
    double threshold = 0;
    bool isLeft = false;
    SimpleBitSet cats;
};
struct Node {
    int id = SB_INVALID_INDEX;
    bool hasSplit = false;
    Split split;
    std::unique_ptr<Node> left, right;
    bool hasPredict = false;
    Predict predict;
    double gain = 0;  // stored as float in the reference (dt/Node.java:384-387)
    double impurity = 0, leftImpurity = 0, rightImpurity = 0;
    Predict leftPredict, rightPredict;
    double wgtCnt = 0, wgtCntRatio = 0;
    bool isLeaf = true;
};
inline int indexToLevel(int nodeIndex) {  // dt/Node.java:459-461
    int hob = 0;
    for (int b = 30; b >= 0; --b)
        if (nodeIndex & (1 << b)) { hob = b; break; }
    return hob + 1;
}
Node* getNode(Node* root, int id) {  // dt/Node.java:496-520 (search by heap path)
    if (id == root->id) return root;
    int depth = indexToLevel(id) - 1;
    Node* cur = root;
    for (int d = depth - 1; d >= 0 && cur; --d) cur = ((id >> d) & 1) ? cur->right.get() : cur->left.get();
    return cur;
}

struct GainInfo {  // dt/GainInfo.java:27-43
    double gain, impurity;
    Predict predict;
    double leftImpurity, rightImpurity;
    Predict leftPredict, rightPredict;
    Split split;
    double wgtCnt;
    int splitBin;  // i of the winning prefix (for sb_gain.split_bin); not a reference field
};

// dt/GainInfo.java:200-214
int maxGainIndex(const std::vector<GainInfo>& list) {
    double maxGain = JAVA_DOUBLE_MIN_VALUE;
    int idx = -1;
    for (size_t i = 0; i < list.size(); ++i) {
        if (list[i].gain > maxGain) { maxGain = list[i].gain; idx = (int)i; }
    }
    return idx;
}

// Double.compareTo: numeric order, -0.0 < +0.0, NaN greatest.
inline bool javaDoubleLess(double a, double b) {
    if (a < b) return true;
    if (a > b) return false;
    if (a == b) {
        if (a == 0.0) return std::signbit(a) && !std::signbit(b);
        return false;
    }
    bool an = std::isnan(a), bn = std::isnan(b);
    return !an && bn;
}

struct Column {
    int columnNum;
    bool categorical;
    int nBins;                        // stats bins
    std::vector<double> binBoundary;  // numeric only
    int nCats() const { return nBins - 1; }  // binCategory.size()
};

// ---------------------------------------------------------------- dt/Impurity.java
struct Impurity {
    int kind;
    int statsSize;
    int minInstancesPerNode;
    double minInfoGain;

    // featureUpdate: Variance :238-242, Entropy :537-540, Gini :719-722
    inline void featureUpdate(double* s, int bin, float label, float significance, float weight) const {
        if (kind == SB_IMPURITY_VARIANCE || kind == SB_IMPURITY_FRIEDMANMSE) {
            s[bin * 3] += (double)(float)(significance * weight);
            s[bin * 3 + 1] += (double)(float)((float)(label * significance) * weight);
            s[bin * 3 + 2] += (double)(float)((float)((float)(label * label) * significance) * weight);
        } else {
            s[bin * statsSize + (int)(float)(label + 0.000001f)] += (double)(float)(significance * weight);
        }
    }

    static double varImpurity(double count, double sum, double sumSquare) {  // :233-235
        return (count != 0.0) ? ((sumSquare - (sum * sum) / count) / count) : 0.0;
    }

    // categorical split bookkeeping shared by all four computeImpurity bodies (:169-198 and copies)
    static void makeCategoricalSplit(const Column& c, int index, SimpleBitSet& leftCategories, int& leftSize,
                                     Split& out) {
        int nCats = c.nCats();
        if (index >= nCats) leftCategories.add(nCats);
        else leftCategories.add(index);
        leftSize += 1;
        bool isLeft = true;
        SimpleBitSet rightCategories;
        if (nCats + 1 <= leftSize * 2) {
            isLeft = false;
            rightCategories = SimpleBitSet(nCats + 1);
            for (int j = 0; j < nCats + 1; ++j)
                if (!leftCategories.contains(j)) rightCategories.add(j);
        }
        out.featureType = SB_SPLIT_CATEGORICAL;
        out.threshold = 0.0;
        out.isLeft = isLeft;
        out.cats = isLeft ? leftCategories : rightCategories;  // copy (:194-196)
    }

    struct Pair { int index; double value; };
    static std::vector<Pair> sortedOrder(std::vector<Pair> v) {  // Collections.sort is a stable merge sort
        std::stable_sort(v.begin(), v.end(), [](const Pair& a, const Pair& b) { return javaDoubleLess(a.value, b.value); });
        return v;
    }

    // Variance :120-211, FriedmanMSE :269-360
    bool computeVariance(const double* stats, int binSize, const Column& c, int feature, bool friedman,
                         GainInfo& best) const {
        double count = 0, sum = 0, sumSquare = 0;
        for (int i = 0; i < binSize; ++i) {
            count += stats[i * 3];
            sum += stats[i * 3 + 1];
            sumSquare += stats[i * 3 + 2];
        }
        double impurity = varImpurity(count, sum, sumSquare);
        Predict predict;
        predict.predict = count == 0.0 ? 0.0 : sum / count;

        double leftCount = 0, leftSum = 0, leftSumSquare = 0, rightCount, rightSum, rightSumSquare;
        std::vector<GainInfo> list;
        SimpleBitSet leftCategories = c.categorical ? SimpleBitSet(c.nCats() + 1) : SimpleBitSet();
        std::vector<Pair> order;
        if (c.categorical) {  // getCategoricalOrderList :213-231
            std::vector<Pair> v((size_t)binSize);
            for (int i = 0; i < binSize; ++i) {
                double binPredict = JAVA_DOUBLE_MIN_VALUE;
                if (stats[i * 3] != 0.0) binPredict = stats[i * 3 + 1] / stats[i * 3];
                v[i] = Pair{i, binPredict};
            }
            order = sortedOrder(std::move(v));
        }
        int leftSize = 0;
        for (int i = 0; i < binSize - 1; ++i) {
            int index = c.categorical ? order[i].index : i;
            leftCount += stats[index * 3];
            leftSum += stats[index * 3 + 1];
            leftSumSquare += stats[index * 3 + 2];
            rightCount = count - leftCount;
            rightSum = sum - leftSum;
            rightSumSquare = sumSquare - leftSumSquare;
            if (leftCount <= minInstancesPerNode || rightCount <= minInstancesPerNode) continue;
            double leftImpurity = varImpurity(leftCount, leftSum, leftSumSquare);
            double rightImpurity = varImpurity(rightCount, rightSum, rightSumSquare);
            double gain;
            if (!friedman) {
                double leftWeight = leftCount / count;
                double rightWeight = rightCount / count;
                gain = impurity - leftWeight * leftImpurity - rightWeight * rightImpurity;
            } else {
                double diff = rightCount * leftSum - leftCount * rightSum;
                gain = (diff * diff) / (leftCount * rightCount * (leftCount + rightCount));
            }
            if (gain <= minInfoGain) continue;
            GainInfo g;
            g.split.feature = feature;
            if (c.categorical) makeCategoricalSplit(c, index, leftCategories, leftSize, g.split);
            else {
                g.split.featureType = SB_SPLIT_CONTINUOUS;
                g.split.threshold = c.binBoundary[(size_t)index + 1];
                g.split.isLeft = false;
            }
            g.gain = gain; g.impurity = impurity; g.predict = predict;
            g.leftImpurity = leftImpurity; g.rightImpurity = rightImpurity;
            g.leftPredict.predict = leftCount == 0.0 ? 0.0 : leftSum / leftCount;
            g.rightPredict.predict = rightCount == 0.0 ? 0.0 : rightSum / rightCount;
            g.wgtCnt = count;
            g.splitBin = i;
            list.push_back(std::move(g));
        }
        int mi = maxGainIndex(list);
        if (mi < 0) return false;
        best = std::move(list[(size_t)mi]);
        return true;
    }

    struct ClsInfo { double sumAll; double indexOfLargestElement; double impurity; };
    // Entropy.getEntropyInterInfo :499-522, Gini.getGiniInfo :683-704
    ClsInfo clsInfo(const std::vector<double>& s) const {
        double sumAll = 0;
        for (double v : s) sumAll += v;
        double impurity = 0.0;
        int idx = -1;
        if (sumAll != 0.0) {
            double maxElement = JAVA_DOUBLE_MIN_VALUE;
            for (size_t i = 0; i < s.size(); ++i) {
                double rate = s[i] / sumAll;
                if (kind == SB_IMPURITY_ENTROPY) {
                    if (rate != 0.0) {
                        impurity -= rate * (std::log(rate) / std::log(2.0));  // log2 :542-544
                        if (s[i] > maxElement) { maxElement = s[i]; idx = (int)i; }
                    }
                } else {
                    impurity -= rate * rate;
                    if (s[i] > maxElement) { maxElement = s[i]; idx = (int)i; }
                }
            }
        }
        return ClsInfo{sumAll, (double)idx, impurity};
    }

    // Entropy :378-477, Gini :563-661 (identical control flow)
    bool computeClass(const double* stats, int statsLen, const Column& c, int feature, GainInfo& best) const {
        int numClasses = statsSize;
        int binSize = statsLen / numClasses;
        std::vector<double> statsByClasses((size_t)numClasses, 0.0);
        for (int i = 0; i < binSize; ++i)
            for (int j = 0; j < numClasses; ++j) statsByClasses[j] += stats[i * numClasses + j];
        std::vector<Pair> order;
        if (c.categorical) {  // :479-497, :663-681: classes 0 and 1 only
            std::vector<Pair> v((size_t)binSize);
            for (int i = 0; i < binSize; ++i) {
                double s = stats[i * numClasses] + stats[i * numClasses + 1];
                double binPredict = 0.0;
                if (s != 0.0) binPredict = stats[i * numClasses + 1] / s;
                v[i] = Pair{i, binPredict};
            }
            order = sortedOrder(std::move(v));
        }
        ClsInfo info = clsInfo(statsByClasses);
        Predict predict;
        predict.predict = info.sumAll == 0.0 ? 0.0 : (statsByClasses[1] / info.sumAll);
        predict.classValue = (int8_t)info.indexOfLargestElement;

        std::vector<double> leftS((size_t)numClasses, 0.0), rightS((size_t)numClasses, 0.0);
        std::vector<GainInfo> list;
        SimpleBitSet leftCategories = c.categorical ? SimpleBitSet(c.nCats() + 1) : SimpleBitSet();
        int leftSize = 0;
        for (int i = 0; i < binSize - 1; ++i) {
            int index = c.categorical ? order[i].index : i;
            for (int j = 0; j < numClasses; ++j) leftS[j] += stats[index * numClasses + j];
            ClsInfo leftInfo = clsInfo(leftS);
            Predict leftPredict;
            leftPredict.predict = leftInfo.sumAll == 0.0 ? 0.0 : (leftS[1] / leftInfo.sumAll);
            leftPredict.classValue = (int8_t)leftInfo.indexOfLargestElement;
            for (int j = 0; j < numClasses; ++j) rightS[j] = statsByClasses[j] - leftS[j];
            ClsInfo rightInfo = clsInfo(rightS);
            if (leftInfo.sumAll <= minInstancesPerNode || rightInfo.sumAll <= minInstancesPerNode) continue;
            Predict rightPredict;
            rightPredict.predict = rightInfo.sumAll == 0.0 ? 0.0 : (rightS[1] / rightInfo.sumAll);
            rightPredict.classValue = (int8_t)rightInfo.indexOfLargestElement;
            double leftWeight = info.sumAll == 0.0 ? 0.0 : (leftInfo.sumAll / info.sumAll);
            double rightWeight = info.sumAll == 0.0 ? 0.0 : (rightInfo.sumAll / info.sumAll);
            double gain = info.impurity - leftWeight * leftInfo.impurity - rightWeight * rightInfo.impurity;
            if (gain <= minInfoGain) continue;
            GainInfo g;
            g.split.feature = feature;
            if (c.categorical) makeCategoricalSplit(c, index, leftCategories, leftSize, g.split);
            else {
                g.split.featureType = SB_SPLIT_CONTINUOUS;
                g.split.threshold = c.binBoundary[(size_t)index + 1];
                g.split.isLeft = false;
            }
            g.gain = gain; g.impurity = info.impurity; g.predict = predict;
            g.leftImpurity = leftInfo.impurity; g.rightImpurity = rightInfo.impurity;
            g.leftPredict = leftPredict; g.rightPredict = rightPredict;
            g.wgtCnt = info.sumAll;
            g.splitBin = i;
            list.push_back(std::move(g));
        }
        int mi = maxGainIndex(list);
        if (mi < 0) return false;
        best = std::move(list[(size_t)mi]);
        return true;
    }

    // masterFriedman=false reproduces the master-side quirk: "friedmanmse" falls through to Variance
    // (dt/DTMaster.java:1026-1032), so distributed training never runs FriedmanMSE.computeImpurity.
    bool computeImpurity(const double* stats, int statsLen, const Column& c, int feature, bool masterQuirk,
                         GainInfo& best) const {
        if (kind == SB_IMPURITY_VARIANCE) return computeVariance(stats, statsLen / 3, c, feature, false, best);
        if (kind == SB_IMPURITY_FRIEDMANMSE) return computeVariance(stats, statsLen / 3, c, feature, !masterQuirk, best);
        return computeClass(stats, statsLen, c, feature, best);
    }
};

// ---------------------------------------------------------------- dt/Loss.java:36-100
inline float lossGradient(int kind, float predict, float label) {
    switch (kind) {
        case SB_LOSS_ABSOLUTE: {  // Float.compare(label, predict) < 0 ? 1f : -1f
            bool less;
            if (label < predict) less = true;
            else if (label > predict) less = false;
            else {
                // Float.compare tie-breaks: -0.0f < 0.0f, NaN greatest
                uint32_t a, b;
                std::memcpy(&a, &label, 4);
                std::memcpy(&b, &predict, 4);
                if (std::isnan(label)) a = 0x7fc00000u;
                if (std::isnan(predict)) b = 0x7fc00000u;
                less = (int32_t)a < (int32_t)b;
            }
            return less ? 1.0f : -1.0f;
        }
        case SB_LOSS_SQUARED: return 2.0f * (float)(predict - label);
        case SB_LOSS_HALFGRADSQUARED: return (float)(predict - label);
        case SB_LOSS_LOG: {
            float num = (float)(2.0f - (float)(4.0f * label));
            float arg = (float)((float)((float)(4.0f * label) * predict) - (float)(2.0f * predict));
            return num / (float)std::exp((double)arg);
        }
    }
    return 0.0f;
}
inline float lossError(int kind, float predict, float label) {
    switch (kind) {
        case SB_LOSS_ABSOLUTE: return std::fabs((float)(label - predict));
        case SB_LOSS_SQUARED:
        case SB_LOSS_HALFGRADSQUARED: {
            float e = (float)(predict - label);
            return (float)(e * e);
        }
        case SB_LOSS_LOG: {
            float arg = (float)((float)(2.0f * predict) - (float)((float)(4.0f * predict) * label));
            return (float)std::log1p(1.0 + std::exp((double)arg));
        }
    }
    return 0.0f;
}

// ---------------------------------------------------------------- dt/DTWorker.java:1001-1034
int getBinIndex(float value, const double* bb, int n) {
    if (n <= 1) return 0;
    if (value == std::numeric_limits<float>::infinity()) return n - 1;
    if (value == -std::numeric_limits<float>::infinity()) return 0;
    int low = 0, high = n - 1;
    while (low <= high) {
        int mid = (int)(((unsigned)low + (unsigned)high) >> 1);
        double lowThreshold = bb[mid];
        double highThreshold = mid == n - 1 ? std::numeric_limits<double>::max() : bb[mid + 1];
        if ((double)value >= lowThreshold && (double)value < highThreshold) return mid;
        if ((double)value >= highThreshold) low = mid + 1;
        else high = mid - 1;
    }
    return -1;
}

// Iteration order of a java.util.HashMap<Integer,?> built with (n, 1f) and filled in `keys` order
// (dt/DTWorker.java:910, dt/DTWorkerParams.java:235; visited at dt/DTMaster.java:324).
std::vector<int> hashMapOrder(const std::vector<int>& keys) {
    int n = (int)keys.size();
    int cap = 1;
    while (cap < n) cap <<= 1;  // tableSizeFor(n) (n == 0 -> 1)
    std::vector<int> pos((size_t)n);
    for (int i = 0; i < n; ++i) pos[i] = i;
    auto bucket = [cap](int k) {
        uint32_t h = (uint32_t)k;
        h ^= (h >> 16);
        return (int)(h & (uint32_t)(cap - 1));
    };
    std::stable_sort(pos.begin(), pos.end(), [&](int a, int b) { return bucket(keys[a]) < bucket(keys[b]); });
    return pos;  // positions into `keys`, in visit order
}

struct Rows {
    int64_t n = 0;
    std::vector<int16_t> inputs;  // row-major [n][F]   (Data.inputs, dt/DTWorker.java:1593)
    std::vector<float> label, output, predict, significance;
    std::vector<float> weights;   // row-major [n][wpr] (Data.subsampleWeights)
};

struct Oracle {
    sb_config cfg;
    std::vector<Column> cols;
    Impurity imp;
    Rows train, valid;
    std::vector<int> allFeatures;  // getAllValidFeatures :965-990 (slots with >1 numeric bins or >0 categories)
};

std::unique_ptr<Node> buildNode(const sb_tree& t, int idx) {
    const sb_node& s = t.nodes[idx];
    auto n = std::make_unique<Node>();
    n->id = s.id;
    n->hasSplit = s.feature >= 0;
    if (n->hasSplit) {
        n->split.feature = s.feature;
        n->split.featureType = s.split_type;
        n->split.threshold = s.threshold;
        n->split.isLeft = s.is_left != 0;
        if (s.split_type == SB_SPLIT_CATEGORICAL && s.cat_bytes > 0)
            n->split.cats.words.assign(t.mask_pool + s.cat_offset, t.mask_pool + s.cat_offset + s.cat_bytes);
    }
    n->hasPredict = s.has_predict != 0;
    n->predict.predict = s.predict;
    if (s.left >= 0) n->left = buildNode(t, s.left);
    if (s.right >= 0) n->right = buildNode(t, s.right);
    return n;
}

// dt/DTWorker.java:1036-1096
const Node* predictNodeIndex(const Oracle& o, const Node* node, const int16_t* inputs) {
    const Node* cur = node;
    for (;;) {
        if (!cur->hasSplit || (!cur->left && !cur->right)) return cur;
        const Split& sp = cur->split;
        const Column& c = o.cols[(size_t)sp.feature];
        const Node* next = nullptr;
        if (!c.categorical) {
            int16_t binIndex = inputs[sp.feature];
            double lowest = c.binBoundary[(size_t)binIndex];
            next = (lowest < sp.threshold) ? cur->left.get() : cur->right.get();
        } else {
            int16_t nCats = (int16_t)c.nCats();
            int16_t indexValue = nCats;
            if (inputs[sp.feature] >= 0 && inputs[sp.feature] < nCats) indexValue = inputs[sp.feature];
            bool has = sp.cats.contains(indexValue);
            if (sp.isLeft) next = has ? cur->left.get() : cur->right.get();
            else next = has ? cur->right.get() : cur->left.get();
        }
        if (!next) return nullptr;  // IllegalStateException in the reference
        cur = next;
    }
}

struct TodoStats {
    int key, treeId, nodeId;
    std::vector<int> features;           // slots, list order (empty list already expanded to all valid)
    std::vector<int64_t> featOffset;     // offset of each feature block in this node's stats
    int64_t offset;                      // offset of the node in the flat stats buffer
    int64_t len;
};

int64_t layoutTodo(const Oracle& o, const sb_iter_in* in, std::vector<TodoStats>& todo) {
    int64_t off = 0;
    todo.clear();
    for (int i = 0; i < in->n_todo; ++i) {
        const sb_todo_node& t = in->todo[i];
        TodoStats ts;
        ts.key = t.key; ts.treeId = t.tree_id; ts.nodeId = t.node_id; ts.offset = off;
        if (t.n_features == 0) ts.features = o.allFeatures;
        else ts.features.assign(t.features, t.features + t.n_features);
        int64_t l = 0;
        for (int f : ts.features) {
            ts.featOffset.push_back(l);
            l += (int64_t)o.cols[(size_t)f].nBins * o.imp.statsSize;  // :915,:919
        }
        ts.len = l;
        off += l;
        todo.push_back(std::move(ts));
    }
    return off;
}

}  // namespace

extern "C" {

SB_API const char* orc_last_error(void) { return g_err.c_str(); }

SB_API int orc_create(const sb_config* cfg, void** out) {
    auto o = new Oracle();
    o->cfg = *cfg;
    const double* bb = cfg->bin_boundary;
    for (int f = 0; f < cfg->n_features; ++f) {
        Column c;
        c.columnNum = cfg->column_num ? cfg->column_num[f] : f;
        c.categorical = cfg->is_categorical && cfg->is_categorical[f];
        c.nBins = cfg->n_bins[f];
        if (!c.categorical) {
            if (bb) { c.binBoundary.assign(bb, bb + c.nBins); bb += c.nBins; }
            else { c.binBoundary.resize((size_t)c.nBins); for (int b = 0; b < c.nBins; ++b) c.binBoundary[b] = b == 0 ? -std::numeric_limits<double>::infinity() : (double)b; }
        }
        // getAllValidFeatures: numeric binBoundary.size() > 1, categorical binCategory.size() > 0
        if ((!c.categorical && c.nBins > 1) || (c.categorical && c.nCats() > 0)) o->allFeatures.push_back(f);
        o->cols.push_back(std::move(c));
    }
    o->imp.kind = cfg->impurity;
    o->imp.statsSize = (cfg->impurity == SB_IMPURITY_VARIANCE || cfg->impurity == SB_IMPURITY_FRIEDMANMSE) ? 3 : cfg->num_classes;
    o->imp.minInstancesPerNode = cfg->min_instances_per_node;
    o->imp.minInfoGain = cfg->min_info_gain;
    o->cfg.n_bins = nullptr; o->cfg.is_categorical = nullptr; o->cfg.column_num = nullptr; o->cfg.bin_boundary = nullptr;
    *out = o;
    return 0;
}
SB_API void orc_destroy(void* p) { delete (Oracle*)p; }

// Rows in the reference's own layout: inputs row-major int16 [n][F], weights row-major [n][wpr].
SB_API int orc_load_rows(void* p, int is_validation, int64_t n, const int16_t* inputs, const float* label,
                         const float* significance, const float* weights) {
    Oracle& o = *(Oracle*)p;
    Rows& r = is_validation ? o.valid : o.train;
    int F = o.cfg.n_features, W = o.cfg.weights_per_row;
    r.inputs.insert(r.inputs.end(), inputs, inputs + n * F);
    r.label.insert(r.label.end(), label, label + n);
    r.output.insert(r.output.end(), label, label + n);   // output = predict = label (:1264-1273)
    r.predict.insert(r.predict.end(), label, label + n);
    if (significance) r.significance.insert(r.significance.end(), significance, significance + n);
    else r.significance.insert(r.significance.end(), (size_t)n, 1.0f);
    if (weights) r.weights.insert(r.weights.end(), weights, weights + n * W);
    else r.weights.insert(r.weights.end(), (size_t)(n * W), 1.0f);
    r.n += n;
    return 0;
}

SB_API int orc_read_rows(void* p, int is_validation, float* predict, float* output, float* weights) {
    Oracle& o = *(Oracle*)p;
    Rows& r = is_validation ? o.valid : o.train;
    if (predict) std::memcpy(predict, r.predict.data(), r.n * sizeof(float));
    if (output) std::memcpy(output, r.output.data(), r.n * sizeof(float));
    if (weights) std::memcpy(weights, r.weights.data(), r.weights.size() * sizeof(float));
    return 0;
}

SB_API int64_t orc_stats_len(void* p, const sb_iter_in* in) {
    std::vector<TodoStats> todo;
    return layoutTodo(*(Oracle*)p, in, todo);
}

/*
 * DTWorker.doCompute, dt/DTWorker.java:558-891.  n_threads = workerThreadCount.  Thread partials are
 * merged in thread-index order (the reference merges in completion order, :865-883, which is
 * non-deterministic; index order is one of its possible outcomes).
 */
SB_API int orc_compute(void* p, const sb_iter_in* in, sb_iter_out* out, int n_threads) {
    Oracle& o = *(Oracle*)p;
    const bool isGBDT = o.cfg.is_gbdt != 0, isRF = !isGBDT;
    const int F = o.cfg.n_features, W = o.cfg.weights_per_row, lossKind = o.cfg.loss;
    std::vector<std::unique_ptr<Node>> trees;
    std::vector<int> treeIds;
    for (int t = 0; t < in->n_trees; ++t) {
        if (in->trees[t].n_nodes == 0) { auto n = std::make_unique<Node>(); n->id = SB_INVALID_INDEX; trees.push_back(std::move(n)); }
        else trees.push_back(buildNode(in->trees[t], 0));
        treeIds.push_back(in->trees[t].tree_id);
    }
    std::vector<TodoStats> todo;
    int64_t statsLen = layoutTodo(o, in, todo);
    if (out->stats && out->stats_len < statsLen) { g_err = "stats buffer too small"; return SB_ERR_INVALID; }

    double trainError = 0, validationError = 0, weightedTrainCount = 0, weightedValidationCount = 0;
    const bool sw = in->switch_to_next_tree != 0, first = in->is_first_tree != 0;

    // ---- phase A (:582-686) and A' (:689-751): single thread, row order
    for (int pass = 0; pass < 2; ++pass) {
        Rows& r = pass == 0 ? o.train : o.valid;
        const bool isTrain = pass == 0;
        for (int64_t i = 0; i < r.n; ++i) {
            const int16_t* inputs = &r.inputs[(size_t)(i * F)];
            if (isRF) {
                for (size_t t = 0; t < trees.size(); ++t) {
                    if (trees[t]->id == SB_INVALID_INDEX) continue;
                    const Node* pn = predictNodeIndex(o, trees[t].get(), inputs);
                    if (!pn) { g_err = "NextNode is null"; return SB_ERR_STATE; }
                    if (pn->hasPredict) {
                        float err = lossError(lossKind, (float)pn->predict.predict, r.label[i]);
                        if (isTrain) {
                            float weight = r.weights[(size_t)(i * W + treeIds[t])];  // [treeId], no modulo (:592)
                            if (weight == 0.0f && !std::signbit(weight)) {             // Float.compare(weight, 0f) == 0
                                validationError += (double)(float)(r.significance[i] * err);
                                weightedValidationCount += (double)r.significance[i];
                            } else {
                                trainError += (double)(float)((float)(weight * r.significance[i]) * err);
                                weightedTrainCount += (double)(float)(weight * r.significance[i]);
                            }
                        } else {
                            validationError += (double)(float)(r.significance[i] * err);
                            weightedValidationCount += (double)r.significance[i];
                        }
                    }
                }
            }
            if (isGBDT) {
                int currTreeIndex = (int)trees.size() - 1;
                if (sw && currTreeIndex >= 1) {
                    const Node* pn = predictNodeIndex(o, trees[(size_t)currTreeIndex - 1].get(), inputs);
                    if (!pn) { g_err = "NextNode is null"; return SB_ERR_STATE; }
                    if (pn->hasPredict) {
                        double predict = pn->predict.predict;
                        if (first) r.predict[i] = (float)predict;
                        else {
                            bool drop = isTrain && in->dropout && in->dropout[i];
                            if (!drop) r.predict[i] = (float)(r.predict[i] + (float)(o.cfg.learning_rate * predict));
                        }
                        r.output[i] = -1.0f * lossGradient(lossKind, r.predict[i], r.label[i]);
                    }
                    if (isTrain && in->new_weights)  // :643-668, randomness injected
                        r.weights[(size_t)(i * W + currTreeIndex % W)] = in->new_weights[i];
                }
                double& err = isTrain ? trainError : validationError;
                double& cnt = isTrain ? weightedTrainCount : weightedValidationCount;
                if (first && !sw) {
                    const Node* pn = predictNodeIndex(o, trees[(size_t)currTreeIndex].get(), inputs);
                    if (!pn) { g_err = "NextNode is null"; return SB_ERR_STATE; }
                    if (pn->hasPredict) {
                        err += (double)(float)(r.significance[i] * lossError(lossKind, (float)pn->predict.predict, r.label[i]));
                        cnt += (double)r.significance[i];
                    }
                } else {
                    err += (double)(float)(r.significance[i] * lossError(lossKind, r.predict[i], r.label[i]));
                    cnt += (double)r.significance[i];
                }
            }
        }
    }

    // ---- phase B (:762-863): workerThreadCount threads over contiguous row ranges (:769-787)
    int64_t realRecords = o.train.n;
    int realThreads = (int)std::min<int64_t>(std::max(1, n_threads), std::max<int64_t>(realRecords, 1));
    std::vector<int64_t> lows((size_t)realThreads), highs((size_t)realThreads);
    if (realRecords > 0) {
        int64_t stepCount = realRecords / realThreads;
        if (realRecords % realThreads != 0) stepCount += (realRecords % realThreads) / stepCount;
        for (int i = 0; i < realThreads; ++i) {
            lows[i] = i * stepCount;
            highs[i] = (i != realThreads - 1) ? lows[i] + stepCount - 1 : realRecords - 1;
        }
    }
    std::vector<std::vector<double>> local((size_t)realThreads);
    std::atomic<int> failed{0};
    auto work = [&](int th) {
        std::vector<double>& st = local[(size_t)th];
        st.assign((size_t)statsLen, 0.0);
        std::vector<int> nodeIndexes(trees.size());
        for (int64_t j = lows[th]; j <= highs[th]; ++j) {
            const int16_t* inputs = &o.train.inputs[(size_t)(j * F)];
            if (isRF) {
                for (size_t t = 0; t < trees.size(); ++t) {
                    if (trees[t]->id == SB_INVALID_INDEX) nodeIndexes[t] = SB_INVALID_INDEX;
                    else {
                        const Node* pn = predictNodeIndex(o, trees[t].get(), inputs);
                        if (!pn) { failed = 1; return; }
                        nodeIndexes[t] = pn->id;
                    }
                }
            } else {
                const Node* pn = predictNodeIndex(o, trees.back().get(), inputs);
                if (!pn) { failed = 1; return; }
                nodeIndexes[0] = pn->id;
            }
            for (const TodoStats& td : todo) {
                int currPredictIndex = isRF ? nodeIndexes[(size_t)td.treeId] : nodeIndexes[0];
                if (td.nodeId != currPredictIndex) continue;
                float weight = o.train.weights[(size_t)(j * W + td.treeId % W)];
                if (weight == 0.0f && !std::signbit(weight)) continue;  // Float.compare(weight, 0f) != 0 (:848)
                for (size_t k = 0; k < td.features.size(); ++k) {
                    int f = td.features[k];
                    o.imp.featureUpdate(&st[(size_t)(td.offset + td.featOffset[k])], inputs[f], o.train.output[j],
                                        o.train.significance[j], weight);
                }
            }
        }
    };
    if (realRecords > 0) {
        if (realThreads == 1) work(0);
        else {
            std::vector<std::thread> ths;
            for (int t = 0; t < realThreads; ++t) ths.emplace_back(work, t);
            for (auto& t : ths) t.join();
        }
    }
    if (failed) { g_err = "NextNode is null"; return SB_ERR_STATE; }
    if (out->stats) {
        // merge (:865-883, mergeNodeStats :893-901): statistics = first map; += the others
        if (realRecords > 0) std::memcpy(out->stats, local[0].data(), (size_t)statsLen * sizeof(double));
        else std::memset(out->stats, 0, (size_t)statsLen * sizeof(double));
        for (int t = 1; t < realThreads && realRecords > 0; ++t)
            for (int64_t i = 0; i < statsLen; ++i) out->stats[i] += local[(size_t)t][(size_t)i];
    }
    out->train_count = weightedTrainCount;
    out->validation_count = weightedValidationCount;
    out->train_error = trainError;
    out->validation_error = validationError;
    return 0;
}

static void fillGain(const GainInfo& g, sb_gain* out, uint8_t* cat_bits) {
    std::memset(out, 0, sizeof(*out));
    out->found = 1;
    out->feature = g.split.feature;
    out->split_type = g.split.featureType;
    out->split_bin = g.splitBin;
    out->is_left = g.split.isLeft ? 1 : 0;
    out->gain = g.gain; out->impurity = g.impurity; out->predict = g.predict.predict;
    out->left_impurity = g.leftImpurity; out->right_impurity = g.rightImpurity;
    out->left_predict = g.leftPredict.predict; out->right_predict = g.rightPredict.predict;
    out->left_class = g.leftPredict.classValue; out->right_class = g.rightPredict.classValue;
    out->node_class = g.predict.classValue;
    out->wgt_cnt = g.wgtCnt;
    if (g.split.featureType == SB_SPLIT_CATEGORICAL) {
        out->cat_bytes = (int32_t)g.split.cats.words.size();
        if (cat_bits) std::memcpy(cat_bits, g.split.cats.words.data(), g.split.cats.words.size());
    }
}

// Impurity.computeImpurity for one feature's double[bins*statsSize].  master_quirk != 0 applies
// dt/DTMaster.java:1026-1032 (friedmanmse -> Variance).
SB_API int orc_compute_impurity(void* p, int feature, const double* stats, int master_quirk, sb_gain* out,
                                uint8_t* cat_bits) {
    Oracle& o = *(Oracle*)p;
    const Column& c = o.cols[(size_t)feature];
    GainInfo g;
    bool ok = o.imp.computeImpurity(stats, c.nBins * o.imp.statsSize, c, feature, master_quirk != 0, g);
    if (!ok) { std::memset(out, 0, sizeof(*out)); return 0; }
    fillGain(g, out, cat_bits);
    return 0;
}

/*
 * The per-node body of DTMaster.doCompute (dt/DTMaster.java:316-340): computeImpurity for every feature of
 * every todo node, visited in HashMap order, then getGainInfoByMaxGain.  `stats` is the flat buffer
 * orc_compute produced for the same sb_iter_in.
 */
SB_API int orc_best_splits(void* p, const sb_iter_in* in, const double* stats, sb_gain* gains, uint8_t* cat_bits) {
    Oracle& o = *(Oracle*)p;
    std::vector<TodoStats> todo;
    layoutTodo(o, in, todo);
    for (size_t n = 0; n < todo.size(); ++n) {
        const TodoStats& td = todo[n];
        std::vector<int> keys;
        for (int f : td.features) keys.push_back(o.cols[(size_t)f].columnNum);
        std::vector<int> order = hashMapOrder(keys);
        std::vector<GainInfo> gainList;
        for (int pos : order) {
            int f = td.features[(size_t)pos];
            const Column& c = o.cols[(size_t)f];
            GainInfo g;
            if (o.imp.computeImpurity(stats + td.offset + td.featOffset[(size_t)pos], c.nBins * o.imp.statsSize, c, f,
                                      true, g))
                gainList.push_back(std::move(g));
        }
        int mi = maxGainIndex(gainList);
        if (mi < 0) std::memset(&gains[n], 0, sizeof(sb_gain));
        else fillGain(gainList[(size_t)mi], &gains[n], cat_bits ? cat_bits + n * SB_MAX_CAT_BYTES : nullptr);
    }
    return 0;
}

// Visit order (positions into the list) of a feature list, for tests of the product's host code.
SB_API int orc_hashmap_order(const int32_t* keys, int n, int32_t* order_out) {
    std::vector<int> k(keys, keys + n);
    std::vector<int> o = hashMapOrder(k);
    for (int i = 0; i < n; ++i) order_out[i] = o[(size_t)i];
    return 0;
}

SB_API int orc_get_bin_index(float value, const double* bb, int n) { return getBinIndex(value, bb, n); }
SB_API float orc_loss_gradient(int kind, float predict, float label) { return lossGradient(kind, predict, label); }
SB_API float orc_loss_error(int kind, float predict, float label) { return lossError(kind, predict, label); }
SB_API void orc_feature_update(int kind, int stats_size, double* stats, int bin, float label, float significance,
                               float weight) {
    Impurity imp{kind, stats_size, 1, 0.0};
    imp.featureUpdate(stats, bin, label, significance, weight);
}
/*
 * IndependentTreeModel.compute(double[] data) over a block of rows (dt/IndependentTreeModel.java:352-446), with
 * predictNode (:516-560) and the score converts (:471-507), on the flattened model of include/shifu_b200.h.
 * Returns 0, or -1 when a walk meets a null child (the reference throws a NullPointerException there).
 */
static int scoreCompareDouble(double a, double b) {  // Double.compare
    if (a < b) return -1;
    if (a > b) return 1;
    int64_t x, y;
    memcpy(&x, &a, 8);
    memcpy(&y, &b, 8);
    if (std::isnan(a)) x = 0x7ff8000000000000LL;  // doubleToLongBits collapses NaNs
    if (std::isnan(b)) y = 0x7ff8000000000000LL;
    return x == y ? 0 : (x < y ? -1 : 1);
}
static int scorePredictNode(const sb_score_model* m, int node, const double* data, bool* npe) {
    while (true) {
        const sb_score_node& cur = m->nodes[node];
        if (cur.column < 0 || (cur.left < 0 && cur.right < 0)) return node;  // split == null || isRealLeaf()
        const double value = data[cur.column];
        int next;
        if (cur.split_type == SB_SPLIT_CONTINUOUS) {
            next = (value < cur.threshold) ? cur.left : cur.right;  // :523-529
        } else {
            const int categoricalSize = m->categorical_size[cur.column];
            int16_t indexValue;
            if (scoreCompareDouble(value, 0.0) < 0 || scoreCompareDouble(value, (double)categoricalSize) >= 0) {
                indexValue = (int16_t)categoricalSize;  // :533-534
            } else {
                const double v = value + 0.1;            // :537, (short) cast = d2i (saturating, NaN -> 0) then i2s
                int32_t i;
                if (std::isnan(v)) i = 0;
                else if (v >= 2147483647.0) i = 2147483647;
                else if (v <= -2147483648.0) i = (int32_t)(-2147483647 - 1);
                else i = (int32_t)v;
                indexValue = (int16_t)i;
            }
            bool contains = false;  // SimpleBitSet.contains (dt/SimpleBitSet.java:89-101)
            if (indexValue >= 0 && (indexValue >> 3) < cur.cat_bytes)
                contains = ((m->cat_pool[cur.cat_offset + (indexValue >> 3)] >> (indexValue % 8)) & 1) != 0;
            if (cur.is_left) next = contains ? cur.left : cur.right;   // :540-546
            else next = contains ? cur.right : cur.left;               // :547-553
        }
        if (next < 0) { *npe = true; return node; }
        node = next;
    }
}
SB_API int orc_score_rows(const sb_score_model* m, int64_t n_rows, const double* data, double* scores) {
    bool npe = false;
    const int out_per_row = m->is_classification ? m->n_trees : 1;
    for (int64_t r = 0; r < n_rows; ++r) {
        const double* row = data + r * m->n_columns;
        double* out = scores + r * out_per_row;
        if (m->is_classification) {  // computeClassificationScore :365-379
            for (int t = 0; t < m->n_trees; ++t) out[t] = (double)m->nodes[scorePredictNode(m, m->tree_root[t], row, &npe)].class_value;
            continue;
        }
        const int bags = m->n_bags;
        double finalPredict = 0.0;
        for (int i = 0; i < bags; ++i) {
            if (m->is_gbdt) {  // :390-420
                double predict = 0.0;
                for (int j = m->bag_begin[i]; j < m->bag_begin[i + 1]; ++j) {
                    const double score = m->nodes[scorePredictNode(m, m->tree_root[j], row, &npe)].predict;
                    predict += score * m->weights[j];
                }
                if (m->convert == SB_CONVERT_OLD_SIGMOID) predict = 1 / (1 + std::min(1.0E19, std::exp(-predict)));
                else if (m->convert == SB_CONVERT_SIGMOID) predict = 1 / (1 + std::min(1.0E19, std::exp(-20 * predict)));
                else if (m->convert == SB_CONVERT_CUTOFF) predict = predict < 0.0 ? 0.0 : (predict > 1.0 ? 1.0 : predict);
                finalPredict += predict;
            } else {           // :423-441
                double predictSum = 0.0, weightSum = 0.0;
                for (int j = m->bag_begin[i]; j < m->bag_begin[i + 1]; ++j) {
                    const double score = m->nodes[scorePredictNode(m, m->tree_root[j], row, &npe)].predict;
                    const double weight = m->weights[j];
                    weightSum += weight;
                    predictSum += score * weight;
                }
                finalPredict += (predictSum / weightSum);
            }
        }
        out[0] = finalPredict / bags;
    }
    return npe ? -1 : 0;
}

SB_API void orc_java_random_doubles(int64_t seed, int n, double* out) {
    JavaRandom r(seed);
    for (int i = 0; i < n; ++i) out[i] = r.nextDouble();
}
SB_API int orc_index_to_level(int id) { return indexToLevel(id); }

/*
 * Categorical binning at load, restated (f1):
 *   CommonUtils.flattenCatValGrp (util/CommonUtils.java:1767-1775): isNotBlank(group) ? Splitter.on("@^").split(group) : [];
 *   DTWorker.init (dt/DTWorker.java:362-376): tmpMap.put(cval, i) for every member of binCategory.get(i), in order;
 *   DTWorker.load (dt/DTWorker.java:1152-1169): "" -> binCategory.size(); not in the map -> binCategory.size().
 * The map is not built: the groups are scanned in order and the LAST one holding the value wins, which is what a
 * sequence of HashMap.put calls leaves behind.
 */
static bool javaIsBlank(const std::string& s) {
    for (unsigned char c : s)
        if (!(c == ' ' || (c >= 0x09 && c <= 0x0d) || (c >= 0x1c && c <= 0x1f))) return false;
    return true;
}
static std::vector<std::string> flattenCatValGrp(const std::string& categoricalValGrp) {
    std::vector<std::string> catVals;
    if (!javaIsBlank(categoricalValGrp)) {
        size_t start = 0;
        while (true) {
            const size_t hit = categoricalValGrp.find("@^", start);
            if (hit == std::string::npos) {
                catVals.push_back(categoricalValGrp.substr(start));
                break;
            }
            catVals.push_back(categoricalValGrp.substr(start, hit - start));
            start = hit + 2;
        }
    }
    return catVals;
}
/* pieces of one group, '\n'-joined into out (cap bytes); returns the number of pieces */
SB_API int orc_flatten_cat_val_grp(const char* group, char* out, int cap) {
    const std::vector<std::string> v = flattenCatValGrp(group);
    std::string joined;
    for (size_t i = 0; i < v.size(); ++i) joined += (i ? "\n" : "") + v[i];
    if ((int)joined.size() + 1 > cap) return -1;
    std::memcpy(out, joined.c_str(), joined.size() + 1);
    return (int)v.size();
}
SB_API int orc_category_index(const char* const* bin_category, int n_categories, const char* input) {
    short shortValue = (short)n_categories;
    if (input == nullptr || input[0] == '\0') return shortValue;
    int categoricalIndex = -1;  // null
    for (int i = 0; i < n_categories; ++i)
        for (const std::string& cval : flattenCatValGrp(bin_category[i]))
            if (cval == input) categoricalIndex = i;
    shortValue = categoricalIndex < 0 ? (short)-1 : (short)categoricalIndex;
    if (shortValue == -1) shortValue = (short)n_categories;
    return shortValue;
}

}  // extern "C"
