/*
 * DTMessageFlattener -- DTMasterParams -> the wire format of sb_iter_in, and the statistics buffer -> NodeStats
 * (package ml.shifu.shifu.core.dtrain.dt).  NOT COMPILED in this repository's build image: there is no JDK here.
 * The layout is documented in include/shifu_b200.h ("Wire format of sb_iter_in"); tests/wire_check.c writes the
 * same bytes from plain C and feeds them to sb_msg_unpack and sb_compute, which pins the format without a JVM.
 *
 * With this class the body of DTWorker.doCompute (dt/DTWorker.java:558-891) becomes
 *
 *   ByteBuffer msg = DTMessageFlattener.flatten(context.getLastMasterResult(), columnConfigList, inputIndexMap,
 *           newWeights, dropout);                       // newWeights / dropout: the draws of :643-668 / :634-638, or null
 *   ByteBuffer stats = statsBuffer(ShifuB200.statsLen(handle, msg));
 *   double[] s = ShifuB200.compute(handle, msg, stats);
 *   return new DTWorkerParams(s[0], s[1], s[2], s[3],
 *           DTMessageFlattener.unflatten(stats, context.getLastMasterResult().getTodoNodes(), columnConfigList,
 *                   getAllValidFeatures(), impurity.getStatsSize()));
 */
package ml.shifu.shifu.core.dtrain.dt;

import java.io.ByteArrayOutputStream;
import java.io.DataOutputStream;
import java.io.IOException;
import java.nio.ByteBuffer;
import java.nio.ByteOrder;
import java.util.ArrayList;
import java.util.HashMap;
import java.util.List;
import java.util.Map;
import java.util.TreeMap;

import ml.shifu.shifu.container.obj.ColumnConfig;
import ml.shifu.shifu.core.dtrain.dt.DTWorkerParams.NodeStats;

final class DTMessageFlattener {

    private static final int MAGIC = 0x314D4253; // "SBM1"
    private static final int NODE_BYTES = 64; // sizeof(sb_node)

    private DTMessageFlattener() {
    }

    /** One flattened tree: sb_node records in pre-order plus the byte pool their masks live in. */
    private static final class FlatTree {
        final int treeId;
        final List<byte[]> nodes = new ArrayList<byte[]>();
        final ByteArrayOutputStream pool = new ByteArrayOutputStream();

        FlatTree(int treeId) {
            this.treeId = treeId;
        }
    }

    private static int pad8(int n) {
        return (n + 7) & ~7;
    }

    /** SimpleBitSet keeps its words private; its own write() gives them back: int length, then the bytes. */
    private static byte[] wordsOf(SimpleBitSet<Short> set) throws IOException {
        ByteArrayOutputStream bos = new ByteArrayOutputStream();
        set.write(new DataOutputStream(bos));
        byte[] all = bos.toByteArray();
        byte[] words = new byte[all.length - 4];
        System.arraycopy(all, 4, words, 0, words.length);
        return words;
    }

    /**
     * left_mask of a split node: bit b set when DTWorker.predictNodeIndex (dt/DTWorker.java:1052-1088) sends a row whose
     * bin index is b to the LEFT child.
     */
    private static byte[] leftMask(Split split, ColumnConfig config) {
        boolean numerical = config.isNumerical();
        int bins = numerical ? config.getBinBoundary().size() : config.getBinCategory().size() + 1;
        byte[] mask = new byte[bins / 8 + 1];
        for(int b = 0; b < bins; b++) {
            boolean left;
            if(numerical) {
                left = config.getBinBoundary().get(b) < split.getThreshold();
            } else {
                int nCats = config.getBinCategory().size();
                short indexValue = (short) ((b >= 0 && b < nCats) ? b : nCats);
                boolean contains = split.getLeftOrRightCategories().contains(indexValue);
                left = split.isLeft() ? contains : !contains;
            }
            if(left) {
                mask[b / 8] |= (byte) (1 << (b % 8));
            }
        }
        return mask;
    }

    /** Appends `node` and its subtree in pre-order; returns the index of `node` in the tree's node list. */
    @SuppressWarnings("unchecked")
    private static int emit(Node node, FlatTree tree, List<ColumnConfig> columnConfigList,
            Map<Integer, Integer> inputIndexMap) throws IOException {
        int me = tree.nodes.size();
        ByteBuffer rec = ByteBuffer.allocate(NODE_BYTES).order(ByteOrder.LITTLE_ENDIAN);
        tree.nodes.add(rec.array());
        Split split = node.getSplit();
        int feature = -1, splitType = Split.CONTINUOUS, maskOff = 0, maskLen = 0, catOff = 0, catLen = 0;
        double threshold = 0d;
        boolean isLeft = false;
        if(split != null) {
            ColumnConfig config = columnConfigList.get(split.getColumnNum());
            feature = inputIndexMap.get(split.getColumnNum());
            splitType = split.getFeatureType();
            threshold = split.getThreshold();
            isLeft = split.isLeft();
            byte[] mask = leftMask(split, config);
            maskOff = tree.pool.size();
            maskLen = mask.length;
            tree.pool.write(mask);
            if(split.getFeatureType() == Split.CATEGORICAL && split.getLeftOrRightCategories() != null) {
                byte[] words = wordsOf((SimpleBitSet<Short>) split.getLeftOrRightCategories());
                catOff = tree.pool.size();
                catLen = words.length;
                tree.pool.write(words);
            }
        }
        int left = node.getLeft() == null ? -1 : emit(node.getLeft(), tree, columnConfigList, inputIndexMap);
        int right = node.getRight() == null ? -1 : emit(node.getRight(), tree, columnConfigList, inputIndexMap);
        rec.putInt(node.getId()).putInt(left).putInt(right).putInt(feature).putInt(splitType).putInt(maskOff)
                .putInt(maskLen).putInt(node.getPredict() == null ? 0 : 1);
        rec.putDouble(node.getPredict() == null ? 0d : node.getPredict().getPredict()).putDouble(threshold);
        rec.putInt(isLeft ? 1 : 0).putInt(catOff).putInt(catLen).putInt(0);
        return me;
    }

    /**
     * @param newWeights
     *            replacement of subsampleWeights[currTreeIndex % len] for every training row, or null
     * @param dropout
     *            per training row 1 = skip the finished tree's contribution, or null
     * @return a direct little-endian buffer, position 0, limit = message size
     */
    static ByteBuffer flatten(DTMasterParams params, List<ColumnConfig> columnConfigList,
            Map<Integer, Integer> inputIndexMap, float[] newWeights, byte[] dropout) throws IOException {
        List<FlatTree> trees = new ArrayList<FlatTree>();
        for(TreeNode treeNode: params.getTrees()) {
            FlatTree flat = new FlatTree(treeNode.getTreeId());
            // a placeholder tree (root id INVALID_INDEX, dt/DTMaster.java:500-504) travels with zero nodes
            if(treeNode.getNode() != null && treeNode.getNode().getId() != Node.INVALID_INDEX) {
                emit(treeNode.getNode(), flat, columnConfigList, inputIndexMap);
            }
            trees.add(flat);
        }
        // todo nodes in ascending key order: the order DTWorker.initTodoNodeStats and DTMaster.doCompute see them in
        Map<Integer, TreeNode> todo = new TreeMap<Integer, TreeNode>(params.getTodoNodes());
        long nRows = newWeights != null ? newWeights.length : (dropout != null ? dropout.length : 0);

        int size = 40;
        for(FlatTree t: trees) {
            size += 16 + t.nodes.size() * NODE_BYTES + pad8(t.pool.size());
        }
        for(TreeNode t: todo.values()) {
            size += 16 + pad8(4 * (t.getFeatures() == null ? 0 : t.getFeatures().size()));
        }
        if(newWeights != null) {
            size += pad8(4 * newWeights.length);
        }
        if(dropout != null) {
            size += pad8(dropout.length);
        }

        ByteBuffer out = ByteBuffer.allocateDirect(size).order(ByteOrder.LITTLE_ENDIAN);
        out.putInt(MAGIC).putInt(size).putInt(trees.size()).putInt(todo.size());
        out.putInt(params.isSwitchToNextTree() ? 1 : 0).putInt(params.isFirstTree() ? 1 : 0);
        out.putInt(params.isContinuousRunningStart() ? 1 : 0);
        out.putInt((newWeights != null ? 1 : 0) | (dropout != null ? 2 : 0)).putLong(nRows);
        for(FlatTree t: trees) {
            byte[] pool = t.pool.toByteArray();
            out.putInt(t.treeId).putInt(t.nodes.size()).putInt(pool.length).putInt(0);
            for(byte[] rec: t.nodes) {
                out.put(rec);
            }
            out.put(pool);
            out.position(out.position() + pad8(pool.length) - pool.length);
        }
        for(Map.Entry<Integer, TreeNode> e: todo.entrySet()) {
            TreeNode t = e.getValue();
            List<Integer> features = t.getFeatures() == null ? new ArrayList<Integer>() : t.getFeatures();
            out.putInt(e.getKey()).putInt(t.getTreeId()).putInt(t.getNode().getId()).putInt(features.size());
            for(Integer columnNum: features) {
                out.putInt(inputIndexMap.get(columnNum)); // TreeNode.features holds column numbers, the device takes slots
            }
            out.position(out.position() + pad8(4 * features.size()) - 4 * features.size());
        }
        if(newWeights != null) {
            for(float w: newWeights) {
                out.putFloat(w);
            }
            out.position(out.position() + pad8(4 * newWeights.length) - 4 * newWeights.length);
        }
        if(dropout != null) {
            out.put(dropout);
            out.position(out.position() + pad8(dropout.length) - dropout.length);
        }
        out.flip();
        return out;
    }

    /**
     * The statistics buffer of ShifuB200.compute back into the map DTWorkerParams carries (dt/DTWorker.java:903-929
     * initTodoNodeStats gives the shapes): todo nodes in ascending key order; inside a node the features in list order
     * (the node's own list, or every valid feature when it is empty); each feature block is the reference's
     * double[bins * statsSize].
     */
    static Map<Integer, NodeStats> unflatten(ByteBuffer stats, Map<Integer, TreeNode> todoNodes,
            List<ColumnConfig> columnConfigList, List<Integer> allValidFeatures, int statsSize) {
        stats.order(ByteOrder.LITTLE_ENDIAN);
        int at = 0;
        Map<Integer, NodeStats> statistics = new HashMap<Integer, NodeStats>(todoNodes.size(), 1f);
        for(Map.Entry<Integer, TreeNode> entry: new TreeMap<Integer, TreeNode>(todoNodes).entrySet()) {
            List<Integer> features = entry.getValue().getFeatures();
            if(features == null || features.isEmpty()) {
                features = allValidFeatures;
            }
            Map<Integer, double[]> featureStatistics = new HashMap<Integer, double[]>(features.size(), 1f);
            for(Integer columnNum: features) {
                ColumnConfig config = columnConfigList.get(columnNum);
                int bins = config.isNumerical() ? config.getBinBoundary().size() : config.getBinCategory().size() + 1;
                double[] block = new double[bins * statsSize];
                for(int i = 0; i < block.length; i++) {
                    block[i] = stats.getDouble(at);
                    at += 8;
                }
                featureStatistics.put(columnNum, block);
            }
            statistics.put(entry.getKey(),
                    new NodeStats(entry.getValue().getTreeId(), entry.getValue().getNode().getId(), featureStatistics));
        }
        return statistics;
    }
}
