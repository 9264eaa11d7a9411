/*
 * ShifuB200 -- the native methods DTWorker calls instead of its own loops (package ml.shifu.shifu.core.dtrain.dt).
 * NOT COMPILED in this repository's build image: there is no JDK here.  The C side is jni/shifu_b200_jni.c; the
 * message layout both sides share is documented in include/shifu_b200.h ("Wire format of sb_iter_in") and pinned
 * without a JVM by tests/wire_check.c.
 *
 * Which reference code each method stands in for:
 *   create      DTWorker.init       dt/DTWorker.java:349-550   (column layout and train params, flattened)
 *   loadRows    DTWorker.load       dt/DTWorker.java:1099-1297 (already-binned rows, one call from postLoad)
 *   compute     DTWorker.doCompute  dt/DTWorker.java:558-891
 *   bestSplits  per-node body of DTMaster.doCompute dt/DTMaster.java:316-340 (optional: the master may keep its own)
 *   destroy     the completion callback that shuts the worker down, dt/DTWorker.java:404-414
 */
package ml.shifu.shifu.core.dtrain.dt;

import java.nio.ByteBuffer;

final class ShifuB200 {

    static {
        System.loadLibrary("shifu_b200_jni"); // links libshifu_b200.so
    }

    private ShifuB200() {
    }

    /** Impurity / loss codes of include/shifu_b200.h. */
    static final int IMPURITY_VARIANCE = 0, IMPURITY_FRIEDMANMSE = 1, IMPURITY_ENTROPY = 2, IMPURITY_GINI = 3;
    static final int LOSS_SQUARED = 0, LOSS_HALFGRADSQUARED = 1, LOSS_ABSOLUTE = 2, LOSS_LOG = 3;

    /**
     * @param nBins
     *            per feature slot: binBoundary.size() (numeric) or binCategory.size() + 1 (categorical)
     * @param binBoundary
     *            concatenated binBoundary lists of the numeric slots, slot order
     * @return an opaque handle; a failure is thrown as RuntimeException with sb_last_error() as message
     */
    static native long create(int[] nBins, byte[] isCategorical, int[] columnNum, double[] binBoundary, int impurity,
            int numClasses, int loss, boolean isGBDT, int treeNum, int weightsPerRow, int maxDepth,
            int minInstancesPerNode, double minInfoGain, double learningRate, int device);

    static native void destroy(long handle);

    /**
     * All buffers are direct and in native byte order.  binColumns[slot] holds nRows bin indices, one byte each when
     * nBins[slot] <= 256, else two (short).  weights is tree-major [weightsPerRow][nRows]; significance / weights may
     * be null (all 1f).
     */
    static native void loadRows(long handle, boolean validation, long nRows, ByteBuffer[] binColumns, ByteBuffer label,
            ByteBuffer significance, ByteBuffer weights);

    /**
     * @param message
     *            DTMasterParams flattened by DTMessageFlattener.flatten (direct buffer, position 0, limit = size)
     * @param statsOut
     *            direct buffer of at least 8 * statsLen(message) bytes; receives the double[] statistics of every todo
     *            node, feature list order, bin-major
     * @return {trainCount, validationCount, trainError, validationError}
     */
    static native double[] compute(long handle, ByteBuffer message, ByteBuffer statsOut);

    /** Number of doubles compute() will write for this message. */
    static native long statsLen(long handle, ByteBuffer message);

    /**
     * Best split of every todo node of the last compute(), from the merged statistics on the device.
     *
     * @param gainsOut
     *            direct buffer of n_todo * 104 bytes (sb_gain records)
     * @param catBitsOut
     *            direct buffer of n_todo * 4096 bytes, or null when no feature is categorical
     */
    static native void bestSplits(long handle, ByteBuffer gainsOut, ByteBuffer catBitsOut);
}
