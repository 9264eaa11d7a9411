/*
 * shifu_b200_jni.c -- JNI side of ml.shifu.shifu.core.dtrain.dt.ShifuB200 (jni/ShifuB200.java): argument marshalling
 * only, every call goes straight into the C ABI of include/shifu_b200.h.
 *
 * NOT COMPILED in this repository's build image: there is no JDK (no jni.h).  Where a JDK exists:
 *   gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -Iinclude \
 *       -o libshifu_b200_jni.so jni/shifu_b200_jni.c -Lshifu_b200 -lshifu_b200
 * The part that can be checked without a JVM -- the message layout and its unpacking -- is tests/wire_check.c.
 *
 * Errors: every failing sb_* call becomes a RuntimeException carrying sb_last_error(), the way DTWorker reports its own
 * failures (dt/DTWorker.java:359, 878: unchecked exceptions that fail the Guagua task).
 */
#include <jni.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "shifu_b200.h"

#define JNI_FN(name) Java_ml_shifu_shifu_core_dtrain_dt_ShifuB200_##name

static void throw_msg(JNIEnv* env, const char* msg) {
    jclass cls = (*env)->FindClass(env, "java/lang/RuntimeException");
    if (cls) (*env)->ThrowNew(env, cls, msg);
}
static void throw_last(JNIEnv* env) { throw_msg(env, sb_last_error()); }

/* address of a direct buffer, or NULL for a null reference; *bytes = its capacity */
static void* direct(JNIEnv* env, jobject buf, jlong* bytes) {
    if (!buf) {
        if (bytes) *bytes = 0;
        return NULL;
    }
    if (bytes) *bytes = (*env)->GetDirectBufferCapacity(env, buf);
    return (*env)->GetDirectBufferAddress(env, buf);
}

/* Unpacks a flattened DTMasterParams; on failure throws and returns NULL.  The view points into the buffer. */
static sb_msg* unpack_message(JNIEnv* env, jobject message) {
    jlong bytes = 0;
    void* p = direct(env, message, &bytes);
    sb_msg* m = NULL;
    if (!p) {
        throw_msg(env, "message must be a direct ByteBuffer");
        return NULL;
    }
    if (sb_msg_unpack(p, (int64_t)bytes, &m) != SB_OK) {
        throw_last(env);
        return NULL;
    }
    return m;
}

JNIEXPORT jlong JNICALL JNI_FN(create)(JNIEnv* env, jclass cls, jintArray nBins, jbyteArray isCategorical,
                                       jintArray columnNum, jdoubleArray binBoundary, jint impurity, jint numClasses,
                                       jint loss, jboolean isGBDT, jint treeNum, jint weightsPerRow, jint maxDepth,
                                       jint minInstancesPerNode, jdouble minInfoGain, jdouble learningRate, jint device) {
    (void)cls;
    sb_config cfg;
    sb_handle* h = NULL;
    memset(&cfg, 0, sizeof(cfg));
    cfg.n_features = (int32_t)(*env)->GetArrayLength(env, nBins);
    jint* bins = (*env)->GetIntArrayElements(env, nBins, NULL);
    jbyte* cat = isCategorical ? (*env)->GetByteArrayElements(env, isCategorical, NULL) : NULL;
    jint* cols = columnNum ? (*env)->GetIntArrayElements(env, columnNum, NULL) : NULL;
    jdouble* bb = binBoundary ? (*env)->GetDoubleArrayElements(env, binBoundary, NULL) : NULL;
    cfg.n_bins = (const int32_t*)bins;
    cfg.is_categorical = (const uint8_t*)cat;
    cfg.column_num = (const int32_t*)cols;
    cfg.bin_boundary = (const double*)bb;
    cfg.impurity = impurity;
    cfg.num_classes = numClasses;
    cfg.loss = loss;
    cfg.is_gbdt = isGBDT ? 1 : 0;
    cfg.tree_num = treeNum;
    cfg.weights_per_row = weightsPerRow;
    cfg.max_depth = maxDepth;
    cfg.min_instances_per_node = minInstancesPerNode;
    cfg.min_info_gain = minInfoGain;
    cfg.learning_rate = learningRate;
    cfg.device = device;
    const int rc = sb_create(&cfg, &h); /* copies what it keeps: the arrays can be released right away */
    (*env)->ReleaseIntArrayElements(env, nBins, bins, JNI_ABORT);
    if (cat) (*env)->ReleaseByteArrayElements(env, isCategorical, cat, JNI_ABORT);
    if (cols) (*env)->ReleaseIntArrayElements(env, columnNum, cols, JNI_ABORT);
    if (bb) (*env)->ReleaseDoubleArrayElements(env, binBoundary, bb, JNI_ABORT);
    if (rc != SB_OK) {
        throw_last(env);
        return 0;
    }
    return (jlong)(intptr_t)h;
}

JNIEXPORT void JNICALL JNI_FN(destroy)(JNIEnv* env, jclass cls, jlong handle) {
    (void)env;
    (void)cls;
    sb_destroy((sb_handle*)(intptr_t)handle);
}

JNIEXPORT void JNICALL JNI_FN(loadRows)(JNIEnv* env, jclass cls, jlong handle, jboolean validation, jlong nRows,
                                        jobjectArray binColumns, jobject label, jobject significance, jobject weights) {
    (void)cls;
    const jsize n_cols = (*env)->GetArrayLength(env, binColumns);
    const void** cols = (const void**)calloc((size_t)(n_cols > 0 ? n_cols : 1), sizeof(void*));
    if (!cols) {
        throw_msg(env, "out of memory");
        return;
    }
    for (jsize f = 0; f < n_cols; ++f) {
        jobject b = (*env)->GetObjectArrayElement(env, binColumns, f);
        cols[f] = direct(env, b, NULL);
        (*env)->DeleteLocalRef(env, b);
        if (!cols[f]) {
            free((void*)cols);
            throw_msg(env, "every bin column must be a direct ByteBuffer");
            return;
        }
    }
    const int rc = sb_load_rows((sb_handle*)(intptr_t)handle, validation ? 1 : 0, (int64_t)nRows, cols,
                                (const float*)direct(env, label, NULL), (const float*)direct(env, significance, NULL),
                                (const float*)direct(env, weights, NULL));
    free((void*)cols);
    if (rc != SB_OK) throw_last(env);
}

JNIEXPORT jlong JNICALL JNI_FN(statsLen)(JNIEnv* env, jclass cls, jlong handle, jobject message) {
    (void)cls;
    sb_msg* m = unpack_message(env, message);
    if (!m) return 0;
    const int64_t n = sb_stats_len((sb_handle*)(intptr_t)handle, sb_msg_iter_in(m));
    sb_msg_free(m);
    if (n < 0) throw_last(env);
    return (jlong)n;
}

JNIEXPORT jdoubleArray JNICALL JNI_FN(compute)(JNIEnv* env, jclass cls, jlong handle, jobject message, jobject statsOut) {
    (void)cls;
    sb_msg* m = unpack_message(env, message);
    if (!m) return NULL;
    sb_iter_out out;
    jlong cap = 0;
    memset(&out, 0, sizeof(out));
    out.stats = (double*)direct(env, statsOut, &cap);
    out.stats_len = (int64_t)(cap / 8);
    const int rc = sb_compute((sb_handle*)(intptr_t)handle, sb_msg_iter_in(m), &out);
    sb_msg_free(m);
    if (rc != SB_OK) {
        throw_last(env);
        return NULL;
    }
    const jdouble s[4] = {out.train_count, out.validation_count, out.train_error, out.validation_error};
    jdoubleArray r = (*env)->NewDoubleArray(env, 4);
    if (r) (*env)->SetDoubleArrayRegion(env, r, 0, 4, s);
    return r;
}

JNIEXPORT void JNICALL JNI_FN(bestSplits)(JNIEnv* env, jclass cls, jlong handle, jobject gainsOut, jobject catBitsOut) {
    (void)cls;
    sb_gain* gains = (sb_gain*)direct(env, gainsOut, NULL);
    if (!gains) {
        throw_msg(env, "gainsOut must be a direct ByteBuffer");
        return;
    }
    if (sb_best_splits((sb_handle*)(intptr_t)handle, gains, (uint8_t*)direct(env, catBitsOut, NULL)) != SB_OK) throw_last(env);
}
