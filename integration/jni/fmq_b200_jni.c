/* UNCOMPILED SOURCE — jni.h does not exist in the build image.  The whole JNI layer is pointer
 * unwrapping around the C ABI of include/fmq_b200.h; nothing here computes anything. */
#include <jni.h>
#include <string.h>
#include "fmq_b200.h"

#define ADDR(buf) ((*env)->GetDirectBufferAddress(env, (buf)))

JNIEXPORT jlong JNICALL Java_me_zjy_GpuWindowAggOperator_nCreate(JNIEnv* env, jclass c, jobject cfg) {
    fmq_handle* h = NULL;
    fmq_config conf;
    memcpy(&conf, ADDR(cfg), sizeof(conf));                      /* the Java side serialises the POD */
    return fmq_create(&conf, &h) == FMQ_OK ? (jlong)(intptr_t)h : 0;
}
JNIEXPORT jint JNICALL Java_me_zjy_GpuWindowAggOperator_nPushBatch(JNIEnv* env, jclass c, jlong h, jobject ts, jobject key, jobject val,
                                                                   jint n, jlong wm) {
    return fmq_push_batch((fmq_handle*)(intptr_t)h, (const int64_t*)ADDR(ts), (const uint64_t*)ADDR(key), ADDR(val), (size_t)n, wm, FMQ_MEM_HOST);
}
JNIEXPORT jint JNICALL Java_me_zjy_GpuWindowAggOperator_nFlush(JNIEnv* env, jclass c, jlong h, jlong wm) {
    return fmq_flush((fmq_handle*)(intptr_t)h, wm);
}
JNIEXPORT jint JNICALL Java_me_zjy_GpuWindowAggOperator_nPollWindowAlerts(JNIEnv* env, jclass c, jlong h, jobject key, jobject wend, jobject value,
                                                                          jint cap, jlongArray outLeft) {
    fmq_walert_cols cols = {(uint64_t*)ADDR(key), (int64_t*)ADDR(wend), (double*)ADDR(value)};
    size_t n_out = 0, n_left = 0;
    int32_t st = fmq_poll_window_alerts((fmq_handle*)(intptr_t)h, &cols, (size_t)cap, FMQ_MEM_HOST, &n_out, &n_left);
    jlong two[2] = {(jlong)n_out, (jlong)n_left};
    (*env)->SetLongArrayRegion(env, outLeft, 0, 2, two);
    return st;
}
JNIEXPORT jlong JNICALL Java_me_zjy_GpuWindowAggOperator_nLateDropped(JNIEnv* env, jclass c, jlong h) {
    fmq_counters ctr;
    return fmq_get_counters((fmq_handle*)(intptr_t)h, &ctr) == FMQ_OK ? (jlong)ctr.late_dropped : -1;
}
JNIEXPORT jstring JNICALL Java_me_zjy_GpuWindowAggOperator_nLastError(JNIEnv* env, jclass c, jlong h) {
    return (*env)->NewStringUTF(env, fmq_last_error((fmq_handle*)(intptr_t)h));
}
JNIEXPORT void JNICALL Java_me_zjy_GpuWindowAggOperator_nDestroy(JNIEnv* env, jclass c, jlong h) {
    fmq_destroy((fmq_handle*)(intptr_t)h);
}
