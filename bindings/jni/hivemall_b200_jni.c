/* hivemall_b200_jni.c -- JNI glue between hivemall.gpu.HivemallB200 (HivemallB200.java) and the C-ABI of
 * include/hivemall_b200.h.  Build where a JDK exists:
 *
 *   gcc -shared -fPIC -I$JAVA_HOME/include -I$JAVA_HOME/include/linux -I../../include hivemall_b200_jni.c \
 *       -L../../hivemall_b200 -lhivemall_b200 -o libhivemall_b200_jni.so
 *
 * This repository's image has no JDK: tests/test_jni_glue.py compiles this file against
 * compile_check/jni.h, a stand-in that declares only the JNI types and calls used here (NOT binary compatible
 * with a JVM), to keep the glue in step with the library header, and checks that every native method of the
 * Java class has its function here.
 */
#include <jni.h>
#include <stdint.h>
#include "hivemall_b200.h"

#define HANDLE(h) ((hm_udtf *)(intptr_t)(h))
#define ADDR(type, buf) ((type *)((buf) ? (*env)->GetDirectBufferAddress(env, (buf)) : 0))

/* HM_ERR_INVALID_ARGUMENT is the UDFArgumentException role, everything else HiveException (hivemall_b200.h). */
static void throw_for(JNIEnv *env, hm_status s)
{
    const char *cls = s == HM_ERR_INVALID_ARGUMENT ? "org/apache/hadoop/hive/ql/exec/UDFArgumentException"
                                                   : "org/apache/hadoop/hive/ql/metadata/HiveException";
    jclass c = (*env)->FindClass(env, cls);
    if (c) (*env)->ThrowNew(env, c, hm_last_error());
}

JNIEXPORT jlong JNICALL Java_hivemall_gpu_HivemallB200_create(JNIEnv *env, jclass cls, jstring fn, jstring opts, jint dev)
{
    (void)cls;
    const char *f = (*env)->GetStringUTFChars(env, fn, 0);
    const char *o = (*env)->GetStringUTFChars(env, opts, 0);
    hm_udtf *h = 0;
    hm_status s = (f && o) ? hm_udtf_create(f, o, dev, &h) : HM_ERR_OOM;
    if (f) (*env)->ReleaseStringUTFChars(env, fn, f);
    if (o) (*env)->ReleaseStringUTFChars(env, opts, o);
    if (s != HM_OK) { throw_for(env, s); return 0; }
    return (jlong)(intptr_t)h;
}

JNIEXPORT jobject JNICALL Java_hivemall_gpu_HivemallB200_hostAlloc(JNIEnv *env, jclass cls, jlong bytes)
{
    (void)cls;
    void *p = 0;
    hm_status s = hm_host_alloc(&p, bytes);
    if (s != HM_OK) { throw_for(env, s); return 0; }
    return (*env)->NewDirectByteBuffer(env, p, bytes);
}

JNIEXPORT void JNICALL Java_hivemall_gpu_HivemallB200_hostFree(JNIEnv *env, jclass cls, jobject buf)
{
    (void)cls;
    hm_status s = hm_host_free(ADDR(void, buf));
    if (s != HM_OK) throw_for(env, s);
}

JNIEXPORT void JNICALL Java_hivemall_gpu_HivemallB200_process(JNIEnv *env, jclass cls, jlong h, jlong n, jobject rowptr,
                                                              jobject feature, jobject value, jobject label)
{
    (void)cls;
    hm_status s = hm_udtf_process(HANDLE(h), n, ADDR(const int64_t, rowptr), ADDR(const int32_t, feature),
                                  ADDR(const float, value) /* NULL: index-only rows */, ADDR(const float, label));
    if (s != HM_OK) throw_for(env, s);
}

JNIEXPORT void JNICALL Java_hivemall_gpu_HivemallB200_processSplit(JNIEnv *env, jclass cls, jlong h, jlong n, jobject q_rowptr,
                                                                   jobject q_feature, jobject q_value, jobject c_rowptr,
                                                                   jobject c_feature, jobject label)
{
    (void)cls;
    hm_status s = hm_udtf_process_split(HANDLE(h), n, ADDR(const int64_t, q_rowptr), ADDR(const int32_t, q_feature),
                                        ADDR(const float, q_value), ADDR(const int64_t, c_rowptr),
                                        ADDR(const int32_t, c_feature), ADDR(const float, label));
    if (s != HM_OK) throw_for(env, s);
}

JNIEXPORT void JNICALL Java_hivemall_gpu_HivemallB200_processColumns(JNIEnv *env, jclass cls, jlong h, jlong n, jint n_q,
                                                                     jobject q_feature, jobject q_value, jint n_c,
                                                                     jobject c_feature, jobject label)
{
    (void)cls;
    hm_status s = hm_udtf_process_columns(HANDLE(h), n, n_q, ADDR(const int32_t, q_feature), ADDR(const float, q_value), n_c,
                                          ADDR(const int32_t, c_feature), ADDR(const float, label));
    if (s != HM_OK) throw_for(env, s);
}

JNIEXPORT void JNICALL Java_hivemall_gpu_HivemallB200_processStrings(JNIEnv *env, jclass cls, jlong h, jlong n, jobject rowptr,
                                                                     jobject bytes, jobject offsets, jobject label)
{
    (void)cls;
    hm_status s = hm_udtf_process_strings(HANDLE(h), n, ADDR(const int64_t, rowptr), ADDR(const uint8_t, bytes),
                                          ADDR(const int64_t, offsets), ADDR(const float, label));
    if (s != HM_OK) throw_for(env, s);
}

JNIEXPORT jlong JNICALL Java_hivemall_gpu_HivemallB200_close(JNIEnv *env, jclass cls, jlong h)
{
    (void)cls;
    int64_t rows = 0;
    hm_status s = hm_udtf_close(HANDLE(h), &rows);
    if (s != HM_OK) { throw_for(env, s); return 0; }
    return (jlong)rows;
}

JNIEXPORT jlong JNICALL Java_hivemall_gpu_HivemallB200_forward(JNIEnv *env, jclass cls, jlong h, jobject feature, jobject weight,
                                                               jobject vif, jlong cap)
{
    (void)cls;
    int64_t written = 0;
    hm_status s = hm_udtf_forward(HANDLE(h), ADDR(int32_t, feature), ADDR(float, weight), ADDR(float, vif), cap, &written);
    if (s != HM_OK) { throw_for(env, s); return 0; }
    return (jlong)written;
}

static jlong table_io(JNIEnv *env, jlong h, jstring path, jint format, jchar delimiter, int save)
{
    const char *p = (*env)->GetStringUTFChars(env, path, 0);
    int64_t rows = 0;
    hm_status s = !p ? HM_ERR_OOM
                     : (save ? hm_udtf_save_model_table(HANDLE(h), p, format, (char)delimiter, &rows)
                             : hm_udtf_load_model_table(HANDLE(h), p, format, (char)delimiter, &rows));
    if (p) (*env)->ReleaseStringUTFChars(env, path, p);
    if (s != HM_OK) { throw_for(env, s); return 0; }
    return (jlong)rows;
}

JNIEXPORT jlong JNICALL Java_hivemall_gpu_HivemallB200_saveModelTable(JNIEnv *env, jclass cls, jlong h, jstring path, jint format,
                                                                      jchar delimiter)
{
    (void)cls;
    return table_io(env, h, path, format, delimiter, 1);
}

JNIEXPORT jlong JNICALL Java_hivemall_gpu_HivemallB200_loadModelTable(JNIEnv *env, jclass cls, jlong h, jstring path, jint format,
                                                                      jchar delimiter)
{
    (void)cls;
    return table_io(env, h, path, format, delimiter, 0);
}

JNIEXPORT void JNICALL Java_hivemall_gpu_HivemallB200_destroy(JNIEnv *env, jclass cls, jlong h)
{
    (void)cls;
    hm_status s = hm_udtf_destroy(HANDLE(h));
    if (s != HM_OK) throw_for(env, s);
}

JNIEXPORT void JNICALL Java_hivemall_gpu_HivemallB200_mhash(JNIEnv *env, jclass cls, jobject bytes, jobject offsets, jlong n,
                                                            jint seed, jint num_features, jobject out, jint device)
{
    (void)cls;
    hm_status s = hm_mhash(ADDR(const uint8_t, bytes), ADDR(const int64_t, offsets), n, (uint32_t)seed, num_features,
                           ADDR(int32_t, out), device);
    if (s != HM_OK) throw_for(env, s);
}

JNIEXPORT void JNICALL Java_hivemall_gpu_HivemallB200_featureHashing(JNIEnv *env, jclass cls, jobject bytes, jobject offsets,
                                                                     jlong n, jint seed, jint num_features, jobject out_index,
                                                                     jobject out_value, jint device)
{
    (void)cls;
    hm_status s = hm_feature_hashing(ADDR(const uint8_t, bytes), ADDR(const int64_t, offsets), n, (uint32_t)seed, num_features,
                                     ADDR(int32_t, out_index), ADDR(float, out_value), device);
    if (s != HM_OK) throw_for(env, s);
}
