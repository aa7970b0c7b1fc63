/* driver.c -- exercises hivemall_b200_jni.c WITHOUT a JVM: a hand-made JNIEnv (the six calls of
 * compile_check/jni.h) whose strings are C strings and whose ByteBuffers are {address, capacity} structs.
 * Trains a small model through the Java_* functions exactly as HivemallB200.java would and prints the model
 * table; tests/test_jni_glue.py compares it with the same training through the Python binding.
 * Test infrastructure; says nothing about a real JVM's behaviour beyond what the JNI specification promises. */
#include <jni.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef struct { void *addr; jlong cap; } FakeBuffer;
static char g_thrown_class[256], g_thrown_msg[1024];

static jclass f_FindClass(JNIEnv *env, const char *name) { (void)env; return (jclass)name; }
static jint f_ThrowNew(JNIEnv *env, jclass c, const char *msg)
{
    (void)env;
    snprintf(g_thrown_class, sizeof g_thrown_class, "%s", (const char *)c);
    snprintf(g_thrown_msg, sizeof g_thrown_msg, "%s", msg);
    return 0;
}
static const char *f_GetStringUTFChars(JNIEnv *env, jstring s, jboolean *copy) { (void)env; if (copy) *copy = 0; return (const char *)s; }
static void f_ReleaseStringUTFChars(JNIEnv *env, jstring s, const char *c) { (void)env; (void)s; (void)c; }
static jobject f_NewDirectByteBuffer(JNIEnv *env, void *addr, jlong cap)
{
    (void)env;
    FakeBuffer *b = (FakeBuffer *)malloc(sizeof *b);
    b->addr = addr; b->cap = cap;
    return b;
}
static void *f_GetDirectBufferAddress(JNIEnv *env, jobject buf) { (void)env; return ((FakeBuffer *)buf)->addr; }

static const struct JNINativeInterface_ g_table = {f_FindClass, f_ThrowNew, f_GetStringUTFChars, f_ReleaseStringUTFChars,
                                                   f_NewDirectByteBuffer, f_GetDirectBufferAddress};

/* the functions under test (hivemall_b200_jni.c) */
jlong Java_hivemall_gpu_HivemallB200_create(JNIEnv *, jclass, jstring, jstring, jint);
jobject Java_hivemall_gpu_HivemallB200_hostAlloc(JNIEnv *, jclass, jlong);
void Java_hivemall_gpu_HivemallB200_hostFree(JNIEnv *, jclass, jobject);
void Java_hivemall_gpu_HivemallB200_process(JNIEnv *, jclass, jlong, jlong, jobject, jobject, jobject, jobject);
jlong Java_hivemall_gpu_HivemallB200_close(JNIEnv *, jclass, jlong);
jlong Java_hivemall_gpu_HivemallB200_forward(JNIEnv *, jclass, jlong, jobject, jobject, jobject, jlong);
void Java_hivemall_gpu_HivemallB200_destroy(JNIEnv *, jclass, jlong);

int main(int argc, char **argv)
{
    JNIEnv envp = &g_table, *env = &envp;
    const char *options = argc > 1 ? argv[1] : "-opt adagrad -mini_batch 64 -dims 1024";
    /* 1. a bad option string must surface as UDFArgumentException with the library's message */
    jlong bad = Java_hivemall_gpu_HivemallB200_create(env, 0, (jstring) "train_classifier", (jstring) "-frobnicate 1", 0);
    printf("bad_create handle=%lld class=%s msg=%s\n", (long long)bad, g_thrown_class, g_thrown_msg);
    g_thrown_class[0] = 0;
    /* 2. rows from a fixed linear congruential stream: 1000 rows x 6 features in [0, 1024) */
    const int64_t n = 1000, per = 6;
    jlong h = Java_hivemall_gpu_HivemallB200_create(env, 0, (jstring) "train_classifier", (jstring)options, 0);
    if (!h) { printf("create failed: %s\n", g_thrown_msg); return 1; }
    jobject b_rp = Java_hivemall_gpu_HivemallB200_hostAlloc(env, 0, (n + 1) * 8);
    jobject b_f = Java_hivemall_gpu_HivemallB200_hostAlloc(env, 0, n * per * 4);
    jobject b_v = Java_hivemall_gpu_HivemallB200_hostAlloc(env, 0, n * per * 4);
    jobject b_y = Java_hivemall_gpu_HivemallB200_hostAlloc(env, 0, n * 4);
    int64_t *rp = (int64_t *)((FakeBuffer *)b_rp)->addr;
    int32_t *f = (int32_t *)((FakeBuffer *)b_f)->addr;
    float *v = (float *)((FakeBuffer *)b_v)->addr, *y = (float *)((FakeBuffer *)b_y)->addr;
    uint32_t s = 12345u;
    for (int64_t i = 0; i < n; i++) {
        rp[i] = i * per;
        int acc = 0;
        for (int64_t k = 0; k < per; k++) {
            s = s * 1664525u + 1013904223u;
            f[i * per + k] = (int32_t)((s >> 8) & 1023u);
            v[i * per + k] = (float)((s >> 20) & 7u) * 0.25f + 0.25f;
            acc += f[i * per + k] & 1;
        }
        y[i] = acc >= 3 ? 1.0f : 0.0f;
    }
    rp[n] = n * per;
    Java_hivemall_gpu_HivemallB200_process(env, 0, h, n, b_rp, b_f, b_v, b_y);
    if (g_thrown_class[0]) { printf("process threw %s: %s\n", g_thrown_class, g_thrown_msg); return 1; }
    jlong rows = Java_hivemall_gpu_HivemallB200_close(env, 0, h);
    jobject o_f = Java_hivemall_gpu_HivemallB200_hostAlloc(env, 0, rows * 4), o_w = Java_hivemall_gpu_HivemallB200_hostAlloc(env, 0, rows * 4);
    jlong got = Java_hivemall_gpu_HivemallB200_forward(env, 0, h, o_f, o_w, 0, rows);
    printf("rows=%lld written=%lld\n", (long long)rows, (long long)got);
    const int32_t *of = (const int32_t *)((FakeBuffer *)o_f)->addr;
    const float *ow = (const float *)((FakeBuffer *)o_w)->addr;
    for (jlong i = 0; i < got; i++) printf("w %d %.9g\n", of[i], ow[i]);
    Java_hivemall_gpu_HivemallB200_destroy(env, 0, h);
    jobject all[] = {b_rp, b_f, b_v, b_y, o_f, o_w};
    for (int i = 0; i < 6; i++) Java_hivemall_gpu_HivemallB200_hostFree(env, 0, all[i]);
    printf("done thrown=%s\n", g_thrown_class[0] ? g_thrown_class : "none");
    return 0;
}
