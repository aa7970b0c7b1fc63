/* jni.h -- COMPILE-CHECK STAND-IN, not the JDK's header.
 *
 * Declares only the JNI types and the JNIEnv calls that bindings/jni/hivemall_b200_jni.c uses, so that the glue
 * can be type-checked against include/hivemall_b200.h in an image without a JDK (tests/test_jni_glue.py).  The
 * function table below is NOT laid out like a JVM's: an object built against this file must never be loaded
 * into a JVM.  Build the real library with the JDK's own jni.h (see the comment in hivemall_b200_jni.c).
 */
#ifndef HM_JNI_COMPILE_CHECK_H
#define HM_JNI_COMPILE_CHECK_H
#include <stdint.h>

#define JNIEXPORT __attribute__((visibility("default")))
#define JNICALL

typedef int32_t jint;
typedef int64_t jlong;
typedef uint16_t jchar;
typedef void *jobject;
typedef jobject jclass;
typedef jobject jstring;
typedef uint8_t jboolean;

struct JNINativeInterface_;
typedef const struct JNINativeInterface_ *JNIEnv;

struct JNINativeInterface_ {
    jclass (*FindClass)(JNIEnv *env, const char *name);
    jint (*ThrowNew)(JNIEnv *env, jclass clazz, const char *msg);
    const char *(*GetStringUTFChars)(JNIEnv *env, jstring str, jboolean *is_copy);
    void (*ReleaseStringUTFChars)(JNIEnv *env, jstring str, const char *chars);
    jobject (*NewDirectByteBuffer)(JNIEnv *env, void *address, jlong capacity);
    void *(*GetDirectBufferAddress)(JNIEnv *env, jobject buf);
};
#endif
