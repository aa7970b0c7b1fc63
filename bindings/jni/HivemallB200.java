package hivemall.gpu;

import java.nio.ByteBuffer;

/**
 * JNI face of libhivemall_b200 (include/hivemall_b200.h): what a Hive UDTF (train_classifier,
 * train_regressor, train_fm and the online learners) binds to.  One handle per task and GPU; not thread-safe.
 * Buffers are direct ByteBuffers in native byte order, best obtained from hostAlloc (pinned memory).
 * Errors surface as UDFArgumentException (bad options / arguments) or HiveException (everything else).
 *
 * NOT COMPILED IN THIS REPOSITORY'S IMAGE (no JDK): tests/test_jni_glue.py checks that every native method
 * below has its C function in hivemall_b200_jni.c and that the C file compiles against the library header.
 */
public final class HivemallB200 {
    static { System.loadLibrary("hivemall_b200_jni"); }

    private HivemallB200() {}

    public static native long create(String function, String options, int device);
    public static native ByteBuffer hostAlloc(long bytes);
    public static native void hostFree(ByteBuffer buf);
    public static native void process(long h, long nRows, ByteBuffer rowptr, ByteBuffer feature,
                                      ByteBuffer value, ByteBuffer label);
    public static native void processSplit(long h, long nRows, ByteBuffer qRowptr, ByteBuffer qFeature,
                                           ByteBuffer qValue, ByteBuffer cRowptr, ByteBuffer cFeature,
                                           ByteBuffer label);
    public static native void processColumns(long h, long nRows, int nQ, ByteBuffer qFeature, ByteBuffer qValue,
                                             int nC, ByteBuffer cFeature, ByteBuffer label);
    public static native void processStrings(long h, long nRows, ByteBuffer rowptr, ByteBuffer bytes,
                                             ByteBuffer offsets, ByteBuffer label);
    public static native long close(long h);
    public static native long forward(long h, ByteBuffer feature, ByteBuffer weight, ByteBuffer vif, long cap);
    public static native long saveModelTable(long h, String path, int format, char delimiter);
    public static native long loadModelTable(long h, String path, int format, char delimiter);
    public static native void destroy(long h);
    public static native void mhash(ByteBuffer bytes, ByteBuffer offsets, long n, int seed, int numFeatures,
                                    ByteBuffer out, int device);
    public static native void featureHashing(ByteBuffer bytes, ByteBuffer offsets, long n, int seed,
                                             int numFeatures, ByteBuffer outIndex, ByteBuffer outValue, int device);
}
