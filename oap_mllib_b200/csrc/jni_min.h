// Minimal JNI declarations (there is no jni.h / JDK in this image).  Only what jni_shim.cpp uses:
// the primitive typedefs and the slot numbers of the JNINativeInterface function table, which are
// fixed by the JNI specification (Java SE "JNI Functions", Interface Function Table).
#pragma once
#include <stdint.h>

typedef int32_t jint;
typedef int64_t jlong;
typedef int8_t jbyte;
typedef uint8_t jboolean;
typedef double jdouble;
typedef float jfloat;
typedef int32_t jsize;
typedef void *jobject;
typedef jobject jclass;
typedef jobject jstring;
typedef jobject jarray;
typedef jarray jdoubleArray;
typedef jarray jbyteArray;
typedef jarray jfloatArray;
typedef jarray jlongArray;
typedef jarray jintArray;
typedef jarray jobjectArray;
typedef void *jfieldID;

#define JNI_ABORT 2
#define JNIEXPORT __attribute__((visibility("default")))
#define JNICALL

// JNIEnv is a pointer to a pointer to the function table (C view of jni.h)
typedef void *const *JNIEnvTable;
typedef JNIEnvTable *JNIEnvPtr;  // == JNIEnv*

enum JniSlot {
    kJniFindClass = 6,
    kJniThrowNew = 14,
    kJniGetObjectClass = 31,
    kJniGetFieldID = 94,
    kJniSetIntField = 109,
    kJniSetLongField = 110,
    kJniSetDoubleField = 112,
    kJniGetStringUTFChars = 169,
    kJniReleaseStringUTFChars = 170,
    kJniGetArrayLength = 171,
    kJniGetObjectArrayElement = 173,
    kJniNewByteArray = 176,
    kJniNewLongArray = 180,
    kJniNewDoubleArray = 182,
    kJniGetByteArrayRegion = 200,
    kJniSetByteArrayRegion = 208,
    kJniSetIntArrayRegion = 211,
    kJniSetLongArrayRegion = 212,
    kJniSetDoubleArrayRegion = 214,
    kJniGetPrimitiveArrayCritical = 222,
    kJniReleasePrimitiveArrayCritical = 223,
    kJniTableSize = 235
};
