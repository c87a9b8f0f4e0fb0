/*
 * Stand-in for the JDK's jni.h — TEST INFRASTRUCTURE. This image has no JDK, so jni/prestogpu_jni.c cannot be built for real;
 * this header declares the JNI types and exactly the JNIEnv / JavaVM entry points the shim uses, with the JNI specification's
 * signatures, so that the CPU suite can at least compile the shim (tests/test_jni_shim.py) and check that every native declared
 * in PrestoGpuNative.java is defined. It is NOT ABI compatible with a real JVM (the function tables are not in JNI order).
 */
#ifndef PGPU_JNI_STUB_H
#define PGPU_JNI_STUB_H
#include <stdint.h>

#define JNIEXPORT __attribute__((visibility("default")))
#define JNICALL
#define JNI_OK 0
#define JNI_ABORT 2

typedef int32_t jint;
typedef int64_t jlong;
typedef int8_t jbyte;
typedef uint8_t jboolean;
typedef jint jsize;
struct _jobject;
typedef struct _jobject* jobject;
typedef jobject jclass, jstring, jarray, jthrowable, jobjectArray, jbyteArray, jintArray, jlongArray, jbooleanArray;
struct _jmethodID;
typedef struct _jmethodID* jmethodID;

struct JNINativeInterface_;
typedef const struct JNINativeInterface_* JNIEnv;
struct JNIInvokeInterface_;
typedef const struct JNIInvokeInterface_* JavaVM;

struct JNINativeInterface_ {
    jclass (*FindClass)(JNIEnv*, const char*);
    jclass (*GetObjectClass)(JNIEnv*, jobject);
    jmethodID (*GetMethodID)(JNIEnv*, jclass, const char*, const char*);
    jmethodID (*GetStaticMethodID)(JNIEnv*, jclass, const char*, const char*);
    jobject (*CallStaticObjectMethod)(JNIEnv*, jclass, jmethodID, ...);
    void (*CallVoidMethod)(JNIEnv*, jobject, jmethodID, ...);
    jobject (*NewObject)(JNIEnv*, jclass, jmethodID, ...);
    jint (*Throw)(JNIEnv*, jthrowable);
    jstring (*NewStringUTF)(JNIEnv*, const char*);
    jobject (*NewGlobalRef)(JNIEnv*, jobject);
    void (*DeleteGlobalRef)(JNIEnv*, jobject);
    jsize (*GetArrayLength)(JNIEnv*, jarray);
    jobjectArray (*NewObjectArray)(JNIEnv*, jsize, jclass, jobject);
    jobject (*GetObjectArrayElement)(JNIEnv*, jobjectArray, jsize);
    void (*SetObjectArrayElement)(JNIEnv*, jobjectArray, jsize, jobject);
    jbyteArray (*NewByteArray)(JNIEnv*, jsize);
    jintArray (*NewIntArray)(JNIEnv*, jsize);
    jlongArray (*NewLongArray)(JNIEnv*, jsize);
    jbooleanArray (*NewBooleanArray)(JNIEnv*, jsize);
    jbyte* (*GetByteArrayElements)(JNIEnv*, jbyteArray, jboolean*);
    void (*ReleaseByteArrayElements)(JNIEnv*, jbyteArray, jbyte*, jint);
    void (*GetIntArrayRegion)(JNIEnv*, jintArray, jsize, jsize, jint*);
    void (*GetLongArrayRegion)(JNIEnv*, jlongArray, jsize, jsize, jlong*);
    void (*SetByteArrayRegion)(JNIEnv*, jbyteArray, jsize, jsize, const jbyte*);
    void (*SetIntArrayRegion)(JNIEnv*, jintArray, jsize, jsize, const jint*);
    void (*SetLongArrayRegion)(JNIEnv*, jlongArray, jsize, jsize, const jlong*);
    void (*SetBooleanArrayRegion)(JNIEnv*, jbooleanArray, jsize, jsize, const jboolean*);
    void* (*GetPrimitiveArrayCritical)(JNIEnv*, jarray, jboolean*);
    void (*ReleasePrimitiveArrayCritical)(JNIEnv*, jarray, void*, jint);
    jint (*GetJavaVM)(JNIEnv*, JavaVM**);
};

struct JNIInvokeInterface_ {
    jint (*AttachCurrentThreadAsDaemon)(JavaVM*, void**, void*);
    jint (*DetachCurrentThread)(JavaVM*);
};
#endif
