/*
 * Declaration-only subset of the Java Native Interface header, written for this repository's tests: just the types and
 * the JNIEnv member functions jni/ftf_jni.cc uses, with the signatures the JNI specification gives them.  It exists so
 * that the shim can be type-checked (g++ -fsyntax-only) on a machine without a JDK.  It is NOT a JNI implementation and
 * nothing links against it.
 */
#ifndef FTF_TEST_JNI_H_
#define FTF_TEST_JNI_H_
#include <stdint.h>

#define JNIEXPORT __attribute__((visibility("default")))
#define JNICALL
#define JNI_FALSE 0
#define JNI_TRUE 1

typedef int32_t jint;
typedef int64_t jlong;
typedef uint8_t jboolean;
typedef float jfloat;
typedef jint jsize;

class _jobject {};
class _jclass : public _jobject {};
class _jstring : public _jobject {};
class _jarray : public _jobject {};
class _jobjectArray : public _jarray {};
class _jintArray : public _jarray {};
class _jlongArray : public _jarray {};
typedef _jobject* jobject;
typedef _jclass* jclass;
typedef _jstring* jstring;
typedef _jarray* jarray;
typedef _jobjectArray* jobjectArray;
typedef _jintArray* jintArray;
typedef _jlongArray* jlongArray;

struct JNIEnv_ {
    jclass FindClass(const char* name);
    jint ThrowNew(jclass clazz, const char* msg);
    const char* GetStringUTFChars(jstring str, jboolean* isCopy);
    void ReleaseStringUTFChars(jstring str, const char* chars);
    jstring NewStringUTF(const char* utf);
    jsize GetArrayLength(jarray array);
    jobjectArray NewObjectArray(jsize len, jclass clazz, jobject init);
    jobject GetObjectArrayElement(jobjectArray array, jsize index);
    void SetObjectArrayElement(jobjectArray array, jsize index, jobject val);
    jlongArray NewLongArray(jsize len);
    void GetIntArrayRegion(jintArray array, jsize start, jsize len, jint* buf);
    void SetIntArrayRegion(jintArray array, jsize start, jsize len, const jint* buf);
    void GetLongArrayRegion(jlongArray array, jsize start, jsize len, jlong* buf);
    void SetLongArrayRegion(jlongArray array, jsize start, jsize len, const jlong* buf);
    jobject NewDirectByteBuffer(void* address, jlong capacity);
    void* GetDirectBufferAddress(jobject buf);
    jlong GetDirectBufferCapacity(jobject buf);
    void DeleteLocalRef(jobject obj);
};
typedef JNIEnv_ JNIEnv;

#endif
