/* Stand-in for <jni.h>: declares exactly the JNI names jni/ytk_jni.cpp uses, with the JDK's signatures, so
 * that the glue can be parsed and type-checked (g++ -fsyntax-only) on a machine without a JDK.  Never
 * shipped and never linked: a real build uses $JAVA_HOME/include/jni.h. */
#ifndef YTK_STUB_JNI_H
#define YTK_STUB_JNI_H
#include <stdint.h>

#define JNIEXPORT __attribute__((visibility("default")))
#define JNICALL
#define JNI_ABORT 2

typedef int32_t jint;
typedef int64_t jlong;
typedef int8_t jbyte;
typedef uint8_t jboolean;
typedef float jfloat;
typedef double jdouble;
typedef jint jsize;

class _jobject {};
class _jclass : public _jobject {};
class _jarray : public _jobject {};
class _jintArray : public _jarray {};
class _jlongArray : public _jarray {};
class _jbyteArray : public _jarray {};
class _jfloatArray : public _jarray {};
class _jdoubleArray : public _jarray {};
typedef _jobject* jobject;
typedef _jclass* jclass;
typedef _jarray* jarray;
typedef _jintArray* jintArray;
typedef _jlongArray* jlongArray;
typedef _jbyteArray* jbyteArray;
typedef _jfloatArray* jfloatArray;
typedef _jdoubleArray* jdoubleArray;

struct JNIEnv {
    jclass FindClass(const char* name);
    jint ThrowNew(jclass clazz, const char* msg);
    jsize GetArrayLength(jarray array);
    void* GetPrimitiveArrayCritical(jarray array, jboolean* isCopy);
    void ReleasePrimitiveArrayCritical(jarray array, void* carray, jint mode);
    void SetDoubleArrayRegion(jdoubleArray array, jsize start, jsize len, const jdouble* buf);
};
#endif
