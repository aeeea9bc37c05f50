/* minimal stand-in for syntax checking the JNI shim where no JDK exists */
#include <stdint.h>
typedef int32_t jint; typedef int64_t jlong; typedef int8_t jbyte; typedef uint8_t jboolean; typedef float jfloat; typedef double jdouble; typedef jint jsize;
typedef void* jobject; typedef jobject jclass; typedef jobject jstring; typedef jobject jarray; typedef jarray jobjectArray; typedef jarray jbyteArray; typedef jarray jintArray;
struct JNINativeInterface_; typedef const struct JNINativeInterface_* JNIEnv;
struct JNINativeInterface_ {
  jclass (*FindClass)(JNIEnv*, const char*);
  jsize (*GetArrayLength)(JNIEnv*, jarray);
  jbyte* (*GetByteArrayElements)(JNIEnv*, jbyteArray, jboolean*);
  void* (*GetDirectBufferAddress)(JNIEnv*, jobject);
  jlong (*GetDirectBufferCapacity)(JNIEnv*, jobject);
  jint* (*GetIntArrayElements)(JNIEnv*, jintArray, jboolean*);
  jobject (*GetObjectArrayElement)(JNIEnv*, jobjectArray, jsize);
  const char* (*GetStringUTFChars)(JNIEnv*, jstring, jboolean*);
  jstring (*NewStringUTF)(JNIEnv*, const char*);
  void (*ReleaseByteArrayElements)(JNIEnv*, jbyteArray, jbyte*, jint);
  void (*ReleaseIntArrayElements)(JNIEnv*, jintArray, jint*, jint);
  void (*ReleaseStringUTFChars)(JNIEnv*, jstring, const char*);
  jint (*ThrowNew)(JNIEnv*, jclass, const char*);
};
#define JNIEXPORT
#define JNICALL
#define JNI_ABORT 2
