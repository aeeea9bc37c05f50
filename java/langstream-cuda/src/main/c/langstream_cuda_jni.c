/* JNI shim: unwrap direct buffers, call the C ABI, map a negative return code to a Java exception
 * carrying lsc_last_error(). Needs <jni.h> from a JDK (absent in the build image: not compiled here). */
#include <jni.h>
#include <stdint.h>

#include "langstream_cuda.h"

static void throw_lsc(JNIEnv* env, int rc) {
    const char* cls = (rc == LSC_E_INVALID || rc == LSC_E_TOO_LONG) ? "java/lang/IllegalArgumentException"
                                                                    : "java/lang/RuntimeException";
    (*env)->ThrowNew(env, (*env)->FindClass(env, cls), lsc_last_error());
}
#define ADDR(buf) ((*env)->GetDirectBufferAddress(env, (buf)))
#define CHECK(rc) do { int _rc = (rc); if (_rc != LSC_OK) throw_lsc(env, _rc); } while (0)

JNIEXPORT jstring JNICALL Java_ai_langstream_cuda_LangstreamCuda_version(JNIEnv* env, jclass c) {
    return (*env)->NewStringUTF(env, lsc_version());
}
JNIEXPORT jint JNICALL Java_ai_langstream_cuda_LangstreamCuda_deviceCount(JNIEnv* env, jclass c) {
    return lsc_device_count();
}
JNIEXPORT jlong JNICALL Java_ai_langstream_cuda_LangstreamCuda_encoderCreate(
        JNIEnv* env, jclass c, jint layers, jint hidden, jint heads, jint ffn, jint vocab, jint maxPos,
        jint typeVocab, jfloat eps, jobject weights, jint device) {
    lsc_encoder_config cfg = {layers, hidden, heads, ffn, vocab, maxPos, typeVocab, eps, 0, {0}};
    lsc_encoder* enc = NULL;
    CHECK(lsc_encoder_create(&cfg, ADDR(weights), (size_t)(*env)->GetDirectBufferCapacity(env, weights), device, &enc));
    return (jlong)(intptr_t)enc;
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_encoderEmbed(
        JNIEnv* env, jclass c, jlong enc, jobject ids, jobject cu, jint n, jobject out) {
    CHECK(lsc_encoder_embed((lsc_encoder*)(intptr_t)enc, (const int32_t*)ADDR(ids), (const int32_t*)ADDR(cu), n,
                            (float*)ADDR(out)));
}
JNIEXPORT jint JNICALL Java_ai_langstream_cuda_LangstreamCuda_encoderDim(JNIEnv* env, jclass c, jlong enc) {
    return lsc_encoder_dim((const lsc_encoder*)(intptr_t)enc);
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_encoderDestroy(JNIEnv* env, jclass c, jlong enc) {
    lsc_encoder_destroy((lsc_encoder*)(intptr_t)enc);
}
JNIEXPORT jlong JNICALL Java_ai_langstream_cuda_LangstreamCuda_indexCreate(
        JNIEnv* env, jclass c, jint dim, jlong capacity, jint device) {
    lsc_index* ix = NULL;
    CHECK(lsc_index_create(dim, capacity, device, &ix));
    return (jlong)(intptr_t)ix;
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_indexUpsert(
        JNIEnv* env, jclass c, jlong ix, jobject ids, jobject vecs, jint n) {
    CHECK(lsc_index_upsert((lsc_index*)(intptr_t)ix, (const int64_t*)ADDR(ids), (const float*)ADDR(vecs), n));
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_indexDelete(
        JNIEnv* env, jclass c, jlong ix, jobject ids, jint n) {
    CHECK(lsc_index_delete((lsc_index*)(intptr_t)ix, (const int64_t*)ADDR(ids), n));
}
JNIEXPORT jlong JNICALL Java_ai_langstream_cuda_LangstreamCuda_indexSize(JNIEnv* env, jclass c, jlong ix) {
    return lsc_index_size((const lsc_index*)(intptr_t)ix);
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_indexSearch(
        JNIEnv* env, jclass c, jlong ix, jobject q, jint b, jint k, jobject outIds, jobject outScores) {
    CHECK(lsc_index_search((lsc_index*)(intptr_t)ix, (const float*)ADDR(q), b, k, (int64_t*)ADDR(outIds),
                           (float*)ADDR(outScores)));
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_indexSave(JNIEnv* env, jclass c, jlong ix, jstring path) {
    const char* p = (*env)->GetStringUTFChars(env, path, NULL);
    int rc = lsc_index_save((lsc_index*)(intptr_t)ix, p);
    (*env)->ReleaseStringUTFChars(env, path, p);
    CHECK(rc);
}
JNIEXPORT jlong JNICALL Java_ai_langstream_cuda_LangstreamCuda_indexLoad(JNIEnv* env, jclass c, jstring path, jint device) {
    const char* p = (*env)->GetStringUTFChars(env, path, NULL);
    lsc_index* ix = NULL;
    int rc = lsc_index_load(p, device, &ix);
    (*env)->ReleaseStringUTFChars(env, path, p);
    CHECK(rc);
    return (jlong)(intptr_t)ix;
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_indexDestroy(JNIEnv* env, jclass c, jlong ix) {
    lsc_index_destroy((lsc_index*)(intptr_t)ix);
}

JNIEXPORT jint JNICALL Java_ai_langstream_cuda_LangstreamCuda_rerankMmr(
        JNIEnv* env, jclass c, jobject vecs, jint n, jint dim, jobject tf, jobject len, jint terms, jint max_docs,
        jdouble lambda, jdouble k1, jdouble b, jobject out) {
    int32_t count = 0;
    int rc = lsc_rerank_mmr((const float*)(*env)->GetDirectBufferAddress(env, vecs), n, dim,
                            (const int32_t*)(*env)->GetDirectBufferAddress(env, tf),
                            (const int32_t*)(*env)->GetDirectBufferAddress(env, len), terms, max_docs, lambda, k1, b,
                            (int32_t*)(*env)->GetDirectBufferAddress(env, out), &count);
    if (rc != LSC_OK) throw_lsc(env, rc);
    return count;
}
