/* JNI shim: unwrap direct buffers, call the C ABI, map a negative return code to a Java exception
 * carrying lsc_last_error(). Needs <jni.h> from a JDK (absent in the build image: not compiled here). */
#include <jni.h>
#include <stdint.h>
#include <stdlib.h>

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
        jint typeVocab, jfloat eps, jint arch, jint relBuckets, jint relMaxDistance, jobject weights, jint device) {
    lsc_encoder_config cfg = {layers, hidden, heads, ffn, vocab, maxPos, typeVocab, eps, 0, arch, relBuckets, relMaxDistance, {0}};
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

JNIEXPORT jint JNICALL Java_ai_langstream_cuda_LangstreamCuda_rerankTermStatistics(
        JNIEnv* env, jclass c, jobjectArray docs, jstring query, jint max_terms, jobject tf, jobject len) {
    const jsize n = (*env)->GetArrayLength(env, docs);
    const char** texts = (const char**)calloc((size_t)(n > 0 ? n : 1), sizeof(char*));
    jstring* refs = (jstring*)calloc((size_t)(n > 0 ? n : 1), sizeof(jstring));
    for (jsize i = 0; i < n; ++i) {   /* modified UTF-8: identical to UTF-8 for the white-space split and byte equality of tokens */
        refs[i] = (jstring)(*env)->GetObjectArrayElement(env, docs, i);
        texts[i] = refs[i] ? (*env)->GetStringUTFChars(env, refs[i], NULL) : NULL;
    }
    const char* q = query ? (*env)->GetStringUTFChars(env, query, NULL) : NULL;
    int32_t n_terms = 0;
    int rc = lsc_rerank_term_statistics(texts, NULL, n, q, -1, max_terms, (int32_t*)(*env)->GetDirectBufferAddress(env, tf),
                                        (int32_t*)(*env)->GetDirectBufferAddress(env, len), &n_terms);
    if (q) (*env)->ReleaseStringUTFChars(env, query, q);
    for (jsize i = 0; i < n; ++i)
        if (refs[i]) (*env)->ReleaseStringUTFChars(env, refs[i], texts[i]);
    free(refs);
    free((void*)texts);
    if (rc != LSC_OK) throw_lsc(env, rc);
    return n_terms;
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_rerankMmrBatchDev(
        JNIEnv* env, jclass c, jlong vecs, jint nq, jint n, jint dim, jlong tf, jlong len, jint terms, jint max_docs,
        jdouble lambda, jdouble k1, jdouble b, jlong order, jint order_stride, jlong count, jint count_stride, jlong stream) {
    CHECK(lsc_rerank_mmr_batch_dev((const float*)(intptr_t)vecs, nq, n, dim, (const int32_t*)(intptr_t)tf,
                                   (const int32_t*)(intptr_t)len, terms, max_docs, lambda, k1, b, (int32_t*)(intptr_t)order,
                                   order_stride, (int32_t*)(intptr_t)count, count_stride, (void*)(intptr_t)stream));
}

/* ---- asynchronous encoder calls, tokenizer, sharded index (same conventions as above) ---- */
JNIEXPORT jlong JNICALL Java_ai_langstream_cuda_LangstreamCuda_encoderSubmit(
        JNIEnv* env, jclass c, jlong enc, jobject ids, jobject cu, jint n, jobject out) {
    int64_t ticket = -1;
    CHECK(lsc_encoder_submit((lsc_encoder*)(intptr_t)enc, (const int32_t*)ADDR(ids), (const int32_t*)ADDR(cu), n,
                             (float*)ADDR(out), &ticket));
    return (jlong)ticket;
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_encoderWait(JNIEnv* env, jclass c, jlong enc, jlong ticket) {
    CHECK(lsc_encoder_wait((lsc_encoder*)(intptr_t)enc, (int64_t)ticket));
}
JNIEXPORT jlong JNICALL Java_ai_langstream_cuda_LangstreamCuda_tokenizerCreate(
        JNIEnv* env, jclass c, jbyteArray vocab, jboolean lower, jint maxLength, jint nThreads) {
    jsize n = (*env)->GetArrayLength(env, vocab);
    jbyte* p = (*env)->GetByteArrayElements(env, vocab, NULL);
    lsc_tokenizer* tok = NULL;
    int rc = lsc_tokenizer_create((const char*)p, (size_t)n, lower ? 1 : 0, maxLength, nThreads, &tok);
    (*env)->ReleaseByteArrayElements(env, vocab, p, JNI_ABORT);
    CHECK(rc);
    return (jlong)(intptr_t)tok;
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_tokenizerSetSpecial(
        JNIEnv* env, jclass c, jlong tok, jstring cls, jstring sep, jstring unk) {
    const char* a = cls ? (*env)->GetStringUTFChars(env, cls, NULL) : NULL;
    const char* b = sep ? (*env)->GetStringUTFChars(env, sep, NULL) : NULL;
    const char* u = unk ? (*env)->GetStringUTFChars(env, unk, NULL) : NULL;
    int rc = lsc_tokenizer_set_special((lsc_tokenizer*)(intptr_t)tok, a, b, u);
    if (a) (*env)->ReleaseStringUTFChars(env, cls, a);
    if (b) (*env)->ReleaseStringUTFChars(env, sep, b);
    if (u) (*env)->ReleaseStringUTFChars(env, unk, u);
    CHECK(rc);
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_tokenizerEncode(
        JNIEnv* env, jclass c, jlong tok, jobject utf8, jobject nbytes, jint nTexts, jobject ids, jobject cu) {
    /* the texts lie back to back: build the pointer table the C ABI takes */
    const char* base = (const char*)ADDR(utf8);
    const int32_t* len = (const int32_t*)ADDR(nbytes);
    const char** ptrs = (const char**)malloc(sizeof(char*) * (size_t)(nTexts > 0 ? nTexts : 1));
    size_t off = 0;
    for (jint i = 0; i < nTexts; ++i) {
        ptrs[i] = base + off;
        off += (size_t)len[i];
    }
    int rc = lsc_tokenizer_encode((lsc_tokenizer*)(intptr_t)tok, ptrs, len, nTexts, (int32_t*)ADDR(ids),
                                  (int64_t)((*env)->GetDirectBufferCapacity(env, ids)), (int32_t*)ADDR(cu));
    free(ptrs);
    CHECK(rc);
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_tokenizerDestroy(JNIEnv* env, jclass c, jlong tok) {
    lsc_tokenizer_destroy((lsc_tokenizer*)(intptr_t)tok);
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_indexReserve(JNIEnv* env, jclass c, jlong ix, jint b, jint k) {
    CHECK(lsc_index_reserve((lsc_index*)(intptr_t)ix, b, k));
}
JNIEXPORT jlong JNICALL Java_ai_langstream_cuda_LangstreamCuda_shardedIndexCreate(
        JNIEnv* env, jclass c, jint dim, jlong capacity, jintArray devices) {
    jsize n = (*env)->GetArrayLength(env, devices);
    jint* d = (*env)->GetIntArrayElements(env, devices, NULL);
    lsc_sharded_index* s = NULL;
    int rc = lsc_sharded_index_create(dim, capacity, (const int*)d, n, &s);
    (*env)->ReleaseIntArrayElements(env, devices, d, JNI_ABORT);
    CHECK(rc);
    return (jlong)(intptr_t)s;
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_shardedIndexUpsert(
        JNIEnv* env, jclass c, jlong s, jobject ids, jobject vecs, jint n) {
    CHECK(lsc_sharded_index_upsert((lsc_sharded_index*)(intptr_t)s, (const int64_t*)ADDR(ids), (const float*)ADDR(vecs), n));
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_shardedIndexDelete(
        JNIEnv* env, jclass c, jlong s, jobject ids, jint n) {
    CHECK(lsc_sharded_index_delete((lsc_sharded_index*)(intptr_t)s, (const int64_t*)ADDR(ids), n));
}
JNIEXPORT jlong JNICALL Java_ai_langstream_cuda_LangstreamCuda_shardedIndexSize(JNIEnv* env, jclass c, jlong s) {
    return lsc_sharded_index_size((const lsc_sharded_index*)(intptr_t)s);
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_shardedIndexSearch(
        JNIEnv* env, jclass c, jlong s, jobject q, jint b, jint k, jobject outIds, jobject outScores) {
    CHECK(lsc_sharded_index_search((lsc_sharded_index*)(intptr_t)s, (const float*)ADDR(q), b, k, (int64_t*)ADDR(outIds),
                                   (float*)ADDR(outScores)));
}
JNIEXPORT void JNICALL Java_ai_langstream_cuda_LangstreamCuda_shardedIndexDestroy(JNIEnv* env, jclass c, jlong s) {
    lsc_sharded_index_destroy((lsc_sharded_index*)(intptr_t)s);
}
