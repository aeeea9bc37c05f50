package ai.langstream.cuda;

import java.nio.ByteBuffer;
import java.nio.FloatBuffer;
import java.nio.IntBuffer;
import java.nio.LongBuffer;

/**
 * One-to-one JNI surface over include/langstream_cuda.h. All buffers must be DIRECT buffers in native
 * byte order; handles are opaque pointers carried as long. Every method throws
 * IllegalArgumentException (LSC_E_INVALID, LSC_E_TOO_LONG) or RuntimeException (anything else) with the
 * text of lsc_last_error().
 */
public final class LangstreamCuda {
    static {
        NativeLoader.loadOnce("langstream_cuda", "langstream_cuda_jni");
    }

    private LangstreamCuda() {}

    public static native String version();

    public static native int deviceCount();

    // ---- encoder (replaces DJL Criteria.loadModel / Predictor.predict) ----
    /** arch: 0 = BERT, 1 = MPNet (typeVocab 0, relBuckets / relMaxDistance from config.json); see lsc_encoder_config. */
    public static native long encoderCreate(
            int layers, int hidden, int heads, int ffn, int vocab, int maxPosition, int typeVocab,
            float lnEps, int arch, int relBuckets, int relMaxDistance, ByteBuffer weightsFp32, int device);

    public static native void encoderEmbed(
            long encoder, IntBuffer tokenIds, IntBuffer cuSeqlens, int nSeq, FloatBuffer out);

    /**
     * lsc_encoder_submit / lsc_encoder_wait: the asynchronous form (EmbeddingsService returns a CompletableFuture).
     * The three buffers must stay untouched until encoderWait(ticket) has returned; up to two tickets are in
     * flight per handle (upload of batch i+1 and download of batch i-1 overlap the kernels of batch i).
     */
    public static native long encoderSubmit(
            long encoder, IntBuffer tokenIds, IntBuffer cuSeqlens, int nSeq, FloatBuffer out);

    public static native void encoderWait(long encoder, long ticket);

    public static native int encoderDim(long encoder);

    // ---- host tokenizer (replaces DJL's HuggingFaceTokenizer inside predictor.predict) ----
    /** lsc_tokenizer_set_special: template tokens other than [CLS] / [SEP] / [UNK] (MPNet: "<s>", "</s>"); null keeps one. */
    public static native void tokenizerSetSpecial(long tokenizer, String cls, String sep, String unk);

    public static native long tokenizerCreate(byte[] vocabTxt, boolean doLowerCase, int maxLength, int nThreads);

    /**
     * lsc_tokenizer_encode: texts as UTF-8 bytes back to back in `utf8` with their lengths in `textNbytes`;
     * writes packed ids and cuSeqlens (nTexts + 1 entries) - the layout encoderEmbed / encoderSubmit take.
     */
    public static native void tokenizerEncode(
            long tokenizer, ByteBuffer utf8, IntBuffer textNbytes, int nTexts, IntBuffer tokenIds, IntBuffer cuSeqlens);

    public static native void tokenizerDestroy(long tokenizer);

    public static native void encoderDestroy(long encoder);

    // ---- vector index (replaces the HerdDB table and its cosine ORDER BY scan) ----
    public static native long indexCreate(int dim, long capacity, int device);

    public static native void indexUpsert(long index, LongBuffer ids, FloatBuffer vectors, int n);

    public static native void indexDelete(long index, LongBuffer ids, int n);

    public static native long indexSize(long index);

    public static native void indexSearch(
            long index, FloatBuffer queries, int b, int k, LongBuffer outIds, FloatBuffer outScores);

    /** lsc_index_reserve: allocate the scratch of later searches (up to b queries, k results) at start-up. */
    public static native void indexReserve(long index, int b, int k);

    // ---- the table spread over several GPUs behind one handle (lsc_sharded_index_*) ----
    public static native long shardedIndexCreate(int dim, long capacityPerShard, int[] devices);

    public static native void shardedIndexUpsert(long sharded, LongBuffer ids, FloatBuffer vectors, int n);

    public static native void shardedIndexDelete(long sharded, LongBuffer ids, int n);

    public static native long shardedIndexSize(long sharded);

    public static native void shardedIndexSearch(
            long sharded, FloatBuffer queries, int b, int k, LongBuffer outIds, FloatBuffer outScores);

    public static native void shardedIndexDestroy(long sharded);

    public static native void indexSave(long index, String path);

    public static native long indexLoad(String path, int device);

    public static native void indexDestroy(long index);

    /**
     * lsc_rerank_mmr: MMR re-rank (ReRankAgent.rerankMMR). termFreq is [nDocs][nTerms] (frequency of each
     * query token per document), docLen the token count per document; returns how many indices were
     * written to outOrder.
     */
    public static native int rerankMmr(
            java.nio.FloatBuffer docVecs, int nDocs, int dim, java.nio.IntBuffer termFreq,
            java.nio.IntBuffer docLen, int nTerms, int maxDocs, double lambda, double k1, double b,
            java.nio.IntBuffer outOrder);

    /**
     * lsc_rerank_term_statistics: ReRankAgent.tokenise (split on \\s+) and the term counts natively - fills termFreq
     * [nDocs][maxTerms] and docLen [nDocs] (direct buffers) and returns the number of query tokens.
     */
    public static native int rerankTermStatistics(
            String[] docTexts, String queryText, int maxTerms, java.nio.IntBuffer termFreq, java.nio.IntBuffer docLen);

    /**
     * lsc_rerank_mmr_batch_dev: a batch of queries (nDocs documents each, device pointers) re-ranked in one launch;
     * outOrder / outCount are device addresses with the given strides (in ints).
     */
    public static native void rerankMmrBatchDev(
            long dDocVecs, int nQueries, int nDocs, int dim, long dTermFreq, long dDocLen, int nTerms, int maxDocs,
            double lambda, double k1, double b, long dOutOrder, int orderStride, long dOutCount, int countStride, long stream);
}
