/*
 * langstream_cuda.h — C ABI of the langstream-cuda native module (B200 / sm_100a).
 *
 * This is the drop-in boundary for LangStream's one data-parallel hot path:
 *   - the local-model `compute-ai-embeddings` agent (a BERT-family sentence encoder), and
 *   - the brute-force cosine top-k behind `query-vector-db` (+ the `vector-db-sink` write side).
 *
 * The reference has no native code: its Java calls DJL (libtorch CPU through JNI) and HerdDB
 * (JDBC).  Each entry point below cites the reference interface whose bottom layer it replaces
 * (paths relative to the reference tree; AI/ = langstream-agents/langstream-ai-agents/src/main/java,
 * VEC/ = langstream-agents/langstream-vector-agents/src/main/java/ai/langstream/agents/vector,
 * API/ = langstream-api/src/main/java/ai/langstream/api).
 *
 * Conventions
 *   - plain C: opaque handles, pointers and sizes only; no exceptions cross the boundary;
 *   - every call returns 0 on success and a negative LSC_E_* code on failure; the message is
 *     available from lsc_last_error() (thread-local, valid until the next failing call on the
 *     same thread);
 *   - the caller owns every input and output buffer; nothing returned aliases native memory;
 *   - one handle drives one GPU; use one handle (and one process or thread) per GPU;
 *   - calls on the same handle are serialised internally (thread-safe, not concurrent), on the GPU as well: the
 *     *_dev / submit calls return without synchronising, but a pass enqueued on one stream waits (through an event)
 *     for the handle's previous pass on any other stream, because all passes share the handle's scratch;
 *   - there is no CPU fallback: without a usable sm_100 device every create call fails.
 */
#ifndef LANGSTREAM_CUDA_H
#define LANGSTREAM_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LSC_OK 0
#define LSC_E_INVALID (-1)    /* bad argument (IllegalArgumentException on the Java side) */
#define LSC_E_CUDA (-2)       /* CUDA runtime / driver error */
#define LSC_E_NOMEM (-3)      /* host or device allocation failed */
#define LSC_E_UNSUPPORTED (-4)/* shape outside what the kernels support */
#define LSC_E_IO (-5)         /* save / load failed */
#define LSC_E_TOO_LONG (-6)   /* a sequence exceeds max_position (reference: TranslateException
                                 "Illegal input, maybe the number of tokens is too high",
                                 AI/com/datastax/oss/streaming/ai/embeddings/AbstractHuggingFaceEmbeddingService.java:197-203) */

typedef struct lsc_encoder lsc_encoder;
typedef struct lsc_index lsc_index;
typedef struct lsc_exchange lsc_exchange;
typedef struct lsc_sharded_index lsc_sharded_index;

/* ---- library ---------------------------------------------------------------------------- */

/* Version string of the native module, e.g. "langstream-cuda 0.1 sm_100a". */
const char* lsc_version(void);
/* Last error message of the calling thread ("" if none). */
const char* lsc_last_error(void);
/* Number of CUDA devices visible; <0 on error. */
int lsc_device_count(void);

/* ---- encoder: replaces the DJL/libtorch forward pass ------------------------------------- */

/*
 * Architecture of the BERT-family encoder (post-LN, absolute positions, exact-erf GELU,
 * LayerNorm eps as given).  MiniLM-L6: layers 6, hidden 384, heads 12, ffn 1536;
 * BGE-base: 12 / 768 / 12 / 3072.  hidden / heads must be 32 or 64.
 */
typedef struct lsc_encoder_config {
    int32_t layers;
    int32_t hidden;
    int32_t heads;
    int32_t ffn;
    int32_t vocab;
    int32_t max_position;
    int32_t type_vocab;
    float ln_eps;
    int32_t gemm_impl; /* 0 = tcgen05/TMEM/TMA (product path); 1 = CUDA-core validation GEMM
                          (debug only: used by the tests to cross-check the tensor-core path) */
    int32_t arch;      /* 0 = BERT (BertModel: all-MiniLM, BGE, e5); 1 = MPNet (MPNetModel: all-mpnet-base-v2) - no
                          token-type embeddings (type_vocab must be 0), position ids start at 2, and a relative-position
                          bias shared by all layers is added to the attention scores; needs a head dimension of 64 */
    int32_t rel_buckets;      /* arch 1: relative_attention_num_buckets (32) */
    int32_t rel_max_distance; /* arch 1: distance at which the buckets saturate (128; at most 128) */
    int32_t reserved[4];
} lsc_encoder_config;

/*
 * Number of bytes of the fp32 weight blob lsc_encoder_create expects for `cfg`.
 * Blob layout (all fp32, little endian, row-major, PyTorch shapes), in this order:
 *   word_embeddings [vocab,hidden], position_embeddings [max_position,hidden],
 *   token_type_embeddings [type_vocab,hidden], embeddings.LayerNorm weight[hidden], bias[hidden],
 *   then per layer:
 *     query.weight [hidden,hidden], query.bias, key.weight, key.bias, value.weight, value.bias,
 *     attention.output.dense.weight [hidden,hidden], .bias, attention.output.LayerNorm weight, bias,
 *     intermediate.dense.weight [ffn,hidden], .bias, output.dense.weight [hidden,ffn], .bias,
 *     output.LayerNorm weight, bias.
 * arch 1 (MPNetModel): no token_type_embeddings; encoder.relative_attention_bias.weight [rel_buckets,heads] follows the
 * embeddings LayerNorm; per layer attention.attn.q / k / v / o and attention.LayerNorm take the places of query / key /
 * value / attention.output.dense and attention.output.LayerNorm.
 * This is the state_dict of the HuggingFace BertModel that DJL traces and runs
 * (AI/com/datastax/oss/streaming/ai/embeddings/AbstractHuggingFaceEmbeddingService.java:132-161).
 */
size_t lsc_encoder_weights_nbytes(const lsc_encoder_config* cfg);

/*
 * Load a model onto GPU `device`.
 * Replaces: AbstractHuggingFaceEmbeddingService constructor → Criteria.loadModel()
 * (AI/com/datastax/oss/streaming/ai/embeddings/AbstractHuggingFaceEmbeddingService.java:110-172).
 */
int lsc_encoder_create(const lsc_encoder_config* cfg, const void* weights, size_t nbytes,
                       int device, lsc_encoder** out);

/*
 * Embed n_seq token sequences given in packed (variable-length) form:
 * token_ids holds cu_seqlens[n_seq] int32 ids, sequence i is token_ids[cu_seqlens[i] .. cu_seqlens[i+1]).
 * Writes n_seq × hidden fp32, row i = L2-normalised masked mean of sequence i's last hidden
 * states.  An empty sequence yields a zero row.  Host pointers; blocks until the result is in `out`.
 * Replaces: the per-text loop `predictor.predict(in)` — tokenised input → BERT forward → mean-pool
 * → L2-normalise → float[hidden]
 * (AI/com/datastax/oss/streaming/ai/embeddings/AbstractHuggingFaceEmbeddingService.java:183-207);
 * the caller widens to List<Double> as
 * AI/com/datastax/oss/streaming/ai/embeddings/HuggingFaceEmbeddingService.java:52-63 does.
 */
int lsc_encoder_embed(lsc_encoder* enc, const int32_t* token_ids, const int32_t* cu_seqlens,
                      int32_t n_seq, float* out);

/*
 * Same computation with DEVICE pointers, enqueued on `stream` (a cudaStream_t, may be NULL for
 * the default stream); returns without synchronising.  `total_tokens` = cu_seqlens[n_seq] and
 * `max_seqlen` = the longest sequence, both known to the host caller.
 * Same reference function as lsc_encoder_embed; used when token ids already live in HBM.
 */
int lsc_encoder_embed_dev(lsc_encoder* enc, const int32_t* d_token_ids, const int32_t* d_cu_seqlens,
                          int32_t n_seq, int32_t total_tokens, int32_t max_seqlen, float* d_out,
                          void* stream);

/*
 * Asynchronous form of lsc_encoder_embed - the shape of the reference's contract, which returns a
 * CompletableFuture (EmbeddingsService.computeEmbeddings, EmbeddingsService.java:24-37).
 * lsc_encoder_submit validates, enqueues upload, kernels and download and returns a ticket at once;
 * lsc_encoder_wait(ticket) blocks until that batch's embeddings are in `out`.  Up to two submissions are in
 * flight per handle (a third submit first waits for the oldest): the upload of batch i+1 and the download of
 * batch i-1 run on their own streams beside the kernels of batch i, so a caller that keeps two batches in
 * flight sees the device-resident rate.  token_ids, cu_seqlens and out must stay valid (and untouched) until
 * the ticket has been waited for; pinned host memory is what makes the copies asynchronous.  Batches are
 * processed in submission order.  A batch that needs more than one pass (> 262,144 tokens) is served
 * synchronously inside submit.
 */
int lsc_encoder_submit(lsc_encoder* enc, const int32_t* token_ids, const int32_t* cu_seqlens, int32_t n_seq,
                       float* out, int64_t* ticket);
int lsc_encoder_wait(lsc_encoder* enc, int64_t ticket);

/*
 * Measurement aid (no reference counterpart): when enabled, every kernel launch of this handle is
 * bracketed by CUDA events on its stream.  lsc_encoder_read_profile waits for them and returns the
 * milliseconds and launch counts accumulated since the previous read, per kernel class:
 * 0 embedding gather+LayerNorm, 1 GEMM, 2 attention, 3 LayerNorm, 4 mean-pool+L2, 5 fused feed-forward block.
 */
int lsc_encoder_set_profiling(lsc_encoder* enc, int enabled);
int lsc_encoder_read_profile(lsc_encoder* enc, float* ms_by_class, int64_t* launches_by_class,
                             int32_t nclass);

/* Hidden size of the loaded model (the embedding dimension). */
int lsc_encoder_dim(const lsc_encoder* enc);
/* Number of kernels this handle has launched since creation (bench accounting). */
int64_t lsc_encoder_launch_count(const lsc_encoder* enc);
/* Replaces: EmbeddingsService.close() (AI/com/datastax/oss/streaming/ai/embeddings/EmbeddingsService.java:34). */
void lsc_encoder_destroy(lsc_encoder* enc);

/* ---- vector index: replaces the HerdDB table + cosine ORDER BY scan ----------------------- */

/*
 * Create an empty index of `dim`-dimensional vectors on GPU `device` with room for `capacity`
 * rows (grows on demand).  dim must be a multiple of 64 and <= 512.
 * Replaces: JdbcDataSourceImpl.initialize / the `CREATE TABLE … FLOATA` asset
 * (AI/ai/langstream/ai/agents/datasource/impl/JdbcDataSourceProvider.java:61-65,
 *  examples/applications/docker-chatbot/crawler.yaml:33).
 */
int lsc_index_create(int32_t dim, int64_t capacity, int device, lsc_index** out);

/*
 * Insert or replace n rows (host fp32, n × dim).  Rows are L2-normalised in fp32 and stored as
 * bf16; a zero vector stays zero and scores 0 against every query, as
 * ReRankAgent.cosineSimilarity does (AI/ai/langstream/ai/agents/rerank/ReRankAgent.java:262-273).
 * Replaces: JdbcVectorDatabaseWriter.upsert — UPDATE, INSERT when no row matched
 * (VEC/jdbc/JdbcWriter.java:139-179).
 */
int lsc_index_upsert(lsc_index* idx, const int64_t* ids, const float* vecs, int32_t n);

/* Device-pointer variant of upsert for rows already in HBM (ids stay on the host). */
int lsc_index_upsert_dev(lsc_index* idx, const int64_t* ids, const float* d_vecs, int32_t n,
                         void* stream);

/*
 * Remove rows by id; unknown ids are ignored.
 * Replaces: the DELETE branch of JdbcVectorDatabaseWriter.upsert when record.value()==null
 * (VEC/jdbc/JdbcWriter.java:139-179; contract at API/database/VectorDatabaseWriter.java:30-36).
 */
int lsc_index_delete(lsc_index* idx, const int64_t* ids, int32_t n);

/* Number of live rows. */
int64_t lsc_index_size(const lsc_index* idx);

/*
 * Brute-force cosine top-k: for each of B queries (host fp32, B × dim) write the k best rows,
 * best first, as ids (B × k int64, -1 when fewer than k rows exist) and scores (B × k fp32 cosine).
 * Order: score descending, ties by lower storage slot (= insertion order while nothing was deleted).
 * Replaces: JdbcDataSourceImpl.fetchData running
 * `SELECT … ORDER BY cosine_similarity(col, CAST(? AS FLOAT ARRAY)) DESC LIMIT k` on HerdDB
 * (AI/ai/langstream/ai/agents/datasource/impl/JdbcDataSourceProvider.java:103-124;
 *  SQL at examples/applications/docker-chatbot/chatbot.yaml:44).  1 <= k <= 128.
 */
int lsc_index_search(lsc_index* idx, const float* queries, int32_t B, int32_t k, int64_t* out_ids,
                     float* out_scores);

/*
 * Device-pointer variant: queries, out_ids and out_scores are device pointers; enqueued on
 * `stream`, no synchronisation.  Used by the sharded multi-GPU path, which all-gathers the per-shard
 * (score,id) candidates and then calls lsc_topk_merge_dev.
 */
int lsc_index_search_dev(lsc_index* idx, const float* d_queries, int32_t B, int32_t k,
                         int64_t* d_out_ids, float* d_out_scores, void* stream);

/*
 * Debug / verification only: exact fp32 dot products of EVERY stored row with every query, by a
 * plain CUDA-core kernel (no tensor cores): d_scores[b * size + slot], slots in storage order,
 * queries normalised and rounded to bf16 exactly as the search does.  The tests use it as the
 * on-device checker at full size.
 */
int lsc_index_scores_naive_dev(lsc_index* idx, const float* d_queries, int32_t B, float* d_scores,
                               void* stream);

/*
 * Merge `parts` candidate lists per query (device pointers, layout [parts][B][k], each list sorted
 * best first, id -1 = empty) into one top-k per query ([B][k]).  This is the step after the NCCL
 * all-gather in the sharded search; it has no counterpart in the single-node reference.
 */
int lsc_topk_merge_dev(const int64_t* d_ids, const float* d_scores, int32_t parts, int32_t B,
                       int32_t k, int64_t* d_out_ids, float* d_out_scores, void* stream);

/*
 * Top-k search that allocates nothing (no counterpart in the reference: it exists for the way this
 * library is deployed, not for anything LangStream's Java does).
 * Meant for callers that cannot let the library call cudaMalloc at query time: agents sharing the GPU
 * with a framework that reserved nearly all of HBM at start-up (a late allocation would fail), and
 * latency-bound query threads (cudaMalloc / cudaFree synchronise the whole device).
 * The caller passes everything as device pointers: `d_rows_bf16` = n_rows x dim L2-normalised bf16
 * rows (16-byte aligned, e.g. produced once at load time), `d_row_ids` = n_rows int64 ids or NULL to
 * get row numbers back, `d_queries` = B x dim fp32, outputs B x k.  Same kernels, ordering and scores
 * as lsc_index_search_dev; the scratch they need lives in a fixed ~5.4 MB array inside the library
 * (held by every GPU that loads the library until the process exits), hence the limits
 * B <= 128, k <= 32, dim a multiple of 64 and <= 512; anything larger is rejected with
 * LSC_E_UNSUPPORTED.  The first call on a device resolves the array's address and configures the
 * kernels: make it at start-up, before any stream capture on that device.  All calls share the array
 * (it keeps the last call's contents; nothing needs to be zeroed between calls), so calls whose GPU
 * work overlaps — different streams or threads — race: keep them on one stream.  Enqueues on `stream`
 * and returns without synchronising.
 */
int lsc_topk_search_static_dev(const void* d_rows_bf16, const int64_t* d_row_ids, int64_t n_rows,
                               int32_t dim, const float* d_queries, int32_t B, int32_t k,
                               int64_t* d_out_ids, float* d_out_scores, void* stream);

/* ---- row-sharded corpus: one shard per GPU ------------------------------------------------------ */

/*
 * No counterpart in the single-node reference (HerdDB answers `ORDER BY cosine_similarity(...) LIMIT k` inside one
 * JVM, AI/ai/langstream/ai/agents/datasource/impl/JdbcDataSourceProvider.java:103-124); this is the multi-GPU
 * form of the same query: the table's rows are spread over the GPUs, every GPU scans its shard and the per-shard
 * top-k lists are merged.
 *
 * lsc_exchange = one GPU's end of the step in between: a receive buffer other GPUs of the NVLink domain store
 * their (score, id) records into.  The scan kernel of lsc_index_search_sharded_dev writes its shard's lists
 * straight into every peer's buffer (peer-to-peer stores), raises a flag per query, and merges the peers' lists
 * as their flags arrive - one kernel launch per pass over the corpus, no collective call.  world <= 8,
 * max_queries = largest slice of a batch exchanged at once (larger batches are cut into slices).
 *   - GPUs driven by ONE process: create one endpoint per GPU, then lsc_exchange_connect_local (or simply use
 *     lsc_sharded_index_*, which does all of this).
 *   - one process per GPU (one agent pod / torchrun rank per GPU): every rank creates its endpoint, publishes
 *     the LSC_EXCHANGE_HANDLE_BYTES from lsc_exchange_ipc_handle to the others over any channel they share and
 *     calls lsc_exchange_connect_ipc with all ranks' handles in rank order (its own slot is ignored).
 */
#define LSC_EXCHANGE_HANDLE_BYTES 64
int lsc_exchange_create(int device, int32_t rank, int32_t world, int32_t max_queries, lsc_exchange** out);
int lsc_exchange_ipc_handle(lsc_exchange* x, void* handle_out);
int lsc_exchange_connect_ipc(lsc_exchange* x, const void* all_handles);
int lsc_exchange_connect_local(lsc_exchange** endpoints, int32_t world);
void lsc_exchange_destroy(lsc_exchange* x);

/*
 * Search this GPU's shard AND merge with the other shards: every rank of the group calls it with the same
 * queries (device fp32 B x dim), B and k (<= 32), in the same order of calls; every rank receives the same
 * merged [B][k] ids / scores.  Equal scores come back in (rank, position inside the rank's list) order, which is
 * the storage order of one unsharded index when rank r holds the r-th contiguous range of rows.  Enqueued on
 * `stream`; does not synchronise.  A rank that never makes the matching call leaves its peers waiting: after 30 s
 * they give up and the next lsc_index_check / host-buffer search on that handle reports it.
 */
int lsc_index_search_sharded_dev(lsc_index* idx, lsc_exchange* x, const float* d_queries, int32_t B, int32_t k,
                                 int64_t* d_out_ids, float* d_out_scores, void* stream);
/*
 * Allocate now whatever scratch a later search of up to B queries and up to k results needs, so that the search
 * itself makes no allocation (a growing buffer costs a device-wide synchronisation, which a latency-bound caller
 * wants at start-up, and which - between shards that share ONE GPU - would stall the shard whose peers already
 * wait for it).  No reference counterpart (HerdDB sizes its own buffers).
 */
int lsc_index_reserve(lsc_index* idx, int32_t B, int32_t k);
/* Synchronises `stream` and reports (once) whether a wait inside an earlier search on this handle gave up. */
int lsc_index_check(lsc_index* idx, void* stream);

/*
 * The whole sharded table behind one handle, for a host that drives several GPUs from one process - the shape a
 * Java datasource needs: QueryStepDataSource.fetchData is called from one JVM thread
 * (AI/com/datastax/oss/streaming/ai/datasource/QueryStepDataSource.java:22-37).  The library keeps one worker
 * thread, stream, lsc_index shard and lsc_exchange endpoint per device; calls are forwarded to all of them and
 * return when every shard is done.  New ids are placed on the shard with the fewest rows; an existing id is
 * replaced where it lives.  Same semantics otherwise as lsc_index_upsert / _delete / _search (host buffers,
 * blocking); k <= 32.  `devices` may name the same GPU more than once (tests on a single GPU).
 * Replaces: the JDBC datasource + writer pair (JdbcDataSourceProvider.java:103-124,
 * VEC/jdbc/JdbcWriter.java:139-179) for a table that does not fit, or should not be scanned by, one GPU.
 */
int lsc_sharded_index_create(int32_t dim, int64_t capacity_per_shard, const int* devices, int32_t ndev,
                             lsc_sharded_index** out);
int lsc_sharded_index_upsert(lsc_sharded_index* s, const int64_t* ids, const float* vecs, int32_t n);
int lsc_sharded_index_delete(lsc_sharded_index* s, const int64_t* ids, int32_t n);
int64_t lsc_sharded_index_size(const lsc_sharded_index* s);
int lsc_sharded_index_search(lsc_sharded_index* s, const float* queries, int32_t B, int32_t k, int64_t* out_ids,
                             float* out_scores);
int32_t lsc_sharded_index_shards(const lsc_sharded_index* s);
/* The shard living on the i-th device of the handle (borrowed: do not destroy), e.g. for lsc_index_upsert_dev. */
lsc_index* lsc_sharded_index_shard(lsc_sharded_index* s, int32_t i);
void lsc_sharded_index_destroy(lsc_sharded_index* s);

/* Same measurement aid as lsc_encoder_set_profiling; classes: 0 normalise (queries / upserted rows),
 * 1 scan, 2 merge. */
int lsc_index_set_profiling(lsc_index* idx, int enabled);
int lsc_index_read_profile(lsc_index* idx, float* ms_by_class, int64_t* launches_by_class,
                           int32_t nclass);

/* Number of kernels this handle has launched since creation. */
int64_t lsc_index_launch_count(const lsc_index* idx);

/*
 * Persist / restore the index (ids + bf16 rows) — the local stand-in for the external database's
 * durability (the agent's persistent state directory, API/runner/code/AgentContext.java:64-66).
 */
int lsc_index_save(lsc_index* idx, const char* path);
int lsc_index_load(const char* path, int device, lsc_index** out);

/* Replaces: QueryStepDataSource.close() / VectorDatabaseWriter.close()
 * (AI/com/datastax/oss/streaming/ai/datasource/QueryStepDataSource.java:36,
 *  API/database/VectorDatabaseWriter.java:38). */
void lsc_index_destroy(lsc_index* idx);

/* ---- host tokenizer (feeds lsc_encoder_embed; no GPU work) -------------------------------- */

typedef struct lsc_tokenizer lsc_tokenizer;

/*
 * Replaces: the tokenizer half of DJL's TextEmbeddingTranslator, i.e. ai.djl.huggingface.tokenizers.
 * HuggingFaceTokenizer = the HuggingFace `tokenizers` BERT pipeline (BertNormalizer, BertPreTokenizer,
 * WordPiece, "[CLS] A [SEP]"), installed by TextEmbeddingTranslatorFactory at
 * AI/com/datastax/oss/streaming/ai/embeddings/AbstractHuggingFaceEmbeddingService.java:20,155 and run
 * inside predictor.predict (:194-196) one text at a time.
 * vocab_txt / nbytes: the contents of the model's vocab.txt (one token per line, id = line number; must
 * contain [UNK], [CLS], [SEP]).  do_lower_case also strips accents, as the library does when strip_accents
 * is unset.  max_length: sequences are truncated to it, specials included (the model's max positions).
 * n_threads <= 0: one per hardware thread.
 */
int lsc_tokenizer_create(const char* vocab_txt, size_t nbytes, int do_lower_case, int32_t max_length,
                         int32_t n_threads, lsc_tokenizer** out);

/*
 * Tokenise n_texts UTF-8 texts (text_nbytes[i] bytes each, or NUL-terminated when text_nbytes is NULL) on
 * the handle's threads and write them packed: token_ids[cu_seqlens[i] .. cu_seqlens[i+1]) are the ids of
 * text i, cu_seqlens has n_texts + 1 entries - the layout lsc_encoder_embed takes.  capacity = room in
 * token_ids (n_texts * max_length always suffices).  Thread-safe: the handle is read-only after creation.
 */
int lsc_tokenizer_encode(lsc_tokenizer* tok, const char* const* texts, const int32_t* text_nbytes,
                         int32_t n_texts, int32_t* token_ids, int64_t capacity, int32_t* cu_seqlens);

/*
 * Other template tokens than [CLS] A [SEP]: MPNet's tokenizer (all-mpnet-base-v2) is the same BERT pipeline with the
 * template "<s> A </s>" (tokenizer.json: TemplateProcessing) and [UNK] as the unknown token.  Each argument is a token
 * of the vocabulary, or NULL to keep the current one.  Call before the first lsc_tokenizer_encode.
 */
int lsc_tokenizer_set_special(lsc_tokenizer* tok, const char* cls_token, const char* sep_token, const char* unk_token);

int32_t lsc_tokenizer_vocab_size(const lsc_tokenizer* tok);
void lsc_tokenizer_destroy(lsc_tokenizer* tok);

/* ---- re-rank (the step after query-vector-db in the reference's chatbot pipeline) --------- */

/*
 * Replaces: ReRankAgent.rerankMMR and what it calls (langstream-agents/langstream-ai-agents/src/main/java/
 * ai/langstream/ai/agents/rerank/ReRankAgent.java:155-252 MMR loop + diversity, :254-273 and :361-372
 * cosine / norm / dot product, :275-343 BM25 and IDF), algorithm "MMR".
 * Picks up to max_docs of n_docs documents: each round takes the document with the highest
 * lambda * BM25(query, doc | documents still in the pool) - (1 - lambda) * mean cosine(doc, picked),
 * first maximum wins.  The host does the text side (ReRankAgent.tokenise = split on \s+, :314-316) and
 * passes, per document, its length in tokens (doc_len[n_docs]) and the frequency of each QUERY token
 * (term_freq[n_docs][n_terms], one column per query token in query order, duplicates included);
 * doc_vecs[n_docs][dim] are the document embeddings (fp32).  Arithmetic follows the reference's types and
 * evaluation order (fp32 sequential sums for cosine, fp64 for BM25 / diversity / score), so the order is
 * the reference's, not an approximation of it.  out_order[0 .. *out_count) receives indices into the input
 * list.  Limits: n_docs <= 128 (the reference's default `max` is 100), n_terms <= 64.
 * Documents must have non-empty text and embeddings (the reference filters those out and then indexes the
 * unfiltered list with the filtered position, :198-217 — not reproduced).
 * Host buffers; uses the current device; blocks until done.
 */
int lsc_rerank_mmr(const float* doc_vecs, int32_t n_docs, int32_t dim, const int32_t* term_freq,
                   const int32_t* doc_len, int32_t n_terms, int32_t max_docs, double lambda, double k1,
                   double b, int32_t* out_order, int32_t* out_count);

/* Same with device pointers (e.g. rows gathered from an index on the device); enqueues one kernel on
 * `stream` and returns.  *d_out_count < 0 reports the reference's "topDocument is null" condition. */
int lsc_rerank_mmr_dev(const float* d_doc_vecs, int32_t n_docs, int32_t dim, const int32_t* d_term_freq,
                       const int32_t* d_doc_len, int32_t n_terms, int32_t max_docs, double lambda, double k1,
                       double b, int32_t* d_out_order, int32_t* d_out_count, void* stream);

/* A batch of queries in ONE launch (one CTA per query): query q brings n_docs documents of its own -
 * d_doc_vecs[n_queries][n_docs][dim], d_term_freq[n_queries][n_docs][n_terms], d_doc_len[n_queries][n_docs] - and
 * receives d_out_order[q * order_stride .. + count) and d_out_count[q * count_stride].  What a query-vector-db step
 * that answers a batch of records hands to the re-rank step; same arithmetic as lsc_rerank_mmr_dev per query. */
int lsc_rerank_mmr_batch_dev(const float* d_doc_vecs, int32_t n_queries, int32_t n_docs, int32_t dim,
                             const int32_t* d_term_freq, const int32_t* d_doc_len, int32_t n_terms, int32_t max_docs,
                             double lambda, double k1, double b, int32_t* d_out_order, int32_t order_stride,
                             int32_t* d_out_count, int32_t count_stride, void* stream);

/*
 * The text side of the re-rank step, natively (host code): ReRankAgent.tokenise (ReRankAgent.java:314-316, Java's
 * split("\\s+"): a leading empty token is kept, trailing ones are dropped) on the query and on every document, then per
 * document its token count (doc_len[n_docs]) and the frequency of every query token (term_freq[n_docs][max_terms],
 * one column per query token in query order, duplicates included, zero-padded) - exactly the inputs of lsc_rerank_mmr.
 * *n_terms receives the number of query tokens; more than max_terms is LSC_E_UNSUPPORTED.  doc_nbytes may be NULL
 * (NUL-terminated texts), query_nbytes < 0 likewise.
 */
int lsc_rerank_term_statistics(const char* const* doc_texts, const int32_t* doc_nbytes, int32_t n_docs,
                               const char* query_text, int32_t query_nbytes, int32_t max_terms, int32_t* term_freq,
                               int32_t* doc_len, int32_t* n_terms);

/* ---- debug / verification hooks (not part of the product path) --------------------------- */

/*
 * C[M,N] (bf16) = epilogue(A[M,K] (bf16) · W[N,K]^T (bf16) + bias[N] (fp32)), device pointers.
 * epilogue: 0 = bias, 1 = bias + exact-erf GELU, 2 = bias + residual[M,N] (bf16).
 * impl: 0 = tcgen05 kernel, 1 = CUDA-core validation kernel.  Lets the tests check the GEMM alone.
 */
int lsc_debug_gemm_dev(const void* d_A, const void* d_W, const float* d_bias, const void* d_residual,
                       void* d_C, int32_t M, int32_t N, int32_t K, int32_t epilogue, int32_t impl,
                       void* stream);

/*
 * The packed variable-length attention kernel alone (debug / verification): d_qkv is [T, 3*hidden]
 * 16-bit (Q | K | V along the feature axis, heads of hidden/heads features each; Q and K bf16, V fp16
 * — the layout the QKV GEMM writes), d_ctx receives [T, hidden] bf16.  Lets the tests check it against an fp32 reference.
 */
int lsc_debug_attention_dev(const void* d_qkv, const int32_t* d_cu_seqlens, int32_t n_seq,
                            int32_t total_tokens, int32_t hidden, int32_t heads, int32_t max_seqlen,
                            void* d_ctx, void* stream);

/*
 * The attention tile list of a ragged batch (debug / verification): runs the kernel that composes 128-row attention
 * tiles out of 32-row blocks of sequences (csrc/attention_tc_blocks.cu) and leaves in d_out (8 + 9 * n_seq int32)
 * the tile count at [0] and from [8] one record of 8 ints per tile: the first global row of each of its four blocks
 * (-1: unused) and, per block, first key block | key blocks << 4 | sequence length << 8 | block within the sequence
 * << 16 | used blocks << 20.  Lets the tests check the packing on its own.
 */
int lsc_debug_attention_blocks_dev(const int32_t* d_cu_seqlens, int32_t n_seq, int32_t* d_out, void* stream);

/*
 * The relative-position bias table an MPNet encoder hands to its attention kernel (debug / verification, host code):
 * out[heads][384] fp32, entry 191 + r = weight[bucket(r)][h] * log2(e) for r = key position - query position in
 * [-191, 191] (bucket = transformers' MPNetEncoder.relative_position_bucket), weight = [rel_buckets][heads] fp32.
 */
int lsc_debug_mpnet_bias_table(const float* weight, int32_t rel_buckets, int32_t rel_max_distance, int32_t heads, float* out);

#ifdef __cplusplus
}
#endif
#endif /* LANGSTREAM_CUDA_H */
