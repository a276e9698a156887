/*
 * dab_b200.h - C-ABI of the B200-native retrieval hot path for django-assistant-bot.
 *
 * This is the drop-in boundary: plain C types, device pointers and sizes, no torch
 * types.  Each entry point names the reference interface it stands behind
 * (paths relative to the reference repository root).  Shipped as libdab_b200.so
 * (CUDA, sm_100a only).  There is no CPU fallback: every call needs a CUDA device.
 *
 * Conventions
 *   - return 0 on success, a DAB_E_* code otherwise; dab_last_error() gives the
 *     message for the calling thread.  Nothing aborts, nothing falls back to the CPU.
 *   - all *_dev pointers are device pointers owned by the caller, 16-byte aligned,
 *     contiguous (row-major).  `stream` is a cudaStream_t passed as void*; every call
 *     is asynchronous with respect to it unless stated otherwise.
 *   - the library owns only what dab_store_create allocates (rows, inverse norms, ids,
 *     per-store scratch) and frees it in dab_store_destroy.
 *   - thread safety: calls on different stores are independent; calls on one store are
 *     serialised by a per-store mutex and every search keeps its own verdict (the reference reaches this code from an asgiref
 *     worker thread, assistant/rag/services/search_service.py:191, and from uvicorn
 *     workers, gpu_service/gunicorn_conf.py:9).
 */
#ifndef DAB_B200_H
#define DAB_B200_H

#include <stdint.h>

#if defined(__GNUC__)
#define DAB_API __attribute__((visibility("default")))
#else
#define DAB_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dab_store dab_store;

/* element types */
#define DAB_F32 0
#define DAB_BF16 1
/* attention-mask element types (HF tokenizers emit int64) */
#define DAB_MASK_U8 0
#define DAB_MASK_I32 1
#define DAB_MASK_I64 2

/* error codes */
#define DAB_OK 0
#define DAB_E_INVALID 1     /* bad argument (null, misaligned, unsupported size) */
#define DAB_E_CUDA 2        /* a CUDA runtime call failed */
#define DAB_E_DIM 3         /* "different vector dimensions", like pgvector raises */
#define DAB_E_CAPACITY 4    /* store full / size beyond a stated maximum */
#define DAB_E_INCOMPLETE 5  /* a search did not finish its scan rounds (see dab_store_search_status) */
#define DAB_E_INEXACT 6     /* tensor path: for some queries more rows lie inside the rescoring margin than it keeps;
                             * exactly those queries come back poisoned (score NaN, id -1), every other query of the
                             * call is final; repeat the poisoned ones with dab_store_set_path(s, 1) */

/* queryset filters evaluated on the device (dab_store_set_column / dab_store_build_allow) */
#define DAB_MAX_COLUMNS 8        /* int64 filter columns per store */
#define DAB_MAX_PREDICATES 8     /* predicates AND-ed in one dab_store_build_allow call */
#define DAB_COL_ID (-1)          /* pseudo-column: the row id (Django pk) */
#define DAB_OP_EQ 0
#define DAB_OP_NE 1
#define DAB_OP_LT 2
#define DAB_OP_LE 3
#define DAB_OP_GT 4
#define DAB_OP_GE 5
#define DAB_OP_IN 6              /* value in set_dev[0..set_n), an ASCENDING device array */
#define DAB_OP_NOT_IN 7

/* limits */
#define DAB_MAX_K 2048           /* largest n of a search; the reference uses 1, 5, 10, 250 */
#define DAB_MAX_QUERY_BATCH 256  /* queries scored per corpus pass; larger Q is chunked */
#define DAB_MAX_ROW_BYTES 8192   /* dim*sizeof(elem): 2048 fp32 / 4096 bf16 */

/* Library version string, e.g. "dab_b200 0.1 sm_100a". */
DAB_API const char *dab_version(void);

/* Message of the last failing call on this thread (never NULL). */
DAB_API const char *dab_last_error(void);

/*
 * Fused masked mean-pool + L2 normalise of encoder hidden states.
 * Replaces: `outputs.last_hidden_state.mean(dim=1).squeeze().tolist()`
 *           assistant/ai/embedders/transformers.py:25 (inside TransformersEmbedder.embeddings, :15-29),
 *           generalised to a padded batch: e[b] = sum_t m[b,t] h[b,t,:] / max(sum_t m[b,t], 1),
 *           out[b] = e[b] / max(|e[b]|, 1e-12).
 *   hidden        [B,T,D] hidden_dtype (DAB_F32 | DAB_BF16)
 *   mask          [B,T]   mask_dtype, non-zero = token counts; NULL = all tokens count
 *   out           [B,D]   fp32, unit-norm rows (all-masked row -> zeros)
 *   out_bf16      [B,D]   bf16 copy of `out` (RNE), or NULL
 *   out_invnorm   [B]     the factor applied, 1/max(|e|,1e-12): out/out_invnorm is the
 *                         reference's un-normalised mean; or NULL
 * D must be a multiple of 4 (fp32) / 8 (bf16).
 */
DAB_API int dab_pool_l2norm(const void *hidden_dev, int hidden_dtype, const void *mask_dev, int mask_dtype,
                    float *out_dev, void *out_bf16_dev_or_null, float *out_invnorm_dev_or_null,
                    int64_t B, int64_t T, int64_t D, void *stream);

/*
 * Device-resident chunk-embedding store: one shard of the table behind
 * assistant/storage/models.py:19-29 (BaseEmbeddingModel.embedding, vector(768) fp32, nullable)
 * and :13 (Document.content_embedding).  Only non-NULL embeddings are stored.
 *   dtype: DAB_F32 (what pgvector stores) or DAB_BF16.  Synchronous (allocates).
 */
/* capacity_rows: 1 .. 2^31-1 */
DAB_API int dab_store_create(dab_store **out, int device, int64_t capacity_rows, int dim, int dtype);
DAB_API void dab_store_destroy(dab_store *s);

/* rows currently held / capacity / dim / dtype */
DAB_API int64_t dab_store_size(const dab_store *s);
DAB_API int64_t dab_store_capacity(const dab_store *s);
DAB_API int dab_store_dim(const dab_store *s);
DAB_API int dab_store_dtype(const dab_store *s);
/* forget all rows (keeps the allocation) */
DAB_API int dab_store_reset(dab_store *s);

/*
 * Append n rows.  Replaces the write side of the embedding column:
 * `Sentence.objects.bulk_update(sentences, ['embedding'])`
 * assistant/processing/documents/steps/embeddings.py:35-39,65-69.
 *   rows_dev  [n,dim] rows_dtype (fp32 rows are rounded RNE into a bf16 store)
 *   ids_dev   [n] int64 row ids (Django pks) or NULL = running row index (+ id_base)
 * Also computes and keeps 1/|row| from the STORED values (0 for an all-zero row, which
 * is then never returned - pgvector gives NaN distance for it and PostgreSQL sorts NaN last).
 * Ties between equal scores are broken by store row order, i.e. by id when ids ascend
 * with insertion order.
 */
DAB_API int dab_store_append(dab_store *s, const void *rows_dev, int rows_dtype, const int64_t *ids_dev_or_null,
                     int64_t n, void *stream);
/*
 * Ingest: pool + normalise encoder hidden states and append the rows to the store in ONE kernel - the write path
 * `embedder.embeddings(texts)` -> `bulk_update(['embedding'])` of
 * assistant/processing/documents/steps/embeddings.py:27-39,57-69 without a round trip through host memory or a second
 * pass over the rows: the pool kernel's outputs are the store's own rows [size, size+B) (bf16 or fp32), their 1/|row|
 * (from the stored values) and their ids (id_base + row; stores with explicit ids are refused).
 *   hidden [B,T,dim] hidden_dtype, mask [B,T] mask_dtype or NULL: as dab_pool_l2norm.
 */
DAB_API int dab_store_ingest(dab_store *s, const void *hidden_dev, int hidden_dtype, const void *mask_dev, int mask_dtype,
                             int64_t B, int64_t T, void *stream);
/*
 * Delete rows (tombstones): the rows stay in place but are never returned again - their stored
 * inverse norm becomes -0.0, which every scan treats like a zero vector.  Replaces the deletes of
 * document finalisation (older documents of a wiki removed, assistant/processing/tasks.py:67-74).
 *   rows_dev [n] int64 STORE ROW indices (not ids); out-of-range entries are ignored.
 */
DAB_API int dab_store_delete_rows(dab_store *s, const int64_t *rows_dev, int64_t n, void *stream);
/*
 * Physically remove the tombstones left by dab_store_delete_rows: surviving rows keep their order and
 * move down, their ids travel with them (explicit ids) or stay "id_base + row" (implicit ids, i.e. the
 * rows are renumbered - a caller that keeps host columns aligned with store rows compacts those with the
 * same survivors).  Genuine zero vectors are kept.  What VACUUM does for the table after the deletes of
 * document finalisation (assistant/processing/tasks.py:67-74); without it every scan keeps streaming the
 * dead rows.  Synchronous: waits for searches already enqueued on this store, allocates temporaries
 * (8 bytes per row + <= 256 MiB of staging), returns when the store has its new size.
 *   removed_out  number of rows dropped (may be NULL)
 */
DAB_API int dab_store_compact(dab_store *s, int64_t *removed_out, void *stream);
/*
 * Copy rows [row0, row0+n) out of the store, in the store's dtype, with their inverse norms
 * (0 = deleted / zero row) and ids; any output may be NULL.  For snapshots of a shard.
 */
DAB_API int dab_store_export(dab_store *s, int64_t row0, int64_t n, void *rows_out_dev, float *inv_norm_out_dev_or_null,
                             int64_t *ids_out_dev_or_null, void *stream);
/* id given to row i when ids are implicit: id_base + i (a shard's first global row index) */
DAB_API int dab_store_set_id_base(dab_store *s, int64_t id_base);

/*
 * Exact cosine top-k of Q queries over the store.
 * Replaces: `qs.annotate(distance=CosineDistance(field, q)).order_by('distance')[:n]`
 *           assistant/rag/services/search_service.py:185-196 (_objects_embedding_search),
 *           i.e. pgvector `<=>` + ORDER BY ... LIMIT n inside PostgreSQL.
 *   q_dev        [Q,dim] fp32 queries (any norm)
 *   k            n of the reference call, 1..DAB_MAX_K
 *   allow_bitmap [ceil(size/32)] uint32, bit r set = row r passes the queryset filter; NULL = all
 *   out_scores   [Q,k] fp32 cosine similarity, descending (distance = 1 - score)
 *   out_ids      [Q,k] int64 row ids; (score desc, row order asc) ; unused slots: id -1, score -inf
 * Asynchronous and sync-free: the scan rounds are enqueued optimistically.  If a round
 * overflowed its candidate buffer more often than the spare rounds allow, the outputs are
 * filled with id -1 / NaN and dab_store_search_status reports DAB_E_INCOMPLETE.
 */
DAB_API int dab_store_search(dab_store *s, const float *q_dev, int Q, int k, const uint32_t *allow_bitmap_dev_or_null,
                     float *out_scores_dev, int64_t *out_ids_dev, void *stream);
/*
 * dab_store_search with the scoring path and the number of spare rounds given for THIS call (-1 = the store's
 * setting), instead of through dab_store_set_path / dab_store_set_spare_rounds, which change the store for
 * every thread.  What a checked wrapper uses to repeat a search after DAB_E_INEXACT (path 1) or
 * DAB_E_INCOMPLETE (spare rounds) while other threads keep searching the same store.
 */
DAB_API int dab_store_search_ex(dab_store *s, const float *q_dev, int Q, int k, const uint32_t *allow_bitmap_dev_or_null,
                                float *out_scores_dev, int64_t *out_ids_dev, int path, int spare_rounds, void *stream);
/*
 * The same search with HOST buffers, complete and checked: what the reference-facing call uses
 * (`_objects_embedding_search` receives the query as a Python list, search_service.py:186).
 * Copies the queries to the device (pinned staging owned by the store, allocated on first use),
 * searches, copies scores/ids back, synchronises `stream`, and repeats internally if the device
 * reported DAB_E_INCOMPLETE (spare rounds) or DAB_E_INEXACT (the poisoned queries alone, on the CUDA-core path).  Synchronous.
 */
DAB_API int dab_store_search_host(dab_store *s, const float *q_host, int Q, int k, const uint32_t *allow_bitmap_dev_or_null,
                                  float *out_scores_host, int64_t *out_ids_host, void *stream);
/* Synchronises `stream`; DAB_OK, DAB_E_INCOMPLETE or DAB_E_INEXACT for the last search call THE CALLING THREAD
 * made on `s` (all of its query chunks).  Verdicts are kept per call (a ring of 256 per store, written by the
 * search's last kernel into pinned memory), so searches other threads enqueue on the same store in between do
 * not disturb it. */
DAB_API int dab_store_search_status(dab_store *s, void *stream);
/* extra scan rounds enqueued per search for overflow retries (default 0) */
DAB_API int dab_store_set_spare_rounds(dab_store *s, int rounds);
/* scoring path: 0 = auto, 1 = CUDA-core fp32 dot products, 2 = tcgen05 tensor tiles (bf16 stores: kind::f16,
 * fp32 stores: kind::tf32; any row length the store accepts) + exact fp32 rescoring.  Auto takes the tensor
 * path whenever the row layout allows it, for any number of queries; after a DAB_E_INEXACT verdict on a search
 * of <= 4 queries the next 64 such searches use the CUDA cores directly. */
DAB_API int dab_store_set_path(dab_store *s, int path);

/*
 * Queryset filters on the device.  Every reference call searches a FILTERED queryset:
 *   Question.objects.filter(document__wiki__bot=bot, document__processing__status=COMPLETED)
 *       assistant/bot/services/context_service/steps/embeddings.py:26-29
 *   Question.objects.filter(document__id__lt=document.id, ...)
 *       assistant/processing/documents/steps/questions.py:124
 * PostgreSQL evaluates those WHERE clauses per row inside the scan; here the columns they test live in HBM
 * next to the rows (one int64 per row and column: document id, bot id, status code, ...) and ONE kernel turns a
 * conjunction of predicates into the allow bitmap dab_store_search takes - nothing per row happens on the host.
 *   dab_store_set_column  values_dev [n] int64 for store rows [row0, row0+n); columns are allocated on first use
 *                         (capacity * 8 bytes each), follow dab_store_compact, and read 0 where never set
 *   dab_store_build_allow bitmap_out_dev [ceil(size/32)] uint32: bit r = row r is alive (not a tombstone) AND
 *                         passes all n_pred predicates (n_pred = 0: alive rows).  A queryset given as a list of
 *                         pks is the predicate {DAB_COL_ID, DAB_OP_IN, sorted pks}.
 */
typedef struct dab_predicate {
    int32_t column;          /* 0..DAB_MAX_COLUMNS-1 or DAB_COL_ID */
    int32_t op;              /* DAB_OP_* */
    int64_t value;           /* operand of the comparison ops */
    const int64_t *set_dev;  /* DAB_OP_IN / DAB_OP_NOT_IN: ascending int64 device array */
    int64_t set_n;
} dab_predicate;
DAB_API int dab_store_set_column(dab_store *s, int column, const int64_t *values_dev, int64_t row0, int64_t n, void *stream);
DAB_API int dab_store_build_allow(dab_store *s, const dab_predicate *preds, int n_pred, uint32_t *bitmap_out_dev, void *stream);
/* The allow bitmap of a queryset given as a LIST of store rows (a Django queryset arrives as its pks; the caller maps
 * pk -> row): the bitmap is cleared and one bit is set per listed row - O(list), not O(rows), for the reference's
 * usual small result sets.  Tombstoned and out-of-range rows are skipped; duplicates are harmless. */
DAB_API int dab_store_allow_from_rows(dab_store *s, const int64_t *rows_dev, int64_t n, uint32_t *bitmap_out_dev, void *stream);

/*
 * k-way merge of per-shard top-k lists (after the all-gather).  New in the build: the
 * reference has a single PostgreSQL table; the sharded store needs the global ORDER BY.
 *   gathered_scores [R,Q,k] fp32, gathered_ids [R,Q,k] int64 (id -1 = empty slot)
 *   out_scores [Q,k], out_ids [Q,k]: (score desc, id asc).  R*k <= 8192.
 */
DAB_API int dab_merge_topk(const float *gathered_scores_dev, const int64_t *gathered_ids_dev, int R, int Q, int k,
                   float *out_scores_dev, int64_t *out_ids_dev, void *stream);

/*
 * Document-level aggregation of the hits of a search, on the device.
 * Replaces: the grouping loop of `embedding_search`,
 *           assistant/rag/services/search_service.py:133-150 - hits grouped by document_id in rank
 *           order, documents with fewer than max_scores_n hits dropped, score
 *           1 - sum(first max_scores_n distances)/max_scores_n, best top_n (ties: first appearance).
 *   scores_dev/rows_dev  [Q,n] output of dab_store_search on a store with implicit ids (id = row)
 *   doc_of_row_dev       [n_rows] int64 document id of every store row
 *   out_doc_ids_dev      [Q,top_n] int64 (-1 = fewer documents qualified), out_doc_scores_dev [Q,top_n] fp32
 * n <= 4096.
 */
DAB_API int dab_doc_aggregate(const float *scores_dev, const int64_t *rows_dev, const int64_t *doc_of_row_dev,
                              int64_t n_rows, int Q, int n, int max_scores_n, int top_n, int64_t *out_doc_ids_dev,
                              float *out_doc_scores_dev, void *stream);

/*
 * The exchange step of the sharded search as ONE kernel over NVLink peer memory (no NCCL call):
 * publish "my local top-k is ready" to every peer, wait for all peers, pull their lists with
 * peer loads and merge - all ranks launch it once per search (SPMD).  New in the build, like
 * dab_merge_topk.  Every rank owns one staging block of dab_exchange_bytes(R, max_q, max_k) bytes,
 * zero-initialised, mapped into every peer (e.g. torch symmetric memory); peer_bases_dev is a
 * DEVICE array of the R mapped base addresses, in rank order.
 *   dab_exchange_slot     where search number `epoch` (1, 2, 3, ...) must write this rank's lists:
 *                         pass these as out_scores/out_ids of dab_store_search
 *   dab_store_status_dev  device address of the store's {incomplete, inexact} verdict (2 ints)
 *   dab_exchange_merge_topk  out_scores/out_ids [Q,k] merged (score desc, id asc); out_status (2 ints,
 *                         device, may be NULL) = OR over all ranks of the verdicts, so that every rank
 *                         takes the same retry decision.  R*k <= 8192.  A peer that never
 *                         publishes makes the kernel trap after a bounded spin instead of hanging.
 */
DAB_API int64_t dab_exchange_bytes(int R, int max_q, int max_k);
DAB_API int dab_exchange_slot(void *base, int R, int max_q, int max_k, uint32_t epoch, float **scores, int64_t **ids);
DAB_API const int *dab_store_status_dev(dab_store *s);   /* verdict slot of the calling thread's last search on s */
DAB_API int dab_exchange_merge_topk(void *const *peer_bases_dev, int rank, int R, int max_q, int max_k, int Q, int k,
                                    uint32_t epoch, const int *local_status_dev_or_null, float *out_scores_dev,
                                    int64_t *out_ids_dev, int *out_status_dev_or_null, void *stream);

/*
 * The row-sharded store behind ONE process: what a Django / Celery / FastAPI worker can call without a launcher
 * (the reference searches from an asgiref worker thread, assistant/rag/services/search_service.py:191, and serves
 * embeddings from gunicorn workers, gpu_service/gunicorn_conf.py:9).  New in the build, like dab_merge_topk: the
 * reference has one PostgreSQL table.  Shard g lives on devices[g] and holds rows [g*N/G, (g+1)*N/G), ids = global
 * row index; the handle owns a stream, a peer-mapped staging block (cudaDeviceEnablePeerAccess) and the query /
 * result staging per device.  A device may be listed more than once (several shards on one GPU; used by tests).
 *   dab_sharded_shard        the dab_store of shard g (append / set_column / delete through the store API, with
 *                            device_out current and stream_out or any stream of that device); any output may be NULL
 *   dab_sharded_search_host  host queries in, merged host results out: every shard scans on its own device and
 *                            stream, dab_exchange_merge_topk runs on every device, shard 0's merged lists and the
 *                            OR-ed verdict return in one copy; inexact / incomplete verdicts are retried inside
 *                            (exact path / spare rounds).  Synchronous; one search at a time per handle.
 */
typedef struct dab_sharded dab_sharded;
DAB_API int dab_sharded_create(dab_sharded **out, const int *devices, int n_devices, int64_t total_rows, int dim, int dtype,
                               int max_q, int max_k);
DAB_API void dab_sharded_destroy(dab_sharded *sh);
DAB_API int dab_sharded_num_shards(const dab_sharded *sh);
DAB_API dab_store *dab_sharded_shard(dab_sharded *sh, int g, int *device_out, int64_t *row_lo_out, int64_t *row_hi_out,
                                     void **stream_out);
DAB_API int dab_sharded_search_host(dab_sharded *sh, const float *q_host, int Q, int k, float *out_scores_host,
                                    int64_t *out_ids_host);

/*
 * Exact cosine top-k over rows the CALLER owns, without allocating anything.
 * No counterpart in the reference (there the rows live in PostgreSQL and the executor
 * allocates as it pleases); it exists for hosts of this library that cannot call cudaMalloc:
 * code running inside the embedding service after the encoder's allocator has reserved the
 * GPU, request handlers that cannot afford cudaMalloc's device-wide synchronisation, and
 * callers ranking embeddings they keep in their own buffers.  Same result as
 * dab_store_append + dab_store_search on the same rows (CUDA-core fp32 path).
 *   rows_dev      [n_rows, dim] DAB_F32 or DAB_BF16, 16-byte aligned
 *   inv_norm_dev  [n_rows] fp32, 1/|row| (0 = never return this row); must be readable up to
 *                 n_rows rounded up to a multiple of 4
 *   id_base       ids written are id_base + row index
 *   q_dev, k, allow_bitmap, out_scores_dev, out_ids_dev: as dab_store_search
 * The scratch a search needs (unit queries, 8192 candidate keys per query, thresholds, round
 * control) lives in one static __device__ array of DAB_NOALLOC_SCRATCH_BYTES that every GPU
 * loading the library holds until exit.  Hence the limits: Q <= DAB_NOALLOC_MAX_QUERIES,
 * dim <= DAB_NOALLOC_MAX_DIM, else DAB_E_CAPACITY.  All calls on a device share that array:
 * each call resets the state it reads, but calls that overlap on the GPU (two streams) race -
 * keep them on one stream or order them with events.  One call is a sequence of kernel launches;
 * the library holds a per-device host mutex while it enqueues them, so host threads that share a
 * stream cannot interleave two calls' launches - threads on different streams must still order
 * those streams themselves (the mutex does not wait for the GPU).  The first call on a device resolves the
 * array's address (cudaGetSymbolAddress): issue it before any stream capture.  Asynchronous;
 * makes no cudaMalloc/cudaFree.  DAB_NOALLOC_SPARE_ROUNDS retry rounds are always enqueued; a
 * search that still overflowed writes id -1 / NaN (check the first id).
 */
#define DAB_NOALLOC_MAX_QUERIES 8
#define DAB_NOALLOC_MAX_DIM 1024
#define DAB_NOALLOC_SCRATCH_BYTES (560 * 1024)
#define DAB_NOALLOC_SPARE_ROUNDS 2
DAB_API int dab_search_noalloc(const void *rows_dev, int rows_dtype, const float *inv_norm_dev, int64_t n_rows, int dim,
                               int64_t id_base, const float *q_dev, int Q, int k,
                               const uint32_t *allow_bitmap_dev_or_null, float *out_scores_dev, int64_t *out_ids_dev,
                               void *stream);

/*
 * Measurement hooks (no counterpart in the reference, which only has the wall-clock
 * TimeDebugger of assistant/utils/debug.py:5-30 around the whole search).
 *   dab_launch_count      kernels this library has launched in this process
 *   dab_store_set_profile on: every scan launch and every search call is bracketed by
 *                         cudaEvents on the caller's stream
 *   dab_store_get_profile synchronises the device, returns the summed event times since the
 *                         last call and clears them
 *   dab_store_describe_plan  what a search of Q queries / top-k over the store's current rows would do:
 *                         path (1 CUDA cores, 2 tcgen05, 3 tcgen05 4-CTA cluster kernel), rows per tile,
 *                         number of sampling levels, growth factor and the tile stride of each level
 *                         (strides[8], 0 past the last level).  Any output pointer may be NULL.
 */
DAB_API uint64_t dab_launch_count(void);
DAB_API int dab_store_describe_plan(dab_store *s, int Q, int k, int *path, int *tile_rows, int *num_levels, int *growth,
                                    int *strides);
DAB_API int dab_store_set_profile(dab_store *s, int on);
DAB_API int dab_store_get_profile(dab_store *s, float *scan_ms, int *scan_launches, float *search_ms, int *searches);

#ifdef __cplusplus
}
#endif
#endif /* DAB_B200_H */
