/* ctr_engine.h -- C ABI of the B200-native CTR scoring engine.
 *
 * This is the drop-in boundary for ONE path of neuzxy/Distributed-TF-Serving: the client's
 * "split candidates -> marshal -> score on N model replicas -> ordered merge" request
 * (DCNClient.java:137-203).  Everything below replaces a piece of that path; each entry
 * point cites the reference interface it stands in for.  DC = src/main/java/tensorflow/
 * serving/DCNClient.java under the reference root.
 *
 * Conventions
 *   - plain C types only (pointers, sizes, ints); no CUDA or torch types in any signature
 *     (streams and device pointers cross as void*)
 *   - every function returns CTR_OK (0) or a negative ctr_status; ctr_last_error() gives
 *     the message of the last failure on the calling thread
 *   - all entry points are thread-safe; the reference calls its path from 6 driver threads
 *     fanning into a 16-thread pool (DC:31, DC:42, DC:213-224)
 *   - layout is the reference's: ids int64 / wts float32, row-major [n_cand, F]
 *     (DC:76-81, DC:97-108); scores float32 [n_cand] in input order (DC:161-164)
 *   - there is no CPU fallback: without a usable CUDA device every call fails with
 *     CTR_ERR_CUDA
 */
#ifndef CTR_ENGINE_H_
#define CTR_ENGINE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CTR_ABI_VERSION 1
#define CTR_MAX_DEVICES 8
#define CTR_MAX_HIDDEN 4
#define CTR_MAX_FIELDS 48

typedef enum ctr_status {
    CTR_OK = 0,
    CTR_ERR_INVALID = -1,   /* bad argument / shape (the RPC's INVALID_ARGUMENT) */
    CTR_ERR_CUDA = -2,      /* CUDA runtime / driver failure, message has the detail */
    CTR_ERR_IO = -3,        /* weight file unreadable or malformed */
    CTR_ERR_STATE = -4,     /* weights not loaded, engine shut down, unknown ticket */
    CTR_ERR_TIMEOUT = -5,   /* ctr_wait ran out of time; the ticket stays valid */
    CTR_ERR_CAPACITY = -6   /* request larger than the configured capacity */
} ctr_status;

/* Model family served behind sendRequest.  The reference addresses its model by the
 * strings ("DCN", "serving_default") (DC:33-34, DC:83-88); here the family is an enum and
 * there is no string lookup on the hot path. */
typedef enum ctr_model_kind {
    CTR_MODEL_LR = 0, CTR_MODEL_WIDE_DEEP = 1, CTR_MODEL_DEEPFM = 2, CTR_MODEL_DCN_V2 = 3
} ctr_model_kind;

typedef enum ctr_id_mode { CTR_ID_DIRECT = 0, CTR_ID_HASHED = 1 } ctr_id_mode;
typedef enum ctr_table_dtype { CTR_TABLE_FP32 = 0, CTR_TABLE_F16 = 1 } ctr_table_dtype;
typedef enum ctr_shard_mode {
    CTR_TABLES_REPLICATED = 0,  /* candidates sharded over devices, tables on every device */
    CTR_TABLES_ROW_SHARDED = 1  /* rows sharded (row mod n_devices); peers read over NVLink */
} ctr_shard_mode;

/* Replaces the reference's compile-time constants (DC:25-42): FIELD_NUM, candidateNum,
 * hostsList (-> device_ids), the pool size. */
typedef struct ctr_config {
    int32_t struct_size;          /* sizeof(ctr_config), for ABI checking */
    int32_t n_devices;            /* number of GPUs this engine fans out to (hostsList.size(), DC:38) */
    int32_t device_ids[CTR_MAX_DEVICES];
    int32_t max_candidates;       /* capacity of ONE shard (candidateNum / hosts, DC:29): a request of n candidates cut P
                                   * ways is accepted when its last shard, n - (P-1)*floor(n/P) (it takes the remainder,
                                   * DC:46-55), fits; with P = n_devices that holds for every
                                   * n <= n_devices * (max_candidates - n_devices + 1) */
    int32_t workers_per_device;   /* host threads (each with its own stream) per device; 0 = default 2 */
    int32_t shard_mode;           /* ctr_shard_mode */
    int32_t min_shard_candidates; /* a request is spread over floor(n / this) GPUs, at most n_devices, at least 1
                                   * (the starting GPU rotates); 0 = default 1024; 1 = always cut n_devices ways
                                   * as DC:148 does.  Scores do not depend on it. */
    int32_t host_pack;            /* the drop-in request path (ctr_predict_async & co. with the reference's int64 id + float
                                   * weight lists) can narrow a shard on its worker thread before it crosses the host link:
                                   * uint32 rows + weights only for the fields that ever had one != 1.0 (the packed layout of
                                   * ctr_predict_packed_async, 208 instead of 468 bytes per candidate for the Criteo shape;
                                   * direct-id tables only; same scores, bit for bit).  It trades host CPU (~0.45 ms of one core
                                   * per 10 000 candidates) for link capacity: 0 = when 5 or more shards are in flight on the GPU
                                   * and workers_per_device >= 8 (fewer shards share the link faster than a core narrows, and
                                   * the cores have to be there: budget ~10 per GPU, else set 1),
                                   * 1 = never, 2 = always */
    int32_t reserved[6];
} ctr_config;

typedef struct ctr_engine ctr_engine;
typedef int64_t ctr_ticket;

/* Completion callback (optional): runs on an engine thread when the request finishes,
 * status is CTR_OK or the failure.  The Java shim completes its CompletableFuture here
 * (normally or exceptionally), mirroring DC:151-159. */
typedef void (*ctr_callback)(void* user, ctr_ticket ticket, int32_t status);

/* ---- lifecycle: replaces getChannels()/shutDownChannels() (DC:118-135) ---- */
int32_t ctr_engine_create(const ctr_config* cfg, ctr_engine** out);
int32_t ctr_engine_destroy(ctr_engine* e);

/* Loads a flat weight file (format: DESIGN.md "Weight file") on every device of the
 * engine.  Stands in for TF-Serving loading the SavedModel named at DC:33-34.  The model
 * family, field count, layer sizes and table description come from the file header. */
int32_t ctr_engine_load_weights(ctr_engine* e, const char* path);

/* Model description after load (F is what the reference calls FIELD_NUM, DC:25). */
typedef struct ctr_model_info {
    int32_t model_kind, num_fields, emb_dim, id_mode, table_dtype, n_hidden;
    int32_t hidden[CTR_MAX_HIDDEN];
    int32_t n_cross, d_padded;
    int64_t num_rows;
    int64_t table_bytes_per_device;
    int32_t chunk_rows;           /* candidates the tower kernels process per launch */
    int32_t reserved;
} ctr_model_info;
int32_t ctr_engine_model_info(ctr_engine* e, ctr_model_info* out);

/* ---- the request: replaces processARequest (DC:137-203) minus the sort ----
 * Splits n_cand candidates over P of the engine's devices (P from ctr_config.
 * min_shard_candidates, at most n_devices) by the reference rule (range = n_cand / P,
 * remainder to the last shard; DC:46-55 applied to candidates), scores every shard on its device (sendRequest, DC:91-115) and writes
 * scores_out[j] for candidate j (ordered merge, DC:161-164).  Does not block on the GPU:
 * returns a ticket at once.  ids / wts / scores_out stay owned by the caller and must
 * stay valid until the ticket completes; ids/wts may be shared read-only between
 * concurrent requests (as the reference shares its subList views, DC:219).  Buffers from
 * ctr_alloc_pinned are used in place (no staging copy). */
int32_t ctr_predict_async(ctr_engine* e, const int64_t* ids, const float* wts, int64_t n_cand,
                          float* scores_out, ctr_callback cb, void* user, ctr_ticket* ticket);

/* Shared-context request (additive).  The reference ships every user /
 * context field again for every candidate (DC:61-63, DC:70-72).  Here the first n_ctx_fields of
 * the F fields are given once (ctx_ids / ctx_wts, n_ctx_fields entries) and cand_ids / cand_wts
 * hold only the remaining F - n_ctx_fields fields per candidate, row-major
 * [n_cand, F - n_ctx_fields].  Scores are exactly those of ctr_predict_async on the expanded
 * [n_cand, F] lists; the host link carries (F - n_ctx_fields) * 12 bytes per candidate instead
 * of F * 12.  The context lists are read when the shards are staged: keep them valid until the
 * ticket completes. */
int32_t ctr_predict_shared_async(ctr_engine* e, const int64_t* ctx_ids, const float* ctx_wts, int32_t n_ctx_fields,
                                 const int64_t* cand_ids, const float* cand_wts, int64_t n_cand, float* scores_out,
                                 ctr_callback cb, void* user, ctr_ticket* ticket);

/* Packed request (additive): the marshaller of sendRequest (DC:97-108) ships an int64 id and a float weight
 * for every field slot, 12 bytes each, 468 bytes per candidate at 39 fields -- and the host link, not the GPU,
 * bounds a request (PCIe Gen5 x16 carries ~107 M such candidates/s).  In the packed layout a per-candidate
 * field costs a 32-bit table ROW index (the id itself for direct tables; mix64(id) & (rows-1), i.e. what the
 * kernel would compute, for hashed ones) and a weight only if the field is flagged in weighted_mask (the
 * others weigh 1.0, as every sparse field of the reference's own inputs does, DC:67-74):
 * 26*4 + 13*8 = 208 bytes per candidate for the Criteo shape -- and 26*4 + 13*4 = 156 when the dense fields, whose id is
 * the same for every candidate (only their value changes), are flagged in fixed_id_mask: such a field has no id
 * column, its id is given once per request behind the context ids.  It composes with the shared context: the
 * first n_ctx_fields are given once (original int64 ids + weights, exactly as for ctr_predict_shared_async)
 * and cand_rows / cand_wts hold only the remaining fields: cand_rows [n_cand, Fk] uint32 with Fk = the per-candidate
 * fields that are not in fixed_id_mask, cand_wts [n_cand, Wk] float with Wk = number of weighted fields among the
 * per-candidate ones, both in field order (cand_wts may be NULL when Wk == 0).  ctx_ids holds n_ctx_fields context
 * ids followed by the ids of the fixed-id fields in field order; ctx_wts holds n_ctx_fields weights.  Scores are bit-identical to ctr_predict_async on the expanded lists.
 * Needs a table of at most 2^32 - 1 rows.  Buffers stay owned by the caller until the ticket completes.  A caller that
 * lays a request out in one slab -- cand_wts starting at the byte where cand_rows ends -- has it shipped as ONE copy. */
typedef struct ctr_packed_layout {
    int32_t struct_size;      /* sizeof(ctr_packed_layout) */
    int32_t n_ctx_fields;     /* leading fields given once per request; 0 = none */
    uint64_t weighted_mask;   /* bit f: field f carries a per-candidate weight (bits of context fields are ignored) */
    uint64_t fixed_id_mask;   /* bit f: per-candidate field f has the same id in every candidate of a request: no id column,
                               * the id travels once (after the context ids in ctx_ids).  A struct of the first 16 bytes only
                               * (struct_size 16) means 0 */
} ctr_packed_layout;
int32_t ctr_predict_packed_async(ctr_engine* e, const ctr_packed_layout* layout, const int64_t* ctx_ids, const float* ctx_wts,
                                 const uint32_t* cand_rows, const float* cand_wts, int64_t n_cand, float* scores_out,
                                 ctr_callback cb, void* user, ctr_ticket* ticket);

/* The packer for callers that hold the reference's lists (ids int64 / wts float, [n_cand, F]; DC:97-108): writes
 * the packed lists ctr_predict_packed_async takes (and the context + fixed ids of candidate 0 into ctx_*_out:
 * ctx_ids_out needs n_ctx_fields + popcount(fixed_id_mask) entries, ctx_wts_out n_ctx_fields).  Pure host code.  Fails with CTR_ERR_INVALID if a field the layout declares
 * unweighted has a weight other than 1.0, or a context field / a fixed id differs between candidates -- the packed
 * request would not score what the lists say.  Direct ids the table cannot hold are kept recognisable and
 * fail the request itself, as they do in the reference layout. */
int32_t ctr_pack_candidates(ctr_engine* e, const ctr_packed_layout* layout, const int64_t* ids, const float* wts, int64_t n_cand,
                            uint32_t* cand_rows_out, float* cand_wts_out, int64_t* ctx_ids_out, float* ctx_wts_out);

/* Blocks like CompletableFuture.join (DC:158-159).  timeout_ms < 0 waits for ever.  On
 * CTR_OK or a failure status the ticket is released; on CTR_ERR_TIMEOUT it stays valid.
 * A request submitted with a callback needs no ctr_wait: the engine releases its ticket by
 * itself (it keeps the 4096 most recently called-back tickets, so a ctr_wait issued soon
 * after the callback still succeeds). */
int32_t ctr_wait(ctr_engine* e, ctr_ticket ticket, int32_t timeout_ms);

/* Synchronous convenience: ctr_predict_async + ctr_wait. */
int32_t ctr_predict(ctr_engine* e, const int64_t* ids, const float* wts, int64_t n_cand, float* scores_out);

/* The same request followed by the reference's ascending sort of the merged scores
 * (Collections.sort(fullCtrList), DC:195; Float.compareTo order: -0.0 < +0.0, NaN last), done on
 * the device.  order_out (may be NULL) receives the candidate index of every sorted score --
 * the identity the reference's sort throws away; ties keep candidate order (stable). */
int32_t ctr_predict_sorted(ctr_engine* e, const int64_t* ids, const float* wts, int64_t n_cand,
                           float* sorted_scores_out, int32_t* order_out);

/* The RPC itself (stub.predict(request.build()), DC:111-112): a serialised
 * tensorflow.serving.PredictRequest exactly as DCNClient builds it (inputs "feat_ids" DT_INT64 and
 * "feat_wts" DT_FLOAT of shape [n, FIELD_NUM], as int64_val / float_val or tensor_content) in, a
 * serialised PredictResponse with outputs["prediction_node"] (DT_FLOAT [n, 1], float_val; DC:35,
 * DC:162) out.  A malformed or inconsistent request fails with CTR_ERR_INVALID, as the server
 * would answer INVALID_ARGUMENT.  If response_cap is too small the call fails with
 * CTR_ERR_CAPACITY and *response_len holds the size needed. */
int32_t ctr_predict_wire(ctr_engine* e, const void* request, int64_t request_len, void* response,
                         int64_t response_cap, int64_t* response_len);

/* The two halves of ctr_predict_wire as pure host functions (no engine, no GPU): parse a
 * PredictRequest into flat ids / weights (cap_elems = capacity of each output in elements), and
 * serialise scores as the PredictResponse the reference client reads. */
int32_t ctr_wire_parse_request(const void* request, int64_t request_len, int32_t field_num, int64_t* ids_out,
                               float* wts_out, int64_t cap_elems, int64_t* n_cand);
int32_t ctr_wire_build_response(const float* scores, int64_t n_cand, void* response, int64_t response_cap,
                                int64_t* response_len);

/* ---- pinned host memory for the zero-copy packer (north star (a)) ---- */
int32_t ctr_alloc_pinned(ctr_engine* e, int64_t bytes, void** out);
int32_t ctr_free_pinned(ctr_engine* e, void* p);

/* ---- device-resident scoring: sendRequest (DC:91-115) with tensors already in HBM ----
 * d_ids / d_wts / d_scores are device pointers on device_ids[device_index]; stream is a
 * cudaStream_t (NULL = the engine's own stream for that device, and the call then returns
 * after the work has finished, with out-of-range ids reported as CTR_ERR_INVALID; with a caller
 * stream the call only enqueues and such ids score as weight 0).  Calls on one device are
 * serialised while they enqueue; do not overlap calls that use different caller streams.  Used by
 * bench.py's device-resident figure and by callers that produce features on the GPU. */
int32_t ctr_score_device(ctr_engine* e, int32_t device_index, const void* d_ids, const void* d_wts,
                         int64_t n_cand, void* d_scores, void* stream);

/* ---- row-sharded tables with an NCCL all-to-all (BASELINE config 4; no counterpart in the reference, whose fan-out
 * DC:148-156 only knows replicated models) ----
 * For an engine created with CTR_TABLES_ROW_SHARDED over G >= 2 devices.  One call scores a batch on EVERY device of
 * the engine (array index = device index; n_cand[d] may be 0): per chunk, each device sends the row indices it needs
 * to their owners (all-to-all #1, 4 bytes per lookup), the owners copy the rows out of their shard at local HBM
 * speed, the rows travel back (all-to-all #2, 128 / 64 bytes per lookup) and the ordinary gather + tower kernels
 * score them -- the alternative to ctr_score_device on such an engine, whose gather reads remote rows directly
 * through NVLink peer mappings.  Same scores, bit for bit.  Blocks until every device is done.  Needs libnccl.so.2
 * at run time (bound with dlopen; CTR_ERR_STATE if absent).  A shard can be asked for at most twice its even share
 * of the lookups of a chunk (CTR_ERR_CAPACITY beyond that). */
int32_t ctr_score_devices_exchange(ctr_engine* e, const void* const* d_ids, const void* const* d_wts, const int64_t* n_cand,
                                   void* const* d_scores);

/* ---- device-resident scoring without any allocation (no counterpart in the reference) ----
 * Same contract and same results as ctr_score_device, for callers that must not allocate GPU
 * memory or enter the driver's allocator on the request path: processes whose HBM is already
 * owned in full by another pool, and latency-critical threads for which a cudaMalloc (which can
 * stall every stream of the device) is not acceptable.  The intermediate activations live in
 * one fixed __device__ array inside the library image (about 10.5 MB, resident on every GPU that
 * loads the library until the process exits), so a call handles at most
 * CTR_STATIC_MAX_CANDIDATES candidates (models with d_padded <= 768, hidden widths <= 1024);
 * larger requests fail with CTR_ERR_CAPACITY.  The call itself performs no cudaMalloc / cudaFree.
 * The first call on a GPU looks up the array's device address (once per GPU); make that first
 * call before starting any stream capture.  All calls on a GPU share the array, which keeps the
 * previous call's contents (what has to start at zero is cleared on `stream` at the start of each
 * call): calls whose GPU work overlaps in time race -- order them on one stream or with events.
 * stream == NULL runs on the engine's stream and returns when the scores are ready.  Host threads may call
 * concurrently: the engine serialises their enqueues per GPU (and, for stream == NULL, the whole call), so
 * two calls on one stream never interleave; calls on different caller streams still need the ordering above. */
#define CTR_STATIC_MAX_CANDIDATES 1024
int32_t ctr_score_device_static(ctr_engine* e, int32_t device_index, const void* d_ids, const void* d_wts,
                                int64_t n_cand, void* d_scores, void* stream);

/* ---- split rule, exported so a host can mirror the reference's partitionList ----
 * DC:46-55: begin/end for part i of n elements in part_num parts. */
int32_t ctr_partition_bounds(int64_t n, int32_t part_num, int64_t* begin, int64_t* end);

/* ---- introspection ---- */
const char* ctr_last_error(void);
int32_t ctr_abi_version(void);
/* Kernels launched by this engine since creation (all devices). */
int64_t ctr_engine_launch_count(ctr_engine* e);
/* Bytes the request path (ctr_predict* and friends) has copied host -> device / device -> host since creation: what really
 * crossed the link, after any host narrowing. */
int64_t ctr_engine_h2d_bytes(ctr_engine* e);
int64_t ctr_engine_d2h_bytes(ctr_engine* e);
/* Stage timing.  ctr_engine_set_timing(e, 1) starts (or restarts) an averaging window: each
 * following ctr_score_device call records CUDA events on the launching stream around every
 * kernel of its FIRST chunk (up to 32 calls per window).  After the caller has synchronised,
 * ctr_engine_stage_ms gives the mean device time in milliseconds of one stage over the window:
 * stage 0 = gather/pool, 1..n = n-th tower GEMM of the plan, 100 = head kernel, -1 = the whole
 * first chunk.  Negative when nothing was recorded. */
int32_t ctr_engine_set_timing(ctr_engine* e, int32_t enabled);
double ctr_engine_stage_ms(ctr_engine* e, int32_t device_index, int32_t stage);
/* Test hook: copies table rows (embedding as float32 [n,K], first-order weight [n]) back
 * from device memory of device_index, after the table dtype has been applied. */
int32_t ctr_debug_read_rows(ctr_engine* e, int32_t device_index, const int64_t* rows, int64_t n,
                            float* emb_out, float* w1_out);

/* Test hook: one tower layer in isolation through the tcgen05 GEMM kernel.
 * out[m, n] = max(sum_k half(A[m,k]) * half(W[n,k]) + bias[n], 0) rounded to half; host pointers,
 * A [M,K], W [N,K] (the [out, in] layout of the weight file), out [M,N]. */
int32_t ctr_debug_gemm(ctr_engine* e, int32_t device_index, const float* A, const float* W, const float* bias,
                       int32_t M, int32_t N, int32_t K, int32_t relu, float* out);
/* Test hook: the gather/pool kernel's outputs for n candidates (host pointers): x_out float32
 * [n, F*16] (the half activations widened), wide_out [n] (lin, + FM for DeepFM); either may be NULL. */
int32_t ctr_debug_gather(ctr_engine* e, int32_t device_index, const int64_t* ids, const float* wts, int64_t n,
                         float* x_out, float* wide_out);

#ifdef __cplusplus
}
#endif
#endif /* CTR_ENGINE_H_ */
