/* libccfd_rf — C ABI of the B200 random-forest scoring path.
 *
 * This is the drop-in boundary for ONE path of the credit-card-fraud demo: the
 * fraud model's batched RandomForestClassifier.predict_proba behind the Seldon
 * model node `modelfull` (reference deploy/model/modelfull.json:20-44, reached
 * at SELDON_URL + SELDON_ENDPOINT, reference deploy/router.yaml:65-68).  The
 * reference ships that node as an opaque Python image, so there is no FFI in
 * the reference to copy; the entry points below are what a Python model class
 * (ctypes) binds in place of scikit-learn's calls.  SK/ is
 * scikit-learn 1.9.0's source tree, where the replaced arithmetic lives.
 *
 * Conventions: every function returns CCFD_OK (0) or a negative CCFD_E* code,
 * never throws, never aborts; ccfd_last_error() returns a thread-local message
 * for the last failure on the calling thread.  Handles are opaque and owned by
 * the library; bulk buffers are owned by the caller unless stated.  Plain
 * pointers and sizes only: no C++ or torch types cross this line.
 */
#ifndef CCFD_RF_H
#define CCFD_RF_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define CCFD_API __attribute__((visibility("default")))
#else
#define CCFD_API
#endif

#define CCFD_OK 0
#define CCFD_EINVAL (-1)       /* bad argument or malformed forest  -> Python ValueError */
#define CCFD_ECUDA (-2)        /* CUDA runtime failure (message carries cudaGetErrorString) */
#define CCFD_ENOMEM (-3)       /* host or device allocation failed */
#define CCFD_EUNSUPPORTED (-4) /* valid input this build does not handle (e.g. n_classes != 2) */
#define CCFD_ESTATE (-5)       /* call made in the wrong state (e.g. wait on an idle slot) */

/* flags for ccfd_predict_proba / ccfd_predict_decision */
#define CCFD_DEVICE_PTRS 0x1u  /* X and out are device pointers on the forest's device (default: host) */
#define CCFD_ASYNC 0x2u        /* device pointers only: enqueue on `stream` and return without waiting */
/* kernel choice; AUTO picks by forest and batch size.  The others exist so
 * tests and the bench can pin a kernel; they fail with CCFD_EUNSUPPORTED when
 * the forest does not fit that kernel. */
#define CCFD_KERNEL_AUTO 0x000u
#define CCFD_KERNEL_STREAM 0x100u /* throughput: rows resident in shared memory, trees streamed through it */
#define CCFD_KERNEL_SPLIT 0x200u  /* latency: (row tile x tree chunk) grid, per-tree scratch, ordered reduce */
#define CCFD_KERNEL_GLOBAL 0x300u /* any forest: nodes read from global memory / L2 */
#define CCFD_KERNEL_MASK 0xF00u

typedef struct ccfd_forest ccfd_forest;   /* immutable after create; shareable across threads */
typedef struct ccfd_session ccfd_session; /* single producer / single consumer */

/* A fitted forest as plain structure-of-arrays, all host pointers, all copied
 * by ccfd_forest_create.  Replaces what predict_proba reads from
 * RandomForestClassifier.estimators_[t].tree_ (SK/tree/_tree.pyx:656-765,
 * Node fields SK/tree/_tree.pxd:15-25).  Node indices are tree-local. */
typedef struct ccfd_forest_desc {
    int32_t n_trees;
    int32_t n_features;            /* 1..255; the demo's model has 30 (Time, V1..V28, Amount) */
    int32_t n_classes;             /* this build: 2 (non-fraud, fraud) */
    int32_t reserved;              /* must be 0 */
    const int64_t* tree_offsets;   /* [n_trees+1] node range of each tree in the arrays below */
    const int32_t* feature;        /* [total] split feature; ignored at leaves */
    const float* threshold;        /* [total] LARGEST fp32 <= sklearn's fp64 threshold (see converter) */
    const int32_t* left;           /* [total] left child, -1 at leaves  (x <= threshold, SK/tree/_tree.pyx:989) */
    const int32_t* right;          /* [total] right child, -1 at leaves */
    const uint8_t* missing_go_to_left; /* [total] NaN routing (SK/tree/_tree.pyx:984-988) */
    const double* value;           /* [total * n_classes] class fractions at every node; leaves are read */
} ccfd_forest_desc;

typedef struct ccfd_forest_info {
    int32_t n_trees, n_features, n_classes, max_depth;
    int64_t n_nodes;       /* reachable nodes kept */
    int64_t n_leaves;
    int64_t model_bytes;   /* packed device bytes: what one launch reads of the model */
    int64_t max_tree_bytes;
    int32_t device;
    int32_t stream_kernel_ok; /* 1 if every tree fits the shared-memory stage of the STREAM/SPLIT kernels */
} ccfd_forest_info;

/* Number of CUDA devices, or a negative CCFD_E* code. */
CCFD_API int ccfd_device_count(void);
/* Thread-local message of the last failure on this thread ("" if none). */
CCFD_API const char* ccfd_last_error(void);

/* Validate, re-lay-out (8-byte nodes, the two children of a node adjacent; breadth-first for trees
 * that fit shared memory, three-level blocks for larger ones) and upload.
 * Replaces: unpickling the forest inside the model image (reference
 * deploy/model/modelfull.json:24) — the library's equivalent of model load. */
CCFD_API int ccfd_forest_create(const ccfd_forest_desc* desc, int device, ccfd_forest** out);
CCFD_API int ccfd_forest_destroy(ccfd_forest* forest);
/* The same validation and re-layout without a GPU: writes the packed image the kernels read
 * (all tree blobs back to back), 8 uint32 per tree {offset/16, bytes, leaf-value offset, depth,
 * first position whose leaf slot is not stored (= node count when all are), value index there,
 * 0 (filled at create time), node-encoding flag (bit 2 set: wide encoding, the forest has a tree
 * beyond a shared-memory stage; clear: compact encoding)} and, if top_nodes is not NULL,
 * 14 uint32 per tree: the seven nodes of the top three levels as the STREAM kernel receives them.
 * Call with blob == NULL and tree_recs == NULL to learn *blob_bytes first.  Lets a deployment
 * build step (and the CPU tests) check a forest before it reaches a GPU; has no counterpart in
 * scikit-learn, whose trees are scored in their fitted layout. */
CCFD_API int ccfd_forest_pack(const ccfd_forest_desc* desc, void* blob, int64_t blob_capacity, uint32_t* tree_recs,
                              uint32_t* top_nodes, int64_t* blob_bytes);
CCFD_API int ccfd_forest_get_info(const ccfd_forest* forest, ccfd_forest_info* info);

/* out[i, 0..1] = mean over trees of the leaf class fractions of row i, float64,
 * summed in tree order 0..T-1 and then divided by T.
 * Replaces ForestClassifier.predict_proba (SK/ensemble/_forest.py:921-967) with
 * everything under it: Tree._apply_dense (SK/tree/_tree.pyx:954-996), the leaf
 * gather (SK/tree/_tree.pyx:939-945) and _accumulate_prediction
 * (SK/ensemble/_forest.py:704-717).
 *   X    n rows of n_features float32, row stride `ld` floats (ld >= n_features).
 *        NaN is routed by missing_go_to_left; +-inf is the caller's to reject
 *        (sklearn raises ValueError before scoring, SK/ensemble/_forest.py:606-625).
 *   out  n * 2 float64, row-major.
 *   stream  a cudaStream_t (NULL = the default stream).
 * Host pointers: the call copies in chunks, overlapping copies with scoring,
 * and returns when `out` is complete (pinned memory makes the copies fast).
 * Re-entrant for device pointers; host-pointer calls on one forest serialise. */
CCFD_API int ccfd_predict_proba(ccfd_forest* forest, const float* X, int64_t n, int64_t ld, double* out,
                       uint32_t flags, void* stream);

/* The request matrix as the Seldon wrapper hands it to predict(): n rows of n_features values in ordinary
 * (pageable) host memory, float64 (x_is_f64 != 0; what JSON decoding produces) or float32, row stride `ld`
 * ELEMENTS.  One call does what ForestClassifier._validate_X_predict (SK/ensemble/_forest.py:606-625: cast
 * to float32, reject +-inf and float32 overflow, allow NaN) and predict_proba (:921-967) do, pipelined:
 * the library's host threads narrow chunk k+1 into pinned memory while the GPU copies in, scores and copies
 * out chunk k.  Input holding +-inf: CCFD_EINVAL with scikit-learn's message, *has_inf = 1 (may be NULL),
 * `out` then partly written.  `flags`: a CCFD_KERNEL_* selector only.  Calls on one forest serialise.
 * (ccfd_predict_proba with host pointers takes the same pipeline by itself when X is pageable float32.) */
CCFD_API int ccfd_predict_proba_rows(ccfd_forest* forest, const void* X, int32_t x_is_f64, int64_t n, int64_t ld,
                                     double* out, uint32_t flags, int32_t* has_inf);

/* Which kernel CCFD_KERNEL_AUTO resolves to for a batch of n rows (*kernel = CCFD_KERNEL_STREAM, _SPLIT or
 * _GLOBAL) and whether a device-pointer call of that size orders its rows first (*rows_presorted, may be
 * NULL; see csrc/sort.cu).  For reports and tests; no counterpart in scikit-learn, which has one code path. */
CCFD_API int ccfd_kernel_for(const ccfd_forest* forest, int64_t n, uint32_t* kernel, int32_t* rows_presorted);

/* Same scoring, plus the router's rule fused into the epilogue: decision[i] = 1
 * iff p_fraud (op) threshold, op being `>` (ge == 0) or `>=` (ge != 0).
 * Replaces the Drools rule on the returned probability against
 * FRAUD_THRESHOLD = 0.5 (reference deploy/router.yaml:69-70, README.md:427).
 * `out` may be NULL to emit decisions only (bulk rescoring: 1 byte per row). */
CCFD_API int ccfd_predict_decision(ccfd_forest* forest, const float* X, int64_t n, int64_t ld, double* out,
                          uint8_t* decision, double threshold, int ge, uint32_t flags, void* stream);

/* Scoring for callers that must not allocate.  No counterpart in the reference or in
 * scikit-learn: it exists for hosts where the other entry points' cudaMallocAsync / cudaMalloc
 * is off limits — all device memory already belongs to another allocator, or an allocation's
 * device-wide synchronisation on the request thread is unacceptable.  Same results as
 * ccfd_predict_proba with CCFD_DEVICE_PTRS | CCFD_KERNEL_SPLIT, but the per-(tree,row) scratch is
 * a fixed 16 MiB array inside the library image (resident on every GPU that loads the library,
 * until exit), so the call performs no allocation of any kind.
 *   X, out   device pointers on the forest's device; the call returns when `out` is complete.
 *   limit    n * n_trees <= CCFD_NOALLOC_MAX_TREE_ROWS, else CCFD_EINVAL (e.g. 10,485 rows of the
 *            100-tree model); forests whose trees do not fit shared memory: CCFD_EUNSUPPORTED.
 *   sharing  one scratch per GPU, shared by all calls and left as the last call wrote it.  No
 *            clearing is needed (each call writes every slot it later reads), but calls that
 *            overlap on one GPU race on it: issue them one after another.
 *   set-up   the first call on each GPU resolves the array's device address and caches it (which,
 *            with lazy module loading, is also when the GPU loads the library image and the 16 MiB
 *            appear); make that call at start-up, before any CUDA stream capture.  The cached address
 *            dies with the GPU's primary context: do not call again after cudaDeviceReset(). */
#define CCFD_NOALLOC_MAX_TREE_ROWS (1 << 20)
CCFD_API int ccfd_predict_proba_noalloc(ccfd_forest* forest, const float* X, int64_t n, int64_t ld, double* out,
                                        void* stream);

/* Latency path: `n_slots` pinned host slots of `max_batch` rows, each with its
 * own device buffers, scratch, a highest-priority stream and one pre-captured CUDA graph per
 * power-of-two request size (copy in, score, result; requests of up to 4,096 rows have their last
 * kernel write the result straight into the pinned buffer, larger ones copy out).
 * Replaces the per-request work of the Seldon wrapper around predict()
 * (requests arriving at reference deploy/router.yaml:65-68): the caller writes
 * rows straight into the slot and no allocation or Python happens per request. */
CCFD_API int ccfd_session_create(ccfd_forest* forest, int64_t max_batch, int32_t n_slots, ccfd_session** out);
CCFD_API int ccfd_session_destroy(ccfd_session* session);
/* Library-owned pinned buffers of slot `slot`: x (max_batch * n_features fp32,
 * dense rows) to write, p (max_batch * 2 fp64) to read after wait. */
CCFD_API int ccfd_session_buffers(ccfd_session* session, int32_t slot, float** x, double** p);
/* Score the first n_rows rows of the slot (1 <= n_rows <= max_batch); returns at once. */
CCFD_API int ccfd_session_submit(ccfd_session* session, int32_t slot, int64_t n_rows);
/* Block until the slot's results are visible in its p buffer. */
CCFD_API int ccfd_session_wait(ccfd_session* session, int32_t slot);
/* submit + wait in one call, for callers that have nothing to overlap (one foreign-function call
 * per request instead of two). */
CCFD_API int ccfd_session_run(ccfd_session* session, int32_t slot, int64_t n_rows);

/* Micro-batcher over a session (BASELINE configs[1]: "1,024-transaction micro-batches replayed from a
 * Kafka-shaped producer stream").  Replaces the per-message model call of the demo's router for the
 * transactions the producer puts on topic odh-demo (reference deploy/kafka/ProducerDeployment.yaml:88-97,
 * README.md:547-548).  Records are n_features float32 each (120 bytes for the demo's model).  The
 * session's pinned slots are the ring: push copies records into the slot being filled, a slot that
 * reaches batch_rows is submitted (one CUDA-graph launch) and the next one fills while it is in flight
 * (the session needs >= 2 slots); a collector thread waits for submitted slots in order.  A batch's
 * latency runs from the arrival of its last record to its results being visible in host memory.
 * One producer thread (push / flush / replay) and one consumer thread (poll) at a time; push blocks
 * while every slot is filling, in flight or not yet polled. */
typedef struct ccfd_batcher ccfd_batcher;
CCFD_API int ccfd_batcher_create(ccfd_session* session, int64_t batch_rows, ccfd_batcher** out);
CCFD_API int ccfd_batcher_destroy(ccfd_batcher* batcher); /* drains what is in flight */
/* n records, record stride `ld` floats (ld >= n_features).  A record holding +-inf is refused with
 * CCFD_EINVAL (scikit-learn refuses it before scoring, SK/ensemble/_forest.py:606-625); NaN is allowed. */
CCFD_API int ccfd_batcher_push(ccfd_batcher* batcher, const float* records, int64_t n, int64_t ld);
CCFD_API int ccfd_batcher_flush(ccfd_batcher* batcher); /* submit a partly filled slot (end of stream, idle timeout) */
/* Next completed batch, in submission order: returns 1 and fills *first_row (index of its first record
 * in the stream), *n_rows, proba (n_rows x 2 float64 copied out of the slot, may be NULL) and
 * *latency_us; returns 0 if none completed within timeout_ms (0 = do not wait).  Polling frees the slot. */
CCFD_API int ccfd_batcher_poll(ccfd_batcher* batcher, int32_t timeout_ms, int64_t* first_row, int64_t* n_rows, double* proba,
                               int64_t cap_rows, double* latency_us);
/* Replay tool (the demo's producer replays creditcard.csv from S3): pushes `n` records from memory at
 * records_per_s (<= 0: as fast as the ring takes them), `burst` records at a time as a consumer poll
 * would deliver them, busy-waiting on the steady clock for each burst's due time, then flushes. */
CCFD_API int ccfd_batcher_replay(ccfd_batcher* batcher, const float* records, int64_t n, int64_t ld, double records_per_s,
                                 int32_t burst);
CCFD_API int ccfd_batcher_stats(ccfd_batcher* batcher, int64_t* batches_done, int64_t* records_in);

/* A whole request in one call: rows as the wrapper holds them (float64 if x_is_f64 != 0, else float32; row
 * stride `ld` elements) are narrowed into slot `slot`'s pinned buffer with scikit-learn's infinity check
 * (SK/ensemble/_forest.py:606-625: CCFD_EINVAL and *has_inf = 1, may be NULL), the slot's graph runs, and the
 * n x 2 probabilities are copied to `out`.  Replaces, for requests of up to max_batch rows, the
 * cast + predict_proba pair behind the Seldon wrapper's predict() (requests arriving at reference
 * deploy/router.yaml:65-68).  One caller per slot at a time. */
CCFD_API int ccfd_session_score(ccfd_session* session, int32_t slot, const void* X, int32_t x_is_f64, int64_t n, int64_t ld,
                                double* out, int32_t* has_inf);

/* Request text -> float32 rows, written straight into `dst` (typically a session slot's pinned x
 * buffer, so the bytes the GPU copies are produced in one pass).  Replaces the JSON decoding of the
 * Seldon Python wrapper in front of predict() for requests arriving at reference
 * deploy/router.yaml:65-68, and the CSV handling of the Kafka consumer side of reference
 * deploy/kafka/ProducerDeployment.yaml:88-97.  Accepts Seldon JSON ("ndarray": a matrix whose inner
 * arrays are the rows; "tensor": {"shape":[n, cols_in], "values":[flat]}; nothing after the payload is
 * read), JSON lines (one array or object per line; object keys are not cells), bare nested arrays, and
 * CSV rows (a line of words only is a header and is skipped; quoted numbers accepted).  The decoder is
 * structure-aware: every row must hold exactly `cols_in` numbers, of which the first `cols_keep` are
 * kept (31 -> 30 drops the CSV's Class column); a row of another width, a non-numeric or boolean cell,
 * a 1-D ndarray, or a tensor whose shape disagrees with its values is CCFD_EINVAL -- the payloads the
 * Seldon wrapper + scikit-learn answer with a ValueError.  Cells are whole tokens ("Info" is a word,
 * not inf).  Values take the same text -> float64 -> float32 roundings as json.loads + numpy.
 * NaN/null -> NaN, Infinity -> inf (for the caller to reject, as sklearn does).  Newline-delimited
 * inputs above 256 KB are parsed on the library's host worker pool.  Pure host code: no GPU is touched. */
CCFD_API int ccfd_parse_rows(const char* text, int64_t len, int32_t cols_in, int32_t cols_keep, float* dst,
                             int64_t cap_rows, int64_t* n_rows);

/* A float64 (src_is_f64 != 0) or float32 matrix of n rows x n_cols, row stride src_ld elements, ->
 * dense float32 rows in dst, and *has_inf = 1 if any narrowed value is +-inf (NaN is allowed), in one
 * pass (the library's persistent host worker pool for inputs above ~1 M values, the calling thread below).  Replaces the cast and the finiteness scan of
 * ForestClassifier._validate_X_predict (SK/ensemble/_forest.py:606-625); the narrowing is
 * bit-identical to numpy's astype(float32).  dst == NULL with float32 input scans only (rows that are
 * already dense float32 need no copy).  Pure host code. */
CCFD_API int ccfd_cast_rows(const void* src, int32_t src_is_f64, int64_t n, int64_t n_cols, int64_t src_ld, float* dst,
                            int32_t* has_inf);

/* The other direction: n_rows x 2 float64 probabilities -> "[[p0,p1],...]", the "ndarray" value of a
 * Seldon response, using the shortest decimal strings that read back to exactly the same float64
 * (what json.dumps produces, several times faster).  `cap` >= 68 * n_rows + 2 always suffices.
 * Replaces the response encoding of the Seldon Python wrapper.  Pure host code. */
CCFD_API int ccfd_format_proba(const double* p, int64_t n_rows, char* dst, int64_t cap, int64_t* len);

#ifdef __cplusplus
}
#endif
#endif /* CCFD_RF_H */
