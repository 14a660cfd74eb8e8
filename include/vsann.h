/*
 * vsann.h — C ABI of the B200-native ANN index (drop-in for the usearch calls of
 * the reference's index actor).
 *
 * Every entry point names the reference interface it replaces. Citations are
 * relative to the reference tree (swasik/vector-store-text):
 *   src/index/usearch.rs  — the orphaned usearch backend (the hot path)
 *   src/lib.rs            — domain newtypes (Dimensions, Connectivity, ...)
 *
 * Conventions
 *   - plain pointers + sizes, no C++/torch types; status codes, never exceptions
 *   - all pointers are HOST memory unless the function name ends in `_device`
 *   - caller keeps ownership of every buffer; inputs are copied/staged before return
 *   - thread-safe: any number of host threads may call into one index concurrently
 *     (the reference holds the library under Arc<RwLock<_>>, usearch.rs:98,201-221,245,276,309);
 *     searches do not wait for a running insertion batch: they read the part of the graph that has landed
 *   - the GPU library has NO CPU fallback: if no CUDA device is usable,
 *     vs_index_create fails with VS_ERR_CUDA.
 *
 * Two libraries export this ABI:
 *   libvsann.so          (vector-store-text_b200/csrc)  — the product, sm_100a CUDA
 *   liboracle_vsann.so   (oracle/)                       — test-only CPU restatement,
 *                                                          same signatures with an `o` prefix (ovs_*)
 */
#ifndef VSANN_H
#define VSANN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vs_index vs_index;

typedef enum vs_status {
    VS_OK = 0,
    VS_ERR_INVALID = 1,     /* null pointer / bad option / k == 0                      */
    VS_ERR_DIMS = 2,        /* wrong embedding length  (usearch.rs:259-272)             */
    VS_NOT_FOUND = 3,       /* remove of an unknown key — non-fatal (usearch.rs:245-247)*/
    VS_ERR_CAPACITY = 4,    /* reserve failed (usearch.rs:206-210)                      */
    VS_ERR_CUDA = 5,        /* CUDA runtime error / no device                           */
    VS_ERR_NOMEM = 6,
    VS_ERR_UNSUPPORTED = 7
} vs_status;

/* Distance definitions (smaller = closer):
 *   COS  = 1 - a.b / (|a| |b|)     L2SQ = sum (a-b)^2  (no sqrt)     IP = 1 - a.b
 * The reference never overrides the metric (usearch.rs:89-96 => crate default, cosine);
 * the ABI makes it explicit. */
typedef enum vs_metric { VS_METRIC_COS = 0, VS_METRIC_L2SQ = 1, VS_METRIC_IP = 2 } vs_metric;

/* Mirrors IndexOptions built at usearch.rs:89-96 + reserve(RESERVE_INCREMENT) at :99.
 * connectivity / expansion_add / expansion_search == 0 => library defaults 16 / 128 / 64
 * (lib.rs:164,182,200: the newtypes default to 0). */
typedef struct vs_options {
    size_t dims;              /* Dimensions (lib.rs:147); must be > 0                       */
    int32_t metric;           /* vs_metric                                                  */
    size_t connectivity;      /* M: level-0 rows hold 2M links, upper levels M. 2..16 (0 => 16): a level-0 row is
                               * at most 32 links, one per lane of the traversing warp; larger values are refused
                               * with VS_ERR_UNSUPPORTED (the reference passes its newtype through unbounded,
                               * lib.rs:164, but only ever uses the crate default)                              */
    size_t expansion_add;     /* ef used while inserting                                    */
    size_t expansion_search;  /* default ef for vs_search (effective ef = max(ef, k))       */
    size_t capacity_hint;     /* initial reserve; 0 => 1 000 000 (usearch.rs:62,99)         */
    int32_t device;           /* CUDA device ordinal (single-device index, n_devices <= 1)  */
    size_t max_batch;         /* insertion batch cap; 0 => library default                  */
    /* One index over several GPUs of one box, inside ONE process (the reference's actor holds one
     * Arc<RwLock<usearch::Index>> - usearch.rs:98,105-136): n_devices in 2..8 => the index owns one shard per
     * device_ids[i], shard = key % n_devices; adds and removes are routed, searches fan out
     * on every device's own stream and a merge kernel on device_ids[0] reads the shards' result lists over
     * NVLink peer mappings. 0 or 1 => a single shard on `device`. (These 32 bytes were `reserved[4]`.) */
    uint8_t n_devices;
    uint8_t device_ids[8];
    uint8_t reserved[23];     /* must be zero                                               */
} vs_options;

/* Work counters; device-side counters are collected only when enabled with
 * vs_set_stats(index, 1) (they cost atomics). */
typedef struct vs_stats {
    uint64_t count;              /* live vectors (adds - removes)                           */
    uint64_t slots;              /* slots ever used (live + tombstoned)                     */
    uint64_t capacity;
    uint64_t pending;            /* staged adds not yet in the graph                        */
    uint64_t search_queries;     /* since creation / last vs_reset_stats                    */
    uint64_t search_dist_evals;
    uint64_t search_expansions;  /* beam pops that fetched a neighbour row                  */
    uint64_t build_vectors;
    uint64_t build_dist_evals;
    uint64_t build_batches;
    double   build_seconds;      /* device time spent inside insertion batches              */
    uint64_t device_bytes;       /* HBM held by this index                                  */
    uint32_t top_level;
    uint32_t dims;                /* embedding dimensions (what a loaded index was saved with) */
    uint64_t exact_tc_queries;   /* exact-mode queries answered by the tensor-core scan + FP32 re-rank */
    uint64_t exact_tc_fallback;  /* ... of which the FP32 tile kernel had to answer (overflow / < k)  */
    uint64_t exact_tc_second_scan; /* ... of which needed a second tensor scan with a proven threshold */
    uint64_t exact_tc_scans;     /* exact-mode calls answered by the tensor-core path                  */
    double   exact_tc_scan_seconds; /* device time of their full scans (only while vs_set_stats is on)  */
    uint64_t exact_stream_queries; /* exact-mode queries answered by the HBM-streaming scan (batches of <= 16 queries) */
    uint64_t search_host_queries;  /* queries that came through vs_search / vs_search_batch (host buffers) ...              */
    uint64_t search_host_launches; /* ... and the launches that answered them: concurrent vs_search calls share a launch      */
    uint64_t build_expansions;     /* beam pops of the insertion searches (device counter, with vs_set_stats)                 */
    uint64_t reserve_calls;        /* capacity growths (usearch.rs:201-212) ...                                               */
    double   reserve_seconds;      /* ... and the wall time they took                                                         */
    uint64_t compactions;          /* rebuilds that dropped tombstoned slots (slot reuse, usearch.rs:214-221,245) ...         */
    double   compact_seconds;      /* ... and the wall time they took                                                         */
    uint32_t tables_grow_in_place; /* 1: tables are VMM-backed and grow without a copy or a new pointer                       */
    uint32_t n_devices;            /* shards (GPUs) behind this index                                                         */
} vs_stats;

/* usearch::Index::new + reserve — usearch.rs:82-99 (called from create_index, :38-53). */
vs_status vs_index_create(const vs_options* options, vs_index** out);
/* Dropping the last Sender<Index> ends the actor and frees the index — usearch.rs:120,133. */
void vs_index_destroy(vs_index* index);

/* usearch::Index::add(key, &embedding) — usearch.rs:221. Adding a key that is present
 * replaces it (remove + add under the same key — usearch.rs:214-225). Adds are staged
 * and land in the graph in batches; vs_count/vs_search/vs_flush make them land. */
vs_status vs_add(vs_index* index, uint64_t key, const float* vec);
vs_status vs_add_batch(vs_index* index, size_t n, const uint64_t* keys, const float* vecs /* n x dims, row-major */);
/* Same, rows already resident in HBM on the index's device (keys stay on the host). */
vs_status vs_add_batch_device(vs_index* index, size_t n, const uint64_t* keys, const float* d_vecs);

/* usearch::Index::remove(key) — usearch.rs:215,245. Tombstone; count drops by one. Does not flush staged adds
 * (a staged add of the key is dropped). Once tombstones exceed a quarter of the slots the index compacts itself:
 * live rows are gathered to the front and the graph is rebuilt, so slots (and HBM) stay bounded under the
 * reference's replace-heavy ingest (usearch frees and reuses the slot, usearch.rs:214-221). */
vs_status vs_remove(vs_index* index, uint64_t key);

/* usearch::Index::search(&embedding, limit) -> Matches{keys, distances} — usearch.rs:276,288,298.
 * Results ascending by (distance, key); *found <= k (fewer if the index is smaller).
 * ef_override == 0 => expansion_search; exact != 0 => brute force over all live vectors.
 * keys_out / dist_out hold nq x k entries; unused tail entries are UINT64_MAX / +inf.
 * vs_search calls that arrive from several threads while a launch is in flight are coalesced into the next launch
 * (the reference runs up to 2 x T of them at once on its rayon pool, usearch.rs:118-131,275-276). */
vs_status vs_search(vs_index* index, const float* query, size_t k,
                    uint64_t* keys_out, float* dist_out, size_t* found);
vs_status vs_search_batch(vs_index* index, size_t nq, const float* queries, size_t k,
                          size_t ef_override, int exact,
                          uint64_t* keys_out, float* dist_out, size_t* found /* nq entries, may be NULL */);
/* Device-resident variant: d_queries, d_keys_out, d_dist_out are device pointers on the
 * index's device; enqueues on `stream` (a cudaStream_t, NULL = the index's own stream)
 * and returns without synchronising when stream != NULL. */
vs_status vs_search_batch_device(vs_index* index, size_t nq, const float* d_queries, size_t k,
                                 size_t ef_override, int exact,
                                 uint64_t* d_keys_out, float* d_dist_out, void* stream);
/* The same graph search with every query's result list written to n_out (1..9) buffers: d_keys_out[0] / d_dist_out[0] is
 * the index's device, the others may be ANY device-writable memory - other GPUs' buffers mapped over NVLink / NVSwitch in a
 * sharded deployment (one index per GPU): a shard's lists reach every peer while its traversal is still
 * running, and the exchange step that follows (vs_merge_topk_flags_device) reads local memory only. Graph search on a
 * single-device index only (exact mode and multi-device indexes: VS_ERR_UNSUPPORTED). No counterpart in the reference, whose
 * index lives in one process (usearch.rs:98). */
vs_status vs_search_batch_fanout_device(vs_index* index, size_t nq, const float* d_queries, size_t k, size_t ef_override,
                                        size_t n_out, uint64_t* const* d_keys_out, float* const* d_dist_out, void* stream);

/* usearch::Index::size / capacity / reserve — usearch.rs:201-206,309. */
size_t vs_count(vs_index* index);
size_t vs_capacity(vs_index* index);
vs_status vs_reserve(vs_index* index, size_t capacity);

/* New: force staged adds into the graph (the reference's tests poll count() instead —
 * usearch.rs:352-358). */
vs_status vs_flush(vs_index* index);

vs_status vs_get_stats(vs_index* index, vs_stats* out);
vs_status vs_set_stats(vs_index* index, int enabled);
vs_status vs_reset_stats(vs_index* index);

/* Sharded search: merge `parts` per-shard result lists
 * (each nq x k, ascending) laid out [part][query][k] into the global top-k per query.
 * Device pointers on `device`; runs on `stream`. Follows the NCCL all-gather. */
vs_status vs_merge_topk_device(int device, size_t parts, size_t nq, size_t k,
                               const uint64_t* d_keys_in, const float* d_dist_in,
                               uint64_t* d_keys_out, float* d_dist_out, void* stream);

/* Fused gather + merge over peer memory: part p's lists (nq x k, ascending) are read in place from d_part_keys[p] /
 * d_part_dist[p] - device pointers that may address a PEER GPU's memory mapped over NVLink / NVSwitch (CUDA IPC,
 * symmetric memory) - so the exchange needs no all-gather and no gathered copy. d_part_keys / d_part_dist are DEVICE
 * arrays of `parts` pointers. The caller orders the call after every shard's search (a cross-rank barrier). */
vs_status vs_merge_topk_peers_device(int device, size_t parts, size_t nq, size_t k,
                                     const uint64_t* const* d_part_keys, const float* const* d_part_dist,
                                     uint64_t* d_keys_out, float* d_dist_out, void* stream);

/* The same with the cross-rank barrier INSIDE the merge kernel (one process per GPU, lists and flag words in symmetric
 * memory): d_part_flags[p] is rank p's flag word as mapped on this device. The call follows the rank's own search on
 * `stream`; the kernel publishes flag[my_part] = epoch (system-scope release) and waits until every peer's flag has
 * reached `epoch` (system-scope acquire over NVLink) before it reads that peer's lists. Epochs increase by one per step;
 * two alternating list buffers make one flag per rank sufficient. No separate barrier kernel, no all-gather. */
vs_status vs_merge_topk_flags_device(int device, size_t parts, size_t nq, size_t k,
                                     const uint64_t* const* d_part_keys, const float* const* d_part_dist,
                                     uint32_t* const* d_part_flags, uint32_t my_part, uint32_t epoch,
                                     uint64_t* d_keys_out, float* d_dist_out, void* stream);

/* ---- allocation-free brute force -------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). Exact top-k of
 * every query over a caller-held device corpus - the arithmetic and the ordering of vs_search_batch(exact = 1):
 * one sequential fmaf chain per (query, row), results ascending by (distance, key) - with its scratch (per-chunk
 * partial top-k lists) in a static device array of VS_NOALLOC_SCRATCH_BYTES that every GPU loading the library
 * holds until exit. The call allocates nothing: no cudaMalloc / cudaFree / allocating library routine; the first
 * call on a device asks the array's address once (cudaGetSymbolAddress) - make that call outside a capture.
 *   d_vecs     n x dims f32 row-major, device memory on `device`; d_keys: n u64 keys or NULL (key of row r = r)
 *   d_queries  nq x dims f32; d_keys_out / d_dist_out: nq x k, unused tail entries UINT64_MAX / +inf
 *   stream     a cudaStream_t; NULL = the default stream, synchronised before returning
 * Limits (VS_ERR_UNSUPPORTED / VS_ERR_DIMS with a message naming the limit): k <= VS_NOALLOC_MAX_K,
 * dims <= VS_NOALLOC_MAX_DIMS, nq x roundup32(k) <= VS_NOALLOC_MAX_QUERY_K (e.g. 512 queries at k <= 32,
 * 64 at k = 256), nq <= 65535, n < 2^31; no tombstones (the corpus is the caller's).
 * Every call shares the one array and finds the previous call's contents there; nothing in it needs clearing
 * (a call writes every word it later reads). Calls that OVERLAP on the GPU race on it: serialise them. */
#define VS_NOALLOC_MAX_K 256
#define VS_NOALLOC_MAX_DIMS 4096
#define VS_NOALLOC_MAX_QUERY_K 16384
#define VS_NOALLOC_SCRATCH_BYTES (8u << 20)
vs_status vs_exact_search_noalloc_device(int device, int32_t metric, size_t dims, size_t n,
                                         const float* d_vecs, const uint64_t* d_keys,
                                         size_t nq, const float* d_queries, size_t k,
                                         uint64_t* d_keys_out, float* d_dist_out, void* stream);

/* Debug/verification: copy the level-`level` neighbour table to the host.
 * links_out holds n_rows x row_width u32 slot ids (UINT32_MAX = empty);
 * slot_keys_out (may be NULL) holds the u64 key of each of the `slots` slots.
 * Query the sizes first with links_out == NULL. */
vs_status vs_export_level(vs_index* index, uint32_t level, uint32_t* links_out,
                          size_t* n_rows, size_t* row_width, uint64_t* slot_keys_out, size_t* slots);

/* ---- persistence -------------------------------------------------------------
 * The reference has none: the index is RAM-only and is rebuilt from a full table scan whenever it
 * is (re)created (src/db_index.rs:108-117,217-237; no usearch save/load/view call anywhere). These
 * two calls dump and restore the landed state - vectors, norms, every link table, slot keys and
 * tombstones - so that a restart costs one file read instead of a rebuild. The restored index
 * answers every query with the same ids and distances as the saved one and accepts further
 * add / remove calls.
 *
 * vs_index_save flushes staged adds first. File: 128-byte header ("VSANNIDX", format version,
 * dims, metric, connectivity, expansions, slot count, live count), the arrays in a fixed order,
 * and a 64-bit checksum of everything before it. */
vs_status vs_index_save(vs_index* index, const char* path);
/* `overrides` may be NULL; only device, capacity_hint and max_batch are read from it (the index
 * geometry comes from the file). Errors: VS_ERR_INVALID (unreadable / not an index file /
 * truncated / checksum mismatch), VS_ERR_UNSUPPORTED (newer format), VS_ERR_CUDA, VS_ERR_CAPACITY. */
vs_status vs_index_load(const char* path, const vs_options* overrides, vs_index** out);

/* Thread-local message of the last failing call on this thread. */
const char* vs_last_error(void);
/* "vsann-b200 <version> sm_100a" for the GPU library, "vsann-oracle ..." for the oracle. */
const char* vs_version(void);

#ifdef __cplusplus
}
#endif
#endif /* VSANN_H */
