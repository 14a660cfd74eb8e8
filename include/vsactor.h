/* vsactor.h — C ABI of the index-actor shim.
 *
 * The reference drives usearch through an actor: `new` (src/index/usearch.rs:82-139) spawns a task
 * that receives `Index::{AddOrReplace, Remove, Ann, Count}` messages over a bounded channel of 10
 * (:102-103), keeps at most 2 x T of them in flight (semaphore, :118) and runs the usearch calls on a
 * pool of T threads (rayon; T = SCYLLA_USEARCH_BACKGROUND_THREADS, README.md:14-15). It owns the
 * PrimaryKey <-> u64 key bimap and the key allocator (:110-113, :181-196). This library is that actor
 * in C++ over ANY implementation of include/vsann.h handed over as a table of function pointers
 * (libvsann.so on the GPU; the CPU oracle in the tests), so the message semantics are testable
 * without a GPU and identical on both.
 *
 * Primary keys are opaque byte strings (the reference's `PrimaryKey` is a Vec<CqlValue>; the
 * caller serialises it). Same ownership rules as the reference's messages: payloads are copied
 * before the call returns; AddOrReplace / Remove are fire-and-forget (no completion signal,
 * usearch.rs:149-158) and become visible through Count / Ann.
 */
#ifndef VSACTOR_H
#define VSACTOR_H

#include "vsann.h"

#ifdef __cplusplus
extern "C" {
#endif

/* The calls the actor makes into the index library (usearch.rs:98-99,201-221,245,276,309). */
typedef struct vs_backend {
    vs_status (*index_create)(const vs_options*, vs_index**);
    void (*index_destroy)(vs_index*);
    vs_status (*add)(vs_index*, uint64_t key, const float* vec);
    vs_status (*remove)(vs_index*, uint64_t key);
    vs_status (*search)(vs_index*, const float* q, size_t k, uint64_t* keys_out, float* dist_out, size_t* found);
    size_t (*count)(vs_index*);
    size_t (*capacity)(vs_index*);
    vs_status (*reserve)(vs_index*, size_t capacity);
    const char* (*last_error)(void);
    int add_replaces;   /* non-zero: add() of a present key replaces it (vsann.h), so the explicit remove of usearch.rs:214-220 is skipped */
} vs_backend;

typedef struct vs_actor vs_actor;
typedef struct vs_matches vs_matches;   /* AnnR = (Vec<PrimaryKey>, Vec<Distance>) — usearch.rs:302 */

/* `new` — usearch.rs:82-139. index_id is only a label ("{keyspace}.{index}"). threads == 0 =>
 * SCYLLA_USEARCH_BACKGROUND_THREADS, else all cores. */
vs_status vs_actor_spawn(const vs_backend* backend, const char* index_id, const vs_options* options, size_t threads, vs_actor** out);
/* Dropping the last Sender<Index>: queued messages are processed, then the index is freed — usearch.rs:120,133. */
void vs_actor_close(vs_actor* actor);

/* Index::AddOrReplace — usearch.rs:149-154,174-233. Blocks only while the channel (10) is full. */
vs_status vs_actor_add_or_replace(vs_actor* actor, const void* pk, size_t pk_len, const float* embedding, size_t dims);
/* Index::Remove — usearch.rs:156-158,235-249. */
vs_status vs_actor_remove(vs_actor* actor, const void* pk, size_t pk_len);
/* Index::Ann — usearch.rs:160-166,251-306. Waits for the reply. On error returns != VS_OK and
 * vs_actor_last_error() holds the reference's message ("ann: wrong embedding dimensions: 2 != 3", ...). */
vs_status vs_actor_ann(vs_actor* actor, const float* embedding, size_t dims, size_t limit, vs_matches** out);
/* Index::Count — usearch.rs:168-170,308-311. */
vs_status vs_actor_count(vs_actor* actor, size_t* out);

size_t vs_matches_len(const vs_matches* m);
const void* vs_matches_pk(const vs_matches* m, size_t i, size_t* len);
float vs_matches_distance(const vs_matches* m, size_t i);
void vs_matches_free(vs_matches* m);

/* Counters for tests and monitoring. */
typedef struct vs_actor_stats {
    uint64_t messages;          /* messages processed                                    */
    uint64_t max_in_flight;     /* high-water mark of concurrently processed messages (<= 2T) */
    uint64_t max_queued;        /* high-water mark of the channel (<= 10)                */
    uint64_t rollbacks;         /* failed adds whose key was taken out of the bimap again (usearch.rs:230-232) */
    uint64_t keys;              /* entries in the bimap                                   */
    uint64_t next_key;          /* the key allocator                                      */
    uint64_t threads;
} vs_actor_stats;
vs_status vs_actor_get_stats(vs_actor* actor, vs_actor_stats* out);
const char* vs_actor_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
