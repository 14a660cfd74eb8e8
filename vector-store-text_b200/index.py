"""Host-side mirror of the reference's index actor for the vector path.

Mirrors, name for name, what callers of the reference see:
  * `IndexFactory.create_index(id, dimensions, connectivity, expansion_add, expansion_search)`
    — `UsearchIndexFactory::create_index`, src/index/usearch.rs:38-53 (5-argument form).
  * `IndexActor.add_or_replace / remove / ann / count` — the vector form of `trait IndexExt`
    (src/index/actor.rs:29-59 in its text form; used as at src/index/usearch.rs:334,353,361,401 and
    src/monitor_items.rs:75,77), served by `process` (src/index/usearch.rs:141-172).
  * primary keys are opaque hashables; the actor allocates dense u64 keys and keeps the
    PrimaryKey <-> Key bimap on the host (src/index/usearch.rs:110-113,181-196,240,285-296).
The arithmetic is behind the C ABI (include/vsann.h); this file holds no vector math.
"""
import threading

import numpy as np

from ._ffi import Metric, VsError, VsIndex, VS_ERR_DIMS


class AnnError(RuntimeError):
    """The anyhow::Error strings of `ann` (src/index/usearch.rs:259-296)."""


class IndexActor:
    def __init__(self, index_id, dimensions, connectivity=0, expansion_add=0, expansion_search=0,
                 metric=Metric.COS, device=0, capacity_hint=0, max_batch=0, lib=None):
        if dimensions <= 0:
            raise ValueError("dimensions must be non-zero (lib.rs:147 Dimensions(NonZeroUsize))")
        self.id = index_id
        self.dimensions = int(dimensions)
        self._ix = VsIndex(self.dimensions, metric=metric, connectivity=connectivity, expansion_add=expansion_add,
                           expansion_search=expansion_search, capacity_hint=capacity_hint, device=device,
                           max_batch=max_batch, lib=lib)
        self._lock = threading.RLock()          # keys: Arc<RwLock<BiMap>>  (usearch.rs:110)
        self._pk_to_key = {}
        self._key_to_pk = {}
        self._next_key = 0                      # usearch_key: AtomicU64     (usearch.rs:113)

    # Index::AddOrReplace — usearch.rs:174-233. Fire-and-forget: no reply, failures roll the bimap back.
    def add_or_replace(self, primary_key, embedding):
        emb = np.ascontiguousarray(embedding, dtype=np.float32)
        with self._lock:
            key = self._pk_to_key.get(primary_key)
            fresh = key is None
            if fresh:                            # insert_no_overwrite succeeded (usearch.rs:183-189)
                key = self._next_key
                self._next_key += 1
                self._pk_to_key[primary_key] = key
                self._key_to_pk[key] = primary_key
            try:
                self._ix.add(key, emb)           # replace = remove + add under the same key (usearch.rs:214-225)
            except VsError:
                if fresh:                        # keys.remove_by_right(&key) (usearch.rs:230-232)
                    del self._pk_to_key[primary_key]
                    del self._key_to_pk[key]

    # Index::Remove — usearch.rs:235-249.
    def remove(self, primary_key):
        with self._lock:
            key = self._pk_to_key.pop(primary_key, None)
            if key is None:
                return
            del self._key_to_pk[key]
            self._ix.remove(key)

    # Index::Ann — usearch.rs:251-306. Returns (primary_keys, distances), ascending by distance.
    def ann(self, embedding, limit=1):
        emb = np.ascontiguousarray(embedding, dtype=np.float32).reshape(-1)
        if emb.size == 0:
            raise AnnError("ann: embedding dimensions == 0")
        if emb.size != self.dimensions:
            raise AnnError(f"ann: wrong embedding dimensions: {emb.size} != {self.dimensions}")
        if limit <= 0:
            raise ValueError("limit must be non-zero (lib.rs:235 Limit(NonZeroUsize))")
        try:
            keys, dists = self._ix.search(emb, int(limit))
        except VsError as e:
            if e.status == VS_ERR_DIMS:
                raise AnnError(f"ann: wrong embedding dimensions: {emb.size} != {self.dimensions}")
            raise AnnError(f"ann: search failed: {e.message}")
        with self._lock:
            pks = []
            for key in keys.tolist():
                pk = self._key_to_pk.get(key)
                if pk is None:
                    raise AnnError(f"not defined primary key column {key}")
                pks.append(pk)
        return pks, dists.tolist()

    # Index::Count — usearch.rs:308-311.
    def count(self):
        return self._ix.count()

    def flush(self):
        self._ix.flush()

    @property
    def raw(self):
        return self._ix

    def close(self):
        self._ix.close()


class IndexFactory:
    """`trait IndexFactory` (src/index/factory.rs:10-12) in its vector form (usearch.rs:38-45)."""

    def __init__(self, metric=Metric.COS, device=0, capacity_hint=0, lib=None):
        self.metric, self.device, self.capacity_hint, self.lib = metric, device, capacity_hint, lib

    def create_index(self, index_id, dimensions, connectivity=0, expansion_add=0, expansion_search=0):
        return IndexActor(index_id, dimensions, connectivity, expansion_add, expansion_search, metric=self.metric,
                          device=self.device, capacity_hint=self.capacity_hint, lib=self.lib)
