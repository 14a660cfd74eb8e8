"""ctypes binding of include/vsactor.h: the C++ index actor (csrc/actor.cpp) over a vsann backend.

`NativeActor` has the surface of `IndexActor` (index.py) — add_or_replace / remove / ann / count — but the
queue (10), the 2 x T in-flight limit, the worker pool and the PrimaryKey <-> key bimap live in C++, as in
src/index/usearch.rs:82-311. Primary keys cross the boundary as bytes (pickled here)."""
import ctypes as C
import os
import pickle

import numpy as np

from ._ffi import VsOptions, load_library, Metric, VS_OK
from .index import AnnError

_FN = {
    "index_create": C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p),
    "index_destroy": C.CFUNCTYPE(None, C.c_void_p),
    "add": C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_uint64, C.c_void_p),
    "remove": C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_uint64),
    "search": C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p),
    "count": C.CFUNCTYPE(C.c_size_t, C.c_void_p),
    "capacity": C.CFUNCTYPE(C.c_size_t, C.c_void_p),
    "reserve": C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_size_t),
    "last_error": C.CFUNCTYPE(C.c_char_p),
}


class VsBackend(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in _FN] + [("add_replaces", C.c_int)]


class VsActorStats(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("messages", "max_in_flight", "max_queued", "rollbacks", "keys", "next_key", "threads")]


ACTOR_SYMBOLS = ["vs_actor_spawn", "vs_actor_close", "vs_actor_add_or_replace", "vs_actor_remove", "vs_actor_ann", "vs_actor_count",
                 "vs_matches_len", "vs_matches_pk", "vs_matches_distance", "vs_matches_free", "vs_actor_get_stats", "vs_actor_last_error"]

_alib = None


def actor_lib_path():
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libvsactor.so")


def load_actor_library():
    global _alib
    if _alib is None:
        if not os.path.exists(actor_lib_path()):
            raise FileNotFoundError(f"{actor_lib_path()} is missing - run `python vector-store-text_b200/build.py`")
        a = C.CDLL(actor_lib_path())
        a.vs_actor_spawn.restype = C.c_int
        a.vs_actor_spawn.argtypes = [C.POINTER(VsBackend), C.c_char_p, C.POINTER(VsOptions), C.c_size_t, C.POINTER(C.c_void_p)]
        a.vs_actor_close.restype = None; a.vs_actor_close.argtypes = [C.c_void_p]
        a.vs_actor_add_or_replace.restype = C.c_int
        a.vs_actor_add_or_replace.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t, C.c_void_p, C.c_size_t]
        a.vs_actor_remove.restype = C.c_int; a.vs_actor_remove.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
        a.vs_actor_ann.restype = C.c_int
        a.vs_actor_ann.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.POINTER(C.c_void_p)]
        a.vs_actor_count.restype = C.c_int; a.vs_actor_count.argtypes = [C.c_void_p, C.POINTER(C.c_size_t)]
        a.vs_matches_len.restype = C.c_size_t; a.vs_matches_len.argtypes = [C.c_void_p]
        a.vs_matches_pk.restype = C.c_void_p; a.vs_matches_pk.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
        a.vs_matches_distance.restype = C.c_float; a.vs_matches_distance.argtypes = [C.c_void_p, C.c_size_t]
        a.vs_matches_free.restype = None; a.vs_matches_free.argtypes = [C.c_void_p]
        a.vs_actor_get_stats.restype = C.c_int; a.vs_actor_get_stats.argtypes = [C.c_void_p, C.POINTER(VsActorStats)]
        a.vs_actor_last_error.restype = C.c_char_p
        _alib = a
    return _alib


def backend_of(lib):
    """The function table of a loaded vsann library (GPU `vs_` or oracle `ovs_`)."""
    b = VsBackend()
    for name in _FN:
        setattr(b, name, C.cast(getattr(lib, lib._vs_prefix + name), C.c_void_p).value)
    b.add_replaces = 1      # vsann.h: adding a present key replaces it
    return b


class NativeActor:
    def __init__(self, index_id, dimensions, connectivity=0, expansion_add=0, expansion_search=0, metric=Metric.COS,
                 device=0, capacity_hint=0, max_batch=0, threads=0, lib=None):
        self.a = load_actor_library()
        self.backend_lib = lib or load_library()
        self.id, self.dimensions = index_id, int(dimensions)
        self._be = backend_of(self.backend_lib)
        opt = VsOptions(dims=self.dimensions, metric=int(metric), connectivity=connectivity, expansion_add=expansion_add,
                        expansion_search=expansion_search, capacity_hint=capacity_hint, device=device, max_batch=max_batch)
        h = C.c_void_p()
        st = self.a.vs_actor_spawn(C.byref(self._be), str(index_id).encode(), C.byref(opt), threads, C.byref(h))
        if st != VS_OK:
            raise RuntimeError(self.a.vs_actor_last_error().decode())
        self._h = h

    @staticmethod
    def _pk(primary_key):
        return pickle.dumps(primary_key, protocol=4)

    def add_or_replace(self, primary_key, embedding):
        e = np.ascontiguousarray(embedding, dtype=np.float32)
        pk = self._pk(primary_key)
        self.a.vs_actor_add_or_replace(self._h, pk, len(pk), e.ctypes.data_as(C.c_void_p), e.size)   # fire-and-forget

    def remove(self, primary_key):
        pk = self._pk(primary_key)
        self.a.vs_actor_remove(self._h, pk, len(pk))

    def ann(self, embedding, limit):
        e = np.ascontiguousarray(embedding, dtype=np.float32)
        if limit <= 0:
            raise ValueError("limit must be non-zero (lib.rs:235 Limit(NonZeroUsize))")
        m = C.c_void_p()
        st = self.a.vs_actor_ann(self._h, e.ctypes.data_as(C.c_void_p), e.size, int(limit), C.byref(m))
        if st != VS_OK:
            raise AnnError(self.a.vs_actor_last_error().decode())
        try:
            n = self.a.vs_matches_len(m)
            pks, dist = [], []
            for i in range(n):
                ln = C.c_size_t()
                p = self.a.vs_matches_pk(m, i, C.byref(ln))
                pks.append(pickle.loads(C.string_at(p, ln.value)))
                dist.append(self.a.vs_matches_distance(m, i))
            return pks, dist
        finally:
            self.a.vs_matches_free(m)

    def count(self):
        n = C.c_size_t()
        st = self.a.vs_actor_count(self._h, C.byref(n))
        if st != VS_OK:
            raise RuntimeError(self.a.vs_actor_last_error().decode())
        return n.value

    def stats(self):
        s = VsActorStats()
        self.a.vs_actor_get_stats(self._h, C.byref(s))
        return {k: getattr(s, k) for k, _ in s._fields_}

    def close(self):
        if self._h is not None:
            self.a.vs_actor_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
