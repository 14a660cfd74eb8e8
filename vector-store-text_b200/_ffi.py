"""ctypes binding of include/vsann.h. The same class binds the GPU library (prefix "vs_") and,
for tests only, the CPU oracle (prefix "ovs_", oracle/liboracle_vsann.so)."""
import ctypes as C
import enum
import os

import numpy as np

VS_OK, VS_ERR_INVALID, VS_ERR_DIMS, VS_NOT_FOUND, VS_ERR_CAPACITY, VS_ERR_CUDA, VS_ERR_NOMEM, VS_ERR_UNSUPPORTED = range(8)

ABI_SYMBOLS = [
    "vs_index_create", "vs_index_destroy", "vs_add", "vs_add_batch", "vs_add_batch_device", "vs_remove",
    "vs_search", "vs_search_batch", "vs_search_batch_device", "vs_search_batch_fanout_device", "vs_count", "vs_capacity", "vs_reserve",
    "vs_flush", "vs_get_stats", "vs_set_stats", "vs_reset_stats", "vs_merge_topk_device", "vs_merge_topk_peers_device", "vs_export_level",
    "vs_merge_topk_flags_device", "vs_exact_search_noalloc_device", "vs_index_save", "vs_index_load", "vs_last_error", "vs_version",
]


class Metric(enum.IntEnum):
    COS = 0
    L2SQ = 1
    IP = 2


class VsOptions(C.Structure):
    _fields_ = [("dims", C.c_size_t), ("metric", C.c_int32), ("connectivity", C.c_size_t),
                ("expansion_add", C.c_size_t), ("expansion_search", C.c_size_t), ("capacity_hint", C.c_size_t),
                ("device", C.c_int32), ("max_batch", C.c_size_t), ("n_devices", C.c_uint8), ("device_ids", C.c_uint8 * 8),
                ("reserved", C.c_uint8 * 23)]


class VsStats(C.Structure):
    _fields_ = [("count", C.c_uint64), ("slots", C.c_uint64), ("capacity", C.c_uint64), ("pending", C.c_uint64),
                ("search_queries", C.c_uint64), ("search_dist_evals", C.c_uint64), ("search_expansions", C.c_uint64),
                ("build_vectors", C.c_uint64), ("build_dist_evals", C.c_uint64), ("build_batches", C.c_uint64),
                ("build_seconds", C.c_double), ("device_bytes", C.c_uint64), ("top_level", C.c_uint32),
                ("dims", C.c_uint32), ("exact_tc_queries", C.c_uint64), ("exact_tc_fallback", C.c_uint64),
                ("exact_tc_second_scan", C.c_uint64), ("exact_tc_scans", C.c_uint64), ("exact_tc_scan_seconds", C.c_double),
                ("exact_stream_queries", C.c_uint64), ("search_host_queries", C.c_uint64), ("search_host_launches", C.c_uint64),
                ("build_expansions", C.c_uint64), ("reserve_calls", C.c_uint64), ("reserve_seconds", C.c_double),
                ("compactions", C.c_uint64), ("compact_seconds", C.c_double), ("tables_grow_in_place", C.c_uint32), ("n_devices", C.c_uint32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class VsError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"vsann status {status}: {message}")
        self.status = status
        self.message = message


def lib_path() -> str:
    if os.environ.get("VSANN_LIB"):          # developer hook: A/B a differently compiled libvsann.so
        return os.environ["VSANN_LIB"]
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libvsann.so")


_libs = {}


def load_library(path: str = None, prefix: str = "vs_"):
    """Loads a library exporting the vsann ABI. The GPU library must exist: there is no fallback."""
    path = path or lib_path()
    key = (path, prefix)
    if key in _libs:
        return _libs[key]
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} is missing - run `python vector-store-text_b200/build.py` "
                                "(the CUDA library is mandatory; there is no CPU fallback)")
    lib = C.CDLL(path)
    p = prefix
    f = lambda name: getattr(lib, p + name)  # noqa: E731
    vp, u64p, f32p, szp = C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p
    sig = {
        "index_create": (C.c_int, [C.POINTER(VsOptions), C.POINTER(vp)]),
        "index_destroy": (None, [vp]),
        "add": (C.c_int, [vp, C.c_uint64, f32p]),
        "add_batch": (C.c_int, [vp, C.c_size_t, u64p, f32p]),
        "remove": (C.c_int, [vp, C.c_uint64]),
        "search": (C.c_int, [vp, f32p, C.c_size_t, u64p, f32p, szp]),
        "search_batch": (C.c_int, [vp, C.c_size_t, f32p, C.c_size_t, C.c_size_t, C.c_int, u64p, f32p, szp]),
        "count": (C.c_size_t, [vp]),
        "capacity": (C.c_size_t, [vp]),
        "reserve": (C.c_int, [vp, C.c_size_t]),
        "flush": (C.c_int, [vp]),
        "get_stats": (C.c_int, [vp, C.POINTER(VsStats)]),
        "set_stats": (C.c_int, [vp, C.c_int]),
        "reset_stats": (C.c_int, [vp]),
        "last_error": (C.c_char_p, []),
        "version": (C.c_char_p, []),
    }
    if prefix == "vs_":
        sig.update({
            "add_batch_device": (C.c_int, [vp, C.c_size_t, u64p, vp]),
            "search_batch_device": (C.c_int, [vp, C.c_size_t, vp, C.c_size_t, C.c_size_t, C.c_int, vp, vp, vp]),
            "search_batch_fanout_device": (C.c_int, [vp, C.c_size_t, vp, C.c_size_t, C.c_size_t, C.c_size_t, vp, vp, vp]),
            "merge_topk_device": (C.c_int, [C.c_int, C.c_size_t, C.c_size_t, C.c_size_t, vp, vp, vp, vp, vp]),
            "merge_topk_peers_device": (C.c_int, [C.c_int, C.c_size_t, C.c_size_t, C.c_size_t, vp, vp, vp, vp, vp]),
            "merge_topk_flags_device": (C.c_int, [C.c_int, C.c_size_t, C.c_size_t, C.c_size_t, vp, vp, vp, C.c_uint32, C.c_uint32, vp, vp, vp]),
            "exact_search_noalloc_device": (C.c_int, [C.c_int, C.c_int32, C.c_size_t, C.c_size_t, vp, vp, C.c_size_t, vp, C.c_size_t, vp, vp, vp]),
            "export_level": (C.c_int, [vp, C.c_uint32, vp, szp, szp, vp, szp]),
            "index_save": (C.c_int, [vp, C.c_char_p]),
            "index_load": (C.c_int, [C.c_char_p, C.POINTER(VsOptions), C.POINTER(vp)]),
        })
    else:
        sig.update({
            "set_threads": (C.c_int, [vp, C.c_uint]),
            "set_seed": (C.c_int, [vp, C.c_uint64]),
            "set_simd": (C.c_int, [vp, C.c_int]),
            "distance": (C.c_float, [C.c_int, f32p, f32p, C.c_size_t]),
            "graph_search": (C.c_int, [C.c_int, C.c_size_t, C.c_size_t, f32p, vp, C.c_size_t, vp, C.c_uint32, f32p,
                                       C.c_size_t, C.c_size_t, vp, f32p, szp, szp, szp]),
        })
    for name, (res, args) in sig.items():
        fn = f(name)
        fn.restype = res
        fn.argtypes = args
    lib._vs_prefix = prefix
    _libs[key] = lib
    return lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class VsIndex:
    """One index = one GPU shard (or one oracle instance). Thin, typed wrapper over the C ABI."""

    def __init__(self, dims, metric=Metric.COS, connectivity=0, expansion_add=0, expansion_search=0,
                 capacity_hint=0, device=0, max_batch=0, lib=None, devices=None):
        """devices: list of CUDA ordinals => ONE index over several GPUs in this process (vs_options.n_devices / device_ids)."""
        self.lib = lib or load_library()
        self.p = self.lib._vs_prefix
        self.dims = int(dims)
        opt = VsOptions(dims=self.dims, metric=int(metric), connectivity=connectivity, expansion_add=expansion_add,
                        expansion_search=expansion_search, capacity_hint=capacity_hint, device=device, max_batch=max_batch)
        if devices:
            opt.n_devices = len(devices)
            for i, dv in enumerate(devices):
                opt.device_ids[i] = int(dv)
        h = C.c_void_p()
        self._h = None
        self._check(self._fn("index_create")(C.byref(opt), C.byref(h)))
        self._h = h

    def _fn(self, name):
        return getattr(self.lib, self.p + name)

    def _check(self, status, ok=(VS_OK,)):
        if status not in ok:
            raise VsError(status, self._fn("last_error")().decode())
        return status

    def close(self):
        if self._h is not None:
            self._fn("index_destroy")(self._h)
            self._h = None

    # -- persistence (include/vsann.h: vs_index_save / vs_index_load; absent in the reference) ----
    def save(self, path):
        self._check(self._fn("index_save")(self._h, os.fsencode(path)))

    @classmethod
    def load(cls, path, device=0, capacity_hint=0, max_batch=0, lib=None):
        self = cls.__new__(cls)
        self.lib = lib or load_library()
        self.p = self.lib._vs_prefix
        self._h = None
        opt = VsOptions(dims=0, metric=0, connectivity=0, expansion_add=0, expansion_search=0,
                        capacity_hint=capacity_hint, device=device, max_batch=max_batch)
        h = C.c_void_p()
        self._check(self._fn("index_load")(os.fsencode(path), C.byref(opt), C.byref(h)))
        self._h = h
        st = VsStats()
        self._check(self._fn("get_stats")(self._h, C.byref(st)))
        self.dims = int(st.dims)
        return self

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- mutation ---------------------------------------------------------------------------
    def add(self, key, vec):
        v = _f32(vec)
        if v.shape != (self.dims,):
            raise VsError(VS_ERR_DIMS, f"add: wrong embedding dimensions: {v.size} != {self.dims}")
        self._check(self._fn("add")(self._h, int(key), _ptr(v)))

    def add_batch(self, keys, vecs):
        k = np.ascontiguousarray(keys, dtype=np.uint64)
        v = _f32(vecs)
        if v.ndim != 2 or v.shape[1] != self.dims or v.shape[0] != k.shape[0]:
            raise VsError(VS_ERR_DIMS, f"add_batch: expected ({k.shape[0]}, {self.dims}) got {v.shape}")
        self._check(self._fn("add_batch")(self._h, k.shape[0], _ptr(k), _ptr(v)))

    def add_batch_device(self, keys, dev_ptr, n):
        k = np.ascontiguousarray(keys, dtype=np.uint64)
        self._check(self._fn("add_batch_device")(self._h, int(n), _ptr(k), C.c_void_p(int(dev_ptr))))

    def remove(self, key):
        """Returns False when the key is unknown (VS_NOT_FOUND is non-fatal)."""
        return self._check(self._fn("remove")(self._h, int(key)), ok=(VS_OK, VS_NOT_FOUND)) == VS_OK

    def flush(self):
        self._check(self._fn("flush")(self._h))

    def reserve(self, capacity):
        self._check(self._fn("reserve")(self._h, int(capacity)))

    # -- queries ----------------------------------------------------------------------------
    def search(self, queries, k, ef=0, exact=False):
        q = _f32(queries)
        single = q.ndim == 1
        q2 = q.reshape(1, -1) if single else q
        if q2.shape[1] != self.dims:
            raise VsError(VS_ERR_DIMS, f"ann: wrong embedding dimensions: {q2.shape[1]} != {self.dims}")
        nq = q2.shape[0]
        keys = np.empty((nq, k), dtype=np.uint64)
        dist = np.empty((nq, k), dtype=np.float32)
        found = np.zeros(nq, dtype=np.uintp)
        self._check(self._fn("search_batch")(self._h, nq, _ptr(q2), k, ef, 1 if exact else 0, _ptr(keys), _ptr(dist), _ptr(found)))
        if single:
            return keys[0, :found[0]], dist[0, :found[0]]
        return keys, dist, found

    def search_one(self, query, k):
        """vs_search: one query, the library's default ef; concurrent callers share launches (include/vsann.h)."""
        q = _f32(query)
        if q.shape != (self.dims,):
            raise VsError(VS_ERR_DIMS, f"ann: wrong embedding dimensions: {q.size} != {self.dims}")
        keys = np.empty(k, dtype=np.uint64)
        dist = np.empty(k, dtype=np.float32)
        found = C.c_size_t()
        self._check(self._fn("search")(self._h, _ptr(q), k, _ptr(keys), _ptr(dist), C.byref(found)))
        return keys[:found.value], dist[:found.value]

    def search_device(self, nq, q_ptr, k, keys_ptr, dist_ptr, ef=0, exact=False, stream=0):
        self._check(self._fn("search_batch_device")(self._h, int(nq), C.c_void_p(int(q_ptr)), int(k), int(ef), 1 if exact else 0,
                                                    C.c_void_p(int(keys_ptr)), C.c_void_p(int(dist_ptr)),
                                                    C.c_void_p(int(stream)) if stream else None))

    def search_fanout_device(self, nq, q_ptr, k, keys_ptrs, dist_ptrs, ef=0, stream=0):
        """Graph search whose result lists go to every (keys, dists) buffer pair: keys_ptrs / dist_ptrs are sequences of device-
        writable addresses (the first pair on the index's device, the others e.g. peer GPUs' memory)."""
        n = len(keys_ptrs)
        ka = (C.c_void_p * n)(*[int(p) for p in keys_ptrs]); da = (C.c_void_p * n)(*[int(p) for p in dist_ptrs])
        self._check(self._fn("search_batch_fanout_device")(self._h, int(nq), C.c_void_p(int(q_ptr)), int(k), int(ef), n, ka, da,
                                                           C.c_void_p(int(stream)) if stream else None))

    def count(self):
        return int(self._fn("count")(self._h))

    def capacity(self):
        return int(self._fn("capacity")(self._h))

    def stats(self):
        s = VsStats()
        self._check(self._fn("get_stats")(self._h, C.byref(s)))
        return s.as_dict()

    def set_stats(self, on=True):
        self._check(self._fn("set_stats")(self._h, 1 if on else 0))

    def reset_stats(self):
        self._check(self._fn("reset_stats")(self._h))

    def export_level(self, level=0):
        nr, w, ns = C.c_size_t(), C.c_size_t(), C.c_size_t()
        self._check(self._fn("export_level")(self._h, level, None, C.byref(nr), C.byref(w), None, C.byref(ns)))
        links = np.empty((nr.value, w.value), dtype=np.uint32)
        keys = np.empty(ns.value, dtype=np.uint64)
        self._check(self._fn("export_level")(self._h, level, _ptr(links), C.byref(nr), C.byref(w), _ptr(keys), C.byref(ns)))
        return links, keys

    # oracle-only helpers
    def set_threads(self, t):
        self._check(self._fn("set_threads")(self._h, int(t)))

    def set_simd(self, on=True):
        """Oracle only: SIMD distances in its HNSW paths (bench.py's timed CPU baseline; parity tests keep the sequential chain)."""
        self._check(self._fn("set_simd")(self._h, 1 if on else 0))

    def set_seed(self, s):
        self._check(self._fn("set_seed")(self._h, int(s)))

    def version(self):
        return self._fn("version")().decode()


def merge_topk_device(device, parts, nq, k, keys_in_ptr, dist_in_ptr, keys_out_ptr, dist_out_ptr, stream=0, lib=None):
    lib = lib or load_library()
    st = lib.vs_merge_topk_device(int(device), int(parts), int(nq), int(k), C.c_void_p(int(keys_in_ptr)), C.c_void_p(int(dist_in_ptr)),
                                  C.c_void_p(int(keys_out_ptr)), C.c_void_p(int(dist_out_ptr)), C.c_void_p(int(stream)) if stream else None)
    if st != VS_OK:
        raise VsError(st, lib.vs_last_error().decode())


def merge_topk_flags_device(device, parts, nq, k, part_keys_ptr, part_dist_ptr, part_flags_ptr, my_part, epoch, keys_out_ptr, dist_out_ptr, stream=0, lib=None):
    """Fused barrier + gather + merge: as merge_topk_peers_device, with each rank's flag word (a DEVICE array of `parts` pointers)."""
    lib = lib or load_library()
    st = lib.vs_merge_topk_flags_device(int(device), int(parts), int(nq), int(k), C.c_void_p(int(part_keys_ptr)), C.c_void_p(int(part_dist_ptr)),
                                        C.c_void_p(int(part_flags_ptr)), int(my_part), int(epoch) & 0xFFFFFFFF,
                                        C.c_void_p(int(keys_out_ptr)), C.c_void_p(int(dist_out_ptr)), C.c_void_p(int(stream)) if stream else None)
    if st != VS_OK:
        raise VsError(st, lib.vs_last_error().decode())


def merge_topk_peers_device(device, parts, nq, k, part_keys_ptr, part_dist_ptr, keys_out_ptr, dist_out_ptr, stream=0, lib=None):
    """Fused gather + merge: part_keys_ptr / part_dist_ptr are DEVICE arrays of `parts` pointers (peer memory allowed)."""
    lib = lib or load_library()
    st = lib.vs_merge_topk_peers_device(int(device), int(parts), int(nq), int(k), C.c_void_p(int(part_keys_ptr)), C.c_void_p(int(part_dist_ptr)),
                                        C.c_void_p(int(keys_out_ptr)), C.c_void_p(int(dist_out_ptr)), C.c_void_p(int(stream)) if stream else None)
    if st != VS_OK:
        raise VsError(st, lib.vs_last_error().decode())
