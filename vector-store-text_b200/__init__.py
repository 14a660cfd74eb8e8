"""vector-store-text_b200 — B200-native ANN index behind the reference's index-actor API.

Only the hot path lives here: `csrc/` (sm_100a CUDA kernels + the C ABI of include/vsann.h) and
the host-side mirror of the reference's interface for that path (`index.py`: the vector form of
`enum Index` / `trait IndexExt`, src/index/actor.rs:13-59 + src/index/usearch.rs:141-311).
"""
from ._ffi import (Metric, VsError, VsIndex, lib_path, load_library, merge_topk_device, merge_topk_peers_device,  # noqa: F401
                   VS_OK, VS_ERR_INVALID, VS_ERR_DIMS, VS_NOT_FOUND, VS_ERR_CAPACITY, VS_ERR_CUDA,
                   VS_ERR_NOMEM, VS_ERR_UNSUPPORTED, ABI_SYMBOLS)
from .index import IndexActor, IndexFactory, AnnError  # noqa: F401
from .actor_native import NativeActor, load_actor_library, actor_lib_path, ACTOR_SYMBOLS  # noqa: F401
from .service import DbEmbedding, ItemsMonitor, Engine, HttpServer  # noqa: F401

__all__ = ["Metric", "VsError", "VsIndex", "IndexActor", "IndexFactory", "AnnError", "lib_path",
           "load_library", "merge_topk_device", "merge_topk_peers_device", "NativeActor", "load_actor_library", "DbEmbedding", "ItemsMonitor",
           "Engine", "HttpServer"]
