"""Builds libvsann.so (sm_100a only) in-tree with nvcc. No torch, no JIT cache: the .so travels with the repo snapshot."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.environ.get("VSANN_LIB_OUT") or os.path.join(HERE, "lib", "libvsann.so")
SOURCES = ["index.cu", "shard.cu", "graph_kernels.cu", "exact_kernels.cu", "exact_tc.cu", "noalloc.cu"]
HEADERS = ["common.cuh", "beam.cuh", "rbeam.cuh", "tbeam.cuh", "tmap.h", "kernels.h", "shard.cuh", "vmm.h", "flatmap.h", os.path.join("..", "..", "include", "vsann.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = os.environ.get("VSANN_NVCC_FLAGS", "").split() + ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC,-O2,-Wall", "--threads", "4"]


ACTOR_OUT = os.path.join(HERE, "lib", "libvsactor.so")
ACTOR_SRC = os.path.join(CSRC, "actor.cpp")
ACTOR_DEPS = [ACTOR_SRC, os.path.join(HERE, "..", "include", "vsactor.h"), os.path.join(HERE, "..", "include", "vsann.h")]


def build_actor(force: bool = False) -> str:
    """libvsactor.so: the index-actor shim (host C++ only, no CUDA) over any include/vsann.h backend."""
    if not force and os.path.exists(ACTOR_OUT) and all(os.path.getmtime(d) <= os.path.getmtime(ACTOR_OUT) for d in ACTOR_DEPS):
        return ACTOR_OUT
    os.makedirs(os.path.dirname(ACTOR_OUT), exist_ok=True)
    cmd = [os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-fPIC", "-Wall", "-Wextra", "-pthread", "-shared", "-o", ACTOR_OUT, ACTOR_SRC]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("g++ failed:\n" + r.stdout + r.stderr)
    return ACTOR_OUT


LOAD_OUT = os.path.join(HERE, "lib", "actor_load")
LOAD_SRC = os.path.join(CSRC, "actor_load.cpp")


def build_actor_load(force: bool = False) -> str:
    """lib/actor_load: native load generator for the actor (links libvsactor.so, dlopen-s the backend)."""
    deps = [LOAD_SRC, ACTOR_OUT] + ACTOR_DEPS
    if not force and os.path.exists(LOAD_OUT) and all(os.path.getmtime(d) <= os.path.getmtime(LOAD_OUT) for d in deps):
        return LOAD_OUT
    cmd = [os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-Wall", "-pthread", "-o", LOAD_OUT, LOAD_SRC,
           "-L" + os.path.dirname(ACTOR_OUT), "-lvsactor", "-ldl", "-Wl,-rpath,$ORIGIN"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("g++ failed:\n" + r.stdout + r.stderr)
    return LOAD_OUT


def needs_build() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    build_actor(force)
    build_actor_load(force)
    if not force and not needs_build():
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    cmd = [NVCC, *FLAGS, "-shared", "-o", OUT, *[os.path.join(CSRC, s) for s in SOURCES], "-lcudart"]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stderr)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
