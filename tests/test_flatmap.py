"""CPU test of csrc/flatmap.h (the key -> slot map behind vs_add / vs_remove, src/index/usearch.rs:181-196,214-225,240): a fuzz against
std::unordered_map - put / get / erase with backward-shift deletion, growth, update_each - compiled with g++ at test time."""
import os
import subprocess

from conftest import ROOT

SRC = r"""
#include <cstdio>
#include <random>
#include <unordered_map>
#include "flatmap.h"
using vsann::FlatMap;
int main() {
    std::mt19937_64 rng(12345);
    for (int round = 0; round < 6; ++round) {
        FlatMap m; std::unordered_map<uint64_t, uint32_t> ref;
        const uint64_t space = round < 2 ? 64 : (round < 4 ? 5000 : 2000000);   // tiny key spaces force long probe chains and many erasures
        const int ops = round < 4 ? 200000 : 600000;
        for (int i = 0; i < ops; ++i) {
            const uint64_t k = (rng() % space) * (round % 2 ? 0x9E3779B97F4A7C15ull : 1ull);   // odd rounds: scattered 64-bit keys
            const uint32_t v = (uint32_t)(rng() & 0x3FFFFFFF);
            switch (rng() % 4) {
                case 0: case 1: {
                    const uint32_t old = m.put(k, v);
                    auto it = ref.find(k);
                    const uint32_t want = it == ref.end() ? FlatMap::NO_SLOT : it->second;
                    if (old != want) { printf("put mismatch round %d op %d\n", round, i); return 1; }
                    ref[k] = v; break;
                }
                case 2: {
                    const uint32_t got = m.erase(k);
                    auto it = ref.find(k);
                    const uint32_t want = it == ref.end() ? FlatMap::NO_SLOT : it->second;
                    if (got != want) { printf("erase mismatch round %d op %d\n", round, i); return 1; }
                    if (it != ref.end()) ref.erase(it);
                    break;
                }
                default: {
                    auto it = ref.find(k);
                    if (m.get(k) != (it == ref.end() ? FlatMap::NO_SLOT : it->second)) { printf("get mismatch round %d op %d\n", round, i); return 1; }
                }
            }
            if (m.size() != ref.size()) { printf("size mismatch round %d op %d\n", round, i); return 1; }
        }
        m.update_each([](uint64_t, uint32_t s) { return s ^ 1u; });           // compaction's remap
        size_t seen = 0; bool ok = true;
        m.for_each([&](uint64_t k, uint32_t s) { ++seen; auto it = ref.find(k); ok = ok && it != ref.end() && (it->second ^ 1u) == s; });
        if (!ok || seen != ref.size()) { printf("for_each / update_each mismatch round %d\n", round); return 1; }
        for (auto& kv : ref) if (m.get(kv.first) != (kv.second ^ 1u)) { printf("final get mismatch round %d\n", round); return 1; }
    }
    printf("ok\n");
    return 0;
}
"""


def test_flatmap_against_unordered_map(tmp_path):
    src = tmp_path / "flatmap_fuzz.cpp"
    src.write_text(SRC)
    exe = tmp_path / "flatmap_fuzz"
    inc = os.path.join(ROOT, "vector-store-text_b200", "csrc")
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", inc, "-o", str(exe), str(src)], check=True, capture_output=True)
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == "ok", r.stdout + r.stderr
