// Stand-alone probe of the per-lane cp.async.bulk + mbarrier pattern used by tbeam.cuh.
#include <cstdio>
#include <vector>
#include "../../vector-store-text_b200/csrc/tbeam.cuh"
using namespace vsann;
__global__ void k_probe(const float* vecs, uint32_t dpad, uint32_t nrows, uint32_t rows, float* out, int mode) {
    extern __shared__ __align__(128) unsigned char smem[];
    const unsigned warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const size_t per_warp = tsmem_bytes(dpad, 1024, rows);
    WarpSmem S; WarpStage T;
    carve_tsmem(smem + warp * per_warp, dpad, 1024, rows, S, T);
    if (mode == 0) return;
    const uint32_t n = min(rows, 32u);
    for (int it = 0; it < 3; ++it) {
        if (lane == 0) mbar_arrive_expect_tx(T.bar, n * T.row_bytes);
        if (lane < n) {
            const uint32_t rid = (blockIdx.x * 131 + warp * 17 + lane * 7 + it) % nrows;
            if (mode >= 2) bulk_g2s(T.buf_s + lane * T.row_bytes, vecs + (size_t)rid * dpad, T.row_bytes, T.bar);
        }
        if (mode == 1) { if (lane == 0) asm volatile("mbarrier.complete_tx.shared::cta.b64 [%0], %1;" ::"r"(T.bar), "r"(n * T.row_bytes) : "memory"); }
        mbar_wait(T.bar, T.parity); T.parity ^= 1u;
        float s = 0.f;
        if (lane < n && mode >= 2) { const float* r = (const float*)(T.buf + lane * T.row_bytes); for (uint32_t c = 0; c < dpad; ++c) s += r[c]; }
        if (lane < n) out[((blockIdx.x * 4 + warp) * 3 + it) * 32 + lane] = s;
        __syncwarp();
    }
}
int main(int argc, char** argv) {
    const uint32_t dpad = argc > 1 ? atoi(argv[1]) : 128, nrows = 5000, rows = argc > 2 ? atoi(argv[2]) : 16;
    std::vector<float> h((size_t)nrows * dpad); for (size_t i = 0; i < h.size(); ++i) h[i] = (float)((i * 2654435761u) % 1000) / 1000.f;
    float *d, *o; cudaMalloc(&d, h.size() * 4); cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    const int blocks = 64; cudaMalloc(&o, blocks * 4 * 3 * 32 * 4); cudaMemset(o, 0, blocks * 4 * 3 * 32 * 4);
    const size_t smem = tsmem_bytes(dpad, 1024, rows) * 4;
    cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    for (int mode = 0; mode <= 2; ++mode) {
        k_probe<<<blocks, 128, smem>>>(d, dpad, nrows, rows, o, mode);
        cudaError_t e = cudaDeviceSynchronize();
        printf("mode %d smem %zu: %s\n", mode, smem, cudaGetErrorString(e));
        if (e != cudaSuccess) return 1;
    }
    std::vector<float> ho(blocks * 4 * 3 * 32); cudaMemcpy(ho.data(), o, ho.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int b = 0; b < blocks; ++b) for (int w = 0; w < 4; ++w) for (int it = 0; it < 3; ++it) for (uint32_t l = 0; l < std::min(rows, 32u); ++l) {
        const uint32_t rid = (b * 131 + w * 17 + l * 7 + it) % nrows; float s = 0; for (uint32_t c = 0; c < dpad; ++c) s += h[(size_t)rid * dpad + c];
        if (s != ho[((b * 4 + w) * 3 + it) * 32 + l]) ++bad;
    }
    printf("mismatches %d\n", bad);
    return bad != 0;
}
