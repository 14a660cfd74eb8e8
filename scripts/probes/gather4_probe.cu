// Probe: semantics of cp.async.bulk.tensor.2d ... tile::gather4 on sm_100a (which box the tensor map needs,
// where the four rows land in shared memory).
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <vector>
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
__global__ void k(const __grid_constant__ CUtensorMap map, int r0, int r1, int r2, int r3, float* out, uint32_t bytes, int n_out) {
    extern __shared__ __align__(128) unsigned char sm[];
    __shared__ __align__(8) unsigned long long bar;
    const uint32_t b = (uint32_t)__cvta_generic_to_shared(&bar), d = (uint32_t)__cvta_generic_to_shared(sm);
    for (int i = threadIdx.x; i < n_out; i += 32) ((float*)sm)[i] = -1.f;
    if (threadIdx.x == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b)); asm volatile("fence.mbarrier_init.release.cluster;"); }
    __syncwarp();
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
                     ::"r"(d), "l"(&map), "r"(b), "r"(0), "r"(r0), "r"(r1), "r"(r2), "r"(r3) : "memory");
    }
    uint32_t ok = 0; int spins = 0;
    do { asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(b), "r"(0) : "memory"); } while (!ok && ++spins < 2000000);
    __syncwarp();
    if (threadIdx.x == 0) out[n_out] = ok ? 1.f : 0.f;
    for (int i = threadIdx.x; i < n_out; i += 32) out[i] = ((float*)sm)[i];
}
int main() {
    const int d = 128, n = 1000;
    std::vector<float> h((size_t)n * d); for (int r = 0; r < n; ++r) for (int c = 0; c < d; ++c) h[(size_t)r * d + c] = r + c * 0.001f;
    float* g; cudaMalloc(&g, h.size() * 4); cudaMemcpy(g, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    void* p = nullptr; cudaDriverEntryPointQueryResult q; cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    EncodeTiledFn enc = (EncodeTiledFn)p;
    const int n_out = 4 * d; float* o; cudaMalloc(&o, (n_out + 1) * 4);
    for (int boxrows : {1, 4}) {
        CUtensorMap m; cuuint64_t dims[2] = {(cuuint64_t)d, (cuuint64_t)n}, strides[1] = {(cuuint64_t)d * 4}; cuuint32_t box[2] = {(cuuint32_t)d, (cuuint32_t)boxrows}, es[2] = {1, 1};
        CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, g, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        printf("box rows %d: encode %d\n", boxrows, (int)r);
        if (r != CUDA_SUCCESS) continue;
        cudaMemset(o, 0, (n_out + 1) * 4);
        k<<<1, 32, 4 * d * 4>>>(m, 5, 100, 7, 33, o, 4 * d * 4, n_out);
        cudaError_t e = cudaDeviceSynchronize();
        std::vector<float> ho(n_out + 1); cudaMemcpy(ho.data(), o, ho.size() * 4, cudaMemcpyDeviceToHost);
        printf("  run: %s, completed %g; row starts: %g %g %g %g | [1]=%g [127]=%g\n", cudaGetErrorString(e), ho[n_out], ho[0], ho[d], ho[2 * d], ho[3 * d], ho[1], ho[127]);
        if (e != cudaSuccess) return 1;
    }
    return 0;
}
