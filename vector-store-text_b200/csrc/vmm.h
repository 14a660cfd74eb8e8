// vmm.h — device arrays that grow IN PLACE (CUDA virtual memory management).
//
// The reference grows its index with usearch::Index::reserve(+1 000 000) every time the free space drops below a third of
// the increment (src/index/usearch.rs:60-66,201-212) under a WRITE lock: searches stall while the tables are re-allocated.
// Here every table is one virtual address range reserved up front for the largest index the GPU could ever hold (virtual
// space is free; 180 GB of HBM bounds the physical side), and growth maps fresh physical chunks behind the mapped part:
// no copy, no new pointer, no stall - kernels that are running keep running, a GraphView taken before the growth stays valid.
// Driver entry points come through cudaGetDriverEntryPoint, so libvsann.so does not link libcuda.
// Fallback (VSANN_VMM=0, or any driver call failing at init): cudaMalloc + device copy, pointer changes, callers must hold
// the table lock exclusively.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdlib>
#include <vector>

namespace vsann {

struct VmmApi {
    CUresult (*reserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*afree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*create)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
    CUresult (*release)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*unmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*set_access)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
    CUresult (*granularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
    bool ok = false;
};

inline const VmmApi& vmm_api() {
    static VmmApi api = [] {
        VmmApi a;
        if (const char* e = getenv("VSANN_VMM")) if (atoi(e) == 0) return a;
        auto get = [](const char* name) -> void* {
            void* p = nullptr; cudaDriverEntryPointQueryResult q;
            if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) { cudaGetLastError(); return nullptr; }
            return p;
        };
        a.reserve = (decltype(a.reserve))get("cuMemAddressReserve");
        a.afree = (decltype(a.afree))get("cuMemAddressFree");
        a.create = (decltype(a.create))get("cuMemCreate");
        a.release = (decltype(a.release))get("cuMemRelease");
        a.map = (decltype(a.map))get("cuMemMap");
        a.unmap = (decltype(a.unmap))get("cuMemUnmap");
        a.set_access = (decltype(a.set_access))get("cuMemSetAccess");
        a.granularity = (decltype(a.granularity))get("cuMemGetAllocationGranularity");
        a.ok = a.reserve && a.afree && a.create && a.release && a.map && a.unmap && a.set_access && a.granularity;
        return a;
    }();
    return api;
}

// One growable array of bytes on one device.
struct VArr {
    unsigned char* p = nullptr;     // stable for the array's lifetime in VMM mode
    size_t bytes = 0;               // usable (mapped / allocated) bytes
    size_t va_bytes = 0;            // reserved virtual range (VMM mode)
    size_t gran = 0;
    int device = 0;
    bool vmm = false;
    std::vector<std::pair<CUmemGenericAllocationHandle, size_t>> chunks;   // (handle, bytes) in mapping order

    // Reserves the range (VMM) - nothing is mapped yet. max_bytes bounds every later ensure().
    cudaError_t init(int dev, size_t max_bytes, bool want_vmm) {
        device = dev; p = nullptr; bytes = 0; vmm = false;
        const VmmApi& a = vmm_api();
        if (!want_vmm || !a.ok || max_bytes == 0) return cudaSuccess;
        CUmemAllocationProp prop{}; prop.type = CU_MEM_ALLOCATION_TYPE_PINNED; prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE; prop.location.id = dev;
        size_t g = 0;
        if (a.granularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || g == 0) return cudaSuccess;   // plain mode
        gran = g;
        const size_t want = ((max_bytes + g - 1) / g) * g;
        CUdeviceptr base = 0;
        if (a.reserve(&base, want, 0, 0, 0) != CUDA_SUCCESS) return cudaSuccess;                                          // plain mode
        p = reinterpret_cast<unsigned char*>(base); va_bytes = want; vmm = true;
        return cudaSuccess;
    }

    // Makes at least `want` bytes usable. New bytes are filled with `fill` (>= 0) on `st`. In plain mode the old contents are
    // copied (`keep` bytes of them) and the pointer changes: *moved is set.
    cudaError_t ensure(size_t want, int fill, size_t keep, cudaStream_t st, bool* moved) {
        if (moved) *moved = false;
        if (want <= bytes) return cudaSuccess;
        if (vmm) {
            const VmmApi& a = vmm_api();
            size_t target = ((want + gran - 1) / gran) * gran;
            if (target > va_bytes) return cudaErrorMemoryAllocation;
            const size_t add = target - bytes;
            CUmemAllocationProp prop{}; prop.type = CU_MEM_ALLOCATION_TYPE_PINNED; prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE; prop.location.id = device;
            CUmemGenericAllocationHandle h;
            if (a.create(&h, add, &prop, 0) != CUDA_SUCCESS) return cudaErrorMemoryAllocation;
            if (a.map((CUdeviceptr)(p + bytes), add, 0, h, 0) != CUDA_SUCCESS) { a.release(h); return cudaErrorMemoryAllocation; }
            CUmemAccessDesc ad{}; ad.location.type = CU_MEM_LOCATION_TYPE_DEVICE; ad.location.id = device; ad.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
            if (a.set_access((CUdeviceptr)(p + bytes), add, &ad, 1) != CUDA_SUCCESS) { a.unmap((CUdeviceptr)(p + bytes), add); a.release(h); return cudaErrorMemoryAllocation; }
            chunks.emplace_back(h, add);
            cudaError_t e = cudaSuccess;
            if (fill >= 0) e = cudaMemsetAsync(p + bytes, fill, add, st);
            bytes = target;
            return e;
        }
        unsigned char* np = nullptr;
        cudaError_t e = cudaMalloc(&np, want);
        if (e != cudaSuccess) return e;
        if (fill >= 0) { e = cudaMemsetAsync(np, fill, want, st); if (e != cudaSuccess) { cudaFree(np); return e; } }
        if (p && keep) { e = cudaMemcpyAsync(np, p, keep < bytes ? keep : bytes, cudaMemcpyDeviceToDevice, st); if (e != cudaSuccess) { cudaFree(np); return e; } }
        e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { cudaFree(np); return e; }
        if (p) cudaFree(p);
        p = np; bytes = want;
        if (moved) *moved = true;
        return cudaSuccess;
    }

    void release() {
        if (vmm) {
            const VmmApi& a = vmm_api();
            size_t off = 0;
            for (auto& c : chunks) { a.unmap((CUdeviceptr)(p + off), c.second); a.release(c.first); off += c.second; }
            chunks.clear();
            if (p) a.afree((CUdeviceptr)p, va_bytes);
        } else if (p) {
            cudaFree(p);
        }
        p = nullptr; bytes = 0; va_bytes = 0; vmm = false;
    }
};

}  // namespace vsann
