// Several GPUs behind ONE call of the C ABI (the reference's user calls one function, R/melissa_vb.R:86-90; its only
// parallel knob is no_cores, R/utils.R:30-43).  One process, one host thread and one stream per GPU, cells sharded
// in contiguous blocks.  The only exchange of an iteration -- the r-weighted sufficient statistics, M K R + K + 1
// doubles per GPU -- is not a library collective: every GPU's reduce kernel reads the partial sums of all GPUs
// straight out of their memory over NVLink (peer loads) and adds them in rank order, so the result is the same
// bit pattern on every GPU and on every run.  Partial sums are double-buffered by iteration parity, which leaves one
// cross-GPU synchronisation point per iteration (an event per GPU, waited for by every peer's stream).
//
// Included from melissa.cu inside its anonymous namespace.
#pragma once

constexpr int MULTI_MAX_GPUS = 16;

// The exchange buffers are the only memory another GPU ever reads, so they alone are made visible to the peers:
// one block per GPU from the driver's virtual-memory API (cuMemCreate / cuMemMap), opened to every GPU of the job with
// cuMemSetAccess.  Switching peer access on for whole devices (cudaDeviceEnablePeerAccess) costs far more: it maps
// every allocation of the device into every peer, once when it is switched on (0.2 s after a C4 pack on 4 GPUs) and
// again for every later allocation (a pack 3x slower).  The entry points are looked up at run time so that the
// library has no link-time dependency on libcuda (it must load on a machine without a driver).
struct DrvApi {
    CUresult (*GetAllocationGranularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags) = nullptr;
    CUresult (*Create)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long) = nullptr;
    CUresult (*AddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*Map)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*SetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t) = nullptr;
    CUresult (*Unmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*AddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*Release)(CUmemGenericAllocationHandle) = nullptr;
    bool ok = false;
    DrvApi() {
        auto get = [](const char *name, void **fn) {
            cudaDriverEntryPointQueryResult q;
            *fn = nullptr;
            return cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess && *fn;
        };
        ok = get("cuMemGetAllocationGranularity", (void **)&GetAllocationGranularity) && get("cuMemCreate", (void **)&Create) &&
             get("cuMemAddressReserve", (void **)&AddressReserve) && get("cuMemMap", (void **)&Map) &&
             get("cuMemSetAccess", (void **)&SetAccess) && get("cuMemUnmap", (void **)&Unmap) &&
             get("cuMemAddressFree", (void **)&AddressFree) && get("cuMemRelease", (void **)&Release);
        cudaGetLastError();
    }
};

struct SharedBuf {               // device memory of one GPU that the listed GPUs may read and write
    CUdeviceptr ptr = 0;
    size_t size = 0;
    CUmemGenericAllocationHandle hnd = 0;
    bool created = false, reserved = false, mapped = false;
    void free(const DrvApi &api) {
        if (mapped) api.Unmap(ptr, size);
        if (reserved) api.AddressFree(ptr, size);
        if (created) api.Release(hnd);
        *this = SharedBuf();
    }
    // false: the caller falls back to whole-device peer access
    bool alloc(const DrvApi &api, int device, size_t bytes, const int *devices, int n_devices) {
        if (!api.ok) return false;
        CUmemAllocationProp prop = {};
        prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        prop.location.id = device;
        size_t gran = 0;
        if (api.GetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_MINIMUM) != CUDA_SUCCESS || gran == 0) return false;
        size = (bytes + gran - 1) / gran * gran;
        if (api.Create(&hnd, size, &prop, 0) != CUDA_SUCCESS) { *this = SharedBuf(); return false; }
        created = true;
        if (api.AddressReserve(&ptr, size, 0, 0, 0) != CUDA_SUCCESS) { free(api); return false; }
        reserved = true;
        if (api.Map(ptr, size, 0, hnd, 0) != CUDA_SUCCESS) { free(api); return false; }
        mapped = true;
        CUmemAccessDesc acc[MULTI_MAX_GPUS];
        int n = 0;
        for (int i = 0; i < n_devices; ++i) {
            bool seen = false;
            for (int j = 0; j < n; ++j) seen = seen || acc[j].location.id == devices[i];
            if (seen) continue;
            acc[n].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
            acc[n].location.id = devices[i];
            acc[n].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
            ++n;
        }
        if (api.SetAccess(ptr, size, acc, (size_t)n) != CUDA_SUCCESS) { free(api); return false; }
        return true;
    }
};

struct PartPtrs {
    const double *p[MULTI_MAX_GPUS];
    int G;
};

// out[i] = sum over GPUs (rank order) of their partial buffer; the addresses in pp are peer-mapped device pointers.
__global__ void __launch_bounds__(256) reduce_parts_kernel(PartPtrs pp, int64_t n, double *__restrict__ out) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t n2 = n >> 1;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
        double2 s = reinterpret_cast<const double2 *>(pp.p[0])[i];
        for (int g = 1; g < pp.G; ++g) {
            const double2 v = reinterpret_cast<const double2 *>(pp.p[g])[i];
            s.x += v.x;
            s.y += v.y;
        }
        reinterpret_cast<double2 *>(out)[i] = s;
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
        double s = pp.p[0][n - 1];
        for (int g = 1; g < pp.G; ++g) s += pp.p[g][n - 1];
        out[n - 1] = s;
    }
}

// Sense-reversing barrier of the per-GPU host threads; wait() returns false once any thread has failed.
struct HostBarrier {
    std::atomic<int> count{0}, gen{0}, failed{0};
    int n = 1;
    bool wait() {
        const int g = gen.load(std::memory_order_acquire);
        if (count.fetch_add(1, std::memory_order_acq_rel) + 1 == n) {
            count.store(0, std::memory_order_relaxed);
            gen.fetch_add(1, std::memory_order_acq_rel);
        } else {
            while (gen.load(std::memory_order_acquire) == g) {
                if (failed.load(std::memory_order_relaxed)) return false;
                std::this_thread::yield();
            }
        }
        return !failed.load(std::memory_order_relaxed);
    }
    void fail_all() { failed.store(1); }
};
