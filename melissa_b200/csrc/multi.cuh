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
