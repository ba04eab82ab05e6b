// Contraction on FP64 tensor cores with the operand staged by the bulk-copy engine.
//
//     acc[m, k, :] = sum over the pairs of region m of  r[cell, k] * stats[pair, :]      (R/melissa_vb.R:368-373 sums)
//
// Same mapping as contract_mma_kernel (sliced.cuh): one CTA per region, a warp walks its share of the region's
// slices and feeds them to mma.sync.m8n8k4.f64, four pairs at a time.  What changed: a slice's record block is ONE
// contiguous run of R * 256 bytes (plus 128 bytes of cell indices), so instead of plain loads issued tile by tile --
// ncu: 18 of 25 stall cycles per issue on the long scoreboard, 0.66 of the copy bandwidth -- one lane hands the whole
// block to the TMA unit (cp.async.bulk, global -> shared, completion on an mbarrier) NST - 1 slices ahead of the
// arithmetic.  The B tiles are then read from shared memory, where a tile is 32 consecutive doubles (conflict-free).
//
// Included from melissa.cu inside its anonymous namespace, after sliced.cuh.
#pragma once

__device__ __forceinline__ void mbar_init(uint32_t bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t phase) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar), "r"(phase)
        : "memory");
}
// global -> this CTA's shared memory; size and both addresses are multiples of 16
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(bar)
                 : "memory");
}

#ifndef CTMA_NST
#define CTMA_NST 2            // stages per warp: NST - 1 slices on their way while one is multiplied (2 CTAs x 8 warps per SM)
#endif

template <int MT, int NT>
__global__ void __launch_bounds__(256, 2)
contract_tma_kernel(const int64_t *__restrict__ reg_slice_ptr, const int32_t *__restrict__ reg_slice,
                    const int32_t *__restrict__ lane_cell, const double *__restrict__ stats,
                    const double *__restrict__ r, double *__restrict__ acc_out, int K, int R) {
    constexpr int NST = CTMA_NST;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int m = blockIdx.x;
    const int nw = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int quad = lane >> 2, tid4 = lane & 3;
    const uint32_t rec_bytes = (uint32_t)R * 256u, stage_bytes = rec_bytes + 128u;
    unsigned char *my_ring = smem_raw + (size_t)warp * NST * stage_bytes;
    const uint32_t ring_addr = (uint32_t)__cvta_generic_to_shared(my_ring);
    const uint32_t bar_addr = (uint32_t)__cvta_generic_to_shared(smem_raw + (size_t)nw * NST * stage_bytes) + (uint32_t)warp * NST * 8u;
    if (lane == 0) {
#pragma unroll
        for (int s = 0; s < NST; ++s) mbar_init(bar_addr + s * 8, 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    const int64_t q0 = reg_slice_ptr[m], q1 = reg_slice_ptr[m + 1];
    // this warp's slices: q0 + warp, q0 + warp + nw, ...
    const int cnt = (int)((q1 - q0 - warp + nw - 1) / nw) > 0 ? (int)((q1 - q0 - warp + nw - 1) / nw) : 0;
    double c[MT][NT][2];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) c[mt][nt][0] = c[mt][nt][1] = 0.0;
    // slice ids, 32 at a time (one per lane), one batch ahead
    auto load_ids = [&](int batch) -> int {
        const int i = batch * 32 + lane;
        return i < cnt ? reg_slice[q0 + warp + (int64_t)i * nw] : 0;
    };
    int ids_cur = load_ids(0), ids_next = load_ids(1);
    int cur_batch = 0;
    auto id_at = [&](int i) -> int64_t {
        const int b = i >> 5;
        const int src = b == cur_batch ? ids_cur : ids_next;
        return (int64_t)__shfl_sync(0xffffffffu, src, i & 31);
    };
    auto issue = [&](int i) {                                   // slice i of this warp -> stage i mod NST
        const int64_t s = id_at(i);
        if (lane == 0) {
            const int st = i % NST;
            const uint32_t dst = ring_addr + (uint32_t)st * stage_bytes, bar = bar_addr + (uint32_t)st * 8u;
            mbar_expect_tx(bar, stage_bytes);
            bulk_g2s(dst, stats + s * ((int64_t)R * 32), rec_bytes, bar);
            bulk_g2s(dst + rec_bytes, lane_cell + s * 32, 128u, bar);
        }
    };
    for (int i = 0; i < NST - 1 && i < cnt; ++i) issue(i);
    for (int i = 0; i < cnt; ++i) {
        if ((i >> 5) != cur_batch) {                            // entering the next batch of ids
            cur_batch = i >> 5;
            ids_cur = ids_next;
            ids_next = load_ids(cur_batch + 1);
        }
        __syncwarp();                                           // every lane is done with the stage that is refilled now
        if (i + NST - 1 < cnt) issue(i + NST - 1);
        const int st = i % NST;
        mbar_wait(bar_addr + (uint32_t)st * 8u, (uint32_t)((i / NST) & 1));
        const double *rec = reinterpret_cast<const double *>(my_ring + (size_t)st * stage_bytes);
        const int32_t *cells = reinterpret_cast<const int32_t *>(my_ring + (size_t)st * stage_bytes + rec_bytes);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
            const int cell = cells[4 * ks + tid4];
            double a[MT], b[NT];
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) {
                const int cl = mt * 8 + quad;
                a[mt] = (cell >= 0 && cl < K) ? r[(int64_t)cell * K + cl] : 0.0;
            }
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                const int j = nt * 8 + quad;
                b[nt] = j < R ? rec[(ks * R + j) * 4 + tid4] : 0.0;   // 32 consecutive doubles per tile
            }
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) dmma_m8n8k4(c[mt][nt][0], c[mt][nt][1], a[mt], b[nt]);
        }
    }
    // C fragment: lane holds rows quad (cluster), columns 2*tid4 + {0,1} (component) of each tile
    __syncthreads();                                            // the rings are free: reuse them for the cross-warp sum
    double *s_red = reinterpret_cast<double *>(smem_raw);
    double *mine = s_red + (size_t)warp * (MT * NT * 64);
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            mine[(mt * NT + nt) * 64 + quad * 8 + 2 * tid4] = c[mt][nt][0];
            mine[(mt * NT + nt) * 64 + quad * 8 + 2 * tid4 + 1] = c[mt][nt][1];
        }
    __syncthreads();
    for (int i = threadIdx.x; i < MT * NT * 64; i += blockDim.x) {
        double v = 0.0;
        for (int w = 0; w < nw; ++w) v += s_red[(size_t)w * (MT * NT * 64) + i];
        const int tile = i >> 6, row = (i >> 3) & 7, col = i & 7;
        const int cl = (tile / NT) * 8 + row, j = (tile % NT) * 8 + col;
        if (cl < K && j < R) acc_out[((int64_t)m * K + cl) * R + j] = v;
    }
}
