// The sliced layout built ON THE DEVICE (replaces the reference's list walking, R/melissa_vb.R:119-131 and
// :254-259, and round 1's host-thread slice builder): everything after the index lists have reached the GPU
// is sorts, scans and scatters, so it overlaps the (much longer) upload of x / y instead of preceding it.
//
//   1. cells ordered by CpG count (stable, most covered first) and cut into blocks of Tn;
//   2. every pair gets the key  (cell block | region | 65535 - length | position of the cell in its block)
//      and the pairs are radix-sorted by it: the pairs of one (block, region) become adjacent, longest first --
//      exactly the order the host builder produced with its per-region stable sorts;
//   3. a group of (block, region) pairs is cut into slices of 32 lanes; slice ids, lanes and the x / y offsets
//      are prefix sums;
//   4. the region-major list of slices (the reference's cell_ind, :121) is one more stable sort.
//
// CUB (part of the CUDA toolkit) provides the radix sorts and scans: pack-time plumbing, not the hot path.
// Included from melissa.cu inside its anonymous namespace.
#pragma once
// (<cub/cub.cuh> is included at the top of melissa.cu, outside its anonymous namespace)

constexpr int PK_J_BITS = 12, PK_LEN_BITS = 16, PK_REG_BITS = 20;      // key fields, low to high; the block id takes the rest
constexpr int PK_GROUP_SHIFT = PK_J_BITS + PK_LEN_BITS;                // key >> this = (block, region)

// A shard's lists as a window of a larger job's lists: offsets and cell numbers made local after the upload.
__global__ void pk_rebase_kernel(int64_t *__restrict__ pair_ptr, int32_t *__restrict__ pair_cell, int64_t P, int64_t ptr_base,
                                 int32_t cell_base) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= P; i += (int64_t)gridDim.x * blockDim.x) {
        pair_ptr[i] -= ptr_base;
        if (i < P) pair_cell[i] -= cell_base;
    }
}

__global__ void pk_validate_kernel(const int64_t *__restrict__ pair_ptr, const int32_t *__restrict__ pair_cell,
                                   const int32_t *__restrict__ pair_region, int64_t P, int n_cells, int n_regions,
                                   unsigned long long *first_bad, int *reg_seen) {
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
        const int n = pair_cell[p], m = pair_region[p];
        int code = 0;
        if (n < 0 || n >= n_cells || m < 0 || m >= n_regions) code = 1;
        else {
            const int pn = p ? pair_cell[p - 1] : -1, pm = p ? pair_region[p - 1] : -1;
            if (n < pn) code = 2;
            else if (n > pn + 1) code = 3;
            else if (n == pn && m <= pm) code = 4;
            else if (pair_ptr[p + 1] <= pair_ptr[p]) code = 5;
            else if (pair_ptr[p + 1] - pair_ptr[p] > 65535) code = 6;
        }
        if (code) atomicMin(first_bad, ((unsigned long long)p << 3) | (unsigned)code);
        else reg_seen[m] = 1;
    }
}

__global__ void pk_cell_ptr_kernel(const int32_t *__restrict__ pair_cell, int64_t P, int N, int64_t *cell_ptr) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p > P) return;
    if (p == P) { cell_ptr[N] = P; return; }
    if (p == 0 || pair_cell[p] != pair_cell[p - 1]) cell_ptr[pair_cell[p]] = p;
}

__global__ void pk_cell_keys_kernel(const int64_t *__restrict__ pair_ptr, const int64_t *__restrict__ cell_ptr, int N,
                                    unsigned long long *key, int32_t *val) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    key[n] = ~(unsigned long long)(pair_ptr[cell_ptr[n + 1]] - pair_ptr[cell_ptr[n]]);      // descending coverage
    val[n] = n;
}

__global__ void pk_cell_pos_kernel(const int32_t *__restrict__ order, int N, int NBTn, int32_t *pos_of_cell, int32_t *blk_cells) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= NBTn) return;
    if (i < N) {
        pos_of_cell[order[i]] = i;
        blk_cells[i] = order[i];
    } else {
        blk_cells[i] = -1;
    }
}

__global__ void pk_pair_keys_kernel(const int64_t *__restrict__ pair_ptr, const int32_t *__restrict__ pair_cell,
                                    const int32_t *__restrict__ pair_region, const int32_t *__restrict__ pos_of_cell,
                                    int64_t P, int Tn, unsigned long long *key, int32_t *val) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const int pos = pos_of_cell[pair_cell[p]];
    const unsigned long long cb = (unsigned)(pos / Tn), j = (unsigned)(pos % Tn);
    const unsigned long long len = (unsigned long long)(pair_ptr[p + 1] - pair_ptr[p]);
    key[p] = (cb << (PK_GROUP_SHIFT + PK_REG_BITS)) | ((unsigned long long)pair_region[p] << PK_GROUP_SHIFT) |
             ((65535ull - len) << PK_J_BITS) | j;
    val[p] = (int32_t)p;
}

__global__ void pk_head_kernel(const unsigned long long *__restrict__ key, int64_t P, int32_t *head_idx) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    head_idx[i] = (i == 0 || (key[i] >> PK_GROUP_SHIFT) != (key[i - 1] >> PK_GROUP_SHIFT)) ? (int32_t)i : 0;
}

__global__ void pk_slice_flag_kernel(const int32_t *__restrict__ group_start, int64_t P, int32_t *flag) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    flag[i] = (((int32_t)i - group_start[i]) & 31) == 0;
}

// one thread per sorted pair: its lane of its slice
__global__ void pk_scatter_kernel(const unsigned long long *__restrict__ key, const int32_t *__restrict__ val,
                                  const int32_t *__restrict__ group_start, const int32_t *__restrict__ slice_id,
                                  const int64_t *__restrict__ pair_ptr, const int32_t *__restrict__ pair_cell, int64_t P,
                                  int32_t *lane_cell, int32_t *lane_pair, int32_t *lane_meta, int64_t *lane_src,
                                  int32_t *slice_region, int32_t *slice_cb, int32_t *slice_len, int64_t *slice_xlen,
                                  int64_t *slice_ylen) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const unsigned long long k = key[i];
    const int p = val[i];
    const int lane = ((int32_t)i - group_start[i]) & 31;
    const int64_t s = slice_id[i] - 1;
    const int len = 65535 - (int)((k >> PK_J_BITS) & 0xffff), j = (int)(k & ((1u << PK_J_BITS) - 1));
    const int64_t t = s * 32 + lane;
    lane_cell[t] = pair_cell[p];
    lane_pair[t] = p;
    lane_meta[t] = len | (j << 16);
    lane_src[t] = pair_ptr[p];
    if (lane == 0) {                                            // the longest pair of the slice
        slice_region[s] = (int)((k >> PK_GROUP_SHIFT) & ((1u << PK_REG_BITS) - 1));
        slice_cb[s] = (int)(k >> (PK_GROUP_SHIFT + PK_REG_BITS));
        slice_len[s] = len;
        slice_xlen[s] = (int64_t)len * 32;
        slice_ylen[s] = (int64_t)((len + 31) / 32) * 32;
    }
}

__global__ void pk_iota_kernel(int32_t *v, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = (int32_t)i;
}

// ptr[m] = first position in the sorted key list with key >= m  (m = 0 .. M)
__global__ void pk_lower_bound_kernel(const int32_t *__restrict__ sorted, int64_t n, int M, int64_t *ptr) {
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m > M) return;
    int64_t lo = 0, hi = n;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (sorted[mid] < m) lo = mid + 1; else hi = mid;
    }
    ptr[m] = lo;
}

struct MaxOp {
    __device__ __forceinline__ int32_t operator()(int32_t a, int32_t b) const { return a > b ? a : b; }
};

// Chunks of the sweep: runs of consecutive slices that stay inside one cell block (blk_first[cb] = first slice of
// block cb, blk_first[NB] = S).  A CTA claims chunks in list order and meets at a barrier at the end of each (the
// block sums are flushed there), where the warps that finished early wait for the last one: about one slice time
// per chunk.  So chunks should be few and long while there is plenty of work left and short only at the end, where
// they decide how evenly the CTAs finish: guided self-scheduling, chunk = what is left / (2 x the number of CTAs),
// between 64 slices and what the limbs can absorb.  (`taper_steps`: the earlier three-step list -- full size for the
// first half of the slices, a half up to 80 %, a quarter after that -- kept for comparison, option "sweep_taper".)
std::vector<int32_t> make_chunks(const std::vector<int64_t> &blk_first, int64_t S, int sm_count, int chunk_opt,
                                 bool taper_steps = false) {
    // (a chunk adds to a shared accumulator at most once per slice; LIMB_MAX_ADDS is what its limbs can absorb)
    const int64_t cap = 2048;
    const int64_t base = std::min<int64_t>(LIMB_MAX_ADDS, chunk_opt > 0 ? chunk_opt
                                               : std::max<int64_t>(256, std::min<int64_t>(cap, S / ((int64_t)sm_count * 8))));
    std::vector<int32_t> chunk_ptr;
    const size_t NB = blk_first.size() - 1;
    for (size_t cb = 0; cb < NB; ++cb) {
        int64_t q = blk_first[cb];
        const int64_t e = blk_first[cb + 1];
        while (q < e) {
            int64_t target = base;
            if (chunk_opt <= 0 && taper_steps) {
                if (q * 10 >= S * 8) target = std::max<int64_t>(48, base / 4);
                else if (q * 2 >= S) target = std::max<int64_t>(96, base / 2);
            } else if (chunk_opt <= 0) {
                target = std::max<int64_t>(64, std::min<int64_t>(cap, (S - q) / (2 * (int64_t)sm_count)));
            }
            const int64_t left = e - q, pieces = (left + target - 1) / target;
            chunk_ptr.push_back((int32_t)q);
            q += (left + pieces - 1) / pieces;
        }
    }
    chunk_ptr.push_back((int32_t)S);
    return chunk_ptr;
}

// Everything index-related of the sliced layout, from DEVICE index lists, on `st`.  x / y are not touched here.
// On return sl holds S, the slice / lane arrays, xbase / ybase, x_len / y_len, blk_cells, reg_slice(_ptr),
// chunk_ptr and d_src_out (the CpG offset of every lane in the flat x / y, for the fill kernel; caller frees).
int build_slice_index_device(melissa_handle *h, Sliced *sl, const int64_t *d_ptr, const int32_t *d_cell, const int32_t *d_reg,
                             cudaStream_t st, int64_t **d_src_out, int sweep_chunk_opt) {
    const int N = h->N, M = h->M, Tn = sl->Tn;
    const int64_t P = h->P;
    const int NB = sl->NB;
    if (Tn > (1 << PK_J_BITS) || M > (1 << PK_REG_BITS) || (int64_t)NB >= (1ll << (64 - PK_GROUP_SHIFT - PK_REG_BITS)))
        return fail(MELISSA_ERR_ARG, "sizes beyond the device packer's key (cell_block <= 4096, regions <= 2^20)");
    int rc = MELISSA_OK;
    std::vector<void *> tmp;                                    // scratch freed on every way out
    auto scratch = [&](size_t bytes) -> void * {
        void *p = nullptr;
        if (cudaMalloc(&p, std::max<size_t>(bytes, 16)) != cudaSuccess) { rc = fail(MELISSA_ERR_CUDA, "out of device memory while packing"); return nullptr; }
        tmp.push_back(p);
        return p;
    };
    auto done = [&](int code) { for (void *p : tmp) cudaFree(p); return code; };
    const unsigned gP = (unsigned)((P + 255) / 256);
    // ---- 1. cells by coverage
    int64_t *cell_ptr = (int64_t *)scratch(((size_t)N + 1) * 8);
    unsigned long long *ck_in = (unsigned long long *)scratch((size_t)N * 8), *ck_out = (unsigned long long *)scratch((size_t)N * 8);
    int32_t *cv_in = (int32_t *)scratch((size_t)N * 4), *order = (int32_t *)scratch((size_t)N * 4);
    int32_t *pos_of_cell = (int32_t *)scratch((size_t)N * 4);
    if (rc) return done(rc);
    pk_cell_ptr_kernel<<<(unsigned)((P + 256) / 256), 256, 0, st>>>(d_cell, P, N, cell_ptr);
    pk_cell_keys_kernel<<<(N + 255) / 256, 256, 0, st>>>(d_ptr, cell_ptr, N, ck_in, cv_in);
    size_t tb = 0, tb2 = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tb, ck_in, ck_out, cv_in, order, N, 0, 64, st);
    // ---- 2. pairs by (block, region, length, position)
    unsigned long long *k_in = (unsigned long long *)scratch((size_t)P * 8), *k_out = (unsigned long long *)scratch((size_t)P * 8);
    int32_t *v_in = (int32_t *)scratch((size_t)P * 4), *v_out = (int32_t *)scratch((size_t)P * 4);
    int nb_bits = 1;
    while ((1 << nb_bits) < NB) ++nb_bits;
    const int end_bit = PK_GROUP_SHIFT + PK_REG_BITS + nb_bits;
    cub::DeviceRadixSort::SortPairs(nullptr, tb2, k_in, k_out, v_in, v_out, (int)P, 0, end_bit, st);
    size_t tb3 = 0, tb4 = 0;
    int32_t *head_idx = v_in;                                   // v_in is free once the sort has run
    int32_t *group_start = (int32_t *)scratch((size_t)P * 4), *slice_id = (int32_t *)scratch((size_t)P * 4);
    cub::DeviceScan::InclusiveScan(nullptr, tb3, head_idx, group_start, MaxOp(), (int)P, st);
    cub::DeviceScan::InclusiveSum(nullptr, tb4, head_idx, slice_id, (int)P, st);
    const size_t tbytes = std::max(std::max(tb, tb2), std::max(tb3, tb4));
    void *d_temp = scratch(tbytes);
    if (rc) return done(rc);
    size_t tbb = tbytes;
    cub::DeviceRadixSort::SortPairs(d_temp, tbb, ck_in, ck_out, cv_in, order, N, 0, 64, st);
    if ((rc = dev_alloc(&sl->blk_cells, (int64_t)NB * Tn))) return done(rc);
    pk_cell_pos_kernel<<<(NB * Tn + 255) / 256, 256, 0, st>>>(order, N, NB * Tn, pos_of_cell, sl->blk_cells);
    pk_pair_keys_kernel<<<gP, 256, 0, st>>>(d_ptr, d_cell, d_reg, pos_of_cell, P, Tn, k_in, v_in);
    tbb = tbytes;
    cub::DeviceRadixSort::SortPairs(d_temp, tbb, k_in, k_out, v_in, v_out, (int)P, 0, end_bit, st);
    // ---- 3. slices
    pk_head_kernel<<<gP, 256, 0, st>>>(k_out, P, head_idx);
    tbb = tbytes;
    cub::DeviceScan::InclusiveScan(d_temp, tbb, head_idx, group_start, MaxOp(), (int)P, st);
    int32_t *flag = head_idx;
    pk_slice_flag_kernel<<<gP, 256, 0, st>>>(group_start, P, flag);
    tbb = tbytes;
    cub::DeviceScan::InclusiveSum(d_temp, tbb, flag, slice_id, (int)P, st);
    int32_t S32 = 0;
    CU(cudaMemcpyAsync(&S32, slice_id + (P - 1), sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const int64_t S = S32;
    sl->S = S;
    int64_t *d_src = nullptr, *xlen = nullptr, *ylen = nullptr;
    if ((rc = dev_alloc(&sl->slice_xbase, S + 1)) || (rc = dev_alloc(&sl->slice_ybase, S + 1)) ||
        (rc = dev_alloc(&sl->slice_len, S)) || (rc = dev_alloc(&sl->slice_region, S)) || (rc = dev_alloc(&sl->slice_cb, S)) ||
        (rc = dev_alloc(&sl->lane_cell, S * 32)) || (rc = dev_alloc(&sl->lane_meta, S * 32)) ||
        (rc = dev_alloc(&sl->lane_pair, S * 32)) || (rc = dev_alloc(&d_src, S * 32)) ||
        (rc = dev_alloc(&sl->reg_slice_ptr, (int64_t)M + 1)) || (rc = dev_alloc(&sl->reg_slice, S)))
        { cudaFree(d_src); return done(rc); }
    *d_src_out = d_src;
    xlen = (int64_t *)scratch(((size_t)S + 1) * 8);
    ylen = (int64_t *)scratch(((size_t)S + 1) * 8);
    int32_t *rk_in = (int32_t *)scratch((size_t)S * 4), *rk_out = (int32_t *)scratch((size_t)S * 4), *rv_in = (int32_t *)scratch((size_t)S * 4);
    int32_t *blk_first = (int32_t *)scratch(((size_t)NB + 1) * 4);
    if (rc) return done(rc);
    CU(cudaMemsetAsync(sl->lane_cell, 0xff, (size_t)S * 32 * 4, st));          // padding lanes: cell -1, pair -1, no CpG
    CU(cudaMemsetAsync(sl->lane_pair, 0xff, (size_t)S * 32 * 4, st));
    CU(cudaMemsetAsync(sl->lane_meta, 0, (size_t)S * 32 * 4, st));
    CU(cudaMemsetAsync(d_src, 0, (size_t)S * 32 * 8, st));
    CU(cudaMemsetAsync(xlen + S, 0, 8, st));
    CU(cudaMemsetAsync(ylen + S, 0, 8, st));
    pk_scatter_kernel<<<gP, 256, 0, st>>>(k_out, v_out, group_start, slice_id, d_ptr, d_cell, P, sl->lane_cell, sl->lane_pair,
                                          sl->lane_meta, d_src, sl->slice_region, sl->slice_cb, sl->slice_len, xlen, ylen);
    {   // x / y offsets: exclusive sums over S + 1 entries, so that entry S is the total
        size_t need = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, need, xlen, sl->slice_xbase, (int)(S + 1), st);
        void *t2 = need > tbytes ? scratch(need) : d_temp;
        if (rc) return done(rc);
        size_t nb = std::max(need, tbytes);
        cub::DeviceScan::ExclusiveSum(t2, nb, xlen, sl->slice_xbase, (int)(S + 1), st);
        nb = std::max(need, tbytes);
        cub::DeviceScan::ExclusiveSum(t2, nb, ylen, sl->slice_ybase, (int)(S + 1), st);
    }
    // ---- 4. region-major list of slices (ascending slice id inside a region, as the host builder had it)
    {
        pk_iota_kernel<<<(unsigned)((S + 255) / 256), 256, 0, st>>>(rv_in, S);
        CU(cudaMemcpyAsync(rk_in, sl->slice_region, (size_t)S * 4, cudaMemcpyDeviceToDevice, st));
        size_t need = 0;
        int rbits = 1;
        while ((1 << rbits) < M) ++rbits;
        cub::DeviceRadixSort::SortPairs(nullptr, need, rk_in, rk_out, rv_in, sl->reg_slice, (int)S, 0, rbits, st);
        void *t2 = need > tbytes ? scratch(need) : d_temp;
        if (rc) return done(rc);
        size_t nb = std::max(need, tbytes);
        cub::DeviceRadixSort::SortPairs(t2, nb, rk_in, rk_out, rv_in, sl->reg_slice, (int)S, 0, rbits, st);
        pk_lower_bound_kernel<<<(M + 256) / 256, 256, 0, st>>>(rk_out, S, M, sl->reg_slice_ptr);
        // first slice of every cell block (slice_cb is ascending), for the chunk table
        pk_lower_bound_kernel<<<(NB + 256) / 256, 256, 0, st>>>(sl->slice_cb, S, NB, (int64_t *)scratch(((size_t)NB + 1) * 8));
        if (rc) return done(rc);
    }
    std::vector<int64_t> blk64((size_t)NB + 1), tot(2);
    CU(cudaMemcpyAsync(blk64.data(), tmp.back(), ((size_t)NB + 1) * 8, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(&tot[0], sl->slice_xbase + S, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(&tot[1], sl->slice_ybase + S, 8, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    (void)blk_first;
    sl->x_len = tot[0];
    sl->y_len = tot[1];
    // ---- chunks of the sweep
    const std::vector<int32_t> chunk_ptr = make_chunks(blk64, S, h->sm_count, sweep_chunk_opt, g_opt_sweep_taper != 0);
    sl->NC = (int)chunk_ptr.size() - 1;
    if ((rc = dev_alloc(&sl->chunk_ptr, (int64_t)chunk_ptr.size()))) return done(rc);
    CU(cudaMemcpyAsync(sl->chunk_ptr, chunk_ptr.data(), chunk_ptr.size() * 4, cudaMemcpyHostToDevice, st));
    CU(cudaStreamSynchronize(st));
    return done(MELISSA_OK);
}
