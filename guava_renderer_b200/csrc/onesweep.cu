// Global-sort binning ("sort_mode" = 1): the reference's own pipeline, rebuilt.
//
//   cub::DeviceScan::InclusiveSum(tiles_touched)   rasterizer_impl.cu:280   \  one kernel: chained
//   duplicateWithKeys                              rasterizer_impl.cu:70-111 > (decoupled look-back) scan
//   + the digit histograms of every sort pass                                /  + key emission + histograms
//   cub::DeviceRadixSort::SortPairs, bits [0,32+bit) rasterizer_impl.cu:303-311  onesweep LSD passes
//   cudaMemset(ranges) + identifyTileRanges        rasterizer_impl.cu:313-320   identify_tile_ranges_kernel
//
// Keys are the reference's 64-bit (tile << 32 | depth_bits), values the Gaussian
// index; emission order is the reference's (Gaussian index ascending, rect walked
// row-major), the sort is a stable least-significant-digit radix sort, so keys,
// values and ranges are bit-identical to the reference AND to the default
// tile-bucket path (binning.cu) -- tests/test_gpu_parity.py checks both.
//
// Onesweep (Adinets & Merrill 2022): every pass reads each pair ONCE.  The digit
// histograms of all passes are taken up front (here: inside the emission kernel,
// so the keys are never re-read for it); a pass kernel ranks a 2048-pair
// partition in shared memory, publishes its per-digit counts, and obtains the
// number of equal-digit pairs in all earlier partitions by looking back at the
// published (aggregate | inclusive-prefix) words of its predecessors instead of
// waiting for a separate scan kernel.  Partitions are handed out by an atomic
// ticket, so a partition's predecessors are always already running: the spin can
// not deadlock whatever the block scheduling order.
//
// This mode exists for tile lists that do not fit on chip (the bucket path sorts
// each tile inside shared memory, <= 16 384 instances) and as the like-for-like
// replacement of the reference's CUB chain; at the GUAVA shapes the bucket path
// moves 20 B per instance instead of ~150 B and stays the default.
#include "internal.cuh"
#include "tma.cuh"

namespace guava {

constexpr int OS_THREADS = 256;
constexpr int OS_WARPS = OS_THREADS / 32;
constexpr int OS_ITEMS = 8;
constexpr int OS_PART = OS_THREADS * OS_ITEMS;   // pairs per partition
constexpr int OS_MAX_PASSES = 8;
// ---------------------------------------------------------------------------------------------
// Scan + duplicateWithKeys + histograms.  One thread per (view, Gaussian); blocks are ordered by
// ticket.  point_offsets gets the reference's inclusive sums (GeometryState::point_offsets).
__global__ void __launch_bounds__(256)
duplicate_with_keys_kernel(const Inputs in, const ushort4* __restrict__ rects, const float* __restrict__ depths,
                           const uint32_t* __restrict__ num_rendered, uint32_t* __restrict__ ticket,
                           uint32_t* __restrict__ scan_state, uint32_t* __restrict__ hist, int passes, int total_bits,
                           uint64_t* __restrict__ keys, uint32_t* __restrict__ vals,
                           uint32_t* __restrict__ point_offsets, unsigned long long capacity)
{
    __shared__ uint32_t s_hist[OS_MAX_PASSES * 256];
    __shared__ uint32_t s_wsum[8];
    __shared__ uint32_t s_bid, s_base;
    if (num_rendered[1]) return;   // capacity overflow: the host re-runs with a larger buffer
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_bid = atomicAdd(ticket, 1u);
    for (int k = tid; k < passes * 256; k += 256) s_hist[k] = 0u;
    __syncthreads();
    const int bid = (int)s_bid;
    const size_t n = (size_t)in.V * in.P;
    const size_t gid = (size_t)bid * 256 + tid;

    uint32_t rx = 0, ry = 0, w = 0, cnt = 0, v = 0, idx = 0, dbits = 0;
    if (gid < n) {
        const ushort4 rc = rects[gid];
        if (rc.z > rc.x && rc.w > rc.y) {
            rx = rc.x; ry = rc.y; w = (uint32_t)(rc.z - rc.x);
            cnt = w * (uint32_t)(rc.w - rc.y);
            v = (uint32_t)(gid / in.P);
            idx = (uint32_t)(gid - (size_t)v * in.P);
            dbits = __float_as_uint(depths[gid]);
        }
    }
    uint32_t incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_wsum[warp] = incl;
    __syncthreads();
    uint32_t woff = 0, btotal = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const uint32_t x = s_wsum[k];
        if (k < warp) woff += x;
        btotal += x;
    }
    if (warp == 0) {
        if (lane == 0) st_volatile(scan_state + bid, (btotal & LB_VAL) | (bid == 0 ? LB_INC : LB_AGG));
        const uint32_t base = lookback_exclusive_warp(scan_state, bid, lane);
        if (lane == 0) {
            if (bid > 0) st_volatile(scan_state + bid, ((base + btotal) & LB_VAL) | LB_INC);
            s_base = base;
        }
    }
    __syncthreads();
    const uint32_t wbase = s_base + woff;
    if (gid < n) point_offsets[gid] = wbase + incl;

    // emission: the warp's instances are consecutive in the output, lane t writes instance base+t
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    const uint32_t excl = incl - cnt;
    for (uint32_t base = 0; base < total; base += 32) {
        const uint32_t t = base + lane;
        int owner = 0;
#pragma unroll
        for (int step = 16; step >= 1; step >>= 1) {
            const uint32_t e = __shfl_sync(0xffffffffu, excl, owner + step);
            if (e <= t) owner += step;
        }
        const uint32_t o_excl = __shfl_sync(0xffffffffu, excl, owner);
        const uint32_t o_w = __shfl_sync(0xffffffffu, w, owner);
        const uint32_t o_rx = __shfl_sync(0xffffffffu, rx, owner);
        const uint32_t o_ry = __shfl_sync(0xffffffffu, ry, owner);
        const uint32_t o_v = __shfl_sync(0xffffffffu, v, owner);
        const uint32_t o_idx = __shfl_sync(0xffffffffu, idx, owner);
        const uint32_t o_d = __shfl_sync(0xffffffffu, dbits, owner);
        if (t < total) {
            const uint32_t local = t - o_excl;
            const uint32_t yy = local / o_w, xx = local - yy * o_w;   // y outer, x inner: rasterizer_impl.cu:98-108
            const uint32_t tile = o_v * (uint32_t)in.tiles + (o_ry + yy) * (uint32_t)in.gx + (o_rx + xx);
            const uint64_t key = ((uint64_t)tile << 32) | o_d;
            const size_t pos = (size_t)wbase + t;
            if (pos < capacity) {
                keys[pos] = key;
                vals[pos] = o_idx;
            }
            for (int p = 0; p < passes; p++) {
                const int nb = min(8, total_bits - 8 * p);
                atomicAdd(&s_hist[p * 256 + (uint32_t)((key >> (8 * p)) & ((1u << nb) - 1u))], 1u);
            }
        }
    }
    __syncthreads();
    for (int k = tid; k < passes * 256; k += 256) {
        const uint32_t c = s_hist[k];
        if (c) atomicAdd(hist + k, c);
    }
}

// ---------------------------------------------------------------------------------------------
// One onesweep pass over digit (key >> shift) & mask.  Items are held warp-striped: warp w owns the
// contiguous segment [w*256, (w+1)*256) of its partition, item i of lane l is element i*32 + l, so
// "ascending (warp, i, lane)" is the input order and ranking in that order keeps the sort stable.
__global__ void __launch_bounds__(OS_THREADS)
onesweep_pass_kernel(const uint64_t* __restrict__ kin, const uint32_t* __restrict__ vin,
                     uint64_t* __restrict__ kout, uint32_t* __restrict__ vout,
                     const uint32_t* __restrict__ num_rendered, uint32_t* __restrict__ ticket,
                     const uint32_t* __restrict__ hist, uint32_t* __restrict__ state, int shift, uint32_t mask)
{
    __shared__ __align__(128) uint64_t s_keys[OS_PART];
    __shared__ __align__(128) uint32_t s_vals[OS_PART];
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ uint32_t s_whist[OS_WARPS][256];
    __shared__ uint32_t s_goff[256];    // global position of the digit's first pair of this partition, minus s_bexcl
    __shared__ uint32_t s_bexcl[256];   // first local (sorted) position of the digit
    __shared__ uint2 s_wsum[OS_WARPS];
    __shared__ uint32_t s_part;

    if (num_rendered[1]) return;
    const uint32_t R = num_rendered[0];
    const uint32_t nparts = (R + OS_PART - 1) / OS_PART;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) {
        s_part = atomicAdd(ticket, 1u);
        tma::mbar_init(&s_bar, 1);
        tma::fence_mbar_init();
    }
#pragma unroll
    for (int k = 0; k < OS_WARPS; k++) s_whist[k][tid] = 0u;
    __syncthreads();
    const uint32_t part = s_part;
    if (part >= nparts) return;   // block-uniform
    const uint32_t base = part * OS_PART;
    const uint32_t nvalid = min((uint32_t)OS_PART, R - base);
    // The partition is one contiguous, 16-byte aligned run of keys and one of values: two TMA bulk copies
    // (cp.async.bulk, completion on an mbarrier transaction count) bring it into the shared-memory buffers
    // that later hold the sorted order; sizes are rounded up to 16 bytes (the padding lies inside the
    // 256-byte aligned allocation and is never used).
    if (tid == 0) {   // issued first: the copies run under the histogram load and the constant-digit vote below
        const uint32_t kb = ((nvalid * 8u) + 15u) & ~15u, vb = ((nvalid * 4u) + 15u) & ~15u;
        tma::mbar_arrive_expect_tx(&s_bar, kb + vb);
        tma::bulk_g2s(s_keys, kin + base, kb, &s_bar);
        tma::bulk_g2s(s_vals, vin + base, vb, &s_bar);
    }

    // A pass whose digit is the same for every key (e.g. the exponent byte of the depths) is the
    // identity permutation: copy the partition straight through, no ranking, no look-back.
    const uint32_t my_hist = hist[tid];
    if (__syncthreads_or(my_hist == R)) {
        tma::mbar_wait(&s_bar, 0u);
#pragma unroll
        for (int i = 0; i < OS_ITEMS; i++) {
            const uint32_t j = (uint32_t)(i * OS_THREADS + tid);
            if (j < nvalid) {
                kout[base + j] = s_keys[j];
                vout[base + j] = s_vals[j];
            }
        }
        return;
    }

    tma::mbar_wait(&s_bar, 0u);
    uint64_t key[OS_ITEMS];
    uint32_t val[OS_ITEMS], rank[OS_ITEMS];
#pragma unroll
    for (int i = 0; i < OS_ITEMS; i++) {
        const uint32_t j = (uint32_t)(warp * (32 * OS_ITEMS) + i * 32 + lane);
        key[i] = j < nvalid ? s_keys[j] : ~0ull;
        val[i] = j < nvalid ? s_vals[j] : 0u;
    }
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint32_t* wh = s_whist[warp];
#pragma unroll
    for (int i = 0; i < OS_ITEMS; i++) {
        const uint32_t j = (uint32_t)(warp * (32 * OS_ITEMS) + i * 32 + lane);
        const bool valid = j < nvalid;
        const uint32_t d = (uint32_t)(key[i] >> shift) & mask;
        const uint32_t peers = __match_any_sync(0xffffffffu, valid ? d : (0x100u + (uint32_t)lane));
        uint32_t pre = 0;
        if (valid) pre = wh[d];
        __syncwarp();
        if (valid && (peers & lt_mask) == 0u) wh[d] = pre + (uint32_t)__popc(peers);
        __syncwarp();
        rank[i] = pre + (uint32_t)__popc(peers & lt_mask);
    }
    __syncthreads();

    // thread d: counts of digit d per warp -> exclusive over warps; partition count
    uint32_t count = 0;
#pragma unroll
    for (int k = 0; k < OS_WARPS; k++) {
        const uint32_t c = s_whist[k][tid];
        s_whist[k][tid] = count;
        count += c;
    }
    st_volatile(state + (size_t)part * 256 + tid, count | (part == 0 ? LB_INC : LB_AGG));

    // exclusive scans over the 256 digits of (partition count, global histogram)
    uint32_t ic = count, ih = my_hist;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t a = __shfl_up_sync(0xffffffffu, ic, o);
        const uint32_t b = __shfl_up_sync(0xffffffffu, ih, o);
        if (lane >= o) { ic += a; ih += b; }
    }
    if (lane == 31) s_wsum[warp] = make_uint2(ic, ih);
    __syncthreads();
    uint32_t oc = 0, oh = 0;
#pragma unroll
    for (int k = 0; k < OS_WARPS; k++) {
        const uint2 x = s_wsum[k];
        if (k < warp) { oc += x.x; oh += x.y; }
    }
    const uint32_t bexcl = oc + ic - count;      // pairs of this partition with a smaller digit
    const uint32_t gexcl = oh + ih - my_hist;    // pairs of the whole input with a smaller digit

    // decoupled look-back: pairs with this digit in earlier partitions
    uint32_t before = 0;
    if (part > 0) {
        for (int q = (int)part - 1; q >= 0; q--) {
            uint32_t s;
            do { s = ld_volatile(state + (size_t)q * 256 + tid); } while ((s >> 30) == 0u);
            before += s & LB_VAL;
            if ((s >> 30) == 2u) break;
        }
        st_volatile(state + (size_t)part * 256 + tid, ((before + count) & LB_VAL) | LB_INC);
    }
    s_bexcl[tid] = bexcl;
    s_goff[tid] = gexcl + before - bexcl;
    __syncthreads();

    // rank -> sorted position inside the partition; stage in shared memory so that the global
    // writes below are runs of consecutive addresses per digit
#pragma unroll
    for (int i = 0; i < OS_ITEMS; i++) {
        const uint32_t j = (uint32_t)(warp * (32 * OS_ITEMS) + i * 32 + lane);
        if (j < nvalid) {
            const uint32_t d = (uint32_t)(key[i] >> shift) & mask;
            const uint32_t lp = s_bexcl[d] + wh[d] + rank[i];
            s_keys[lp] = key[i];
            s_vals[lp] = val[i];
        }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < OS_ITEMS; i++) {
        const uint32_t j = (uint32_t)(i * OS_THREADS + tid);
        if (j < nvalid) {
            const uint64_t k = s_keys[j];
            const uint32_t d = (uint32_t)(k >> shift) & mask;
            const uint32_t dst = s_goff[d] + j;
            kout[dst] = k;
            vout[dst] = s_vals[j];
        }
    }
}

// ---------------------------------------------------------------------------------------------
// identifyTileRanges (rasterizer_impl.cu:116-138) over the sorted keys; `ranges` was zeroed.
__global__ void __launch_bounds__(256)
identify_tile_ranges_kernel(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ num_rendered,
                            uint2* __restrict__ ranges)
{
    if (num_rendered[1]) return;
    const uint32_t R = num_rendered[0];
    for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < R; i += gridDim.x * 256u) {
        const uint32_t tile = (uint32_t)(keys[i] >> 32);
        if (i == 0) ranges[tile].x = 0u;
        else {
            const uint32_t prev = (uint32_t)(keys[i - 1] >> 32);
            if (prev != tile) {
                ranges[prev].y = i;
                ranges[tile].x = i;
            }
        }
        if (i == R - 1) ranges[tile].y = R;
    }
}

int onesweep_passes(int total_tiles, int* total_bits)
{
    int bit = 0;
    while (bit < 32 && ((unsigned)total_tiles >> bit) != 0u) bit++;   // == getHigherMsb, rasterizer_impl.cu:35-50
    *total_bits = 32 + bit;
    return (32 + bit + 7) / 8;
}

size_t onesweep_scratch_bytes(size_t n, size_t capacity, int total_tiles)
{
    int bits;
    const int passes = onesweep_passes(total_tiles, &bits);
    const size_t dup_blocks = (n + 255) / 256;
    const size_t nparts = (capacity + OS_PART - 1) / OS_PART;
    return 4 * (64 + dup_blocks + (size_t)OS_MAX_PASSES * 256 + (size_t)passes * nparts * 256);
}

// keys0/vals0 receive the sorted result (vals0 = point_list); keys1/vals1 are the alternates.
void launch_onesweep_binning(const Inputs& in, const GeomPtrs& g, const ImagePtrs& img, uint32_t* point_offsets,
                             uint64_t* keys0, uint32_t* vals0, uint64_t* keys1, uint32_t* vals1, uint32_t* scratch,
                             size_t capacity, cudaStream_t s)
{
    const size_t n = (size_t)in.V * in.P;
    const int total_tiles = in.V * in.tiles;
    if (n == 0 || total_tiles == 0) return;
    int bits;
    const int passes = onesweep_passes(total_tiles, &bits);
    const size_t dup_blocks = (n + 255) / 256;
    const size_t nparts = (capacity + OS_PART - 1) / OS_PART;
    uint32_t* tickets = scratch;                 // [0] emission, [1 + p] pass p
    uint32_t* scan_state = scratch + 64;
    uint32_t* hist = scan_state + dup_blocks;
    uint32_t* state = hist + OS_MAX_PASSES * 256;
    cudaMemsetAsync(scratch, 0, onesweep_scratch_bytes(n, capacity, total_tiles), s);
    cudaMemsetAsync(img.ranges, 0, (size_t)total_tiles * sizeof(uint2), s);
    // after `passes` ping-pongs the result must sit in (keys0, vals0)
    uint64_t* kb[2] = {keys0, keys1};
    uint32_t* vb[2] = {vals0, vals1};
    int cur = passes & 1;
    duplicate_with_keys_kernel<<<(unsigned)dup_blocks, 256, 0, s>>>(in, g.rects, g.depths, img.num_rendered, tickets, scan_state,
                                                                     hist, passes, bits, kb[cur], vb[cur], point_offsets,
                                                                     (unsigned long long)capacity);
    for (int p = 0; p < passes; p++) {
        const int nb = bits - 8 * p < 8 ? bits - 8 * p : 8;
        onesweep_pass_kernel<<<(unsigned)(nparts ? nparts : 1), OS_THREADS, 0, s>>>(
            kb[cur], vb[cur], kb[cur ^ 1], vb[cur ^ 1], img.num_rendered, tickets + 1 + p, hist + p * 256,
            state + (size_t)p * nparts * 256, 8 * p, (1u << nb) - 1u);
        cur ^= 1;
    }
    identify_tile_ranges_kernel<<<sm_count() * 4, 256, 0, s>>>(kb[cur], img.num_rendered, img.ranges);
    count_launch(passes + 2);
}

}  // namespace guava
