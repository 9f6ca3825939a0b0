// Instance binning: tile scan -> bucket scatter -> per-tile depth sort.
//
// Replaces, with identical results, the reference chain
//   cub::DeviceScan::InclusiveSum            rasterizer_impl.cu:280
//   cudaMemcpy(num_rendered) (host sync)     rasterizer_impl.cu:284
//   duplicateWithKeys                        rasterizer_impl.cu:70-111,292
//   cub::DeviceRadixSort::SortPairs          rasterizer_impl.cu:303-311
//   cudaMemset(ranges) + identifyTileRanges  rasterizer_impl.cu:313-320,116-138
//
// Why the results are identical although the algorithm is not an LSD sort of
// 64-bit keys: the reference key is (tile << 32 | depth_bits) and its sort is
// stable over an emission order that is ascending in Gaussian index; a Gaussian
// touches a tile at most once, so the reference's final order is exactly the
// strict total order (tile, depth_bits, gaussian_index).  Any algorithm that
// sorts by that triple produces the same point_list, the same keys and the same
// ranges.  Here the most significant digit (the tile id) is resolved by a
// counting pass whose histogram was accumulated by the preprocess kernel and
// whose scan *is* identifyTileRanges; the remaining (depth_bits, index) order is
// established per tile inside shared memory.  HBM traffic per instance drops
// from 152 B (1 histogram + 6 onesweep passes over 12-byte pairs) to 8 B
// written by the scatter + 8 B read and 4 B written by the tile sort.
#include <algorithm>
#include <atomic>
#include <cstdlib>

#include "internal.cuh"

namespace guava {

// ------------------------------------------------------------ scan over tiles
// One 1024-thread block.  Tiles are visited in chunks of 1024 (thread t <-> tile chunk*1024 + t), so the
// loads of a chunk are independent and in flight together; a running carry links the chunks.
__device__ __forceinline__ uint32_t block_exclusive_scan_1024(uint32_t x, uint32_t* warp_sums, int tid, uint32_t& total)
{
    uint32_t incl = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((tid & 31) >= o) incl += t;
    }
    __syncthreads();   // warp_sums may still be read from the previous call
    if ((tid & 31) == 31) warp_sums[tid >> 5] = incl;
    __syncthreads();
    if (tid < 32) {
        uint32_t w = warp_sums[tid];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
            if (tid >= o) w += t;
        }
        warp_sums[tid] = w;
    }
    __syncthreads();
    total = warp_sums[31];
    return incl - x + ((tid >> 5) ? warp_sums[(tid >> 5) - 1] : 0u);
}

// Scan scratch (zeroed with the tile counters, which it follows in the image buffer): ticket, longest list,
// 256 length bins, 256 bin cursors, one look-back word per 1024-tile block.
constexpr int SCAN_TICKET = 0, SCAN_LONGEST = 1, SCAN_DONE = 2, SCAN_BINS = 32, SCAN_BINCUR = 288, SCAN_STATE = 544;
size_t tile_scan_scratch_bytes(size_t total_tiles) { return 4 * (SCAN_STATE + (total_tiles + 1023) / 1024 + 1); }

// Zero-fill duty of the scan launches' spare blocks (guava_call_options::zero_fill and the backward's grad_geo scratch):
// `nthreads` threads, this one is number `first`; 16-byte vectors with a scalar head and tail.
__device__ __forceinline__ void zero_fill_slice(uint32_t* p, size_t words, size_t first, size_t nthreads)
{
    if (p == nullptr) return;
    const size_t head = min(words, (size_t)((16u - ((uintptr_t)p & 15u)) & 15u) / 4);   // words up to 16-byte alignment
    const size_t nvec = (words - head) / 4;
    uint4* const vec = reinterpret_cast<uint4*>(p + head);
    for (size_t k = first; k < nvec; k += nthreads) vec[k] = make_uint4(0u, 0u, 0u, 0u);
    for (size_t k = first; k < head; k += nthreads) p[k] = 0u;
    for (size_t k = head + 4 * nvec + first; k < words; k += nthreads) p[k] = 0u;
}

// One block per 1024 tiles, chained by a decoupled look-back (blocks ordered by ticket): exclusive scan of the
// per-tile counts = tile ranges (identifyTileRanges) and bucket cursors, the histogram of list lengths, the longest
// list, and -- by the last block of the chain -- R and the overflow flag.  The block that FINISHES last then writes
// the work items in longest-list-first order (counting sort on the length bin), in which the persistent compositing
// kernels pull their blocks so that the long lists -- each a serial chain through one warp -- start first and the
// short ones fill the tail.  (Until round 2 that was a second, dependent launch: 6 us of latency per step.)
__global__ void __launch_bounds__(1024)
tile_scan_kernel(int n, uint32_t* __restrict__ cursor, uint2* __restrict__ ranges, uint32_t* __restrict__ scratch,
                 uint32_t* __restrict__ num_rendered, uint32_t* __restrict__ host_R, unsigned long long capacity,
                 uint32_t list_limit, uint32_t call_flags, uint4* __restrict__ tile_order, int scan_blocks,
                 uint32_t* __restrict__ zero_fill, size_t zero_words, uint32_t* __restrict__ zero_fill2, size_t zero_words2)
{
    if ((int)blockIdx.x >= scan_blocks) {   // spare blocks: zero-fill duty (see tile_scan_single_kernel)
        const size_t nthreads = (size_t)(gridDim.x - scan_blocks) * 1024;
        const size_t first = (size_t)(blockIdx.x - scan_blocks) * 1024 + threadIdx.x;
        zero_fill_slice(zero_fill, zero_words, first, nthreads);
        zero_fill_slice(zero_fill2, zero_words2, first, nthreads);
        return;
    }
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t bins[256];
    __shared__ uint32_t bin_base[256];
    __shared__ uint32_t s_bid, s_carry, s_last;
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid < 256) bins[tid] = 0u;
    if (tid == 0) s_bid = atomicAdd(scratch + SCAN_TICKET, 1u);
    __syncthreads();
    const int bid = (int)s_bid;
    const int nblocks = scan_blocks;
    const int k = bid * 1024 + tid;
    const uint32_t c = k < n ? cursor[(size_t)k * CURSOR_STRIDE] : 0u;
    // length bins (255 - len/8) and the longest list; warp-aggregated: most tiles of a frame are empty
    {
        const uint32_t bin = k < n ? 255u - min(c >> 3, 255u) : 256u;
        const uint32_t peers = __match_any_sync(0xffffffffu, bin);
        if (lane == __ffs(peers) - 1 && bin < 256u) atomicAdd(&bins[bin], (uint32_t)__popc(peers));
        const uint32_t wmax = __reduce_max_sync(0xffffffffu, c);
        if (lane == 0 && wmax != 0u) atomicMax(scratch + SCAN_LONGEST, wmax);
    }
    uint32_t chunk_total;
    const uint32_t local = block_exclusive_scan_1024(c, warp_sums, tid, chunk_total);   // syncs: bins complete
    if (tid < 256 && bins[tid] != 0u) atomicAdd(scratch + SCAN_BINS + tid, bins[tid]);
    __threadfence();   // the bins and the maximum are visible before this block's word is published
    __syncthreads();
    if (tid < 32) {
        uint32_t* state = scratch + SCAN_STATE;
        if (lane == 0) st_volatile(state + bid, (chunk_total & LB_VAL) | (bid == 0 ? LB_INC : LB_AGG));
        const uint32_t carry = lookback_exclusive_warp(state, bid, lane);
        if (lane == 0) {
            if (bid > 0) st_volatile(state + bid, ((carry + chunk_total) & LB_VAL) | LB_INC);
            s_carry = carry;
        }
    }
    __syncthreads();
    const uint32_t run = s_carry + local;
    if (k < n) {
        // untouched tiles keep (0,0), as the reference's memset leaves them
        ranges[k] = c ? make_uint2(run, run + c) : make_uint2(0u, 0u);
        cursor[(size_t)k * CURSOR_STRIDE] = run;
    }
    if (bid == nblocks - 1 && tid == 0) {   // every other block has published, hence fenced its bins / maximum
        const uint32_t R = s_carry + chunk_total;
        const uint32_t longest = ld_volatile(scratch + SCAN_LONGEST);
        // overflow: more instances than the binning buffer holds, or a list longer than the sort kernels
        // that were launched can handle (the long-bucket sort is skipped while lists stay short)
        const bool overflow = (unsigned long long)R > capacity || longest > list_limit;
        num_rendered[0] = R;
        num_rendered[1] = overflow ? 1u : 0u;
        num_rendered[2] = 0u;   // work counters of the persistent compositing kernels
        num_rendered[3] = 0u;
        num_rendered[4] = longest;   // longest tile list: lets the long-bucket sort exit at once
        num_rendered[5] = call_flags;                      // what this call resolved its options to, for the backward
        num_rendered[6] = (uint32_t)capacity;              // call and for guava_buffer_layout_of_call
        num_rendered[7] = (uint32_t)(capacity >> 32);
        // one aligned 8-byte store: (R, longest) reach the pinned host words together, no system fences
        if (host_R != nullptr)
            *reinterpret_cast<volatile unsigned long long*>(host_R) =
                (unsigned long long)R | ((unsigned long long)(longest | (overflow ? 0x80000000u : 0u)) << 32);
    }

    // ---- work order: by the block that finishes last (all ranges and all length bins are then complete)
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(scratch + SCAN_DONE, 1u) == (uint32_t)nblocks - 1u ? 1u : 0u;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    {
        uint32_t dummy;
        const uint32_t mine = tid < 256 ? __ldcg(scratch + SCAN_BINS + tid) : 0u;
        const uint32_t ex = block_exclusive_scan_1024(mine, warp_sums, tid, dummy);
        if (tid < 256) { bin_base[tid] = ex; bins[tid] = 0u; }   // bins: now the cursors of the counting sort
    }
    __syncthreads();
    for (int k0 = 0; k0 < n; k0 += 1024) {
        const int kk = k0 + tid;
        const uint2 r = kk < n ? __ldcg(ranges + kk) : make_uint2(0u, 0u);
        const uint32_t cc = r.y - r.x;
        const uint32_t bin = kk < n ? 255u - min(cc >> 3, 255u) : 256u;   // 256: no tile
        const uint32_t peers = __match_any_sync(0xffffffffu, bin);
        const int leader = __ffs(peers) - 1;
        uint32_t pos = 0;
        if (lane == leader && bin < 256u) pos = bin_base[bin] + atomicAdd(&bins[bin], (uint32_t)__popc(peers));
        pos = __shfl_sync(0xffffffffu, pos, leader);
        if (bin < 256u) tile_order[pos + __popc(peers & ((1u << lane) - 1u))] = make_uint4((uint32_t)kk, r.x, r.y, 0u);
    }
}

// The same results from ONE block when the frame has few tiles (up to 1024 * ITEMS = 2048: every GUAVA training shape --
// 2 views of 512 x 512 are 2048 tiles): thread t owns ITEMS consecutive tiles, so the scan is a serial sum per thread
// plus one block scan, and the work order is written from registers.  No ticket, no look-back, no "last block" election,
// no second read of the ranges: one global round trip (the counters) instead of seven (13.3 -> 9.6 us under ncu).
template <int ITEMS>
__global__ void __launch_bounds__(1024)
tile_scan_single_kernel(int n, uint32_t* __restrict__ cursor, uint2* __restrict__ ranges,
                        uint32_t* __restrict__ num_rendered, uint32_t* __restrict__ host_R, unsigned long long capacity,
                        uint32_t list_limit, uint32_t call_flags, uint4* __restrict__ tile_order,
                        uint32_t* __restrict__ zero_fill, size_t zero_words, uint32_t* __restrict__ zero_fill2,
                        size_t zero_words2)
{
    // Blocks 1 .. gridDim.x-1 (guava_call_options::zero_fill): the scan is one block of latency on one SM; the other
    // SMs zero-fill a caller's buffer meanwhile -- the summed backward's [P,43] output, 17 MB that were a 5 us fill
    // kernel of their own in front of the step.
    if (blockIdx.x != 0) {
        const size_t nthreads = (size_t)(gridDim.x - 1) * 1024, first = (size_t)(blockIdx.x - 1) * 1024 + threadIdx.x;
        zero_fill_slice(zero_fill, zero_words, first, nthreads);     // the caller's buffer (guava_call_options::zero_fill)
        zero_fill_slice(zero_fill2, zero_words2, first, nthreads);   // the backward's per-Gaussian scratch (GeomPtrs::grad_geo)
        return;
    }
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t bins[256];
    __shared__ uint32_t bin_base[256];
    __shared__ uint32_t s_longest;
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid < 256) bins[tid] = 0u;
    if (tid == 0) s_longest = 0u;
    uint32_t c[ITEMS], sum = 0u, mx = 0u;
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const int k = tid * ITEMS + i;
        c[i] = k < n ? cursor[(size_t)k * CURSOR_STRIDE] : 0u;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const int k = tid * ITEMS + i;
        sum += c[i];
        mx = max(mx, c[i]);
        const uint32_t bin = k < n ? 255u - min(c[i] >> 3, 255u) : 256u;
        const uint32_t peers = __match_any_sync(0xffffffffu, bin);
        if (lane == __ffs(peers) - 1 && bin < 256u) atomicAdd(&bins[bin], (uint32_t)__popc(peers));
    }
    const uint32_t wmax = __reduce_max_sync(0xffffffffu, mx);
    if (lane == 0 && wmax != 0u) atomicMax(&s_longest, wmax);
    uint32_t R;
    uint32_t run = block_exclusive_scan_1024(sum, warp_sums, tid, R);   // syncs: bins and s_longest complete
    uint2 rg[ITEMS];
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const int k = tid * ITEMS + i;
        rg[i] = c[i] ? make_uint2(run, run + c[i]) : make_uint2(0u, 0u);   // untouched tiles keep (0,0)
        if (k < n) {
            ranges[k] = rg[i];
            cursor[(size_t)k * CURSOR_STRIDE] = run;
        }
        run += c[i];
    }
    if (tid == 0) {
        const uint32_t longest = s_longest;
        const bool overflow = (unsigned long long)R > capacity || longest > list_limit;
        num_rendered[0] = R;
        num_rendered[1] = overflow ? 1u : 0u;
        num_rendered[2] = 0u;
        num_rendered[3] = 0u;
        num_rendered[4] = longest;
        num_rendered[5] = call_flags;
        num_rendered[6] = (uint32_t)capacity;
        num_rendered[7] = (uint32_t)(capacity >> 32);
        if (host_R != nullptr)
            *reinterpret_cast<volatile unsigned long long*>(host_R) =
                (unsigned long long)R | ((unsigned long long)(longest | (overflow ? 0x80000000u : 0u)) << 32);
    }
    {
        uint32_t dummy;
        const uint32_t mine = tid < 256 ? bins[tid] : 0u;
        const uint32_t ex = block_exclusive_scan_1024(mine, warp_sums, tid, dummy);   // first sync: every `mine` is read
        if (tid < 256) { bin_base[tid] = ex; bins[tid] = 0u; }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < ITEMS; i++) {
        const int k = tid * ITEMS + i;
        const uint32_t bin = k < n ? 255u - min(c[i] >> 3, 255u) : 256u;
        const uint32_t peers = __match_any_sync(0xffffffffu, bin);
        const int leader = __ffs(peers) - 1;
        uint32_t pos = 0;
        if (lane == leader && bin < 256u) pos = bin_base[bin] + atomicAdd(&bins[bin], (uint32_t)__popc(peers));
        pos = __shfl_sync(0xffffffffu, pos, leader);
        if (bin < 256u) tile_order[pos + __popc(peers & ((1u << lane) - 1u))] = make_uint4((uint32_t)k, rg[i].x, rg[i].y, 0u);
    }
}

void launch_tile_scan(int total_tiles, const ImagePtrs& img, uint32_t* host_mapped_R, size_t capacity,
                      uint32_t list_limit, uint32_t call_flags, void* zero_fill, size_t zero_bytes, void* zero_fill2,
                      size_t zero_bytes2, cudaStream_t s)
{
    uint32_t* const zf = zero_bytes ? static_cast<uint32_t*>(zero_fill) : nullptr;
    const size_t zw = zero_bytes / 4;
    uint32_t* const zf2 = zero_bytes2 ? static_cast<uint32_t*>(zero_fill2) : nullptr;
    const size_t zw2 = zero_bytes2 / 4;
    static const int single = [] { const char* e = getenv("GUAVA_SCAN_SINGLE"); return e ? atoi(e) : 1; }();   // tuning knob
    // + fill blocks on the SMs the scan leaves idle, 16 KB of zeros per block and pass
    const size_t fill_bytes = (zf ? zero_bytes : 0) + (zf2 ? zero_bytes2 : 0);
    const unsigned nblocks = (unsigned)((total_tiles + 1023) / 1024);
    if (single && total_tiles <= 2048) {   // (eight tiles per thread for frames of up to 8192 tiles measured slower than the chain: 32 against 19 us at C2)
        const unsigned fill_blocks = (unsigned)std::min<size_t>((size_t)sm_count() - 1, (fill_bytes + 16383) / 16384);
        tile_scan_single_kernel<2><<<1 + fill_blocks, 1024, 0, s>>>(total_tiles, img.tile_cursor, img.ranges, img.num_rendered,
                                                                    host_mapped_R, (unsigned long long)capacity, list_limit,
                                                                    call_flags, img.tile_order, zf, zw, zf2, zw2);
        count_launch();
        return;
    }
    // larger frames: one scan block per 1024 tiles (they start first: blocks are dispatched in index order, and the
    // look-back is ordered by ticket, not by block index), the rest of the machine fills
    const unsigned spare = (unsigned)sm_count() > nblocks ? (unsigned)sm_count() - nblocks : 0u;
    unsigned fill_blocks = (unsigned)std::min<size_t>(spare, (fill_bytes + 16383) / 16384);
    if (fill_bytes && spare < 32u) {   // frames of > 100 k tiles: the scan fills the machine itself, plain memsets
        if (zf) cudaMemsetAsync(zf, 0, zero_bytes, s);
        if (zf2) cudaMemsetAsync(zf2, 0, zero_bytes2, s);
        fill_blocks = 0;
    }
    uint32_t* scratch = img.tile_cursor + (size_t)total_tiles * CURSOR_STRIDE;
    tile_scan_kernel<<<nblocks + fill_blocks, 1024, 0, s>>>(total_tiles, img.tile_cursor, img.ranges, scratch, img.num_rendered,
                                                            host_mapped_R, (unsigned long long)capacity, list_limit, call_flags,
                                                            img.tile_order, (int)nblocks, zf, zw, zf2, zw2);
    count_launch();
}

// -------------------------------------------------------------------- scatter
// Warp-cooperative: the 32 (view, Gaussian) entries of a warp are expanded to
// their (entry, tile) instances with a warp prefix sum, and every lane claims
// ONE bucket slot per round, so all of a warp's atomics are in flight together
// (the reference's duplicateWithKeys walks each rect serially in one thread,
// rasterizer_impl.cu:98-108).  Slot order inside a bucket is arbitrary
// (atomics); the tile sort fixes it.
__global__ void __launch_bounds__(256)
scatter_kernel(const Inputs in, const GeomPtrs g, const uint32_t* __restrict__ num_rendered,
               uint32_t* __restrict__ cursor, uint64_t* __restrict__ keys, unsigned long long capacity)
{
    if (num_rendered[1]) return;   // capacity overflow: the host re-runs with a larger buffer
    const size_t n = (size_t)in.V * in.P;
    const size_t gid = (size_t)blockIdx.x * 256 + threadIdx.x;
    const int lane = threadIdx.x & 31;
    uint32_t rx = 0, ry = 0, w = 0, cnt = 0, v = 0, klo = 0, khi = 0;
    if (gid < n) {
        const ushort4 rc = g.rects[gid];
        if (rc.z > rc.x && rc.w > rc.y) {
            rx = rc.x; ry = rc.y; w = (uint32_t)(rc.z - rc.x);
            cnt = w * (uint32_t)(rc.w - rc.y);
            v = (uint32_t)(gid / in.P);
            klo = (uint32_t)(gid - (size_t)v * in.P);
            khi = __float_as_uint(g.depths[gid]);
        }
    }
    uint32_t incl = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
    const uint32_t excl = incl - cnt;
    // (Four rounds per iteration, their atomics in flight together, measured the same 19 us: profiles/r02_v16_scan_single.txt.)
    for (uint32_t base = 0; base < total; base += 32) {
        const uint32_t t = base + lane;
        // owner = last lane whose exclusive offset is <= t (binary search over the warp)
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
        const uint32_t o_klo = __shfl_sync(0xffffffffu, klo, owner);
        const uint32_t o_khi = __shfl_sync(0xffffffffu, khi, owner);
        if (t < total) {
            const uint32_t local = t - o_excl;
            const uint32_t yy = local / o_w, xx = local - yy * o_w;
            const size_t tile = (size_t)o_v * in.tiles + (size_t)(o_ry + yy) * in.gx + (o_rx + xx);
            const uint32_t pos = atomicAdd(cursor + tile * CURSOR_STRIDE, 1u);
            if (pos < capacity) keys[pos] = ((uint64_t)o_khi << 32) | o_klo;
        }
    }
}

void launch_scatter(const Inputs& in, const GeomPtrs& g, const ImagePtrs& img, const BinningPtrs& b,
                    const int* radii, cudaStream_t s)
{
    (void)radii;
    const size_t n = (size_t)in.V * in.P;
    if (n == 0) return;
    scatter_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(in, g, img.num_rendered, img.tile_cursor,
                                                                b.keys_unsorted, (unsigned long long)b.capacity);
    count_launch();
}

// ------------------------------------------------------------------ tile sort
// In-place bitonic network in its "all comparators ascending" form (flip +
// disperse steps).  Because every comparator moves the minimum to the lower
// index, elements beyond n behave as +inf without being stored, so any n is
// sorted in place without padding.
template <typename Ptr>
__device__ __forceinline__ void bitonic_sort_ascending(Ptr s, uint32_t n, uint32_t npow2, int tid, int nthreads)
{
    for (uint32_t k = 2; k <= npow2; k <<= 1) {
        // flip: i ^ (k-1)
        __syncthreads();
        const uint32_t half = k >> 1;
        for (uint32_t p = tid; p < (npow2 >> 1); p += nthreads) {
            const uint32_t blk = p / half, t = p - blk * half;
            const uint32_t i = blk * k + t, j = blk * k + (k - 1 - t);
            if (j < n) {
                const uint64_t a = s[i], b = s[j];
                if (a > b) { s[i] = b; s[j] = a; }
            }
        }
        for (uint32_t d = k >> 2; d > 0; d >>= 1) {
            __syncthreads();
            for (uint32_t p = tid; p < (npow2 >> 1); p += nthreads) {
                const uint32_t i = 2 * p - (p & (d - 1)), j = i + d;
                if (j < n) {
                    const uint64_t a = s[i], b = s[j];
                    if (a > b) { s[i] = b; s[j] = a; }
                }
            }
        }
    }
    __syncthreads();
}

__device__ __forceinline__ uint32_t next_pow2(uint32_t n)
{
    return n <= 1 ? 1u : 1u << (32 - __clz(n - 1));
}

// Buckets up to SORT_SMALL keys: bitonic network with the keys in REGISTERS.  256 threads hold
// E consecutive keys each (N = 256*E slots, missing keys = +inf).  A compare-exchange at distance j
// is register-to-register for j < E, a warp shuffle for j < 32*E, and goes through shared memory
// (transposed layout, conflict-free) only for the few steps with j >= 32*E -- 6 of 55 steps at
// N = 1024 -- where the first version synchronised the block after every step.
template <int E, int T = 256>
__device__ __forceinline__ void tile_sort_registers(const uint64_t* __restrict__ src, uint32_t n, uint64_t* skeys,
                                                    uint32_t* __restrict__ point_list, uint64_t* __restrict__ keys_sorted,
                                                    uint64_t tile_hi, int tid)
{
    constexpr int N = T * E;
    uint64_t key[E];
#pragma unroll
    for (int e = 0; e < E; e++) {
        const uint32_t i = (uint32_t)(tid * E + e);
        key[e] = i < n ? src[i] : ~0ull;
    }
#pragma unroll
    for (int k = 2; k <= N; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j >= 1; j >>= 1) {
            if (j < E) {
#pragma unroll
                for (int e = 0; e < E; e++) {
                    if ((e & j) == 0) {
                        const bool asc = (((tid * E + e) & k) == 0);
                        const uint64_t a = key[e], b = key[e | j];
                        const bool sw = (a > b) == asc;
                        key[e] = sw ? b : a;
                        key[e | j] = sw ? a : b;
                    }
                }
            } else if (j < 32 * E) {
                // one 64-bit compare per element: keys are distinct (the Gaussian index is part of the
                // key), and between two equal +inf pads either choice stores the same value
                const int tj = j / E;   // partner thread distance, within the warp
                // k >= 2j >= 2E here, so the direction bit (index & k) is the same for all E keys of a thread
                const bool keep_max = ((tid & tj) == 0) != (((tid * E) & k) == 0);
#pragma unroll
                for (int e = 0; e < E; e++) {
                    const uint64_t other = __shfl_xor_sync(0xffffffffu, key[e], tj);
                    const bool take_other = (other < key[e]) != keep_max;
                    key[e] = take_other ? other : key[e];
                }
            } else {
                const int tj = j / E;
                const bool keep_max = ((tid & tj) == 0) != (((tid * E) & k) == 0);
                __syncthreads();
#pragma unroll
                for (int e = 0; e < E; e++) skeys[e * T + tid] = key[e];
                __syncthreads();
#pragma unroll
                for (int e = 0; e < E; e++) {
                    const uint64_t other = skeys[e * T + (tid ^ tj)];
                    const bool take_other = (other < key[e]) != keep_max;
                    key[e] = take_other ? other : key[e];
                }
            }
        }
    }
#pragma unroll
    for (int e = 0; e < E; e++) {
        const uint32_t i = (uint32_t)(tid * E + e);
        if (i < n) {
            point_list[i] = (uint32_t)key[e];
            if (keys_sorted != nullptr) keys_sorted[i] = tile_hi | (key[e] >> 32);
        }
    }
}

__global__ void __launch_bounds__(256)
tile_sort_small_kernel(const uint4* __restrict__ tile_order, const uint32_t* __restrict__ num_rendered,
                       const uint64_t* __restrict__ keys, uint32_t* __restrict__ point_list,
                       uint64_t* __restrict__ keys_sorted, int tiles_per_view)
{
    __shared__ uint64_t skeys[2048];
    if (num_rendered[1]) return;
    // blocks take the tiles longest list first (the compositors' work order): all non-empty buckets are
    // resident in one wave at the GUAVA shapes, and this deals the long sorts evenly over the SMs
    // instead of leaving them clustered where the image centre falls in row-major tile order
    const uint4 item = tile_order[blockIdx.x];
    const uint2 r = make_uint2(item.y, item.z);
    const uint32_t n = r.y - r.x;
    if (n == 0 || n > 2048) return;   // block-uniform; longer buckets: tile_sort_large_kernel
    const int tid = threadIdx.x;
    const uint64_t tile_hi = (uint64_t)(item.x % (uint32_t)tiles_per_view) << 32;
    uint64_t* ks = keys_sorted != nullptr ? keys_sorted + r.x : nullptr;
    if (n <= 256) tile_sort_registers<1>(keys + r.x, n, skeys, point_list + r.x, ks, tile_hi, tid);
    else if (n <= 512) tile_sort_registers<2>(keys + r.x, n, skeys, point_list + r.x, ks, tile_hi, tid);
    else if (n <= 1024) tile_sort_registers<4>(keys + r.x, n, skeys, point_list + r.x, ks, tile_hi, tid);
    else tile_sort_registers<8>(keys + r.x, n, skeys, point_list + r.x, ks, tile_hi, tid);
}

// Buckets longer than SORT_SMALL are rare (none at the GUAVA shapes): one persistent
// 1024-thread CTA per SM walks the tile table and sorts them -- in 128 KB of shared
// memory up to SORT_LARGE keys, beyond that with the same network in place in global
// memory (L2-resident).  When there is nothing to do the launch costs a few microseconds.
static constexpr uint32_t SORT_SMALL = 2048;    // 16 KB of keys, 256 threads
static constexpr uint32_t SORT_MID = 4096;      // 512 threads x 8 keys in registers (the real GUAVA training size, P = 202 k, has lists of 2-3 k)
static constexpr uint32_t SORT_LARGE = 16384;   // 128 KB of keys, 1024 threads

__global__ void __launch_bounds__(1024)
tile_sort_large_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ num_rendered,
                       uint64_t* __restrict__ keys, uint32_t* __restrict__ point_list,
                       uint64_t* __restrict__ keys_sorted, int tiles_per_view, int total_tiles)
{
    extern __shared__ uint64_t skeys[];
    if (num_rendered[1] || num_rendered[4] <= SORT_MID) return;
    const int tid = threadIdx.x;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const uint2 r = ranges[tile];
        const uint32_t n = r.y - r.x;
        if (n <= SORT_SMALL) continue;   // block-uniform
        const uint64_t tile_hi = (uint64_t)(tile % tiles_per_view) << 32;
        if (n <= SORT_MID) continue;   // block-uniform: tile_sort_mid_kernel's lists
        if (n <= SORT_LARGE) {
            for (uint32_t k = tid; k < n; k += 1024) skeys[k] = keys[r.x + k];
            bitonic_sort_ascending(skeys, n, next_pow2(n), tid, 1024);
            for (uint32_t k = tid; k < n; k += 1024) {
                const uint64_t key = skeys[k];
                point_list[r.x + k] = (uint32_t)key;
                if (keys_sorted != nullptr) keys_sorted[r.x + k] = tile_hi | (key >> 32);
            }
        } else {
            volatile uint64_t* seg = keys + r.x;
            bitonic_sort_ascending(seg, n, next_pow2(n), tid, 1024);
            for (uint32_t k = tid; k < n; k += 1024) {
                const uint64_t key = seg[k];
                point_list[r.x + k] = (uint32_t)key;
                if (keys_sorted != nullptr) keys_sorted[r.x + k] = tile_hi | (key >> 32);
            }
        }
        __syncthreads();   // skeys is reused by the next tile
    }
}

// Buckets of SORT_SMALL + 1 .. SORT_MID keys -- the real GUAVA training size (P = 202 k) has lists of 2-3 k: the same
// register-resident network as the small buckets, 512 threads x 8 keys (the steps at distance >= 256 go through shared
// memory).  A kernel of its own with 32 KB of static shared memory and 48 registers, so that two to three CTAs fit an
// SM: inside the long-bucket kernel (128 KB of dynamic shared memory, one 1024-thread CTA per SM) these sorts took
// 0.040 ms of the P = 202 k step.  The tile
// order lists the tiles longest first in bins of 8 keys, every list of 2040 keys or more in the first bin: the CTAs
// stride over that head of the order and stop at the first shorter list.
__global__ void __launch_bounds__(512)
tile_sort_mid_kernel(const uint4* __restrict__ tile_order, const uint32_t* __restrict__ num_rendered,
                     const uint64_t* __restrict__ keys, uint32_t* __restrict__ point_list,
                     uint64_t* __restrict__ keys_sorted, int tiles_per_view, int total_tiles)
{
    __shared__ uint64_t skeys[SORT_MID];
    if (num_rendered[1] || num_rendered[4] <= SORT_SMALL) return;
    const int tid = threadIdx.x;
    for (int w = blockIdx.x; w < total_tiles; w += gridDim.x) {
        const uint4 item = tile_order[w];
        const uint32_t n = item.z - item.y;
        if (n < 2040u) break;                          // block-uniform: the rest of the order is shorter
        if (n <= SORT_SMALL || n > SORT_MID) continue;   // block-uniform: the small / long-bucket kernels' lists
        const uint64_t tile_hi = (uint64_t)(item.x % (uint32_t)tiles_per_view) << 32;
        tile_sort_registers<8, 512>(keys + item.y, n, skeys, point_list + item.y,
                                     keys_sorted != nullptr ? keys_sorted + item.y : nullptr, tile_hi, tid);
        __syncthreads();   // skeys is reused by the next tile
    }
}

uint32_t tile_sort_small_limit() { return SORT_SMALL; }

void launch_tile_sort(const Inputs& in, const GeomPtrs& g, const ImagePtrs& img, const BinningPtrs& b, bool with_large,
                      cudaStream_t s)
{
    const int total_tiles = in.V * in.tiles;
    if (total_tiles == 0) return;
    // function attributes are per device: one flag per device ordinal (setting it twice is harmless)
    static std::atomic<bool> attr_set[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || !attr_set[dev].load(std::memory_order_acquire)) {
        cudaFuncSetAttribute(tile_sort_large_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)(SORT_LARGE * sizeof(uint64_t)));
        if (dev >= 0 && dev < 64) attr_set[dev].store(true, std::memory_order_release);
    }
    tile_sort_small_kernel<<<total_tiles, 256, 0, s>>>(img.tile_order, img.num_rendered, b.keys_unsorted, b.point_list,
                                                       b.keys_sorted, in.tiles);
    count_launch();
    if (!with_large) return;
    const int mid_grid = total_tiles < 3 * sm_count() ? total_tiles : 3 * sm_count();
    tile_sort_mid_kernel<<<mid_grid, 512, 0, s>>>(img.tile_order, img.num_rendered, b.keys_unsorted, b.point_list,
                                                   b.keys_sorted, in.tiles, total_tiles);
    count_launch();
    const int big_grid = total_tiles < sm_count() ? total_tiles : sm_count();
    tile_sort_large_kernel<<<big_grid, 1024, SORT_LARGE * sizeof(uint64_t), s>>>(
        img.ranges, img.num_rendered, b.keys_unsorted, b.point_list, b.keys_sorted, in.tiles, total_tiles);
    count_launch();
}

}  // namespace guava
