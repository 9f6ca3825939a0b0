// Backward compositing: every 8x4 pixel block of a 16x16 tile (one warp = one CTA) walks the
// tile's depth-sorted list back to front and produces the per-Gaussian screen-space gradients.
//
// Reference behaviour: DGR32/cuda_rasterizer/backward.cu:452-638 (renderCUDA),
// which issues 39 global atomicAdd per contributing (pixel, Gaussian) pair
// (:593,603,626-635) and keeps accum_rec[32] + last_color[32] + dL_dpixel[32]
// live per thread (173 registers).
//
// B200 design:
//  * same asynchronously staged, double-buffered instance batches as the forward kernel
//    (SplatRec + 128-byte feature row per instance), walked in reverse;
//  * warp-footprint culling (splat_may_touch, internal.cuh): each warp owns an
//    8x4 pixel block and walks only the staged splats that can reach
//    alpha >= 1/255 somewhere in that block;
//  * the reference's per-channel recurrence
//        accum_rec[ch] = last_alpha*last_color[ch] + (1-last_alpha)*accum_rec[ch]
//        dL_dalpha += (c[ch] - accum_rec[ch]) * dL_dpixel[ch]
//    is linear in the channel index, so only its contraction with dL_dpixel is
//    carried:  A = last_alpha*last_d + (1-last_alpha)*A,  d = <c, dL_dpixel>.
//    That removes 64 live registers per thread;
//  * dL_dcolor[g][ch] = sum over pixels of w(pixel,g) * dL_dpixel[pixel][ch] is
//    a (32 pixels) x (32 channels) matrix-vector product per (warp, Gaussian)
//    whose matrix -- the warp's upstream gradients -- never changes.  Each lane
//    keeps column `lane` of that matrix in registers (transposed once per tile
//    through shared memory); the 32 blending weights of the warp are exchanged
//    through 128 bytes of shared memory and every lane does 32 FFMAs.  The
//    reference's 32 atomics per PIXEL become one 128-byte reduction per WARP,
//    without the shuffle tree a per-pixel formulation needs;
//  * the seven geometric partials (moments of s = dL/dG*G about the pixel offset,
//    G*dL/dalpha, w*dL/dinvdepth) are reduced in registers with a transposing
//    butterfly (7 shuffles per splat; lanes 4k..4k+3 end with four partial sums
//    of moment k and each adds its own) into a packed per-Gaussian record
//    (GeomPtrs::grad_geo) that the backward preprocess kernel consumes.  The shared-memory data pipe is the busiest unit
//    of this kernel: only what every lane needs in full -- feature
//    rows, the 32 blending weights -- goes through shared memory.
#include "internal.cuh"
#include "tma.cuh"
#include <cuda.h>
#include <cstdlib>
#include <cstring>

namespace guava {

// Survivor records staged PAIR-INTERLEAVED as in the forward kernel (FwdStage): pair k = survivors 2k, 2k+1, so one
// LDS.128 delivers a quantity of both as the halves of packed f32x2 operands.
struct __align__(128) BwdStage {
    float feat[32][NUM_CHANNELS];   // 4 KB
    float4 xy[16];                  // x0 x1 y0 y1
    float4 cab[16];                 // conic.x0 conic.x1 -conic.y0 -conic.y1
    float4 cop[16];                 // conic.z0 conic.z1 opacity0 opacity1
    float2 rd[16];                  // rdepth0 rdepth1
};
static_assert(sizeof(BwdStage) >= 32 * 32 * sizeof(float), "a stage doubles as the transpose scratch");

// Transposing butterfly over two 8-vectors at once: afterwards lanes 4k .. 4k+3 hold in a[0] / b[0] four partial
// sums (over 8 lanes each) of element k of a / b.  7 shuffles per vector (4 + 2 + 1, each step exchanges half of what
// is left).  The four partials are not combined here: each of those lanes adds its own to the same word of the
// packed record -- two more shuffle steps on the data pipe measured 2.8 % slower than the 21 extra L2 reductions
// per splat (profiles/r01_v46).
__device__ __forceinline__ void warp_transpose_reduce8x2(float (&a)[8], float (&b)[8], int lane)
{
    int mask = 16;
#pragma unroll
    for (int half = 4; half >= 1; half >>= 1, mask >>= 1) {
        const bool upper = (lane & mask) != 0;
#pragma unroll
        for (int i = 0; i < half; i++) {
            const float sa = upper ? a[i] : a[i + half], ka = upper ? a[i + half] : a[i];
            const float sb = upper ? b[i] : b[i + half], kb = upper ? b[i + half] : b[i];
            a[i] = ka + __shfl_xor_sync(0xffffffffu, sa, mask);
            b[i] = kb + __shfl_xor_sync(0xffffffffu, sb, mask);
        }
    }
}

// One warp (= one CTA) per 8x4 pixel block, as in the forward kernel: no block
// barrier, private double-buffered stage of culled survivors.
// Gradients leave the SM as 4-byte reductions: one 128-byte request from all 32 lanes per (warp, Gaussian) for
// dL_dcolors plus 28 lanes for the packed record.  (16-byte vector reductions, red.global.add.v4.f32 from 8 lanes,
// cost three gathering shuffles per splat on the data pipe and measured 2 % slower, profiles/r01_v43; removed.)
// MASKS = 1 (option "taken_masks"): see render_forward.cu -- the forward has marked, per block, the list entries it got
// nothing from; they are skipped here without a footprint test, record gather or staging.
// TMA = 1 (option "tma_staging"): feature rows staged by cp.async.bulk.tensor gather4.
// INV = 0: no upstream gradient for the inverse depth (GUAVA's losses never read the depth output,
// utils/loss_utils.py:89-138): the per-splat 1/depth operand, its term of the dot product and the seventh moment drop out.
// NST: number of staging buffers.  2 = chunk c+1 is staged while chunk c is blended; 3 = chunk c+2 is (a chunk with
// one or two survivors is blended in a few hundred cycles, less than the L2 round trip of the next chunk's rows: the
// wait behind cp.async.wait_group held 7 % of the NST = 2 kernel's stall samples, profiles/r02_v6).
// MINB: resident one-warp CTAs per SM the kernel is compiled for (register cap 65536 / (32 MINB)).
template <int MASKS, int TMA, int INV, int NST, int MINB = 12>
__global__ void __launch_bounds__(32, MINB)
render_backward_kernel(const Inputs in, const __grid_constant__ CUtensorMap fmap, const SplatRec* __restrict__ recs,
                       const uint2* __restrict__ ranges, const uint4* __restrict__ tile_order,
                       const uint32_t* __restrict__ point_list, const uint8_t* __restrict__ taken_marks,
                       const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                       const float* __restrict__ dL_dpix, const float* __restrict__ dL_dinvdepth_pix,
                       float* __restrict__ dL_dcolor, uint32_t color_rs, size_t color_vs,
                       float* __restrict__ grad_geo, uint32_t* __restrict__ num_rendered)
{
    // dL_dcolor: row `id` of view v starts at dL_dcolor + v * color_vs + id * color_rs (floats).  [V,P,32] is
    // (32, 32 P); a sum over the views into columns of a packed [P,43] row is (43, 0) -- the reductions are atomic,
    // so the views add up in place (guava_rasterizer_backward_batch_summed).
    __shared__ BwdStage stage[NST];
    __shared__ __align__(16) float ex[4][32];   // blending weights of a survivor, one per pixel: two pairs in flight
    __shared__ __align__(16) uint32_t srows[NST][32];   // TMA: tensor-map row of the survivor in each slot
    __shared__ __align__(8) uint64_t bars[NST];         // TMA: one transaction barrier per stage

    // The forward's binning overflowed (a replayed capture whose capacity was outgrown): its instance kernels
    // skipped themselves, the lists this kernel would walk were never written.  Gradients stay zero; the caller
    // learns of it through the status words / num_rendered[1].
    if (num_rendered[1]) return;
    uint32_t* const work_counter = num_rendered + 3;
    // MASKS: the forward's marks lie behind point_list (api.cu compute_layout), whose length is in the call's header;
    // a forward that ran without marking says so there, and every entry is then walked (slow, correct)
    const uint8_t* marks = nullptr;
    size_t marks_plane = 0;
    if (MASKS && ((num_rendered[5] >> 8) & 1u) != 0u) {
        const size_t cap = (size_t)num_rendered[6] | ((size_t)num_rendered[7] << 32);
        marks = reinterpret_cast<const uint8_t*>(point_list) + ((cap * 4 + 255) / 256) * 256;
        marks_plane = (cap + 255) / 256 * 256;
    }
    (void)taken_marks;
    const int lane = threadIdx.x;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const size_t HW = (size_t)in.H * in.W;
    uint32_t parity = 0;   // ex buffers (parity, parity+1) & 3 are used by the next pair
    float* const exbase = &ex[0][0];
    uint32_t phase = 0;    // TMA: bit s = parity the next wait on bars[s] uses
    if (TMA) {
        if (lane == 0) {
            for (int q = 0; q < NST; q++) tma::mbar_init(&bars[q], 1);
            tma::fence_mbar_init();
        }
        __syncwarp();
    }

    // persistent warps pulling 8x4 pixel blocks from a work counter (see render_forward.cu)
    const uint32_t blocks_per_view = (uint32_t)in.tiles * 8u;
    const uint32_t total_work = blocks_per_view * (uint32_t)in.V;
    auto claim_work = [&]() -> uint32_t { return lane == 0 ? atomicAdd(work_counter, 1u) : 0u; };

  uint32_t claimed = claim_work();
  for (uint32_t work = __shfl_sync(0xffffffffu, claimed, 0); work < total_work; work = __shfl_sync(0xffffffffu, claimed, 0)) {
    claimed = claim_work();   // the next block's claim: broadcast when that block begins (see render_forward.cu)
    const uint4 item = tile_order[work >> 3];   // longest lists first: (tile, range.x, range.y)
    const uint32_t gtile = item.x;
    const int sub = (int)(work & 7u);
    const int v = (int)(gtile / (uint32_t)in.tiles);
    const int tile = (int)(gtile - (uint32_t)v * (uint32_t)in.tiles);
    const int tile_x = tile % in.gx, tile_y = tile / in.gx;
    const uint32_t bx = tile_x * TILE_X + (sub & 1) * 8, by = tile_y * TILE_Y + (sub >> 1) * 4;
    if (bx >= (uint32_t)in.W || by >= (uint32_t)in.H) continue;
    const uint32_t px = bx + (lane & 7), py = by + (lane >> 3);
    const bool inside = px < (uint32_t)in.W && py < (uint32_t)in.H;
    const float pxf = (float)px, pyf = (float)py;
    const float wx0 = (float)bx, wy0 = (float)by;
    const size_t pix = (size_t)py * in.W + px;

    const uint32_t last_contributor = inside ? n_contrib[(size_t)v * HW + pix] : 0u;
    const int total = (int)__reduce_max_sync(0xffffffffu, last_contributor);   // list entries this block walks
    if (total == 0) continue;
    const int nchunks = (total + 31) >> 5;
    const uint2 range = make_uint2(item.y, item.z);

    // upstream gradients of this pixel, all channels (row of the block's 32x32 matrix) ...
    float2 g[NUM_CHANNELS / 2];   // channel pairs, for packed FFMA2 (sm_100 fma.rn.f32x2)
    float g_inv = 0.0f, bgdot = 0.0f;
    const float T_final = inside ? final_Ts[(size_t)v * HW + pix] : 0.0f;
    {
        const float* __restrict__ bg = in.background + (size_t)v * NUM_CHANNELS;
        const float* __restrict__ gp = dL_dpix + (size_t)v * NUM_CHANNELS * HW + pix;
#pragma unroll
        for (int ch = 0; ch < NUM_CHANNELS / 2; ch++) {
            g[ch].x = inside ? gp[(size_t)(2 * ch) * HW] : 0.0f;
            g[ch].y = inside ? gp[(size_t)(2 * ch + 1) * HW] : 0.0f;
            bgdot = fmaf(bg[2 * ch], g[ch].x, bgdot);
            bgdot = fmaf(bg[2 * ch + 1], g[ch].y, bgdot);
        }
        if (INV && dL_dinvdepth_pix != nullptr && inside) g_inv = dL_dinvdepth_pix[(size_t)v * HW + pix];
    }
    // ... and channel `lane` of all 32 pixels (column), through an XOR-swizzled 32x32 tile
    // in the not-yet-used staging memory (conflict-free both ways).
    float2 gT[16];   // gT[k] = channel `lane` of pixels 2k, 2k+1
    {
        float* scratch = &stage[0].feat[0][0];
#pragma unroll
        for (int c = 0; c < NUM_CHANNELS / 2; c++) {
            scratch[lane * 32 + ((2 * c) ^ lane)] = g[c].x;
            scratch[lane * 32 + ((2 * c + 1) ^ lane)] = g[c].y;
        }
        __syncwarp();
#pragma unroll
        for (int p = 0; p < 16; p++) {
            gT[p].x = scratch[(2 * p) * 32 + (lane ^ (2 * p))];
            gT[p].y = scratch[(2 * p + 1) * 32 + (lane ^ (2 * p + 1))];
        }
    }
    __syncwarp();

    const SplatRec* __restrict__ vrecs = recs + (size_t)v * in.P;
    const float* __restrict__ vfeat = in.colors + (size_t)v * in.colors_vs;
    const uint32_t vrow0 = (uint32_t)(((size_t)v * in.colors_vs) / NUM_CHANNELS);   // TMA: first tensor-map row of this view

    constexpr uint32_t NO_ID = 0xffffffffu;
    // chunk c, lane L  <->  list position  total-1-(32c+L)   (back to front); same software
    // pipeline as the forward kernel (ids two chunks ahead, SplatRecs one chunk ahead).
    auto load_id = [&](int c) -> uint32_t {
        const int k = (c << 5) + lane;
        if (k >= total) return NO_ID;
        const uint32_t p = range.x + (uint32_t)(total - 1 - k);
        const uint32_t id = point_list[p];
        // MASKS: an entry the forward marked for this block is no entry at all (both loads are in flight together)
        if (MASKS && marks != nullptr) return marks[(size_t)sub * marks_plane + p] != 0 ? NO_ID : id;
        return id;
    };
    // Unconditional (lanes past the end of the list re-read record 0 and are masked by `keep`): a predicated
    // load into loop-carried registers is compiled as load-to-temporary + predicated moves, and the first of
    // those moves stalled the warp for the whole L2 round trip right after the load was issued
    // (profiles/r01_v8: 7.7 % / 5 % of the forward / backward kernel's stall samples on that one instruction).
    auto load_rec = [&](uint32_t id, float4& c0, float4& c1) {
        const uint32_t r = id != NO_ID ? id : 0u;
        const float4* __restrict__ rp = reinterpret_cast<const float4*>(vrecs + r);
        c0 = rp[0];
        c1 = rp[1];
    };
    auto stage_chunk = [&](int sq, uint32_t id, const float4& c0, const float4& c1) -> uint32_t {   // sq: staging buffer
        const bool keep = id != NO_ID &&
                          (MASKS || splat_may_touch(c0.x, c0.y, c0.z, c0.w, c1.x, c1.w, wx0, wx0 + 7.0f, wy0, wy0 + 3.0f));
        const uint32_t mask = __ballot_sync(0xffffffffu, keep);
        if (mask != 0u) {
            BwdStage& st = stage[sq];
            if (keep) {
                const int slot = __popc(mask & lt_mask);
                const int o = (slot >> 1) * 4 + (slot & 1);
                float* q;
                q = &st.xy[0].x + o;   q[0] = c0.x; q[2] = c0.y;
                q = &st.cab[0].x + o;  q[0] = c0.z; q[2] = -c0.w;
                q = &st.cop[0].x + o;  q[0] = c1.x; q[2] = c1.y;
                if (INV) (&st.rd[0].x)[slot] = c1.z;
                if (TMA) {
                    srows[sq][slot] = vrow0 + id;
                } else {
                    // the lane's own 128-byte feature row as eight 16-byte asynchronous copies, piece order rotated by
                    // the slot so that the lanes of one instruction fall on distinct bank groups (render_forward.cu)
                    const char* src = reinterpret_cast<const char*>(vfeat + (size_t)id * NUM_CHANNELS);
                    char* dstf = reinterpret_cast<char*>(&st.feat[slot][0]);
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        const int piece = ((j + slot) & 7) * 16;
                        tma::cp_async16(dstf + piece, src + piece);
                    }
                }
            }
            if (TMA) {   // four survivor rows per gather4, last group padded with the first survivor's row (render_forward.cu)
                const int nsurv = __popc(mask);
                const int ngroups = (nsurv + 3) >> 2;
                const uint32_t first_row = vrow0 + __shfl_sync(0xffffffffu, id, __ffs(mask) - 1);
                if (lane < 3 && nsurv + lane < 4 * ngroups) srows[sq][nsurv + lane] = first_row;
                __syncwarp();
                if (lane == 0) {
                    tma::mbar_arrive_expect_tx(&bars[sq], (uint32_t)ngroups * 512u);
                    for (int gq = 0; gq < ngroups; gq++) {
                        const uint4 r = *reinterpret_cast<const uint4*>(&srows[sq][4 * gq]);
                        tma::tensor_gather4_g2s(&st.feat[4 * gq][0], &fmap, 0, (int)r.x, (int)r.y, (int)r.z, (int)r.w, &bars[sq]);
                    }
                }
            }
        }
        if (!TMA) tma::cp_async_commit();   // one group per chunk, empty or not: the waits count groups
        return mask;
    };

    float T = T_final;
    float A = 0.0f, last_alpha = 0.0f, last_d = 0.0f;
    const size_t gbase = (size_t)v * in.P;
    float* const colbase = dL_dcolor + (size_t)v * color_vs;
    float* const geobase = grad_geo + gbase * 8;

    // Software pipeline, AHEAD = NST - 1 chunks deep: while chunk c is blended, chunks c+1 .. c+AHEAD are staged (rows in
    // flight), the records of chunk c+AHEAD+1 are in flight into a0/a1 and the ids of chunk c+AHEAD+2 into id_q[AHEAD+2].
    // id_q[k]: ids of chunk c+k (id_q[0] gives the gradient addresses of the chunk being blended).
    constexpr int AHEAD = NST - 1;
    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
    uint32_t id_q[AHEAD + 3], mask_q[AHEAD + 1];
#pragma unroll
    for (int k = 0; k < AHEAD + 2; k++) id_q[k] = load_id(k);
#pragma unroll
    for (int k = 0; k < AHEAD; k++) {   // chunks 0 .. AHEAD-1 staged before the loop
        load_rec(id_q[k], a0, a1);
        if (k < nchunks) mask_q[k] = stage_chunk(k, id_q[k], a0, a1);
        else { mask_q[k] = 0u; if (!TMA) tma::cp_async_commit(); }
    }
    load_rec(id_q[AHEAD], a0, a1);
    int s = 0;   // c % NST
    for (int c = 0; c < nchunks; c++) {
        int sq = s + AHEAD; if (sq >= NST) sq -= NST;   // (c + AHEAD) % NST
        mask_q[AHEAD] = 0u;
        if (c + AHEAD < nchunks) mask_q[AHEAD] = stage_chunk(sq, id_q[AHEAD], a0, a1);
        else if (!TMA) tma::cp_async_commit();
        load_rec(id_q[AHEAD + 1], a0, a1);
        id_q[AHEAD + 2] = load_id(c + AHEAD + 2);
        const uint32_t mask_cur = mask_q[0];
        const uint32_t id_cur = id_q[0];
        if (mask_cur != 0u) {
            if (TMA) {
                tma::mbar_wait(&bars[s], (phase >> s) & 1u);
                phase ^= 1u << s;
            } else {
                tma::cp_async_wait<AHEAD>();   // everything but the newest AHEAD groups has landed
            }
            __syncwarp();
            const BwdStage& stg = stage[s];
            int kp = 0;
            const float4* __restrict__ featp = reinterpret_cast<const float4*>(&stage[s].feat[0][0]);
            const uint32_t pos0 = (uint32_t)(total - 1 - (c << 5));   // list position of chunk lane 0
            uint32_t m = mask_cur;
            // Survivors are walked two at a time.  Everything that does not depend on the running
            // transmittance -- record loads, alpha, the moments' products -- is evaluated for both at once
            // with packed f32x2 instructions, the two 32-channel dot products side by side (independent
            // latency chains); the T / accum recurrences then run in list order; the exchange, the two
            // matrix-vector products and the reductions are again independent.
            for (; m != 0u; kp++, featp += 2 * (NUM_CHANNELS / 4)) {
                const int jl0 = __ffs(m) - 1;
                m &= m - 1;
                const bool has1 = m != 0u;
                const int jl1 = has1 ? __ffs(m) - 1 : jl0;
                m &= m - 1;
                // both alpha chains as packed f32x2 instructions: per half the very operations, in the very order, of
                // the forward kernel's eval_pair (bit-identical alpha); the second half is stale data if !has1
                const float4 pa = stg.xy[kp], pb = stg.cab[kp], pc = stg.cop[kp];
                const float2 rdp = INV ? stg.rd[kp] : make_float2(0.0f, 0.0f);
                const float2 dxp = __fadd2_rn(f2(pa.x, pa.y), f2(-pxf));
                const float2 dyp = __fadd2_rn(f2(pa.z, pa.w), f2(-pyf));
                const float2 qd = __ffma2_rn(dxp, __fmul2_rn(dxp, f2(pb.x, pb.y)), __fmul2_rn(dyp, __fmul2_rn(dyp, f2(pc.x, pc.y))));
                const float2 pwp = __ffma2_rn(qd, f2(-0.5f), __fmul2_rn(dyp, __fmul2_rn(dxp, f2(pb.z, pb.w))));   // pb.zw = -conic.y
                const float2 Gp = expf_pair(pwp);
                const float2 agp = __fmul2_rn(f2(pc.z, pc.w), Gp);
                const float pw0 = pwp.x, pw1 = pwp.y, G0 = Gp.x, G1 = Gp.y;
                const float alpha0 = fminf(0.99f, agp.x);
                const float alpha1 = fminf(0.99f, agp.y);
                const float rdepth0 = rdp.x, rdepth1 = rdp.y;
                const bool take0 = ((pos0 - (uint32_t)jl0) < last_contributor) && !(pw0 > 0.0f) && !(alpha0 < 1.0f / 255.0f);
                const bool take1 = has1 && ((pos0 - (uint32_t)jl1) < last_contributor) && !(pw1 > 0.0f) &&
                                   !(alpha1 < 1.0f / 255.0f);
                const bool any0 = __any_sync(0xffffffffu, take0);
                const bool any1 = __any_sync(0xffffffffu, take1);
                if (!(any0 || any1)) continue;

                // <features, dL/dpixel> of both splats
                float2 da = make_float2(0.0f, 0.0f), db = da, dc = da, dd = da;
#pragma unroll
                for (int q = 0; q < NUM_CHANNELS / 4; q++) {
                    const float4 f = featp[q];
                    const float4 h = featp[NUM_CHANNELS / 4 + q];
                    da = __ffma2_rn(make_float2(f.x, f.y), g[2 * q + 0], da);
                    db = __ffma2_rn(make_float2(f.z, f.w), g[2 * q + 1], db);
                    dc = __ffma2_rn(make_float2(h.x, h.y), g[2 * q + 0], dc);
                    dd = __ffma2_rn(make_float2(h.z, h.w), g[2 * q + 1], dd);
                }
                const float2 ds0 = __fadd2_rn(da, db), ds1 = __fadd2_rn(dc, dd);
                const float d0 = INV ? fmaf(rdepth0, g_inv, ds0.x + ds0.y) : ds0.x + ds0.y;
                const float d1 = INV ? fmaf(rdepth1, g_inv, ds1.x + ds1.y) : ds1.x + ds1.y;

                // Per-pixel partials: the blending weight w and moments of s = dL/dG * G about the
                // pixel offset d (the conic / mean2D / opacity gradients are linear in them;
                // backward.cu:612-635 is finished per Gaussian by the backward preprocess kernel).
                // In list order: first splat ...
                float w0 = 0.0f, mo0 = 0.0f, w1 = 0.0f, mo1 = 0.0f;
                const float2 om2 = __ffma2_rn(f2(alpha0, alpha1), f2(-1.0f), f2(1.0f));   // 1 - alpha of both
                {
                    const float inv_om = __fdividef(1.0f, om2.x);
                    const float Tn = T * inv_om;
                    const float An = fmaf(last_alpha, last_d - A, A);   // last_alpha*last_d + (1-last_alpha)*A
                    const float dLda = fmaf(d0 - An, Tn, -(T_final * inv_om) * bgdot);
                    if (take0) {
                        T = Tn; A = An; last_d = d0; last_alpha = alpha0;
                        w0 = alpha0 * Tn;
                        mo0 = G0 * dLda;             // -> dL/dopacity
                    }
                }
                // ... then the second
                {
                    const float inv_om = __fdividef(1.0f, om2.y);
                    const float Tn = T * inv_om;
                    const float An = fmaf(last_alpha, last_d - A, A);
                    const float dLda = fmaf(d1 - An, Tn, -(T_final * inv_om) * bgdot);
                    if (take1) {
                        T = Tn; A = An; last_d = d1; last_alpha = alpha1;
                        w1 = alpha1 * Tn;
                        mo1 = G1 * dLda;
                    }
                }
                // the moments of both splats as packed products (dxp / dyp hold the offsets of both)
                const float2 s2 = __fmul2_rn(f2(pc.z, pc.w), f2(mo0, mo1));
                const float2 sx2 = __fmul2_rn(s2, dxp), sy2 = __fmul2_rn(s2, dyp);   // s*dx, s*dy
                const float2 sxx2 = __fmul2_rn(sx2, dxp), sxy2 = __fmul2_rn(sx2, dyp), syy2 = __fmul2_rn(sy2, dyp);
                const float2 wg2 = INV ? __fmul2_rn(f2(w0, w1), f2(g_inv)) : make_float2(0.0f, 0.0f);
                float* exb0 = exbase + (parity & 3u) * 32;
                float* exb1 = exbase + ((parity + 1u) & 3u) * 32;
                parity += 2u;
                // Only the blending weights go through shared memory (every lane needs all 32 of them for the
                // colour-gradient mat-vec); the seven moments are reduced in registers by a transposing butterfly:
                // 14 shuffles per pair (18 before the last two steps were left to the L2) instead of 14 STS + 4 non-broadcast LDS.128 (16 wavefronts) + 4 shuffles on the
                // shared-memory data pipe, the busiest unit of this kernel (+1.7 % step throughput, profiles/r01_v42).
                exb0[lane] = w0;
                exb1[lane] = w1;
                __syncwarp();
                float mv0[8] = {sx2.x, sy2.x, sxx2.x, sxy2.x, syy2.x, mo0, wg2.x, 0.0f};
                float mv1[8] = {sx2.y, sy2.y, sxx2.y, sxy2.y, syy2.y, mo1, wg2.y, 0.0f};
                warp_transpose_reduce8x2(mv0, mv1, lane);
                const float mom0 = mv0[0], mom1 = mv1[0];   // lanes 4k..4k+3: partial sums of moment k
                // (a) lane c accumulates channel c of dL/dcolor: 32 weights x its column of g
                const float4* __restrict__ wa = reinterpret_cast<const float4*>(exb0);
                const float4* __restrict__ wb = reinterpret_cast<const float4*>(exb1);
                float2 ca = make_float2(0.0f, 0.0f), cb = ca, cc = ca, cd = ca;
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    const float4 u = wa[q];
                    const float4 x = wb[q];
                    ca = __ffma2_rn(make_float2(u.x, u.y), gT[2 * q + 0], ca);
                    cb = __ffma2_rn(make_float2(u.z, u.w), gT[2 * q + 1], cb);
                    cc = __ffma2_rn(make_float2(x.x, x.y), gT[2 * q + 0], cc);
                    cd = __ffma2_rn(make_float2(x.z, x.w), gT[2 * q + 1], cd);
                }
                const float2 cs0 = __fadd2_rn(ca, cb), cs1 = __fadd2_rn(cc, cd);
                const float col0 = cs0.x + cs0.y;
                const float col1 = cs1.x + cs1.y;

                const uint32_t gi0 = __shfl_sync(0xffffffffu, id_cur, jl0);
                const uint32_t gi1 = __shfl_sync(0xffffffffu, id_cur, jl1);
                // Lanes 0..27 add their partial of moment lane/4 to the packed record (word 7 is unused).
                const bool geo_lane = lane < (INV ? 28 : 24);   // lanes 4k..4k+3: partials of moment k (7 is unused, 6 = invdepth)
                if (any0) atomicAdd(colbase + (size_t)gi0 * color_rs + lane, col0);
                if (any1) atomicAdd(colbase + (size_t)gi1 * color_rs + lane, col1);
                if (any0 && geo_lane) atomicAdd(geobase + (size_t)gi0 * 8 + (lane >> 2), mom0);
                if (any1 && geo_lane) atomicAdd(geobase + (size_t)gi1 * 8 + (lane >> 2), mom1);
            }
        }
#pragma unroll
        for (int k = 0; k < AHEAD; k++) mask_q[k] = mask_q[k + 1];
#pragma unroll
        for (int k = 0; k < AHEAD + 2; k++) id_q[k] = id_q[k + 1];
        if (++s == NST) s = 0;
        __syncwarp();   // every lane is done reading this stage before a later chunk overwrites it
    }
    if (!TMA) tma::cp_async_wait<0>();
    __syncwarp();   // nothing in flight when the next block reuses the stages
  }
}

bool make_feature_tensor_map(const Inputs& in, CUtensorMap* out);   // render_forward.cu

void launch_render_backward(const Inputs& in, const GeomPtrs& g, const ImagePtrs& img, const BinningPtrs& b,
                            const float* dL_dpix, const float* dL_dinvdepth_pix, float* dL_dcolor,
                            size_t color_row_stride, size_t color_view_stride,
                            int taken_masks, int tma_staging, cudaStream_t s)
{
    const uint32_t crs = (uint32_t)color_row_stride;
    if (in.tiles == 0 || in.V == 0) return;
    CUtensorMap fmap;
    memset(&fmap, 0, sizeof(fmap));
    if (tma_staging && !make_feature_tensor_map(in, &fmap)) tma_staging = 0;
    const long long total = (long long)in.tiles * 8 * in.V;
    static const int minb = [] { const char* e = getenv("GUAVA_BWD_MINB"); return e ? atoi(e) : 12; }();          // tuning knob: 12, 14 or 16
    static const int ctas_per_sm = [] { const char* e = getenv("GUAVA_BWD_CTAS"); return e ? atoi(e) : minb; }();   // tuning knob
    const long long full = (long long)sm_count() * ctas_per_sm;
    const unsigned grid = (unsigned)(total < full ? total : full);
    // the counter was zeroed by the forward's tile scan; preprocess_backward_kernel, which follows, re-arms it for a repeated backward
    static const int stages = [] { const char* e = getenv("GUAVA_BWD_STAGES"); return e ? atoi(e) : 2; }();   // tuning knob
#define GUAVA_LAUNCH_BWD2(MK, TM, IV, NS)                                                                                 \
    render_backward_kernel<MK, TM, IV, NS><<<grid, 32, 0, s>>>(in, fmap, g.rec, img.ranges, img.tile_order, b.point_list,  \
                                                               b.taken_marks, img.final_T, img.n_contrib, dL_dpix,        \
                                                               dL_dinvdepth_pix, dL_dcolor, crs, color_view_stride, g.grad_geo, img.num_rendered)
#define GUAVA_LAUNCH_BWD(MK, TM) do { if (dL_dinvdepth_pix != nullptr) GUAVA_LAUNCH_BWD2(MK, TM, 1, 2); else GUAVA_LAUNCH_BWD2(MK, TM, 0, 2); } while (0)
#define GUAVA_LAUNCH_BWD_MINB(MB)                                                                                          \
    do {                                                                                                                  \
        if (dL_dinvdepth_pix != nullptr)                                                                                  \
            render_backward_kernel<1, 0, 1, 2, MB><<<grid, 32, 0, s>>>(in, fmap, g.rec, img.ranges, img.tile_order,       \
                b.point_list, b.taken_marks, img.final_T, img.n_contrib, dL_dpix, dL_dinvdepth_pix, dL_dcolor, crs, color_view_stride, \
                g.grad_geo, img.num_rendered);                                                                                        \
        else                                                                                                              \
            render_backward_kernel<1, 0, 0, 2, MB><<<grid, 32, 0, s>>>(in, fmap, g.rec, img.ranges, img.tile_order,       \
                b.point_list, b.taken_marks, img.final_T, img.n_contrib, dL_dpix, dL_dinvdepth_pix, dL_dcolor, crs, color_view_stride, \
                g.grad_geo, img.num_rendered);                                                                                        \
    } while (0)
    if (taken_masks && !tma_staging && stages == 2 && minb == 14) GUAVA_LAUNCH_BWD_MINB(14);
    else if (taken_masks && !tma_staging && stages == 2 && minb == 16) GUAVA_LAUNCH_BWD_MINB(16);
    else if (taken_masks && !tma_staging && stages == 3) {   // the deeper pipeline exists for the default variant only
        if (dL_dinvdepth_pix != nullptr) GUAVA_LAUNCH_BWD2(1, 0, 1, 3); else GUAVA_LAUNCH_BWD2(1, 0, 0, 3);
    } else if (taken_masks) { if (tma_staging) GUAVA_LAUNCH_BWD(1, 1); else GUAVA_LAUNCH_BWD(1, 0); }
    else             { if (tma_staging) GUAVA_LAUNCH_BWD(0, 1); else GUAVA_LAUNCH_BWD(0, 0); }
#undef GUAVA_LAUNCH_BWD
#undef GUAVA_LAUNCH_BWD2
#undef GUAVA_LAUNCH_BWD_MINB
    count_launch();
}

}  // namespace guava
