// Backward compositing: one CTA per 16x16 tile walks its depth-sorted list back
// to front and produces the per-Gaussian screen-space gradients.
//
// Reference behaviour: DGR32/cuda_rasterizer/backward.cu:452-638 (renderCUDA),
// which issues 39 global atomicAdd per contributing (pixel, Gaussian) pair
// (:593,603,626-635) and keeps accum_rec[32] + last_color[32] + dL_dpixel[32]
// live per thread (173 registers).
//
// B200 design:
//  * same TMA-staged, double-buffered instance batches as the forward kernel
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
//  * the seven geometric partials (mean2D.xy, conic.xyw, opacity, invdepth) are
//    reduced with a transposing butterfly and leave the SM as one 32-byte
//    reduction into a packed per-Gaussian record (GeomPtrs::grad_geo) that the
//    backward preprocess kernel consumes.
#include "internal.cuh"
#include "tma.cuh"
#include <cstdlib>

namespace guava {

struct __align__(128) BwdStage {
    float feat[32][NUM_CHANNELS];   // 4 KB
    SplatRec rec[32];               // 1 KB
};
static_assert(sizeof(BwdStage) >= 32 * 32 * sizeof(float), "a stage doubles as the transpose scratch");
// row stride of the exchange buffer: 32 pixels + 4 floats of padding, so that the 8-lane phases of
// the row-sum LDS.128 (lanes 4k..4k+3 read row k) fall on distinct banks
constexpr int EX_STRIDE = 36;

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d)
{
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}

// One warp (= one CTA) per 8x4 pixel block, as in the forward kernel: no block
// barrier, private double-buffered stage of culled survivors.
// VEC_RED: 1 = 16-byte vector reductions (red.global.add.v4.f32, 8 + 2 lanes per
// (warp, Gaussian)); 0 = 4-byte reductions (32 + 8 lanes), for unaligned outputs.
template <int VEC_RED>
__global__ void __launch_bounds__(32, 16)
render_backward_kernel(const Inputs in, const SplatRec* __restrict__ recs,
                       const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order,
                       const uint32_t* __restrict__ point_list, const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                       const float* __restrict__ dL_dpix, const float* __restrict__ dL_dinvdepth_pix,
                       float* __restrict__ dL_dcolor, float* __restrict__ grad_geo, uint32_t* __restrict__ work_counter)
{
    __shared__ BwdStage stage[2];
    __shared__ __align__(8) uint64_t full_bar[2];
    __shared__ __align__(16) float ex[2][8][EX_STRIDE];   // per-survivor exchange, double buffered

    const int lane = threadIdx.x;
    if (lane == 0) {
        tma::mbar_init(&full_bar[0], 1);
        tma::mbar_init(&full_bar[1], 1);
        tma::fence_mbar_init();
    }
    __syncwarp();
    const uint32_t lt_mask = (1u << lane) - 1u;
    const size_t HW = (size_t)in.H * in.W;
    uint32_t parity = 0;   // which half of ex the next exchange uses
    uint32_t phase = 0;    // bit s = parity to wait for on full_bar[s]

    // persistent warps pulling 8x4 pixel blocks from a work counter (see render_forward.cu)
    const uint32_t blocks_per_view = (uint32_t)in.tiles * 8u;
    const uint32_t total_work = blocks_per_view * (uint32_t)in.V;
    auto fetch_work = [&]() -> uint32_t {
        uint32_t wk = 0;
        if (lane == 0) wk = atomicAdd(work_counter, 1u);
        return __shfl_sync(0xffffffffu, wk, 0);
    };

  uint32_t work = fetch_work();
  while (work < total_work) {
    const uint32_t work_next = fetch_work();
    const uint32_t gtile = tile_order[work >> 3];   // longest lists first
    const int sub = (int)(work & 7u);
    work = work_next;
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
    const uint2 range = ranges[(size_t)v * in.tiles + tile];

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
        if (dL_dinvdepth_pix != nullptr && inside) g_inv = dL_dinvdepth_pix[(size_t)v * HW + pix];
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
    // generic-proxy accesses to the scratch are ordered before the async-proxy (TMA) writes
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();

    const SplatRec* __restrict__ vrecs = recs + (size_t)v * in.P;
    const float* __restrict__ vfeat = in.colors + (size_t)v * in.colors_vs;

    constexpr uint32_t NO_ID = 0xffffffffu;
    // chunk c, lane L  <->  list position  total-1-(32c+L)   (back to front); same software
    // pipeline as the forward kernel (ids two chunks ahead, SplatRecs one chunk ahead).
    auto load_id = [&](int c) -> uint32_t {
        const int k = (c << 5) + lane;
        return k < total ? point_list[range.x + (uint32_t)(total - 1 - k)] : NO_ID;
    };
    auto load_rec = [&](uint32_t id, float4& c0, float4& c1) {
        if (id != NO_ID) {
            const float4* __restrict__ rp = reinterpret_cast<const float4*>(vrecs + id);
            c0 = rp[0];
            c1 = rp[1];
        }
    };
    auto stage_chunk = [&](int c, uint32_t id, const float4& c0, const float4& c1) -> uint32_t {
        const bool keep = id != NO_ID &&
                          splat_may_touch(c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, wx0, wx0 + 7.0f, wy0, wy0 + 3.0f);
        const uint32_t mask = __ballot_sync(0xffffffffu, keep);
        if (mask != 0u) {
            BwdStage& st = stage[c & 1];
            uint64_t* bar = &full_bar[c & 1];
            if (lane == 0) tma::mbar_arrive_expect_tx(bar, (uint32_t)__popc(mask) * (NUM_CHANNELS * 4u));
            __syncwarp();
            if (keep) {
                const int slot = __popc(mask & lt_mask);
                float4* dst = reinterpret_cast<float4*>(&st.rec[slot]);
                dst[0] = c0;
                dst[1] = c1;
                tma::bulk_g2s(&st.feat[slot][0], vfeat + (size_t)id * NUM_CHANNELS, NUM_CHANNELS * sizeof(float), bar);
            }
        }
        return mask;
    };

    float T = T_final;
    float A = 0.0f, last_alpha = 0.0f, last_d = 0.0f;
    const size_t gbase = (size_t)v * in.P;

    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
    uint32_t id_cur = load_id(0);
    uint32_t id_next = load_id(1);
    load_rec(id_cur, a0, a1);
    uint32_t mask_cur = stage_chunk(0, id_cur, a0, a1);
    load_rec(id_next, a0, a1);
    uint32_t id_nn = load_id(2);
    for (int c = 0; c < nchunks; c++) {
        // id_cur: ids of chunk c (for the gradient addresses); id_next / a0,a1: chunk c+1
        const uint32_t mask_next = (c + 1 < nchunks) ? stage_chunk(c + 1, id_next, a0, a1) : 0u;
        const uint32_t id_after = id_next;
        id_next = id_nn;
        load_rec(id_next, a0, a1);
        id_nn = load_id(c + 3);
        const int s = c & 1;
        if (mask_cur != 0u) {
            tma::mbar_wait(&full_bar[s], (phase >> s) & 1u);
            phase ^= 1u << s;
            __syncwarp();
            const float4* __restrict__ recp = reinterpret_cast<const float4*>(&stage[s].rec[0]);
            const float4* __restrict__ featp = reinterpret_cast<const float4*>(&stage[s].feat[0][0]);
            const uint32_t pos0 = (uint32_t)(total - 1 - (c << 5));   // list position of chunk lane 0
            uint32_t m = mask_cur;
            for (; m != 0u; recp += 2, featp += NUM_CHANNELS / 4) {
                const int jl = __ffs(m) - 1;
                m &= m - 1;
                const uint32_t pos = pos0 - (uint32_t)jl;
                const float4 r0 = recp[0];   // x, y, cx, cy
                const float4 r1 = recp[1];   // cz, op, rdepth, aux
                const float dx = fsub(r0.x, pxf), dy = fsub(r0.y, pyf);
                const float t1 = fmul(dy, fmul(dy, r1.x));
                const float sq = ffma(dx, fmul(dx, r0.z), t1);
                const float power = ffma(sq, -0.5f, -fmul(dy, fmul(dx, r0.w)));
                const float G = expf(power);
                const float alpha = fminf(0.99f, fmul(r1.y, G));
                const bool take = (pos < last_contributor) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                if (!__any_sync(0xffffffffu, take)) continue;

                // Per-pixel partials of this splat: the blending weight w and six moments of
                // s = dL/dG * G about the pixel offset d (the conic / mean2D / opacity gradients
                // are linear in them; backward.cu:612-635 is finished per Gaussian by the
                // backward preprocess kernel).
                float w = 0.0f, m_x = 0.0f, m_y = 0.0f, m_xx = 0.0f, m_xy = 0.0f, m_yy = 0.0f, m_o = 0.0f;
                if (take) {
                    const float inv_om = __fdividef(1.0f, 1.0f - alpha);
                    T *= inv_om;
                    w = alpha * T;
                    float2 d01 = make_float2(0.0f, 0.0f), d23 = d01;
#pragma unroll
                    for (int q = 0; q < NUM_CHANNELS / 4; q++) {
                        const float4 f = featp[q];
                        d01 = __ffma2_rn(make_float2(f.x, f.y), g[2 * q + 0], d01);
                        d23 = __ffma2_rn(make_float2(f.z, f.w), g[2 * q + 1], d23);
                    }
                    const float d = fmaf(r1.z, g_inv, (d01.x + d01.y) + (d23.x + d23.y));
                    A = fmaf(last_alpha, last_d - A, A);   // last_alpha*last_d + (1-last_alpha)*A
                    last_d = d;
                    last_alpha = alpha;
                    const float dL_dalpha = fmaf(d - A, T, -(T_final * inv_om) * bgdot);
                    m_o = G * dL_dalpha;             // -> dL/dopacity
                    const float sg = r1.y * m_o;     // dL/dG * G
                    m_x = sg * dx;
                    m_y = sg * dy;
                    m_xx = m_x * dx;
                    m_xy = m_x * dy;
                    m_yy = m_y * dy;
                }
                // Exchange through shared memory: row k of ex holds partial k of all 32 pixels.
                float* exb = &ex[parity][0][0];
                parity ^= 1u;
                exb[0 * EX_STRIDE + lane] = m_x;
                exb[1 * EX_STRIDE + lane] = m_y;
                exb[2 * EX_STRIDE + lane] = m_xx;
                exb[3 * EX_STRIDE + lane] = m_xy;
                exb[4 * EX_STRIDE + lane] = m_yy;
                exb[5 * EX_STRIDE + lane] = m_o;
                exb[6 * EX_STRIDE + lane] = w * g_inv;
                exb[7 * EX_STRIDE + lane] = w;
                __syncwarp();
                // (a) lane c accumulates channel c of dL/dcolor: 32 weights x its column of g
                const float4* __restrict__ w4 = reinterpret_cast<const float4*>(exb + 7 * EX_STRIDE);
                float2 c01 = make_float2(0.0f, 0.0f), c23 = c01;
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    const float4 ww = w4[q];
                    c01 = __ffma2_rn(make_float2(ww.x, ww.y), gT[2 * q + 0], c01);
                    c23 = __ffma2_rn(make_float2(ww.z, ww.w), gT[2 * q + 1], c23);
                }
                const float col = (c01.x + c01.y) + (c23.x + c23.y);
                // (b) lanes 4k..4k+3 sum row k (8 pixels each), two shuffles finish it
                const float4* __restrict__ e4 = reinterpret_cast<const float4*>(exb + (lane >> 2) * EX_STRIDE + (lane & 3) * 8);
                const float4 ea = e4[0], eb = e4[1];
                float mom = ((ea.x + ea.y) + (ea.z + ea.w)) + ((eb.x + eb.y) + (eb.z + eb.w));
                mom += __shfl_xor_sync(0xffffffffu, mom, 1);
                mom += __shfl_xor_sync(0xffffffffu, mom, 2);

                const size_t gi = gbase + __shfl_sync(0xffffffffu, id_cur, jl);
                if (VEC_RED) {
                    const float a1 = __shfl_down_sync(0xffffffffu, col, 1);
                    const float a2 = __shfl_down_sync(0xffffffffu, col, 2);
                    const float a3 = __shfl_down_sync(0xffffffffu, col, 3);
                    if ((lane & 3) == 0) red_add_v4(dL_dcolor + gi * NUM_CHANNELS + lane, col, a1, a2, a3);
                } else {
                    atomicAdd(dL_dcolor + gi * NUM_CHANNELS + lane, col);
                }
                if ((lane & 3) == 0 && lane < 28) atomicAdd(grad_geo + gi * 8 + (lane >> 2), mom);
            }
        }
        mask_cur = mask_next;
        id_cur = id_after;
        __syncwarp();   // every lane is done reading stage[s] before chunk c+2 overwrites it
    }
  }
}

void launch_render_backward(const Inputs& in, const GeomPtrs& g, const ImagePtrs& img, const BinningPtrs& b,
                            const float* dL_dpix, const float* dL_dinvdepth_pix, float* dL_dcolor, int vec_red,
                            cudaStream_t s)
{
    if (in.tiles == 0 || in.V == 0) return;
    const long long total = (long long)in.tiles * 8 * in.V;
    static const int ctas_per_sm = [] { const char* e = getenv("GUAVA_BWD_CTAS"); return e ? atoi(e) : 16; }();   // tuning knob
    const unsigned grid = (unsigned)(total < 148LL * ctas_per_sm ? total : 148LL * ctas_per_sm);
    // the counter was zeroed by the forward's tile scan and is re-armed below for a repeated backward
    if (vec_red)
        render_backward_kernel<1><<<grid, 32, 0, s>>>(in, g.rec, img.ranges, img.tile_order, b.point_list, img.final_T, img.n_contrib,
                                                      dL_dpix, dL_dinvdepth_pix, dL_dcolor, g.grad_geo,
                                                      img.num_rendered + 3);
    else
        render_backward_kernel<0><<<grid, 32, 0, s>>>(in, g.rec, img.ranges, img.tile_order, b.point_list, img.final_T, img.n_contrib,
                                                      dL_dpix, dL_dinvdepth_pix, dL_dcolor, g.grad_geo,
                                                      img.num_rendered + 3);
    cudaMemsetAsync(img.num_rendered + 3, 0, sizeof(uint32_t), s);
    count_launch();
}

}  // namespace guava
