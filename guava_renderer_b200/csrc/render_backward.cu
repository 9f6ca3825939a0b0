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
//  * the reference's per-channel recurrence
//        accum_rec[ch] = last_alpha*last_color[ch] + (1-last_alpha)*accum_rec[ch]
//        dL_dalpha += (c[ch] - accum_rec[ch]) * dL_dpixel[ch]
//    is linear in the channel index, so only its contraction with dL_dpixel is
//    carried:  A = last_alpha*last_d + (1-last_alpha)*A,  d = <c, dL_dpixel>.
//    That removes 64 live registers per thread;
//  * per-Gaussian gradients are reduced across the 32 pixels of a warp with a
//    transposing butterfly (31 shuffles put channel `lane` of the warp total in
//    lane `lane`) and leave the SM as ONE coalesced 128-byte red.global per
//    (warp, Gaussian) plus 7 scalar reds, instead of 39 per pixel.
#include "internal.cuh"
#include "tma.cuh"

namespace guava {

constexpr int BWD_NB = 128;
constexpr uint32_t BWD_BYTES_PER_INST = sizeof(SplatRec) + NUM_CHANNELS * sizeof(float);

struct __align__(128) BwdStage {
    float feat[BWD_NB][NUM_CHANNELS];
    SplatRec rec[BWD_NB];
    uint32_t id[BWD_NB];
};

// After the call lane l holds in v[0] the warp-wide sum of element l (N == 32),
// or of element (l >> (5 - log2 N)) for smaller N once `finish` has run.
template <int N>
__device__ __forceinline__ void warp_transpose_reduce(float (&v)[N], int lane)
{
    int mask = 16;
#pragma unroll
    for (int half = N / 2; half >= 1; half >>= 1, mask >>= 1) {
        const bool upper = (lane & mask) != 0;
#pragma unroll
        for (int i = 0; i < half; i++) {
            const float send = upper ? v[i] : v[i + half];
            const float keep = upper ? v[i + half] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, mask);
        }
    }
}

__global__ void __launch_bounds__(TILE_PIXELS)
render_backward_kernel(const Inputs in, const SplatRec* __restrict__ recs,
                       const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                       const float* __restrict__ final_Ts, const uint32_t* __restrict__ n_contrib,
                       const float* __restrict__ dL_dpix, const float* __restrict__ dL_dinvdepth_pix,
                       float* __restrict__ dL_dmean2D, float* __restrict__ dL_dconic,
                       float* __restrict__ dL_dopacity, float* __restrict__ dL_dcolor,
                       float* __restrict__ dL_dinvdepth)
{
    __shared__ BwdStage stage[2];
    __shared__ __align__(8) uint64_t full_bar[2];
    __shared__ uint32_t warp_max_s[8];

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int v = blockIdx.y;
    const int tile = blockIdx.x;
    const int tile_x = tile % in.gx, tile_y = tile / in.gx;
    const uint32_t px = tile_x * TILE_X + (warp & 1) * 8 + (lane & 7);
    const uint32_t py = tile_y * TILE_Y + (warp >> 1) * 4 + (lane >> 3);
    const bool inside = px < (uint32_t)in.W && py < (uint32_t)in.H;
    const float pxf = (float)px, pyf = (float)py;
    const size_t HW = (size_t)in.H * in.W;
    const size_t pix = (size_t)py * in.W + px;

    const uint2 range = ranges[(size_t)v * in.tiles + tile];

    const uint32_t last_contributor = inside ? n_contrib[(size_t)v * HW + pix] : 0u;
    const uint32_t warp_max = __reduce_max_sync(0xffffffffu, last_contributor);
    if (lane == 0) warp_max_s[warp] = warp_max;
    if (tid == 0) {
        tma::mbar_init(&full_bar[0], 1);
        tma::mbar_init(&full_bar[1], 1);
        tma::fence_mbar_init();
    }
    __syncthreads();
    uint32_t total = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) total = max(total, warp_max_s[w]);
    const int nbatches = (int)((total + BWD_NB - 1) / BWD_NB);
    if (nbatches == 0) return;

    const SplatRec* __restrict__ vrecs = recs + (size_t)v * in.P;
    const float* __restrict__ vfeat = in.colors + (size_t)v * in.colors_vs;

    // batch b, slot j  <->  list position  total-1-(b*NB+j)
    auto issue = [&](int b) {
        const int s = b & 1;
        const int nb = min(BWD_NB, (int)total - b * BWD_NB);
        if (tid == 0) tma::mbar_arrive_expect_tx(&full_bar[s], (uint32_t)nb * BWD_BYTES_PER_INST);
        if (tid < nb) {
            const uint32_t id = point_list[range.x + (total - 1 - (uint32_t)(b * BWD_NB + tid))];
            stage[s].id[tid] = id;
            tma::bulk_g2s(&stage[s].rec[tid], vrecs + id, sizeof(SplatRec), &full_bar[s]);
            tma::bulk_g2s(&stage[s].feat[tid][0], vfeat + (size_t)id * NUM_CHANNELS,
                          NUM_CHANNELS * sizeof(float), &full_bar[s]);
        }
    };
    issue(0);

    float g[NUM_CHANNELS];
    float g_inv = 0.0f, bgdot = 0.0f;
    const float T_final = inside ? final_Ts[(size_t)v * HW + pix] : 0.0f;
    {
        const float* __restrict__ bg = in.background + (size_t)v * NUM_CHANNELS;
        const float* __restrict__ gp = dL_dpix + (size_t)v * NUM_CHANNELS * HW + pix;
#pragma unroll
        for (int ch = 0; ch < NUM_CHANNELS; ch++) {
            g[ch] = inside ? gp[(size_t)ch * HW] : 0.0f;
            bgdot = fmaf(bg[ch], g[ch], bgdot);
        }
        if (dL_dinvdepth_pix != nullptr && inside) g_inv = dL_dinvdepth_pix[(size_t)v * HW + pix];
    }
    const float ddelx_dx = 0.5f * in.W, ddely_dy = 0.5f * in.H;

    float T = T_final;
    float A = 0.0f, last_alpha = 0.0f, last_d = 0.0f;
    const size_t gbase = (size_t)v * in.P;

    for (int b = 0; b < nbatches; b++) {
        if (b + 1 < nbatches) issue(b + 1);
        const int s = b & 1;
        tma::mbar_wait(&full_bar[s], (uint32_t)((b >> 1) & 1));
        const int nb = min(BWD_NB, (int)total - b * BWD_NB);
        const float4* __restrict__ rec4 = reinterpret_cast<const float4*>(&stage[s].rec[0]);
        for (int j = 0; j < nb; j++) {
            const uint32_t pos = total - 1 - (uint32_t)(b * BWD_NB + j);
            if (pos >= warp_max) continue;                      // warp-uniform
            const float4 r0 = rec4[2 * j];      // x, y, cx, cy
            const float4 r1 = rec4[2 * j + 1];  // cz, op, rdepth, aux
            const float dx = fsub(r0.x, pxf), dy = fsub(r0.y, pyf);
            const float t1 = fmul(dy, fmul(dy, r1.x));
            const float sq = ffma(dx, fmul(dx, r0.z), t1);
            const float power = ffma(sq, -0.5f, -fmul(dy, fmul(dx, r0.w)));
            const float G = expf(power);
            const float alpha = fminf(0.99f, fmul(r1.y, G));
            const bool take = (pos < last_contributor) && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
            if (!__any_sync(0xffffffffu, take)) continue;

            float vals[NUM_CHANNELS];
            float geo[8];
            if (take) {
                const float om = 1.0f - alpha;
                T = T / om;
                const float w = alpha * T;
                const float4* __restrict__ f4 = reinterpret_cast<const float4*>(&stage[s].feat[j][0]);
                float d = 0.0f;
#pragma unroll
                for (int q = 0; q < NUM_CHANNELS / 4; q++) {
                    const float4 f = f4[q];
                    d = fmaf(f.x, g[4 * q + 0], d);
                    d = fmaf(f.y, g[4 * q + 1], d);
                    d = fmaf(f.z, g[4 * q + 2], d);
                    d = fmaf(f.w, g[4 * q + 3], d);
                    vals[4 * q + 0] = w * g[4 * q + 0];
                    vals[4 * q + 1] = w * g[4 * q + 1];
                    vals[4 * q + 2] = w * g[4 * q + 2];
                    vals[4 * q + 3] = w * g[4 * q + 3];
                }
                d = fmaf(r1.z, g_inv, d);
                A = last_alpha * last_d + (1.0f - last_alpha) * A;
                last_d = d;
                float dL_dalpha = (d - A) * T;
                last_alpha = alpha;
                dL_dalpha += (-T_final / om) * bgdot;
                const float dL_dG = r1.y * dL_dalpha;
                const float gdx = G * dx, gdy = G * dy;
                const float dG_ddelx = -gdx * r0.z - gdy * r0.w;
                const float dG_ddely = -gdy * r1.x - gdx * r0.w;
                geo[0] = dL_dG * dG_ddelx * ddelx_dx;
                geo[1] = dL_dG * dG_ddely * ddely_dy;
                geo[2] = -0.5f * gdx * dx * dL_dG;
                geo[3] = -0.5f * gdx * dy * dL_dG;
                geo[4] = -0.5f * gdy * dy * dL_dG;
                geo[5] = G * dL_dalpha;
                geo[6] = w * g_inv;
                geo[7] = 0.0f;
            } else {
#pragma unroll
                for (int ch = 0; ch < NUM_CHANNELS; ch++) vals[ch] = 0.0f;
#pragma unroll
                for (int k = 0; k < 8; k++) geo[k] = 0.0f;
            }
            warp_transpose_reduce<NUM_CHANNELS>(vals, lane);
            warp_transpose_reduce<8>(geo, lane);
            geo[0] += __shfl_xor_sync(0xffffffffu, geo[0], 2);
            geo[0] += __shfl_xor_sync(0xffffffffu, geo[0], 1);

            const size_t gi = gbase + stage[s].id[j];
            atomicAdd(dL_dcolor + gi * NUM_CHANNELS + lane, vals[0]);
            if ((lane & 3) == 0) {
                const int k = lane >> 2;
                float* dst = nullptr;
                if (k < 2) dst = dL_dmean2D + gi * 3 + k;
                else if (k == 2) dst = dL_dconic + gi * 4 + 0;
                else if (k == 3) dst = dL_dconic + gi * 4 + 1;
                else if (k == 4) dst = dL_dconic + gi * 4 + 3;
                else if (k == 5) dst = dL_dopacity + gi;
                else if (k == 6 && dL_dinvdepth != nullptr) dst = dL_dinvdepth + gi;
                if (dst != nullptr) atomicAdd(dst, geo[0]);
            }
        }
        __syncthreads();   // all warps are done with stage s before it is refilled
    }
}

void launch_render_backward(const Inputs& in, const GeomPtrs& g, const ImagePtrs& img, const BinningPtrs& b,
                            const float* dL_dpix, const float* dL_dinvdepth_pix,
                            float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                            float* dL_dcolor, float* dL_dinvdepth, cudaStream_t s)
{
    if (in.tiles == 0 || in.V == 0) return;
    dim3 grid(in.tiles, in.V);
    render_backward_kernel<<<grid, TILE_PIXELS, 0, s>>>(in, g.rec, img.ranges, b.point_list, img.final_T,
                                                        img.n_contrib, dL_dpix, dL_dinvdepth_pix, dL_dmean2D,
                                                        dL_dconic, dL_dopacity, dL_dcolor, dL_dinvdepth);
    count_launch();
}

}  // namespace guava
