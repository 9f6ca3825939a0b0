// Forward compositing: one CTA per 16x16 tile, front-to-back alpha blending of
// 32 feature channels + inverse depth.
//
// Reference behaviour: DGR32/cuda_rasterizer/forward.cu:274-397 (renderCUDA).
// Bit-exact with it: the per-pair alpha / transmittance chain (and therefore
// final_T and n_contrib).  Colour accumulation uses one FFMA per channel with
// w = alpha*T instead of the reference's FMUL+FFMA pair (same value to 1 ulp per
// term; gate is 1e-5 absolute).
//
// B200 design:
//  * instances are staged into shared memory by TMA bulk copies
//    (cp.async.bulk global->shared, mbarrier transaction counting): per instance
//    one 32-byte SplatRec and the 128-byte feature row, double buffered so the
//    next batch streams in while the current one is blended.  The reference
//    re-reads the 32 features from global memory for every (pixel, Gaussian)
//    pair (forward.cu:371-372).
//  * warps own 8x4 pixel blocks; a warp-ballot skips the 32 FFMAs + 8 LDS.128
//    when no lane of the warp takes the Gaussian, and warps / the CTA leave the
//    list as soon as all their pixels are saturated.
#include "internal.cuh"
#include "tma.cuh"

namespace guava {

constexpr int FWD_NB = 128;   // instances per stage
constexpr uint32_t FWD_BYTES_PER_INST = sizeof(SplatRec) + NUM_CHANNELS * sizeof(float);

struct __align__(128) FwdStage {
    float feat[FWD_NB][NUM_CHANNELS];   // 16 KB
    SplatRec rec[FWD_NB];               //  4 KB
};

__global__ void __launch_bounds__(TILE_PIXELS)
render_forward_kernel(const Inputs in, const SplatRec* __restrict__ recs,
                      const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                      const uint32_t* __restrict__ num_rendered,
                      float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                      float* __restrict__ out_color, float* __restrict__ out_invdepth)
{
    __shared__ FwdStage stage[2];
    __shared__ __align__(8) uint64_t full_bar[2];

    if (num_rendered[1]) return;   // binning overflowed; the host re-runs
    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int v = blockIdx.y;
    const int tile = blockIdx.x;
    const int tile_x = tile % in.gx, tile_y = tile / in.gx;
    const uint32_t px = tile_x * TILE_X + (warp & 1) * 8 + (lane & 7);
    const uint32_t py = tile_y * TILE_Y + (warp >> 1) * 4 + (lane >> 3);
    const bool inside = px < (uint32_t)in.W && py < (uint32_t)in.H;
    const float pxf = (float)px, pyf = (float)py;
    // this warp's 8x4 pixel block (pixel centres wx0..wx0+7, wy0..wy0+3)
    const float wx0 = (float)(tile_x * TILE_X + (warp & 1) * 8), wy0 = (float)(tile_y * TILE_Y + (warp >> 1) * 4);

    const uint2 range = ranges[(size_t)v * in.tiles + tile];
    const int len = (int)(range.y - range.x);
    const int nbatches = (len + FWD_NB - 1) / FWD_NB;

    if (tid == 0) {
        tma::mbar_init(&full_bar[0], 1);
        tma::mbar_init(&full_bar[1], 1);
        tma::fence_mbar_init();
    }
    __syncthreads();

    const SplatRec* __restrict__ vrecs = recs + (size_t)v * in.P;
    const float* __restrict__ vfeat = in.colors + (size_t)v * in.colors_vs;

    auto issue = [&](int b) {
        const int s = b & 1;
        const int nb = min(FWD_NB, len - b * FWD_NB);
        if (tid == 0) tma::mbar_arrive_expect_tx(&full_bar[s], (uint32_t)nb * FWD_BYTES_PER_INST);
        if (tid < nb) {
            const uint32_t id = point_list[range.x + b * FWD_NB + tid];
            tma::bulk_g2s(&stage[s].rec[tid], vrecs + id, sizeof(SplatRec), &full_bar[s]);
            tma::bulk_g2s(&stage[s].feat[tid][0], vfeat + (size_t)id * NUM_CHANNELS,
                          NUM_CHANNELS * sizeof(float), &full_bar[s]);
        }
    };

    float T = 1.0f;
    float C[NUM_CHANNELS];
#pragma unroll
    for (int ch = 0; ch < NUM_CHANNELS; ch++) C[ch] = 0.0f;
    float D = 0.0f;
    uint32_t last_contributor = 0;
    bool done = !inside;
    bool warp_done = __all_sync(0xffffffffu, done);

    if (nbatches > 0) issue(0);
    for (int b = 0; b < nbatches; b++) {
        if (b + 1 < nbatches) issue(b + 1);
        const int s = b & 1;
        tma::mbar_wait(&full_bar[s], (uint32_t)((b >> 1) & 1));
        const int nb = min(FWD_NB, len - b * FWD_NB);
        if (!warp_done) {
            const float4* __restrict__ rec4 = reinterpret_cast<const float4*>(&stage[s].rec[0]);
            // Warp-footprint culling: lane L tests staged splat 32*r+L against this warp's
            // 8x4 pixel block; only the survivors are walked (in list order) below.
            uint32_t live[FWD_NB / 32];
#pragma unroll
            for (int r = 0; r < FWD_NB / 32; r++) {
                const int j = 32 * r + lane;
                bool keep = false;
                if (j < nb) {
                    const float4 c0 = rec4[2 * j], c1 = rec4[2 * j + 1];
                    keep = splat_may_touch(c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, wx0, wx0 + 7.0f, wy0, wy0 + 3.0f);
                }
                live[r] = __ballot_sync(0xffffffffu, keep);
            }
#pragma unroll
            for (int r = 0; r < FWD_NB / 32; r++) {
                uint32_t m = live[r];
                while (m != 0u && !warp_done) {
                    const int j = 32 * r + __ffs(m) - 1;
                    m &= m - 1;
                    const float4 r0 = rec4[2 * j];      // x, y, cx, cy
                    const float4 r1 = rec4[2 * j + 1];  // cz, op, rdepth, aux
                    // SASS order of forward.cu:350-353
                    const float dx = fsub(r0.x, pxf), dy = fsub(r0.y, pyf);
                    const float t1 = fmul(dy, fmul(dy, r1.x));
                    const float sq = ffma(dx, fmul(dx, r0.z), t1);
                    const float power = ffma(sq, -0.5f, -fmul(dy, fmul(dx, r0.w)));
                    bool take = !done && !(power > 0.0f);
                    const float alpha = fminf(fmul(r1.y, expf(power)), 0.99f);
                    take = take && !(alpha < 1.0f / 255.0f);
                    const float test_T = fmul(T, fsub(1.0f, alpha));
                    if (take && test_T < 0.0001f) { done = true; take = false; }
                    if (__any_sync(0xffffffffu, take)) {
                        const float w = take ? fmul(alpha, T) : 0.0f;
                        const float4* __restrict__ f4 = reinterpret_cast<const float4*>(&stage[s].feat[j][0]);
#pragma unroll
                        for (int q = 0; q < NUM_CHANNELS / 4; q++) {
                            const float4 f = f4[q];
                            C[4 * q + 0] = fmaf(w, f.x, C[4 * q + 0]);
                            C[4 * q + 1] = fmaf(w, f.y, C[4 * q + 1]);
                            C[4 * q + 2] = fmaf(w, f.z, C[4 * q + 2]);
                            C[4 * q + 3] = fmaf(w, f.w, C[4 * q + 3]);
                        }
                        D = fmaf(w, r1.z, D);
                        if (take) { T = test_T; last_contributor = (uint32_t)(b * FWD_NB + j + 1); }
                    }
                    if (__all_sync(0xffffffffu, done)) warp_done = true;
                }
            }
        }
        // also orders this batch's shared-memory reads before the refill of stage s
        const int block_done = __syncthreads_and(warp_done ? 1 : 0);
        if (block_done) {
            if (b + 1 < nbatches)   // never exit with a bulk copy still in flight
                tma::mbar_wait(&full_bar[(b + 1) & 1], (uint32_t)(((b + 1) >> 1) & 1));
            break;
        }
    }

    if (inside) {
        const size_t HW = (size_t)in.H * in.W;
        const size_t pix = (size_t)py * in.W + px;
        final_T[(size_t)v * HW + pix] = T;
        n_contrib[(size_t)v * HW + pix] = last_contributor;
        const float* __restrict__ bg = in.background + (size_t)v * NUM_CHANNELS;
        float* __restrict__ oc = out_color + (size_t)v * NUM_CHANNELS * HW + pix;
#pragma unroll
        for (int ch = 0; ch < NUM_CHANNELS; ch++) oc[(size_t)ch * HW] = fmaf(T, bg[ch], C[ch]);
        if (out_invdepth != nullptr) out_invdepth[(size_t)v * HW + pix] = D;
    }
}

void launch_render_forward(const Inputs& in, const GeomPtrs& g, const ImagePtrs& img, const BinningPtrs& b,
                           float* out_color, float* out_invdepth, cudaStream_t s)
{
    if (in.tiles == 0 || in.V == 0) return;
    dim3 grid(in.tiles, in.V);
    render_forward_kernel<<<grid, TILE_PIXELS, 0, s>>>(in, g.rec, img.ranges, b.point_list, img.num_rendered,
                                                       img.final_T, img.n_contrib, out_color, out_invdepth);
    count_launch();
}

}  // namespace guava
