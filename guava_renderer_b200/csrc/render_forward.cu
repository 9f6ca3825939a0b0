// Forward compositing: front-to-back alpha blending of 32 feature channels +
// inverse depth over the depth-sorted per-tile lists.
//
// Reference behaviour: DGR32/cuda_rasterizer/forward.cu:274-397 (renderCUDA).
// Bit-exact with it: the per-pair alpha / transmittance chain (and therefore
// final_T and n_contrib).  Colour accumulation uses one FFMA per channel with
// w = alpha*T instead of the reference's FMUL+FFMA pair (same value to 1 ulp per
// term; gate is 1e-5 absolute).
//
// B200 design -- one WARP per CTA, eight CTAs per 16x16 tile:
//  * The reference runs one 256-thread CTA per tile and synchronises it twice
//    per 256 staged splats; measured on B200 the first version of this kernel
//    (same CTA shape) spent 36 % of its warp-cycles in `barrier` stalls
//    (profiles/r01_v2_*): the eight warps of a tile see very different numbers
//    of splats.  Here every 8x4 pixel block is its own 32-thread CTA with its
//    own pipeline: no block barrier exists, a block that saturates early frees
//    its SM slot at once, and up to 20+ such CTAs are resident per SM.
//  * Warp-footprint culling (splat_may_touch, internal.cuh): the 32 lanes test
//    32 list entries at a time against the 8x4 block; only survivors are
//    staged and blended.  The tile lists themselves (keys, order, ranges) stay
//    exactly the reference's, so n_contrib keeps its meaning (list position).
//  * Survivors are compacted into a double-buffered shared-memory stage: the
//    record by the lane that tested it (pair-interleaved, see FwdStage), the
//    128-byte feature row by eight 16-byte asynchronous copies of that lane
//    (cp.async / LDGSTS, one commit group per chunk), so the next chunk's rows
//    stream in while the current one is blended.  (First version: one TMA bulk
//    copy per row; see tma.cuh for why that lost.)  The reference re-reads the
//    32 features from global memory for every (pixel, Gaussian) pair
//    (forward.cu:371-372).
#include "internal.cuh"
#include "tma.cuh"
#include <cuda.h>
#include <cuda_bf16.h>
#include <cstdlib>
#include <cstring>

namespace guava {

// Survivor records are staged PAIR-INTERLEAVED: pair k = survivors 2k, 2k+1 of the chunk, so that one
// LDS.128 delivers a quantity of both as the two halves of packed f32x2 operands.
struct __align__(128) FwdStage {
    float feat[32][NUM_CHANNELS];   // 4 KB
    float4 xy[16];                  // x0 x1 y0 y1
    float4 cab[16];                 // conic.x0 conic.x1 -conic.y0 -conic.y1
    float4 cop[16];                 // conic.z0 conic.z1 opacity0 opacity1
    float4 dpos[16];                // rdepth0 rdepth1 list-position0 list-position1 (positions as bits, 1-based)
};

// MASKS = 1 (option "taken_masks"): every list entry this block got nothing from -- rejected by the footprint test, or
// a survivor that no pixel took -- is marked in this block's plane of BinningPtrs::taken_marks (one byte store per lane
// and 32-entry chunk -- 32 consecutive bytes -- no atomics), for the backward compositor to skip.
// TMA = 1 (option "tma_staging"): the survivors' 128-byte feature rows are fetched by the TMA unit -- one elected
// lane issues cp.async.bulk.tensor...tile::gather4 over a [rows, 32] tensor map of colors_precomp, four survivor rows
// per instruction, completion on a per-stage mbarrier -- instead of eight 16-byte LDGSTS per survivor lane.
template <int MASKS, int TMA>
__global__ void __launch_bounds__(32, 16)
render_forward_kernel(const Inputs in, const __grid_constant__ CUtensorMap fmap, const SplatRec* __restrict__ recs,
                      const uint2* __restrict__ ranges, const uint4* __restrict__ tile_order,
                      const uint32_t* __restrict__ point_list, uint8_t* __restrict__ taken_marks, size_t marks_plane,
                      uint32_t* __restrict__ num_rendered,
                      float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                      float* __restrict__ out_color, float* __restrict__ out_invdepth)
{
    __shared__ FwdStage stage[2];
    __shared__ __align__(16) uint32_t srows[2][32];   // TMA: tensor-map row of the survivor in each slot
    __shared__ __align__(8) uint64_t bars[2];         // TMA: one transaction barrier per stage

    if (num_rendered[1]) return;   // binning overflowed; the host re-runs
    const int lane = threadIdx.x;
    const uint32_t lt_mask = (1u << lane) - 1u;
    uint32_t phase = 0;            // TMA: bit s = parity the next wait on bars[s] uses
    if (TMA) {
        if (lane == 0) {
            tma::mbar_init(&bars[0], 1);
            tma::mbar_init(&bars[1], 1);
            tma::fence_mbar_init();
        }
        __syncwarp();
    }

    // Persistent warps: the grid is sized to fill the machine once and every warp pulls
    // 8x4 pixel blocks from a global work counter (num_rendered[2], zeroed by the tile
    // scan), so SM slots never idle between blocks and empty tiles cost one atomic.
    const uint32_t blocks_per_view = (uint32_t)in.tiles * 8u;
    const uint32_t total_work = blocks_per_view * (uint32_t)in.V;
    // lane 0 claims; the claim is broadcast only when the block it is for begins (the loop's increment), so the
    // atomic's round trip to the L2 lies under the block before (the shuffle right behind the atomic held 2.5 % of
    // this kernel's stall samples, profiles/r02_v6)
    auto claim_work = [&]() -> uint32_t { return lane == 0 ? atomicAdd(&num_rendered[2], 1u) : 0u; };
    // backgrounds of the first views in shared memory: the epilogue's 32 FFMAs otherwise each wait for a global load
    constexpr int BG_VIEWS = 8;
    __shared__ float s_bg[BG_VIEWS][NUM_CHANNELS];
    for (int vv = 0; vv < min(in.V, BG_VIEWS); vv++) s_bg[vv][lane] = in.background[(size_t)vv * NUM_CHANNELS + lane];
    __syncwarp();

  uint32_t claimed = claim_work();
  for (uint32_t work = __shfl_sync(0xffffffffu, claimed, 0); work < total_work; work = __shfl_sync(0xffffffffu, claimed, 0)) {
    claimed = claim_work();   // the next block's claim, in flight while this block is blended
    // one 16-byte work item (tile, list range) instead of two dependent loads (tile order -> range); claiming
    // two blocks ahead to hide this load as well measured 10 % slower (coarser dynamic balancing)
    const uint4 item = tile_order[work >> 3];   // longest lists first
    const uint32_t gtile = item.x;
    const uint2 range = make_uint2(item.y, item.z);
    const int sub = (int)(work & 7u);
    const int v = (int)(gtile / (uint32_t)in.tiles);
    const int tile = (int)(gtile - (uint32_t)v * (uint32_t)in.tiles);
    const int tile_x = tile % in.gx, tile_y = tile / in.gx;
    const uint32_t bx = tile_x * TILE_X + (sub & 1) * 8, by = tile_y * TILE_Y + (sub >> 1) * 4;
    if (bx >= (uint32_t)in.W || by >= (uint32_t)in.H) continue;   // block entirely outside the image
    const uint32_t px = bx + (lane & 7), py = by + (lane >> 3);
    const bool inside = px < (uint32_t)in.W && py < (uint32_t)in.H;
    const float pxf = (float)px, pyf = (float)py;
    const float wx0 = (float)bx, wy0 = (float)by;

    const int len = (int)(range.y - range.x);
    const int nchunks = (len + 31) >> 5;

    const SplatRec* __restrict__ vrecs = recs + (size_t)v * in.P;
    const float* __restrict__ vfeat = in.colors + (size_t)v * in.colors_vs;
    const uint32_t vrow0 = (uint32_t)(((size_t)v * in.colors_vs) / NUM_CHANNELS);   // TMA: first tensor-map row of this view

    constexpr uint32_t NO_ID = 0xffffffffu;
    // Software pipeline over 32-entry chunks of the tile list.  While chunk c is blended,
    // the SplatRecs of chunk c+1 (gathered through ids loaded one iteration earlier) and
    // the ids of chunk c+2 are in flight; chunk c+1 is tested and its survivors' feature
    // rows are requested at the top of the next iteration.
    auto load_id = [&](int c) -> uint32_t {
        const int k = (c << 5) + lane;
        return k < len ? point_list[range.x + k] : NO_ID;
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
    // Test chunk c (one list entry per lane) and stage the survivors into stage[c & 1].
    // Returns the survivor mask (bit = lane = list offset within the chunk).
    // MASKS: bit = lane of the entries of the chunk being blended / being staged that the footprint test rejected;
    // `untaken`: bit = slot of the chunk's survivors that no pixel of the block took (filled by the blend loop)
    uint32_t marks_cur = 0u, marks_next = 0u;
    auto flush_marks = [&](int c, uint32_t culled, uint32_t survivors, uint32_t untaken) {   // one byte store per lane
        if (!MASKS) return;
        const int k = (c << 5) + lane;
        if (k >= len) return;
        const int slot = __popc(survivors & lt_mask);
        const bool mine = ((culled >> lane) & 1u) || ((((survivors >> lane) & 1u) != 0u) && (((untaken >> slot) & 1u) != 0u));
        taken_marks[(size_t)sub * marks_plane + (size_t)range.x + (size_t)k] = mine ? 1 : 0;   // 32 consecutive bytes per chunk
    };
    auto stage_chunk = [&](int c, uint32_t id, const float4& c0, const float4& c1) -> uint32_t {
        const bool keep = id != NO_ID &&
                          splat_may_touch(c0.x, c0.y, c0.z, c0.w, c1.x, c1.w, wx0, wx0 + 7.0f, wy0, wy0 + 3.0f);
        const uint32_t mask = __ballot_sync(0xffffffffu, keep);
        if (MASKS) marks_next = __ballot_sync(0xffffffffu, id != NO_ID) & ~mask;   // rejected by the footprint test
        if (mask != 0u) {
            FwdStage& st = stage[c & 1];
            if (keep) {
                const int slot = __popc(mask & lt_mask);
                const int o = (slot >> 1) * 4 + (slot & 1);
                float* q;
                q = &st.xy[0].x + o;   q[0] = c0.x; q[2] = c0.y;
                q = &st.cab[0].x + o;  q[0] = c0.z; q[2] = -c0.w;
                q = &st.cop[0].x + o;  q[0] = c1.x; q[2] = c1.y;
                q = &st.dpos[0].x + o; q[0] = c1.z; q[2] = __uint_as_float((uint32_t)(c << 5) + (uint32_t)lane + 1u);
                if (TMA) {
                    srows[c & 1][slot] = vrow0 + id;
                } else {
                    // the lane's own 128-byte feature row as eight 16-byte asynchronous copies; the piece order is
                    // rotated by the slot so that the 32 lanes of one instruction fall on distinct bank groups
                    const char* src = reinterpret_cast<const char*>(vfeat + (size_t)id * NUM_CHANNELS);
                    char* dstf = reinterpret_cast<char*>(&st.feat[slot][0]);
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        const int piece = ((j + slot) & 7) * 16;
                        tma::cp_async16(dstf + piece, src + piece);
                    }
                }
            }
            if (TMA) {
                // four survivor rows per instruction; the last group is padded with the first survivor's row (its
                // copies land in unused slots), so every instruction completes exactly 512 bytes
                const int nsurv = __popc(mask);
                const int ngroups = (nsurv + 3) >> 2;
                const uint32_t first_row = vrow0 + __shfl_sync(0xffffffffu, id, __ffs(mask) - 1);
                if (lane < 3 && nsurv + lane < 4 * ngroups) srows[c & 1][nsurv + lane] = first_row;
                __syncwarp();
                if (lane == 0) {
                    tma::mbar_arrive_expect_tx(&bars[c & 1], (uint32_t)ngroups * 512u);
                    for (int gq = 0; gq < ngroups; gq++) {
                        const uint4 r = *reinterpret_cast<const uint4*>(&srows[c & 1][4 * gq]);
                        tma::tensor_gather4_g2s(&st.feat[4 * gq][0], &fmap, 0, (int)r.x, (int)r.y, (int)r.z, (int)r.w, &bars[c & 1]);
                    }
                }
            }
        }
        if (!TMA) tma::cp_async_commit();   // one group per chunk, empty or not: the waits below count groups
        return mask;
    };
    // the rows of a staged chunk have landed (and, LDGSTS: every older group)
    auto wait_rows = [&](int s) {
        if (TMA) {
            tma::mbar_wait(&bars[s], (phase >> s) & 1u);
            phase ^= 1u << s;
        } else {
            tma::cp_async_wait<1>();   // everything but the newest group (chunk c+1) has landed
        }
    };

    float T = 1.0f;
    float2 C[NUM_CHANNELS / 2];   // channel pairs: blended with packed FFMA2 (sm_100 fma.rn.f32x2)
#pragma unroll
    for (int ch = 0; ch < NUM_CHANNELS / 2; ch++) C[ch] = make_float2(0.0f, 0.0f);
    float D = 0.0f;
    uint32_t last_contributor = 0;
    bool done = !inside;
    bool warp_done = false;

    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
    uint32_t id_next = load_id(0);
    uint32_t id_nn = load_id(1);
    load_rec(id_next, a0, a1);
    uint32_t mask_cur = stage_chunk(0, id_next, a0, a1);
    marks_cur = marks_next;
    id_next = id_nn;
    load_rec(id_next, a0, a1);
    id_nn = load_id(2);
    for (int c = 0; c < nchunks; c++) {
        uint32_t mask_next = 0u;
        marks_next = 0u;
        if (c + 1 < nchunks) mask_next = stage_chunk(c + 1, id_next, a0, a1);
        else if (!TMA) tma::cp_async_commit();
        id_next = id_nn;
        load_rec(id_next, a0, a1);
        id_nn = load_id(c + 3);
        const int s = c & 1;
        uint32_t untaken = 0u, ubit = 1u;   // MASKS
        if (mask_cur != 0u) {
            wait_rows(s);
            __syncwarp();              // ... and the staging lanes' stores and copies are visible to every lane
            const FwdStage& st = stage[s];
            const float4* __restrict__ featp = reinterpret_cast<const float4*>(&st.feat[0][0]);
            const int nsurv = __popc(mask_cur);
            // Survivors are walked two at a time.  The state-independent half of both alpha chains (offsets,
            // power, exp, alpha, 1 - alpha) is evaluated with packed f32x2 instructions -- per half the very
            // operations, in the very order, of forward.cu:350-358 as compiled -- the transmittance updates
            // then run strictly in list order, so every pixel sees exactly the reference's sequence.
            // ... and it is software-pipelined: the chain of pair k+1 (LDS -> offsets -> power -> exp -> alpha) is
            // issued before the transmittance updates and the blends of pair k, whose FFMA2 stream hides its latency.
            float2 pw_n, al_n, om_n;
            float4 pd_n;
            auto eval_pair = [&](int k) {
                const float4 pa = st.xy[k], pb = st.cab[k], pc = st.cop[k];
                pd_n = st.dpos[k];
                const float2 dx = __fadd2_rn(f2(pa.x, pa.y), f2(-pxf));
                const float2 dy = __fadd2_rn(f2(pa.z, pa.w), f2(-pyf));
                const float2 qd = __ffma2_rn(dx, __fmul2_rn(dx, f2(pb.x, pb.y)), __fmul2_rn(dy, __fmul2_rn(dy, f2(pc.x, pc.y))));
                pw_n = __ffma2_rn(qd, f2(-0.5f), __fmul2_rn(dy, __fmul2_rn(dx, f2(pb.z, pb.w))));   // pb.zw = -conic.y
                const float2 ag = __fmul2_rn(f2(pc.z, pc.w), expf_pair(pw_n));
                al_n = f2(fminf(ag.x, 0.99f), fminf(ag.y, 0.99f));
                om_n = __ffma2_rn(al_n, f2(-1.0f), f2(1.0f));                                        // 1 - alpha
            };
            eval_pair(0);
            for (int k = 0; 2 * k < nsurv && !warp_done; k++, featp += 2 * (NUM_CHANNELS / 4)) {
                const bool has1 = 2 * k + 1 < nsurv;
                const float2 pw = pw_n, om = om_n;
                const float alpha0 = al_n.x, alpha1 = al_n.y;
                const float4 pd = pd_n;
                eval_pair(min(k + 1, 15));   // past the last pair: a harmless re-evaluation of staged (or stale) data
                const bool ok0 = !(pw.x > 0.0f) && !(alpha0 < 1.0f / 255.0f);
                const bool ok1 = has1 && !(pw.y > 0.0f) && !(alpha1 < 1.0f / 255.0f);
                // in list order
                bool take0 = !done && ok0;
                const float test_T0 = fmul(T, om.x);
                if (take0 && test_T0 < 0.0001f) { done = true; take0 = false; }
                const float w0 = take0 ? fmul(alpha0, T) : 0.0f;
                if (take0) { T = test_T0; last_contributor = __float_as_uint(pd.z); }
                bool take1 = !done && ok1;
                const float test_T1 = fmul(T, om.y);
                if (take1 && test_T1 < 0.0001f) { done = true; take1 = false; }
                const float w1 = take1 ? fmul(alpha1, T) : 0.0f;
                if (take1) { T = test_T1; last_contributor = __float_as_uint(pd.w); }
                const bool any0 = __any_sync(0xffffffffu, take0);
                const bool any1 = __any_sync(0xffffffffu, take1);
                if (MASKS) {   // a survivor no pixel of the block takes contributes nothing in the backward either
                    untaken |= any0 ? 0u : ubit;
                    untaken |= (has1 && !any1) ? (ubit << 1) : 0u;
                    ubit <<= 2;
                }
                if (any0) {
                    const float2 w2 = make_float2(w0, w0);
#pragma unroll
                    for (int q = 0; q < NUM_CHANNELS / 4; q++) {
                        const float4 f = featp[q];
                        C[2 * q + 0] = __ffma2_rn(w2, make_float2(f.x, f.y), C[2 * q + 0]);
                        C[2 * q + 1] = __ffma2_rn(w2, make_float2(f.z, f.w), C[2 * q + 1]);
                    }
                    D = fmaf(w0, pd.x, D);
                }
                if (any1) {
                    const float2 w2 = make_float2(w1, w1);
#pragma unroll
                    for (int q = 0; q < NUM_CHANNELS / 4; q++) {
                        const float4 f = featp[NUM_CHANNELS / 4 + q];
                        C[2 * q + 0] = __ffma2_rn(w2, make_float2(f.x, f.y), C[2 * q + 0]);
                        C[2 * q + 1] = __ffma2_rn(w2, make_float2(f.z, f.w), C[2 * q + 1]);
                    }
                    D = fmaf(w1, pd.y, D);
                }
                if (__all_sync(0xffffffffu, done)) warp_done = true;
            }
        }
        flush_marks(c, marks_cur, mask_cur, untaken);
        marks_cur = marks_next;
        if (warp_done) {   // never leave the block with a copy still in flight
            if (TMA) { if (mask_next != 0u) wait_rows(s ^ 1); }
            else tma::cp_async_wait<0>();
            break;
        }
        mask_cur = mask_next;
        __syncwarp();   // every lane is done reading stage[s] before chunk c+2 overwrites it
    }

    if (inside) {
        const size_t HW = (size_t)in.H * in.W;
        const size_t pix = (size_t)py * in.W + px;
        final_T[(size_t)v * HW + pix] = T;
        n_contrib[(size_t)v * HW + pix] = last_contributor;
        const float* bg = v < BG_VIEWS ? &s_bg[v][0] : in.background + (size_t)v * NUM_CHANNELS;
        float* __restrict__ oc = out_color + (size_t)v * NUM_CHANNELS * HW + pix;
#pragma unroll
        for (int ch = 0; ch < NUM_CHANNELS / 2; ch++) {
            oc[(size_t)(2 * ch) * HW] = fmaf(T, bg[2 * ch], C[ch].x);
            oc[(size_t)(2 * ch + 1) * HW] = fmaf(T, bg[2 * ch + 1], C[ch].y);
        }
        if (out_invdepth != nullptr) out_invdepth[(size_t)v * HW + pix] = D;
        // Refiner hand-off (gaussian_render.py:70-73): the same 32 values once more as one 64-byte
        // channels-last bf16 pixel, so the refiner's first convolution reads its preferred layout
        // without a separate permute + cast pass over the image.
        if (in.aux_nhwc_bf16 != nullptr) {
            uint4* ap = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(in.aux_nhwc_bf16) +
                                                 ((size_t)v * HW + pix) * NUM_CHANNELS);
#pragma unroll
            for (int q = 0; q < NUM_CHANNELS / 8; q++) {
                uint32_t w[4];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const int ch = 4 * q + k;
                    const __nv_bfloat162 h = __floats2bfloat162_rn(fmaf(T, bg[2 * ch], C[ch].x),
                                                                   fmaf(T, bg[2 * ch + 1], C[ch].y));
                    w[k] = *reinterpret_cast<const uint32_t*>(&h);
                }
                ap[q] = make_uint4(w[0], w[1], w[2], w[3]);
            }
        }
    }
    if (!TMA) tma::cp_async_wait<0>();
    __syncwarp();   // the stages are reused by the next block
  }
}

// Self-test hook (guava_selftest_cull): one thread per (splat, 8x4 block) sample evaluates splat_may_touch exactly
// as the compositors do (threshold from preprocess's __logf) AND the per-pixel alpha test at all 32 pixel centres of
// the block with the compositors' own arithmetic (expf_pair; forward.cu:350-362).  rejected = the footprint test
// says "cannot touch"; passing = number of pixels at which the splat passes the alpha test.  The culling is
// output-neutral iff no sample has rejected != 0 && passing != 0.
__global__ void __launch_bounds__(256)
selftest_cull_kernel(int n, const float* __restrict__ splats, const float* __restrict__ origins,
                     uint8_t* __restrict__ rejected, uint8_t* __restrict__ passing)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    const float mx = splats[6 * i], my = splats[6 * i + 1], A = splats[6 * i + 2], B = splats[6 * i + 3],
                C = splats[6 * i + 4], op = splats[6 * i + 5];
    const float x0 = origins[2 * i], y0 = origins[2 * i + 1];
    const float lthr = __logf(255.0f * op) + 0.01f;   // SplatRec::lthr, preprocess.cu
    rejected[i] = splat_may_touch(mx, my, A, B, C, lthr, x0, x0 + 7.0f, y0, y0 + 3.0f) ? 0 : 1;
    int pass = 0;
    for (int p = 0; p < 32; p += 2) {
        const float pxa = x0 + (float)(p & 7), pya = y0 + (float)(p >> 3);
        const float pxb = x0 + (float)((p + 1) & 7), pyb = y0 + (float)((p + 1) >> 3);
        const float2 dx = __fadd2_rn(f2(mx), f2(-pxa, -pxb));
        const float2 dy = __fadd2_rn(f2(my), f2(-pya, -pyb));
        const float2 qd = __ffma2_rn(dx, __fmul2_rn(dx, f2(A)), __fmul2_rn(dy, __fmul2_rn(dy, f2(C))));
        const float2 pw = __ffma2_rn(qd, f2(-0.5f), __fmul2_rn(dy, __fmul2_rn(dx, f2(-B))));
        const float2 ag = __fmul2_rn(f2(op), expf_pair(pw));
        const float a0 = fminf(ag.x, 0.99f), a1 = fminf(ag.y, 0.99f);
        pass += (!(pw.x > 0.0f) && !(a0 < 1.0f / 255.0f)) ? 1 : 0;
        pass += (!(pw.y > 0.0f) && !(a1 < 1.0f / 255.0f)) ? 1 : 0;
    }
    passing[i] = (uint8_t)pass;
}

void launch_selftest_cull(int n, const float* splats, const float* origins, uint8_t* rejected, uint8_t* passing, cudaStream_t s)
{
    if (n <= 0) return;
    selftest_cull_kernel<<<(n + 255) / 256, 256, 0, s>>>(n, splats, origins, rejected, passing);
    count_launch();
}

// Tensor map of colors_precomp as a [rows, 32] fp32 matrix with a {32, 1} box: what tile::gather4 indexes by row.
// Encoded on the host per call (the base address is the caller's tensor); cuTensorMapEncodeTiled is taken from the
// driver through the runtime, so the library does not link against libcuda.
bool make_feature_tensor_map(const Inputs& in, CUtensorMap* out)
{
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn encode = [] {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            fn = nullptr;
        return reinterpret_cast<EncodeFn>(fn);
    }();
    if (encode == nullptr || (in.colors_vs % NUM_CHANNELS) != 0) return false;
    const cuuint64_t rows = (cuuint64_t)(in.V > 0 ? in.V - 1 : 0) * (in.colors_vs / NUM_CHANNELS) + (cuuint64_t)in.P;
    const cuuint64_t dims[2] = {NUM_CHANNELS, rows};
    const cuuint64_t strides[1] = {NUM_CHANNELS * sizeof(float)};
    const cuuint32_t box[2] = {NUM_CHANNELS, 1};
    const cuuint32_t estr[2] = {1, 1};
    return encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(in.colors), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

void launch_render_forward(const Inputs& in, const GeomPtrs& g, const ImagePtrs& img, const BinningPtrs& b,
                           float* out_color, float* out_invdepth, int taken_masks, int tma_staging, cudaStream_t s)
{
    if (in.tiles == 0 || in.V == 0) return;
    CUtensorMap fmap;
    memset(&fmap, 0, sizeof(fmap));
    if (tma_staging && !make_feature_tensor_map(in, &fmap)) tma_staging = 0;   // no driver entry point: LDGSTS staging
    const long long total = (long long)in.tiles * 8 * in.V;
    static const int ctas_per_sm = [] { const char* e = getenv("GUAVA_FWD_CTAS"); return e ? atoi(e) : 16; }();   // tuning knob
    const long long full = (long long)sm_count() * ctas_per_sm;
    const unsigned grid = (unsigned)(total < full ? total : full);
#define GUAVA_LAUNCH_FWD(MK, TM)                                                                                        \
    render_forward_kernel<MK, TM><<<grid, 32, 0, s>>>(in, fmap, g.rec, img.ranges, img.tile_order, b.point_list,             \
                                                      b.taken_marks, (b.capacity + 255) / 256 * 256, img.num_rendered,        \
                                                      img.final_T, img.n_contrib, out_color, out_invdepth)
    if (taken_masks) { if (tma_staging) GUAVA_LAUNCH_FWD(1, 1); else GUAVA_LAUNCH_FWD(1, 0); }
    else             { if (tma_staging) GUAVA_LAUNCH_FWD(0, 1); else GUAVA_LAUNCH_FWD(0, 0); }
#undef GUAVA_LAUNCH_FWD
    count_launch();
}

}  // namespace guava
