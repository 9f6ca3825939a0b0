// C ABI (include/guava_rasterizer.h) and host-side orchestration.
//
// Replaces DGR32/cuda_rasterizer/rasterizer_impl.cu:141-450
// (CudaRasterizer::Rasterizer::{markVisible, forward, backward}).
//
// Launch chain of one forward call (single view or V batched views):
//   memset(tile counters) -> preprocess(+count) -> tile scan -> scatter ->
//   tile sort (3 size classes) -> composite
// There is no device->host dependency inside the chain.  The reference blocks
// on a cudaMemcpy of num_rendered (rasterizer_impl.cu:284) before it can size
// the binning buffer; here the buffer is sized from a running capacity estimate,
// the whole chain is enqueued, and only then the host reads num_rendered from a
// pinned word the scan kernel wrote (by then long finished).  If the estimate
// was too small the kernels skipped themselves (overflow flag) and the tail of
// the chain is re-enqueued with the exact size.
#include "internal.cuh"
#include "../../include/guava_rasterizer.h"

#include <atomic>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>

namespace guava {

static std::atomic<unsigned long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

static thread_local std::string t_last_error;
static std::atomic<long long> g_opt_emit_sorted_keys{0};
static std::atomic<long long> g_opt_exact_binning{0};
static std::atomic<long long> g_opt_capacity_hint{0};
static std::atomic<long long> g_opt_tma_staging{0};   // compositors stage feature rows with cp.async.bulk.tensor gather4 (1) or LDGSTS (0)
static std::atomic<long long> g_opt_taken_masks{1};   // see BLOCK_MASK_SHIFT (internal.cuh); 0 = the round-1 compositors, kept for A/B
static std::atomic<long long> g_opt_os_identify_ranges{0};   // sort mode 1: recompute the tile ranges from the sorted keys (identifyTileRanges restated) although the tile scan already wrote them
static std::atomic<long long> g_opt_sort_mode{-1};   // 0 = tile buckets (binning.cu), 1 = global onesweep sort (onesweep.cu)

// Number of SMs of the current device (grids of the persistent kernels are multiples of it); cached per ordinal.
int sm_count()
{
    static std::atomic<int> cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
    int n = cached[dev].load(std::memory_order_relaxed);
    if (n == 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev].store(n, std::memory_order_relaxed);
    }
    return n;
}

static int fail(int code, const std::string& msg)
{
    t_last_error = msg;
    return code;
}

int report_error(int code, const char* msg) { return fail(code, msg ? msg : ""); }

#define GUAVA_CUDA(expr)                                                                        \
    do {                                                                                        \
        cudaError_t e__ = (expr);                                                               \
        if (e__ != cudaSuccess)                                                                 \
            return fail(GUAVA_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__));   \
    } while (0)

// reference: CHECK_CUDA (auxiliary.h:178-185): in debug mode synchronise after
// every stage and report the failing stage
#define GUAVA_STAGE(what, debug)                                                                \
    do {                                                                                        \
        cudaError_t e__ = cudaGetLastError();                                                   \
        if (e__ == cudaSuccess && (debug)) e__ = cudaStreamSynchronize(stream);                 \
        if (e__ != cudaSuccess)                                                                 \
            return fail(GUAVA_ERR_CUDA, std::string("[CUDA ERROR] in ") + what + ": " +         \
                                            cudaGetErrorString(e__));                           \
    } while (0)

static inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// ------------------------------------------------------------ stage profiling
// Option "profile_stages": bracket every kernel of a call with CUDA events on the
// launching stream and accumulate per-stage device time (read back with
// guava_stage_time).  Used by bench.py for the roofline object; off by default.
static std::atomic<long long> g_opt_profile_stages{0};
static const char* const kStageNames[] = {"preprocess", "tile_scan", "scatter", "tile_sort", "render_forward",
                                          "render_backward", "preprocess_backward"};
constexpr int kNumStages = 7;
static double g_stage_ms[kNumStages];
static unsigned long long g_stage_calls[kNumStages];
static std::mutex g_stage_mu;

struct StageTimer {
    bool on;
    cudaStream_t stream;
    struct Rec { int stage; cudaEvent_t a, b; };
    Rec recs[16];
    int n = 0;
    StageTimer(cudaStream_t s) : on(g_opt_profile_stages.load() != 0), stream(s) {}
    void begin(int stage)
    {
        if (!on || n >= 16) return;
        recs[n].stage = stage;
        cudaEventCreate(&recs[n].a);
        cudaEventCreate(&recs[n].b);
        cudaEventRecord(recs[n].a, stream);
    }
    void end()
    {
        if (!on || n >= 16) return;
        cudaEventRecord(recs[n].b, stream);
        n++;
    }
    ~StageTimer()
    {
        if (!on || n == 0) return;
        cudaEventSynchronize(recs[n - 1].b);
        std::lock_guard<std::mutex> lk(g_stage_mu);
        for (int i = 0; i < n; i++) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, recs[i].a, recs[i].b) == cudaSuccess) {
                g_stage_ms[recs[i].stage] += ms;
                g_stage_calls[recs[i].stage] += 1;
            }
            cudaEventDestroy(recs[i].a);
            cudaEventDestroy(recs[i].b);
        }
    }
};

static void compute_layout(int V, int P, int W, int H, size_t capacity, bool emit_keys, guava_layout* L, int sort_mode = 0,
                           bool taken_masks = false)
{
    const size_t n = (size_t)V * P;
    const size_t gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
    const size_t tiles = (size_t)V * gx * gy;
    const size_t npix = (size_t)V * H * W;
    size_t o = 0;
    L->splat_rec = o;     o = align_up(o + n * sizeof(SplatRec));
    L->depths = o;        o = align_up(o + n * 4);
    L->rects = o;         o = align_up(o + n * 8);
    L->tiles_touched = o; o = align_up(o + n * 4);
    L->grad_geo = o;      o = align_up(o + n * 32);
    L->point_offsets = o; o = align_up(o + (sort_mode == 1 ? n * 4 : 0));
    L->geom_bytes = o;
    o = 0;
    L->final_T = o;      o = align_up(o + npix * 4);
    L->n_contrib = o;    o = align_up(o + npix * 4);
    L->ranges = o;       o = align_up(o + tiles * 8);
    L->tile_cursor = o;  o = align_up(o + tiles * 4 * CURSOR_STRIDE + tile_scan_scratch_bytes(tiles));   // + scan scratch
    L->tile_order = o;   o = align_up(o + tiles * 16);
    L->num_rendered = o; o = align_up(o + 32);
    L->image_bytes = o;
    o = 0;
    L->point_list = o;    o = align_up(o + capacity * 4);
    L->taken_marks = o;   o = align_up(o + (taken_masks ? align_up(capacity) * 8 : 0));   // right behind point_list: the backward finds it from the capacity in the call's header
    L->keys_unsorted = o; o = align_up(o + capacity * 8);
    L->keys_sorted = o;   o = align_up(o + ((emit_keys || sort_mode == 1) ? capacity * 8 : 0));
    L->os_vals_alt = o;   o = align_up(o + (sort_mode == 1 ? capacity * 4 : 0));
    L->os_scratch = o;    o = align_up(o + (sort_mode == 1 ? onesweep_scratch_bytes(n, capacity, (int)tiles) : 0));
    L->binning_bytes = o;
}

// ------------------------------------------------ num_rendered readback slots
struct HostSlots {
    uint32_t* host = nullptr;   // pinned, mapped
    uint32_t* dev = nullptr;
    std::atomic<unsigned> next{0};
    static constexpr unsigned N = 256;
};
static HostSlots g_slots;
static std::once_flag g_slots_once;
static cudaError_t g_slots_err = cudaSuccess;

static cudaError_t ensure_slots()
{
    std::call_once(g_slots_once, [] {
        void* h = nullptr;
        g_slots_err = cudaHostAlloc(&h, HostSlots::N * 2 * sizeof(uint32_t), cudaHostAllocMapped | cudaHostAllocPortable);
        if (g_slots_err != cudaSuccess) return;
        g_slots.host = (uint32_t*)h;
        void* d = nullptr;
        g_slots_err = cudaHostGetDevicePointer(&d, h, 0);
        g_slots.dev = (uint32_t*)d;
    });
    return g_slots_err;
}

// --------------------------------------------------------- capacity estimates
struct CapEntry { size_t n; int W, H; size_t cap; uint32_t longest; };
static CapEntry g_caps[16];
static unsigned g_caps_next = 0;
static std::mutex g_caps_mu;

static size_t capacity_estimate(size_t n, int W, int H)
{
    const long long forced = g_opt_capacity_hint.load();
    if (forced > 0) return (size_t)forced;
    std::lock_guard<std::mutex> lk(g_caps_mu);
    for (auto& e : g_caps)
        if (e.n == n && e.W == W && e.H == H && e.cap) return e.cap;
    size_t c = 16 * n;
    if (c < (1u << 20)) c = 1u << 20;
    return c;
}
// longest tile list the last call of this shape produced (0 = shape not seen yet)
static uint32_t longest_estimate(size_t n, int W, int H)
{
    std::lock_guard<std::mutex> lk(g_caps_mu);
    for (auto& e : g_caps)
        if (e.n == n && e.W == W && e.H == H && e.cap) return e.longest;
    return 0;
}
static void capacity_record(size_t n, int W, int H, size_t R, uint32_t longest)
{
    std::lock_guard<std::mutex> lk(g_caps_mu);
    const size_t want = R + R / 4 + 4096;
    for (auto& e : g_caps)
        if (e.n == n && e.W == W && e.H == H && e.cap) {
            // grow immediately, shrink slowly
            e.cap = want > e.cap ? want : (e.cap * 15 + want) / 16;
            e.longest = longest;
            return;
        }
    g_caps[g_caps_next++ % 16] = CapEntry{n, W, H, want, longest};
}

__global__ void reset_cursor_kernel(int n, const uint2* __restrict__ ranges, uint32_t* __restrict__ cursor,
                                    uint32_t* __restrict__ num_rendered, unsigned long long capacity)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) cursor[(size_t)i * CURSOR_STRIDE] = ranges[i].x;
    if (i == 0) {
        num_rendered[1] = ((unsigned long long)num_rendered[0] > capacity) ? 1u : 0u;
        num_rendered[2] = 0u;
        num_rendered[6] = (uint32_t)capacity;           // header: the capacity this call's binning buffer has now
        num_rendered[7] = (uint32_t)(capacity >> 32);
    }
}

// One warp per packed [43] row, four rows in flight: lane l sums feature column l over the V views
// (one coalesced 128-byte load per view and row) and lanes 0..10 the eleven geometry columns, so the
// kernel has no index arithmetic beyond the row stride (the first version, one thread per output
// element with a division by 43, was instruction-bound at 2.6 TB/s).  The views are added in index
// order starting from 0.0f.  V is a template parameter for the common batch sizes so that the loads
// of a row are independent and in flight together (VT == 0: runtime loop).
template <int VT>
__global__ void __launch_bounds__(256)
pack_view_gradients_kernel(int V, int P, const float* __restrict__ g_xyz, const float* __restrict__ g_sc,
                           const float* __restrict__ g_rot, const float* __restrict__ g_op,
                           const float* __restrict__ g_feat, float* __restrict__ packed)
{
    constexpr int ROWS = 4;
    const int lane = threadIdx.x & 31;
    const size_t warp = ((size_t)blockIdx.x * 256 + threadIdx.x) >> 5;
    const size_t nwarps = (size_t)gridDim.x * 8;
    const float* gsrc = g_op;
    int gw = 1, gc = 0;
    if (lane < 3) { gsrc = g_xyz; gw = 3; gc = lane; }
    else if (lane < 6) { gsrc = g_sc; gw = 3; gc = lane - 3; }
    else if (lane < 10) { gsrc = g_rot; gw = 4; gc = lane - 6; }
    const bool geo = lane < 11;
    const int nv = VT > 0 ? VT : V;
    for (size_t p0 = warp * ROWS; p0 < (size_t)P; p0 += nwarps * ROWS) {
        float f[ROWS], q[ROWS];
#pragma unroll
        for (int r = 0; r < ROWS; r++) { f[r] = 0.0f; q[r] = 0.0f; }
        if (VT > 0) {
            float xf[ROWS][VT > 0 ? VT : 1], xq[ROWS][VT > 0 ? VT : 1];
#pragma unroll
            for (int r = 0; r < ROWS; r++) {
                const size_t p = p0 + r < (size_t)P ? p0 + r : (size_t)P - 1;
#pragma unroll
                for (int v = 0; v < VT; v++) {
                    xf[r][v] = __ldcs(g_feat + ((size_t)v * P + p) * 32 + lane);
                    xq[r][v] = geo ? __ldcs(gsrc + ((size_t)v * P + p) * gw + gc) : 0.0f;
                }
            }
#pragma unroll
            for (int r = 0; r < ROWS; r++)
#pragma unroll
                for (int v = 0; v < VT; v++) { f[r] += xf[r][v]; q[r] += xq[r][v]; }
        } else {
            for (int v = 0; v < nv; v++)
#pragma unroll
                for (int r = 0; r < ROWS; r++) {
                    const size_t p = p0 + r < (size_t)P ? p0 + r : (size_t)P - 1;
                    f[r] += __ldcs(g_feat + ((size_t)v * P + p) * 32 + lane);
                    if (geo) q[r] += __ldcs(gsrc + ((size_t)v * P + p) * gw + gc);
                }
        }
#pragma unroll
        for (int r = 0; r < ROWS; r++) {
            const size_t p = p0 + r;
            if (p < (size_t)P) {
                packed[p * 43 + 11 + lane] = f[r];
                if (geo) packed[p * 43 + lane] = q[r];
            }
        }
    }
}

// <a, b> with a deliberately small footprint: `IMAGE_DOT_CTAS` CTAs stream both arrays while a
// persistent compositor owns the rest of the machine (a full-grid library reduction launched next to the
// backward compositor delayed its first wave; launched after it, it was exposed at the end of the step).
// Deterministic: per-thread partial sums over a fixed stride, tree reduction per CTA, and the CTA that
// takes the last ticket adds the per-CTA partials in index order.
constexpr int IMAGE_DOT_CTAS = 16;
__global__ void __launch_bounds__(512)
image_dot_kernel(const float* __restrict__ a, const float* __restrict__ b, size_t n, float* __restrict__ out)
{
    __shared__ float warp_part[16];
    __shared__ bool last;
    const size_t n4 = n >> 2;
    const float4* __restrict__ a4 = reinterpret_cast<const float4*>(a);
    const float4* __restrict__ b4 = reinterpret_cast<const float4*>(b);
    const size_t stride = (size_t)gridDim.x * 512;
    float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    size_t i = (size_t)blockIdx.x * 512 + threadIdx.x;
    for (; i + 3 * stride < n4; i += 4 * stride) {
        float4 x[4], y[4];
#pragma unroll
        for (int u = 0; u < 4; u++) { x[u] = __ldcs(a4 + i + u * stride); y[u] = __ldcs(b4 + i + u * stride); }
#pragma unroll
        for (int u = 0; u < 4; u++)
            acc[u] += x[u].x * y[u].x + x[u].y * y[u].y + x[u].z * y[u].z + x[u].w * y[u].w;
    }
    for (; i < n4; i += stride) {
        const float4 x = __ldcs(a4 + i), y = __ldcs(b4 + i);
        acc[0] += x.x * y.x + x.y * y.y + x.z * y.z + x.w * y.w;
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) acc[1] += a[(n4 << 2) + threadIdx.x] * b[(n4 << 2) + threadIdx.x];
    float v = (acc[0] + acc[1]) + (acc[2] + acc[3]);
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) warp_part[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
        for (int w = 0; w < 16; w++) t += warp_part[w];
        out[1 + blockIdx.x] = t;
        __threadfence();
        const unsigned ticket = atomicAdd(reinterpret_cast<unsigned*>(out) + 63, 1u);
        last = ticket == gridDim.x - 1;
    }
    __syncthreads();
    if (last && threadIdx.x == 0) {
        __threadfence();
        float t = 0.0f;
        for (unsigned c = 0; c < gridDim.x; c++) t += reinterpret_cast<volatile float*>(out)[1 + c];
        out[0] = t;
        reinterpret_cast<unsigned*>(out)[63] = 0u;   // ready for the next call
    }
}

// ------------------------------------------------------------------- forward
static int forward_impl(guava_alloc_fn geometryBuffer, void* geometry_user,
                        guava_alloc_fn binningBuffer, void* binning_user,
                        guava_alloc_fn imageBuffer, void* image_user,
                        Inputs in, float* out_color, float* out_invdepth, int* radii,
                        int debug, cudaStream_t stream, int* num_rendered, const guava_call_options* co)
{
    if (num_rendered) *num_rendered = 0;
    if (co != nullptr && co->struct_size != sizeof(guava_call_options))
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "guava_call_options::struct_size does not match this library");
    in.aux_nhwc_bf16 = co ? co->refiner_output_nhwc_bf16 : nullptr;
    const int block_masks = g_opt_taken_masks.load() != 0 && !(co != nullptr && co->forward_only != 0);
    uint32_t* const status_host = co ? (uint32_t*)co->status_words : nullptr;
    if (((uintptr_t)status_host & 7) != 0)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "status words must be 8-byte aligned");
    if (((uintptr_t)in.aux_nhwc_bf16 & 15) != 0)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "refiner output must be 16-byte aligned");
    if (co != nullptr && co->zero_fill != nullptr && ((((uintptr_t)co->zero_fill) | co->zero_fill_bytes) & 3) != 0)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "zero_fill must be 4-byte aligned and a multiple of 4 bytes long");
    if (in.V < 0 || in.P < 0 || in.W <= 0 || in.H <= 0)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "negative size");
    if ((size_t)in.V * in.P == 0) {   // rasterize_points.cu:88: outputs stay as allocated
        if (co != nullptr && co->zero_fill != nullptr && co->zero_fill_bytes)
            GUAVA_CUDA(cudaMemsetAsync(co->zero_fill, 0, co->zero_fill_bytes, stream));
        return GUAVA_OK;
    }
    if (in.colors == nullptr)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "For non-RGB, provide precomputed Gaussian colors!");
    if (in.means3D == nullptr || in.opacities == nullptr || in.viewmatrix == nullptr || in.projmatrix == nullptr ||
        in.background == nullptr || out_color == nullptr || radii == nullptr)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "null required pointer");
    if (in.cov3D_precomp == nullptr && (in.scales == nullptr || in.rotations == nullptr))
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "need scales+rotations or cov3D_precomp");
    if (((uintptr_t)in.colors & 15) || ((in.colors_vs * sizeof(float)) & 15))
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "colors_precomp must be 16-byte aligned");
    in.gx = (in.W + TILE_X - 1) / TILE_X;
    in.gy = (in.H + TILE_Y - 1) / TILE_Y;
    in.tiles = in.gx * in.gy;
    if (in.gx > 65535 || in.gy > 65535) return fail(GUAVA_ERR_INVALID_ARGUMENT, "image too large");

    const size_t n = (size_t)in.V * in.P;
    const bool emit_keys = g_opt_emit_sorted_keys.load() != 0;
    const bool exact = g_opt_exact_binning.load() != 0;
    const bool async_fwd = co != nullptr && co->no_host_wait != 0 && !exact;
    // "sort_mode": 0 = tile buckets, 1 = global onesweep sort, -1 (default) = by the longest tile list the
    // previous call of this shape produced: lists beyond the register-resident bucket sorts (4096) are where
    // the onesweep passes win (measured: profiles/r02_v2_sort_vs_cub.jsonl -- at P = 202 k, lists of 2-3 k, the
    // buckets take 0.128 ms against 0.158 ms; at C5, lists of 3-8 k, 1.26 against 0.72).  Results are identical.
    int sort_mode = (int)g_opt_sort_mode.load();
    if (sort_mode < 0) sort_mode = longest_estimate((size_t)in.V * in.P, in.W, in.H) > 4096u ? 1 : 0;
    GUAVA_CUDA(ensure_slots());

    guava_layout L;
    size_t cap = capacity_estimate(n, in.W, in.H);
    compute_layout(in.V, in.P, in.W, in.H, cap, emit_keys, &L, sort_mode, block_masks != 0);
    char* geom = geometryBuffer(geometry_user, L.geom_bytes);
    char* img = imageBuffer(image_user, L.image_bytes);
    if (!geom || !img) return fail(GUAVA_ERR_ALLOC, "allocator returned NULL");
    if (((uintptr_t)geom & 127) || ((uintptr_t)img & 127))
        return fail(GUAVA_ERR_ALLOC, "allocator must return 128-byte aligned memory");

    GeomPtrs g;
    g.rec = (SplatRec*)(geom + L.splat_rec);
    g.depths = (float*)(geom + L.depths);
    g.rects = (ushort4*)(geom + L.rects);
    g.tiles_touched = (uint32_t*)(geom + L.tiles_touched);
    g.grad_geo = (float*)(geom + L.grad_geo);
    ImagePtrs im;
    im.final_T = (float*)(img + L.final_T);
    im.n_contrib = (uint32_t*)(img + L.n_contrib);
    im.ranges = (uint2*)(img + L.ranges);
    im.tile_cursor = (uint32_t*)(img + L.tile_cursor);
    im.tile_order = (uint4*)(img + L.tile_order);
    im.num_rendered = (uint32_t*)(img + L.num_rendered);

    // [0] = R, [1] = longest tile list | overflow << 31: the library's rotating pinned slots, or the
    // caller's own two words (guava_rasterizer_set_status_words)
    volatile uint32_t* hslot;
    uint32_t* hslot_dev;
    if (status_host != nullptr) {
        void* d = nullptr;
        if (cudaHostGetDevicePointer(&d, status_host, 0) != cudaSuccess || d == nullptr) {
            cudaGetLastError();
            return fail(GUAVA_ERR_INVALID_ARGUMENT, "status words must be pinned, device-mapped host memory");
        }
        hslot = status_host;
        hslot_dev = (uint32_t*)d;
    } else {
        const unsigned slot = g_slots.next.fetch_add(1) % HostSlots::N;
        hslot = g_slots.host + 2 * slot;
        hslot_dev = g_slots.dev + 2 * slot;
    }
    if (!async_fwd || status_host == nullptr) *hslot = 0xFFFFFFFFu;   // a captured step's words are only written by its replays

    const int total_tiles = in.V * in.tiles;
    StageTimer timer(stream);
    timer.begin(0);
    GUAVA_CUDA(cudaMemsetAsync(im.tile_cursor, 0,
                               (size_t)total_tiles * 4 * CURSOR_STRIDE + tile_scan_scratch_bytes((size_t)total_tiles), stream));
    launch_preprocess_forward(in, g, im, radii, stream);
    timer.end();
    GUAVA_STAGE("preprocess", debug);
    timer.begin(1);
    // The long-bucket sort (lists beyond the register-resident network) is an empty 4 us launch at the
    // GUAVA shapes: it is left out while the previous call of this shape stayed well below the limit.  A
    // list that outgrows the limit anyway is reported by the tile scan exactly like a capacity overflow
    // (every later kernel skips itself) and the call is redone with the long-bucket sort.
    bool with_large = true;
    if (sort_mode == 0 && !exact) {
        const uint32_t est = longest_estimate(n, in.W, in.H);
        if (est != 0 && est <= tile_sort_small_limit() * 3 / 4) with_large = false;
    }
    launch_tile_scan(total_tiles, im, hslot_dev, exact ? (size_t)-1 : cap,
                     with_large ? 0xFFFFFFFFu : tile_sort_small_limit(),
                     (uint32_t)sort_mode | ((uint32_t)block_masks << 8) | ((uint32_t)emit_keys << 16),
                     co ? co->zero_fill : nullptr, co && co->zero_fill ? co->zero_fill_bytes : 0,
                     g.grad_geo, (size_t)in.V * in.P * 32, stream);   // grad_geo: see backward_impl
    timer.end();
    GUAVA_STAGE("tile_scan", debug);

    auto wait_R = [&](uint32_t& R) -> int {
        unsigned spins = 0;
        while (*hslot == 0xFFFFFFFFu) {
            if ((++spins & 0x3ff) == 0) {
                cudaError_t q = cudaStreamQuery(stream);
                if (q == cudaSuccess) {
                    if (*hslot == 0xFFFFFFFFu) return fail(GUAVA_ERR_CUDA, "num_rendered was never written");
                    break;
                }
                if (q != cudaErrorNotReady)
                    return fail(GUAVA_ERR_CUDA, std::string("waiting for num_rendered: ") + cudaGetErrorString(q));
            }
        }
        R = *hslot;
        return GUAVA_OK;
    };

    uint32_t R = 0;
    bool have_R = false;
    if (exact) {
        int rc = wait_R(R);
        if (rc) return rc;
        have_R = true;
        cap = R;
        compute_layout(in.V, in.P, in.W, in.H, cap, emit_keys, &L, sort_mode, block_masks != 0);
        const unsigned long long cap64 = (unsigned long long)cap;   // header words [6], [7] (the scan saw "unbounded")
        GUAVA_CUDA(cudaMemcpyAsync(im.num_rendered + 6, &cap64, 8, cudaMemcpyHostToDevice, stream));
    }
    char* bin = binningBuffer(binning_user, L.binning_bytes ? L.binning_bytes : 256);
    if (!bin) return fail(GUAVA_ERR_ALLOC, "allocator returned NULL");

    for (int attempt = 0; attempt < 2; attempt++) {
        BinningPtrs b;
        b.point_list = (uint32_t*)(bin + L.point_list);
        b.keys_unsorted = (uint64_t*)(bin + L.keys_unsorted);
        b.keys_sorted = emit_keys ? (uint64_t*)(bin + L.keys_sorted) : nullptr;
        b.taken_marks = block_masks ? (uint8_t*)(bin + L.taken_marks) : nullptr;
        b.capacity = cap;
        if (sort_mode == 1) {
            if (cap >= (1u << 30)) return fail(GUAVA_ERR_INVALID_ARGUMENT, "sort_mode 1 supports < 2^30 instances");
            timer.begin(3);
            launch_onesweep_binning(in, g, im, (uint32_t*)(geom + L.point_offsets), (uint64_t*)(bin + L.keys_sorted),
                                    b.point_list, b.keys_unsorted, (uint32_t*)(bin + L.os_vals_alt),
                                    (uint32_t*)(bin + L.os_scratch), cap, g_opt_os_identify_ranges.load() != 0, stream);
            timer.end();
            GUAVA_STAGE("onesweep", debug);
        } else {
        timer.begin(2);
        launch_scatter(in, g, im, b, radii, stream);
        timer.end();
        GUAVA_STAGE("scatter", debug);
        timer.begin(3);
        launch_tile_sort(in, g, im, b, with_large, stream);
        timer.end();
        GUAVA_STAGE("tile_sort", debug);
        }
        timer.begin(4);
        launch_render_forward(in, g, im, b, out_color, out_invdepth, block_masks, g_opt_tma_staging.load() != 0, stream);
        timer.end();
        GUAVA_STAGE("render_forward", debug);
        if (async_fwd) { R = 0xFFFFFFFFu; break; }
        if (!have_R) {
            int rc = wait_R(R);
            if (rc) return rc;
            have_R = true;
        }
        if (attempt == 1 || (hslot[1] & 0x80000000u) == 0u) break;
        // estimate too small (capacity, or a list beyond the sort that was launched): every kernel after
        // the scan skipped itself
        with_large = true;
        if ((size_t)R > cap) cap = (size_t)R;
        compute_layout(in.V, in.P, in.W, in.H, cap, emit_keys, &L, sort_mode, block_masks != 0);
        bin = binningBuffer(binning_user, L.binning_bytes);
        if (!bin) return fail(GUAVA_ERR_ALLOC, "allocator returned NULL");
        reset_cursor_kernel<<<(total_tiles + 255) / 256, 256, 0, stream>>>(total_tiles, im.ranges, im.tile_cursor,
                                                                           im.num_rendered, (unsigned long long)cap);
        count_launch();
    }
    if (!async_fwd) capacity_record(n, in.W, in.H, R, hslot[1] & 0x7FFFFFFFu);
    if (num_rendered) *num_rendered = async_fwd ? -1 : (int)R;
    return GUAVA_OK;
}

// ------------------------------------------------------------------ backward
static int backward_impl(Inputs in, const int* radii, char* geom, char* bin, char* img,
                         const float* dL_dpix, const float* dL_invdepths,
                         float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                         float* dL_dinvdepth, float* dL_dmean3D, float* dL_dcov3D,
                         float* dL_dscale, float* dL_drot, int debug, cudaStream_t stream,
                         const SummedRows* sum = nullptr)   // non-NULL: outputs are sums over the views, rows of these strides
{
    if ((size_t)in.V * in.P == 0) return GUAVA_OK;
    const int block_masks = g_opt_taken_masks.load() != 0;   // a backward after a forward that did not mark finds that in the header
    if (!geom || !bin || !img || !dL_dpix || !dL_dopacity || !dL_dcolor || !dL_dmean3D || !radii)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "null required pointer");
    if (sum == nullptr && (!dL_dmean2D || !dL_dconic || !dL_dcov3D))
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "null required pointer");
    if (in.cov3D_precomp == nullptr && (!dL_dscale || !dL_drot))
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "null required pointer");
    if (sum == nullptr && dL_invdepths != nullptr && dL_dinvdepth == nullptr)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "dL_dinvdepth buffer missing");
    if (sum == nullptr && ((uintptr_t)dL_drot & 15) != 0)   // written with one 16-byte store per Gaussian
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "dL_drot must be 16-byte aligned");
    if (((uintptr_t)in.colors & 15) || ((in.colors_vs * sizeof(float)) & 15))
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "colors_precomp must be 16-byte aligned");
    in.gx = (in.W + TILE_X - 1) / TILE_X;
    in.gy = (in.H + TILE_Y - 1) / TILE_Y;
    in.tiles = in.gx * in.gy;
    guava_layout L;
    compute_layout(in.V, in.P, in.W, in.H, 0, false, &L);
    GeomPtrs g;
    g.rec = (SplatRec*)(geom + L.splat_rec);
    g.depths = (float*)(geom + L.depths);
    g.rects = (ushort4*)(geom + L.rects);
    g.tiles_touched = (uint32_t*)(geom + L.tiles_touched);
    g.grad_geo = (float*)(geom + L.grad_geo);
    ImagePtrs im;
    im.final_T = (float*)(img + L.final_T);
    im.n_contrib = (uint32_t*)(img + L.n_contrib);
    im.ranges = (uint2*)(img + L.ranges);
    im.tile_cursor = (uint32_t*)(img + L.tile_cursor);
    im.tile_order = (uint4*)(img + L.tile_order);
    im.num_rendered = (uint32_t*)(img + L.num_rendered);
    BinningPtrs b;
    b.point_list = (uint32_t*)(bin + L.point_list);   // offset 0 for every capacity
    b.keys_unsorted = nullptr;
    b.keys_sorted = nullptr;
    b.taken_marks = nullptr;   // located by the kernel: behind point_list, whose length (the capacity) is in the header
    b.capacity = 0;

    StageTimer timer(stream);
    timer.begin(5);
    // GeomPtrs::grad_geo is all zeros here: the forward zero-filled it under its tile scan and every backward's
    // preprocess kernel re-zeroes the records it has consumed (a memset node per backward until round 2).
    launch_render_backward(in, g, im, b, dL_dpix, dL_invdepths, dL_dcolor, sum ? sum->rs_color : (size_t)NUM_CHANNELS,
                           sum ? 0 : (size_t)in.P * NUM_CHANNELS, block_masks, g_opt_tma_staging.load() != 0, stream);
    timer.end();
    GUAVA_STAGE("render_backward", debug);
    timer.begin(6);
    launch_preprocess_backward(in, g, im.num_rendered, radii, dL_invdepths != nullptr, dL_dmean2D, dL_dconic, dL_dinvdepth,
                               dL_dopacity, dL_dmean3D, dL_dcov3D, dL_dscale, dL_drot, sum, stream);
    timer.end();
    GUAVA_STAGE("preprocess_backward", debug);
    return GUAVA_OK;
}

}  // namespace guava

using namespace guava;

extern "C" {

int guava_rasterizer_forward(
    guava_alloc_fn geometryBuffer, void* geometry_user, guava_alloc_fn binningBuffer, void* binning_user,
    guava_alloc_fn imageBuffer, void* image_user, int P, int D, int M, const float* background, int width,
    int height, const float* means3D, const float* shs, const float* colors_precomp, const float* opacities,
    const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
    const float* viewmatrix, const float* projmatrix, const float* cam_pos, float tan_fovx, float tan_fovy,
    int prefiltered, float* out_color, float* out_invdepth, int antialiasing, int* radii, int debug,
    void* stream, int* num_rendered)
{
    (void)D; (void)M; (void)cam_pos;
    if (shs != nullptr)
        return fail(GUAVA_ERR_UNSUPPORTED, "spherical harmonics are not supported by the 32-channel rasterizer");
    if (!geometryBuffer || !binningBuffer || !imageBuffer)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "null allocator");
    Inputs in{};
    in.V = 1; in.P = P; in.W = width; in.H = height;
    in.means3D = means3D; in.colors = colors_precomp; in.opacities = opacities;
    in.scales = scales; in.rotations = rotations; in.cov3D_precomp = cov3D_precomp;
    in.scale_modifier = scale_modifier;
    in.viewmatrix = viewmatrix; in.projmatrix = projmatrix;
    in.tan_fov_dev = nullptr; in.tan_fovx = tan_fovx; in.tan_fovy = tan_fovy;
    in.background = background; in.antialiasing = antialiasing; in.prefiltered = prefiltered;
    return forward_impl(geometryBuffer, geometry_user, binningBuffer, binning_user, imageBuffer, image_user, in,
                        out_color, out_invdepth, radii, debug, (cudaStream_t)stream, num_rendered, nullptr);
}

int guava_rasterizer_backward(
    int P, int D, int M, int R, const float* background, int width, int height, const float* means3D,
    const float* shs, const float* colors_precomp, const float* opacities, const float* scales,
    float scale_modifier, const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
    const float* projmatrix, const float* campos, float tan_fovx, float tan_fovy, const int* radii,
    char* geom_buffer, char* binning_buffer, char* image_buffer, const float* dL_dpix,
    const float* dL_invdepths, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
    float* dL_dinvdepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
    int antialiasing, int debug, void* stream)
{
    (void)D; (void)M; (void)R; (void)campos; (void)dL_dsh;
    if (shs != nullptr)
        return fail(GUAVA_ERR_UNSUPPORTED, "spherical harmonics are not supported by the 32-channel rasterizer");
    Inputs in{};
    in.V = 1; in.P = P; in.W = width; in.H = height;
    in.means3D = means3D; in.colors = colors_precomp; in.opacities = opacities;
    in.scales = scales; in.rotations = rotations; in.cov3D_precomp = cov3D_precomp;
    in.scale_modifier = scale_modifier;
    in.viewmatrix = viewmatrix; in.projmatrix = projmatrix;
    in.tan_fov_dev = nullptr; in.tan_fovx = tan_fovx; in.tan_fovy = tan_fovy;
    in.background = background; in.antialiasing = antialiasing; in.prefiltered = 0;
    if (colors_precomp == nullptr)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "For non-RGB, provide precomputed Gaussian colors!");
    return backward_impl(in, radii, geom_buffer, binning_buffer, image_buffer, dL_dpix, dL_invdepths, dL_dmean2D,
                         dL_dconic, dL_dopacity, dL_dcolor, dL_dinvdepth, dL_dmean3D, dL_dcov3D, dL_dscale,
                         dL_drot, debug, (cudaStream_t)stream);
}

int guava_rasterizer_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                                  uint8_t* present, void* stream)
{
    (void)projmatrix;
    if (P < 0) return fail(GUAVA_ERR_INVALID_ARGUMENT, "negative size");
    if (P == 0) return GUAVA_OK;
    if (!means3D || !viewmatrix || !present) return fail(GUAVA_ERR_INVALID_ARGUMENT, "null required pointer");
    launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(GUAVA_ERR_CUDA, cudaGetErrorString(e));
    return GUAVA_OK;
}

int guava_rasterizer_forward_batch(
    guava_alloc_fn geometryBuffer, void* geometry_user, guava_alloc_fn binningBuffer, void* binning_user,
    guava_alloc_fn imageBuffer, void* image_user, int V, int P, const float* background, int width, int height,
    const float* means3D, size_t means3D_view_stride, const float* colors_precomp, size_t colors_view_stride,
    const float* opacities, size_t opacities_view_stride, const float* scales, size_t scales_view_stride,
    float scale_modifier, const float* rotations, size_t rotations_view_stride, const float* viewmatrix,
    const float* projmatrix, const float* tan_fov, float* out_color, float* out_invdepth, int antialiasing,
    int* radii, int debug, void* stream, int* num_rendered, const guava_call_options* options)
{
    if (!geometryBuffer || !binningBuffer || !imageBuffer)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "null allocator");
    if (tan_fov == nullptr) return fail(GUAVA_ERR_INVALID_ARGUMENT, "tan_fov must be a device array [V,2]");
    Inputs in{};
    in.V = V; in.P = P; in.W = width; in.H = height;
    in.means3D = means3D; in.means3D_vs = means3D_view_stride;
    in.colors = colors_precomp; in.colors_vs = colors_view_stride;
    in.opacities = opacities; in.opacities_vs = opacities_view_stride;
    in.scales = scales; in.scales_vs = scales_view_stride;
    in.rotations = rotations; in.rotations_vs = rotations_view_stride;
    in.cov3D_precomp = nullptr;
    in.scale_modifier = scale_modifier;
    in.viewmatrix = viewmatrix; in.projmatrix = projmatrix;
    in.tan_fov_dev = tan_fov; in.tan_fovx = 0.f; in.tan_fovy = 0.f;
    in.background = background; in.antialiasing = antialiasing; in.prefiltered = 0;
    return forward_impl(geometryBuffer, geometry_user, binningBuffer, binning_user, imageBuffer, image_user, in,
                        out_color, out_invdepth, radii, debug, (cudaStream_t)stream, num_rendered, options);
}

int guava_rasterizer_backward_batch(
    int V, int P, int R, const float* background, int width, int height, const float* means3D,
    size_t means3D_view_stride, const float* colors_precomp, size_t colors_view_stride, const float* opacities,
    size_t opacities_view_stride, const float* scales, size_t scales_view_stride, float scale_modifier,
    const float* rotations, size_t rotations_view_stride, const float* viewmatrix, const float* projmatrix,
    const float* tan_fov, const int* radii, char* geom_buffer, char* binning_buffer, char* image_buffer,
    const float* dL_dpix, const float* dL_invdepths, float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
    float* dL_dcolor, float* dL_dinvdepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dscale, float* dL_drot,
    int antialiasing, int debug, void* stream)
{
    (void)R;
    if (tan_fov == nullptr) return fail(GUAVA_ERR_INVALID_ARGUMENT, "tan_fov must be a device array [V,2]");
    if (colors_precomp == nullptr)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "For non-RGB, provide precomputed Gaussian colors!");
    Inputs in{};
    in.V = V; in.P = P; in.W = width; in.H = height;
    in.means3D = means3D; in.means3D_vs = means3D_view_stride;
    in.colors = colors_precomp; in.colors_vs = colors_view_stride;
    in.opacities = opacities; in.opacities_vs = opacities_view_stride;
    in.scales = scales; in.scales_vs = scales_view_stride;
    in.rotations = rotations; in.rotations_vs = rotations_view_stride;
    in.cov3D_precomp = nullptr;
    in.scale_modifier = scale_modifier;
    in.viewmatrix = viewmatrix; in.projmatrix = projmatrix;
    in.tan_fov_dev = tan_fov;
    in.background = background; in.antialiasing = antialiasing; in.prefiltered = 0;
    return backward_impl(in, radii, geom_buffer, binning_buffer, image_buffer, dL_dpix, dL_invdepths, dL_dmean2D,
                         dL_dconic, dL_dopacity, dL_dcolor, dL_dinvdepth, dL_dmean3D, dL_dcov3D, dL_dscale,
                         dL_drot, debug, (cudaStream_t)stream);
}

int guava_rasterizer_backward_batch_summed(
    int V, int P, int R, const float* background, int width, int height, const float* means3D,
    size_t means3D_view_stride, const float* colors_precomp, size_t colors_view_stride, const float* opacities,
    size_t opacities_view_stride, const float* scales, size_t scales_view_stride, float scale_modifier,
    const float* rotations, size_t rotations_view_stride, const float* viewmatrix, const float* projmatrix,
    const float* tan_fov, const int* radii, char* geom_buffer, char* binning_buffer, char* image_buffer,
    const float* dL_dpix, const float* dL_invdepths, float* dL_dmean3D, float* dL_dscale, float* dL_drot,
    float* dL_dopacity, float* dL_dcolor, size_t out_row_stride, int antialiasing, int debug, void* stream)
{
    (void)R;
    if (tan_fov == nullptr) return fail(GUAVA_ERR_INVALID_ARGUMENT, "tan_fov must be a device array [V,2]");
    if (colors_precomp == nullptr)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "For non-RGB, provide precomputed Gaussian colors!");
    if (scales == nullptr || rotations == nullptr)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "the batched entry takes scales and rotations");
    if (out_row_stride != 0 && out_row_stride < 43)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "out_row_stride is 0 (dense [P,w] outputs) or at least 43 floats");
    if (out_row_stride > 0xffffffffull) return fail(GUAVA_ERR_INVALID_ARGUMENT, "out_row_stride too large");
    Inputs in{};
    in.V = V; in.P = P; in.W = width; in.H = height;
    in.means3D = means3D; in.means3D_vs = means3D_view_stride;
    in.colors = colors_precomp; in.colors_vs = colors_view_stride;
    in.opacities = opacities; in.opacities_vs = opacities_view_stride;
    in.scales = scales; in.scales_vs = scales_view_stride;
    in.rotations = rotations; in.rotations_vs = rotations_view_stride;
    in.cov3D_precomp = nullptr;
    in.scale_modifier = scale_modifier;
    in.viewmatrix = viewmatrix; in.projmatrix = projmatrix;
    in.tan_fov_dev = tan_fov;
    in.background = background; in.antialiasing = antialiasing; in.prefiltered = 0;
    const uint32_t rs = (uint32_t)out_row_stride;
    const SummedRows sum = rs ? SummedRows{rs, rs, rs, rs, rs, -1} : SummedRows{3u, 3u, 4u, 1u, (uint32_t)NUM_CHANNELS, -1};
    return backward_impl(in, radii, geom_buffer, binning_buffer, image_buffer, dL_dpix, dL_invdepths, nullptr, nullptr,
                         dL_dopacity, dL_dcolor, nullptr, dL_dmean3D, nullptr, dL_dscale, dL_drot, debug,
                         (cudaStream_t)stream, &sum);
}

int guava_pack_view_gradients(int V, int P, const float* dL_dmeans3D, const float* dL_dscales,
                              const float* dL_drotations, const float* dL_dopacity, const float* dL_dcolors,
                              float* packed, void* stream)
{
    if (V < 0 || P < 0) return fail(GUAVA_ERR_INVALID_ARGUMENT, "negative size");
    if (P == 0) return GUAVA_OK;
    if (!dL_dmeans3D || !dL_dscales || !dL_drotations || !dL_dopacity || !dL_dcolors || !packed)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "null required pointer");
    const size_t want = ((size_t)P + 31) / 32;                     // 8 warps x 4 rows per block and pass
    const size_t full = (size_t)sm_count() * 8u;
    const unsigned grid = (unsigned)(want < full ? want : full);
    cudaStream_t st = (cudaStream_t)stream;
#define GUAVA_PACK(VT) pack_view_gradients_kernel<VT><<<grid, 256, 0, st>>>(V, P, dL_dmeans3D, dL_dscales, dL_drotations, \
                                                                            dL_dopacity, dL_dcolors, packed)
    if (V == 1) GUAVA_PACK(1);
    else if (V == 2) GUAVA_PACK(2);
    else if (V == 4) GUAVA_PACK(4);
    else if (V == 8) GUAVA_PACK(8);
    else GUAVA_PACK(0);
#undef GUAVA_PACK
    count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(GUAVA_ERR_CUDA, cudaGetErrorString(e));
    return GUAVA_OK;
}

int guava_image_dot(const float* a, const float* b, size_t n, float* out64, void* stream)
{
    if (!a || !b || !out64) return fail(GUAVA_ERR_INVALID_ARGUMENT, "null required pointer");
    if (((uintptr_t)a & 15) || ((uintptr_t)b & 15)) return fail(GUAVA_ERR_INVALID_ARGUMENT, "inputs must be 16-byte aligned");
    static const int ctas = [] { const char* e = getenv("GUAVA_DOT_CTAS"); int c = e ? atoi(e) : IMAGE_DOT_CTAS; return c < 1 ? 1 : (c > 62 ? 62 : c); }();   // tuning knob
    image_dot_kernel<<<ctas, 512, 0, (cudaStream_t)stream>>>(a, b, n, out64);
    count_launch();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(GUAVA_ERR_CUDA, cudaGetErrorString(e));
    return GUAVA_OK;
}

int guava_buffer_layout(int V, int P, int width, int height, size_t capacity, guava_layout* out)
{
    if (!out || V < 0 || P < 0 || width <= 0 || height <= 0) return fail(GUAVA_ERR_INVALID_ARGUMENT, "bad layout query");
    compute_layout(V, P, width, height, capacity, g_opt_emit_sorted_keys.load() != 0, out, (int)g_opt_sort_mode.load());
    return GUAVA_OK;
}

int guava_buffer_layout_of_call(const char* image_buffer, int V, int P, int width, int height, void* stream,
                                guava_layout* out, int* sort_mode, int* block_masks, size_t* capacity)
{
    if (!out || !image_buffer || V < 0 || P < 0 || width <= 0 || height <= 0)
        return fail(GUAVA_ERR_INVALID_ARGUMENT, "bad layout query");
    guava_layout L0;
    compute_layout(V, P, width, height, 0, false, &L0, 0);   // the image buffer's layout does not depend on the rest
    uint32_t hdr[8];
    GUAVA_CUDA(cudaMemcpyAsync(hdr, image_buffer + L0.num_rendered, sizeof(hdr), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    GUAVA_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    const int mode = (int)(hdr[5] & 0xffu);
    const size_t cap = (size_t)hdr[6] | ((size_t)hdr[7] << 32);
    // the sorted keys exist in mode 1 always, in mode 0 when the call ran with "emit_sorted_keys" (header bit 16)
    compute_layout(V, P, width, height, cap, (hdr[5] >> 16) & 1u, out, mode, ((hdr[5] >> 8) & 1u) != 0);
    if (sort_mode) *sort_mode = mode;
    if (block_masks) *block_masks = (int)((hdr[5] >> 8) & 1u);
    if (capacity) *capacity = cap;
    return GUAVA_OK;
}

int guava_selftest_cull(int n, const float* splats, const float* block_origins, uint8_t* rejected, uint8_t* passing,
                        void* stream)
{
    if (n < 0) return fail(GUAVA_ERR_INVALID_ARGUMENT, "negative size");
    if (n == 0) return GUAVA_OK;
    if (!splats || !block_origins || !rejected || !passing) return fail(GUAVA_ERR_INVALID_ARGUMENT, "null required pointer");
    launch_selftest_cull(n, splats, block_origins, rejected, passing, (cudaStream_t)stream);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(GUAVA_ERR_CUDA, cudaGetErrorString(e));
    return GUAVA_OK;
}

static std::atomic<long long>* find_option(const char* name)
{
    if (!name) return nullptr;
    if (!strcmp(name, "emit_sorted_keys")) return &g_opt_emit_sorted_keys;
    if (!strcmp(name, "exact_binning")) return &g_opt_exact_binning;
    if (!strcmp(name, "capacity_hint")) return &g_opt_capacity_hint;
    if (!strcmp(name, "profile_stages")) return &g_opt_profile_stages;
    if (!strcmp(name, "tma_staging")) return &g_opt_tma_staging;
    if (!strcmp(name, "sort_mode")) return &g_opt_sort_mode;
    if (!strcmp(name, "taken_masks")) return &g_opt_taken_masks;
    if (!strcmp(name, "os_identify_ranges")) return &g_opt_os_identify_ranges;
    return nullptr;
}

int guava_set_option(const char* name, long long value)
{
    auto* o = find_option(name);
    if (!o) return fail(GUAVA_ERR_INVALID_ARGUMENT, "unknown option");
    o->store(value);
    return GUAVA_OK;
}

long long guava_get_option(const char* name)
{
    auto* o = find_option(name);
    return o ? o->load() : -1;
}

unsigned long long guava_launch_count(void) { return g_launches.load(); }

int guava_stage_time(const char* stage, double* total_ms, unsigned long long* calls)
{
    if (!stage) return fail(GUAVA_ERR_INVALID_ARGUMENT, "null stage name");
    for (int i = 0; i < kNumStages; i++)
        if (!strcmp(stage, kStageNames[i])) {
            std::lock_guard<std::mutex> lk(g_stage_mu);
            if (total_ms) *total_ms = g_stage_ms[i];
            if (calls) *calls = g_stage_calls[i];
            return GUAVA_OK;
        }
    return fail(GUAVA_ERR_INVALID_ARGUMENT, "unknown stage");
}

void guava_stage_time_reset(void)
{
    std::lock_guard<std::mutex> lk(g_stage_mu);
    for (int i = 0; i < kNumStages; i++) { g_stage_ms[i] = 0.0; g_stage_calls[i] = 0; }
}

const char* guava_last_error(void) { return t_last_error.c_str(); }

const char* guava_version(void) { return "guava_rasterizer_b200 0.1 (sm_100a)"; }

}  // extern "C"
