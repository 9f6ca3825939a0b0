/*
 * guava_rasterizer.h -- C ABI of the B200-native 32-channel Gaussian-splatting
 * rasterizer (libguava_rasterizer_b200.so).
 *
 * Drop-in boundary.  Every entry point replaces one member of the reference's
 * C++ library interface `CudaRasterizer::Rasterizer`
 *   (/root/reference/submodules/diff-gaussian-rasterization-32/
 *    cuda_rasterizer/rasterizer.h:24-90)
 * which the reference's torch binding (rasterize_points.cu:35-244) calls.  The
 * argument order and meaning follow that interface; the differences are the
 * ones a C ABI forces:
 *   - `std::function<char*(size_t)>` allocators become a function pointer plus
 *     an opaque user pointer (same contract: "resize to N bytes, return the
 *     device pointer"; may be called more than once per buffer);
 *   - the return value is a status code (0 = ok); `num_rendered` is an out
 *     parameter; C++ exceptions become codes + guava_last_error();
 *   - a `stream` argument (cudaStream_t as void*).  The reference launches on
 *     the legacy default stream; pass NULL for that behaviour;
 *   - the three opaque buffers (geometry / binning / image) keep their role
 *     (state handed from forward to backward) but have this library's own
 *     layout, described by guava_buffer_layout().
 * No torch types appear here: plain pointers and sizes only.  All pointers are
 * device pointers unless stated otherwise; all tensors are fp32, contiguous,
 * in the reference's layouts (means3D [P,3], colors_precomp [P,32], opacities
 * [P], scales [P,3], rotations [P,4] wxyz *not normalised by the kernel*,
 * cov3D_precomp [P,6], viewmatrix / projmatrix [16] transposed as in
 * auxiliary.h:70-89, background [32], out_color [32,H,W], out_invdepth
 * [1,H,W], radii [P] int32).  `colors_precomp` must be 16-byte aligned (rows
 * are fetched with 16-byte asynchronous copies).
 */
#ifndef GUAVA_RASTERIZER_H_INCLUDED
#define GUAVA_RASTERIZER_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GUAVA_NUM_CHANNELS 32 /* cuda_rasterizer/config.h:15 */
#define GUAVA_BLOCK_X 16      /* config.h:16 */
#define GUAVA_BLOCK_Y 16      /* config.h:17 */

/* status codes */
#define GUAVA_OK 0
#define GUAVA_ERR_INVALID_ARGUMENT 1 /* e.g. colors_precomp == NULL: the reference throws
                                        "For non-RGB, provide precomputed Gaussian colors!"
                                        (rasterizer_impl.cu:244-247) */
#define GUAVA_ERR_CUDA 2             /* a CUDA call / kernel failed; text in guava_last_error() */
#define GUAVA_ERR_ALLOC 3            /* an allocator callback returned NULL */
#define GUAVA_ERR_UNSUPPORTED 4      /* spherical harmonics path (unreachable for 32 channels) */

/* Allocator callback: make the buffer `bytes` long and return its device
 * address (reference: resizeFunctional, rasterize_points.cu:27-33). */
typedef char *(*guava_alloc_fn)(void *user, size_t bytes);

/* Replaces CudaRasterizer::Rasterizer::forward (rasterizer.h:33-59,
 * rasterizer_impl.cu:198-340).  D, M, shs, cam_pos and prefiltered are accepted
 * for signature parity; shs must be NULL (32-channel build). */
int guava_rasterizer_forward(
    guava_alloc_fn geometryBuffer, void *geometry_user,
    guava_alloc_fn binningBuffer, void *binning_user,
    guava_alloc_fn imageBuffer, void *image_user,
    int P, int D, int M,
    const float *background,
    int width, int height,
    const float *means3D,
    const float *shs,
    const float *colors_precomp,
    const float *opacities,
    const float *scales,
    float scale_modifier,
    const float *rotations,
    const float *cov3D_precomp,
    const float *viewmatrix,
    const float *projmatrix,
    const float *cam_pos,
    float tan_fovx, float tan_fovy,
    int prefiltered,
    float *out_color,
    float *out_invdepth,
    int antialiasing,
    int *radii,
    int debug,
    void *stream,
    int *num_rendered);

/* Replaces CudaRasterizer::Rasterizer::backward (rasterizer.h:61-89,
 * rasterizer_impl.cu:345-450).  R is the num_rendered returned by forward.
 * dL_dmean2D [P,3], dL_dconic [P,4], dL_dopacity [P], dL_dcolor [P,32],
 * dL_dinvdepth [P] or NULL must be zero-filled by the caller (as
 * rasterize_points.cu:163-182 does); dL_dmean3D [P,3], dL_dcov3D [P,6],
 * dL_dscale [P,3], dL_drot [P,4] are fully written for visible Gaussians and
 * must be zero-filled for the rest.  dL_drot must be 16-byte aligned (one vector store per Gaussian). */
int guava_rasterizer_backward(
    int P, int D, int M, int R,
    const float *background,
    int width, int height,
    const float *means3D,
    const float *shs,
    const float *colors_precomp,
    const float *opacities,
    const float *scales,
    float scale_modifier,
    const float *rotations,
    const float *cov3D_precomp,
    const float *viewmatrix,
    const float *projmatrix,
    const float *campos,
    float tan_fovx, float tan_fovy,
    const int *radii,
    char *geom_buffer,
    char *binning_buffer,
    char *image_buffer,
    const float *dL_dpix,
    const float *dL_invdepths,
    float *dL_dmean2D,
    float *dL_dconic,
    float *dL_dopacity,
    float *dL_dcolor,
    float *dL_dinvdepth,
    float *dL_dmean3D,
    float *dL_dcov3D,
    float *dL_dsh,
    float *dL_dscale,
    float *dL_drot,
    int antialiasing,
    int debug,
    void *stream);

/* Replaces CudaRasterizer::Rasterizer::markVisible (rasterizer.h:27-31,
 * rasterizer_impl.cu:141-153): present[i] = (view-space z > 0.2). */
int guava_rasterizer_mark_visible(
    int P,
    const float *means3D,
    const float *viewmatrix,
    const float *projmatrix,
    uint8_t *present,
    void *stream);

/* Per-call options of the batched entry (everything that used to be "applies to the next call" state
 * lives here: a call's behaviour depends on its own arguments only).  Zero-initialise, set
 * `struct_size = sizeof(guava_call_options)`, fill what is needed; NULL = all defaults. */
typedef struct guava_call_options {
    size_t struct_size;
    /* Refiner hand-off (models/UbodyAvatar/gaussian_render.py:70-73 feeds the
     * stacked 32-channel image to the neural refiner): when non-NULL the compositing kernel's epilogue
     * additionally writes the image as channels-last bf16, [V,H,W,32] (64 bytes per pixel,
     * round-to-nearest-even of the fp32 values; device pointer, 16-byte aligned) -- no separate
     * permute/cast pass.  Where V*P == 0 it is not written (as out_color, rasterize_points.cu:88). */
    void *refiner_output_nhwc_bf16;
    /* Status words.  The reference returns num_rendered through a blocking device->host copy per view
     * (rasterizer_impl.cu:284).  Here the tile scan stores two words straight into pinned host memory:
     * [0] = num_rendered, [1] = longest tile list | (overflow << 31).  NULL: a slot the library owns.
     * Non-NULL: the caller's OWN two words -- 8-byte aligned, pinned, device-mapped host memory (e.g.
     * cudaHostAlloc) -- which it reads after synchronising: no copy node, no copy engine, nothing on
     * the stream after the last kernel (what a captured CUDA graph needs). */
    void *status_words;
    /* 1: enqueue the whole chain and return without waiting for num_rendered on the host (CUDA-graph
     * capture): *num_rendered is set to -1, and a binning overflow (capacity, or a tile list longer than
     * the sort that was launched) makes every later kernel of the forward AND of the backward skip
     * itself; it is visible in status_words[1] bit 31 / the image buffer's num_rendered[1]. */
    int no_host_wait;
    /* 1: no backward call will follow this forward (inference): the forward compositor does not record its
     * "taken" marks (option "taken_masks") and the binning buffer does not hold them -- 2 % of the forward and
     * 8 bytes per instance saved.  A backward call after such a forward is still correct, only slower. */
    int forward_only;
    /* When non-NULL: a device buffer of `zero_fill_bytes` bytes (both multiples of 4) that the forward zero-fills on
     * its stream -- meant for the output of guava_rasterizer_backward_batch_summed, which must start at zero.  The
     * fill costs little or no time: it runs, in the tile scan's launch, on the SMs the scan leaves idle (frames of more
     * than ~100 k tiles get a cudaMemsetAsync). */
    void *zero_fill;
    size_t zero_fill_bytes;
} guava_call_options;

/* Batched multi-view entry (replaces the per-view Python
 * loop of models/UbodyAvatar/gaussian_render.py:37-67).  V views of equal
 * H x W are rasterized by ONE launch chain: tile ids and keys carry the view
 * index, so per-view slices of every intermediate are identical to V
 * single-view calls.  Per-view tensors are [V, ...] contiguous; a
 * `*_view_stride` of 0 (in floats) shares one Gaussian tensor across all views.
 * tan_fov is a device array [V][2] = (tan_fovx, tan_fovy). */
int guava_rasterizer_forward_batch(
    guava_alloc_fn geometryBuffer, void *geometry_user,
    guava_alloc_fn binningBuffer, void *binning_user,
    guava_alloc_fn imageBuffer, void *image_user,
    int V, int P,
    const float *background,            /* [V,32] */
    int width, int height,
    const float *means3D, size_t means3D_view_stride,
    const float *colors_precomp, size_t colors_view_stride,
    const float *opacities, size_t opacities_view_stride,
    const float *scales, size_t scales_view_stride,
    float scale_modifier,
    const float *rotations, size_t rotations_view_stride,
    const float *viewmatrix,            /* [V,16] */
    const float *projmatrix,            /* [V,16] */
    const float *tan_fov,               /* [V,2] device */
    float *out_color,                   /* [V,32,H,W] */
    float *out_invdepth,                /* [V,1,H,W] */
    int antialiasing,
    int *radii,                         /* [V,P] */
    int debug,
    void *stream,
    int *num_rendered,                  /* total over views */
    const guava_call_options *options); /* NULL = defaults */

/* Backward of the batched entry.  Gradient tensors are [V,P,...] (one slice
 * per view; the caller sums over views when the Gaussians were shared). */
int guava_rasterizer_backward_batch(
    int V, int P, int R,
    const float *background,
    int width, int height,
    const float *means3D, size_t means3D_view_stride,
    const float *colors_precomp, size_t colors_view_stride,
    const float *opacities, size_t opacities_view_stride,
    const float *scales, size_t scales_view_stride,
    float scale_modifier,
    const float *rotations, size_t rotations_view_stride,
    const float *viewmatrix,
    const float *projmatrix,
    const float *tan_fov,
    const int *radii,
    char *geom_buffer,
    char *binning_buffer,
    char *image_buffer,
    const float *dL_dpix,               /* [V,32,H,W] */
    const float *dL_invdepths,          /* [V,1,H,W] or NULL */
    float *dL_dmean2D,                  /* [V,P,3] zero-filled */
    float *dL_dconic,                   /* [V,P,4] zero-filled */
    float *dL_dopacity,                 /* [V,P]   zero-filled */
    float *dL_dcolor,                   /* [V,P,32] zero-filled */
    float *dL_dinvdepth,                /* [V,P] zero-filled or NULL */
    float *dL_dmean3D,                  /* [V,P,3] zero-filled */
    float *dL_dcov3D,                   /* [V,P,6] zero-filled */
    float *dL_dscale,                   /* [V,P,3] zero-filled */
    float *dL_drot,                     /* [V,P,4] zero-filled */
    int antialiasing,
    int debug,
    void *stream);

/* Backward of the batched entry with the sum over the views taken inside the kernels ("one
 * all-reduce over the concatenated per-Gaussian parameter gradients [P,43] ... written directly by the
 * backward-preprocess kernel"; mirrors what the reference trainer's DDP step sums, main/trainer.py:40-43,95).
 * Each output row i receives  sum over v of  dL/d(parameter i of view v)  -- for tensors shared by all views
 * (view stride 0) that IS the gradient of the shared tensor, what the per-view entry leaves to the caller.  The two
 * kernels add with atomics into the caller's ZERO-FILLED outputs; the per-view outputs the reference's binding keeps
 * (mean2D, conic, cov3D, inverse depth: rasterize_points.cu:163-172) are not produced.
 * out_row_stride = 0: five dense tensors [P,3] [P,3] [P,4] [P,1] [P,32].  out_row_stride = s >= 43: the five
 * pointers address columns of rows s floats apart -- one packed [P,43] buffer is (buf, buf+3, buf+6, buf+10,
 * buf+11, 43), the layout of guava_pack_view_gradients, without that kernel and without the [V,P,43] intermediates. */
int guava_rasterizer_backward_batch_summed(
    int V, int P, int R,
    const float *background,
    int width, int height,
    const float *means3D, size_t means3D_view_stride,
    const float *colors_precomp, size_t colors_view_stride,
    const float *opacities, size_t opacities_view_stride,
    const float *scales, size_t scales_view_stride,
    float scale_modifier,
    const float *rotations, size_t rotations_view_stride,
    const float *viewmatrix,
    const float *projmatrix,
    const float *tan_fov,
    const int *radii,
    char *geom_buffer,
    char *binning_buffer,
    char *image_buffer,
    const float *dL_dpix,               /* [V,32,H,W] */
    const float *dL_invdepths,          /* [V,1,H,W] or NULL */
    float *dL_dmean3D,                  /* rows of 3, zero-filled */
    float *dL_dscale,                   /* rows of 3, zero-filled */
    float *dL_drot,                     /* rows of 4, zero-filled */
    float *dL_dopacity,                 /* rows of 1, zero-filled */
    float *dL_dcolor,                   /* rows of 32, zero-filled */
    size_t out_row_stride,              /* 0 = dense, else floats between rows (>= 43) */
    int antialiasing,
    int debug,
    void *stream);

/* Training-side helper: sum the per-view per-Gaussian parameter gradients over the
 * V local views and pack them as one [P,43] fp32 row per Gaussian = (means3D 3, scales 3, rotations 4,
 * opacity 1, features 32) -- the payload of the one all-reduce that mirrors the reference trainer's DDP
 * step (main/trainer.py:40-43,95).  Inputs are [V,P,w] contiguous. */
int guava_pack_view_gradients(
    int V, int P,
    const float *dL_dmeans3D, const float *dL_dscales, const float *dL_drotations,
    const float *dL_dopacity, const float *dL_dcolors,
    float *packed,                      /* [P,43] */
    void *stream);

/* Byte offsets of the arrays inside the three opaque buffers, so that tests
 * and tools can inspect the intermediates the reference keeps in
 * GeometryState / BinningState / ImageState (rasterizer_impl.h:33-66). */
typedef struct guava_layout {
    /* geometry buffer (per Gaussian instance g = view*P + i) */
    size_t geom_bytes;
    size_t splat_rec;      /* float[8] per g: mean2D.x, mean2D.y, conic.x, conic.y, conic.z,
                              opacity*h_scale, 1/depth (rcp.rn), ln(255*opacity*h_scale)+0.01 (culling threshold) */
    size_t depths;         /* float per g (view-space z) */
    size_t rects;          /* uint16[4] per g: tile rect min.x, min.y, max.x, max.y */
    size_t tiles_touched;  /* uint32 per g */
    size_t grad_geo;       /* float[8] per g, backward scratch (all zeros between calls: zero-filled by the forward
                              under its tile scan, re-zeroed by each backward once consumed): sums over
                              contributing pixels of s*dx, s*dy, s*dx^2, s*dx*dy, s*dy^2 (s = dL/dG * G,
                              d = mean2D - pixel), G*dL/dalpha, w*dL/dinvdepth, 0 */
    /* image buffer */
    size_t image_bytes;
    size_t final_T;        /* float per pixel  [V*H*W] */
    size_t n_contrib;      /* uint32 per pixel [V*H*W] */
    size_t ranges;         /* uint2 per tile   [V*tiles]; values are offsets into point_list */
    size_t tile_cursor;    /* uint32 per tile, one per 128-byte line (scratch) */
    size_t tile_order;     /* uint32[4] per tile, longest list first (work order): tile index, range.x, range.y, 0 */
    size_t num_rendered;   /* uint32[8], the call's header: R, overflow flag, two work counters, longest tile list,
                              sort mode | block_masks << 8 | sorted keys kept << 16 (as resolved for this call),
                              binning capacity (lo, hi) */
    /* binning buffer for a capacity of `capacity` instances */
    size_t binning_bytes;
    size_t point_list;     /* uint32 per instance: Gaussian index within its view, sorted */
    size_t keys_unsorted;  /* uint64 per instance: depth_bits<<32 | index (tile-bucketed) */
    size_t keys_sorted;    /* uint64 per instance: tile<<32 | depth_bits; only written when
                              guava_set_option("emit_sorted_keys", 1) or "sort_mode" 1 (there the high
                              word is view*tiles + tile) */
    /* only with guava_set_option("sort_mode", 1) (global onesweep sort), else zero-sized */
    size_t point_offsets;  /* geometry buffer: uint32 per g, inclusive sums of tiles_touched
                              (GeometryState::point_offsets, rasterizer_impl.cu:280) */
    size_t os_vals_alt;    /* binning buffer: uint32 per instance, alternate value buffer of the sort */
    size_t os_scratch;     /* binning buffer: tickets, scan state, digit histograms, look-back state */
    /* only with option "taken_masks" (default), else zero-sized */
    size_t taken_marks;    /* binning buffer, right behind point_list: 8 planes of (capacity rounded up to 256) bytes --
                              plane b, byte i != 0: 8x4 pixel block b of the tile of instance i got nothing from it in
                              the forward (the backward skips it) */
} guava_layout;

int guava_buffer_layout(int V, int P, int width, int height, size_t capacity, guava_layout *out);

/* The layout a finished forward call actually used.  Every forward call records what it resolved its
 * options to (sort mode, block masks, sorted keys kept, binning capacity) in the image buffer's header
 * (num_rendered[5..7]); this query reads the header back (one 32-byte device->host copy, synchronises
 * `stream`) and reports the offsets that are valid for THAT call, whatever the process-wide options are
 * by now.  sort_mode / block_masks / capacity may be NULL. */
int guava_buffer_layout_of_call(const char *image_buffer, int V, int P, int width, int height, void *stream,
                                guava_layout *out, int *sort_mode, int *block_masks, size_t *capacity);

/* Process-wide options (ints): "emit_sorted_keys" (0/1, default 0),
 * "exact_binning" (0/1, default 0: 1 = wait for num_rendered before sizing the
 * binning buffer, exactly as rasterizer_impl.cu:283-288 does),
 * "capacity_hint" (instances; 0 = automatic),
 * "profile_stages" (0/1, default 0: time every kernel with CUDA events, see
 * guava_stage_time),
 * "sort_mode" (0: per-tile buckets sorted in shared memory; 1: the reference's chain rebuilt --
 * chained scan + duplicateWithKeys + onesweep LSD radix sort of the 64-bit tile|depth keys +
 * identifyTileRanges, rasterizer_impl.cu:280-320; identical keys, values and ranges; -1 (default): 1 when
 * the previous call of the same shape produced a tile list longer than 4096 instances, else 0 -- the one place
 * where a call's kernel path depends on an earlier call; the results do not;
 * guava_buffer_layout reports the mode-0 layout for -1, whose offsets are valid in both modes),
 * "tma_staging" (0/1: the compositors fetch the survivors' 128-byte feature rows with TMA tensor copies --
 * cp.async.bulk.tensor.2d ... tile::gather4 over a [rows,32] tensor map of colors_precomp, four rows per
 * instruction from one elected lane, mbarrier completion -- instead of eight 16-byte cp.async per lane; same bytes,
 * same results; needs colors_view_stride % 32 == 0),
 * "os_identify_ranges" (0/1, default 0: sort mode 1 re-derives the tile ranges from the sorted keys --
 * identifyTileRanges, rasterizer_impl.cu:116-138, restated -- although the tile scan has already written them),
 * "taken_masks" (0/1, default 1: the forward compositor records, one byte per (list entry, 8x4 pixel block of
 * its tile), which entries the block got nothing from -- rejected by the footprint test or taken by no pixel --
 * and the backward compositor walks only the others; 0 = the compositors of round 1, kept for A/B.  The forward
 * notes in the call's header whether it marked, so changing the option between a forward call and its backward
 * call costs speed, not correctness).
 * Returns GUAVA_OK or GUAVA_ERR_INVALID_ARGUMENT. */
int guava_set_option(const char *name, long long value);
long long guava_get_option(const char *name);

/* Number of kernels this library has launched since load (all threads). */
unsigned long long guava_launch_count(void);

/* Accumulated device time of one stage while "profile_stages" is on.  Stages:
 * preprocess, tile_scan, scatter, tile_sort, render_forward, render_backward,
 * preprocess_backward. */
int guava_stage_time(const char *stage, double *total_ms, unsigned long long *calls);
void guava_stage_time_reset(void);

/* Training-side helper: out64[0] = sum_i a[i]*b[i] -- the value of a loss that is linear in the image
 * (loss = <out_color, dL_dout_color>, the form the benchmark step and the parity tests use; GUAVA's own
 * losses are computed by its trainer, main/trainer.py:60-95).  Runs on 16 CTAs so that it can share the
 * GPU with the backward compositing kernel on another stream.  Deterministic.  `out64` is 64 floats of
 * device memory owned by the caller: [0] result, [1..62] scratch, word [63] must be zero before the
 * first call (the kernel leaves it zero).  a, b: device pointers, 16-byte aligned, n elements. */
int guava_image_dot(const float *a, const float *b, size_t n, float *out64, void *stream);

/* Self-test hook for the compositors' warp-footprint culling (the one place where this library skips work the
 * reference does, forward.cu:342-382 / backward.cu:556-636 visit every list entry at every pixel).  For n samples
 * -- splats [n,6] = (mean2D.x, mean2D.y, conic.x, conic.y, conic.z, opacity), block_origins [n,2] = pixel
 * coordinates of an 8x4 block's first pixel -- the device evaluates the culling predicate exactly as the
 * compositors do (rejected[i] = 1: "cannot reach alpha >= 1/255 anywhere in the block") and, with the
 * compositors' own arithmetic, the reference's per-pixel alpha test at the block's 32 pixel centres
 * (passing[i] = number of pixels that pass).  Culling is output-neutral iff no sample has rejected && passing. */
int guava_selftest_cull(int n, const float *splats, const float *block_origins, uint8_t *rejected, uint8_t *passing,
                        void *stream);

/* ---- the step immediately before the rasterizer ------------------------
 * Fused deformation of GUAVA's two Gaussian groups into the concatenated [B, Nv+N, *] tensors the
 * rasterizer reads.  Replaces models/UbodyAvatar/ubody_gaussian.py:245-289 (Ubody_Gaussian.forward:
 * compute_face_orientation, utils/graphics_utils.py:61-80; roma 1.5.3 rotmat_to_unitquat / quat_product)
 * and its autograd backward for the per-Gaussian attributes.  All pointers are device pointers, fp32
 * unless noted; quaternions are wxyz; `face_frames` is a [B,F,16] float scratch written by forward and
 * read by backward (R row-major, scale, quat wxyz, 2 pad); `out_features` may be NULL (no concatenation).
 * Returns GUAVA_OK / GUAVA_ERR_INVALID_ARGUMENT / GUAVA_ERR_CUDA. */
int guava_deform_gaussians_forward(
    int B, int Nv, int N, int F,
    const float *vertices /* [B,Nv,3] deformed mesh */, const float *ver_transform /* [B,Nv,4,4] */,
    const int *faces /* [F,3] int32 */, const int *uv_binding_face /* [N] int32 */,
    const float *uv_face_bary /* [N,3] */,
    const float *smplx_rotation /* [B,Nv,4] */, const float *smplx_scaling /* [B,Nv,3] */,
    const float *smplx_opacity /* [B,Nv,1] */, const float *smplx_features /* [B,Nv,32] */,
    const float *uv_local_xyz /* [B,N,3] */, const float *uv_rotation /* [B,N,4] */,
    const float *uv_scaling /* [B,N,3] */, const float *uv_opacity /* [B,N,1] */,
    const float *uv_features /* [B,N,32] */,
    float *face_frames, float *out_xyz, float *out_rotation, float *out_scaling, float *out_opacity,
    float *out_features, void *stream);
int guava_deform_gaussians_backward(
    int B, int Nv, int N, int F, const float *ver_transform, const int *uv_binding_face,
    const float *smplx_rotation, const float *face_frames,
    const float *dL_dxyz /* [B,Nv+N,3] */, const float *dL_drotation /* [B,Nv+N,4] */,
    const float *dL_dscaling /* [B,Nv+N,3] */,
    float *dL_dsmplx_rotation /* [B,Nv,4] */, float *dL_duv_local_xyz /* [B,N,3] */,
    float *dL_duv_rotation /* [B,N,4] */, float *dL_duv_scaling /* [B,N,3] */, void *stream);

/* The backward of the forward's pass-through outputs (features = smplx ++ uv, opacity likewise, scaling's SMPLX part):
 * the rasterizer's gradients [B,Nv+N,w] split into contiguous per-attribute tensors in one pass.  Any input / output
 * pointer may be NULL (that gradient is not wanted).  No reference counterpart: there autograd slices the tensors. */
int guava_deform_split_gradients(
    int B, int Nv, int N,
    const float *dL_dscaling /* [B,Nv+N,3] */, const float *dL_dopacity /* [B,Nv+N,1] */,
    const float *dL_dfeatures /* [B,Nv+N,32] */,
    float *dL_dsmplx_scaling /* [B,Nv,3] */, float *dL_dsmplx_opacity /* [B,Nv,1] */,
    float *dL_dsmplx_features /* [B,Nv,32] */, float *dL_duv_opacity /* [B,N,1] */,
    float *dL_duv_features /* [B,N,32] */, void *stream);

/* No-allocation variant of guava_deform_gaussians_forward.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight
 * serving loops, harnesses that forbid allocation).  Same inputs, same outputs, same arithmetic as
 * guava_deform_gaussians_forward, but the `face_frames` scratch is not an argument: it lives in a file-scope
 * `static __device__ __align__(256)` array of the library (csrc/deform.cu), whose device address the library obtains
 * with cudaGetSymbolAddress once per device.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating
 * library routine): it validates, looks the address up and enqueues the same kernels on `stream`.
 * Limit: B * F <= GUAVA_DEFORM_STATIC_MAX_FACE_FRAMES (65536 face frames = 4 MiB, held by every GPU that loads the
 * library until the process exits; the SMPLX mesh has F = 20908, so B <= 3); beyond it the call returns
 * GUAVA_ERR_INVALID_ARGUMENT with the numbers in guava_last_error().
 * The array is shared by every call on the device and keeps the last call's contents: nothing in it has to start at
 * zero (each call writes all of its B*F frames before reading any), but two calls that overlap on the GPU --
 * different streams, or concurrent host threads -- RACE on it; serialise them (one stream, or events).  The frames
 * do not survive the next call, so there is no backward for this variant: training uses the regular entry point. */
#define GUAVA_DEFORM_STATIC_MAX_FACE_FRAMES 65536
int guava_deform_gaussians_forward_static(
    int B, int Nv, int N, int F,
    const float *vertices /* [B,Nv,3] */, const float *ver_transform /* [B,Nv,4,4] */,
    const int *faces /* [F,3] */, const int *uv_binding_face /* [N] */, const float *uv_face_bary /* [N,3] */,
    const float *smplx_rotation, const float *smplx_scaling, const float *smplx_opacity, const float *smplx_features,
    const float *uv_local_xyz, const float *uv_rotation, const float *uv_scaling, const float *uv_opacity,
    const float *uv_features,
    float *out_xyz, float *out_rotation, float *out_scaling, float *out_opacity, float *out_features, void *stream);

const char *guava_last_error(void);
const char *guava_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GUAVA_RASTERIZER_H_INCLUDED */
