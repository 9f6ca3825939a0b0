// Per-Gaussian kernels: forward preprocess (+ per-tile instance counting),
// markVisible, and the fused backward preprocess.
//
// Reference behaviour being reproduced (DGR32 = /root/reference/submodules/
// diff-gaussian-rasterization-32):
//   forward  : DGR32/cuda_rasterizer/forward.cu:151-269 (preprocessCUDA),
//              :74-109 (computeCov2D), :114-148 (computeCov3D),
//              auxiliary.h:40-55,70-89,151-176
//   backward : DGR32/cuda_rasterizer/backward.cu:147-326 (computeCov2DCUDA),
//              :330-393 (computeCov3D), :398-449 (preprocessCUDA)
// The forward rounding sequence is the one nvcc 12.9 + ptxas emit for the
// reference on sm_100a (see oracle/splat_oracle.c, which documents the same
// sequence and is the CPU checker for this file).
#include "internal.cuh"
#include <cstdio>

namespace guava {

// ------------------------------------------------------------------ forward
__device__ __forceinline__ void cov3d_from_scale_rot(const float* __restrict__ sc, float mod,
                                                     const float* __restrict__ q, float* c3)
{
    const float sx = fmul(mod, sc[0]), sy = fmul(mod, sc[1]), sz = fmul(mod, sc[2]);
    const float r = q[0], x = q[1], y = q[2], z = q[3];
    const float xz = fmul(x, z), rx = fmul(r, x), rz = fmul(r, z), yy = fmul(y, y), zz = fmul(z, z);
    const float xz_p_ry = ffma(r, y, xz);
    const float xz_m_ry = ffma(-r, y, xz);
    const float yz_m_rx = ffma(y, z, -rx);
    const float yz_p_rx = ffma(y, z, rx);
    const float xx_p_yy = ffma(x, x, yy);
    const float xy_m_rz = ffma(x, y, -rz);
    const float xy_p_rz = ffma(x, y, rz);
    const float yy_p_zz = fadd(yy, zz);
    const float xx_p_zz = ffma(x, x, zz);

    const float R00 = fsub(1.0f, fadd(yy_p_zz, yy_p_zz));
    const float R01 = fadd(xy_m_rz, xy_m_rz);
    const float R02 = fadd(xz_p_ry, xz_p_ry);
    const float R10 = fadd(xy_p_rz, xy_p_rz);
    const float R11 = fsub(1.0f, fadd(xx_p_zz, xx_p_zz));
    const float R12 = fadd(yz_m_rx, yz_m_rx);
    const float R20 = fadd(xz_m_ry, xz_m_ry);
    const float R21 = fadd(yz_p_rx, yz_p_rx);
    const float R22 = fsub(1.0f, fadd(xx_p_yy, xx_p_yy));

    // M = S * R with the zero entries of S kept (they decide the sign of zero
    // and NaN/Inf propagation exactly as the reference's GLM product does).
    const float Z = 0.0f;
    const float M00 = ffma(R02, Z, ffma(R01, Z, fmul(sx, R00)));
    const float t0 = fmul(R00, Z);
    const float M01 = ffma(R02, Z, ffma(sy, R01, t0));
    const float M02 = ffma(sz, R02, ffma(R01, Z, t0));
    const float t1 = fmul(R11, Z);
    const float M10 = ffma(R12, Z, ffma(sx, R10, t1));
    const float M11 = ffma(R12, Z, ffma(R10, Z, fmul(sy, R11)));
    const float M12 = ffma(sz, R12, ffma(R10, Z, t1));
    const float t2 = fmul(R21, Z);
    const float M20 = ffma(R22, Z, ffma(sx, R20, t2));
    const float M21 = ffma(R22, Z, ffma(R20, Z, fmul(sy, R21)));
    const float M22 = ffma(sz, R22, ffma(R20, Z, t2));

    c3[0] = ffma(M02, M02, ffma(M00, M00, fmul(M01, M01)));
    c3[1] = ffma(M12, M02, ffma(M10, M00, fmul(M11, M01)));
    c3[2] = ffma(M22, M02, ffma(M20, M00, fmul(M21, M01)));
    c3[3] = ffma(M12, M12, ffma(M10, M10, fmul(M11, M11)));
    c3[4] = ffma(M22, M12, ffma(M20, M10, fmul(M21, M11)));
    c3[5] = ffma(M22, M22, ffma(M20, M20, fmul(M21, M21)));
}

__device__ __forceinline__ void cov2d_forward(float mx, float my, float mz, float focal_x, float focal_y,
                                              float tan_fovx, float tan_fovy, const float* c3,
                                              const float* __restrict__ V, float& cxx, float& cxy, float& cyy)
{
    const float tx = xform_row(V, 0, mx, my, mz);
    const float ty = xform_row(V, 1, mx, my, mz);
    const float tz = xform_row(V, 2, mx, my, mz);
    const float limx = fmul(tan_fovx, 1.3f), limy = fmul(tan_fovy, 1.3f);
    const float txtz = __fdiv_rn(tx, tz), tytz = __fdiv_rn(ty, tz);
    const float clx = fminf(limx, fmaxf(-limx, txtz));
    const float cly = fminf(limy, fmaxf(-limy, tytz));
    const float J00 = __fdiv_rn(focal_x, tz);
    const float tz2 = fmul(tz, tz);
    const float J02 = __fdiv_rn(fmul(focal_x, fmul(clx, -tz)), tz2);
    const float J11 = __fdiv_rn(focal_y, tz);
    const float J12 = __fdiv_rn(fmul(focal_y, fmul(cly, -tz)), tz2);

    const float Z = 0.0f;
    const float Ta0 = ffma(V[2], J02, ffma(V[0], J00, fmul(V[1], Z)));
    const float Ta1 = ffma(V[6], J02, ffma(V[4], J00, fmul(V[5], Z)));
    const float Ta2 = ffma(J02, V[10], ffma(V[8], J00, fmul(V[9], Z)));
    const float Tb0 = ffma(V[2], J12, ffma(V[0], Z, fmul(J11, V[1])));
    const float Tb1 = ffma(V[6], J12, ffma(V[4], Z, fmul(J11, V[5])));
    const float Tb2 = ffma(J12, V[10], ffma(V[8], Z, fmul(J11, V[9])));

    const float A0 = ffma(Ta2, c3[2], ffma(Ta0, c3[0], fmul(Ta1, c3[1])));
    const float B0 = ffma(Tb2, c3[2], ffma(Tb0, c3[0], fmul(Tb1, c3[1])));
    const float A1 = ffma(Ta2, c3[4], ffma(Ta0, c3[1], fmul(Ta1, c3[3])));
    const float B1 = ffma(Tb2, c3[4], ffma(Tb0, c3[1], fmul(Tb1, c3[3])));
    const float A2 = ffma(Ta2, c3[5], ffma(Ta0, c3[2], fmul(Ta1, c3[4])));
    const float B2 = ffma(Tb2, c3[5], ffma(Tb0, c3[2], fmul(Tb1, c3[4])));
    cxx = ffma(Ta2, A2, ffma(Ta0, A0, fmul(Ta1, A1)));
    cxy = ffma(Ta2, B2, ffma(Ta0, B0, fmul(Ta1, B1)));
    cyy = ffma(Tb2, B2, ffma(Tb0, B0, fmul(Tb1, B1)));
}

// The caller's [P,3] tensors (means3D, scales) are arrays of 12-byte structures: read thread-per-Gaussian they are
// stride-12 scalar loads.  A block's 256 Gaussians are one contiguous 3072-byte run, so the block fetches it as 192
// coalesced 16-byte vectors into shared memory and every thread picks its three floats from there (stride 3 words:
// conflict-free); [P,4] rotations are one 16-byte load per thread.  Falls back to scalar loads when the run is not
// 16-byte aligned or the block straddles two views.
__device__ __forceinline__ void load_xyz_block(const float* __restrict__ base, size_t vs, int P, size_t gid0, int count,
                                               float* smem, int tid, float& x, float& y, float& z, bool valid)
{
    const int v0 = (int)(gid0 / P), v1 = (int)((gid0 + count - 1) / P);
    const float* __restrict__ src = base + (size_t)v0 * vs + 3 * (gid0 - (size_t)v0 * P);
    const bool vec = v0 == v1 && ((reinterpret_cast<uintptr_t>(src) & 15u) == 0u);   // block-uniform
    if (vec) {
        const int nf = count * 3, nv = nf >> 2;
        if (tid < nv) reinterpret_cast<float4*>(smem)[tid] = __ldg(reinterpret_cast<const float4*>(src) + tid);
        if (tid < nf - 4 * nv) smem[4 * nv + tid] = __ldg(src + 4 * nv + tid);
        __syncthreads();
        if (valid) { x = smem[3 * tid]; y = smem[3 * tid + 1]; z = smem[3 * tid + 2]; }
    } else if (valid) {
        const size_t gid = gid0 + tid;
        const int v = (int)(gid / P);
        const float* __restrict__ p = base + (size_t)v * vs + 3 * (gid - (size_t)v * P);
        x = p[0]; y = p[1]; z = p[2];
    }
}

__global__ void __launch_bounds__(256)
preprocess_forward_kernel(const Inputs in, const GeomPtrs g, uint32_t* __restrict__ tile_count,
                          int* __restrict__ radii)
{
    __shared__ __align__(16) float s_xyz[256 * 3];
    __shared__ __align__(16) float s_scl[256 * 3];
    const size_t n_total = (size_t)in.V * in.P;
    const size_t gid0 = (size_t)blockIdx.x * 256;
    const size_t gid = gid0 + threadIdx.x;
    const int count = (int)min((size_t)256, n_total - gid0);
    const bool valid = gid < n_total;
    // Every input of a Gaussian is requested before the first early exit: the grid is a single wave, so
    // each dependent round trip to HBM behind a branch (scale/rotation after the frustum test, opacity after
    // the rect test) lengthened the whole kernel by one memory latency.  Culled Gaussians read 32 bytes in vain.
    float mx = 0.0f, my = 0.0f, mz = 0.0f;
    float sr[7] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    load_xyz_block(in.means3D, in.means3D_vs, in.P, gid0, count, s_xyz, threadIdx.x, mx, my, mz, valid);
    if (in.cov3D_precomp == nullptr)
        load_xyz_block(in.scales, in.scales_vs, in.P, gid0, count, s_scl, threadIdx.x, sr[0], sr[1], sr[2], valid);
    if (!valid) return;
    const int v = (int)(gid / in.P);
    const int i = (int)(gid - (size_t)v * in.P);

    radii[gid] = 0;
    g.tiles_touched[gid] = 0;
    g.rects[gid] = make_ushort4(0, 0, 0, 0);

    const float* __restrict__ Vm = in.viewmatrix + 16 * v;
    const float* __restrict__ Pm = in.projmatrix + 16 * v;
    if (in.cov3D_precomp == nullptr) {
        const float* __restrict__ qp = in.rotations + (size_t)v * in.rotations_vs + 4 * (size_t)i;
        if ((reinterpret_cast<uintptr_t>(qp) & 15u) == 0u) {   // [P,4]: one 16-byte load per Gaussian
            const float4 q = __ldg(reinterpret_cast<const float4*>(qp));
            sr[3] = q.x; sr[4] = q.y; sr[5] = q.z; sr[6] = q.w;
        } else {
            sr[3] = qp[0]; sr[4] = qp[1]; sr[5] = qp[2]; sr[6] = qp[3];
        }
    }
    const float opacity = in.opacities[(size_t)v * in.opacities_vs + i];
    float tan_fovx = in.tan_fovx, tan_fovy = in.tan_fovy;
    if (in.tan_fov_dev != nullptr) { tan_fovx = in.tan_fov_dev[2 * v]; tan_fovy = in.tan_fov_dev[2 * v + 1]; }

    // in_frustum (auxiliary.h:151-176): only the near-plane test is live
    const float depth = xform_row(Vm, 2, mx, my, mz);
    if (depth <= 0.2f) {
        if (in.prefiltered) {
            printf("Point is filtered although prefiltered is set. This shouldn't happen!");
            __trap();
        }
        return;
    }
    const float hx = xform_row(Pm, 0, mx, my, mz);
    const float hy = xform_row(Pm, 1, mx, my, mz);
    const float hw = xform_row(Pm, 3, mx, my, mz);
    const float p_w = __frcp_rn(fadd(hw, 0.0000001f));
    const float projx = fmul(hx, p_w), projy = fmul(hy, p_w);

    float c3[6];
    if (in.cov3D_precomp != nullptr) {
        const float* cp = in.cov3D_precomp + 6 * (size_t)gid;
#pragma unroll
        for (int k = 0; k < 6; k++) c3[k] = cp[k];
    } else {
        cov3d_from_scale_rot(sr, in.scale_modifier, sr + 3, c3);
    }

    // rasterizer_impl.cu:223-224 (host float arithmetic in the reference)
    const float focal_y = __fdiv_rn(__int2float_rn(in.H), fmul(2.0f, tan_fovy));
    const float focal_x = __fdiv_rn(__int2float_rn(in.W), fmul(2.0f, tan_fovx));

    float cxx, cxy, cyy;
    cov2d_forward(mx, my, mz, focal_x, focal_y, tan_fovx, tan_fovy, c3, Vm, cxx, cxy, cyy);

    const float bb = fmul(cxy, cxy);
    const float det_cov = ffma(cxx, cyy, -bb);
    const float a = fadd(cxx, 0.3f), c = fadd(cyy, 0.3f);
    const float det = ffma(a, c, -bb);
    float h_scale = 1.0f;
    if (in.antialiasing) h_scale = __fsqrt_rn(fmaxf(__fdiv_rn(det_cov, det), 0.000025f));
    if (det == 0.0f) return;
    const float det_inv = __frcp_rn(det);
    const float conx = fmul(c, det_inv), cony = fmul(det_inv, -cxy), conz = fmul(a, det_inv);

    const float mid = fmul(fadd(a, c), 0.5f);
    const float sq = __fsqrt_rn(fmaxf(ffma(mid, mid, -det), 0.1f));
    const float lambda1 = fadd(mid, sq), lambda2 = fsub(mid, sq);
    const float my_radius = ceilf(fmul(__fsqrt_rn(fmaxf(lambda1, lambda2)), 3.0f));
    const float px = ndc2pix(projx, in.W), py = ndc2pix(projy, in.H);
    const int rad = __float2int_rz(my_radius);
    uint32_t minx, miny, maxx, maxy;
    get_rect(px, py, rad, (uint32_t)in.gx, (uint32_t)in.gy, minx, miny, maxx, maxy);
    const uint32_t touched = (maxx - minx) * (maxy - miny);
    if (touched == 0) return;

    SplatRec rec;
    rec.x = px; rec.y = py; rec.cx = conx; rec.cy = cony; rec.cz = conz;
    rec.op = fmul(h_scale, opacity);
    rec.rdepth = __frcp_rn(depth);
    rec.lthr = __logf(255.0f * rec.op) + 0.01f;   // culling threshold ln(255*opacity) + margin (splat_may_touch)
    float4* rp = reinterpret_cast<float4*>(g.rec + gid);
    rp[0] = make_float4(rec.x, rec.y, rec.cx, rec.cy);
    rp[1] = make_float4(rec.cz, rec.op, rec.rdepth, rec.lthr);
    g.depths[gid] = depth;
    radii[gid] = rad;
    g.rects[gid] = make_ushort4((unsigned short)minx, (unsigned short)miny, (unsigned short)maxx, (unsigned short)maxy);
    g.tiles_touched[gid] = touched;

    // Per-tile instance counts (replaces the per-Gaussian InclusiveSum of
    // rasterizer_impl.cu:280: offsets are taken per tile, so tile ranges fall
    // out of one scan over tiles and num_rendered never has to visit the host
    // before the instance buffers can be filled).
    uint32_t* tc = tile_count + (size_t)v * in.tiles * CURSOR_STRIDE;
    for (uint32_t y = miny; y < maxy; y++)
        for (uint32_t x = minx; x < maxx; x++)
            atomicAdd(tc + (size_t)(y * in.gx + x) * CURSOR_STRIDE, 1u);
}

void launch_preprocess_forward(const Inputs& in, const GeomPtrs& g, const ImagePtrs& img, int* radii, cudaStream_t s)
{
    const size_t n = (size_t)in.V * in.P;
    if (n == 0) return;
    preprocess_forward_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(in, g, img.tile_cursor, radii);
    count_launch();
}

// -------------------------------------------------------------- markVisible
__global__ void __launch_bounds__(256)
mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ Vm, uint8_t* __restrict__ present)
{
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P) return;
    const float d = xform_row(Vm, 2, means3D[3 * (size_t)i], means3D[3 * (size_t)i + 1], means3D[3 * (size_t)i + 2]);
    present[i] = d > 0.2f ? 1 : 0;
}

void launch_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present, cudaStream_t s)
{
    if (P == 0) return;
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, viewmatrix, present);
    count_launch();
}

// ----------------------------------------------------------------- backward
// One kernel for backward.cu's computeCov2DCUDA + preprocessCUDA + computeCov3D:
// each visible Gaussian reads its accumulated screen-space gradients once and
// writes dL/dmean3D, dL/dcov3D, dL/dscale, dL/drot.  The 3D covariance is
// recomputed from scale/rotation (12 FFMA-chains) instead of being stored by
// the forward pass and re-read here (24 B/Gaussian each way).
// SUM = 1 (guava_rasterizer_backward_batch_summed): the parameter gradients of Gaussian i are ADDED, over the views, to
// row i of the four output arrays (row strides in `sum`: the columns of one packed [P,43] row, or dense [P,w] tensors);
// the per-view outputs the reference's binding keeps (mean2D, conic, cov3D, inverse depth) are not written.
template <int SUM>
__global__ void __launch_bounds__(256)
preprocess_backward_kernel(const Inputs in, uint32_t* __restrict__ num_rendered, const int* __restrict__ radii,
                           const SplatRec* __restrict__ recs,
                           float* __restrict__ grad_geo,
                           int has_invdepth, float* __restrict__ dL_dmean2D, float* __restrict__ dL_dconic,
                           float* __restrict__ dL_dinvdepth, float* __restrict__ dL_dopacity,
                           float* __restrict__ dL_dmean3D, float* __restrict__ dL_dcov3D,
                           float* __restrict__ dL_dscale, float* __restrict__ dL_drot, const SummedRows sum)
{
    // SUM with a power-of-two view count (sum.vshift >= 0): the V views of a Gaussian sit in adjacent lanes, their
    // eleven values are added with shuffles and lane v = 0 stores the sums -- plain stores instead of 11 V scattered
    // reductions per Gaussian (at V = 2, P = 100 k the atomics cost 7 us of a 12 us kernel).  Every lane then stays to
    // the end: what used to be early returns are predicates.
    const size_t t = (size_t)blockIdx.x * 256 + threadIdx.x;
    const size_t n_total = (size_t)in.V * in.P;
    const bool lanes = SUM && sum.vshift >= 0;
    if (!lanes && t >= n_total) return;
    if (num_rendered[1]) return;   // the forward's binning overflowed: no lists, no gradients (see render_backward.cu); grid-uniform
    if (t == 0) num_rendered[3] = 0u;   // the backward compositor's work counter, re-armed for a repeated backward (it ran before this kernel)
    const bool in_range = t < n_total;
    const size_t gid = !lanes ? t : (in_range ? (t & (size_t)(in.V - 1)) * in.P + (t >> sum.vshift) : 0);
    float acc[11];
#pragma unroll
    for (int k = 0; k < 11; k++) acc[k] = 0.0f;
    // Per-Gaussian sums written by the backward compositing kernel: moments of s = dL/dG * G about
    // the pixel offset d = mean2D - pixel, (sum s*dx, s*dy, s*dx^2, s*dx*dy, s*dy^2), then
    // sum G*dL/dalpha and sum w*dL/dinvdepth.  backward.cu:612-635 in terms of them.
    // All of this Gaussian's inputs are requested together with its radius, before the visibility exit
    // (single-wave grid: a second round trip to HBM behind the branch is paid by the whole kernel).
    const int rad = radii[gid];
    const float4 mo0 = reinterpret_cast<const float4*>(grad_geo)[2 * gid];
    const float4 mo1 = reinterpret_cast<const float4*>(grad_geo)[2 * gid + 1];
    const float4 rc0 = reinterpret_cast<const float4*>(recs)[2 * gid];       // x, y, conic.x, conic.y
    const float4 rc1 = reinterpret_cast<const float4*>(recs)[2 * gid + 1];   // conic.z, ...
    const int v = (int)(gid / in.P);
    const int i = (int)(gid - (size_t)v * in.P);
    const float* __restrict__ mp = in.means3D + (size_t)v * in.means3D_vs + 3 * (size_t)i;
    const float mx = mp[0], my = mp[1], mz = mp[2];
    float sr[7] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    if (in.cov3D_precomp == nullptr) {
        const float* __restrict__ sp = in.scales + (size_t)v * in.scales_vs + 3 * (size_t)i;
        const float* __restrict__ qp = in.rotations + (size_t)v * in.rotations_vs + 4 * (size_t)i;
        sr[0] = sp[0]; sr[1] = sp[1]; sr[2] = sp[2];
        sr[3] = qp[0]; sr[4] = qp[1]; sr[5] = qp[2]; sr[6] = qp[3];
    }
    if (in_range && rad > 0) {
    // consumed: the record is zero again for the next backward of this forward (the forward's scan zero-filled it first)
    reinterpret_cast<float4*>(grad_geo)[2 * gid] = make_float4(0.f, 0.f, 0.f, 0.f);
    reinterpret_cast<float4*>(grad_geo)[2 * gid + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 gg0, gg1;
    gg0.x = -(rc0.z * mo0.x + rc0.w * mo0.y) * (0.5f * in.W);   // dL/dmean2D.x (NDC units, :626)
    gg0.y = -(rc1.x * mo0.y + rc0.w * mo0.x) * (0.5f * in.H);   // dL/dmean2D.y
    gg0.z = -0.5f * mo0.z;                                       // dL/dconic.x
    gg0.w = -0.5f * mo0.w;                                       // dL/dconic.y
    gg1.x = -0.5f * mo1.x;                                       // dL/dconic.w
    gg1.y = mo1.y;                                               // dL/dopacity (before antialias scaling)
    gg1.z = mo1.z;                                               // dL/dinvdepth
    gg1.w = 0.0f;
    // outputs the reference's binding returns / keeps (rasterize_points.cu:163-172)
    if (!SUM) {
        dL_dmean2D[3 * gid] = gg0.x;
        dL_dmean2D[3 * gid + 1] = gg0.y;
        if (dL_dconic != nullptr) {
            dL_dconic[4 * gid] = gg0.z;
            dL_dconic[4 * gid + 1] = gg0.w;
            dL_dconic[4 * gid + 3] = gg1.x;
        }
        if (has_invdepth && dL_dinvdepth != nullptr) dL_dinvdepth[gid] = gg1.z;
    }
    const float* __restrict__ V = in.viewmatrix + 16 * v;
    const float* __restrict__ proj = in.projmatrix + 16 * v;

    float tan_fovx = in.tan_fovx, tan_fovy = in.tan_fovy;
    if (in.tan_fov_dev != nullptr) { tan_fovx = in.tan_fov_dev[2 * v]; tan_fovy = in.tan_fov_dev[2 * v + 1]; }
    const float h_y = in.H / (2.0f * tan_fovy), h_x = in.W / (2.0f * tan_fovx);

    float c3[6];
    if (in.cov3D_precomp != nullptr) {
#pragma unroll
        for (int k = 0; k < 6; k++) c3[k] = in.cov3D_precomp[6 * gid + k];
    } else {
        cov3d_from_scale_rot(sr, in.scale_modifier, sr + 3, c3);
    }

    const float dLc_x = gg0.z, dLc_y = gg0.w, dLc_z = gg1.x;
    float tx = xform_row(V, 0, mx, my, mz);
    float ty = xform_row(V, 1, mx, my, mz);
    const float tz = xform_row(V, 2, mx, my, mz);
    const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
    const float txtz = tx / tz, tytz = ty / tz;
    tx = fminf(limx, fmaxf(-limx, txtz)) * tz;
    ty = fminf(limy, fmaxf(-limy, tytz)) * tz;
    const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.0f : 1.0f;
    const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.0f : 1.0f;

    const float J00 = h_x / tz, J02 = -(h_x * tx) / (tz * tz);
    const float J11 = h_y / tz, J12 = -(h_y * ty) / (tz * tz);
    float T0[3], T1[3];
#pragma unroll
    for (int r = 0; r < 3; r++) {
        T0[r] = V[4 * r + 0] * J00 + V[4 * r + 2] * J02;
        T1[r] = V[4 * r + 1] * J11 + V[4 * r + 2] * J12;
    }
    const float K[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
    float VT0[3], VT1[3];
#pragma unroll
    for (int r = 0; r < 3; r++) {
        VT0[r] = K[r][0] * T0[0] + K[r][1] * T0[1] + K[r][2] * T0[2];
        VT1[r] = K[r][0] * T1[0] + K[r][1] * T1[1] + K[r][2] * T1[2];
    }
    float c_xx = T0[0] * VT0[0] + T0[1] * VT0[1] + T0[2] * VT0[2];
    const float c_xy = T0[0] * VT1[0] + T0[1] * VT1[1] + T0[2] * VT1[2];
    float c_yy = T1[0] * VT1[0] + T1[1] * VT1[1] + T1[2] * VT1[2];

    const float h_var = 0.3f;
    float d_inside_root = 0.0f;
    if (in.antialiasing) {
        const float det_cov = c_xx * c_yy - c_xy * c_xy;
        c_xx += h_var; c_yy += h_var;
        const float det_cov_plus_h_cov = c_xx * c_yy - c_xy * c_xy;
        const float h_scale = sqrtf(fmaxf(0.000025f, det_cov / det_cov_plus_h_cov));
        const float dL_dop_v = gg1.y;
        const float d_h_scale = dL_dop_v * in.opacities[(size_t)v * in.opacities_vs + i];
        if (SUM) acc[10] = dL_dop_v * h_scale;
        else dL_dopacity[gid] = dL_dop_v * h_scale;
        d_inside_root = (det_cov / det_cov_plus_h_cov) <= 0.000025f ? 0.0f : d_h_scale / (2 * h_scale);
    } else {
        c_xx += h_var; c_yy += h_var;
        if (SUM) acc[10] = gg1.y;
        else dL_dopacity[gid] = gg1.y;
    }
    float dL_dc_xx = 0, dL_dc_xy = 0, dL_dc_yy = 0;
    if (in.antialiasing) {
        const float x = c_xx, y = c_yy, z = c_xy, w = h_var;
        const float den = w * w + w * (x + y) + x * y - z * z;
        const float denom_f = d_inside_root / (den * den);
        dL_dc_xx = w * (w * y + y * y + z * z) * denom_f;
        dL_dc_yy = w * (w * x + x * x + z * z) * denom_f;
        dL_dc_xy = -2.0f * w * z * (w + x + y) * denom_f;
    }
    const float denom = c_xx * c_yy - c_xy * c_xy;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    float dcov[6];
    if (denom2inv != 0) {
        dL_dc_xx += denom2inv * (-c_yy * c_yy * dLc_x + 2 * c_xy * c_yy * dLc_y + (denom - c_xx * c_yy) * dLc_z);
        dL_dc_yy += denom2inv * (-c_xx * c_xx * dLc_z + 2 * c_xx * c_xy * dLc_y + (denom - c_xx * c_yy) * dLc_x);
        dL_dc_xy += denom2inv * 2 * (c_xy * c_yy * dLc_x - (denom + 2 * c_xy * c_xy) * dLc_y + c_xx * c_xy * dLc_z);
        dcov[0] = T0[0] * T0[0] * dL_dc_xx + T0[0] * T1[0] * dL_dc_xy + T1[0] * T1[0] * dL_dc_yy;
        dcov[3] = T0[1] * T0[1] * dL_dc_xx + T0[1] * T1[1] * dL_dc_xy + T1[1] * T1[1] * dL_dc_yy;
        dcov[5] = T0[2] * T0[2] * dL_dc_xx + T0[2] * T1[2] * dL_dc_xy + T1[2] * T1[2] * dL_dc_yy;
        dcov[1] = 2 * T0[0] * T0[1] * dL_dc_xx + (T0[0] * T1[1] + T0[1] * T1[0]) * dL_dc_xy + 2 * T1[0] * T1[1] * dL_dc_yy;
        dcov[2] = 2 * T0[0] * T0[2] * dL_dc_xx + (T0[0] * T1[2] + T0[2] * T1[0]) * dL_dc_xy + 2 * T1[0] * T1[2] * dL_dc_yy;
        dcov[4] = 2 * T0[2] * T0[1] * dL_dc_xx + (T0[1] * T1[2] + T0[2] * T1[1]) * dL_dc_xy + 2 * T1[1] * T1[2] * dL_dc_yy;
    } else {
#pragma unroll
        for (int k = 0; k < 6; k++) dcov[k] = 0;
    }
    if (!SUM) {
#pragma unroll
        for (int k = 0; k < 6; k++) dL_dcov3D[6 * gid + k] = dcov[k];
    }

    const float dL_dT00 = 2 * VT0[0] * dL_dc_xx + VT1[0] * dL_dc_xy;
    const float dL_dT01 = 2 * VT0[1] * dL_dc_xx + VT1[1] * dL_dc_xy;
    const float dL_dT02 = 2 * VT0[2] * dL_dc_xx + VT1[2] * dL_dc_xy;
    const float dL_dT10 = 2 * VT1[0] * dL_dc_yy + VT0[0] * dL_dc_xy;
    const float dL_dT11 = 2 * VT1[1] * dL_dc_yy + VT0[1] * dL_dc_xy;
    const float dL_dT12 = 2 * VT1[2] * dL_dc_yy + VT0[2] * dL_dc_xy;
    const float dL_dJ00 = V[0] * dL_dT00 + V[4] * dL_dT01 + V[8] * dL_dT02;
    const float dL_dJ02 = V[2] * dL_dT00 + V[6] * dL_dT01 + V[10] * dL_dT02;
    const float dL_dJ11 = V[1] * dL_dT10 + V[5] * dL_dT11 + V[9] * dL_dT12;
    const float dL_dJ12 = V[2] * dL_dT10 + V[6] * dL_dT11 + V[10] * dL_dT12;
    const float itz = 1.0f / tz, itz2 = itz * itz, itz3 = itz2 * itz;
    const float dL_dtx = x_grad_mul * -h_x * itz2 * dL_dJ02;
    const float dL_dty = y_grad_mul * -h_y * itz2 * dL_dJ12;
    float dL_dtz = -h_x * itz2 * dL_dJ00 - h_y * itz2 * dL_dJ11
                 + (2 * h_x * tx) * itz3 * dL_dJ02 + (2 * h_y * ty) * itz3 * dL_dJ12;
    if (has_invdepth) dL_dtz -= gg1.z / (tz * tz);

    float gmx = V[0] * dL_dtx + V[1] * dL_dty + V[2] * dL_dtz;
    float gmy = V[4] * dL_dtx + V[5] * dL_dty + V[6] * dL_dtz;
    float gmz = V[8] * dL_dtx + V[9] * dL_dty + V[10] * dL_dtz;

    // projection part (backward.cu:420-440)
    const float m_w = 1.0f / (xform_row(proj, 3, mx, my, mz) + 0.0000001f);
    const float mul1 = (proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12]) * m_w * m_w;
    const float mul2 = (proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13]) * m_w * m_w;
    const float g2x = gg0.x, g2y = gg0.y;
    gmx += (proj[0] * m_w - proj[3] * mul1) * g2x + (proj[1] * m_w - proj[3] * mul2) * g2y;
    gmy += (proj[4] * m_w - proj[7] * mul1) * g2x + (proj[5] * m_w - proj[7] * mul2) * g2y;
    gmz += (proj[8] * m_w - proj[11] * mul1) * g2x + (proj[9] * m_w - proj[11] * mul2) * g2y;
    if (SUM) {
        acc[0] = gmx; acc[1] = gmy; acc[2] = gmz;
    } else {
        dL_dmean3D[3 * gid] = gmx; dL_dmean3D[3 * gid + 1] = gmy; dL_dmean3D[3 * gid + 2] = gmz;
    }

    if (in.cov3D_precomp == nullptr) {
        const float r = sr[3], x = sr[4], y = sr[5], z = sr[6];
        const float R[3][3] = {
            {1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
            {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
            {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
        const float s[3] = {in.scale_modifier * sr[0], in.scale_modifier * sr[1], in.scale_modifier * sr[2]};
        float M[3][3];
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int rr = 0; rr < 3; rr++) M[c][rr] = s[rr] * R[c][rr];
        const float dS[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                                {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                                {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
        float dMt[3][3];   // dMt[c][r] = dL_dM[r][c]
#pragma unroll
        for (int c = 0; c < 3; c++)
#pragma unroll
            for (int rr = 0; rr < 3; rr++)
                dMt[rr][c] = 2.0f * (M[0][rr] * dS[c][0] + M[1][rr] * dS[c][1] + M[2][rr] * dS[c][2]);
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const float ds = R[0][k] * dMt[k][0] + R[1][k] * dMt[k][1] + R[2][k] * dMt[k][2];
            if (SUM) acc[3 + k] = ds;
            else dL_dscale[3 * gid + k] = ds;
        }
#pragma unroll
        for (int k = 0; k < 3; k++)
#pragma unroll
            for (int rr = 0; rr < 3; rr++) dMt[k][rr] *= s[k];
        float4 dq;
        dq.x = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
        dq.y = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
        dq.z = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
        dq.w = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
        if (SUM) {
            acc[6] = dq.x; acc[7] = dq.y; acc[8] = dq.z; acc[9] = dq.w;
        } else {
            reinterpret_cast<float4*>(dL_drot)[gid] = dq;
        }
    }
    }   // visible
    if (!SUM) return;
    if (lanes) {
        // lanes 0 .. V-1 of every aligned group hold the views of one Gaussian: butterfly over the group
        for (int o = 1; o < in.V; o <<= 1) {
#pragma unroll
            for (int k = 0; k < 11; k++) acc[k] += __shfl_xor_sync(0xffffffffu, acc[k], o);
        }
        if (!in_range || v != 0) return;
        float* q;
        q = dL_dmean3D + (size_t)i * sum.rs_mean3D;  q[0] = acc[0]; q[1] = acc[1]; q[2] = acc[2];
        q = dL_dscale + (size_t)i * sum.rs_scale;    q[0] = acc[3]; q[1] = acc[4]; q[2] = acc[5];
        q = dL_drot + (size_t)i * sum.rs_rot;        q[0] = acc[6]; q[1] = acc[7]; q[2] = acc[8]; q[3] = acc[9];
        dL_dopacity[(size_t)i * sum.rs_opacity] = acc[10];
    } else if (rad > 0) {   // any view count: the views add up in the zero-filled rows
        float* q;
        q = dL_dmean3D + (size_t)i * sum.rs_mean3D;  atomicAdd(q, acc[0]); atomicAdd(q + 1, acc[1]); atomicAdd(q + 2, acc[2]);
        q = dL_dscale + (size_t)i * sum.rs_scale;    atomicAdd(q, acc[3]); atomicAdd(q + 1, acc[4]); atomicAdd(q + 2, acc[5]);
        q = dL_drot + (size_t)i * sum.rs_rot;
        atomicAdd(q, acc[6]); atomicAdd(q + 1, acc[7]); atomicAdd(q + 2, acc[8]); atomicAdd(q + 3, acc[9]);
        atomicAdd(dL_dopacity + (size_t)i * sum.rs_opacity, acc[10]);
    }
}

void launch_preprocess_backward(const Inputs& in, const GeomPtrs& g, uint32_t* num_rendered, const int* radii, int has_invdepth,
                                float* dL_dmean2D, float* dL_dconic, float* dL_dinvdepth,
                                float* dL_dopacity, float* dL_dmean3D, float* dL_dcov3D,
                                float* dL_dscale, float* dL_drot, const SummedRows* sum, cudaStream_t s)
{
    const size_t n = (size_t)in.V * in.P;
    if (n == 0) return;
    const unsigned grid = (unsigned)((n + 255) / 256);
    if (sum != nullptr) {
        SummedRows sr = *sum;
        sr.vshift = -1;
        if (in.V <= 32 && (in.V & (in.V - 1)) == 0)
            for (int k = 0; k < 6; k++)
                if ((1 << k) == in.V) sr.vshift = k;
        preprocess_backward_kernel<1><<<grid, 256, 0, s>>>(
            in, num_rendered, radii, g.rec, g.grad_geo, has_invdepth, dL_dmean2D, dL_dconic, dL_dinvdepth, dL_dopacity, dL_dmean3D,
            dL_dcov3D, dL_dscale, dL_drot, sr);
    }
    else
        preprocess_backward_kernel<0><<<grid, 256, 0, s>>>(
            in, num_rendered, radii, g.rec, g.grad_geo, has_invdepth, dL_dmean2D, dL_dconic, dL_dinvdepth, dL_dopacity, dL_dmean3D,
            dL_dcov3D, dL_dscale, dL_drot, SummedRows{});
    count_launch();
}

}  // namespace guava
