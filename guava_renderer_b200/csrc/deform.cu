// Fused deformation -> rasterizer inputs: the step immediately before the
// rasterizer in GUAVA.
//
// Reference behaviour: models/UbodyAvatar/ubody_gaussian.py:245-289 (Ubody_Gaussian.forward) --
//   * SMPLX-vertex Gaussians: position = deformed vertex; rotation =
//     normalize(quat(ver_transform[:3,:3]) (x) smplx_rotation)                                (:252-254)
//   * UV Gaussians bound to mesh faces: per-face frame (orientation, scale) from the deformed mesh
//     (utils/graphics_utils.py:61-80 compute_face_orientation), then
//         xyz      = (R_face . local_xyz) * s_face + sum_k bary_k * v_k                       (:260-267)
//         rotation = quat(R_face) (x) uv_rotation                                             (:269-270)
//         scaling  = uv_scaling * s_face                                                      (:271)
//   * concatenation of both groups along the Gaussian axis                                    (:274-279)
// In the reference this is ~25 torch ops that materialise [B,N,3,3] gathers; here one kernel builds
// the per-face frames and one kernel per direction produces / back-propagates all Gaussians, writing
// straight into the concatenated [B, Nv+N, *] tensors the rasterizer reads.
//
// Quaternion helpers restate roma 1.5.3 (requirements.txt:6; the package is not vendored in the
// reference): rotmat_to_unitquat (the SciPy Rotation.from_matrix algorithm: largest of the three
// diagonal entries and the trace decides the branch; XYZW order) and quat_product (Hamilton product).
// Gradients are produced for the per-Gaussian attributes (uv_local_xyz, uv_rotation, uv_scaling,
// smplx_rotation); the mesh (vertices, vertex transforms) is an input without gradient.
#include "internal.cuh"
#include "../../include/guava_rasterizer.h"

#include <atomic>
#include <string>

namespace guava {

struct Q4 { float w, x, y, z; };

__device__ __forceinline__ Q4 quat_mul(const Q4& p, const Q4& q)
{
    // roma.quat_product: v = pw*qv + qw*pv + pv x qv ; w = pw*qw - pv.qv
    Q4 r;
    r.x = p.w * q.x + q.w * p.x + (p.y * q.z - p.z * q.y);
    r.y = p.w * q.y + q.w * p.y + (p.z * q.x - p.x * q.z);
    r.z = p.w * q.z + q.w * p.z + (p.x * q.y - p.y * q.x);
    r.w = p.w * q.w - (p.x * q.x + p.y * q.y + p.z * q.z);
    return r;
}

// d(p (x) q)/dq applied to g (the product is linear in q)
__device__ __forceinline__ Q4 quat_mul_grad_q(const Q4& p, const Q4& g)
{
    Q4 d;
    d.w = (g.x * p.x + g.y * p.y + g.z * p.z) + g.w * p.w;
    d.x = p.w * g.x + (g.y * p.z - g.z * p.y) - g.w * p.x;
    d.y = p.w * g.y + (g.z * p.x - g.x * p.z) - g.w * p.y;
    d.z = p.w * g.z + (g.x * p.y - g.y * p.x) - g.w * p.z;
    return d;
}

// roma.rotmat_to_unitquat; m is row-major m[3*i + j]; returns wxyz
__device__ __forceinline__ Q4 rotmat_to_quat(const float* m)
{
    const float d0 = m[0], d1 = m[4], d2 = m[8];
    const float tr = d0 + d1 + d2;
    float q[4];   // xyzw
    // argmax over (d0, d1, d2, tr), first maximum wins
    int c = 0;
    float best = d0;
    if (d1 > best) { best = d1; c = 1; }
    if (d2 > best) { best = d2; c = 2; }
    if (tr > best) { c = 3; }
    if (c != 3) {
        const int i = c, j = (i + 1) % 3, k = (j + 1) % 3;
        q[i] = 1.0f - tr + 2.0f * m[3 * i + i];
        q[j] = m[3 * j + i] + m[3 * i + j];
        q[k] = m[3 * k + i] + m[3 * i + k];
        q[3] = m[3 * k + j] - m[3 * j + k];
    } else {
        q[0] = m[7] - m[5];
        q[1] = m[2] - m[6];
        q[2] = m[3] - m[1];
        q[3] = 1.0f + tr;
    }
    const float inv = 1.0f / sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    Q4 r;
    r.x = q[0] * inv; r.y = q[1] * inv; r.z = q[2] * inv; r.w = q[3] * inv;
    return r;
}

// graphics_utils.py:55-59: x / sqrt(clamp(dot(x,x), min=1e-20))
__device__ __forceinline__ float safe_len(float x, float y, float z)
{
    return sqrtf(fmaxf(x * x + y * y + z * z, 1e-20f));
}

// ------------------------------------------------------------------------------- face frames
// One thread per (batch, face): 16 floats = R (row-major, columns a0 a1 a2), scale, quat wxyz, pad.
__global__ void __launch_bounds__(256)
face_frames_kernel(int B, int Nv, int F, const float* __restrict__ vertices, const int* __restrict__ faces,
                   float4* __restrict__ frames)
{
    const size_t t = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (t >= (size_t)B * F) return;
    const int b = (int)(t / F), f = (int)(t - (size_t)b * F);
    const float* vb = vertices + (size_t)b * Nv * 3;
    const int i0 = faces[3 * f], i1 = faces[3 * f + 1], i2 = faces[3 * f + 2];
    const float v0x = vb[3 * i0], v0y = vb[3 * i0 + 1], v0z = vb[3 * i0 + 2];
    const float e1x = vb[3 * i1] - v0x, e1y = vb[3 * i1 + 1] - v0y, e1z = vb[3 * i1 + 2] - v0z;
    const float e2x = vb[3 * i2] - v0x, e2y = vb[3 * i2 + 1] - v0y, e2z = vb[3 * i2 + 2] - v0z;
    const float s0 = safe_len(e1x, e1y, e1z);
    const float a0x = e1x / s0, a0y = e1y / s0, a0z = e1z / s0;
    float cx = a0y * e2z - a0z * e2y, cy = a0z * e2x - a0x * e2z, cz = a0x * e2y - a0y * e2x;
    float l = safe_len(cx, cy, cz);
    const float a1x = cx / l, a1y = cy / l, a1z = cz / l;
    cx = a1y * a0z - a1z * a0y; cy = a1z * a0x - a1x * a0z; cz = a1x * a0y - a1y * a0x;
    l = safe_len(cx, cy, cz);
    const float a2x = -(cx / l), a2y = -(cy / l), a2z = -(cz / l);
    const float s1 = fabsf(a2x * e2x + a2y * e2y + a2z * e2z);
    const float scale = (s0 + s1) / 2.0f;
    const float m[9] = {a0x, a1x, a2x, a0y, a1y, a2y, a0z, a1z, a2z};
    const Q4 q = rotmat_to_quat(m);
    float4* o = frames + t * 4;
    o[0] = make_float4(m[0], m[1], m[2], m[3]);
    o[1] = make_float4(m[4], m[5], m[6], m[7]);
    o[2] = make_float4(m[8], scale, q.w, q.x);
    o[3] = make_float4(q.y, q.z, 0.0f, 0.0f);
}

struct DeformArgs {
    int B, Nv, N, F;
    const float* vertices;        // [B,Nv,3]
    const float* ver_transform;   // [B,Nv,4,4]
    const int* faces;             // [F,3]
    const int* binding;           // [N]
    const float* bary;            // [N,3]
    const float* smplx_rotation;  // [B,Nv,4] wxyz
    const float* smplx_scaling;   // [B,Nv,3]
    const float* smplx_opacity;   // [B,Nv,1]
    const float* uv_local_xyz;    // [B,N,3]
    const float* uv_rotation;     // [B,N,4] wxyz
    const float* uv_scaling;      // [B,N,3]
    const float* uv_opacity;      // [B,N,1]
    const float4* frames;         // [B,F,4]
};

__device__ __forceinline__ void load_rot3(const float* T, float* m)
{
    m[0] = T[0]; m[1] = T[1]; m[2] = T[2];
    m[3] = T[4]; m[4] = T[5]; m[5] = T[6];
    m[6] = T[8]; m[7] = T[9]; m[8] = T[10];
}

// One thread per (batch, Gaussian) of the concatenated cloud.
__global__ void __launch_bounds__(256)
deform_forward_kernel(const DeformArgs a, float* __restrict__ out_xyz, float* __restrict__ out_rot,
                      float* __restrict__ out_scale, float* __restrict__ out_opacity)
{
    const int P = a.Nv + a.N;
    const size_t t = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (t >= (size_t)a.B * P) return;
    const int b = (int)(t / P), i = (int)(t - (size_t)b * P);
    float x, y, z, sx, sy, sz, op;
    Q4 r;
    if (i < a.Nv) {
        const size_t s = (size_t)b * a.Nv + i;
        x = a.vertices[3 * s]; y = a.vertices[3 * s + 1]; z = a.vertices[3 * s + 2];
        float m[9];
        load_rot3(a.ver_transform + 16 * s, m);
        const Q4 p = rotmat_to_quat(m);
        const float4 qv = *reinterpret_cast<const float4*>(a.smplx_rotation + 4 * s);
        const Q4 q = {qv.x, qv.y, qv.z, qv.w};
        r = quat_mul(p, q);
        const float inv = 1.0f / fmaxf(sqrtf(r.w * r.w + r.x * r.x + r.y * r.y + r.z * r.z), 1e-12f);   // F.normalize
        r.w *= inv; r.x *= inv; r.y *= inv; r.z *= inv;
        sx = a.smplx_scaling[3 * s]; sy = a.smplx_scaling[3 * s + 1]; sz = a.smplx_scaling[3 * s + 2];
        op = a.smplx_opacity[s];
    } else {
        const int n = i - a.Nv;
        const size_t s = (size_t)b * a.N + n;
        const int f = a.binding[n];
        const float4* fr = a.frames + ((size_t)b * a.F + f) * 4;
        const float4 f0 = fr[0], f1 = fr[1], f2 = fr[2], f3 = fr[3];
        const float scale = f2.y;
        const float lx = a.uv_local_xyz[3 * s], ly = a.uv_local_xyz[3 * s + 1], lz = a.uv_local_xyz[3 * s + 2];
        const float* vb = a.vertices + (size_t)b * a.Nv * 3;
        const int i0 = a.faces[3 * f], i1 = a.faces[3 * f + 1], i2 = a.faces[3 * f + 2];
        const float b0 = a.bary[3 * n], b1 = a.bary[3 * n + 1], b2 = a.bary[3 * n + 2];
        const float cx = b0 * vb[3 * i0] + b1 * vb[3 * i1] + b2 * vb[3 * i2];
        const float cy = b0 * vb[3 * i0 + 1] + b1 * vb[3 * i1 + 1] + b2 * vb[3 * i2 + 1];
        const float cz = b0 * vb[3 * i0 + 2] + b1 * vb[3 * i1 + 2] + b2 * vb[3 * i2 + 2];
        x = (f0.x * lx + f0.y * ly + f0.z * lz) * scale + cx;
        y = (f0.w * lx + f1.x * ly + f1.y * lz) * scale + cy;
        z = (f1.z * lx + f1.w * ly + f2.x * lz) * scale + cz;
        const Q4 p = {f2.z, f2.w, f3.x, f3.y};
        const float4 qv = *reinterpret_cast<const float4*>(a.uv_rotation + 4 * s);
        const Q4 q = {qv.x, qv.y, qv.z, qv.w};
        r = quat_mul(p, q);
        sx = a.uv_scaling[3 * s] * scale; sy = a.uv_scaling[3 * s + 1] * scale; sz = a.uv_scaling[3 * s + 2] * scale;
        op = a.uv_opacity[s];
    }
    out_xyz[3 * t] = x; out_xyz[3 * t + 1] = y; out_xyz[3 * t + 2] = z;
    *reinterpret_cast<float4*>(out_rot + 4 * t) = make_float4(r.w, r.x, r.y, r.z);
    out_scale[3 * t] = sx; out_scale[3 * t + 1] = sy; out_scale[3 * t + 2] = sz;
    out_opacity[t] = op;
}

// gradients w.r.t. the per-Gaussian attributes; one thread per (batch, Gaussian)
__global__ void __launch_bounds__(256)
deform_backward_kernel(const DeformArgs a, const float* __restrict__ g_xyz, const float* __restrict__ g_rot,
                       const float* __restrict__ g_scale, float* __restrict__ d_smplx_rot,
                       float* __restrict__ d_uv_local, float* __restrict__ d_uv_rot, float* __restrict__ d_uv_scale)
{
    const int P = a.Nv + a.N;
    const size_t t = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (t >= (size_t)a.B * P) return;
    const int b = (int)(t / P), i = (int)(t - (size_t)b * P);
    const float4 gv = *reinterpret_cast<const float4*>(g_rot + 4 * t);
    Q4 g = {gv.x, gv.y, gv.z, gv.w};
    if (i < a.Nv) {
        const size_t s = (size_t)b * a.Nv + i;
        float m[9];
        load_rot3(a.ver_transform + 16 * s, m);
        const Q4 p = rotmat_to_quat(m);
        const float4 qv = *reinterpret_cast<const float4*>(a.smplx_rotation + 4 * s);
        const Q4 q = {qv.x, qv.y, qv.z, qv.w};
        const Q4 r = quat_mul(p, q);
        const float len = sqrtf(r.w * r.w + r.x * r.x + r.y * r.y + r.z * r.z);
        const float inv = 1.0f / fmaxf(len, 1e-12f);
        if (len > 1e-12f) {   // d(r/|r|) = (g - o (o.g)) / |r|
            const float ow = r.w * inv, ox = r.x * inv, oy = r.y * inv, oz = r.z * inv;
            const float og = ow * g.w + ox * g.x + oy * g.y + oz * g.z;
            g.w = (g.w - ow * og) * inv; g.x = (g.x - ox * og) * inv;
            g.y = (g.y - oy * og) * inv; g.z = (g.z - oz * og) * inv;
        } else {
            g.w *= inv; g.x *= inv; g.y *= inv; g.z *= inv;
        }
        const Q4 d = quat_mul_grad_q(p, g);
        *reinterpret_cast<float4*>(d_smplx_rot + 4 * s) = make_float4(d.w, d.x, d.y, d.z);
    } else {
        const int n = i - a.Nv;
        const size_t s = (size_t)b * a.N + n;
        const int f = a.binding[n];
        const float4* fr = a.frames + ((size_t)b * a.F + f) * 4;
        const float4 f0 = fr[0], f1 = fr[1], f2 = fr[2], f3 = fr[3];
        const float scale = f2.y;
        const float gx = g_xyz[3 * t] * scale, gy = g_xyz[3 * t + 1] * scale, gz = g_xyz[3 * t + 2] * scale;
        d_uv_local[3 * s] = f0.x * gx + f0.w * gy + f1.z * gz;       // R^T g
        d_uv_local[3 * s + 1] = f0.y * gx + f1.x * gy + f1.w * gz;
        d_uv_local[3 * s + 2] = f0.z * gx + f1.y * gy + f2.x * gz;
        const Q4 p = {f2.z, f2.w, f3.x, f3.y};
        const Q4 d = quat_mul_grad_q(p, g);
        *reinterpret_cast<float4*>(d_uv_rot + 4 * s) = make_float4(d.w, d.x, d.y, d.z);
        d_uv_scale[3 * s] = g_scale[3 * t] * scale;
        d_uv_scale[3 * s + 1] = g_scale[3 * t + 1] * scale;
        d_uv_scale[3 * s + 2] = g_scale[3 * t + 2] * scale;
    }
}

// [B,Nv,32] ++ [B,N,32] -> [B,Nv+N,32], 16 bytes per thread
__global__ void __launch_bounds__(256)
concat_features_kernel(int B, int Nv, int N, const float4* __restrict__ a, const float4* __restrict__ u, float4* __restrict__ out)
{
    const size_t rowq = NUM_CHANNELS / 4;
    const size_t P = (size_t)Nv + N;
    const size_t total = (size_t)B * P * rowq;
    for (size_t t = (size_t)blockIdx.x * 256 + threadIdx.x; t < total; t += (size_t)gridDim.x * 256) {
        const size_t row = t / rowq, c = t - row * rowq;
        const size_t b = row / P, i = row - b * P;
        out[t] = i < (size_t)Nv ? a[(b * Nv + i) * rowq + c] : u[(b * N + (i - Nv)) * rowq + c];
    }
}

// The backward of that concatenation (and of the opacity / SMPLX-scaling pass-through): the rasterizer's gradients
// [B,Nv+N,w] split into the contiguous per-attribute tensors autograd hands to the leaves.  As slices of the
// [B,Nv+N,w] tensors they would each be cloned by torch's gradient accumulation (four copy kernels, 29 us at the
// GUAVA size); this is one pass, 16 bytes per thread for the features.
__global__ void __launch_bounds__(256)
split_gradients_kernel(int B, int Nv, int N, const float* __restrict__ g_scale, const float* __restrict__ g_op,
                       const float4* __restrict__ g_feat, float* __restrict__ d_smplx_scale, float* __restrict__ d_smplx_op,
                       float4* __restrict__ d_smplx_feat, float* __restrict__ d_uv_op, float4* __restrict__ d_uv_feat)
{
    const size_t rowq = NUM_CHANNELS / 4;
    const size_t P = (size_t)Nv + N;
    const size_t first = (size_t)blockIdx.x * 256 + threadIdx.x, step = (size_t)gridDim.x * 256;
    if (g_feat != nullptr) {
        const size_t total = (size_t)B * P * rowq;
        for (size_t t = first; t < total; t += step) {
            const size_t row = t / rowq, c = t - row * rowq;
            const size_t b = row / P, i = row - b * P;
            const float4 v = g_feat[t];
            if (i < (size_t)Nv) { if (d_smplx_feat != nullptr) d_smplx_feat[(b * Nv + i) * rowq + c] = v; }
            else if (d_uv_feat != nullptr) d_uv_feat[(b * N + (i - Nv)) * rowq + c] = v;
        }
    }
    if (g_op != nullptr) {
        for (size_t t = first; t < (size_t)B * P; t += step) {
            const size_t b = t / P, i = t - b * P;
            if (i < (size_t)Nv) { if (d_smplx_op != nullptr) d_smplx_op[b * Nv + i] = g_op[t]; }
            else if (d_uv_op != nullptr) d_uv_op[b * N + (i - Nv)] = g_op[t];
        }
    }
    if (g_scale != nullptr && d_smplx_scale != nullptr) {
        for (size_t t = first; t < (size_t)B * Nv * 3; t += step) {
            const size_t b = t / ((size_t)Nv * 3), k = t - b * (size_t)Nv * 3;
            d_smplx_scale[t] = g_scale[b * P * 3 + k];
        }
    }
}

void launch_deform_forward(const DeformArgs& a, const float* smplx_features, const float* uv_features,
                           float* out_xyz, float* out_rot, float* out_scale, float* out_opacity, float* out_features,
                           cudaStream_t s)
{
    const size_t nf = (size_t)a.B * a.F, np = (size_t)a.B * (a.Nv + a.N);
    if (nf) face_frames_kernel<<<(unsigned)((nf + 255) / 256), 256, 0, s>>>(a.B, a.Nv, a.F, a.vertices, a.faces,
                                                                              const_cast<float4*>(a.frames));
    if (np) deform_forward_kernel<<<(unsigned)((np + 255) / 256), 256, 0, s>>>(a, out_xyz, out_rot, out_scale, out_opacity);
    int n = (nf ? 1 : 0) + (np ? 1 : 0);
    if (np && out_features != nullptr) {
        concat_features_kernel<<<sm_count() * 8, 256, 0, s>>>(a.B, a.Nv, a.N, (const float4*)smplx_features,
                                                       (const float4*)uv_features, (float4*)out_features);
        n++;
    }
    count_launch(n);
}

void launch_deform_backward(const DeformArgs& a, const float* g_xyz, const float* g_rot, const float* g_scale,
                            float* d_smplx_rot, float* d_uv_local, float* d_uv_rot, float* d_uv_scale, cudaStream_t s)
{
    const size_t np = (size_t)a.B * (a.Nv + a.N);
    if (!np) return;
    deform_backward_kernel<<<(unsigned)((np + 255) / 256), 256, 0, s>>>(a, g_xyz, g_rot, g_scale, d_smplx_rot, d_uv_local,
                                                                        d_uv_rot, d_uv_scale);
    count_launch();
}

// ------------------------------------------------------------------------------------ no-allocation variant
// Scratch of guava_deform_gaussians_forward_static: the per-(batch, face) frames (4 x float4 each) that the regular
// entry point takes from its caller.  File-scope device memory: every GPU that loads the library holds these 4 MiB
// until the process exits.  In HOST code `g_static_frames` is only a stand-in for the symbol, not a device address:
// the address a kernel may be given comes from cudaGetSymbolAddress, asked once per device with that device current.
static __device__ __align__(256) float4 g_static_frames[(size_t)GUAVA_DEFORM_STATIC_MAX_FACE_FRAMES * 4];

static cudaError_t static_frames_address(float4** out)
{
    static std::atomic<float4*> cached[64];   // per device ordinal
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    float4* p = (dev >= 0 && dev < 64) ? cached[dev].load(std::memory_order_acquire) : nullptr;
    if (p == nullptr) {
        void* addr = nullptr;
        e = cudaGetSymbolAddress(&addr, g_static_frames);   // the current device's instance of the array
        if (e != cudaSuccess) return e;
        if (addr == nullptr) return cudaErrorInvalidSymbol;
        p = static_cast<float4*>(addr);
        if (dev >= 0 && dev < 64) cached[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return cudaSuccess;
}

}  // namespace guava

using namespace guava;

extern "C" {

int guava_deform_gaussians_forward_static(int B, int Nv, int N, int F, const float* vertices, const float* ver_transform,
                                          const int* faces, const int* uv_binding_face, const float* uv_face_bary,
                                          const float* smplx_rotation, const float* smplx_scaling,
                                          const float* smplx_opacity, const float* smplx_features,
                                          const float* uv_local_xyz, const float* uv_rotation, const float* uv_scaling,
                                          const float* uv_opacity, const float* uv_features, float* out_xyz,
                                          float* out_rotation, float* out_scaling, float* out_opacity,
                                          float* out_features, void* stream)
{
    if (B < 0 || Nv < 0 || N < 0 || F < 0) return report_error(GUAVA_ERR_INVALID_ARGUMENT, "negative size");
    if ((size_t)B * ((size_t)Nv + N) == 0) return GUAVA_OK;
    if (N != 0 && (size_t)B * (size_t)F > (size_t)GUAVA_DEFORM_STATIC_MAX_FACE_FRAMES) {
        const std::string msg = "guava_deform_gaussians_forward_static: B*F = " + std::to_string((size_t)B * (size_t)F) +
                                " face frames exceed the static scratch (maximum " +
                                std::to_string((size_t)GUAVA_DEFORM_STATIC_MAX_FACE_FRAMES) +
                                "); use guava_deform_gaussians_forward with caller-provided face_frames";
        return report_error(GUAVA_ERR_INVALID_ARGUMENT, msg.c_str());
    }
    if (!vertices || !out_xyz || !out_rotation || !out_scaling || !out_opacity)
        return report_error(GUAVA_ERR_INVALID_ARGUMENT, "null required pointer");
    if (Nv && (!ver_transform || !smplx_rotation || !smplx_scaling || !smplx_opacity))
        return report_error(GUAVA_ERR_INVALID_ARGUMENT, "null required pointer");
    if (N && (!faces || !uv_binding_face || !uv_face_bary || !uv_local_xyz || !uv_rotation || !uv_scaling || !uv_opacity))
        return report_error(GUAVA_ERR_INVALID_ARGUMENT, "null required pointer");
    if (out_features && ((Nv && !smplx_features) || (N && !uv_features)))
        return report_error(GUAVA_ERR_INVALID_ARGUMENT, "null required pointer");
    float4* frames = nullptr;
    const cudaError_t e = static_frames_address(&frames);
    if (e != cudaSuccess)
        return report_error(GUAVA_ERR_CUDA, (std::string("cudaGetSymbolAddress(g_static_frames): ") + cudaGetErrorString(e)).c_str());
    // Nothing of the scratch has to start at zero: face_frames_kernel writes all B*F frames of this call before
    // deform_forward_kernel (same stream) reads any; what an earlier call left beyond them is never read.
    DeformArgs a{B, Nv, N, N ? F : 0, vertices, ver_transform, faces, uv_binding_face, uv_face_bary, smplx_rotation,
                 smplx_scaling, smplx_opacity, uv_local_xyz, uv_rotation, uv_scaling, uv_opacity, frames};
    launch_deform_forward(a, smplx_features, uv_features, out_xyz, out_rotation, out_scaling, out_opacity, out_features,
                          (cudaStream_t)stream);
    const cudaError_t le = cudaGetLastError();
    if (le != cudaSuccess) return report_error(GUAVA_ERR_CUDA, cudaGetErrorString(le));
    return GUAVA_OK;
}

int guava_deform_gaussians_forward(int B, int Nv, int N, int F, const float* vertices, const float* ver_transform,
                                   const int* faces, const int* uv_binding_face, const float* uv_face_bary,
                                   const float* smplx_rotation, const float* smplx_scaling, const float* smplx_opacity,
                                   const float* smplx_features, const float* uv_local_xyz, const float* uv_rotation,
                                   const float* uv_scaling, const float* uv_opacity, const float* uv_features,
                                   float* face_frames, float* out_xyz, float* out_rotation, float* out_scaling,
                                   float* out_opacity, float* out_features, void* stream)
{
    if (B < 0 || Nv < 0 || N < 0 || F < 0) return 1;
    if ((size_t)B * ((size_t)Nv + N) == 0) return 0;
    if (!vertices || !out_xyz || !out_rotation || !out_scaling || !out_opacity) return 1;
    if (Nv && (!ver_transform || !smplx_rotation || !smplx_scaling || !smplx_opacity)) return 1;
    if (N && (!faces || !uv_binding_face || !uv_face_bary || !uv_local_xyz || !uv_rotation || !uv_scaling ||
              !uv_opacity || !face_frames))
        return 1;
    if (out_features && ((Nv && !smplx_features) || (N && !uv_features))) return 1;
    DeformArgs a{B, Nv, N, N ? F : 0, vertices, ver_transform, faces, uv_binding_face, uv_face_bary, smplx_rotation,
                 smplx_scaling, smplx_opacity, uv_local_xyz, uv_rotation, uv_scaling, uv_opacity,
                 (const float4*)face_frames};
    a.F = F;
    launch_deform_forward(a, smplx_features, uv_features, out_xyz, out_rotation, out_scaling, out_opacity, out_features,
                          (cudaStream_t)stream);
    return cudaGetLastError() == cudaSuccess ? 0 : 2;
}

int guava_deform_gaussians_backward(int B, int Nv, int N, int F, const float* ver_transform, const int* uv_binding_face,
                                    const float* smplx_rotation, const float* face_frames, const float* dL_dxyz,
                                    const float* dL_drotation, const float* dL_dscaling, float* dL_dsmplx_rotation,
                                    float* dL_duv_local_xyz, float* dL_duv_rotation, float* dL_duv_scaling, void* stream)
{
    if (B < 0 || Nv < 0 || N < 0 || F < 0) return 1;
    if ((size_t)B * ((size_t)Nv + N) == 0) return 0;
    if (!dL_dxyz || !dL_drotation || !dL_dscaling) return 1;
    if (Nv && (!ver_transform || !smplx_rotation || !dL_dsmplx_rotation)) return 1;
    if (N && (!uv_binding_face || !face_frames || !dL_duv_local_xyz || !dL_duv_rotation || !dL_duv_scaling)) return 1;
    DeformArgs a{};
    a.B = B; a.Nv = Nv; a.N = N; a.F = F;
    a.ver_transform = ver_transform; a.binding = uv_binding_face; a.smplx_rotation = smplx_rotation;
    a.frames = (const float4*)face_frames;
    launch_deform_backward(a, dL_dxyz, dL_drotation, dL_dscaling, dL_dsmplx_rotation, dL_duv_local_xyz, dL_duv_rotation,
                           dL_duv_scaling, (cudaStream_t)stream);
    return cudaGetLastError() == cudaSuccess ? 0 : 2;
}

int guava_deform_split_gradients(int B, int Nv, int N, const float* dL_dscaling, const float* dL_dopacity,
                                 const float* dL_dfeatures, float* dL_dsmplx_scaling, float* dL_dsmplx_opacity,
                                 float* dL_dsmplx_features, float* dL_duv_opacity, float* dL_duv_features, void* stream)
{
    if (B < 0 || Nv < 0 || N < 0) return report_error(GUAVA_ERR_INVALID_ARGUMENT, "negative size");
    if ((size_t)B * ((size_t)Nv + N) == 0) return GUAVA_OK;
    if (((uintptr_t)dL_dfeatures | (uintptr_t)dL_dsmplx_features | (uintptr_t)dL_duv_features) & 15)
        return report_error(GUAVA_ERR_INVALID_ARGUMENT, "feature gradients must be 16-byte aligned");
    split_gradients_kernel<<<sm_count() * 8, 256, 0, (cudaStream_t)stream>>>(
        B, Nv, N, dL_dscaling, dL_dopacity, (const float4*)dL_dfeatures, dL_dsmplx_scaling, dL_dsmplx_opacity,
        (float4*)dL_dsmplx_features, dL_duv_opacity, (float4*)dL_duv_features);
    count_launch();
    return cudaGetLastError() == cudaSuccess ? GUAVA_OK : report_error(GUAVA_ERR_CUDA, "split_gradients_kernel launch failed");
}

}  // extern "C"
