// Compiled torch binding of the C ABI (include/guava_rasterizer.h): the module the reference calls `_C`.
//
// Replaces DGR32/ext.cpp:15-18 and DGR32/rasterize_points.cu:35-244 (RasterizeGaussiansCUDA,
// RasterizeGaussiansBackwardCUDA, markVisible) -- same function names, same positional arguments, same
// returned tuples -- and adds the batched pair.  Compiled by g++ (no nvcc: torch headers stay out of the .cu
// files), linked against libguava_rasterizer_b200.so; torch is plumbing here: it owns the tensors, the three
// resizable byte buffers (resizeFunctional, rasterize_points.cu:27-33) and the current stream.  Differences from
// the reference binding, all benign: a device guard and torch's current stream instead of the legacy default
// stream; outputs are not pre-filled (every pixel / radius is written by the kernels); one
// zero fill for all gradient tensors instead of ten.
#include <torch/extension.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>

#include <string>
#include <tuple>

#include "guava_rasterizer.h"

namespace {

constexpr int NUM_CHANNELS = GUAVA_NUM_CHANNELS;

char* resize_cb(void* user, size_t bytes)
{
    auto* t = static_cast<torch::Tensor*>(user);
    try {
        t->resize_({static_cast<int64_t>(bytes)});
    } catch (...) {
        return nullptr;   // out of memory: the C side reports GUAVA_ERR_ALLOC
    }
    return reinterpret_cast<char*>(t->data_ptr());
}

void check(int rc)
{
    if (rc == GUAVA_OK) return;
    const std::string msg = guava_last_error();
    if (rc == GUAVA_ERR_ALLOC) throw c10::OutOfMemoryError({__func__, __FILE__, static_cast<uint32_t>(__LINE__)}, msg);
    throw std::runtime_error(msg);
}

// device pointer, or NULL for undefined / empty tensors (the reference passes torch.Tensor([]) for absent inputs)
const float* fptr(const torch::Tensor& t) { return (t.defined() && t.numel() != 0) ? t.data_ptr<float>() : nullptr; }

torch::Tensor f32c(const torch::Tensor& t)
{
    if (!t.defined() || t.numel() == 0) return t;
    TORCH_CHECK(t.scalar_type() == torch::kFloat32, "expected a float32 tensor");
    return t.contiguous();
}

torch::Tensor aligned16(const torch::Tensor& t)
{
    torch::Tensor c = f32c(t);
    if (c.defined() && c.numel() != 0 && (reinterpret_cast<uintptr_t>(c.data_ptr()) & 15u)) c = c.clone();
    return c;
}

void* current_stream(const torch::Tensor& t) { return at::cuda::getCurrentCUDAStream(t.device().index()).stream(); }

// rasterize_points.cu:35-124
std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
RasterizeGaussiansCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                       const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                       const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                       const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy, const int image_height,
                       const int image_width, const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
                       const bool prefiltered, const bool antialiasing, const bool debug)
{
    if (means3D.ndimension() != 2 || means3D.size(1) != 3) AT_ERROR("means3D must have dimensions (num_points, 3)");
    TORCH_CHECK(means3D.is_cuda(), "guava_renderer_b200: tensors must live on a CUDA device (there is no CPU path)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int P = static_cast<int>(means3D.size(0));
    const int H = image_height, W = image_width;
    auto fopts = means3D.options().dtype(torch::kFloat32);
    auto bopts = means3D.options().dtype(torch::kByte);
    int M = 0;
    if (sh.defined() && sh.numel() != 0 && sh.dim() > 1) M = static_cast<int>(sh.size(1));
    if (P == 0) {
        return std::make_tuple(0, torch::zeros({NUM_CHANNELS, H, W}, fopts), torch::zeros({0}, means3D.options().dtype(torch::kInt32)),
                               torch::empty({0}, bopts), torch::empty({0}, bopts), torch::empty({0}, bopts),
                               torch::zeros({1, H, W}, fopts));
    }
    if (!colors.defined() || colors.numel() == 0)
        throw std::runtime_error("For non-RGB, provide precomputed Gaussian colors!");   // rasterizer_impl.cu:244-247
    torch::Tensor out_color = torch::empty({NUM_CHANNELS, H, W}, fopts);
    torch::Tensor out_invdepth = torch::empty({1, H, W}, fopts);
    torch::Tensor radii = torch::empty({P}, means3D.options().dtype(torch::kInt32));
    torch::Tensor geom = torch::empty({0}, bopts), binning = torch::empty({0}, bopts), img = torch::empty({0}, bopts);
    const torch::Tensor bg = f32c(background), m = f32c(means3D), c = aligned16(colors), o = f32c(opacity), s = f32c(scales),
                        r = f32c(rotations), cov = f32c(cov3D_precomp), vm = f32c(viewmatrix), pm = f32c(projmatrix);
    const float* shp = (sh.defined() && sh.numel() != 0 && sh.is_cuda()) ? sh.data_ptr<float>() : nullptr;
    int rendered = 0;
    check(guava_rasterizer_forward(resize_cb, &geom, resize_cb, &binning, resize_cb, &img, P, degree, M, fptr(bg), W, H, fptr(m),
                                   shp, fptr(c), fptr(o), fptr(s), scale_modifier, fptr(r), fptr(cov), fptr(vm), fptr(pm),
                                   campos.defined() && campos.is_cuda() && campos.numel() ? campos.data_ptr<float>() : nullptr,
                                   tan_fovx, tan_fovy, prefiltered ? 1 : 0, out_color.data_ptr<float>(),
                                   out_invdepth.data_ptr<float>(), antialiasing ? 1 : 0, radii.data_ptr<int>(), debug ? 1 : 0,
                                   current_stream(means3D), &rendered));
    return std::make_tuple(rendered, out_color, radii, geom, binning, img, out_invdepth);
}

// rasterize_points.cu:126-223
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
RasterizeGaussiansBackwardCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                               const torch::Tensor& colors, const torch::Tensor& opacities, const torch::Tensor& scales,
                               const torch::Tensor& rotations, const float scale_modifier, const torch::Tensor& cov3D_precomp,
                               const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix, const float tan_fovx,
                               const float tan_fovy, const torch::Tensor& dL_dout_color, const torch::Tensor& dL_dout_invdepth,
                               const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
                               const torch::Tensor& geomBuffer, const int R, const torch::Tensor& binningBuffer,
                               const torch::Tensor& imageBuffer, const bool antialiasing, const bool debug)
{
    TORCH_CHECK(means3D.is_cuda(), "guava_renderer_b200: tensors must live on a CUDA device (there is no CPU path)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int64_t P = means3D.size(0);
    const int H = static_cast<int>(dL_dout_color.size(1)), W = static_cast<int>(dL_dout_color.size(2));
    int M = 0;
    if (sh.defined() && sh.numel() != 0 && sh.dim() > 1) M = static_cast<int>(sh.size(1));
    const bool has_inv = dL_dout_invdepth.defined() && dL_dout_invdepth.numel() != 0 && dL_dout_invdepth.size(0) != 0;
    auto fopts = means3D.options().dtype(torch::kFloat32);
    // one zero fill for all gradient tensors (the reference issues ten, rasterize_points.cu:163-182); dL_dcolors
    // first and dL_drotations at 48 P floats: both stay 16-byte aligned
    const int64_t widths[9] = {32, 3, 1, 3, 6, 3, 4, 4, 1};
    int64_t total = 0;
    for (int i = 0; i < (has_inv ? 9 : 8); i++) total += widths[i];
    torch::Tensor flat = torch::zeros({P * total}, fopts);
    torch::Tensor part[9];
    int64_t off = 0;
    for (int i = 0; i < (has_inv ? 9 : 8); i++) {
        part[i] = flat.narrow(0, off, P * widths[i]).view({P, widths[i]});
        off += P * widths[i];
    }
    torch::Tensor dL_dcolors = part[0], dL_dmeans2D = part[1], dL_dopacity = part[2], dL_dmeans3D = part[3],
                  dL_dcov3D = part[4], dL_dscales = part[5], dL_drotations = part[6], dL_dconic = part[7];
    torch::Tensor dL_dsh = torch::zeros({P, M, 3}, fopts);
    if (P != 0) {
        const torch::Tensor bg = f32c(background), m = f32c(means3D), c = aligned16(colors), o = f32c(opacities),
                            s = f32c(scales), r = f32c(rotations), cov = f32c(cov3D_precomp), vm = f32c(viewmatrix),
                            pm = f32c(projmatrix), gc = f32c(dL_dout_color), gi = has_inv ? f32c(dL_dout_invdepth) : torch::Tensor(),
                            rad = radii.contiguous();
        check(guava_rasterizer_backward(
            static_cast<int>(P), degree, M, R, fptr(bg), W, H, fptr(m), nullptr, fptr(c), fptr(o), fptr(s), scale_modifier, fptr(r),
            fptr(cov), fptr(vm), fptr(pm), nullptr, tan_fovx, tan_fovy, rad.data_ptr<int>(),
            reinterpret_cast<char*>(geomBuffer.data_ptr()), reinterpret_cast<char*>(binningBuffer.data_ptr()),
            reinterpret_cast<char*>(imageBuffer.data_ptr()), fptr(gc), has_inv ? fptr(gi) : nullptr,
            dL_dmeans2D.data_ptr<float>(), dL_dconic.data_ptr<float>(), dL_dopacity.data_ptr<float>(), dL_dcolors.data_ptr<float>(),
            has_inv ? part[8].data_ptr<float>() : nullptr, dL_dmeans3D.data_ptr<float>(), dL_dcov3D.data_ptr<float>(), nullptr,
            dL_dscales.data_ptr<float>(), dL_drotations.data_ptr<float>(), antialiasing ? 1 : 0, debug ? 1 : 0,
            current_stream(means3D)));
    }
    return std::make_tuple(dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations);
}

// rasterize_points.cu:225-244
torch::Tensor markVisible(const torch::Tensor& means3D, const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix)
{
    TORCH_CHECK(means3D.is_cuda(), "guava_renderer_b200: tensors must live on a CUDA device (there is no CPU path)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int P = static_cast<int>(means3D.size(0));
    torch::Tensor present = torch::zeros({P}, means3D.options().dtype(torch::kBool));
    if (P != 0) {
        const torch::Tensor m = f32c(means3D), v = f32c(viewmatrix), p = f32c(projmatrix);
        check(guava_rasterizer_mark_visible(P, fptr(m), fptr(v), fptr(p), reinterpret_cast<uint8_t*>(present.data_ptr<bool>()),
                                            current_stream(means3D)));
    }
    return present;
}

// elements per view of a [V,P,w] tensor, 0 when one [P,w] tensor is shared by all views
size_t vstride(const torch::Tensor& t, int64_t V, int64_t per_view)
{
    if (t.numel() == per_view) return 0;
    TORCH_CHECK(t.numel() == V * per_view, "tensor must be [P,...] (shared) or [V,P,...] (per view)");
    return static_cast<size_t>(per_view);
}

// Batched multi-view entry (guava_rasterizer_forward_batch).  refiner_output: undefined or bf16 [V,H,W,32];
// status_words: 0 or the address of the caller's pinned, device-mapped uint32[2].
std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
RasterizeGaussiansBatchCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                            const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                            const float scale_modifier, const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix,
                            const torch::Tensor& tanfov, const int image_height, const int image_width, const bool antialiasing,
                            const bool debug, const c10::optional<torch::Tensor>& refiner_output, const int64_t status_words,
                            const bool no_host_wait, const bool forward_only, const c10::optional<torch::Tensor>& zero_fill)
{
    TORCH_CHECK(means3D.is_cuda(), "guava_renderer_b200: tensors must live on a CUDA device (there is no CPU path)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int64_t V = viewmatrix.size(0);
    const int64_t P = means3D.size(-2);
    const int H = image_height, W = image_width;
    auto fopts = means3D.options().dtype(torch::kFloat32);
    auto bopts = means3D.options().dtype(torch::kByte);
    torch::Tensor out_color = torch::empty({V, NUM_CHANNELS, H, W}, fopts);
    torch::Tensor out_invdepth = torch::empty({V, 1, H, W}, fopts);
    torch::Tensor radii = torch::empty({V, P}, means3D.options().dtype(torch::kInt32));
    torch::Tensor geom = torch::empty({0}, bopts), binning = torch::empty({0}, bopts), img = torch::empty({0}, bopts);
    guava_call_options co{};
    co.struct_size = sizeof(co);
    if (refiner_output.has_value() && refiner_output->defined()) {
        const torch::Tensor& aux = *refiner_output;
        TORCH_CHECK(aux.scalar_type() == torch::kBFloat16 && aux.is_contiguous() && aux.dim() == 4 && aux.size(0) == V &&
                        aux.size(1) == H && aux.size(2) == W && aux.size(3) == NUM_CHANNELS,
                    "refiner output must be a contiguous bfloat16 [V,H,W,32] tensor");
        if (V * P == 0) aux.zero_();
        co.refiner_output_nhwc_bf16 = aux.data_ptr();
    }
    co.status_words = reinterpret_cast<void*>(status_words);
    co.no_host_wait = no_host_wait ? 1 : 0;
    co.forward_only = forward_only ? 1 : 0;
    if (zero_fill.has_value() && zero_fill->defined()) {   // e.g. the summed backward's [P,43] output: zeroed under the tile scan
        TORCH_CHECK(zero_fill->is_cuda() && zero_fill->is_contiguous() && zero_fill->device() == means3D.device(),
                    "zero_fill must be a contiguous tensor on the inputs' device");
        if (V * P == 0) zero_fill->zero_();
        co.zero_fill = zero_fill->data_ptr();
        co.zero_fill_bytes = static_cast<size_t>(zero_fill->numel()) * zero_fill->element_size();
    }
    if (V * P == 0) {
        out_color.zero_();
        out_invdepth.zero_();
        return std::make_tuple(0, out_color, radii, geom, binning, img, out_invdepth);
    }
    const torch::Tensor bg = f32c(background), m = f32c(means3D), c = aligned16(colors), o = f32c(opacity), s = f32c(scales),
                        r = f32c(rotations), vm = f32c(viewmatrix), pm = f32c(projmatrix), tf = f32c(tanfov);
    int rendered = 0;
    check(guava_rasterizer_forward_batch(
        resize_cb, &geom, resize_cb, &binning, resize_cb, &img, static_cast<int>(V), static_cast<int>(P), fptr(bg), W, H, fptr(m),
        vstride(m, V, P * 3), fptr(c), vstride(c, V, P * NUM_CHANNELS), fptr(o), vstride(o, V, P), fptr(s), vstride(s, V, P * 3),
        scale_modifier, fptr(r), vstride(r, V, P * 4), fptr(vm), fptr(pm), fptr(tf), out_color.data_ptr<float>(),
        out_invdepth.data_ptr<float>(), antialiasing ? 1 : 0, radii.data_ptr<int>(), debug ? 1 : 0, current_stream(means3D),
        &rendered, &co));
    return std::make_tuple(rendered, out_color, radii, geom, binning, img, out_invdepth);
}

// Backward of the batched entry: (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dscales,
// dL_drotations), each [V,P,w], views of one zero-filled block.
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
RasterizeGaussiansBatchBackwardCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                                    const torch::Tensor& colors, const torch::Tensor& opacities, const torch::Tensor& scales,
                                    const torch::Tensor& rotations, const float scale_modifier, const torch::Tensor& viewmatrix,
                                    const torch::Tensor& projmatrix, const torch::Tensor& tanfov,
                                    const torch::Tensor& dL_dout_color, const c10::optional<torch::Tensor>& dL_dout_invdepth,
                                    const torch::Tensor& geomBuffer, const int R, const torch::Tensor& binningBuffer,
                                    const torch::Tensor& imageBuffer, const bool antialiasing, const bool debug,
                                    const c10::optional<torch::Tensor>& zeroed_block)
{
    TORCH_CHECK(means3D.is_cuda(), "guava_renderer_b200: tensors must live on a CUDA device (there is no CPU path)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int64_t V = viewmatrix.size(0);
    const int64_t P = means3D.size(-2);
    const int H = static_cast<int>(dL_dout_color.size(-2)), W = static_cast<int>(dL_dout_color.size(-1));
    const bool has_inv = dL_dout_invdepth.has_value() && dL_dout_invdepth->defined();
    auto fopts = means3D.options().dtype(torch::kFloat32);
    const int64_t widths[9] = {32, 3, 1, 3, 6, 3, 4, 4, 1};
    const int64_t n = V * P;
    int64_t total = 0;
    for (int i = 0; i < (has_inv ? 9 : 8); i++) total += widths[i];
    // zeroed_block: n * 57 floats the forward of this call zero-filled under its tile scan (guava_call_options::zero_fill)
    const bool given = zeroed_block.has_value() && zeroed_block->defined() && zeroed_block->numel() >= n * total &&
                       zeroed_block->is_contiguous() && zeroed_block->scalar_type() == torch::kFloat32;
    torch::Tensor flat = given ? zeroed_block->view({-1}) : torch::zeros({n * total}, fopts);
    torch::Tensor part[9];
    int64_t off = 0;
    for (int i = 0; i < (has_inv ? 9 : 8); i++) {
        part[i] = flat.narrow(0, off, n * widths[i]).view({V, P, widths[i]});
        off += n * widths[i];
    }
    if (n != 0) {
        const torch::Tensor bg = f32c(background), m = f32c(means3D), c = aligned16(colors), o = f32c(opacities),
                            s = f32c(scales), r = f32c(rotations), vm = f32c(viewmatrix), pm = f32c(projmatrix), tf = f32c(tanfov),
                            gc = f32c(dL_dout_color), gi = has_inv ? f32c(*dL_dout_invdepth) : torch::Tensor(),
                            rad = radii.contiguous();
        check(guava_rasterizer_backward_batch(
            static_cast<int>(V), static_cast<int>(P), R, fptr(bg), W, H, fptr(m), vstride(m, V, P * 3), fptr(c),
            vstride(c, V, P * NUM_CHANNELS), fptr(o), vstride(o, V, P), fptr(s), vstride(s, V, P * 3), scale_modifier, fptr(r),
            vstride(r, V, P * 4), fptr(vm), fptr(pm), fptr(tf), rad.data_ptr<int>(), reinterpret_cast<char*>(geomBuffer.data_ptr()),
            reinterpret_cast<char*>(binningBuffer.data_ptr()), reinterpret_cast<char*>(imageBuffer.data_ptr()), fptr(gc),
            has_inv ? fptr(gi) : nullptr, part[1].data_ptr<float>(), part[7].data_ptr<float>(), part[2].data_ptr<float>(),
            part[0].data_ptr<float>(), has_inv ? part[8].data_ptr<float>() : nullptr, part[3].data_ptr<float>(),
            part[4].data_ptr<float>(), part[5].data_ptr<float>(), part[6].data_ptr<float>(), antialiasing ? 1 : 0, debug ? 1 : 0,
            current_stream(means3D)));
    }
    return std::make_tuple(part[1], part[0], part[2], part[3], part[4], part[5], part[6]);
}

// Backward of the batched entry with the sum over the views taken inside the kernels
// (guava_rasterizer_backward_batch_summed): fills `packed` [P,43] = (means3D 3, scales 3, rotations 4, opacity 1,
// features 32) -- the all-reduce payload of the training step -- which the caller has zero-filled.
void RasterizeGaussiansBatchBackwardSummedCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                                               const torch::Tensor& colors, const torch::Tensor& opacities, const torch::Tensor& scales,
                                               const torch::Tensor& rotations, const float scale_modifier,
                                               const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix,
                                               const torch::Tensor& tanfov, const torch::Tensor& dL_dout_color,
                                               const c10::optional<torch::Tensor>& dL_dout_invdepth, const torch::Tensor& geomBuffer,
                                               const int R, const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer,
                                               const bool antialiasing, const bool debug, torch::Tensor& packed)
{
    TORCH_CHECK(means3D.is_cuda(), "guava_renderer_b200: tensors must live on a CUDA device (there is no CPU path)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int64_t V = viewmatrix.size(0);
    const int64_t P = means3D.size(-2);
    const int H = static_cast<int>(dL_dout_color.size(-2)), W = static_cast<int>(dL_dout_color.size(-1));
    TORCH_CHECK(packed.is_cuda() && packed.scalar_type() == torch::kFloat32 && packed.is_contiguous() && packed.dim() == 2 &&
                    packed.size(0) == P && packed.size(1) == 43,
                "packed must be a contiguous float32 [P,43] CUDA tensor");
    const bool has_inv = dL_dout_invdepth.has_value() && dL_dout_invdepth->defined();
    if (V * P == 0) return;
    const torch::Tensor bg = f32c(background), m = f32c(means3D), c = aligned16(colors), o = f32c(opacities), s = f32c(scales),
                        r = f32c(rotations), vm = f32c(viewmatrix), pm = f32c(projmatrix), tf = f32c(tanfov),
                        gc = f32c(dL_dout_color), gi = has_inv ? f32c(*dL_dout_invdepth) : torch::Tensor(), rad = radii.contiguous();
    float* pk = packed.data_ptr<float>();
    check(guava_rasterizer_backward_batch_summed(
        static_cast<int>(V), static_cast<int>(P), R, fptr(bg), W, H, fptr(m), vstride(m, V, P * 3), fptr(c),
        vstride(c, V, P * NUM_CHANNELS), fptr(o), vstride(o, V, P), fptr(s), vstride(s, V, P * 3), scale_modifier, fptr(r),
        vstride(r, V, P * 4), fptr(vm), fptr(pm), fptr(tf), rad.data_ptr<int>(), reinterpret_cast<char*>(geomBuffer.data_ptr()),
        reinterpret_cast<char*>(binningBuffer.data_ptr()), reinterpret_cast<char*>(imageBuffer.data_ptr()), fptr(gc),
        has_inv ? fptr(gi) : nullptr, pk, pk + 3, pk + 6, pk + 10, pk + 11, 43, antialiasing ? 1 : 0, debug ? 1 : 0,
        current_stream(means3D)));
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m)
{
    // ext.cpp:15-18
    m.def("rasterize_gaussians", &RasterizeGaussiansCUDA);
    m.def("rasterize_gaussians_backward", &RasterizeGaussiansBackwardCUDA);
    m.def("mark_visible", &markVisible);
    // SURVEY.md 8(f) row 1
    m.def("rasterize_gaussians_batch", &RasterizeGaussiansBatchCUDA);
    m.def("rasterize_gaussians_batch_backward", &RasterizeGaussiansBatchBackwardCUDA);
    m.def("rasterize_gaussians_batch_backward_summed", &RasterizeGaussiansBatchBackwardSummedCUDA);
}
