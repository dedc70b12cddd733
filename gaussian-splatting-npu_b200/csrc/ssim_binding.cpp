// ssim_binding.cpp -- torch-extension layer of the loss stage over the C ABI (include/gsr_loss.h).
//
// Keeps the reference's two pybind surfaces for this stage:
//   module fused_ssim_cuda  (FS/ext.cpp:4-7, FS = /root/reference/submodules/fused-ssim):
//       fusedssim(C1, C2, img1, img2, train) -> (ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12)   FS/ssim.cu:373-409
//       fusedssim_backward(C1, C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12)       FS/ssim.cu:411-444
//   diff_gaussian_rasterization._C  (what /root/reference/utils/loss_utils.py:16-19,24-37 imports when the
//   accelerated rasterizer is installed):
//       fusedssim(C1, C2, img1, img2) -> ssim_map ;  fusedssim_backward(C1, C2, img1, img2, dL_dmap) -> dL_dimg1
// plus the fused training loss (train.py:118-126) as l1_ssim_loss_forward / l1_ssim_loss_backward.
// Compiled twice: into _C (bind_ssim called from torch_binding.cpp) and, with -DGSR_SSIM_STANDALONE, as
// the top-level module fused_ssim_cuda.
#include <torch/extension.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <stdexcept>
#include <tuple>
#include "gsr.h"
#include "gsr_loss.h"

namespace {

void check(int status) {
    if (status != GSR_OK) throw std::runtime_error(gsr_last_error());
}

torch::Tensor as_f32_cuda(const torch::Tensor& t, const char* name) {
    TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor (there is no CPU fallback)");
    torch::Tensor c = t.contiguous();
    if (c.scalar_type() != torch::kFloat32) c = c.to(torch::kFloat32);
    return c;
}

struct Dims { int planes, H, W; };
Dims dims_of(const torch::Tensor& img, const char* name) {
    TORCH_CHECK(img.dim() >= 2 && img.dim() <= 4, name, " must be [B,CH,H,W], [CH,H,W] or [H,W]");
    const int64_t H = img.size(-2), W = img.size(-1);
    const int64_t planes = (H * W) ? img.numel() / (H * W) : 0;
    return Dims{(int)planes, (int)H, (int)W};
}

}  // namespace

std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
fusedssim(float C1, float C2, torch::Tensor& img1, torch::Tensor& img2, bool train) {
    torch::Tensor a = as_f32_cuda(img1, "img1"), b = as_f32_cuda(img2, "img2");
    TORCH_CHECK(a.sizes() == b.sizes(), "img1 and img2 must have the same shape");
    const c10::cuda::CUDAGuard guard(a.device());
    const Dims d = dims_of(a, "img1");
    torch::Tensor map = torch::empty_like(a);
    // the reference returns torch::empty(0) (CPU) placeholders when train is false (ssim.cu:388-390)
    torch::Tensor d0 = train ? torch::empty_like(a) : torch::empty(0);
    torch::Tensor d1 = train ? torch::empty_like(a) : torch::empty(0);
    torch::Tensor d2 = train ? torch::empty_like(a) : torch::empty(0);
    check(gsr_ssim_forward(d.planes, d.H, d.W, C1, C2, a.data_ptr<float>(), b.data_ptr<float>(), map.data_ptr<float>(),
                           train ? d0.data_ptr<float>() : nullptr, train ? d1.data_ptr<float>() : nullptr,
                           train ? d2.data_ptr<float>() : nullptr, at::cuda::getCurrentCUDAStream()));
    return std::make_tuple(map, d0, d1, d2);
}

torch::Tensor fusedssim_backward(float C1, float C2, torch::Tensor& img1, torch::Tensor& img2, torch::Tensor& dL_dmap,
                                 torch::Tensor& dm_dmu1, torch::Tensor& dm_dsigma1_sq, torch::Tensor& dm_dsigma12) {
    (void)C1; (void)C2;   // the derivative maps already carry them (as in the reference kernel, which ignores both)
    torch::Tensor a = as_f32_cuda(img1, "img1"), b = as_f32_cuda(img2, "img2"), g = as_f32_cuda(dL_dmap, "dL_dmap");
    torch::Tensor d0 = as_f32_cuda(dm_dmu1, "dm_dmu1"), d1 = as_f32_cuda(dm_dsigma1_sq, "dm_dsigma1_sq"),
                  d2 = as_f32_cuda(dm_dsigma12, "dm_dsigma12");
    TORCH_CHECK(a.numel() == b.numel() && a.numel() == g.numel() && a.numel() == d0.numel() && a.numel() == d1.numel() &&
                a.numel() == d2.numel(), "fusedssim_backward: all tensors must have img1's shape");
    const c10::cuda::CUDAGuard guard(a.device());
    const Dims d = dims_of(a, "img1");
    torch::Tensor out = torch::empty_like(a);
    check(gsr_ssim_backward(d.planes, d.H, d.W, a.data_ptr<float>(), b.data_ptr<float>(), g.data_ptr<float>(),
                            d0.data_ptr<float>(), d1.data_ptr<float>(), d2.data_ptr<float>(), out.data_ptr<float>(),
                            at::cuda::getCurrentCUDAStream()));
    return out;
}

// the 4-/5-argument forms utils/loss_utils.py:24-37 calls (no saved derivative maps: backward recomputes them)
torch::Tensor fusedssim_map_only(float C1, float C2, torch::Tensor& img1, torch::Tensor& img2) {
    return std::get<0>(fusedssim(C1, C2, img1, img2, false));
}
torch::Tensor fusedssim_backward_recompute(float C1, float C2, torch::Tensor& img1, torch::Tensor& img2, torch::Tensor& dL_dmap) {
    torch::Tensor a = as_f32_cuda(img1, "img1"), b = as_f32_cuda(img2, "img2");
    const c10::cuda::CUDAGuard guard(a.device());
    const Dims d = dims_of(a, "img1");
    torch::Tensor d0 = torch::empty_like(a), d1 = torch::empty_like(a), d2 = torch::empty_like(a);
    check(gsr_ssim_forward(d.planes, d.H, d.W, C1, C2, a.data_ptr<float>(), b.data_ptr<float>(), nullptr,
                           d0.data_ptr<float>(), d1.data_ptr<float>(), d2.data_ptr<float>(), at::cuda::getCurrentCUDAStream()));
    return fusedssim_backward(C1, C2, a, b, dL_dmap, d0, d1, d2);
}

// loss[3] = (loss, mean L1, mean ssim) + the three derivative maps (empty when !train)
std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
l1_ssim_loss_forward(torch::Tensor& image, torch::Tensor& gt, float lambda_dssim, float C1, float C2, bool train) {
    torch::Tensor a = as_f32_cuda(image, "image"), b = as_f32_cuda(gt, "gt");
    TORCH_CHECK(a.sizes() == b.sizes(), "image and gt must have the same shape");
    const c10::cuda::CUDAGuard guard(a.device());
    const Dims d = dims_of(a, "image");
    torch::Tensor loss = torch::empty({3}, a.options());
    torch::Tensor d0 = train ? torch::empty_like(a) : torch::empty(0);
    torch::Tensor d1 = train ? torch::empty_like(a) : torch::empty(0);
    torch::Tensor d2 = train ? torch::empty_like(a) : torch::empty(0);
    torch::Tensor partials = torch::empty({(long long)gsr_l1_ssim_partials_bytes(d.planes, d.H, d.W) / 8},
                                          a.options().dtype(torch::kFloat64));
    check(gsr_l1_ssim_loss_forward(d.planes, d.H, d.W, C1, C2, lambda_dssim, a.data_ptr<float>(), b.data_ptr<float>(),
                                   loss.data_ptr<float>(), train ? d0.data_ptr<float>() : nullptr,
                                   train ? d1.data_ptr<float>() : nullptr, train ? d2.data_ptr<float>() : nullptr,
                                   reinterpret_cast<char*>(partials.data_ptr()), at::cuda::getCurrentCUDAStream()));
    return std::make_tuple(loss, d0, d1, d2);
}

torch::Tensor l1_ssim_loss_backward(torch::Tensor& image, torch::Tensor& gt, float lambda_dssim, torch::Tensor& dm_dmu1,
                                    torch::Tensor& dm_dsigma1_sq, torch::Tensor& dm_dsigma12, torch::Tensor& upstream) {
    torch::Tensor a = as_f32_cuda(image, "image"), b = as_f32_cuda(gt, "gt");
    torch::Tensor d0 = as_f32_cuda(dm_dmu1, "dm_dmu1"), d1 = as_f32_cuda(dm_dsigma1_sq, "dm_dsigma1_sq"),
                  d2 = as_f32_cuda(dm_dsigma12, "dm_dsigma12");
    const c10::cuda::CUDAGuard guard(a.device());
    const Dims d = dims_of(a, "image");
    torch::Tensor up;
    const float* up_ptr = nullptr;
    if (upstream.numel() != 0) {
        TORCH_CHECK(upstream.numel() == 1, "upstream must be a scalar");
        up = as_f32_cuda(upstream, "upstream");
        up_ptr = up.data_ptr<float>();
    }
    torch::Tensor out = torch::empty_like(a);
    check(gsr_l1_ssim_loss_backward(d.planes, d.H, d.W, lambda_dssim, a.data_ptr<float>(), b.data_ptr<float>(),
                                    d0.data_ptr<float>(), d1.data_ptr<float>(), d2.data_ptr<float>(), up_ptr,
                                    out.data_ptr<float>(), at::cuda::getCurrentCUDAStream()));
    return out;
}

void bind_ssim(pybind11::module_& m, bool rasterizer_module) {
    if (rasterizer_module) {
        // utils/loss_utils.py:24-37 signatures, plus the fused-ssim ones as overloads
        m.def("fusedssim", &fusedssim_map_only);
        m.def("fusedssim_backward", &fusedssim_backward_recompute);
    }
    m.def("fusedssim", &fusedssim);
    m.def("fusedssim_backward", &fusedssim_backward);
    m.def("l1_ssim_loss_forward", &l1_ssim_loss_forward);
    m.def("l1_ssim_loss_backward", &l1_ssim_loss_backward);
}

#ifdef GSR_SSIM_STANDALONE
PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
    bind_ssim(m, false);
    m.attr("backend") = "gsr_b200 sm_100a";
}
#endif
