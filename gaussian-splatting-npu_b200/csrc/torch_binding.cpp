// torch_binding.cpp -- the thin torch-extension layer over the C ABI (include/gsr.h).
//
// Keeps the reference's pybind surface so diff_gaussian_rasterization/__init__.py (and therefore
// gaussian_renderer/__init__.py and train.py) run unchanged:
//   rasterize_gaussians           <- RasterizeGaussiansCUDA          (DGR/rasterize_points.cu:35-124,  ext.cpp:16)
//   rasterize_gaussians_backward  <- RasterizeGaussiansBackwardCUDA  (DGR/rasterize_points.cu:126-223, ext.cpp:17)
//   mark_visible                  <- markVisible                     (DGR/rasterize_points.cu:225-244, ext.cpp:18)
// (DGR = /root/reference/diff-gaussian-rasterization-npu; the fork renamed the first two to ...NPU
// and commented the last two bindings out -- both spellings are exported here.)
// Same argument order, same return tuples, same error text.  Differences: work is issued on the
// current CUDA stream of means3D's device (the reference uses the legacy default stream and the
// *current* device), and outputs the kernels fully write are torch::empty instead of zero-filled.
#include <torch/extension.h>
#include <ATen/cuda/CUDAContext.h>
#include <ATen/cuda/CUDAEvent.h>
#include <c10/cuda/CUDACachingAllocator.h>
#include <c10/cuda/CUDAGuard.h>
#include <mutex>
#include <stdexcept>
#include <tuple>
#include "gsr.h"

namespace {

char* resize_cb(size_t bytes, void* user) {
    auto* t = static_cast<torch::Tensor*>(user);
    t->resize_({(long long)bytes});
    return reinterpret_cast<char*>(t->data_ptr());
}

// contiguous fp32 data pointer, nullptr for an empty tensor; `keep` owns any temporary copy
const float* fptr(const torch::Tensor& t, std::vector<torch::Tensor>& keep, bool align16 = false) {
    if (t.numel() == 0) return nullptr;
    torch::Tensor c = t.contiguous();
    if (c.scalar_type() != torch::kFloat32) c = c.to(torch::kFloat32);
    if (align16 && (reinterpret_cast<uintptr_t>(c.data_ptr()) & 15)) c = c.clone();
    keep.push_back(c);
    return c.data_ptr<float>();
}

void check(int status) {
    if (status != GSR_OK) throw std::runtime_error(gsr_last_error());
}

}  // namespace

std::tuple<int, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
RasterizeGaussiansCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                       const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                       const float scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                       const torch::Tensor& projmatrix, const float tan_fovx, const float tan_fovy,
                       const int image_height, const int image_width, const torch::Tensor& sh, const int degree,
                       const torch::Tensor& campos, const bool prefiltered, const bool antialiasing, const bool debug) {
    if (means3D.ndimension() != 2 || means3D.size(1) != 3) {
        AT_ERROR("means3D must have dimensions (num_points, 3)");
    }
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (there is no CPU fallback)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    cudaStream_t stream = at::cuda::getCurrentCUDAStream();

    const int P = means3D.size(0);
    const int H = image_height, W = image_width;
    auto float_opts = means3D.options().dtype(torch::kFloat32);
    auto int_opts = means3D.options().dtype(torch::kInt32);
    auto byte_opts = means3D.options().dtype(torch::kByte);

    torch::Tensor out_color = P ? torch::empty({3, H, W}, float_opts) : torch::zeros({3, H, W}, float_opts);
    torch::Tensor out_invdepth = P ? torch::empty({1, H, W}, float_opts) : torch::zeros({1, H, W}, float_opts);
    torch::Tensor radii = P ? torch::empty({P}, int_opts) : torch::zeros({P}, int_opts);
    torch::Tensor geomBuffer = torch::empty({0}, byte_opts);
    torch::Tensor binningBuffer = torch::empty({0}, byte_opts);
    torch::Tensor imgBuffer = torch::empty({0}, byte_opts);

    int rendered = 0;
    if (P != 0) {
        std::vector<torch::Tensor> keep;
        gsr_forward_args a;
        a.P = P; a.D = degree; a.M = sh.size(0) != 0 ? (int)sh.size(1) : 0;
        a.width = W; a.height = H;
        a.background = fptr(background, keep);
        a.means3D = fptr(means3D, keep);
        a.shs = fptr(sh, keep, true);
        a.colors_precomp = fptr(colors, keep);
        a.opacities = fptr(opacity, keep);
        a.scales = fptr(scales, keep);
        a.scale_modifier = scale_modifier;
        a.rotations = fptr(rotations, keep, true);
        a.cov3D_precomp = fptr(cov3D_precomp, keep);
        a.viewmatrix = fptr(viewmatrix, keep);
        a.projmatrix = fptr(projmatrix, keep);
        a.cam_pos = fptr(campos, keep);
        a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy;
        a.prefiltered = prefiltered; a.antialiasing = antialiasing; a.debug = debug;
        a.out_color = out_color.data_ptr<float>();
        a.out_invdepth = out_invdepth.data_ptr<float>();
        a.radii = radii.data_ptr<int>();
        check(gsr_forward(&a, resize_cb, &geomBuffer, resize_cb, &binningBuffer, resize_cb, &imgBuffer, stream, &rendered));
    }
    return std::make_tuple(rendered, out_color, radii, geomBuffer, binningBuffer, imgBuffer, out_invdepth);
}

// Optional gradient sink (extension over the reference API, see diff_gaussian_rasterization.grad_sink):
// the five parameter gradients are written -- or, with accumulate, added -- straight into caller-owned
// tensors (views of an all-reduce bucket) instead of fresh allocations.
struct GradSink { torch::Tensor means3D, sh, opacity, scales, rotations; bool accumulate = false; bool active = false; bool defer = false; };

// Deferred sink (grad_sink(..., defer=True)): autograd hands over the views of a step one backward at a time; each runs its
// backward blend at once (gsr_backward_blend) and is parked here with the tensors the per-Gaussian stage still needs; ONE
// gsr_backward_params_batch per (up to) GSR_MAX_BATCH_VIEWS views then sums their parameter gradients in registers and
// writes the sink once (flush_grad_sink, called when the Python context closes).  Process-wide like the Python-side sink:
// autograd runs backward nodes on its own threads.
struct DeferredView {
    gsr_backward_args a;                       // pointers into the tensors below
    std::vector<torch::Tensor> keep;
    at::cuda::CUDAEvent done;                  // recorded behind this view's blend, on the stream the blend ran on
};
struct DeferredState {
    std::mutex mu;
    std::vector<DeferredView> views;
    bool accumulate = false;                   // of the next flush: the first one follows the sink's flag, later ones add
    c10::Device device = c10::Device(c10::kCPU);
};
DeferredState& deferred() { static DeferredState d; return d; }

// Two views may share one gsr_backward_params_batch launch only if they read the same Gaussians (all five parameter
// arrays, not just means3D / shs: a caller that recomputes sigmoid / exp per render hands over fresh tensors every time)
// and write the same sink.  This is exactly what batch_shared_check (gsr_api.cu) insists on, so a view that passes it
// can never make the flush fail and take the parked views' gradients with it.
static bool same_batch(const gsr_backward_args& f, const gsr_backward_args& a) {
    return f.P == a.P && f.D == a.D && f.M == a.M && f.means3D == a.means3D && f.shs == a.shs && f.opacities == a.opacities &&
           f.scales == a.scales && f.rotations == a.rotations && f.scale_modifier == a.scale_modifier &&
           f.antialiasing == a.antialiasing && f.dL_dmean3D == a.dL_dmean3D && f.dL_dsh == a.dL_dsh &&
           f.dL_dopacity == a.dL_dopacity && f.dL_dscale == a.dL_dscale && f.dL_drot == a.dL_drot;
}

void flush_locked(DeferredState& d) {
    if (d.views.empty()) return;
    const c10::cuda::CUDAGuard guard(d.device);
    // The flush runs on the flushing thread's current stream, which need not be the one the blends ran on (backward()
    // under another stream context, or the context closed elsewhere): wait for every view's blend, and tell the caching
    // allocator that the parked tensors are read on this stream too before they are released below.
    const at::cuda::CUDAStream flush_stream = at::cuda::getCurrentCUDAStream();
    cudaStream_t stream = flush_stream;
    std::vector<gsr_backward_args> arr;
    for (auto& v : d.views) {
        arr.push_back(v.a);
        v.done.block(flush_stream);
        for (const torch::Tensor& t : v.keep)
            if (t.defined() && t.is_cuda() && t.numel()) c10::cuda::CUDACachingAllocator::recordStream(t.storage().data_ptr(), flush_stream);
    }
    arr[0].accumulate = d.accumulate ? 1 : 0;
    const int st = gsr_backward_params_batch((int)arr.size(), arr.data(), stream);
    d.views.clear();
    d.accumulate = true;                       // whatever follows in the same context adds to what was just written
    check(st);
}

typedef std::tuple<torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor> BwdTuple;

static BwdTuple
backward_impl(const GradSink& sink, const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                               const torch::Tensor& colors, const torch::Tensor& opacities, const torch::Tensor& scales,
                               const torch::Tensor& rotations, const float scale_modifier, const torch::Tensor& cov3D_precomp,
                               const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix, const float tan_fovx,
                               const float tan_fovy, const torch::Tensor& dL_dout_color, const torch::Tensor& dL_dout_invdepth,
                               const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
                               const torch::Tensor& geomBuffer, const int R, const torch::Tensor& binningBuffer,
                               const torch::Tensor& imageBuffer, const bool antialiasing, const bool debug) {
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (there is no CPU fallback)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    cudaStream_t stream = at::cuda::getCurrentCUDAStream();

    const int P = means3D.size(0);
    const int H = dL_dout_color.size(1), W = dL_dout_color.size(2);
    const int M = sh.size(0) != 0 ? (int)sh.size(1) : 0;
    auto opts = means3D.options().dtype(torch::kFloat32);
    const bool has_depth = dL_dout_invdepth.size(0) != 0;
    const bool has_scales = scales.numel() != 0;
    const bool live = P != 0;

    auto mk = [&](std::initializer_list<int64_t> shape, bool written) {
        return (live && written) ? torch::empty(shape, opts) : torch::zeros(shape, opts);
    };
    auto sunk = [&](const torch::Tensor& t, std::initializer_list<int64_t> shape, bool written, const char* name) {
        if (!(sink.active && live && written)) return mk(shape, written);
        int64_t n = 1; for (auto d : shape) n *= d;
        TORCH_CHECK(t.defined() && t.is_cuda() && t.device() == means3D.device() && t.scalar_type() == torch::kFloat32 &&
                    t.is_contiguous() && t.numel() == n && (reinterpret_cast<uintptr_t>(t.data_ptr()) & 15) == 0,
                    "grad sink tensor '", name, "' must be a contiguous, 16-byte aligned fp32 CUDA tensor of ", n, " elements");
        return t;
    };
    torch::Tensor dL_dmeans3D = sunk(sink.means3D, {P, 3}, true, "means3D");
    torch::Tensor dL_dmeans2D = mk({P, 3}, true);
    torch::Tensor dL_dcolors = mk({P, 3}, true);
    torch::Tensor dL_dconic = mk({P, 2, 2}, true);
    torch::Tensor dL_dopacity = sunk(sink.opacity, {P, 1}, true, "opacity");
    torch::Tensor dL_dcov3D = mk({P, 6}, true);
    torch::Tensor dL_dsh = sunk(sink.sh, {P, M, 3}, M > 0, "sh");
    torch::Tensor dL_dscales = sunk(sink.scales, {P, 3}, has_scales, "scales");
    torch::Tensor dL_drotations = sunk(sink.rotations, {P, 4}, has_scales, "rotations");
    torch::Tensor dL_dinvdepths = has_depth ? mk({P, 1}, true) : torch::zeros({0, 1}, opts);

    if (live) {
        std::vector<torch::Tensor> keep;
        gsr_backward_args a;
        a.P = P; a.D = degree; a.M = M; a.R = R; a.width = W; a.height = H;
        a.background = fptr(background, keep);
        a.means3D = fptr(means3D, keep);
        a.shs = fptr(sh, keep, true);
        a.colors_precomp = fptr(colors, keep);
        a.opacities = fptr(opacities, keep);
        a.scales = fptr(scales, keep);
        a.scale_modifier = scale_modifier;
        a.rotations = fptr(rotations, keep, true);
        a.cov3D_precomp = fptr(cov3D_precomp, keep);
        a.viewmatrix = fptr(viewmatrix, keep);
        a.projmatrix = fptr(projmatrix, keep);
        a.cam_pos = fptr(campos, keep);
        a.tan_fovx = tan_fovx; a.tan_fovy = tan_fovy;
        torch::Tensor radii_c = radii.contiguous();
        a.radii = radii_c.data_ptr<int>();
        torch::Tensor gb = geomBuffer.contiguous(), bb = binningBuffer.contiguous(), ib = imageBuffer.contiguous();
        a.geom_buffer = reinterpret_cast<char*>(gb.data_ptr());
        a.binning_buffer = reinterpret_cast<char*>(bb.data_ptr());
        a.image_buffer = reinterpret_cast<char*>(ib.data_ptr());
        a.dL_dpix = fptr(dL_dout_color, keep);
        a.dL_dinvdepth_pix = has_depth ? fptr(dL_dout_invdepth, keep) : nullptr;
        a.dL_dmean2D = dL_dmeans2D.data_ptr<float>();
        a.dL_dconic = dL_dconic.data_ptr<float>();
        a.dL_dopacity = dL_dopacity.data_ptr<float>();
        a.dL_dcolor = dL_dcolors.data_ptr<float>();
        a.dL_dinvdepth = has_depth ? dL_dinvdepths.data_ptr<float>() : nullptr;
        a.dL_dmean3D = dL_dmeans3D.data_ptr<float>();
        a.dL_dcov3D = dL_dcov3D.data_ptr<float>();
        a.dL_dsh = M > 0 ? dL_dsh.data_ptr<float>() : nullptr;
        a.dL_dscale = has_scales ? dL_dscales.data_ptr<float>() : nullptr;
        a.dL_drot = has_scales ? dL_drotations.data_ptr<float>() : nullptr;
        a.antialiasing = antialiasing; a.debug = debug;
        a.accumulate = (sink.active && sink.accumulate) ? 1 : 0;            // the reference API always overwrites
        const bool can_defer = sink.active && sink.defer && M == 16 && has_scales && colors.numel() == 0 && cov3D_precomp.numel() == 0;
        if (can_defer) {
            DeferredState& d = deferred();
            std::lock_guard<std::mutex> lock(d.mu);
            // a different set of Gaussians / sinks than the parked views' (or a full batch): write those out first
            if (!d.views.empty()) {
                if (d.views.size() >= GSR_MAX_BATCH_VIEWS || !same_batch(d.views[0].a, a)) flush_locked(d);
            } else {
                d.accumulate = sink.accumulate;
            }
            check(gsr_backward_blend(&a, stream));
            dL_dcov3D.zero_();                     // not produced on this path (it only exists per view); nobody reads it
            d.device = means3D.device();
            DeferredView dv;
            dv.a = a;
            dv.keep = keep;
            dv.done.record(at::cuda::getCurrentCUDAStream());
            for (const torch::Tensor& t : { radii_c, gb, dL_dmeans2D, dL_dconic, dL_dcolors, dL_dinvdepths, dL_dmeans3D, dL_dopacity,
                                             dL_dsh, dL_dscales, dL_drotations }) dv.keep.push_back(t);
            d.views.push_back(std::move(dv));
        } else {
            check(gsr_backward(&a, stream));
        }
    }
    return std::make_tuple(dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations);
}

BwdTuple
RasterizeGaussiansBackwardCUDA(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                               const torch::Tensor& colors, const torch::Tensor& opacities, const torch::Tensor& scales,
                               const torch::Tensor& rotations, const float scale_modifier, const torch::Tensor& cov3D_precomp,
                               const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix, const float tan_fovx,
                               const float tan_fovy, const torch::Tensor& dL_dout_color, const torch::Tensor& dL_dout_invdepth,
                               const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
                               const torch::Tensor& geomBuffer, const int R, const torch::Tensor& binningBuffer,
                               const torch::Tensor& imageBuffer, const bool antialiasing, const bool debug) {
    return backward_impl(GradSink(), background, means3D, radii, colors, opacities, scales, rotations, scale_modifier, cov3D_precomp,
                         viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, dL_dout_invdepth, sh, degree, campos,
                         geomBuffer, R, binningBuffer, imageBuffer, antialiasing, debug);
}

// Same arguments + the sink tensors and the accumulate flag.
BwdTuple
RasterizeGaussiansBackwardInto(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii,
                               const torch::Tensor& colors, const torch::Tensor& opacities, const torch::Tensor& scales,
                               const torch::Tensor& rotations, const float scale_modifier, const torch::Tensor& cov3D_precomp,
                               const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix, const float tan_fovx,
                               const float tan_fovy, const torch::Tensor& dL_dout_color, const torch::Tensor& dL_dout_invdepth,
                               const torch::Tensor& sh, const int degree, const torch::Tensor& campos,
                               const torch::Tensor& geomBuffer, const int R, const torch::Tensor& binningBuffer,
                               const torch::Tensor& imageBuffer, const bool antialiasing, const bool debug,
                               const torch::Tensor& sink_means3D, const torch::Tensor& sink_sh, const torch::Tensor& sink_opacity,
                               const torch::Tensor& sink_scales, const torch::Tensor& sink_rotations, const bool accumulate,
                               const bool defer) {
    GradSink sink;
    sink.means3D = sink_means3D; sink.sh = sink_sh; sink.opacity = sink_opacity; sink.scales = sink_scales;
    sink.rotations = sink_rotations; sink.accumulate = accumulate; sink.active = true; sink.defer = defer;
    return backward_impl(sink, background, means3D, radii, colors, opacities, scales, rotations, scale_modifier, cov3D_precomp,
                         viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, dL_dout_invdepth, sh, degree, campos,
                         geomBuffer, R, binningBuffer, imageBuffer, antialiasing, debug);
}

// Writes out whatever grad_sink(defer=True) has parked; returns how many views that were.
int FlushGradSink() {
    DeferredState& d = deferred();
    std::lock_guard<std::mutex> lock(d.mu);
    const int n = (int)d.views.size();
    flush_locked(d);
    d.accumulate = false;
    return n;
}

torch::Tensor markVisible(torch::Tensor& means3D, torch::Tensor& viewmatrix, torch::Tensor& projmatrix) {
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (there is no CPU fallback)");
    const c10::cuda::CUDAGuard guard(means3D.device());
    const int P = means3D.size(0);
    torch::Tensor present = torch::full({P}, false, means3D.options().dtype(at::kBool));
    if (P != 0) {
        std::vector<torch::Tensor> keep;
        check(gsr_mark_visible(P, fptr(means3D, keep), fptr(viewmatrix, keep), fptr(projmatrix, keep),
                               reinterpret_cast<unsigned char*>(present.data_ptr<bool>()), at::cuda::getCurrentCUDAStream()));
    }
    return present;
}

void bind_ssim(pybind11::module_& m, bool rasterizer_module);   // ssim_binding.cpp

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
    m.def("rasterize_gaussians", &RasterizeGaussiansCUDA);
    m.def("rasterize_gaussians_backward", &RasterizeGaussiansBackwardCUDA);
    m.def("mark_visible", &markVisible);
    m.def("rasterize_gaussians_backward_into", &RasterizeGaussiansBackwardInto);
    m.def("flush_grad_sink", &FlushGradSink);
    // the fork's spellings (DGR/ext.cpp:16, rasterize_points.h:18,41)
    m.def("rasterize_gaussians_npu", &RasterizeGaussiansCUDA);
    m.def("rasterize_gaussians_backward_npu", &RasterizeGaussiansBackwardCUDA);
    // loss stage (utils/loss_utils.py:16-19 imports fusedssim / fusedssim_backward from this module)
    bind_ssim(m, true);
    m.attr("backend") = "gsr_b200 sm_100a";
}
