// torch_binding.cpp -- the thin PyTorch C++ extension over the C ABI of include/gsr.h (module `_gsr_torch`).
//
// It replaces the reference's torch binding: rasterize_points.cu:35-217 (RasterizeGaussiansCUDA,
// RasterizeGaussiansBackwardCUDA, markVisible), bound at ext.cpp:15-19, with the same positional call shapes plus the
// superset arguments of this repo (antialiasing, inverse depth, separate degree-0 SH tensor, in-kernel accumulation
// into caller-owned gradient tensors). torch supplies device memory, the current stream and the device guard; every
// computation happens behind gsr_forward / gsr_backward / gsr_mark_visible in libgsr.so. Compiled by build.py with the
// host compiler only (no CUDA code in this file); diff_gaussian_rasterization/_C.py prefers it over its ctypes binding
// when the module is present, because a call then costs one pybind transition instead of ~60 Python-level operations.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/extension.h>

#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "../../include/gsr.h"

namespace {

const float* fptr(const torch::Tensor& t) { return t.numel() ? t.data_ptr<float>() : nullptr; }

// contiguous fp32 CUDA tensor (rasterize_points.cu:94-111 calls .contiguous() on every input); empty tensors pass
torch::Tensor prep(const torch::Tensor& t, const char* name, std::vector<torch::Tensor>& keep) {
    if (!t.defined() || t.numel() == 0) return t;
    TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor (the rasterizer has no CPU path)");
    torch::Tensor c = t.scalar_type() == torch::kFloat32 ? t : t.to(torch::kFloat32);
    c = c.contiguous();
    keep.push_back(c);
    return c;
}

struct BlobCtx {                 // one scratch blob handed out through the gsr_alloc_fn callback
    torch::Tensor t;
    torch::TensorOptions opt;
};
char* alloc_cb(void* ctx, size_t n) {                       // == resizeFunctional, rasterize_points.cu:27-33
    auto* b = static_cast<BlobCtx*>(ctx);
    try {
        b->t = torch::empty({(long long)n}, b->opt);
    } catch (...) {
        return nullptr;                                     // surfaces as GSR_E_ALLOC
    }
    return reinterpret_cast<char*>(b->t.data_ptr());
}

struct ViewArgs {
    gsr_view v;
    std::vector<torch::Tensor> keep;
};
void fill_view(ViewArgs& a, const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
               const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations, double scale_modifier,
               const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix,
               double tan_fovx, double tan_fovy, int64_t H, int64_t W, const torch::Tensor& sh, int64_t degree,
               const torch::Tensor& campos, bool prefiltered, bool antialiasing, bool debug, const torch::Tensor& dc) {
    TORCH_CHECK(means3D.dim() == 2 && means3D.size(1) == 3, "means3D must have dimensions (num_points, 3)");   // :57-59
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (the rasterizer has no CPU path)");
    int M = sh.numel() ? (int)sh.size(1) : 0;               // rasterize_points.cu:84-88
    const bool has_dc = dc.defined() && dc.numel() != 0;
    if (has_dc) M += 1;
    gsr_view& v = a.v;
    v.P = (int)means3D.size(0); v.D = (int)degree; v.M = M; v.W = (int)W; v.H = (int)H;
    v.tan_fovx = (float)tan_fovx; v.tan_fovy = (float)tan_fovy; v.scale_modifier = (float)scale_modifier;
    v.prefiltered = prefiltered; v.antialiasing = antialiasing; v.debug = debug;
    v.background = fptr(prep(background, "background", a.keep));
    v.means3D = fptr(prep(means3D, "means3D", a.keep));
    v.shs = fptr(prep(sh, "shs", a.keep));
    v.colors_precomp = fptr(prep(colors, "colors_precomp", a.keep));
    v.opacities = fptr(prep(opacity, "opacities", a.keep));
    v.scales = fptr(prep(scales, "scales", a.keep));
    v.rotations = fptr(prep(rotations, "rotations", a.keep));
    v.cov3D_precomp = fptr(prep(cov3D_precomp, "cov3D_precomp", a.keep));
    v.viewmatrix = fptr(prep(viewmatrix, "viewmatrix", a.keep));
    v.projmatrix = fptr(prep(projmatrix, "projmatrix", a.keep));
    v.cam_pos = fptr(prep(campos, "cam_pos", a.keep));
    v.sh_dc = has_dc ? fptr(prep(dc, "sh_dc", a.keep)) : nullptr;
}

void check(int rc, const char* what) { TORCH_CHECK(rc == 0, what, ": ", gsr_error_string(rc), " (code ", rc, ")"); }

}  // namespace

// rasterize_points.cu:35-115 + superset outputs. Returns
// (num_rendered, color[3,H,W], radii[P] int32, geomBuffer, binningBuffer, imgBuffer, invdepth[1,H,W] or empty).
std::tuple<int64_t, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor, torch::Tensor>
rasterize_gaussians(const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& colors,
                    const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations,
                    double scale_modifier, const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix,
                    const torch::Tensor& projmatrix, double tan_fovx, double tan_fovy, int64_t H, int64_t W,
                    const torch::Tensor& sh, int64_t degree, const torch::Tensor& campos, bool prefiltered, bool debug,
                    bool antialiasing, bool want_invdepth, const torch::Tensor& dc) {
    ViewArgs a;
    fill_view(a, background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
              projmatrix, tan_fovx, tan_fovy, H, W, sh, degree, campos, prefiltered, antialiasing, debug, dc);
    c10::cuda::CUDAGuard guard(means3D.device());
    const auto f = torch::TensorOptions().dtype(torch::kFloat32).device(means3D.device());
    const auto bytes = f.dtype(torch::kByte);
    const int64_t P = means3D.size(0);
    torch::Tensor color = torch::empty({3, H, W}, f);
    torch::Tensor radii = torch::empty({P}, f.dtype(torch::kInt32));
    torch::Tensor invdepth = want_invdepth ? torch::empty({1, H, W}, f) : torch::empty({0}, f);
    BlobCtx geom{torch::empty({0}, bytes), bytes}, binning{torch::empty({0}, bytes), bytes}, img{torch::empty({0}, bytes), bytes};
    int rendered = 0;
    const int rc = gsr_forward(&a.v, alloc_cb, &geom, alloc_cb, &binning, alloc_cb, &img, color.data_ptr<float>(),
                               P ? radii.data_ptr<int>() : nullptr, want_invdepth ? invdepth.data_ptr<float>() : nullptr,
                               &rendered, at::cuda::getCurrentCUDAStream().stream());
    check(rc, "gsr_forward");
    return std::make_tuple((int64_t)rendered, color, radii, geom.t, binning.t, img.t, invdepth);
}

namespace {
// The packed gradient accumulator of the backward (48 B per Gaussian) is kept per (device, stream, P) and stays ALL ZERO
// between calls: gsr_backward zeroes the rows it read (gsr_grads.accum_clean), so there is no clear per view.
std::mutex g_acc_mu;
std::map<std::tuple<int, void*, int64_t>, torch::Tensor> g_acc;

torch::Tensor sink_or_new(const c10::optional<torch::Tensor>& sink, at::IntArrayRef shape, const torch::TensorOptions& f,
                          const char* name) {
    if (!sink.has_value()) return torch::empty(shape, f);
    const torch::Tensor& t = *sink;
    int64_t want = 1;
    for (auto d : shape) want *= d;
    TORCH_CHECK(t.is_cuda() && t.device() == f.device() && t.scalar_type() == torch::kFloat32 && t.is_contiguous() &&
                    t.numel() == want,
                "grad_sinks['", name, "'] must be a contiguous fp32 tensor of the input's shape on its device");
    return t;
}
}  // namespace

// rasterize_points.cu:117-196. Returns (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh,
// dL_dscales, dL_drotations, dL_dconic, dL_ddc); entries of inputs that were not supplied -- or that were accumulated
// into a sink -- are undefined tensors (None in Python). `sinks`: 7 optional tensors in the order (means3D, opacities,
// colors_precomp, cov3D_precomp, shs, scales, rotations) + dc; when `accumulate` is set every supplied input needs one.
// `overwrite` (with sinks): the kernel writes the sinks in full instead of adding to them (first view of a batch).
std::vector<c10::optional<torch::Tensor>> rasterize_gaussians_backward(
    const torch::Tensor& background, const torch::Tensor& means3D, const torch::Tensor& radii, const torch::Tensor& colors,
    const torch::Tensor& opacity, const torch::Tensor& scales, const torch::Tensor& rotations, double scale_modifier,
    const torch::Tensor& cov3D_precomp, const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix, double tan_fovx,
    double tan_fovy, const torch::Tensor& dL_dout_color, const c10::optional<torch::Tensor>& dL_dout_invdepth,
    const torch::Tensor& sh, int64_t degree, const torch::Tensor& campos, const torch::Tensor& geomBuffer, int64_t R,
    const torch::Tensor& binningBuffer, const torch::Tensor& imageBuffer, bool antialiasing, bool debug, bool want_conic,
    bool accumulate, const std::vector<c10::optional<torch::Tensor>>& sinks, const torch::Tensor& dc, bool overwrite) {
    const int64_t P = means3D.size(0), H = dL_dout_color.size(1), W = dL_dout_color.size(2);
    ViewArgs a;
    fill_view(a, background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
              projmatrix, tan_fovx, tan_fovy, H, W, sh, degree, campos, false, antialiasing, debug, dc);
    c10::cuda::CUDAGuard guard(means3D.device());
    const auto f = torch::TensorOptions().dtype(torch::kFloat32).device(means3D.device());
    const bool has_dc = dc.defined() && dc.numel() != 0;
    const int64_t M = a.v.M - (has_dc ? 1 : 0);                 // rows of `sh` / dL_dsh
    TORCH_CHECK(!accumulate || sinks.size() == 8, "grad_sinks: expected 8 entries");
    auto sink = [&](int k) -> c10::optional<torch::Tensor> { return accumulate ? sinks[k] : c10::nullopt; };
    auto need = [&](int k, bool supplied, const char* name) {
        TORCH_CHECK(!accumulate || !supplied || sinks[k].has_value(), "grad_sinks: no sink for supplied input '", name,
                    "' (accumulation is all-or-nothing)");
    };
    need(0, true, "means3D"); need(1, true, "opacities"); need(2, colors.numel() != 0, "colors_precomp");
    need(3, cov3D_precomp.numel() != 0, "cov3D_precomp"); need(4, sh.numel() != 0, "shs"); need(5, scales.numel() != 0, "scales");
    need(6, rotations.numel() != 0, "rotations"); need(7, has_dc, "dc");
    torch::Tensor dmeans2D = torch::empty({P, 3}, f);
    torch::Tensor dconic = want_conic ? torch::empty({P, 2, 2}, f) : torch::Tensor();
    torch::Tensor dmeans3D = sink_or_new(sink(0), {P, 3}, f, "means3D");
    torch::Tensor dopacity = sink_or_new(sink(1), {P, 1}, f, "opacities");
    torch::Tensor dcolors = colors.numel() ? sink_or_new(sink(2), {P, 3}, f, "colors_precomp") : torch::Tensor();
    torch::Tensor dcov = cov3D_precomp.numel() ? sink_or_new(sink(3), {P, 6}, f, "cov3D_precomp") : torch::Tensor();
    torch::Tensor dsh = sh.numel() ? sink_or_new(sink(4), {P, M, 3}, f, "shs") : torch::Tensor();
    torch::Tensor dscales = scales.numel() ? sink_or_new(sink(5), {P, 3}, f, "scales") : torch::Tensor();
    torch::Tensor drot = rotations.numel() ? sink_or_new(sink(6), {P, 4}, f, "rotations") : torch::Tensor();
    torch::Tensor ddc = has_dc ? sink_or_new(sink(7), {P, 1, 3}, f, "dc") : torch::Tensor();
    if (P != 0) {
        auto wp = [](const torch::Tensor& t) -> float* { return t.defined() && t.numel() ? t.data_ptr<float>() : nullptr; };
        gsr_grads g{wp(dmeans2D), wp(dopacity), wp(dmeans3D), wp(dcolors), wp(dcov), wp(dsh), wp(dscales), wp(drot), wp(dconic),
                    wp(ddc), (accumulate && !overwrite) ? 1 : 0, 1};   // overwrite: the sinks are written in full
        std::vector<torch::Tensor> keep;
        torch::Tensor dpix = prep(dL_dout_color, "dL_dout_color", keep);
        torch::Tensor dinv = dL_dout_invdepth.has_value() ? prep(*dL_dout_invdepth, "dL_dout_invdepth", keep) : torch::Tensor();
        auto stream = at::cuda::getCurrentCUDAStream();
        torch::Tensor acc;
        const auto key = std::make_tuple((int)means3D.device().index(), (void*)stream.stream(), P);
        {
            std::lock_guard<std::mutex> lk(g_acc_mu);
            auto it = g_acc.find(key);
            if (it == g_acc.end()) {
                if (g_acc.size() >= 8) g_acc.clear();
                it = g_acc.emplace(key, torch::zeros({(long long)gsr_backward_scratch_bytes((int)P)}, f.dtype(torch::kByte))).first;
            }
            acc = it->second;
        }
        const int rc = gsr_backward(&a.v, (int)R, radii.data_ptr<int>(), reinterpret_cast<const char*>(geomBuffer.data_ptr()),
                                    binningBuffer.numel() ? reinterpret_cast<const char*>(binningBuffer.data_ptr()) : nullptr,
                                    reinterpret_cast<const char*>(imageBuffer.data_ptr()), dpix.data_ptr<float>(),
                                    dinv.defined() ? dinv.data_ptr<float>() : nullptr, reinterpret_cast<char*>(acc.data_ptr()), &g,
                                    stream.stream());
        if (rc != 0) {
            std::lock_guard<std::mutex> lk(g_acc_mu);
            g_acc.erase(key);                                   // possibly dirty now
        }
        check(rc, "gsr_backward");
    }
    auto out = [&](const torch::Tensor& t) -> c10::optional<torch::Tensor> {
        if (!t.defined() || accumulate) return c10::nullopt;
        return t;
    };
    std::vector<c10::optional<torch::Tensor>> res;
    res.push_back(dmeans2D);                                    // per view, never accumulated
    res.push_back(out(dcolors)); res.push_back(out(dopacity)); res.push_back(out(dmeans3D)); res.push_back(out(dcov));
    res.push_back(out(dsh)); res.push_back(out(dscales)); res.push_back(out(drot));
    res.push_back(dconic.defined() ? c10::optional<torch::Tensor>(dconic) : c10::nullopt);
    res.push_back(out(ddc));
    return res;
}

// rasterize_points.cu:198-217
torch::Tensor mark_visible(const torch::Tensor& means3D, const torch::Tensor& viewmatrix, const torch::Tensor& projmatrix) {
    TORCH_CHECK(means3D.is_cuda(), "means3D must be a CUDA tensor (the rasterizer has no CPU path)");
    const int64_t P = means3D.size(0);
    torch::Tensor present = torch::zeros({P}, means3D.options().dtype(torch::kBool));
    if (P != 0) {
        c10::cuda::CUDAGuard guard(means3D.device());
        std::vector<torch::Tensor> keep;
        torch::Tensor m = prep(means3D, "means3D", keep), v = prep(viewmatrix, "viewmatrix", keep), p = prep(projmatrix, "projmatrix", keep);
        check(gsr_mark_visible((int)P, m.data_ptr<float>(), v.data_ptr<float>(), p.data_ptr<float>(),
                               reinterpret_cast<uint8_t*>(present.data_ptr()), at::cuda::getCurrentCUDAStream().stream()),
              "gsr_mark_visible");
    }
    return present;
}

void drop_cached_scratch() {
    std::lock_guard<std::mutex> lk(g_acc_mu);
    g_acc.clear();
}

PYBIND11_MODULE(_gsr_torch, m) {
    m.def("rasterize_gaussians", &rasterize_gaussians);
    m.def("rasterize_gaussians_backward", &rasterize_gaussians_backward);
    m.def("mark_visible", &mark_visible);
    m.def("drop_cached_scratch", &drop_cached_scratch);
    m.def("version", []() { return std::string(gsr_version()); });
}
