// gsr_api.cu -- the extern "C" boundary (include/gsr.h) and the host-side stage sequence.
//
// Mirrors CudaRasterizer::Rasterizer::{forward,backward,markVisible}
// (diff-gaussian-rasterization-npu/cuda_rasterizer/rasterizer_impl.cu:141-153, :198-341, :345-450).
// No global mutable state: scratch lives in the three caller-owned buffers, errors and stage
// timings are thread-local.
#include "../../include/gsr.h"
#include "gsr_common.cuh"
#include <cstdio>
#include <cstring>
#include <cstdarg>

static thread_local char tl_error[512] = "";
static thread_local int tl_profiling = 0;
static thread_local float tl_stage_ms[12] = { -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1, -1 };

static int fail(int code, const char* fmt, ...) {
    va_list ap; va_start(ap, fmt);
    vsnprintf(tl_error, sizeof(tl_error), fmt, ap);
    va_end(ap);
    return code;
}

#define CUDA_TRY(expr)                                                                         \
    do { cudaError_t e_ = (expr);                                                              \
         if (e_ != cudaSuccess) return fail(GSR_ERR_CUDA, "[CUDA ERROR] in %s\nLine %d: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); } while (0)

// reference CHECK_CUDA (auxiliary.h:178-185): with debug, synchronise after the stage and report
#define STAGE_CHECK(debug, stream, what)                                                       \
    do { cudaError_t e_ = cudaGetLastError();                                                  \
         if (e_ == cudaSuccess && (debug)) e_ = cudaStreamSynchronize(stream);                 \
         if (e_ != cudaSuccess) return fail(GSR_ERR_CUDA, "[CUDA ERROR] in %s: %s", what, cudaGetErrorString(e_)); } while (0)

static inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

extern "C" const char* gsr_last_error(void) { return tl_error; }
// error channel for the other translation units (ssim.cu, adam.cu): same thread-local message
int gsr_set_error(int code, const char* msg) { return fail(code, "%s", msg); }
extern "C" const char* gsr_version(void) { return "gsr_b200 0.1 (sm_100a)"; }
extern "C" void gsr_set_profiling(int enabled) { tl_profiling = enabled; }
extern "C" float gsr_last_stage_ms(int stage) { return (stage >= 0 && stage < 12) ? tl_stage_ms[stage] : -1.0f; }

// ---- scratch layouts ------------------------------------------------------------------------------
// Internal geometry layout = the public one + the depth-sort ping-pong buffers and scan state.
struct GeomInternal { gsr_geometry_layout pub; size_t gkeys[2], gvals_b, gsort_temp, scan_state; };
static GeomInternal geom_layout(int P) {
    GeomInternal g; gsr_geometry_layout* o = &g.pub;
    size_t off = 0; const size_t p = (size_t)(P > 0 ? P : 0);
    o->depths = off;         off = up256(off + p * 4);
    o->records = off;        off = up256(off + p * GSR_REC_FLOATS * 4);
    o->cov3D = off;          off = up256(off + p * 6 * 4);
    o->clamped = off;        off = up256(off + p);
    o->tiles_touched = off;  off = up256(off + p * 4);
    o->point_offsets = off;  off = up256(off + p * 4);
    o->depth_order = off;    off = up256(off + p * 4);     // = depth-sort values, buffer (a)
    o->sorted_offsets = off; off = up256(off + p * 4);
    g.gkeys[0] = off;        off = up256(off + p * 4);
    g.gkeys[1] = off;        off = up256(off + p * 4);
    g.gvals_b = off;         off = up256(off + p * 4);
    g.gsort_temp = off;      off = up256(off + sort_temp_bytes(p));
    g.scan_state = off;      off = up256(off + gsr_scan_state_words(p) * 4);
    o->total = off + 256;
    return g;
}
extern "C" void gsr_get_geometry_layout(int P, gsr_geometry_layout* o) { *o = geom_layout(P).pub; }
extern "C" void gsr_get_image_layout(int width, int height, gsr_image_layout* o) {
    const size_t hw = (size_t)width * height;
    const size_t tiles = (size_t)((width + GSR_TILE - 1) / GSR_TILE) * ((height + GSR_TILE - 1) / GSR_TILE);
    size_t off = 0;
    o->final_T = off;   off = up256(off + hw * 4);
    o->n_contrib = off; off = up256(off + hw * 4);
    o->ranges = off;    off = up256(off + tiles * 8);
    o->total = off + 256;
}
extern "C" void gsr_get_binning_layout(size_t n, int width, int height, gsr_binning_layout* o) {
    (void)width; (void)height;
    size_t off = 0;
    o->tile_keys = off;       off = up256(off + n * 4);
    o->point_list = off;      off = up256(off + n * 4);
    o->tile_keys_alt = off;   off = up256(off + n * 4);
    o->point_list_alt = off;  off = up256(off + n * 4);
    o->sort_temp = off;       off = up256(off + sort_temp_bytes(n));
    o->total = off + 256;
}
extern "C" size_t gsr_geometry_bytes(int P) { gsr_geometry_layout l; gsr_get_geometry_layout(P, &l); return l.total; }
extern "C" size_t gsr_image_bytes(int w, int h) { gsr_image_layout l; gsr_get_image_layout(w, h, &l); return l.total; }
extern "C" size_t gsr_binning_bytes(size_t n, int w, int h) { gsr_binning_layout l; gsr_get_binning_layout(n, w, h, &l); return l.total; }

static inline char* align_base(char* p) { return (char*)(((uintptr_t)p + 255) & ~(uintptr_t)255); }

static GeomView geom_view(char* base, int P) {
    base = align_base(base);
    const GeomInternal gi = geom_layout(P); const gsr_geometry_layout& l = gi.pub;
    GeomView g;
    g.depths = (float*)(base + l.depths);
    g.records = (float4*)(base + l.records);
    g.cov3D = (float*)(base + l.cov3D);
    g.clamped = (unsigned char*)(base + l.clamped);
    g.tiles_touched = (uint32_t*)(base + l.tiles_touched);
    g.point_offsets = (uint32_t*)(base + l.point_offsets);
    g.depth_order = (uint32_t*)(base + l.depth_order);
    g.sorted_offsets = (uint32_t*)(base + l.sorted_offsets);
    g.gsort_keys[0] = (uint32_t*)(base + gi.gkeys[0]);
    g.gsort_keys[1] = (uint32_t*)(base + gi.gkeys[1]);
    g.gsort_vals[0] = g.depth_order;
    g.gsort_vals[1] = (uint32_t*)(base + gi.gvals_b);
    g.gsort_temp = base + gi.gsort_temp;
    g.scan_state = (uint32_t*)(base + gi.scan_state);
    return g;
}
static ImageView image_view(char* base, int w, int h) {
    base = align_base(base);
    gsr_image_layout l; gsr_get_image_layout(w, h, &l);
    ImageView v;
    v.final_T = (float*)(base + l.final_T);
    v.n_contrib = (uint32_t*)(base + l.n_contrib);
    v.ranges = (uint2*)(base + l.ranges);
    return v;
}
static BinningView binning_view(char* base, size_t n, int w, int h) {
    base = align_base(base);
    gsr_binning_layout l; gsr_get_binning_layout(n, w, h, &l);
    BinningView b;
    b.tile_keys = (uint32_t*)(base + l.tile_keys);
    b.point_list = (uint32_t*)(base + l.point_list);
    b.tile_keys_alt = (uint32_t*)(base + l.tile_keys_alt);
    b.point_list_alt = (uint32_t*)(base + l.point_list_alt);
    b.block_masks = (unsigned char*)b.tile_keys_alt;
    b.sort_temp = base + l.sort_temp;
    return b;
}

// rasterizer_impl.cu:35-50
static uint32_t higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) { step /= 2; if (n >> msb) msb += step; else msb -= step; }
    if (n >> msb) msb++;
    return msb;
}

// ---- stage timers (profiling only) -----------------------------------------------------------------
struct StageTimer {
    cudaEvent_t ev[12]; int n = 0; bool on; cudaStream_t s;
    StageTimer(bool enabled, cudaStream_t st) : on(enabled), s(st) { if (on) for (auto& e : ev) cudaEventCreate(&e); }
    void mark() { if (on && n < 12) cudaEventRecord(ev[n++], s); }
    void finish(const int* stage_ids) {
        if (!on) return;
        cudaStreamSynchronize(s);
        for (int k = 0; k + 1 < n; k++) { float ms = -1; cudaEventElapsedTime(&ms, ev[k], ev[k + 1]); tl_stage_ms[stage_ids[k]] = ms; }
        for (auto& e : ev) cudaEventDestroy(e);
        on = false;
    }
    ~StageTimer() { if (on) for (auto& e : ev) cudaEventDestroy(e); }
};

static int check_aligned16(const void* p, const char* name) {
    if (p && ((uintptr_t)p & 15)) return fail(GSR_ERR_INVALID, "%s must be 16-byte aligned", name);
    return GSR_OK;
}

extern "C" int gsr_forward(const gsr_forward_args* a, gsr_alloc_fn geom_alloc, void* geom_user,
                           gsr_alloc_fn binning_alloc, void* binning_user, gsr_alloc_fn image_alloc,
                           void* image_user, void* stream_, int* num_rendered) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!a || !num_rendered || !geom_alloc || !binning_alloc || !image_alloc) return fail(GSR_ERR_INVALID, "null argument");
    *num_rendered = 0;
    const int P = a->P, W = a->width, H = a->height;
    if (P < 0 || W <= 0 || H <= 0) return fail(GSR_ERR_INVALID, "bad sizes P=%d W=%d H=%d", P, W, H);
    if (!a->out_color) return fail(GSR_ERR_INVALID, "out_color is null");
    if (P == 0) return GSR_OK;     // reference: kernels skipped, outputs stay as allocated (rasterize_points.cu:88)
    if (!a->means3D || !a->opacities || !a->viewmatrix || !a->projmatrix || !a->cam_pos || !a->background || !a->radii)
        return fail(GSR_ERR_INVALID, "missing required input pointer");
    if (!a->shs && !a->colors_precomp)   // reference: NUM_CHANNELS != 3 && colors_precomp == nullptr (rasterizer_impl.cu:244-247)
        return fail(GSR_ERR_INVALID, "For non-RGB, provide precomputed Gaussian colors!");
    if (a->shs && a->M <= 0) return fail(GSR_ERR_INVALID, "shs given but M=%d", a->M);
    if (a->shs && (a->D < 0 || (a->D + 1) * (a->D + 1) > a->M || a->D > 3)) return fail(GSR_ERR_INVALID, "SH degree %d does not fit M=%d", a->D, a->M);
    if (!a->cov3D_precomp && (!a->scales || !a->rotations)) return fail(GSR_ERR_INVALID, "need scales+rotations or cov3D_precomp");
    if (check_aligned16(a->rotations, "rotations") || (a->M == 16 && check_aligned16(a->shs, "shs"))) return GSR_ERR_INVALID;

    const int gx = (W + GSR_TILE - 1) / GSR_TILE, gy = (H + GSR_TILE - 1) / GSR_TILE;

    char* geom_raw = geom_alloc(gsr_geometry_bytes(P), geom_user);
    if (!geom_raw) return fail(GSR_ERR_ALLOC, "geometry allocator returned null");
    char* img_raw = image_alloc(gsr_image_bytes(W, H), image_user);
    if (!img_raw) return fail(GSR_ERR_ALLOC, "image allocator returned null");
    GeomView g = geom_view(geom_raw, P);
    ImageView im = image_view(img_raw, W, H);

    FwdParams p;
    p.P = P; p.D = a->D; p.M = a->shs ? a->M : 0; p.W = W; p.H = H; p.gx = gx; p.gy = gy;
    p.background = a->background; p.means3D = a->means3D; p.shs = a->shs; p.colors_precomp = a->colors_precomp;
    p.opacities = a->opacities; p.scales = a->scales; p.rotations = a->rotations; p.cov3D_precomp = a->cov3D_precomp;
    p.viewmatrix = a->viewmatrix; p.projmatrix = a->projmatrix; p.cam_pos = a->cam_pos;
    p.scale_modifier = a->scale_modifier; p.tan_fovx = a->tan_fovx; p.tan_fovy = a->tan_fovy;
    p.focal_y = H / (2.0f * a->tan_fovy);          // rasterizer_impl.cu:224-225
    p.focal_x = W / (2.0f * a->tan_fovx);
    p.prefiltered = a->prefiltered; p.antialiasing = a->antialiasing;

    StageTimer tm(tl_profiling != 0, stream);
    tm.mark();
    CUDA_TRY(cudaMemsetAsync(g.scan_state, 0, gsr_scan_state_words((size_t)P) * 4, stream));
    launch_preprocess_fwd(p, g, a->radii, stream);
    STAGE_CHECK(a->debug, stream, "preprocess");
    tm.mark();
    // depth sort of the visible Gaussians: 32 key bits = 4 passes, an even count, so starting in the
    // (a) pair makes the sorted indices land in g.depth_order (= gsort_vals[0])
    const int gsort_bits = 32, gsort_passes = (gsort_bits + 7) / 8;
    const int g0 = (gsort_passes & 1) ? 1 : 0;
    launch_scan_compact(P, g, g.gsort_keys[g0], g.gsort_vals[g0], stream);
    STAGE_CHECK(a->debug, stream, "scan");

    // instance / visible counts -> host.  The reference blocks in cudaMemcpy here
    // (rasterizer_impl.cu:283-284); we only record an event and keep the GPU busy with the depth
    // sort (which needs no host-side sizes) while the host waits for the 12 bytes.
    static thread_local uint32_t* pinned = nullptr;
    static thread_local cudaEvent_t counts_ready_dev[64] = {};     // one event per device this thread used
    int dev_id = 0;
    CUDA_TRY(cudaGetDevice(&dev_id));
    if (dev_id < 0 || dev_id >= 64) return fail(GSR_ERR_INVALID, "device ordinal %d out of range", dev_id);
    if (!pinned) CUDA_TRY(cudaHostAlloc((void**)&pinned, 64, cudaHostAllocPortable));
    if (!counts_ready_dev[dev_id]) CUDA_TRY(cudaEventCreateWithFlags(&counts_ready_dev[dev_id], cudaEventDisableTiming));
    cudaEvent_t counts_ready = counts_ready_dev[dev_id];
    CUDA_TRY(cudaMemcpyAsync(pinned, g.scan_state + 1, 12, cudaMemcpyDeviceToHost, stream));
    CUDA_TRY(cudaEventRecord(counts_ready, stream));
    tm.mark();
    {
        int in_b = 0;
        launch_sort_pairs32_pingpong(g.gsort_keys[g0], g.gsort_vals[g0], g.gsort_keys[1 - g0], g.gsort_vals[1 - g0],
                                     (size_t)P, g.scan_state + 3, gsort_bits, g.gsort_temp, stream, &in_b);
        STAGE_CHECK(a->debug, stream, "depth sort");
    }
    tm.mark();
    CUDA_TRY(cudaEventSynchronize(counts_ready));
    const uint32_t trapped = pinned[0];
    const size_t N = pinned[1];
    if (trapped) return fail(GSR_ERR_PREFILTERED, "Point is filtered although prefiltered is set. This shouldn't happen!");
    if (N >= (1u << 30)) return fail(GSR_ERR_INVALID, "num_rendered=%zu exceeds the supported 2^30", N);
    *num_rendered = (int)N;
    tm.mark();      // host wait ends here (stage 8 = device idle gap, if any)

    char* bin_raw = binning_alloc(gsr_binning_bytes(N, W, H), binning_user);
    if (!bin_raw) return fail(GSR_ERR_ALLOC, "binning allocator returned null");
    BinningView b = binning_view(bin_raw, N, W, H);

    const int end_bit = (int)higher_msb((uint32_t)(gx * gy));     // tile-id bits (rasterizer_impl.cu:303)
    const int passes = (end_bit + 7) / 8;
    // ping-pong: start in whichever buffer pair makes the sorted result land in (tile_keys, point_list)
    uint32_t* k_start = (passes & 1) ? b.tile_keys_alt : b.tile_keys;
    uint32_t* v_start = (passes & 1) ? b.point_list_alt : b.point_list;
    uint32_t* k_other = (passes & 1) ? b.tile_keys : b.tile_keys_alt;
    uint32_t* v_other = (passes & 1) ? b.point_list : b.point_list_alt;
    if (N > 0) {
        launch_sorted_scan(P, g, g.depth_order, g.sorted_offsets, stream);
        launch_emit_tiles(P, gx, gy, g, g.depth_order, g.sorted_offsets, a->radii, k_start, v_start, stream);
        STAGE_CHECK(a->debug, stream, "duplicateWithKeys");
        tm.mark();
        int in_b = 0;
        launch_sort_pairs32_pingpong(k_start, v_start, k_other, v_other, N, nullptr, end_bit, b.sort_temp, stream, &in_b);
        STAGE_CHECK(a->debug, stream, "radix sort");
        tm.mark();
    } else { tm.mark(); tm.mark(); }
    launch_tile_ranges(N, b.tile_keys, im.ranges, gx * gy, stream);
    STAGE_CHECK(a->debug, stream, "identifyTileRanges");
    tm.mark();
    launch_render_fwd(W, H, gx, gy, im.ranges, b.point_list, b.block_masks, g.records, a->background, im.final_T, im.n_contrib,
                      a->out_color, a->out_invdepth, stream);
    STAGE_CHECK(a->debug, stream, "render");
    tm.mark();
    // marks: start | preprocess | scan+copy | depth sort | host wait | scan2+emit | tile sort | ranges | render
    static const int ids[] = { 0, 1, 9, 8, 2, 3, 4, 5 };
    tm.finish(ids);
    return GSR_OK;
}

extern "C" int gsr_materialize_keys(int P, int num_rendered, int width, int height, const char* geom_buffer,
                                    const char* binning_buffer, uint64_t* out_keys, void* stream_) {
    if (P < 0 || num_rendered < 0 || width <= 0 || height <= 0) return fail(GSR_ERR_INVALID, "bad sizes");
    if (num_rendered == 0) return GSR_OK;
    if (!geom_buffer || !binning_buffer || !out_keys) return fail(GSR_ERR_INVALID, "null argument");
    GeomView g = geom_view(const_cast<char*>(geom_buffer), P);
    BinningView b = binning_view(const_cast<char*>(binning_buffer), (size_t)num_rendered, width, height);
    launch_materialize_keys((size_t)num_rendered, b.tile_keys, b.point_list, g.depths, out_keys, (cudaStream_t)stream_);
    STAGE_CHECK(0, (cudaStream_t)stream_, "materialize keys");
    return GSR_OK;
}

extern "C" int gsr_backward(const gsr_backward_args* a, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (!a) return fail(GSR_ERR_INVALID, "null argument");
    const int P = a->P, W = a->width, H = a->height;
    if (P < 0 || W <= 0 || H <= 0 || a->R < 0) return fail(GSR_ERR_INVALID, "bad sizes");
    if (P == 0) return GSR_OK;
    if (!a->geom_buffer || !a->binning_buffer || !a->image_buffer || !a->dL_dpix || !a->radii)
        return fail(GSR_ERR_INVALID, "missing buffer");
    if (!a->dL_dmean2D || !a->dL_dconic || !a->dL_dopacity || !a->dL_dcolor || !a->dL_dmean3D || !a->dL_dcov3D)
        return fail(GSR_ERR_INVALID, "missing gradient output");
    if (a->dL_dinvdepth_pix && !a->dL_dinvdepth) return fail(GSR_ERR_INVALID, "dL_dinvdepth output required");
    if (a->shs && a->M > 0 && !a->dL_dsh) return fail(GSR_ERR_INVALID, "dL_dsh output required");
    if (a->scales && (!a->dL_dscale || !a->dL_drot || !a->rotations)) return fail(GSR_ERR_INVALID, "dL_dscale/dL_drot outputs required");
    if (check_aligned16(a->rotations, "rotations") || check_aligned16(a->dL_drot, "dL_drot") ||
        check_aligned16(a->dL_dconic, "dL_dconic") || (a->M == 16 && (check_aligned16(a->shs, "shs") || check_aligned16(a->dL_dsh, "dL_dsh"))))
        return GSR_ERR_INVALID;
    const int gx = (W + GSR_TILE - 1) / GSR_TILE, gy = (H + GSR_TILE - 1) / GSR_TILE;
    GeomView g = geom_view(a->geom_buffer, P);
    ImageView im = image_view(a->image_buffer, W, H);
    BinningView b = binning_view(a->binning_buffer, (size_t)a->R, W, H);

    BwdParams p;
    p.P = P; p.D = a->D; p.M = a->shs ? a->M : 0; p.W = W; p.H = H; p.gx = gx; p.gy = gy;
    p.background = a->background; p.means3D = a->means3D; p.shs = a->shs; p.colors_precomp = a->colors_precomp;
    p.opacities = a->opacities; p.scales = a->scales; p.rotations = a->rotations; p.cov3D_precomp = a->cov3D_precomp;
    p.viewmatrix = a->viewmatrix; p.projmatrix = a->projmatrix; p.cam_pos = a->cam_pos;
    p.scale_modifier = a->scale_modifier; p.tan_fovx = a->tan_fovx; p.tan_fovy = a->tan_fovy;
    p.focal_y = H / (2.0f * a->tan_fovy); p.focal_x = W / (2.0f * a->tan_fovx);
    p.antialiasing = a->antialiasing; p.radii = a->radii;
    p.accumulate = a->accumulate != 0;
    // per-view opacity gradient: the blend accumulates it in scratch (the depth-sort key buffer of the
    // forward, dead by now); preprocess_bwd applies the antialiasing factor and writes / adds the output
    p.dop_view = reinterpret_cast<float*>(g.gsort_keys[0]);
    p.dL_dpix = a->dL_dpix; p.dL_dinvdepth_pix = a->dL_dinvdepth_pix;
    p.dL_dmean2D = a->dL_dmean2D; p.dL_dconic = a->dL_dconic; p.dL_dopacity = a->dL_dopacity; p.dL_dcolor = a->dL_dcolor;
    p.dL_dinvdepth = a->dL_dinvdepth_pix ? a->dL_dinvdepth : nullptr;
    p.dL_dmean3D = a->dL_dmean3D; p.dL_dcov3D = a->dL_dcov3D; p.dL_dsh = a->dL_dsh; p.dL_dscale = a->dL_dscale; p.dL_drot = a->dL_drot;

    StageTimer tm(tl_profiling != 0, stream);
    tm.mark();
    // the five atomically accumulated arrays start from zero; everything else is fully written below
    const size_t sp = (size_t)P;
    CUDA_TRY(cudaMemsetAsync(a->dL_dmean2D, 0, sp * 3 * 4, stream));
    CUDA_TRY(cudaMemsetAsync(a->dL_dconic, 0, sp * 4 * 4, stream));
    CUDA_TRY(cudaMemsetAsync(p.dop_view, 0, sp * 4, stream));
    CUDA_TRY(cudaMemsetAsync(a->dL_dcolor, 0, sp * 3 * 4, stream));
    if (p.dL_dinvdepth) CUDA_TRY(cudaMemsetAsync(p.dL_dinvdepth, 0, sp * 4, stream));
    if (a->R > 0) {
        launch_render_bwd(p, im.ranges, b.point_list, b.block_masks, g.records, im.final_T, im.n_contrib, stream);
        STAGE_CHECK(a->debug, stream, "render backward");
    }
    tm.mark();
    launch_preprocess_bwd(p, g, stream);
    STAGE_CHECK(a->debug, stream, "preprocess backward");
    tm.mark();
    static const int ids[] = { 6, 7 };
    tm.finish(ids);
    return GSR_OK;
}

// ---- the views of a step in one call (include/gsr.h: gsr_forward_batch / gsr_backward_batch) ---------------------
static int batch_shared_check(int n, int P0, int D0, int M0, const float* m0, const float* sh0, const float* op0, const float* sc0,
                              const float* ro0, float mod0, int aa0, int P, int D, int M, const float* m, const float* sh,
                              const float* op, const float* sc, const float* ro, float mod, int aa, const float* colors,
                              const float* cov3D) {
    (void)n;
    if (P != P0 || D != D0 || M != M0 || m != m0 || sh != sh0 || op != op0 || sc != sc0 || ro != ro0 || mod != mod0 || aa != aa0)
        return fail(GSR_ERR_INVALID, "batch: the views must share the Gaussians, SH layout, scale_modifier and antialiasing");
    if (colors || cov3D || !sh || M != 16 || !sc || !ro)
        return fail(GSR_ERR_INVALID, "batch: needs shs with M = 16 and scales + rotations (no colors_precomp / cov3D_precomp)");
    return GSR_OK;
}

// Per-(thread, device) events of the batch entry points: [0] = the batched kernel has run on the caller's stream,
// [1 + v] = view v's work on its own stream (counts copied / chain finished).
static_assert(GSR_MAX_VIEWS == GSR_MAX_BATCH_VIEWS, "the public batch limit (gsr.h) and the kernels' view arrays (gsr_common.cuh) must agree");
static cudaEvent_t* batch_events(int dev_id) {
    static thread_local cudaEvent_t ev[64][1 + GSR_MAX_VIEWS] = {};
    if (!ev[dev_id][0])
        for (int k = 0; k < 1 + GSR_MAX_VIEWS; k++)
            if (cudaEventCreateWithFlags(&ev[dev_id][k], cudaEventDisableTiming) != cudaSuccess) return nullptr;
    return ev[dev_id];
}

extern "C" int gsr_forward_batch(int n_views, const gsr_forward_args* views, gsr_alloc_fn geom_alloc, void* const* geom_users,
                                 gsr_alloc_fn binning_alloc, void* const* binning_users, gsr_alloc_fn image_alloc,
                                 void* const* image_users, void* stream_, void* const* view_streams, int* num_rendered) {
    cudaStream_t stream = (cudaStream_t)stream_;
    // the per-view chains run on view_streams[v] when given (they are independent of each other), else on `stream`
    auto vstream = [&](int v) { return (view_streams && view_streams[v]) ? (cudaStream_t)view_streams[v] : stream; };
    if (!views || !num_rendered || !geom_alloc || !binning_alloc || !image_alloc || !geom_users || !binning_users || !image_users)
        return fail(GSR_ERR_INVALID, "null argument");
    if (n_views < 1 || n_views > GSR_MAX_VIEWS) return fail(GSR_ERR_INVALID, "batch: n_views = %d outside 1..%d", n_views, GSR_MAX_VIEWS);
    const gsr_forward_args& a0 = views[0];
    const int P = a0.P;
    for (int v = 0; v < n_views; v++) {
        const gsr_forward_args& a = views[v];
        num_rendered[v] = 0;
        if (a.P < 0 || a.width <= 0 || a.height <= 0) return fail(GSR_ERR_INVALID, "bad sizes");
        if (!a.out_color) return fail(GSR_ERR_INVALID, "out_color is null");
        if (P == 0) continue;
        if (!a.means3D || !a.opacities || !a.viewmatrix || !a.projmatrix || !a.cam_pos || !a.background || !a.radii)
            return fail(GSR_ERR_INVALID, "missing required input pointer");
        if (a.D < 0 || a.D > 3 || (a.D + 1) * (a.D + 1) > a.M) return fail(GSR_ERR_INVALID, "SH degree %d does not fit M=%d", a.D, a.M);
        if (batch_shared_check(n_views, a0.P, a0.D, a0.M, a0.means3D, a0.shs, a0.opacities, a0.scales, a0.rotations, a0.scale_modifier,
                               a0.antialiasing, a.P, a.D, a.M, a.means3D, a.shs, a.opacities, a.scales, a.rotations, a.scale_modifier,
                               a.antialiasing, a.colors_precomp, a.cov3D_precomp)) return GSR_ERR_INVALID;
        if (a.prefiltered != a0.prefiltered) return fail(GSR_ERR_INVALID, "batch: prefiltered differs between views");
    }
    if (P == 0) return GSR_OK;
    if (check_aligned16(a0.rotations, "rotations") || check_aligned16(a0.shs, "shs")) return GSR_ERR_INVALID;

    GeomView g[GSR_MAX_VIEWS]; ImageView im[GSR_MAX_VIEWS];
    FwdBatchParams bp;
    bp.P = P; bp.D = a0.D; bp.V = n_views; bp.prefiltered = a0.prefiltered; bp.antialiasing = a0.antialiasing;
    bp.scale_modifier = a0.scale_modifier;
    bp.means3D = a0.means3D; bp.shs = a0.shs; bp.opacities = a0.opacities; bp.scales = a0.scales; bp.rotations = a0.rotations;
    for (int v = 0; v < n_views; v++) {
        const gsr_forward_args& a = views[v];
        char* geom_raw = geom_alloc(gsr_geometry_bytes(P), geom_users[v]);
        if (!geom_raw) return fail(GSR_ERR_ALLOC, "geometry allocator returned null");
        char* img_raw = image_alloc(gsr_image_bytes(a.width, a.height), image_users[v]);
        if (!img_raw) return fail(GSR_ERR_ALLOC, "image allocator returned null");
        g[v] = geom_view(geom_raw, P);
        im[v] = image_view(img_raw, a.width, a.height);
        FwdBatchView& w = bp.view[v];
        w.viewmatrix = a.viewmatrix; w.projmatrix = a.projmatrix; w.cam_pos = a.cam_pos;
        w.tan_fovx = a.tan_fovx; w.tan_fovy = a.tan_fovy;
        w.focal_y = a.height / (2.0f * a.tan_fovy); w.focal_x = a.width / (2.0f * a.tan_fovx);
        w.W = a.width; w.H = a.height; w.gx = (a.width + GSR_TILE - 1) / GSR_TILE; w.gy = (a.height + GSR_TILE - 1) / GSR_TILE;
        w.depths = g[v].depths; w.records = g[v].records; w.cov3D = g[v].cov3D; w.clamped = g[v].clamped;
        w.tiles_touched = g[v].tiles_touched; w.scan_state = g[v].scan_state; w.radii = a.radii;
        CUDA_TRY(cudaMemsetAsync(g[v].scan_state, 0, gsr_scan_state_words((size_t)P) * 4, stream));
    }
    launch_preprocess_fwd_batch(bp, stream);
    STAGE_CHECK(a0.debug, stream, "preprocess (batch)");

    // per view: scan + counts on their way to the host, then the depth sort, which runs while the host waits for the counts
    static thread_local uint32_t* pinned = nullptr;
    int dev_id = 0;
    CUDA_TRY(cudaGetDevice(&dev_id));
    if (dev_id < 0 || dev_id >= 64) return fail(GSR_ERR_INVALID, "device ordinal %d out of range", dev_id);
    if (!pinned) CUDA_TRY(cudaHostAlloc((void**)&pinned, 4 * 4 * GSR_MAX_VIEWS, cudaHostAllocPortable));
    cudaEvent_t* ev = batch_events(dev_id);
    if (!ev) return fail(GSR_ERR_CUDA, "batch: cannot create events");
    CUDA_TRY(cudaEventRecord(ev[0], stream));
    // on every exit path -- errors included -- the caller's stream is ordered after whatever the view streams were given:
    // the scratch they touch is the caller's and may be released once this call returns
    struct JoinViews {
        cudaStream_t main; cudaStream_t sv[GSR_MAX_VIEWS]; cudaEvent_t* ev; int n = 0; bool armed = true;
        ~JoinViews() {
            if (!armed) return;
            for (int v = 0; v < n; v++)
                if (sv[v] != main) { cudaEventRecord(ev[1 + v], sv[v]); cudaStreamWaitEvent(main, ev[1 + v], 0); }
        }
    } join_views;
    join_views.main = stream; join_views.ev = ev;
    const int gsort_bits = 32, gsort_passes = (gsort_bits + 7) / 8;
    const int g0 = (gsort_passes & 1) ? 1 : 0;
    for (int v = 0; v < n_views; v++) {
        cudaStream_t sv = vstream(v);
        join_views.sv[v] = sv; join_views.n = v + 1;
        if (sv != stream) CUDA_TRY(cudaStreamWaitEvent(sv, ev[0], 0));
        launch_scan_compact(P, g[v], g[v].gsort_keys[g0], g[v].gsort_vals[g0], sv);
        CUDA_TRY(cudaMemcpyAsync(pinned + 4 * v, g[v].scan_state + 1, 12, cudaMemcpyDeviceToHost, sv));
        CUDA_TRY(cudaEventRecord(ev[1 + v], sv));
        int in_b = 0;
        launch_sort_pairs32_pingpong(g[v].gsort_keys[g0], g[v].gsort_vals[g0], g[v].gsort_keys[1 - g0], g[v].gsort_vals[1 - g0],
                                     (size_t)P, g[v].scan_state + 3, gsort_bits, g[v].gsort_temp, sv, &in_b);
    }
    STAGE_CHECK(a0.debug, stream, "scan + depth sort (batch)");
    for (int v = 0; v < n_views; v++) CUDA_TRY(cudaEventSynchronize(ev[1 + v]));
    for (int v = 0; v < n_views; v++) {
        if (pinned[4 * v]) return fail(GSR_ERR_PREFILTERED, "Point is filtered although prefiltered is set. This shouldn't happen!");
        if (pinned[4 * v + 1] >= (1u << 30)) return fail(GSR_ERR_INVALID, "num_rendered=%u exceeds the supported 2^30", pinned[4 * v + 1]);
        num_rendered[v] = (int)pinned[4 * v + 1];
    }
    for (int v = 0; v < n_views; v++) {
        const gsr_forward_args& a = views[v];
        cudaStream_t sv = vstream(v);
        const size_t N = (size_t)num_rendered[v];
        const int W = a.width, H = a.height, gx = bp.view[v].gx, gy = bp.view[v].gy;
        char* bin_raw = binning_alloc(gsr_binning_bytes(N, W, H), binning_users[v]);
        if (!bin_raw) return fail(GSR_ERR_ALLOC, "binning allocator returned null");
        BinningView b = binning_view(bin_raw, N, W, H);
        const int end_bit = (int)higher_msb((uint32_t)(gx * gy));
        const int passes = (end_bit + 7) / 8;
        uint32_t* k_start = (passes & 1) ? b.tile_keys_alt : b.tile_keys;
        uint32_t* v_start = (passes & 1) ? b.point_list_alt : b.point_list;
        uint32_t* k_other = (passes & 1) ? b.tile_keys : b.tile_keys_alt;
        uint32_t* v_other = (passes & 1) ? b.point_list : b.point_list_alt;
        if (N > 0) {
            launch_sorted_scan(P, g[v], g[v].depth_order, g[v].sorted_offsets, sv);
            launch_emit_tiles(P, gx, gy, g[v], g[v].depth_order, g[v].sorted_offsets, a.radii, k_start, v_start, sv);
            int in_b = 0;
            launch_sort_pairs32_pingpong(k_start, v_start, k_other, v_other, N, nullptr, end_bit, b.sort_temp, sv, &in_b);
        }
        launch_tile_ranges(N, b.tile_keys, im[v].ranges, gx * gy, sv);
        launch_render_fwd(W, H, gx, gy, im[v].ranges, b.point_list, b.block_masks, g[v].records, a.background, im[v].final_T,
                          im[v].n_contrib, a.out_color, a.out_invdepth, sv);
        STAGE_CHECK(a.debug, sv, "render (batch)");
        if (sv != stream) {             // the caller's stream continues after every view's chain
            CUDA_TRY(cudaEventRecord(ev[1 + v], sv));
            CUDA_TRY(cudaStreamWaitEvent(stream, ev[1 + v], 0));
        }
    }
    join_views.armed = false;           // joined view by view above
    return GSR_OK;
}

// stages: 1 = backward blend of every view, 2 = the one per-Gaussian launch over all views, 3 = both
static int backward_batch_impl(int n_views, const gsr_backward_args* views, void* stream_, void* const* view_streams, int stages) {
    cudaStream_t stream = (cudaStream_t)stream_;
    auto vstream = [&](int v) { return (view_streams && view_streams[v]) ? (cudaStream_t)view_streams[v] : stream; };
    // every argument is checked before anything is enqueued: a rejected call leaves the streams untouched
    if (!views) return fail(GSR_ERR_INVALID, "null argument");
    if (n_views < 1 || n_views > GSR_MAX_VIEWS) return fail(GSR_ERR_INVALID, "batch: n_views = %d outside 1..%d", n_views, GSR_MAX_VIEWS);
    const gsr_backward_args& a0 = views[0];
    const int P = a0.P;
    if (P < 0) return fail(GSR_ERR_INVALID, "bad sizes");
    if (P == 0) return GSR_OK;
    if ((stages & 2) && (!a0.dL_dmean3D || !a0.dL_dsh || !a0.dL_dopacity || !a0.dL_dscale || !a0.dL_drot))
        return fail(GSR_ERR_INVALID, "batch: views[0] must carry the five parameter-gradient outputs");
    if (check_aligned16(a0.rotations, "rotations") || check_aligned16(a0.dL_drot, "dL_drot") || check_aligned16(a0.shs, "shs") ||
        check_aligned16(a0.dL_dsh, "dL_dsh")) return GSR_ERR_INVALID;
    for (int v = 0; v < n_views; v++) {
        const gsr_backward_args& a = views[v];
        if (a.width <= 0 || a.height <= 0 || a.R < 0) return fail(GSR_ERR_INVALID, "bad sizes");
        if (a.D < 0 || a.D > 3 || (a.D + 1) * (a.D + 1) > a.M) return fail(GSR_ERR_INVALID, "SH degree %d does not fit M=%d", a.D, a.M);
        if (batch_shared_check(n_views, a0.P, a0.D, a0.M, a0.means3D, a0.shs, a0.opacities, a0.scales, a0.rotations, a0.scale_modifier,
                               a0.antialiasing, a.P, a.D, a.M, a.means3D, a.shs, a.opacities, a.scales, a.rotations, a.scale_modifier,
                               a.antialiasing, a.colors_precomp, a.cov3D_precomp)) return GSR_ERR_INVALID;
        if (!a.means3D || !a.opacities) return fail(GSR_ERR_INVALID, "missing required input pointer");
        if ((stages & 2) && (!a.viewmatrix || !a.projmatrix || !a.cam_pos)) return fail(GSR_ERR_INVALID, "missing camera pointer");
        if ((stages & 1) && !a.background) return fail(GSR_ERR_INVALID, "background is null");
        if (!a.geom_buffer || !a.radii || ((stages & 1) && (!a.binning_buffer || !a.image_buffer || !a.dL_dpix)))
            return fail(GSR_ERR_INVALID, "missing buffer");
        if (!a.dL_dmean2D || !a.dL_dconic || !a.dL_dcolor) return fail(GSR_ERR_INVALID, "missing per-view gradient output");
        if (a.dL_dinvdepth_pix && !a.dL_dinvdepth) return fail(GSR_ERR_INVALID, "dL_dinvdepth output required");
        if (check_aligned16(a.dL_dconic, "dL_dconic")) return GSR_ERR_INVALID;
    }
    int dev_id = 0;
    CUDA_TRY(cudaGetDevice(&dev_id));
    if (dev_id < 0 || dev_id >= 64) return fail(GSR_ERR_INVALID, "device ordinal %d out of range", dev_id);
    cudaEvent_t* ev = batch_events(dev_id);
    if (!ev) return fail(GSR_ERR_CUDA, "batch: cannot create events");
    CUDA_TRY(cudaEventRecord(ev[0], stream));
    BwdBatchParams bp;
    bp.P = P; bp.D = a0.D; bp.V = n_views; bp.antialiasing = a0.antialiasing; bp.accumulate = a0.accumulate != 0;
    bp.scale_modifier = a0.scale_modifier;
    bp.means3D = a0.means3D; bp.shs = a0.shs; bp.opacities = a0.opacities; bp.scales = a0.scales; bp.rotations = a0.rotations;
    bp.dL_dmean3D = a0.dL_dmean3D; bp.dL_dsh = a0.dL_dsh; bp.dL_dopacity = a0.dL_dopacity; bp.dL_dscale = a0.dL_dscale; bp.dL_drot = a0.dL_drot;
    for (int v = 0; v < n_views; v++) {
        const gsr_backward_args& a = views[v];
        const int W = a.width, H = a.height;
        const int gx = (W + GSR_TILE - 1) / GSR_TILE, gy = (H + GSR_TILE - 1) / GSR_TILE;
        GeomView g = geom_view(a.geom_buffer, P);
        BwdParams p;
        memset(&p, 0, sizeof(p));
        p.P = P; p.D = a.D; p.M = a.M; p.W = W; p.H = H; p.gx = gx; p.gy = gy;
        p.background = a.background;
        p.dop_view = reinterpret_cast<float*>(g.gsort_keys[0]);
        p.dL_dpix = a.dL_dpix; p.dL_dinvdepth_pix = a.dL_dinvdepth_pix;
        p.dL_dmean2D = a.dL_dmean2D; p.dL_dconic = a.dL_dconic; p.dL_dcolor = a.dL_dcolor;
        // stage 2 alone has no pixel gradients to look at: a non-null dL_dinvdepth output says the blend filled it
        p.dL_dinvdepth = (a.dL_dinvdepth_pix || !(stages & 1)) ? a.dL_dinvdepth : nullptr;
        const size_t sp = (size_t)P;
        cudaStream_t sv = vstream(v);
        if (stages & 1) {
        ImageView im = image_view(a.image_buffer, W, H);
        BinningView b = binning_view(a.binning_buffer, (size_t)a.R, W, H);
        if (sv != stream) CUDA_TRY(cudaStreamWaitEvent(sv, ev[0], 0));
        CUDA_TRY(cudaMemsetAsync(a.dL_dmean2D, 0, sp * 3 * 4, sv));
        CUDA_TRY(cudaMemsetAsync(a.dL_dconic, 0, sp * 4 * 4, sv));
        CUDA_TRY(cudaMemsetAsync(p.dop_view, 0, sp * 4, sv));
        CUDA_TRY(cudaMemsetAsync(a.dL_dcolor, 0, sp * 3 * 4, sv));
        if (p.dL_dinvdepth) CUDA_TRY(cudaMemsetAsync(p.dL_dinvdepth, 0, sp * 4, sv));
        if (a.R > 0) {
            launch_render_bwd(p, im.ranges, b.point_list, b.block_masks, g.records, im.final_T, im.n_contrib, sv);
            STAGE_CHECK(a.debug, sv, "render backward (batch)");
        }
        if (sv != stream) {             // the one per-Gaussian launch below needs every view's blend gradients
            CUDA_TRY(cudaEventRecord(ev[1 + v], sv));
            CUDA_TRY(cudaStreamWaitEvent(stream, ev[1 + v], 0));
        }
        }
        BwdBatchView& w = bp.view[v];
        w.viewmatrix = a.viewmatrix; w.projmatrix = a.projmatrix; w.cam_pos = a.cam_pos;
        w.tan_fovx = a.tan_fovx; w.tan_fovy = a.tan_fovy;
        w.focal_y = H / (2.0f * a.tan_fovy); w.focal_x = W / (2.0f * a.tan_fovx);
        w.radii = a.radii; w.clamped = g.clamped;
        w.dL_dconic = a.dL_dconic; w.dL_dmean2D = a.dL_dmean2D; w.dL_dcolor = a.dL_dcolor; w.dL_dinvdepth = p.dL_dinvdepth;
        w.dop_view = p.dop_view;
    }
    if (stages & 2) {
        launch_preprocess_bwd_batch(bp, stream);
        STAGE_CHECK(a0.debug, stream, "preprocess backward (batch)");
    }
    return GSR_OK;
}

extern "C" int gsr_backward_batch(int n_views, const gsr_backward_args* views, void* stream, void* const* view_streams) {
    return backward_batch_impl(n_views, views, stream, view_streams, 3);
}
// The two halves of gsr_backward_batch for callers that learn about the views of a step one at a time (an autograd engine
// calls one backward per view): the blend of a view now, the per-Gaussian launch over all of them once the last is in.
extern "C" int gsr_backward_blend(const gsr_backward_args* view, void* stream) {
    return backward_batch_impl(1, view, stream, nullptr, 1);
}
extern "C" int gsr_backward_params_batch(int n_views, const gsr_backward_args* views, void* stream) {
    return backward_batch_impl(n_views, views, stream, nullptr, 2);
}

extern "C" int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                                unsigned char* present, void* stream_) {
    (void)projmatrix;   // the reference's projected-bounds test is commented out (auxiliary.h:166)
    if (P < 0) return fail(GSR_ERR_INVALID, "bad P");
    if (P == 0) return GSR_OK;
    if (!means3D || !viewmatrix || !present) return fail(GSR_ERR_INVALID, "null argument");
    launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream_);
    STAGE_CHECK(0, (cudaStream_t)stream_, "markVisible");
    return GSR_OK;
}

extern "C" size_t gsr_sort_temp_bytes(size_t n) { return sort_temp_bytes(n) + 256; }

extern "C" int gsr_sort_pairs(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out, size_t n,
                              int end_bit, char* temp, void* stream_) {
    if (end_bit < 0 || end_bit > 64) return fail(GSR_ERR_INVALID, "end_bit out of range");
    if (n >= ((size_t)1 << 30)) return fail(GSR_ERR_INVALID, "n exceeds 2^30");
    if (n && (!keys_in || !vals_in || !keys_out || !vals_out || !temp)) return fail(GSR_ERR_INVALID, "null argument");
    launch_sort_pairs(keys_in, vals_in, keys_out, vals_out, n, end_bit, align_base(temp), (cudaStream_t)stream_);
    STAGE_CHECK(0, (cudaStream_t)stream_, "sort");
    return GSR_OK;
}
