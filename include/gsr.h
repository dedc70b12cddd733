/*
 * gsr.h -- C ABI of the B200-native differentiable Gaussian-splatting rasterizer (libgsr_b200.so).
 *
 * This is the drop-in boundary.  Every entry point replaces one static method of the reference's
 * torch-free core class CudaRasterizer::Rasterizer
 * (/root/reference/diff-gaussian-rasterization-npu/cuda_rasterizer/rasterizer.h:24-90), with the
 * same argument meaning, so the reference's binding layer (rasterize_points.cu) can call it
 * unchanged in spirit: plain pointers and sizes, no torch types, no C++ types, no exceptions.
 *
 * Conventions
 *   - all pointers are DEVICE pointers on the current CUDA device unless the name says host;
 *   - all floats are IEEE fp32, all arrays contiguous, layouts as in the reference
 *     (means3D[P,3], scales[P,3], rotations[P,4] (w,x,y,z), opacities[P], shs[P,M,3],
 *      viewmatrix/projmatrix[16] with memory index m[4*col+row], out_color[3,H,W] planar);
 *   - a NULL pointer selects the same code path an empty tensor selects in the reference
 *     (shs==NULL <-> colors_precomp path, scales==NULL <-> cov3D_precomp path,
 *      dL_dout_invdepth==NULL <-> no inverse-depth gradient);
 *   - every function returns GSR_OK (0) or a negative error code and never throws;
 *     gsr_last_error() returns the calling thread's last message;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream, which is what the
 *     reference launches on).  Entry points are re-entrant and keep no global mutable state
 *     (the one exception is gsr_sort_pairs_static, which says so).
 */
#ifndef GSR_H_INCLUDED
#define GSR_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSR_OK 0
#define GSR_ERR_INVALID (-1)   /* bad argument (reference: AT_ERROR / std::runtime_error) */
#define GSR_ERR_CUDA (-2)      /* a CUDA call failed (reference: CHECK_CUDA, auxiliary.h:178-185) */
#define GSR_ERR_PREFILTERED (-3) /* a Gaussian was culled although prefiltered=1 (reference: __trap, auxiliary.h:168-172) */
#define GSR_ERR_ALLOC (-4)     /* an allocator callback returned NULL */

/*
 * Scratch allocator callback: must return a device buffer of at least `bytes` bytes that stays
 * valid until the matching gsr_backward has run.  Replaces std::function<char*(size_t)>
 * (rasterizer.h:36-38; the reference's binding resizes a torch byte tensor, rasterize_points.cu:27-33).
 */
typedef char* (*gsr_alloc_fn)(size_t bytes, void* user);

/* Arguments of Rasterizer::forward, rasterizer.h:35-61. */
typedef struct gsr_forward_args {
    int P;                       /* number of Gaussians */
    int D;                       /* active SH degree (0..3) */
    int M;                       /* SH coefficients per Gaussian in `shs` (0 if shs==NULL) */
    int width, height;
    const float* background;     /* [3] */
    const float* means3D;        /* [P,3] */
    const float* shs;            /* [P,M,3] or NULL */
    const float* colors_precomp; /* [P,3] or NULL */
    const float* opacities;      /* [P] */
    const float* scales;         /* [P,3] or NULL */
    float scale_modifier;
    const float* rotations;      /* [P,4] or NULL */
    const float* cov3D_precomp;  /* [P,6] or NULL */
    const float* viewmatrix;     /* [16] */
    const float* projmatrix;     /* [16] */
    const float* cam_pos;        /* [3] */
    float tan_fovx, tan_fovy;
    int prefiltered;
    int antialiasing;
    int debug;                   /* 1: synchronise + check after every stage (CHECK_CUDA) */
    float* out_color;            /* [3,H,W], fully written */
    float* out_invdepth;         /* [1,H,W] or NULL, fully written */
    int* radii;                  /* [P], fully written */
} gsr_forward_args;

/* Arguments of Rasterizer::backward, rasterizer.h:63-89. */
typedef struct gsr_backward_args {
    int P, D, M;
    int R;                       /* num_rendered returned by the matching forward */
    int width, height;
    const float* background;
    const float* means3D;
    const float* shs;
    const float* colors_precomp;
    const float* opacities;
    const float* scales;
    float scale_modifier;
    const float* rotations;
    const float* cov3D_precomp;
    const float* viewmatrix;
    const float* projmatrix;
    const float* cam_pos;
    float tan_fovx, tan_fovy;
    const int* radii;            /* [P] from forward */
    char* geom_buffer;           /* the three scratch buffers the forward filled */
    char* binning_buffer;
    char* image_buffer;
    const float* dL_dpix;        /* [3,H,W] */
    const float* dL_dinvdepth_pix; /* [1,H,W] or NULL */
    /* Outputs.  All are FULLY written by gsr_backward (rows of invisible Gaussians get zeros), so
     * the caller may pass uninitialised memory; the reference needed ten torch::zeros
     * (rasterize_points.cu:163-172). */
    float* dL_dmean2D;           /* [P,3]  (.x,.y written, .z = 0) */
    float* dL_dconic;            /* [P,4]  (.x,.y,.w written, .z = 0) */
    float* dL_dopacity;          /* [P] */
    float* dL_dcolor;            /* [P,3] */
    float* dL_dinvdepth;         /* [P] or NULL (required iff dL_dinvdepth_pix != NULL) */
    float* dL_dmean3D;           /* [P,3] */
    float* dL_dcov3D;            /* [P,6] */
    float* dL_dsh;               /* [P,M,3] or NULL when M==0 */
    float* dL_dscale;            /* [P,3] or NULL when scales==NULL */
    float* dL_drot;              /* [P,4] or NULL when rotations==NULL */
    int antialiasing;
    int debug;
    /* 0 (the reference's behaviour): the outputs are overwritten.  1: the five PARAMETER gradients
     * (dL_dmean3D, dL_dsh, dL_dopacity, dL_dscale, dL_drot) are ADDED to what the buffers hold, rows
     * of invisible Gaussians are left untouched -- several views of one training step accumulate
     * into one gradient bucket (what autograd does for the reference when a step renders more than
     * one camera) without a per-view buffer and add pass.  The per-view intermediates (dL_dmean2D,
     * dL_dconic, dL_dcolor, dL_dinvdepth, dL_dcov3D) are overwritten in both modes. */
    int accumulate;
} gsr_backward_args;

/*
 * Rasterizer::forward (rasterizer.h:35, rasterizer_impl.cu:198-341).  Calls the three allocators
 * (geometry, binning, image -- in that order, binning after the instance count is known), runs
 * preprocess -> scan -> duplicateWithKeys -> radix sort -> tile ranges -> blend on `stream`, and
 * stores the number of rendered (Gaussian,tile) instances in *num_rendered (host int).  It
 * synchronises `stream` once (the reference's blocking cudaMemcpy, rasterizer_impl.cu:284).
 */
int gsr_forward(const gsr_forward_args* args,
                gsr_alloc_fn geom_alloc, void* geom_user,
                gsr_alloc_fn binning_alloc, void* binning_user,
                gsr_alloc_fn image_alloc, void* image_user,
                void* stream, int* num_rendered);

/* Rasterizer::backward (rasterizer.h:63, rasterizer_impl.cu:345-450).  Asynchronous on `stream`. */
int gsr_backward(const gsr_backward_args* args, void* stream);

/*
 * The views of one step in one call.  The reference renders the cameras of a step -- a
 * training batch, or the frames of render.py:30-46 -- one Rasterizer::forward / backward after another
 * (train.py:74-87, :111), and every call reads every Gaussian again.  Here the per-Gaussian stages run ONCE for all
 * n_views cameras: parameters and the SH row are read once in the forward (preprocess_fwd_batch_kernel) and the 59
 * parameter-gradient floats are written once in the backward (preprocess_bwd_batch_kernel, the views' gradients summed
 * in registers); everything between (scan, sorts, key emission, tile ranges, the two blend kernels) runs per view as in
 * gsr_forward / gsr_backward, and every per-view output is bit-identical to the single-view call's.
 *
 * views[v] are ordinary argument structs that differ in camera, image size and per-view outputs only: P, D, M (= 16),
 * means3D, shs, opacities, scales, rotations, scale_modifier, prefiltered and antialiasing must be the same in all of them,
 * and colors_precomp / cov3D_precomp must be NULL (other configurations: call the single-view entry points).
 * 1 <= n_views <= GSR_MAX_BATCH_VIEWS.  The allocators are called once per view with that view's user pointer.
 * gsr_backward_batch writes (accumulate = 0) or adds (accumulate = 1, taken from views[0]) the SUM over the views into the
 * parameter-gradient outputs of views[0] (dL_dmean3D, dL_dsh, dL_dopacity, dL_dscale, dL_drot); the per-view outputs
 * dL_dmean2D, dL_dconic, dL_dcolor, dL_dinvdepth are those of each view; dL_dcov3D is not produced (it only exists per view).
 * view_streams (optional): the per-view chains are independent of each other, so view v's may run on view_streams[v]
 * (several views may share a stream); the one-launch stages run on `stream`, which is ordered before the chains and
 * continues after all of them.  NULL (or a NULL entry): everything on `stream`.
 */
#define GSR_MAX_BATCH_VIEWS 8   /* the kernels' own name for this limit is GSR_MAX_VIEWS (csrc/gsr_common.cuh); a static_assert keeps them equal */
int gsr_forward_batch(int n_views, const gsr_forward_args* views,
                      gsr_alloc_fn geom_alloc, void* const* geom_users,
                      gsr_alloc_fn binning_alloc, void* const* binning_users,
                      gsr_alloc_fn image_alloc, void* const* image_users,
                      void* stream, void* const* view_streams /* [n_views] or NULL */, int* num_rendered /* [n_views] */);
int gsr_backward_batch(int n_views, const gsr_backward_args* views, void* stream, void* const* view_streams /* or NULL */);
/* gsr_backward_batch in two calls, for callers that are handed the views of a step one at a time (an autograd engine runs
 * one backward per view): gsr_backward_blend does the backward blend of ONE view now (its dL_dmean2D / dL_dconic /
 * dL_dcolor / dL_dinvdepth outputs and the view's geometry buffer must stay alive), gsr_backward_params_batch the one
 * per-Gaussian launch over all of them later (same argument structs; pixel gradients, binning and image buffers are not
 * read any more; a non-NULL dL_dinvdepth means the blend produced it). */
int gsr_backward_blend(const gsr_backward_args* view, void* stream);
int gsr_backward_params_batch(int n_views, const gsr_backward_args* views, void* stream);

/* Rasterizer::markVisible (rasterizer.h:28-33, rasterizer_impl.cu:141-153). present is bool[P] (1 byte each). */
int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     unsigned char* present, void* stream);

/* Scratch sizes (reference: required<GeometryState>(P) etc., rasterizer_impl.h:66-72). */
size_t gsr_geometry_bytes(int P);
size_t gsr_image_bytes(int width, int height);
size_t gsr_binning_bytes(size_t num_rendered, int width, int height);

/*
 * Stable LSD radix sort of n (u64 key, u32 value) pairs on key bits [0, end_bit): the hand-written
 * onesweep that replaces cub::DeviceRadixSort::SortPairs (rasterizer_impl.cu:306-311) as a
 * stand-alone primitive (gsr_forward uses the same kernel template on 32-bit keys, twice).  `temp` must
 * hold gsr_sort_temp_bytes(n) bytes.  keys_in/vals_in are clobbered (ping-pong); the result is in
 * keys_out/vals_out.  Exposed so the sort can be tested and timed alone.
 */
size_t gsr_sort_temp_bytes(size_t n);
int gsr_sort_pairs(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                   size_t n, int end_bit, char* temp, void* stream);

/*
 * gsr_sort_pairs for callers that cannot allocate device memory.  (No counterpart in the reference:
 * its sort always draws temp storage from the torch allocator through the binning-buffer callback.)
 *
 * Who needs it: code that is handed its input and output buffers by someone else and has no device
 * allocator of its own -- a plugin inside a host application that owns the device heap, a frame loop
 * that bans allocator calls after start-up, an FFI binding without torch.  Same sort, same kernels,
 * same result as gsr_sort_pairs (stable, key bits [0, end_bit), inputs clobbered, result in
 * keys_out/vals_out), but the histogram / ticket / look-back status block is a fixed array built into
 * libgsr_b200.so instead of a `temp` argument.  The call itself performs no cudaMalloc/cudaFree and
 * uses no allocating library routine.
 *
 * Limits and rules
 *   - n <= GSR_SORT_STATIC_MAX_PAIRS (also returned by gsr_sort_static_max_pairs()); a larger n is
 *     refused with GSR_ERR_INVALID and a message naming the limit -- use gsr_sort_pairs then.
 *   - The array is about 2.1 MB and is held by every device that loads the library, until exit.
 *   - The first call on each device looks up the array's device address (the device must be current,
 *     as for every entry point) and keeps it; make that first call before starting any stream
 *     capture.  Later calls on that device only enqueue a memset and the sort kernels on `stream`.
 *   - There is ONE array per device and every call uses it.  Each call clears the words it needs
 *     first, so nothing has to be reset between calls, but two calls whose GPU work can overlap
 *     (different streams without an ordering between them, or two host threads) race on it and
 *     produce garbage.  Order them on one stream or with events.
 */
#define GSR_SORT_STATIC_MAX_PAIRS (1u << 20)
size_t gsr_sort_static_max_pairs(void);
int gsr_sort_pairs_static(uint64_t* keys_in, uint32_t* vals_in, uint64_t* keys_out, uint32_t* vals_out,
                          size_t n, int end_bit, void* stream);

/*
 * Introspection of the opaque scratch buffers for the parity tests (the reference's tests would
 * use GeometryState/BinningState/ImageState::fromChunk, rasterizer_impl.h:29-65).  Fills byte
 * offsets of each sub-array relative to the buffer base.
 */
typedef struct gsr_geometry_layout {
    size_t depths;        /* f32[P] */
    size_t records;       /* f32[P,12]: x,y,conic.a,conic.c | conic.b,opacity,r,g | b,1/depth,0,0  (conic = [a b; b c]) */
    size_t cov3D;         /* f32[P,6] */
    size_t clamped;       /* u8[P]: bit c set <=> colour channel c was clamped */
    size_t tiles_touched; /* u32[P] */
    size_t point_offsets; /* u32[P] inclusive scan (index order, as the reference's) */
    size_t depth_order;   /* u32[P], first P_vis valid: visible Gaussians in (depth bits, index) order */
    size_t sorted_offsets;/* u32[P], first P_vis valid: inclusive scan of tiles_touched in depth order */
    size_t total;
} gsr_geometry_layout;
typedef struct gsr_binning_layout {
    size_t tile_keys;     /* u32[N] sorted tile ids = high words of the reference's sorted keys */
    size_t point_list;    /* u32[N] sorted Gaussian indices */
    size_t tile_keys_alt; /* u32[N] ping-pong partners (scratch after the sort) */
    size_t point_list_alt;
    size_t sort_temp;
    size_t total;
} gsr_binning_layout;
typedef struct gsr_image_layout {
    size_t final_T;       /* f32[H*W] */
    size_t n_contrib;     /* u32[H*W] */
    size_t ranges;        /* u32[T,2] */
    size_t total;
} gsr_image_layout;
void gsr_get_geometry_layout(int P, gsr_geometry_layout* out);
void gsr_get_binning_layout(size_t num_rendered, int width, int height, gsr_binning_layout* out);
void gsr_get_image_layout(int width, int height, gsr_image_layout* out);

/*
 * Rebuilds the reference's sorted 64-bit keys (tile id << 32 | float bits of depth,
 * rasterizer_impl.cu:100-104) from the scratch buffers of a forward into out_keys (u64[num_rendered],
 * device).  The hot path never materialises them: it sorts Gaussians by depth once and instances
 * by tile id (see DESIGN.md "binning"); this entry point exists for parity tests and debugging.
 */
int gsr_materialize_keys(int P, int num_rendered, int width, int height, const char* geom_buffer,
                         const char* binning_buffer, uint64_t* out_keys, void* stream);

/* Per-stage device times of the calling thread's last gsr_forward/gsr_backward when profiling is
 * enabled with gsr_set_profiling(1) (adds event records + one extra sync; off by default).
 * Stage ids: 0 preprocess, 1 scan+compact, 2 duplicate (sorted scan + tile emission), 3 tile sort,
 *            4 ranges, 5 render_fwd, 6 render_bwd, 7 preprocess_bwd, 8 host wait for num_rendered
 *            (device idle gap), 9 depth sort of the visible Gaussians.
 * Returns milliseconds or -1. */
void gsr_set_profiling(int enabled);
float gsr_last_stage_ms(int stage);

const char* gsr_last_error(void);
const char* gsr_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GSR_H_INCLUDED */
