/*
 * gsr_loss.h -- C ABI of the loss stage that follows the rasterizer in every training iteration
 * (same library, libgsr_b200.so).
 *
 * Replaces the reference's fused-ssim extension
 *   /root/reference/submodules/fused-ssim/ssim.h:8-26  (fusedssim, fusedssim_backward)
 *   /root/reference/submodules/fused-ssim/ssim.cu:188-444
 * and, as one fused pair of launches, the loss of /root/reference/train.py:118-126
 *   loss = (1 - lambda_dssim) * l1_loss(image, gt) + lambda_dssim * (1 - fused_ssim(image, gt))
 * (l1_loss: /root/reference/utils/loss_utils.py:40-41).
 *
 * Conventions as in gsr.h: device pointers, fp32, contiguous, NCHW; an image batch [B,CH,H,W] is passed
 * as `planes` = B*CH independent [H,W] planes; every function returns GSR_OK or a negative code and
 * never throws; `stream` is a cudaStream_t passed as void*; no global mutable state.
 */
#ifndef GSR_LOSS_H_INCLUDED
#define GSR_LOSS_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/*
 * fusedssim (ssim.cu:373-409): the SSIM map of two image batches under the 11x11 Gaussian window
 * (sigma 1.5) with zero padding, and -- when the three dm_* pointers are non-NULL (train = true) -- the
 * partial derivatives of the map w.r.t. mu1, sigma1^2 and sigma12 that the backward needs.
 * All outputs [planes,H,W], fully written.  ssim_map may be NULL only together with the dm_* outputs
 * being non-NULL (derivatives only).
 */
int gsr_ssim_forward(int planes, int H, int W, float C1, float C2, const float* img1, const float* img2,
                     float* ssim_map, float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12, void* stream);

/*
 * fusedssim_backward (ssim.cu:411-444): dL_dimg1 [planes,H,W] (fully written) from dL_dmap and the three
 * saved derivative maps:  conv(dL*dm_dmu1) + 2*img1*conv(dL*dm_dsigma1_sq) + img2*conv(dL*dm_dsigma12).
 */
int gsr_ssim_backward(int planes, int H, int W, const float* img1, const float* img2, const float* dL_dmap,
                      const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12,
                      float* dL_dimg1, void* stream);

/*
 * The training loss in two launches (train.py:118-126 with padding "same").
 *
 * Forward: one pass over image and ground truth produces the three derivative maps (when non-NULL)
 * and per-CTA partial sums of |image - gt| and of the SSIM map; a one-CTA reduction (fixed order,
 * double accumulation: deterministic) writes loss_out[0] = loss, [1] = mean L1, [2] = mean SSIM.
 * The SSIM map itself is never written.  `partials` must hold gsr_l1_ssim_partials_bytes(...) bytes.
 *
 * Backward: dL_dimage = upstream * [ (1-lambda)/n * sign(image-gt) - lambda/n * d(sum ssim)/d image ],
 * n = planes*H*W, fully written.  `upstream` is a DEVICE scalar (autograd's grad of the loss; NULL = 1),
 * so no host synchronisation is needed between the two launches.
 */
size_t gsr_l1_ssim_partials_bytes(int planes, int H, int W);
int gsr_l1_ssim_loss_forward(int planes, int H, int W, float C1, float C2, float lambda_dssim, const float* image,
                             const float* gt, float* loss_out, float* dm_dmu1, float* dm_dsigma1_sq,
                             float* dm_dsigma12, char* partials, void* stream);
int gsr_l1_ssim_loss_backward(int planes, int H, int W, float lambda_dssim, const float* image, const float* gt,
                              const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12,
                              const float* upstream, float* dL_dimage, void* stream);

/*
 * Render epilogue of /root/reference/gaussian_renderer/__init__.py:112-119 (between the rasterizer and the loss):
 *   if use_trained_exp:  y[c] = sum_k x[k] * exposure[k][c] + exposure[c][3]      (exposure: row-major [3,4])
 *   y = clamp(y, 0, 1)
 * on one [3,H,W] image, as ONE launch (the reference: permute, matmul, permute, add, clamp).  exposure == NULL skips
 * the affine map (use_trained_exp false).  out is fully written.
 */
int gsr_exposure_clamp_forward(int H, int W, const float* image, const float* exposure, float* out, void* stream);

/*
 * Its backward: dL_dimage[k] = sum_c g[c] * exposure[k][c] with g[c] = dL_dout[c] where the unclamped value lies in
 * [0, 1] (torch.clamp's gradient mask) and 0 elsewhere; dL_dexposure[k][c] = sum_pix x[k] g[c],
 * dL_dexposure[c][3] = sum_pix g[c] (12 floats, row-major [3,4]; NULL when exposure is NULL or not trained).
 * The pixel sums are reduced in a fixed order (per-CTA partials in `partials`, gsr_exposure_partials_bytes(H, W)
 * bytes, then one CTA in double): deterministic.
 */
size_t gsr_exposure_partials_bytes(int H, int W);
int gsr_exposure_clamp_backward(int H, int W, const float* image, const float* exposure, const float* dL_dout,
                                float* dL_dimage, float* dL_dexposure, char* partials, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GSR_LOSS_H_INCLUDED */
