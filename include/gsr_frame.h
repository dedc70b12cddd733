/*
 * gsr_frame.h -- C ABI of the frame output that follows a forward-only render (headless
 * "device buffer -> bytes / PNG" in place of the network viewer and torchvision's image writer; same library).
 *
 * Replaces
 *   /root/reference/train.py:81          (torch.clamp(net_image, 0, 1) * 255).byte().permute(1, 2, 0).contiguous()
 *                                        -- the frame network_gui.send() ships to the viewer (GSR_FRAME_TRUNCATE)
 *   /root/reference/render.py:44-45      torchvision.utils.save_image: x.mul(255).add_(0.5).clamp_(0, 255)
 *                                        .permute(1, 2, 0).to(uint8)  (GSR_FRAME_ROUND) -- the pixels of the PNG
 * by ONE launch: planar fp32 [3,H,W] in, interleaved 8-bit [H,W,3] out (five / four torch kernels and an HWC fp32
 * intermediate in the reference).  Integer output: bit-exact against the torch expressions.  NaN pixels give 0.
 */
#ifndef GSR_FRAME_H_INCLUDED
#define GSR_FRAME_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSR_FRAME_TRUNCATE 0 /* floor(clamp(x, 0, 1) * 255)          train.py:81   */
#define GSR_FRAME_ROUND 1    /* floor(clamp(x * 255 + 0.5, 0, 255))  render.py:44  */

/* image: device fp32 [3,H,W] contiguous; rgb8: device bytes [H,W,3], fully written.
 * x_begin / x_end select the column range [x_begin, x_end) of every row (render.py:40-42 keeps the right half when
 * train_test_exp is set); the output is [H, x_end - x_begin, 3].  Pass 0, W for the whole image. */
int gsr_frame_to_rgb8(int H, int W, int x_begin, int x_end, int mode, const float* image, uint8_t* rgb8, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GSR_FRAME_H_INCLUDED */
