/*
 * gsr_optim.h -- C ABI of the step that consumes the rasterizer's gradients in every training iteration
 * (same library, libgsr_b200.so): the Adam update of the Gaussian parameters
 * and the densification statistics.
 *
 * Replaces, for the flat parameter/gradient bucket of gsr_b200/dist.py,
 *   /root/reference/scene/gaussian_model.py:178-201   torch.optim.Adam(l, lr=0.0, eps=1e-15) over six parameter
 *                                                     groups with their own learning rates
 *   /root/reference/train.py:176-186                  optimizer.step(); optimizer.zero_grad()
 *   /root/reference/train.py:163-167, scene/gaussian_model.py:471-473
 *                                                     max_radii2D / xyz_gradient_accum / denom updates
 * torch.optim.Adam is a third-party dependency of the reference (environment.yml pins pytorch 1.12.1; this image has
 * torch 2.11): its published algorithm (Kingma & Ba with bias correction, amsgrad off, weight_decay 0, betas
 * (0.9, 0.999)) is restated in oracle/optim_oracle.py and pinned against torch.optim.Adam itself in the tests.
 *
 * Conventions as in gsr.h: device pointers, fp32, contiguous; GSR_OK or a negative code; `stream` is a cudaStream_t
 * passed as void*; no global mutable state.
 */
#ifndef GSR_OPTIM_H_INCLUDED
#define GSR_OPTIM_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSR_ADAM_MAX_SEGMENTS 8

/* One parameter tensor [rows, cols] (row = Gaussian) with its gradient and Adam state.  Columns < split_col use lr_a,
 * the others lr_b: the SH tensor [P,48] carries the reference's f_dc (first 3 floats, feature_lr) and f_rest (45
 * floats, feature_lr / 20) groups in one row (gaussian_model.py:184-185).  split_col = cols for a single rate. */
typedef struct gsr_adam_segment {
    float* param;            /* [rows*cols] updated in place */
    const float* grad;       /* [rows*cols] */
    float* exp_avg;          /* [rows*cols] first moment, updated in place */
    float* exp_avg_sq;       /* [rows*cols] second moment, updated in place */
    size_t rows;
    int cols;
    int split_col;
    float lr_a, lr_b;
} gsr_adam_segment;

/*
 * One Adam step over all segments in ONE launch (torch.optim.Adam.step for the six groups: ~24 foreach kernels).
 *   m = m + (1-beta1) (g - m);  v = beta2 v + (1-beta2) g g;
 *   p = p - lr/(1-beta1^step) * m / (sqrt(v)/sqrt(1-beta2^step) + eps)
 * `step` is the 1-based step count AFTER the increment (torch increments first).  beta1/beta2/eps are doubles, as
 * in torch: 1-beta and the bias corrections are formed in double and rounded to fp32 once.
 * visible (u8[rows], may be NULL): when given, only rows with visible[row] != 0 are updated -- parameters AND
 * moments of the other rows are left untouched (the "sparse Adam" variant train.py:179-181 asks for through
 * optimizer.step(visible, N); NULL = the reference's dense torch.optim.Adam).  All segments must have the same row
 * count when a mask is given.
 * zero_grad != 0 additionally clears the gradients that were consumed (optimizer.zero_grad(), train.py:186), so the
 * next step's first backward can run in accumulate mode; the gradient pointers must then be writable.
 */
int gsr_adam_step(const gsr_adam_segment* segments, int n_segments, int step, double beta1, double beta2, double eps,
                  const unsigned char* visible, int zero_grad, void* stream);

/*
 * Densification statistics of one rendered view (train.py:166-167, gaussian_model.py:471-473), one launch:
 *   for rows with radii > 0:  max_radii2D = max(max_radii2D, radii);
 *                             xyz_gradient_accum += || grad_means2D[row, :2] ||_2;   denom += 1
 * grad_means2D is the [P,3] viewspace gradient the rasterizer's backward returns (dL_dmeans2D).
 */
int gsr_densification_stats(int P, const float* grad_means2D, const int* radii, float* xyz_gradient_accum,
                            float* denom, float* max_radii2D, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GSR_OPTIM_H_INCLUDED */
