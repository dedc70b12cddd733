/*
 * gsr_dist.h -- C ABI of the multi-GPU exchange step of view-parallel training (same library,
 * libgsr_b200.so).  One process per GPU; every GPU holds a replica of the Gaussians and, after rendering its share of
 * the step's views, a flat bucket of parameter gradients (59 fp32 per Gaussian).  The reference has no multi-GPU path
 * (utils/general_utils.py:133 pins cuda:0); what has to stay equivalent is the single-GPU sequence
 *   /root/reference/train.py:163-186                  statistics -> densify -> optimizer.step() -> zero_grad()
 *   /root/reference/scene/gaussian_model.py:178-201   Adam(lr=0, eps=1e-15) over the six groups
 *   /root/reference/scene/gaussian_model.py:471-473   xyz_gradient_accum / denom (SUM over views), train.py:166 max_radii2D (MAX)
 * applied to the SUM of the gradients of all views of the step.
 *
 * gsr_exchange_step is ONE kernel per rank that does, over NVLink / NVSwitch peer memory and tile by tile,
 *   reduce-scatter  : rank r sums shard r of the gradient bucket over all ranks
 *                     (multimem.ld_reduce through the switch when a multicast mapping is given, else W peer loads),
 *   sharded Adam    : the update of shard r with rank-local moments (1/W of the optimizer state and work per GPU),
 *   all-gather      : the updated parameters of shard r stored into every replica (multimem.st or W peer stores),
 *   statistics      : SUM / MAX all-reduce of the densification-statistic segments in the same pass,
 * with the cross-GPU barriers inside the kernel (one flag exchange per CTA on entry and on exit).  It replaces
 * ncclAllReduce(bucket) + gsr_adam_step(all rows) + 3 small all-reduces.  With GSR_EXCHANGE_SUM segments only it is a
 * plain in-place all-reduce of the bucket.
 *
 * The library never opens a communicator: the caller (gsr_b200/dist.py, through torch.distributed's symmetric memory)
 * maps the peers' buffers and hands over plain device pointers.  world == 1 degenerates to the local Adam step.
 */
#ifndef GSR_DIST_H_INCLUDED
#define GSR_DIST_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSR_MAX_RANKS 16
#define GSR_EXCHANGE_MAX_SEGMENTS 12

enum { GSR_EXCHANGE_ADAM = 0,   /* reduce gradient, Adam on the owner's shard, broadcast the parameters */
       GSR_EXCHANGE_SUM = 1,    /* all-reduce (sum) in place */
       GSR_EXCHANGE_MAX = 2 };  /* all-reduce (max) in place; values must be >= 0 (compared as bit patterns) */

/* A range [begin, end) of floats of the flat symmetric buffers; begin/end multiples of 4.  ADAM segments are
 * [rows, cols] tensors (row = Gaussian): columns < split_col step with lr_a, the others with lr_b (the SH tensor
 * carries f_dc and f_rest in one row, gaussian_model.py:184-185). */
typedef struct gsr_exchange_segment {
    size_t begin, end;
    int kind;                 /* GSR_EXCHANGE_* */
    int cols, split_col;
    float lr_a, lr_b;
} gsr_exchange_segment;

typedef struct gsr_exchange_args {
    int rank, world;
    size_t n;                          /* floats in each flat buffer (multiple of 4) */
    float* grad_peer[GSR_MAX_RANKS];   /* [world] every rank's gradient bucket as mapped into THIS process */
    float* param_peer[GSR_MAX_RANKS];  /* [world] every rank's parameter buffer (same layout as the bucket) */
    float* grad_mc;                    /* multicast mapping of the gradient buckets, or NULL (peer loads / stores) */
    float* param_mc;                   /* multicast mapping of the parameter buffers, or NULL */
    uint32_t* signal_peer[GSR_MAX_RANKS]; /* [world] every rank's signal words: >= gsr_exchange_signal_words() u32, zeroed once */
    float* exp_avg;                    /* rank-local Adam moments of THIS rank's shard: gsr_exchange_shard_floats() each */
    float* exp_avg_sq;
    const gsr_exchange_segment* segments;   /* host array */
    int n_segments;
    int step;                          /* Adam step count, 1-based, after the increment */
    double beta1, beta2, eps;
} gsr_exchange_args;

/* number of u32 signal words every rank must provide (zero-initialised, symmetric) */
size_t gsr_exchange_signal_words(void);
/* the shard of rank `rank`: floats [*begin, *end) of the flat buffers (multiples of 4) */
void gsr_exchange_shard(size_t n, int rank, int world, size_t* begin, size_t* end);

/* Launches the exchange kernel of this rank on `stream`.  Every rank of the group must call it once per step with the
 * same n / segments / step.  When the kernel has finished on a rank, that rank's parameters (ADAM segments) and
 * SUM / MAX segments hold the step's result and no peer still writes to them. *
 * Co-residency: CTA b of a rank waits on flags from CTA b of every peer, so the whole grid (148 SMs x 2 CTAs) has to be
 * resident on every rank while the step runs.  With world > 1 the call refuses (GSR_ERR_INVALID) a device that cannot
 * hold that many CTAs of the kernel.  It cannot see what else is running: do not keep another kernel that occupies SMs
 * for long, or another collective (NCCL included), in flight on the device while the exchange runs -- order them on the
 * same stream or with events.  A flag that does not arrive within 300 s (a dead or missing rank) makes the kernel trap,
 * so the failure surfaces as a CUDA error at the next synchronisation instead of a hang. */
int gsr_exchange_step(const gsr_exchange_args* args, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GSR_DIST_H_INCLUDED */
