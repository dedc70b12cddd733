/*
 * ssim_oracle.c -- CPU oracle for the loss stage next to the rasterizer.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path may import, link or call this file; only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg use it, as the checker / the
 * timed CPU baseline.
 *
 * It restates, in plain C, the algorithm of the reference's fused SSIM extension
 * (FS = /root/reference/submodules/fused-ssim) and of the loss around it:
 *   FS/ssim.cu:10-20      the 11-tap Gaussian window (sigma 1.5, fp32 literals)   -> G[]
 *   FS/ssim.cu:38-44      zero padding outside the image                           -> pix()
 *   FS/ssim.cu:104-164    horizontal 11-tap pass, taps accumulated left to right   -> conv_x_row()
 *   FS/ssim.cu:166-186    vertical 11-tap pass, taps accumulated top to bottom     -> conv_y_at()
 *   FS/ssim.cu:188-292    mu1, sigma1^2, mu2, sigma2^2, sigma12, the SSIM map and
 *                         its three partial derivatives                            -> sso_ssim_forward()
 *   FS/ssim.cu:294-371    dL/dimg1 = conv(dL*dm_dmu1) + 2*img1*conv(dL*dm_dsigma1_sq)
 *                                    + img2*conv(dL*dm_dsigma12)                   -> sso_ssim_backward()
 *   FS/fused_ssim/__init__.py:8-41  "valid" crop and the mean                      -> python side (ssim_oracle.py)
 *   /root/reference/utils/loss_utils.py:40-41, train.py:118-126
 *                         loss = (1-lambda)*mean|img-gt| + lambda*(1-mean(ssim))   -> sso_l1_ssim_loss()
 *
 * Floating-point contract: fp32 throughout.  The reference is CUDA built with -fmad=true; every
 * `val += G_k * p` of its two passes is one FFMA in tap order k = 0..10, and `x - mu*mu` is one FFMA.
 * This file spells those as fmaf() and is compiled with -ffp-contract=off.  The squared / product
 * inputs (do_sq, multiply_shared_mem, FS/ssim.cu:66-86) are rounded to fp32 before the window is
 * applied, as in the reference.  The closing rational expressions follow the FFMA placement of the
 * reference's own sm_100a SASS (see sso_ssim_forward) with IEEE divisions, so on a CPU with a correctly
 * rounded fmaf the four maps and dL_dimg1 are BIT-IDENTICAL to the reference CUDA kernel's
 * (tests/test_ssim_golden.py asserts equality).  Sums over pixels (the mean) are accumulated in double.
 *
 * Parity pin: (1) the reference's own test recipe (FS/tests/test.py:14-55: torch conv2d SSIM, checked
 * with torch.isclose) is run against this file by tests/test_ssim_oracle_cpu.py; (2) golden vectors
 * produced by the UNMODIFIED FS/ssim.cu built for sm_100a (oracle/_ref/fused_ssim_cuda) on a B200 are
 * committed under tests/golden/ssim_*.npz together with tests/golden/make_golden_ssim.py.
 */
#include <math.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static const float G[11] = {
    0.001028380123898387f, 0.0075987582094967365f, 0.036000773310661316f, 0.10936068743467331f,
    0.21300552785396576f,  0.26601171493530273f,   0.21300552785396576f,  0.10936068743467331f,
    0.036000773310661316f, 0.0075987582094967365f, 0.001028380123898387f};

int sso_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* FS/ssim.cu:38-44 */
static inline float pix(const float* plane, int H, int W, int y, int x) {
    return (x >= W || y >= H || x < 0 || y < 0) ? 0.0f : plane[(size_t)y * W + x];
}

/* Horizontal pass of one (zero-padded) source row into out[0..W): FS/ssim.cu:104-164.
 * mode 0: a     mode 1: a*a     mode 2: a*b (both rounded to fp32 first). */
static void conv_x_row(const float* a, const float* b, int H, int W, int y, int mode, float* out) {
    for (int x = 0; x < W; ++x) {
        float val = 0.0f;
        for (int k = 0; k < 11; ++k) {
            float p = pix(a, H, W, y, x + k - 5);
            if (mode == 1) p = p * p;
            else if (mode == 2) p = p * pix(b, H, W, y, x + k - 5);
            val = fmaf(G[k], p, val);
        }
        out[x] = val;
    }
}

/* Full separable window over one plane: out[y][x] = sum_ky G[ky] * rowconv[y+ky-5][x], rows outside the
 * image contribute conv_x of zeros = 0 (FS/ssim.cu:166-186 over the flushed scratch). */
static void window_plane(const float* a, const float* b, int H, int W, int mode, float* tmp, float* out) {
    for (int y = 0; y < H; ++y) conv_x_row(a, b, H, W, y, mode, tmp + (size_t)y * W);
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            float val = 0.0f;
            for (int k = 0; k < 11; ++k) {
                int yy = y + k - 5;
                float p = (yy < 0 || yy >= H) ? 0.0f : tmp[(size_t)yy * W + x];
                val = fmaf(G[k], p, val);
            }
            out[(size_t)y * W + x] = val;
        }
}

/*
 * fusedssim (FS/ssim.cu:188-292, host wrapper :373-409).  img1/img2/ssim_map/dm_*: [planes][H][W] with
 * planes = B*CH (every plane is independent).  dm_* may be NULL (train = false).
 */
void sso_ssim_forward(int planes, int H, int W, float C1, float C2, const float* img1, const float* img2,
                      float* ssim_map, float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12) {
    size_t n = (size_t)H * W;
#pragma omp parallel for schedule(dynamic, 1)
    for (int pl = 0; pl < planes; ++pl) {
        const float* a = img1 + pl * n;
        const float* b = img2 + pl * n;
        float* tmp = (float*)malloc(n * sizeof(float));
        float* w[5];
        for (int i = 0; i < 5; ++i) w[i] = (float*)malloc(n * sizeof(float));
        window_plane(a, NULL, H, W, 0, tmp, w[0]); /* conv(img1)      */
        window_plane(a, NULL, H, W, 1, tmp, w[1]); /* conv(img1^2)    */
        window_plane(b, NULL, H, W, 0, tmp, w[2]); /* conv(img2)      */
        window_plane(b, NULL, H, W, 1, tmp, w[3]); /* conv(img2^2)    */
        window_plane(a, b, H, W, 2, tmp, w[4]);    /* conv(img1*img2) */
        for (size_t i = 0; i < n; ++i) {
            /* FFMA placement read from the reference's own sm_100a SASS (nvcc 12.9.86 on FS/ssim.cu,
             * cuobjdump of oracle/_ref/fused_ssim_cuda_ref.so): mu1*mu2 and mu2*mu2 are rounded products that are
             * reused, so sigma12 and sigma2_sq are FMUL + FADD; sigma1_sq and mu1_sq + mu2_sq are fused. */
            float mu1 = w[0][i], mu2 = w[2][i];
            float mu1_mu2 = mu1 * mu2;
            float mu2_sq = mu2 * mu2;
            float sigma12 = w[4][i] - mu1_mu2;
            float sigma1_sq = fmaf(-mu1, mu1, w[1][i]);
            float sigma2_sq = w[3][i] - mu2_sq;
            float C = fmaf(mu1_mu2, 2.0f, C1);
            float D = fmaf(sigma12, 2.0f, C2);
            float A = fmaf(mu1, mu1, mu2_sq) + C1;
            float B = (sigma1_sq + sigma2_sq) + C2;
            float AB = A * B;
            ssim_map[pl * n + i] = (C * D) / AB;
            if (dm_dmu1) {
                float two_mu2 = mu2 + mu2, two_mu1 = mu1 + mu1;
                float t1 = (D * two_mu2) / AB;
                float t2 = (C * two_mu2) / AB;
                float num = D * (C * two_mu1);
                float t3 = num / (B * (A * A));
                float ABB = B * AB;
                float t4 = num / ABB;
                dm_dmu1[pl * n + i] = ((t1 - t2) - t3) + t4;
                dm_dsigma1_sq[pl * n + i] = (C * -D) / ABB;
                dm_dsigma12[pl * n + i] = (C + C) / AB;
            }
        }
        for (int i = 0; i < 5; ++i) free(w[i]);
        free(tmp);
    }
}

/* fusedssim_backward (FS/ssim.cu:294-371, host wrapper :411-444). */
void sso_ssim_backward(int planes, int H, int W, const float* img1, const float* img2, const float* dL_dmap,
                       const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12, float* dL_dimg1) {
    size_t n = (size_t)H * W;
#pragma omp parallel for schedule(dynamic, 1)
    for (int pl = 0; pl < planes; ++pl) {
        const float* dl = dL_dmap + pl * n;
        float* tmp = (float*)malloc(n * sizeof(float));
        float* w[3];
        for (int i = 0; i < 3; ++i) w[i] = (float*)malloc(n * sizeof(float));
        window_plane(dm_dmu1 + pl * n, dl, H, W, 2, tmp, w[0]);
        window_plane(dm_dsigma1_sq + pl * n, dl, H, W, 2, tmp, w[1]);
        window_plane(dm_dsigma12 + pl * n, dl, H, W, 2, tmp, w[2]);
        for (size_t i = 0; i < n; ++i) {
            /* reference SASS: FADD(0, w0) ; FFMA(w1, p1+p1, .) ; FFMA(w2, p2, .) */
            float p1 = img1[pl * n + i];
            float g = fmaf(w[1][i], p1 + p1, 0.0f + w[0][i]);
            dL_dimg1[pl * n + i] = fmaf(w[2][i], img2[pl * n + i], g);
        }
        for (int i = 0; i < 3; ++i) free(w[i]);
        free(tmp);
    }
}

/*
 * The training loss around the rasterizer (train.py:118-126, utils/loss_utils.py:40-41,
 * FS/fused_ssim/__init__.py:33-41 with padding "same"):
 *     loss = (1 - lambda) * mean|img - gt| + lambda * (1 - mean(ssim_map(img, gt)))
 * and its gradient w.r.t. img for an upstream d(loss) = upstream:
 *     dL/dimg = upstream * [ (1-lambda)/n * sign(img - gt)  -  lambda/n * d(sum ssim)/dimg ]
 * out[0] = loss, out[1] = mean L1, out[2] = mean ssim (doubles).  dL_dimg may be NULL.
 */
void sso_l1_ssim_loss(int planes, int H, int W, float C1, float C2, float lambda, float upstream, const float* img,
                      const float* gt, double* out, float* dL_dimg) {
    size_t n = (size_t)planes * H * W;
    float* map = (float*)malloc(n * sizeof(float));
    float *d0 = NULL, *d1 = NULL, *d2 = NULL, *dl = NULL;
    if (dL_dimg) {
        d0 = (float*)malloc(n * sizeof(float));
        d1 = (float*)malloc(n * sizeof(float));
        d2 = (float*)malloc(n * sizeof(float));
        dl = (float*)malloc(n * sizeof(float));
    }
    sso_ssim_forward(planes, H, W, C1, C2, img, gt, map, d0, d1, d2);
    double s_l1 = 0.0, s_ssim = 0.0;
    for (size_t i = 0; i < n; ++i) {
        s_l1 += fabs((double)(img[i] - gt[i]));
        s_ssim += (double)map[i];
    }
    out[1] = s_l1 / (double)n;
    out[2] = s_ssim / (double)n;
    out[0] = (1.0 - (double)lambda) * out[1] + (double)lambda * (1.0 - out[2]);
    if (dL_dimg) {
        float c_ssim = -(upstream * lambda) / (float)n;
        float c_l1 = (upstream * (1.0f - lambda)) / (float)n;
        for (size_t i = 0; i < n; ++i) dl[i] = c_ssim;
        sso_ssim_backward(planes, H, W, img, gt, dl, d0, d1, d2, dL_dimg);
        for (size_t i = 0; i < n; ++i) {
            float d = img[i] - gt[i];
            float sgn = (d > 0.0f) ? 1.0f : ((d < 0.0f) ? -1.0f : 0.0f);
            dL_dimg[i] = fmaf(c_l1, sgn, dL_dimg[i]);
        }
        free(d0); free(d1); free(d2); free(dl);
    }
    free(map);
}
