/* gsiron_host.c -- the fork's own host renderer, restated without Eigen / OpenCV / XRT.
 *
 * TEST / BASELINE INFRASTRUCTURE ONLY (bench.py's cpu_baseline.gsiron_host row and tests/): the product never links it.
 *
 * Follows the CPU branch (#ifndef __USE_NPU) of /root/reference/GS-IRON/npu-1/base.cpp:104-447 step by step:
 *   :209-222  full_proj * p, divide by (w + 1e-7), view-space xyz, screen = ((ndc + 1) * size - 1) * 0.5
 *   :236-239  skip when view z < 0.2
 *   :242-272  normalised quaternion -> R, M = R S, Sigma = M M^T, J (fx, fy = 1111.1), J_R = J * W[0:3,0:3]
 *   :276-292  cov2D = J_R Sigma J_R^T, +0.3 on the diagonal, inverse (antialiasing is `false`, :28, so the opacity
 *             factor is 1)
 *   :296-312  eigenvalue radius ceil(3 sqrt(max lambda)), getRelatedTiles (util.cpp:8-17), push into every tile of the
 *             rect; the grid is width/16 x height/16 (floor, :120) and tiles are indexed tx * grid_y + ty (:227-234)
 *   :315-349  degree-3 SH colour with the direction cam.pos - xyz (the rasterizer uses xyz - campos), +0.5, max(0, .)
 *   :358-376  per-tile sort by view z, ties by opacity  (std::sort; here ties beyond that keep Gaussian order)
 *   :380-433  per tile, per pixel, front to back at the pixel CENTRE (+0.5, :400-401): skip exponent > 0 and
 *             alpha < 1/255, blend, stop AFTER the Gaussian that takes T to <= 1e-4 (:413-418), clamp to [0,1]
 * Camera: load_camera (camera_loader.cpp:9-52) -- world_to_view is the INVERSE of the hard-coded matrix of
 * base.cpp:453-457, full_proj = getProjMat(100, 0.01, 0.69, 0.69) * world_to_view (util.cpp:20-41), 800x800.
 * The loader's activations (exp of the scales, sigmoid of the opacity, loader.cpp) are the caller's: inputs arrive
 * activated, like the rasterizer's.  Single-threaded by construction, forward only.  The image is returned RGB float
 * [H, W, 3] (base.cpp stores BGR for OpenCV, :429).
 *
 * Not parity-comparable with diff_gaussian_rasterization: half-pixel sample offset, flipped view
 * direction, no 1.3 tan_fov clamp, normalised quaternion, different early-out.  Checked by tests/test_gsiron_host_cpu.py
 * against an independent numpy evaluation of the same formulas on the fork's 1- and 2-Gaussian fixtures.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define GRID 16
static const float SH_C0 = 0.28209479177387814f, SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = { 1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f, -1.0925484305920792f, 0.5462742152960396f };
static const float SH_C3[7] = { -0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f, 0.3731763325901154f,
                                -0.4570457994644658f, 1.445305721320277f, -0.5900435899266435f };

typedef struct {
    float zview, opacity, sx, sy;      /* sort keys, screen centre */
    float ia, ib, ic;                  /* inverse 2D covariance: [ia ib; ib ic] */
    float r, g, b;
} Splat;

static const Splat* g_splats;
static int cmp_idx(const void* pa, const void* pb) {
    const int a = *(const int*)pa, b = *(const int*)pb;
    const Splat *A = g_splats + a, *B = g_splats + b;
    if (A->zview != B->zview) return A->zview < B->zview ? -1 : 1;
    if (A->opacity != B->opacity) return A->opacity < B->opacity ? -1 : 1;
    return a < b ? -1 : (a > b);
}

/* 4x4 inverse (general, row-major) */
static int inv4(const float* m, float* out) {
    double a[4][8];
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) { a[i][j] = m[4 * i + j]; a[i][4 + j] = (i == j); }
    for (int c = 0; c < 4; c++) {
        int piv = c;
        for (int r = c + 1; r < 4; r++) if (fabs(a[r][c]) > fabs(a[piv][c])) piv = r;
        if (fabs(a[piv][c]) < 1e-20) return -1;
        if (piv != c) for (int j = 0; j < 8; j++) { double t = a[c][j]; a[c][j] = a[piv][j]; a[piv][j] = t; }
        const double d = a[c][c];
        for (int j = 0; j < 8; j++) a[c][j] /= d;
        for (int r = 0; r < 4; r++) if (r != c) { const double f = a[r][c]; for (int j = 0; j < 8; j++) a[r][j] -= f * a[c][j]; }
    }
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) out[4 * i + j] = (float)a[i][4 + j];
    return 0;
}

/* Renders P Gaussians with the camera matrix `base_mat` (row-major 4x4, the matrix main() hard-codes; NULL = that one).
 * xyz[P,3], scale[P,3] (activated), rot[P,4] (w,x,y,z), opacity[P] (activated), f_dc[P,3], f_rest[P,45] (channel-major,
 * the PLY's order).  image: float [H,W,3] RGB.  Returns the number of (Gaussian, tile) instances, or -1. */
long gsiron_render(int P, const float* xyz, const float* scale, const float* rot, const float* opacity, const float* f_dc,
                   const float* f_rest, const float* base_mat, int width, int height, float fx, float fy, float fov,
                   float* image) {
    static const float default_mat[16] = {
        -0.9250140190124512f, -0.2748899757862091f, 0.2622683644294739f, -1.0572376251220703f,
        -0.37993317842483526f, 0.6692678928375244f, -0.6385383605957031f, 2.5740303993225098f,
        -0.0f, -0.6903012990951539f, -0.7235219478607177f, 2.9166102409362793f,
        0.f, 0.f, 0.f, 1.f };
    if (!base_mat) base_mat = default_mat;
    /* load_camera: world_to_view = inverse(base_mat); cam.pos = translation of inverse(world_to_view) = of base_mat */
    float W2V[16], PM[16] = { 0 }, FP[16];
    if (inv4(base_mat, W2V)) return -1;
    W2V[12] = W2V[13] = W2V[14] = 0.f; W2V[15] = 1.f;
    {   /* getProjMat(zfar = 100, znear = 0.01, fovX = fovY = fov) */
        const float zfar = 100.f, znear = 0.01f;
        const float top = tanf(fov / 2) * znear, right = tanf(fov / 2) * znear;
        PM[0] = 2.0 * znear / (right + right); PM[5] = 2.0 * znear / (top + top);
        PM[2] = 0.f; PM[6] = 0.f; PM[14] = 1.0f; PM[10] = zfar / (zfar - znear); PM[11] = -(zfar * znear) / (zfar - znear);
    }
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) {
        float s = 0; for (int k = 0; k < 4; k++) s += PM[4 * i + k] * W2V[4 * k + j];
        FP[4 * i + j] = s;
    }
    float W2Vinv[16];
    if (inv4(W2V, W2Vinv)) return -1;
    const float cpx = W2Vinv[3], cpy = W2Vinv[7], cpz = W2Vinv[11];

    const int gx = width / GRID, gy = height / GRID;
    const int ntiles = gx * gy;
    Splat* sp = (Splat*)malloc(sizeof(Splat) * (size_t)(P > 0 ? P : 1));
    int* rect = (int*)malloc(sizeof(int) * 4 * (size_t)(P > 0 ? P : 1));
    long* count = (long*)calloc((size_t)ntiles + 1, sizeof(long));
    if (!sp || !rect || !count) { free(sp); free(rect); free(count); return -1; }

    for (int i = 0; i < P; i++) {
        Splat* s = sp + i;
        rect[4 * i] = rect[4 * i + 1] = rect[4 * i + 2] = rect[4 * i + 3] = 0;
        const float x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
        float pv[4], vv[3];
        for (int r = 0; r < 4; r++) pv[r] = FP[4 * r] * x + FP[4 * r + 1] * y + FP[4 * r + 2] * z + FP[4 * r + 3];
        const float w = pv[3] + 0.0000001f;
        for (int r = 0; r < 3; r++) vv[r] = W2V[4 * r] * x + W2V[4 * r + 1] * y + W2V[4 * r + 2] * z + W2V[4 * r + 3];
        s->sx = (float)(((pv[0] / w + 1.0) * width - 1.0) * 0.5);
        s->sy = (float)(((pv[1] / w + 1.0) * height - 1.0) * 0.5);
        s->zview = vv[2];
        s->opacity = opacity[i];
        if (vv[2] < 0.2f) continue;
        /* quaternion (normalised) -> R ; M = R S ; Sigma = M M^T */
        float qw = rot[4 * i], qx = rot[4 * i + 1], qy = rot[4 * i + 2], qz = rot[4 * i + 3];
        const float qn = sqrtf(qw * qw + qx * qx + qy * qy + qz * qz);
        qw /= qn; qx /= qn; qy /= qn; qz /= qn;
        const float R[9] = { 1 - 2 * qy * qy - 2 * qz * qz, 2 * qx * qy - 2 * qz * qw, 2 * qx * qz + 2 * qy * qw,
                             2 * qx * qy + 2 * qz * qw, 1 - 2 * qx * qx - 2 * qz * qz, 2 * qy * qz - 2 * qx * qw,
                             2 * qx * qz - 2 * qy * qw, 2 * qy * qz + 2 * qx * qw, 1 - 2 * qx * qx - 2 * qy * qy };
        float M[9], Sg[9];
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) M[3 * r + c] = R[3 * r + c] * scale[3 * i + c];
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++)
            Sg[3 * r + c] = M[3 * r] * M[3 * c] + M[3 * r + 1] * M[3 * c + 1] + M[3 * r + 2] * M[3 * c + 2];
        /* J (third row zero) and J_R = J * W2V[0:3,0:3] */
        const float zv = vv[2];
        const float J[6] = { fx / zv, 0.f, -fx * vv[0] / (zv * zv), 0.f, fy / zv, -fy * vv[1] / (zv * zv) };
        float JR[6];
        for (int r = 0; r < 2; r++) for (int c = 0; c < 3; c++)
            JR[3 * r + c] = J[3 * r] * W2V[c] + J[3 * r + 1] * W2V[4 + c] + J[3 * r + 2] * W2V[8 + c];
        float T[6];   /* JR * Sigma */
        for (int r = 0; r < 2; r++) for (int c = 0; c < 3; c++)
            T[3 * r + c] = JR[3 * r] * Sg[c] + JR[3 * r + 1] * Sg[3 + c] + JR[3 * r + 2] * Sg[6 + c];
        float c00 = T[0] * JR[0] + T[1] * JR[1] + T[2] * JR[2];
        const float c10 = T[3] * JR[0] + T[4] * JR[1] + T[5] * JR[2];
        float c11 = T[3] * JR[3] + T[4] * JR[4] + T[5] * JR[5];
        c00 += 0.3f; c11 += 0.3f;
        const float det = c00 * c11 - c10 * c10;
        const float det_inv = 1.f / det;
        s->ia = c11 * det_inv; s->ib = -c10 * det_inv; s->ic = c00 * det_inv;
        const float b = 0.5f * (c00 + c11);
        const float l1 = b + sqrtf(fmaxf(0.1f, b * b - det)), l2 = b - sqrtf(fmaxf(0.1f, b * b - det));
        const int radius = (int)ceilf(3.f * sqrtf(fmaxf(l1, l2)));
        /* getRelatedTiles: C truncation, clamped to the grid */
        int x0 = (int)((s->sx - radius) / GRID), y0 = (int)((s->sy - radius) / GRID);
        int x1 = (int)((s->sx + radius + GRID - 1) / GRID), y1 = (int)((s->sy + radius + GRID - 1) / GRID);
        x0 = x0 < 0 ? 0 : (x0 > gx ? gx : x0); y0 = y0 < 0 ? 0 : (y0 > gy ? gy : y0);
        x1 = x1 < 0 ? 0 : (x1 > gx ? gx : x1); y1 = y1 < 0 ? 0 : (y1 > gy ? gy : y1);
        if ((x1 - x0) * (y1 - y0) == 0) continue;
        rect[4 * i] = x0; rect[4 * i + 1] = y0; rect[4 * i + 2] = x1; rect[4 * i + 3] = y1;
        for (int tx = x0; tx < x1; tx++) for (int ty = y0; ty < y1; ty++) count[tx * gy + ty + 1]++;
        /* colour: degree-3 SH, direction cam.pos - xyz */
        float dx = cpx - x, dy = cpy - y, dz = cpz - z;
        const float dn = sqrtf(dx * dx + dy * dy + dz * dz);
        dx /= dn; dy /= dn; dz /= dn;
        const float xx = dx * dx, yy = dy * dy, zz = dz * dz, xy = dx * dy, yz = dy * dz, xz = dx * dz;
        float col[3];
        for (int c = 0; c < 3; c++) {
            const float* fr = f_rest + 45 * (size_t)i + 15 * c;
            float v = 0.5f;
            v += SH_C0 * f_dc[3 * i + c];
            v += SH_C1 * (-1 * dy * fr[0] + dz * fr[1] - dx * fr[2]);
            v += SH_C2[0] * xy * fr[3] + SH_C2[1] * yz * fr[4] + SH_C2[2] * (2.0f * zz - xx - yy) * fr[5] +
                 SH_C2[3] * xz * fr[6] + SH_C2[4] * (xx - yy) * fr[7];
            v += SH_C3[0] * dy * (3.0f * xx - yy) * fr[8] + SH_C3[1] * xy * dz * fr[9] +
                 SH_C3[2] * dy * (4.0f * zz - xx - yy) * fr[10] + SH_C3[3] * dz * (2.0f * zz - 3.0f * xx - 3.0f * yy) * fr[11] +
                 SH_C3[4] * dx * (4.0f * zz - xx - yy) * fr[12] + SH_C3[5] * dz * (xx - yy) * fr[13] +
                 SH_C3[6] * dx * (xx - 3.0f * yy) * fr[14];
            col[c] = fmaxf(0.0f, v);
        }
        s->r = col[0]; s->g = col[1]; s->b = col[2];
    }
    /* tile lists in push order (Gaussian index), then the per-tile sort */
    for (int t = 0; t < ntiles; t++) count[t + 1] += count[t];
    const long total = count[ntiles];
    int* list = (int*)malloc(sizeof(int) * (size_t)(total > 0 ? total : 1));
    long* fill = (long*)malloc(sizeof(long) * (size_t)(ntiles > 0 ? ntiles : 1));
    if (!list || !fill) { free(sp); free(rect); free(count); free(list); free(fill); return -1; }
    memcpy(fill, count, sizeof(long) * (size_t)ntiles);
    for (int i = 0; i < P; i++)
        for (int tx = rect[4 * i]; tx < rect[4 * i + 2]; tx++)
            for (int ty = rect[4 * i + 1]; ty < rect[4 * i + 3]; ty++) list[fill[tx * gy + ty]++] = i;
    g_splats = sp;
    for (int t = 0; t < ntiles; t++)
        if (count[t + 1] - count[t] > 1) qsort(list + count[t], (size_t)(count[t + 1] - count[t]), sizeof(int), cmp_idx);

    memset(image, 0, sizeof(float) * 3 * (size_t)width * height);
    for (int tx = 0; tx < gx; tx++) for (int ty = 0; ty < gy; ty++) {
        const int t = tx * gy + ty;
        for (int i = 0; i < GRID; i++) for (int j = 0; j < GRID; j++) {
            const int px = tx * GRID + i, py = ty * GRID + j;
            if (px >= width || py >= height) continue;
            float cr = 0.f, cg = 0.f, cb = 0.f, T = 1.0f;
            for (long k = count[t]; k < count[t + 1]; k++) {
                const Splat* s = sp + list[k];
                const float d0 = px + 0.5f - s->sx, d1 = py + 0.5f - s->sy;
                const float e = -0.5f * (d0 * (s->ia * d0 + s->ib * d1) + d1 * (s->ib * d0 + s->ic * d1));
                if (e > 0.f) continue;
                const float alpha = fminf(0.99f, s->opacity * expf(e));
                if (alpha < 1.0f / 255.0f) continue;
                cr += T * alpha * s->r; cg += T * alpha * s->g; cb += T * alpha * s->b;
                T = T * (1.f - alpha);
                if (T <= 0.0001f) break;
            }
            float* o = image + 3 * ((size_t)py * width + px);
            o[0] = fmaxf(0.f, fminf(1.f, cr)); o[1] = fmaxf(0.f, fminf(1.f, cg)); o[2] = fmaxf(0.f, fminf(1.f, cb));
        }
    }
    free(sp); free(rect); free(count); free(list); free(fill);
    return total;
}
