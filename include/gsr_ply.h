/*
 * gsr_ply.h -- C ABI of the trained-model PLY format the callers feed the rasterizer with
 * (same library, libgsr_b200.so).
 *
 * Replaces, without the `plyfile` package,
 *   /root/reference/scene/gaussian_model.py:225-238   construct_list_of_attributes (62 fp32 properties)
 *   /root/reference/scene/gaussian_model.py:240-258   save_ply
 *   /root/reference/scene/gaussian_model.py:265-314   load_ply (properties looked up BY NAME: files written by
 *                                                     other tools order them differently, e.g. the fork's own
 *                                                     GS-IRON/base/single_gaussian.ply, generator gaussian.py:70-92)
 *   /root/reference/GS-IRON/npu-1/loader.cpp:10-193   the fork's C++ loader of the same records
 * and the activations the renderer applies to the stored values before they reach the rasterizer
 *   /root/reference/scene/gaussian_model.py:47-58,107-124  scaling = exp, opacity = sigmoid, rotation = normalize.
 *
 * The header is parsed on the host; the records (one AoS row of fp32 per Gaussian) are transposed into / out of the
 * rasterizer's tensors by one kernel each, on the device, from the raw file bytes.
 */
#ifndef GSR_PLY_H_INCLUDED
#define GSR_PLY_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Index of each of the 62 attributes in gsr_ply_layout.offset[], in the reference's save_ply order
 * (gaussian_model.py:225-238): x y z nx ny nz f_dc_0..2 f_rest_0..44 opacity scale_0..2 rot_0..3 */
#define GSR_PLY_NUM_ATTR 62
#define GSR_PLY_X 0
#define GSR_PLY_NX 3
#define GSR_PLY_FDC 6
#define GSR_PLY_FREST 9
#define GSR_PLY_OPACITY 54
#define GSR_PLY_SCALE 55
#define GSR_PLY_ROT 58

typedef struct gsr_ply_layout {
    uint64_t num_vertices;      /* `element vertex N` */
    uint64_t header_bytes;      /* offset of the first record in the file */
    uint32_t record_bytes;      /* stride of one record (all properties, whatever their type) */
    int32_t n_rest;             /* number of f_rest_* properties found (0, 9, 24 or 45 = SH degree 0..3) */
    int32_t offset[GSR_PLY_NUM_ATTR]; /* BYTE offset of each attribute inside a record, -1 if the file lacks it */
} gsr_ply_layout;

/*
 * Parses the ASCII header of a binary_little_endian PLY held in host memory (at least the header must be inside
 * [data, data+len)).  Attributes must be `property float`; other properties are skipped by their size.  Errors:
 * not a PLY, ascii / big-endian body, no vertex element, list properties inside vertex, a required attribute
 * (x y z opacity scale_0..2 rot_0..3 f_dc_0..2) missing or not float, f_rest count not in {0, 9, 24, 45}
 * (the reference asserts the count matches its SH degree, gaussian_model.py:290).  HOST function, no GPU needed.
 */
int gsr_ply_parse_header(const char* data, size_t len, gsr_ply_layout* out);

/* Writes the reference's header (save_ply order, 62 properties for sh_degree 3; f_rest count follows sh_degree)
 * into buf; returns the header length, or the required length if buf_len is too small (nothing written).  HOST. */
size_t gsr_ply_write_header(uint64_t num_vertices, int sh_degree, char* buf, size_t buf_len);

/*
 * Device: records (raw file payload, DEVICE pointer, any alignment >= 4 bytes) -> tensors.
 *   xyz[P,3], shs[P,M,3] with M = (sh_degree+1)^2 coefficient-major / RGB innermost -- features_dc in coefficient 0,
 *   f_rest_{c*(M-1)+k} in shs[1+k][c]  (the reference's reshape (P,3,M-1) + transpose(1,2), gaussian_model.py:293,309)
 *   -- opacity[P], scales[P,3], rotations[P,4].
 * activate != 0 applies the renderer's activations on the way (exp / sigmoid / quaternion normalisation);
 * activate == 0 stores the raw (optimiser-space) values, as load_ply does.  `sh_degree` is the degree of the OUTPUT
 * tensor; coefficients the file lacks are zero-filled.
 */
int gsr_ply_unpack(const gsr_ply_layout* layout, const char* records, int sh_degree, int activate, float* xyz,
                   float* shs, float* opacity, float* scales, float* rotations, void* stream);

/* Device: tensors (raw, optimiser-space values) -> records in the reference's save_ply order, normals zero.
 * `records` must hold P * (17 + 3*((sh_degree+1)^2 - 1)) floats. */
int gsr_ply_pack(uint64_t P, int sh_degree, const float* xyz, const float* shs, const float* opacity,
                 const float* scales, const float* rotations, char* records, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GSR_PLY_H_INCLUDED */
