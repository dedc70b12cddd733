// ply_io.cu -- the trained-model PLY format (include/gsr_ply.h): host header parser / writer and the two device
// kernels that transpose the AoS records into / out of the rasterizer's tensors.
//
// Replaces GaussianModel.load_ply / save_ply (/root/reference/scene/gaussian_model.py:240-314; on the host, through
// the `plyfile` package and ~70 numpy column copies) and the fork's C++ loader
// (/root/reference/GS-IRON/npu-1/loader.cpp:10-193).  The raw file payload goes to the GPU as it lies on disk; one
// kernel splits the records (attributes located BY NAME through the parsed header) into xyz / SH / opacity / scale /
// rotation, with the SH rows re-ordered from the file's channel-major f_rest to the rasterizer's coefficient-major
// [P,16,3], and optionally applies the renderer's activations (gaussian_model.py:47-58).
// Bound: HBM -- 4*R bytes in (R = floats per record, 62) + 4*59 bytes out per Gaussian; the CTA stages 128 records in
// shared memory with coalesced word loads and writes every output tensor with coalesced stores.
#include "../../include/gsr.h"
#include "../../include/gsr_ply.h"
#include "gsr_common.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

namespace {

constexpr int PLY_RECS = 128;            // records per CTA
constexpr int PLY_MAX_WORDS = 96;        // largest record (in 4-byte words) the staged kernel takes: 48 KB / 128 records

struct PlyDev {
    int32_t off[GSR_PLY_NUM_ATTR];       // byte offsets, -1 = absent
    uint32_t record_bytes;
    int32_t n_rest;
};

__device__ __forceinline__ float load_unaligned(const unsigned char* p) {
    uint32_t w = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
    return __uint_as_float(w);
}

// value of attribute `a` of local record r (staged words or raw global bytes); absent attributes read as 0
template <bool STAGED>
__device__ __forceinline__ float attr(const PlyDev& L, const uint32_t* sw, const unsigned char* rec_base, int r, int a) {
    const int o = L.off[a];
    if (o < 0) return 0.0f;
    if (STAGED) return __uint_as_float(sw[r * (L.record_bytes >> 2) + (o >> 2)]);
    return load_unaligned(rec_base + (size_t)r * L.record_bytes + o);
}

// gaussian_model.py:47-58: scaling_activation = exp, opacity_activation = sigmoid, rotation_activation = normalize
__device__ __forceinline__ float sigmoidf(float x) { return 1.0f / (1.0f + expf(-x)); }

template <bool STAGED>
__global__ void __launch_bounds__(PLY_RECS)
ply_unpack_kernel(const __grid_constant__ PlyDev L, const unsigned char* __restrict__ records, uint64_t P, int M, int activate,
                  float* __restrict__ xyz, float* __restrict__ shs, float* __restrict__ opacity, float* __restrict__ scales,
                  float* __restrict__ rotations) {
    extern __shared__ uint32_t sw[];
    const uint64_t r0 = (uint64_t)blockIdx.x * PLY_RECS;
    const int nrec = (int)min((uint64_t)PLY_RECS, P - r0);
    const unsigned char* base = records + r0 * L.record_bytes;
    if (STAGED) {
        const int words = nrec * (int)(L.record_bytes >> 2);
        const uint32_t* src = reinterpret_cast<const uint32_t*>(base);
        for (int i = threadIdx.x; i < words; i += PLY_RECS) sw[i] = __ldg(src + i);
        __syncthreads();
    }
    // xyz, scales: 3 floats per record, written as one contiguous run per CTA
    for (int e = threadIdx.x; e < nrec * 3; e += PLY_RECS) {
        const int r = e / 3, c = e - 3 * r;
        xyz[r0 * 3 + e] = attr<STAGED>(L, sw, base, r, GSR_PLY_X + c);
        const float s = attr<STAGED>(L, sw, base, r, GSR_PLY_SCALE + c);
        scales[r0 * 3 + e] = activate ? expf(s) : s;
    }
    for (int r = threadIdx.x; r < nrec; r += PLY_RECS) {
        const float o = attr<STAGED>(L, sw, base, r, GSR_PLY_OPACITY);
        opacity[r0 + r] = activate ? sigmoidf(o) : o;
        float q[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) q[c] = attr<STAGED>(L, sw, base, r, GSR_PLY_ROT + c);
        if (activate) {   // torch.nn.functional.normalize: x / max(||x||_2, 1e-12)
            const float n = fmaxf(sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]), 1e-12f);
#pragma unroll
            for (int c = 0; c < 4; ++c) q[c] = q[c] / n;
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) rotations[(r0 + r) * 4 + c] = q[c];
    }
    // SH: file is channel-major (f_rest_{c*(Mf-1)+k}), the tensor is coefficient-major [P,M,3]
    const int row = 3 * M;
    const int Mf1 = L.n_rest / 3;                       // non-DC coefficients per channel in the FILE
    for (int e = threadIdx.x; e < nrec * row; e += PLY_RECS) {
        const int r = e / row, j = e - r * row;
        const int k = j / 3, c = j - 3 * k;
        float v = 0.0f;
        if (k == 0) v = attr<STAGED>(L, sw, base, r, GSR_PLY_FDC + c);
        else if (k - 1 < Mf1) v = attr<STAGED>(L, sw, base, r, GSR_PLY_FREST + c * Mf1 + (k - 1));
        shs[r0 * row + e] = v;
    }
}

// tensors -> records in save_ply order: x y z nx ny nz f_dc(3) f_rest(3*(M-1), channel-major) opacity scale(3) rot(4)
__global__ void __launch_bounds__(256)
ply_pack_kernel(uint64_t P, int M, const float* __restrict__ xyz, const float* __restrict__ shs,
                const float* __restrict__ opacity, const float* __restrict__ scales, const float* __restrict__ rotations,
                float* __restrict__ out) {
    const int R = 17 + 3 * (M - 1);
    const int nrest = 3 * (M - 1);
    const uint64_t total = P * (uint64_t)R;
    for (uint64_t e = (uint64_t)blockIdx.x * 256 + threadIdx.x; e < total; e += (uint64_t)gridDim.x * 256) {
        const uint64_t r = e / R;
        const int a = (int)(e - r * R);
        float v;
        if (a < 3) v = xyz[r * 3 + a];
        else if (a < 6) v = 0.0f;                                   // normals = zeros_like(xyz), gaussian_model.py:244
        else if (a < 9) v = shs[r * 3 * M + (a - 6)];               // features_dc.transpose(1,2).flatten
        else if (a < 9 + nrest) {
            const int i = a - 9, c = i / (M - 1), k = i - c * (M - 1);
            v = shs[r * 3 * M + (1 + k) * 3 + c];                   // features_rest.transpose(1,2).flatten
        } else if (a == 9 + nrest) v = opacity[r];
        else if (a < 13 + nrest) v = scales[r * 3 + (a - 10 - nrest)];
        else v = rotations[r * 4 + (a - 13 - nrest)];
        out[e] = v;
    }
}

int type_size(const std::string& t) {
    if (t == "char" || t == "uchar" || t == "int8" || t == "uint8") return 1;
    if (t == "short" || t == "ushort" || t == "int16" || t == "uint16") return 2;
    if (t == "int" || t == "uint" || t == "float" || t == "int32" || t == "uint32" || t == "float32") return 4;
    if (t == "double" || t == "float64") return 8;
    return 0;
}

// index of a property name in the 62-attribute table, -1 if it is not one of them
int attr_index(const std::string& n) {
    static const char* fixed[] = {"x", "y", "z", "nx", "ny", "nz"};
    for (int i = 0; i < 6; ++i) if (n == fixed[i]) return i;
    auto num = [&](const char* prefix, int lo, int count) -> int {
        const size_t pl = strlen(prefix);
        if (n.compare(0, pl, prefix) != 0 || n.size() == pl) return -1;
        char* end = nullptr;
        const long v = strtol(n.c_str() + pl, &end, 10);
        if (*end != 0 || v < 0 || v >= count) return -1;
        return lo + (int)v;
    };
    int i;
    if ((i = num("f_dc_", GSR_PLY_FDC, 3)) >= 0) return i;
    if ((i = num("f_rest_", GSR_PLY_FREST, 45)) >= 0) return i;
    if (n == "opacity") return GSR_PLY_OPACITY;
    if ((i = num("scale_", GSR_PLY_SCALE, 3)) >= 0) return i;
    if ((i = num("rot_", GSR_PLY_ROT, 4)) >= 0) return i;
    return -1;
}

int ply_fail(const char* msg) { return gsr_set_error(GSR_ERR_INVALID, msg); }

}  // namespace

extern "C" int gsr_ply_parse_header(const char* data, size_t len, gsr_ply_layout* out) {
    if (!data || !out) return ply_fail("ply: NULL argument");
    memset(out, 0, sizeof(*out));
    for (int i = 0; i < GSR_PLY_NUM_ATTR; ++i) out->offset[i] = -1;
    size_t pos = 0;
    bool first = true, binary_le = false, in_vertex = false, seen_vertex = false, done = false;
    uint32_t stride = 0;
    while (pos < len && !done) {
        size_t eol = pos;
        while (eol < len && data[eol] != '\n') ++eol;
        if (eol == len) return ply_fail("ply: header is not terminated by end_header inside the given bytes");
        std::string line(data + pos, eol - pos);
        pos = eol + 1;
        while (!line.empty() && (line.back() == '\r' || line.back() == ' ')) line.pop_back();
        if (first) {
            if (line != "ply") return ply_fail("ply: missing 'ply' magic");
            first = false;
            continue;
        }
        // split into at most 4 tokens
        std::string tok[4];
        int nt = 0;
        size_t i = 0;
        while (i < line.size() && nt < 4) {
            while (i < line.size() && line[i] == ' ') ++i;
            size_t j = i;
            while (j < line.size() && line[j] != ' ') ++j;
            if (j > i) tok[nt++] = line.substr(i, j - i);
            i = j;
        }
        if (nt == 0 || tok[0] == "comment" || tok[0] == "obj_info") continue;
        if (tok[0] == "format") {
            if (nt < 2 || tok[1] != "binary_little_endian") return ply_fail("ply: only format binary_little_endian is supported");
            binary_le = true;
        } else if (tok[0] == "element") {
            if (nt < 3) return ply_fail("ply: malformed element line");
            if (tok[1] == "vertex") {
                if (seen_vertex) return ply_fail("ply: two vertex elements");
                seen_vertex = in_vertex = true;
                out->num_vertices = strtoull(tok[2].c_str(), nullptr, 10);
            } else {
                if (!seen_vertex && strtoull(tok[2].c_str(), nullptr, 10) != 0)
                    return ply_fail("ply: a non-empty element precedes the vertex element");
                in_vertex = false;
            }
        } else if (tok[0] == "property") {
            if (!in_vertex) continue;
            if (nt < 3) return ply_fail("ply: malformed property line");
            if (tok[1] == "list") return ply_fail("ply: list property inside the vertex element");
            const int sz = type_size(tok[1]);
            if (sz == 0) return ply_fail("ply: unknown property type");
            const int a = attr_index(tok[2]);
            if (a >= 0) {
                if (sz != 4 || !(tok[1] == "float" || tok[1] == "float32")) return ply_fail("ply: Gaussian attributes must be 'property float'");
                if (out->offset[a] >= 0) return ply_fail("ply: duplicate attribute");
                out->offset[a] = (int32_t)stride;
            }
            stride += (uint32_t)sz;
        } else if (tok[0] == "end_header") {
            done = true;
        } else {
            return ply_fail("ply: unknown header keyword");
        }
    }
    if (!done) return ply_fail("ply: header is not terminated by end_header inside the given bytes");
    if (!binary_le) return ply_fail("ply: only format binary_little_endian is supported");
    if (!seen_vertex) return ply_fail("ply: no vertex element");
    static const int required[] = {0, 1, 2, GSR_PLY_FDC, GSR_PLY_FDC + 1, GSR_PLY_FDC + 2, GSR_PLY_OPACITY, GSR_PLY_SCALE,
                                   GSR_PLY_SCALE + 1, GSR_PLY_SCALE + 2, GSR_PLY_ROT, GSR_PLY_ROT + 1, GSR_PLY_ROT + 2, GSR_PLY_ROT + 3};
    for (int a : required)
        if (out->offset[a] < 0) return ply_fail("ply: a required attribute (x y z f_dc_* opacity scale_* rot_*) is missing");
    int n_rest = 0;
    for (int i = 0; i < 45; ++i) {
        if (out->offset[GSR_PLY_FREST + i] >= 0) {
            if (i != n_rest) return ply_fail("ply: f_rest_* attributes are not contiguous from 0");
            ++n_rest;
        }
    }
    // gaussian_model.py:290: assert len(extra_f_names) == 3*(max_sh_degree+1)**2 - 3
    if (n_rest != 0 && n_rest != 9 && n_rest != 24 && n_rest != 45) return ply_fail("ply: f_rest count is not 3*(D+1)^2-3 for a degree D in 0..3");
    out->n_rest = n_rest;
    out->record_bytes = stride;
    out->header_bytes = pos;
    return GSR_OK;
}

extern "C" size_t gsr_ply_write_header(uint64_t num_vertices, int sh_degree, char* buf, size_t buf_len) {
    if (sh_degree < 0 || sh_degree > 3) return 0;
    const int M = (sh_degree + 1) * (sh_degree + 1);
    std::string h = "ply\nformat binary_little_endian 1.0\nelement vertex " + std::to_string(num_vertices) + "\n";
    auto prop = [&](const std::string& n) { h += "property float " + n + "\n"; };
    for (const char* n : {"x", "y", "z", "nx", "ny", "nz"}) prop(n);
    for (int i = 0; i < 3; ++i) prop("f_dc_" + std::to_string(i));
    for (int i = 0; i < 3 * (M - 1); ++i) prop("f_rest_" + std::to_string(i));
    prop("opacity");
    for (int i = 0; i < 3; ++i) prop("scale_" + std::to_string(i));
    for (int i = 0; i < 4; ++i) prop("rot_" + std::to_string(i));
    h += "end_header\n";
    if (buf && buf_len >= h.size()) memcpy(buf, h.data(), h.size());
    return h.size();
}

static int ply_launch_status(const char* what) {
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        char msg[256];
        snprintf(msg, sizeof(msg), "[CUDA ERROR] in %s: %s", what, cudaGetErrorString(e));
        return gsr_set_error(GSR_ERR_CUDA, msg);
    }
    return GSR_OK;
}

extern "C" int gsr_ply_unpack(const gsr_ply_layout* layout, const char* records, int sh_degree, int activate, float* xyz,
                              float* shs, float* opacity, float* scales, float* rotations, void* stream) {
    if (!layout) return ply_fail("ply unpack: NULL layout");
    if (sh_degree < 0 || sh_degree > 3) return ply_fail("ply unpack: sh_degree must be 0..3");
    const uint64_t P = layout->num_vertices;
    if (P == 0) return GSR_OK;
    if (!records || !xyz || !shs || !opacity || !scales || !rotations) return ply_fail("ply unpack: NULL argument");
    if (P > ((uint64_t)1 << 31) / PLY_RECS * PLY_RECS) return ply_fail("ply unpack: too many vertices");
    PlyDev L;
    memcpy(L.off, layout->offset, sizeof(L.off));
    L.record_bytes = layout->record_bytes;
    L.n_rest = layout->n_rest;
    for (int a = 0; a < GSR_PLY_NUM_ATTR; ++a)
        if (L.off[a] >= 0 && (uint32_t)L.off[a] + 4 > L.record_bytes) return ply_fail("ply unpack: attribute offset outside the record");
    const int M = (sh_degree + 1) * (sh_degree + 1);
    const unsigned grid = (unsigned)((P + PLY_RECS - 1) / PLY_RECS);
    bool staged = (L.record_bytes % 4 == 0) && (L.record_bytes / 4 <= PLY_MAX_WORDS) && ((uintptr_t)records % 4 == 0);
    for (int a = 0; a < GSR_PLY_NUM_ATTR && staged; ++a)
        if (L.off[a] >= 0 && L.off[a] % 4 != 0) staged = false;
    const unsigned char* rec = reinterpret_cast<const unsigned char*>(records);
    if (staged) {
        const size_t smem = (size_t)PLY_RECS * L.record_bytes;
        ply_unpack_kernel<true><<<grid, PLY_RECS, smem, (cudaStream_t)stream>>>(L, rec, P, M, activate, xyz, shs, opacity, scales, rotations);
    } else {
        ply_unpack_kernel<false><<<grid, PLY_RECS, 0, (cudaStream_t)stream>>>(L, rec, P, M, activate, xyz, shs, opacity, scales, rotations);
    }
    return ply_launch_status("ply_unpack");
}

extern "C" int gsr_ply_pack(uint64_t P, int sh_degree, const float* xyz, const float* shs, const float* opacity,
                            const float* scales, const float* rotations, char* records, void* stream) {
    if (sh_degree < 0 || sh_degree > 3) return ply_fail("ply pack: sh_degree must be 0..3");
    if (P == 0) return GSR_OK;
    if (!records || !xyz || !shs || !opacity || !scales || !rotations) return ply_fail("ply pack: NULL argument");
    if ((uintptr_t)records % 4 != 0) return ply_fail("ply pack: records must be 4-byte aligned");
    const int M = (sh_degree + 1) * (sh_degree + 1);
    const uint64_t total = P * (uint64_t)(17 + 3 * (M - 1));
    const unsigned grid = (unsigned)std::min<uint64_t>((total + 255) / 256, 148 * 64);
    ply_pack_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(P, M, xyz, shs, opacity, scales, rotations, reinterpret_cast<float*>(records));
    return ply_launch_status("ply_pack");
}
