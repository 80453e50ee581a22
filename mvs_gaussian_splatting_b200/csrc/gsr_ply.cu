// gsr_ply.cu — the on-disk format either side of the render path: the body of a Gaussian
// model .ply is a row-major [P, n_props] binary32 matrix (one row per Gaussian: x y z nx ny nz f_dc_* f_rest_*
// opacity scale_* rot_*).  These two kernels move between that matrix and the six parameter arrays the
// rasterizer takes, on the device, so a 6 M-Gaussian checkpoint is one file read + one host-to-device copy
// + one launch instead of per-property Python loops.
//
// Reference: scene/gaussian_model.py:279-309 (save_ply: attribute order, features stored TRANSPOSED —
// f_rest_{c*(M-1)+k} = _features_rest[p, k, c] — normals written as zeros) and :316-358 (load_ply: properties
// looked up by name, f_rest_* / scale_* / rot_* sorted by their numeric suffix, reshape (P, 3, M-1) then
// transpose(1, 2)).  The name -> column resolution is host logic (ply.py); the kernels take the resolved
// column of every output element.  Pure data movement: results are bit-exact by construction.
//
// Layout: one CTA stages 128 consecutive rows (a contiguous span of the file body) in shared memory with
// coalesced 16-byte loads, then writes each parameter array's span of those rows with coalesced stores
// (and the reverse for packing).  Algorithmic bytes: 4 * n_props + 4 * n_values per Gaussian (484 B at SH 3).
#include "gsr_internal.cuh"

namespace gsr {

namespace {

constexpr int kPlyRows = 128;
constexpr int kPlyThreads = 256;

struct PlyArgs {
    long long P;
    int n_props;
    int width[6];             // xyz 3, f_dc 3, f_rest 3(M-1), opacity 1, scale 3, rot 4
    float* arr[6];
    float* rows;
    short col[6][48];         // file column of element j of a row of array a
};

template <bool kPack>
__global__ void __launch_bounds__(kPlyThreads) ply_kernel(const PlyArgs a) {
    extern __shared__ __align__(16) float s_rows[];   // [kPlyRows][n_props]
    __shared__ short s_col[6][48];                    // lanes index the table divergently: shared, not constant, memory
    for (int i = threadIdx.x; i < 6 * 48; i += kPlyThreads) s_col[i / 48][i % 48] = a.col[i / 48][i % 48];
    const long long row0 = (long long)blockIdx.x * kPlyRows;
    const int n = (int)min((long long)kPlyRows, a.P - row0);
    const int span = n * a.n_props;
    float* body = a.rows + row0 * a.n_props;          // 128 * n_props * 4 B per CTA: 16-byte aligned spans
    if (!kPack) {
        if (n == kPlyRows) {
            const float4* src = reinterpret_cast<const float4*>(body);
            for (int i = threadIdx.x; i < span / 4; i += kPlyThreads) reinterpret_cast<float4*>(s_rows)[i] = __ldcs(src + i);
        } else {
            for (int i = threadIdx.x; i < span; i += kPlyThreads) s_rows[i] = __ldcs(body + i);
        }
    } else {
        for (int i = threadIdx.x; i < span; i += kPlyThreads) s_rows[i] = 0.f;   // nx ny nz and unmapped columns
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        const int w = a.width[k];
        if (w == 0) continue;
        float* base = a.arr[k] + row0 * w;
        const float inv_w = 1.f / (float)w;           // i < 128 * 45: (i + 0.5) / w is exact enough to truncate
        for (int i = threadIdx.x; i < n * w; i += kPlyThreads) {
            const int r = __float2int_rz(((float)i + 0.5f) * inv_w), j = i - r * w;
            if (kPack) s_rows[r * a.n_props + s_col[k][j]] = __ldcs(base + i);
            else __stcs(base + i, s_rows[r * a.n_props + s_col[k][j]]);
        }
    }
    if (kPack) {
        __syncthreads();
        if (n == kPlyRows) {
            float4* dst = reinterpret_cast<float4*>(body);
            for (int i = threadIdx.x; i < span / 4; i += kPlyThreads) __stcs(dst + i, reinterpret_cast<float4*>(s_rows)[i]);
        } else {
            for (int i = threadIdx.x; i < span; i += kPlyThreads) __stcs(body + i, s_rows[i]);
        }
    }
}

int fill(PlyArgs& a, int64_t P, const gsr_ply_columns* c, float* rows, float* xyz, float* f_dc, float* f_rest,
         float* opacity, float* scale, float* rot) {
    a.P = P;
    a.n_props = c->n_props;
    const int widths[6] = {3, 3, c->n_rest, 1, 3, 4};
    float* arrs[6] = {xyz, f_dc, f_rest, opacity, scale, rot};
    const int16_t* cols[6] = {c->xyz, c->f_dc, c->f_rest, &c->opacity, c->scale, c->rot};
    unsigned char seen[GSR_PLY_MAX_PROPS] = {0};
    for (int k = 0; k < 6; ++k) {
        a.width[k] = widths[k];
        a.arr[k] = arrs[k];
        if (widths[k] && !arrs[k]) return GSR_ERR_INVALID_ARGUMENT;
        for (int j = 0; j < widths[k]; ++j) {
            const int col = cols[k][j];
            if (col < 0 || col >= c->n_props || seen[col]) return GSR_ERR_INVALID_ARGUMENT;   // each column once
            seen[col] = 1;
            a.col[k][j] = (short)col;
        }
    }
    a.rows = rows;
    return GSR_OK;
}

}  // namespace

int launch_ply(bool pack, int64_t P, const gsr_ply_columns* c, float* rows, float* xyz, float* f_dc, float* f_rest,
               float* opacity, float* scale, float* rot, cudaStream_t s) {
    PlyArgs a{};
    int st = fill(a, P, c, rows, xyz, f_dc, f_rest, opacity, scale, rot);
    if (st != GSR_OK) return st;
    const size_t smem = (size_t)kPlyRows * c->n_props * sizeof(float);
    const unsigned blocks = (unsigned)((P + kPlyRows - 1) / kPlyRows);
    if (pack) {
        if (smem > 48 * 1024) cudaFuncSetAttribute(ply_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        ply_kernel<true><<<blocks, kPlyThreads, smem, s>>>(a);
    } else {
        if (smem > 48 * 1024) cudaFuncSetAttribute(ply_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        ply_kernel<false><<<blocks, kPlyThreads, smem, s>>>(a);
    }
    GSR_AFTER_LAUNCH(s, 0);
    return GSR_OK;
}

}  // namespace gsr
