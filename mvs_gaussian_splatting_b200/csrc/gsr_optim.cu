// gsr_optim.cu — the step after backward: one fused Adam launch over all parameter
// groups of the Gaussian model, and the densification statistics the training loop keeps per Gaussian.
//
// Reference: scene/gaussian_model.py:245-268 builds torch.optim.Adam(l, lr=0.0, eps=1e-15) over six groups
// (xyz, f_dc, f_rest, opacity, scaling, rotation: 59 floats per Gaussian at SH degree 3) and train.py:140
// calls optimizer.step(); torch's implementation walks each group with several elementwise kernels.  Here
// every element is read and written exactly once (g, m, v, p in; m, v, p out: 28 B per float, HBM-bound),
// all groups in one launch (blockIdx.y = group).  Arithmetic follows torch.optim.Adam's single-tensor
// formula operation for operation (lerp, mul+addcmul, sqrt / bias_correction2_sqrt + eps, addcdiv).
// train.py:130-131 / scene/gaussian_model.py:775-777: max_radii2D, xyz_gradient_accum, denom.
#include "gsr_internal.cuh"

namespace gsr {

namespace {

struct AdamParams {
    int n_groups;
    float beta1_w;      // 1 - beta1 (lerp weight)
    float beta2, beta2_w;
    float eps, inv_unused;
    float bc2_sqrt;     // sqrt(1 - beta2^step)
    struct Group {
        float* p; const float* g; float* m; float* v;
        long long n;
        float step_size;  // lr / (1 - beta1^step)
    } grp[GSR_MAX_ADAM_GROUPS];
};

__device__ __forceinline__ void adam_elem(float& p, float g, float& m, float& v, const AdamParams& a, float step_size) {
    m = __fmaf_rn(a.beta1_w, __fadd_rn(g, -m), m);                    // exp_avg.lerp_(grad, 1 - beta1)
    v = __fmaf_rn(__fmul_rn(a.beta2_w, g), g, __fmul_rn(v, a.beta2)); // exp_avg_sq.mul_(beta2).addcmul_(g, g, 1 - beta2)
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), a.bc2_sqrt), a.eps);
    p = __fmaf_rn(-step_size, __fdiv_rn(m, denom), p);                // param.addcdiv_(exp_avg, denom, value=-step_size)
}

__global__ void __launch_bounds__(256) adam_kernel(const AdamParams a) {
    const AdamParams::Group& G = a.grp[blockIdx.y];
    const long long n4 = G.n >> 2;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 p = reinterpret_cast<float4*>(G.p)[i];
        const float4 g = __ldcs(reinterpret_cast<const float4*>(G.g) + i);
        float4 m = reinterpret_cast<float4*>(G.m)[i], v = reinterpret_cast<float4*>(G.v)[i];
        adam_elem(p.x, g.x, m.x, v.x, a, G.step_size);
        adam_elem(p.y, g.y, m.y, v.y, a, G.step_size);
        adam_elem(p.z, g.z, m.z, v.z, a, G.step_size);
        adam_elem(p.w, g.w, m.w, v.w, a, G.step_size);
        reinterpret_cast<float4*>(G.p)[i] = p;
        reinterpret_cast<float4*>(G.m)[i] = m;
        reinterpret_cast<float4*>(G.v)[i] = v;
    }
    // tail (n not a multiple of 4): first block of the group
    if (blockIdx.x == 0) {
        for (long long i = 4 * n4 + threadIdx.x; i < G.n; i += blockDim.x) {
            float p = G.p[i], m = G.m[i], v = G.v[i];
            adam_elem(p, G.g[i], m, v, a, G.step_size);
            G.p[i] = p; G.m[i] = m; G.v[i] = v;
        }
    }
}

__global__ void __launch_bounds__(256) densification_stats_kernel(int P, const int32_t* __restrict__ radii,
                                                                 const float* __restrict__ dL_dmeans2D,
                                                                 float* __restrict__ max_radii2D,
                                                                 float* __restrict__ xyz_gradient_accum,
                                                                 float* __restrict__ denom) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int r = radii[i];
    if (r <= 0) return;  // visibility_filter = radii > 0 (gaussian_renderer/__init__.py:311)
    max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);
    const float gx = dL_dmeans2D[3 * (size_t)i], gy = dL_dmeans2D[3 * (size_t)i + 1];
    xyz_gradient_accum[i] += __fsqrt_rn(__fmaf_rn(gy, gy, __fmul_rn(gx, gx)));
    denom[i] += 1.0f;
}

}  // namespace

int launch_adam(int n_groups, const gsr_adam_group* groups, double beta1, double beta2, double eps, int step,
                cudaStream_t s) {
    if (n_groups < 1 || n_groups > GSR_MAX_ADAM_GROUPS || step < 1) return GSR_ERR_INVALID_ARGUMENT;
    AdamParams a;
    a.n_groups = n_groups;
    a.beta1_w = (float)(1.0 - beta1);   // Python forms 1 - beta in binary64, the tensor op rounds it to binary32
    a.beta2 = (float)beta2;
    a.beta2_w = (float)(1.0 - beta2);
    a.eps = (float)eps;
    a.inv_unused = 0.f;
    const double bc1 = 1.0 - pow(beta1, (double)step), bc2 = 1.0 - pow(beta2, (double)step);
    a.bc2_sqrt = (float)sqrt(bc2);
    long long nmax = 0;
    for (int k = 0; k < n_groups; ++k) {
        const gsr_adam_group& g = groups[k];
        if (g.n < 0 || (g.n > 0 && (!g.param || !g.grad || !g.exp_avg || !g.exp_avg_sq))) return GSR_ERR_INVALID_ARGUMENT;
        const uintptr_t bits = reinterpret_cast<uintptr_t>(g.param) | reinterpret_cast<uintptr_t>(g.grad) |
                               reinterpret_cast<uintptr_t>(g.exp_avg) | reinterpret_cast<uintptr_t>(g.exp_avg_sq);
        if (bits & 15u) return GSR_ERR_INVALID_ARGUMENT;
        a.grp[k].p = g.param; a.grp[k].g = g.grad; a.grp[k].m = g.exp_avg; a.grp[k].v = g.exp_avg_sq;
        a.grp[k].n = g.n;
        a.grp[k].step_size = (float)(g.lr / bc1);
        nmax = g.n > nmax ? g.n : nmax;
    }
    if (nmax == 0) return GSR_OK;
    const long long want = (nmax / 4 + 255) / 256;
    const unsigned bx = (unsigned)(want < 1 ? 1 : (want > 148 * 16 ? 148 * 16 : want));
    adam_kernel<<<dim3(bx, (unsigned)n_groups), 256, 0, s>>>(a);
    GSR_AFTER_LAUNCH(s, 0);
    return GSR_OK;
}

int launch_densification_stats(int P, const int32_t* radii, const float* dL_dmeans2D, float* max_radii2D,
                               float* xyz_gradient_accum, float* denom, cudaStream_t s) {
    densification_stats_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, radii, dL_dmeans2D, max_radii2D, xyz_gradient_accum,
                                                             denom);
    GSR_AFTER_LAUNCH(s, 0);
    return GSR_OK;
}

}  // namespace gsr
