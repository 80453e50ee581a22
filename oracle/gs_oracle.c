/*
 * gs_oracle.c — CPU restatement of the differentiable Gaussian-splatting rasterizer.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under mvs_gaussian_splatting_b200/ may import, link or
 * call this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, and only as the checker or the timed CPU baseline.
 *
 * PARITY STATUS: *** parity unpinned ***.  The arithmetic of this path lives in the
 * third-party module `diff-gaussian-rasterization`
 * (https://github.com/graphdeco-inria/diff-gaussian-rasterization), which the reference
 * lists as an un-vendored git submodule with NO pinned commit (/root/reference/.gitmodules:4-6;
 * zero gitlinks recorded) and which is not installed here.  The reference has no tests or
 * golden vectors for the path.  This file therefore restates the *published*
 * algorithm (Kerbl et al. 2023 §6 + appendix; Zwicker et al. 2001 EWA splatting) as the 2023
 * API generation of that module implements it, anchored on what IS in the reference tree:
 *   - call site and argument meaning: gaussian_renderer/__init__.py:42-57,257-265
 *   - SH basis, constants, sign pattern:  utils/sh_utils.py:26-43,57-112 (+0.5 / clamp:
 *     gaussian_renderer/__init__.py:80-84)          -> pinned by tests/golden/sh_eval.npz
 *   - quaternion (r,x,y,z) -> R, L = R diag(s), Sigma = L L^T, 6-float packing:
 *     utils/general_utils.py:64-110, scene/gaussian_model.py:28-32
 *   - matrix conventions (transposed / flat column-major, znear .01, zfar 100):
 *     utils/graphics_utils.py:38-71, scene/cameras.py:48-57 -> tests/golden/cameras.npz
 *   - how outputs are consumed (dL/dmean2D scale, radii>0): scene/gaussian_model.py:775-777,
 *     train.py:130
 *
 * NUMERICS POLICY (forward geometry is compared BIT-EXACTLY with the CUDA path, so the
 * operation order is part of the spec; build with -ffp-contract=off):
 *   every product/sum below is a single IEEE-754 binary32 rounding unless written FMA();
 *   a left-associated sum of products  p0 + p1 + p2 [+ k]  is evaluated as
 *       FMA(p2a,p2b, FMA(p0a,p0b, RN(p1a*p1b))) [then a plain add of k]
 *   and  A*B - C*D  as  FMA(A,B, -RN(C*D)).  This is what nvcc 12.9 (PTX contraction) followed
 *   by ptxas (-fmad=true) emits for those expression shapes (tools/fma_probe.cu),
 *   i.e. the most likely numerics of a stock build of the third-party module on sm_100.
 *   Division, sqrt, ceil and float->int truncation are the IEEE / CUDA-saturating ones.
 *
 * The same source builds three times: binary32 (suffix _f32, the parity oracle), binary64
 * (suffix _f64, -DORC_F64, used for finite-difference / autograd gradient checks) and, compiled as
 * C++ with -DORC_RE (suffix _re), binary64 values that carry a running rounding-error bound of the
 * binary32 evaluation — the conditioning-aware scale of the gradient tests.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef ORC_RE
/* ---- third build (g++ -x c++ -DORC_RE): RUNNING ERROR ANALYSIS of the same source -----------------
 * `real` carries (v, e): the binary64 value and a first-order bound, in units of the binary32 unit
 * roundoff u = 2^-24, on the absolute error ANY binary32 evaluation of the same expression tree
 * accumulates (Higham, Accuracy and Stability of Numerical Algorithms, §3.3): every operation
 * propagates its operands' bounds through its partial derivatives and adds |result| for its own
 * rounding.  Inputs are exact (e = 0).  Gradient tests use e as the conditioning-aware scale of each
 * element: |cuda - oracle_f64| <= max(1e-4 |oracle|, K u e) for EVERY element (tests/helpers.py).
 * Discrete decisions (skip / stop / clamp selection) follow v. */
#ifndef __cplusplus
#error "the ORC_RE build compiles this file as C++ (operator overloading); see oracle/Makefile"
#endif
struct real {
    double v, e;
    real() = default;
    real(double x) : v(x), e((double)(float)x == x ? 0.0 : (x < 0 ? -x : x)) {}  /* constants round once */
    real(double x, double err) : v(x), e(err) {}
    explicit operator double() const { return v; }
};
static inline double re_abs(double x) { return x < 0 ? -x : x; }
static inline real operator-(real a) { return real(-a.v, a.e); }
static inline real operator+(real a, real b) { double v = a.v + b.v; return real(v, a.e + b.e + re_abs(v)); }
static inline real operator-(real a, real b) { double v = a.v - b.v; return real(v, a.e + b.e + re_abs(v)); }
static inline real operator*(real a, real b) {
    double v = a.v * b.v;
    return real(v, re_abs(a.v) * b.e + re_abs(b.v) * a.e + re_abs(v));
}
static inline real operator/(real a, real b) {
    double v = a.v / b.v;
    return real(v, a.e / re_abs(b.v) + re_abs(v) * b.e / re_abs(b.v) + re_abs(v));
}
static inline real& operator+=(real& a, real b) { a = a + b; return a; }
static inline real& operator*=(real& a, real b) { a = a * b; return a; }
static inline bool operator<(real a, real b) { return a.v < b.v; }
static inline bool operator>(real a, real b) { return a.v > b.v; }
static inline bool operator<=(real a, real b) { return a.v <= b.v; }
static inline bool operator>=(real a, real b) { return a.v >= b.v; }
static inline bool operator==(real a, real b) { return a.v == b.v; }
static inline bool operator!=(real a, real b) { return a.v != b.v; }
static inline real re_fma(real a, real b, real c) {   /* one rounding */
    double v = fma(a.v, b.v, c.v);
    return real(v, re_abs(a.v) * b.e + re_abs(b.v) * a.e + c.e + re_abs(v));
}
static inline real re_sqrt(real a) {
    double v = sqrt(a.v);
    return real(v, (v > 0 ? a.e / (2.0 * v) : 0.0) + v);
}
static inline real re_exp(real a) {   /* library / MUFU exponentials are good to ~2 ulp */
    double v = exp(a.v);
    return real(v, v * a.e + 2.0 * v);
}
static inline real re_ceil(real a) { return real(ceil(a.v), 0.0); }   /* only feeds the integer radius */
#define K(x) real(x)
#define FMA(a, b, c) re_fma((a), (b), (c))
#define SQRT(x) re_sqrt(x)
#define EXP(x) re_exp(x)
#define CEIL(x) re_ceil(x)
#define FMIN(a, b) ((b) < (a) ? (b) : (a))
#define FMAX(a, b) ((b) > (a) ? (b) : (a))
#define FABS(x) real(re_abs((x).v), (x).e)
#define FN(n) n##_re
#define TOD(x) ((x).v)
/* a decision `x < thr` whose operand sits within RE_UNSURE_K u e of the threshold may fall either way in
 * binary32 (measured: binary32 results stay within 0.35 u e of the binary64 ones, tests/test_oracle_golden.py) */
#define RE_UNSURE_K 2.0
#define UNSURE(x, thr) (re_abs((x).v - (thr)) <= RE_UNSURE_K * 5.9604644775390625e-08 * (x).e)
extern "C" {
#elif defined(ORC_F64)
typedef double real;
#define K(x) x
#define FMA(a, b, c) fma((a), (b), (c))
#define SQRT(x) sqrt(x)
#define EXP(x) exp(x)
#define CEIL(x) ceil(x)
#define FMIN(a, b) fmin((a), (b))
#define FMAX(a, b) fmax((a), (b))
#define FABS(x) fabs(x)
#define FN(n) n##_f64
#define TOD(x) ((double)(x))
#define UNSURE(x, thr) 0
#else
typedef float real;
#define K(x) x##f
#define FMA(a, b, c) fmaf((a), (b), (c))
#define SQRT(x) sqrtf(x)
#define EXP(x) expf(x)
#define CEIL(x) ceilf(x)
#define FMIN(a, b) fminf((a), (b))
#define FMAX(a, b) fmaxf((a), (b))
#define FABS(x) fabsf(x)
#define FN(n) n##_f32
#define TOD(x) ((double)(x))
#define UNSURE(x, thr) 0
#endif

#define TILE 16

/* ---- small helpers ------------------------------------------------------------------ */

/* p0 + p1 + p2 under the contraction policy above. */
static inline real dot3(real a0, real b0, real a1, real b1, real a2, real b2) {
    return FMA(a2, b2, FMA(a0, b0, a1 * b1));
}
/* p0 + p1 */
static inline real dot2(real a0, real b0, real a1, real b1) { return FMA(a0, b0, a1 * b1); }

/* float -> int32 with CUDA cvt.rzi.s32.f32 semantics (saturating, NaN -> 0). */
static inline int f2i(real v) {
    if (v != v) return 0;
    if (v >= (real)2147483648.0) return INT32_MAX;
    if (v <= (real)-2147483648.0) return INT32_MIN;
    return (int)TOD(v);
}
static inline int imin(int a, int b) { return a < b ? a : b; }
static inline int imax(int a, int b) { return a > b ? a : b; }

static inline uint32_t depth_bits(real d) {
    float f = (float)TOD(d);
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
}

/* Flat 16-float matrices are column-major (m[4*c+r]); the reference stores the transposed
 * torch matrix, scene/cameras.py:54-56, which reads back as exactly this. */
static inline void xform4x3(const real* m, const real* p, real* o) {
    o[0] = dot3(m[0], p[0], m[4], p[1], m[8], p[2]) + m[12];
    o[1] = dot3(m[1], p[0], m[5], p[1], m[9], p[2]) + m[13];
    o[2] = dot3(m[2], p[0], m[6], p[1], m[10], p[2]) + m[14];
}
static inline void xform4x4(const real* m, const real* p, real* o) {
    xform4x3(m, p, o);
    o[3] = dot3(m[3], p[0], m[7], p[1], m[11], p[2]) + m[15];
}

/* Sigma = R diag(s)^2 R^T from scale*mod and the quaternion AS GIVEN (not re-normalised).
 * Follows utils/general_utils.py:78-110 + scene/gaussian_model.py:28-32 for the maths and
 * packing (xx,xy,xz,yy,yz,zz). Mrow[k][j] = s_k * R[j][k]. */
static void cov3d_from_scale_rot(const real* scale, real mod, const real* q, real* c6) {
    real s0 = mod * scale[0], s1 = mod * scale[1], s2 = mod * scale[2];
    real r = q[0], x = q[1], y = q[2], z = q[3];
    real R00 = K(1.0) - K(2.0) * dot2(y, y, z, z);
    real R01 = K(2.0) * FMA(x, y, -(r * z));
    real R02 = K(2.0) * dot2(x, z, r, y);
    real R10 = K(2.0) * dot2(x, y, r, z);
    real R11 = K(1.0) - K(2.0) * dot2(x, x, z, z);
    real R12 = K(2.0) * FMA(y, z, -(r * x));
    real R20 = K(2.0) * FMA(x, z, -(r * y));
    real R21 = K(2.0) * dot2(y, z, r, x);
    real R22 = K(1.0) - K(2.0) * dot2(x, x, y, y);
    /* M[k][j] = s_k * R[j][k] */
    real M00 = s0 * R00, M01 = s0 * R10, M02 = s0 * R20;
    real M10 = s1 * R01, M11 = s1 * R11, M12 = s1 * R21;
    real M20 = s2 * R02, M21 = s2 * R12, M22 = s2 * R22;
    c6[0] = dot3(M00, M00, M10, M10, M20, M20);
    c6[1] = dot3(M00, M01, M10, M11, M20, M21);
    c6[2] = dot3(M00, M02, M10, M12, M20, M22);
    c6[3] = dot3(M01, M01, M11, M11, M21, M21);
    c6[4] = dot3(M01, M02, M11, M12, M21, M22);
    c6[5] = dot3(M02, M02, M12, M12, M22, M22);
}

/* EWA projection. Returns (a,b,c) of the 2x2 screen covariance incl. the +0.3 dilation.
 * Also returns the two rows of M = J*Rview (2x3) and the clamp multipliers for backward. */
static void cov2d_project(const real* mean, real fx, real fy, real tanx, real tany,
                          const real* c6, const real* v, real* abc, real* Mrows /*6 or NULL*/,
                          real* tclamped /*3 or NULL*/, real* gradmul /*2 or NULL*/) {
    real t[3];
    xform4x3(v, mean, t);
    real limx = K(1.3) * tanx, limy = K(1.3) * tany;
    real txtz = t[0] / t[2], tytz = t[1] / t[2];
    if (gradmul) {
        gradmul[0] = (txtz < -limx || txtz > limx) ? K(0.0) : K(1.0);
        gradmul[1] = (tytz < -limy || tytz > limy) ? K(0.0) : K(1.0);
    }
    t[0] = FMIN(limx, FMAX(-limx, txtz)) * t[2];
    t[1] = FMIN(limy, FMAX(-limy, tytz)) * t[2];
    if (tclamped) { tclamped[0] = t[0]; tclamped[1] = t[1]; tclamped[2] = t[2]; }
    real j00 = fx / t[2];
    real j11 = fy / t[2];
    real j02 = -(fx * t[0]) / (t[2] * t[2]);
    real j12 = -(fy * t[1]) / (t[2] * t[2]);
    /* M0[k] = j00*Rv[0][k] + j02*Rv[2][k];  M1[k] = j11*Rv[1][k] + j12*Rv[2][k],
     * Rv[r][k] = v[4k+r].  The zero entries of J contribute exact zeros. */
    real M0[3], M1[3];
    for (int k = 0; k < 3; ++k) {
        M0[k] = FMA(v[4 * k + 2], j02, v[4 * k + 0] * j00);
        M1[k] = FMA(v[4 * k + 2], j12, v[4 * k + 1] * j11);
    }
    real S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
    /* A_r[c] = sum_k M_r[k] * S[k][c] */
    real A0[3], A1[3];
    for (int c = 0; c < 3; ++c) {
        A0[c] = dot3(M0[0], S[0][c], M0[1], S[1][c], M0[2], S[2][c]);
        A1[c] = dot3(M1[0], S[0][c], M1[1], S[1][c], M1[2], S[2][c]);
    }
    abc[0] = dot3(A0[0], M0[0], A0[1], M0[1], A0[2], M0[2]) + K(0.3);
    abc[1] = dot3(A1[0], M0[0], A1[1], M0[1], A1[2], M0[2]);
    abc[2] = dot3(A1[0], M1[0], A1[1], M1[1], A1[2], M1[2]) + K(0.3);
    if (Mrows) {
        for (int k = 0; k < 3; ++k) { Mrows[k] = M0[k]; Mrows[3 + k] = M1[k]; }
    }
}

/* Real SH, degrees 0..3; constants and signs as utils/sh_utils.py:26-43,57-101. */
static const real SH_C0 = K(0.28209479177387814);
static const real SH_C1 = K(0.4886025119029199);
static const real SH_C2[5] = {K(1.0925484305920792), K(-1.0925484305920792), K(0.31539156525252005),
                              K(-1.0925484305920792), K(0.5462742152960396)};
static const real SH_C3[7] = {K(-0.5900435899266435), K(2.890611442640554), K(-0.4570457994644658),
                              K(0.3731763325901154), K(-0.4570457994644658), K(1.445305721320277),
                              K(-0.5900435899266435)};

/* sh: [M,3] coefficient-major, RGB innermost (scene/gaussian_model.py:176-179). */
static void sh_to_rgb(int deg, const real* pos, const real* campos, const real* sh, real* rgb,
                      uint8_t* clamped) {
    real dx = pos[0] - campos[0], dy = pos[1] - campos[1], dz = pos[2] - campos[2];
    real len = SQRT(dot3(dx, dx, dy, dy, dz, dz));
    real x = dx / len, y = dy / len, z = dz / len;
    for (int c = 0; c < 3; ++c) {
#define SH(k) sh[3 * (k) + c]
        real res = SH_C0 * SH(0);
        if (deg > 0) {
            res = FMA(-(SH_C1 * y), SH(1), res);
            res = FMA(SH_C1 * z, SH(2), res);
            res = FMA(-(SH_C1 * x), SH(3), res);
            if (deg > 1) {
                real xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                res = FMA(SH_C2[0] * xy, SH(4), res);
                res = FMA(SH_C2[1] * yz, SH(5), res);
                res = FMA(SH_C2[2] * (FMA(K(2.0), zz, -xx) - yy), SH(6), res);
                res = FMA(SH_C2[3] * xz, SH(7), res);
                res = FMA(SH_C2[4] * (xx - yy), SH(8), res);
                if (deg > 2) {
                    res = FMA(SH_C3[0] * y * FMA(K(3.0), xx, -yy), SH(9), res);
                    res = FMA(SH_C3[1] * xy * z, SH(10), res);
                    res = FMA(SH_C3[2] * y * (FMA(K(4.0), zz, -xx) - yy), SH(11), res);
                    res = FMA(SH_C3[3] * z * FMA(-K(3.0), yy, FMA(K(2.0), zz, -(K(3.0) * xx))),
                              SH(12), res);
                    res = FMA(SH_C3[4] * x * (FMA(K(4.0), zz, -xx) - yy), SH(13), res);
                    res = FMA(SH_C3[5] * z * (xx - yy), SH(14), res);
                    res = FMA(SH_C3[6] * x * FMA(-K(3.0), yy, xx), SH(15), res);
                }
            }
        }
#undef SH
        res = res + K(0.5);
        clamped[c] = res < K(0.0);
        rgb[c] = FMAX(res, K(0.0));
    }
}

/* Tile rectangle of a splat: [min,max) in tile units, clamped to the grid. */
static void tile_rect(real px, real py, int radius, int gx, int gy, int* r4) {
    real rf = (real)radius;
    r4[0] = imin(gx, imax(0, f2i((px - rf) / K(16.0))));
    r4[1] = imin(gy, imax(0, f2i((py - rf) / K(16.0))));
    r4[2] = imin(gx, imax(0, f2i((((px + rf) + K(16.0)) - K(1.0)) / K(16.0))));
    r4[3] = imin(gy, imax(0, f2i((((py + rf) + K(16.0)) - K(1.0)) / K(16.0))));
}

/* Direct entry points for the two sub-steps the reference restates in its own Python
 * (tests/golden/sh_eval.npz, cov3d.npz pin them). */
void FN(orc_sh_eval)(int P, int D, int M, const real* pos, const real* campos, const real* shs,
                     real* rgb, uint8_t* clamped) {
    for (int i = 0; i < P; ++i)
        sh_to_rgb(D, pos + 3 * (size_t)i, campos, shs + 3 * (size_t)M * i, rgb + 3 * (size_t)i,
                  clamped + 3 * (size_t)i);
}
void FN(orc_cov3d)(int P, const real* scales, real mod, const real* rots, real* cov6) {
    for (int i = 0; i < P; ++i)
        cov3d_from_scale_rot(scales + 3 * (size_t)i, mod, rots + 4 * (size_t)i, cov6 + 6 * (size_t)i);
}

/* ---- K1: preprocess forward ---------------------------------------------------------- */
/* One Gaussian at a time: near-plane cull (view z <= 0.2), cov3D, EWA cov2D, conic, radius,
 * pixel centre, tile rect, SH colour.  Outputs for culled Gaussians are zero. */
void FN(orc_preprocess)(int P, int D, int M, const real* means3D, const real* scales, real mod,
                        const real* rotations, const real* opacities, const real* shs,
                        const real* cov3D_precomp, const real* colors_precomp, const real* view,
                        const real* proj, const real* campos, int W, int H, real tanfovx,
                        real tanfovy, int* radii, real* means2D, real* depths, real* cov3D,
                        real* rgb, real* conic_opacity, uint32_t* tiles_touched, int* rects,
                        uint8_t* clamped, int i_begin, int i_end) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    const real fx = (real)W / (K(2.0) * tanfovx), fy = (real)H / (K(2.0) * tanfovy);
    (void)P;
    for (int i = i_begin; i < i_end; ++i) {
        radii[i] = 0;
        tiles_touched[i] = 0;
        means2D[2 * i] = means2D[2 * i + 1] = 0;
        depths[i] = 0;
        for (int k = 0; k < 6; ++k) cov3D[6 * i + k] = 0;
        for (int k = 0; k < 3; ++k) { rgb[3 * i + k] = 0; clamped[3 * i + k] = 0; }
        for (int k = 0; k < 4; ++k) { conic_opacity[4 * i + k] = 0; rects[4 * i + k] = 0; }

        const real* p = means3D + 3 * i;
        real pv[3];
        xform4x3(view, p, pv);
        if (!(pv[2] > K(0.2))) continue; /* near-plane cull; x/y frustum test is not applied */

        real ph[4];
        xform4x4(proj, p, ph);
        real pw = K(1.0) / (ph[3] + K(0.0000001));
        real ndcx = ph[0] * pw, ndcy = ph[1] * pw;

        real c6[6];
        if (cov3D_precomp) memcpy(c6, cov3D_precomp + 6 * i, sizeof(c6));
        else cov3d_from_scale_rot(scales + 3 * i, mod, rotations + 4 * i, c6);

        real abc[3];
        cov2d_project(p, fx, fy, tanfovx, tanfovy, c6, view, abc, NULL, NULL, NULL);
        real a = abc[0], b = abc[1], c = abc[2];
        real det = FMA(a, c, -(b * b));
        if (det == K(0.0)) continue;
        real det_inv = K(1.0) / det;
        real mid = K(0.5) * (a + c);
        real sq = SQRT(FMAX(K(0.1), FMA(mid, mid, -det)));
        real lam1 = mid + sq, lam2 = mid - sq;
        int rad = f2i(CEIL(K(3.0) * SQRT(FMAX(lam1, lam2))));
        real px = FMA(ndcx + K(1.0), (real)W, -K(1.0)) * K(0.5);
        real py = FMA(ndcy + K(1.0), (real)H, -K(1.0)) * K(0.5);
        int r4[4];
        tile_rect(px, py, rad, gx, gy, r4);
        int area = (r4[2] - r4[0]) * (r4[3] - r4[1]);
        if (area == 0) continue;

        if (colors_precomp) {
            for (int k = 0; k < 3; ++k) rgb[3 * i + k] = colors_precomp[3 * i + k];
        } else {
            sh_to_rgb(D, p, campos, shs + (size_t)3 * M * i, rgb + 3 * i, clamped + 3 * i);
        }
        depths[i] = pv[2];
        radii[i] = rad;
        means2D[2 * i] = px;
        means2D[2 * i + 1] = py;
        for (int k = 0; k < 6; ++k) cov3D[6 * i + k] = c6[k];
        conic_opacity[4 * i + 0] = c * det_inv;
        conic_opacity[4 * i + 1] = -b * det_inv;
        conic_opacity[4 * i + 2] = a * det_inv;
        conic_opacity[4 * i + 3] = opacities[i];
        for (int k = 0; k < 4; ++k) rects[4 * i + k] = r4[k];
        tiles_touched[i] = (uint32_t)area;
    }
}

/* ---- K11: markVisible ------------------------------------------------------------------ */
void FN(orc_mark_visible)(int P, const real* means3D, const real* view, uint8_t* present) {
    for (int i = 0; i < P; ++i) {
        real pv[3];
        xform4x3(view, means3D + 3 * i, pv);
        present[i] = pv[2] > K(0.2);
    }
}

/* ---- K2..K6: binning ------------------------------------------------------------------- */
/* Returns R = sum(tiles_touched). */
int64_t FN(orc_count_instances)(int P, const uint32_t* tiles_touched) {
    int64_t r = 0;
    for (int i = 0; i < P; ++i) r += tiles_touched[i];
    return r;
}

/* Emits (tile<<32 | depth bits, gaussian) for every tile of every visible splat's rect,
 * y-major, in Gaussian order; sorts stably by key (ties keep Gaussian order); writes per-tile
 * [start,end) ranges (untouched tiles stay (0,0)).  keys/vals: [R]; ranges: [tiles,2]. */
void FN(orc_bin)(int P, const int* radii, const int* rects, const real* depths, int W, int H,
                 int64_t R, uint64_t* keys_unsorted, uint32_t* vals_unsorted, uint64_t* keys,
                 uint32_t* vals, uint32_t* ranges) {
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    int64_t off = 0;
    for (int i = 0; i < P; ++i) {
        if (radii[i] <= 0) continue;
        const int* r4 = rects + 4 * i;
        uint32_t db = depth_bits(depths[i]);
        for (int y = r4[1]; y < r4[3]; ++y)
            for (int x = r4[0]; x < r4[2]; ++x) {
                keys_unsorted[off] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | db;
                vals_unsorted[off] = (uint32_t)i;
                ++off;
            }
    }
    /* stable LSD radix sort, 8 passes x 8 bits */
    uint64_t* ka = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)(R > 0 ? R : 1));
    uint32_t* va = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(R > 0 ? R : 1));
    memcpy(keys, keys_unsorted, sizeof(uint64_t) * (size_t)R);
    memcpy(vals, vals_unsorted, sizeof(uint32_t) * (size_t)R);
    uint64_t *ksrc = keys, *kdst = ka;
    uint32_t *vsrc = vals, *vdst = va;
    for (int pass = 0; pass < 8; ++pass) {
        int64_t cnt[257];
        memset(cnt, 0, sizeof(cnt));
        int sh = 8 * pass;
        for (int64_t j = 0; j < R; ++j) cnt[((ksrc[j] >> sh) & 255) + 1]++;
        for (int d = 0; d < 256; ++d) cnt[d + 1] += cnt[d];
        for (int64_t j = 0; j < R; ++j) {
            int64_t dst = cnt[(ksrc[j] >> sh) & 255]++;
            kdst[dst] = ksrc[j];
            vdst[dst] = vsrc[j];
        }
        uint64_t* tk = ksrc; ksrc = kdst; kdst = tk;
        uint32_t* tv = vsrc; vsrc = vdst; vdst = tv;
    }
    /* 8 passes: result is back in keys/vals */
    free(ka);
    free(va);
    memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)gx * gy);
    for (int64_t j = 0; j < R; ++j) {
        uint32_t t = (uint32_t)(keys[j] >> 32);
        if (j == 0 || (uint32_t)(keys[j - 1] >> 32) != t) ranges[2 * t] = (uint32_t)j;
        if (j == R - 1 || (uint32_t)(keys[j + 1] >> 32) != t) ranges[2 * t + 1] = (uint32_t)(j + 1);
    }
}

/* ---- K7: blend forward ------------------------------------------------------------------ */
/* Per pixel, front to back over its tile's list:
 *   power = -1/2 (A dx^2 + C dy^2) - B dx dy ; skip if power > 0
 *   alpha = min(0.99, o * exp(power))        ; skip if alpha < 1/255
 *   stop BEFORE blending when T (1-alpha) < 1e-4
 *   C += rgb * alpha * T ; T *= (1-alpha)
 * n_contrib = 1-based list position of the last blended entry; out = C + T*bg (planar CHW).
 * fragile[pix] = 1 when some decision above sat within `margin` (relative) of its threshold,
 * i.e. a different-but-valid exp() rounding may legally flip it; the running-error build (ORC_RE)
 * also flags decisions whose operand is within 2 u e of the threshold (ill-conditioned pairs of
 * needle splats, whose exponent carries far more than 1e-4 relative error in binary32). */
void FN(orc_blend_fwd)(int W, int H, const uint32_t* ranges, const uint32_t* point_list,
                       const real* means2D, const real* colors, const real* conic_opacity,
                       const real* bg, real* out_color, real* final_T, uint32_t* n_contrib,
                       uint8_t* fragile, real margin, int tile_begin, int tile_end) {
    const int gx = (W + TILE - 1) / TILE;
    for (int tile = tile_begin; tile < tile_end; ++tile) {
        int tx = tile % gx, ty = tile / gx;
        uint32_t s = ranges[2 * tile], e = ranges[2 * tile + 1];
        for (int ly = 0; ly < TILE; ++ly)
            for (int lx = 0; lx < TILE; ++lx) {
                int pxi = tx * TILE + lx, pyi = ty * TILE + ly;
                if (pxi >= W || pyi >= H) continue;
                real pxf = (real)pxi, pyf = (real)pyi;
                real T = K(1.0), C[3] = {0, 0, 0};
                uint32_t contributor = 0, last = 0;
                uint8_t frag = 0;
                for (uint32_t j = s; j < e; ++j) {
                    ++contributor;
                    uint32_t g = point_list[j];
                    real dx = means2D[2 * g] - pxf, dy = means2D[2 * g + 1] - pyf;
                    const real* co = conic_opacity + 4 * g;
                    real q = FMA(co[0] * dx, dx, (co[2] * dy) * dy);
                    real cross = (co[1] * dx) * dy;
                    real power = FMA(q, -K(0.5), -cross);
                    if (FABS(power) < margin * (FABS(q) + FABS(cross)) || UNSURE(power, 0.0)) frag = 1;
                    if (power > K(0.0)) continue;
                    real av = co[3] * EXP(power);
                    real alpha = FMIN(K(0.99), av);
                    if (FABS(alpha - K(1.0) / K(255.0)) <= margin * (K(1.0) / K(255.0)) || UNSURE(alpha, 1.0 / 255.0)) frag = 1;
                    if (alpha < K(1.0) / K(255.0)) continue;
                    real test_T = T * (K(1.0) - alpha);
                    if (FABS(test_T - K(0.0001)) <= K(8.0) * margin * K(0.0001) || UNSURE(test_T, 0.0001)) frag = 1;
                    if (test_T < K(0.0001)) break;
                    for (int ch = 0; ch < 3; ++ch)
                        C[ch] = FMA(colors[3 * g + ch] * alpha, T, C[ch]);
                    T = test_T;
                    last = contributor;
                }
                size_t pix = (size_t)pyi * W + pxi;
                final_T[pix] = T;
                n_contrib[pix] = last;
                if (fragile) fragile[pix] = frag;
                for (int ch = 0; ch < 3; ++ch)
                    out_color[(size_t)ch * H * W + pix] = FMA(T, bg[ch], C[ch]);
            }
    }
}

/* ---- K8: blend backward ----------------------------------------------------------------- */
/* Per pixel, back to front from its last contributor, rebuilding T by division.
 * Processes tiles [tile_begin, tile_end) so callers may thread it with private accumulators.
 * Accumulates (per Gaussian) dL/dmean2D (already scaled to NDC: *W/2, *H/2 — the convention
 * scene/gaussian_model.py:775-777 observes), dL/dconic as TRUE partials of (A,B,C),
 * dL/dopacity, dL/dcolor.  Accumulation here is in pixel order per tile, tiles in order. */
void FN(orc_blend_bwd)(int W, int H, const uint32_t* ranges, const uint32_t* point_list,
                       const real* means2D, const real* colors, const real* conic_opacity,
                       const real* bg, const real* final_T, const uint32_t* n_contrib,
                       const real* dL_dpix, real* dL_dmean2D /*[P,2]*/, real* dL_dconic /*[P,3]*/,
                       real* dL_dopacity /*[P]*/, real* dL_dcolor /*[P,3]*/, int tile_begin,
                       int tile_end) {
    const int gx = (W + TILE - 1) / TILE;
    const real sx = K(0.5) * (real)W, sy = K(0.5) * (real)H;
    for (int tile = tile_begin; tile < tile_end; ++tile) {
        int tx = tile % gx, ty = tile / gx;
        uint32_t s = ranges[2 * tile];
        for (int ly = 0; ly < TILE; ++ly)
            for (int lx = 0; lx < TILE; ++lx) {
                int pxi = tx * TILE + lx, pyi = ty * TILE + ly;
                if (pxi >= W || pyi >= H) continue;
                size_t pix = (size_t)pyi * W + pxi;
                real pxf = (real)pxi, pyf = (real)pyi;
                real Tfin = final_T[pix], T = Tfin;
                real g3[3], behind[3] = {0, 0, 0}, last_col[3] = {0, 0, 0}, last_alpha = 0;
                real bgdot = 0;
                for (int ch = 0; ch < 3; ++ch) {
                    g3[ch] = dL_dpix[(size_t)ch * H * W + pix];
                    bgdot += bg[ch] * g3[ch];
                }
                for (uint32_t k = n_contrib[pix]; k > 0; --k) {
                    uint32_t g = point_list[s + k - 1];
                    real dx = means2D[2 * g] - pxf, dy = means2D[2 * g + 1] - pyf;
                    const real* co = conic_opacity + 4 * g;
                    real q = FMA(co[0] * dx, dx, (co[2] * dy) * dy);
                    real power = FMA(q, -K(0.5), -((co[1] * dx) * dy));
                    if (power > K(0.0)) continue;
                    real G = EXP(power);
                    real alpha = FMIN(K(0.99), co[3] * G);
                    if (alpha < K(1.0) / K(255.0)) continue;
                    T = T / (K(1.0) - alpha);
                    real w = alpha * T;
                    real dL_dalpha = 0;
                    for (int ch = 0; ch < 3; ++ch) {
                        real c = colors[3 * g + ch];
                        behind[ch] = last_alpha * last_col[ch] + (K(1.0) - last_alpha) * behind[ch];
                        last_col[ch] = c;
                        dL_dalpha += (c - behind[ch]) * g3[ch];
                        dL_dcolor[3 * g + ch] += w * g3[ch];
                    }
                    dL_dalpha *= T;
                    last_alpha = alpha;
                    dL_dalpha += (-Tfin / (K(1.0) - alpha)) * bgdot;
                    real dL_dG = co[3] * dL_dalpha; /* the 0.99 cap passes gradient through */
                    real gdx = G * dx, gdy = G * dy;
                    real dG_ddx = -gdx * co[0] - gdy * co[1];
                    real dG_ddy = -gdy * co[2] - gdx * co[1];
                    dL_dmean2D[2 * g] += dL_dG * dG_ddx * sx;
                    dL_dmean2D[2 * g + 1] += dL_dG * dG_ddy * sy;
                    dL_dconic[3 * g + 0] += -K(0.5) * gdx * dx * dL_dG;
                    dL_dconic[3 * g + 1] += -gdx * dy * dL_dG;
                    dL_dconic[3 * g + 2] += -K(0.5) * gdy * dy * dL_dG;
                    dL_dopacity[g] += G * dL_dalpha;
                }
            }
    }
}

/* ---- K9 + K10: preprocess backward ------------------------------------------------------- */
/* Per visible Gaussian (radii > 0): conic -> cov2D -> cov3D/mean3D (EWA, zero through an
 * active frustum clamp), mean2D -> mean3D through the perspective divide, colour -> SH and
 * view direction -> mean3D (clamped channels pass nothing), cov3D -> scale / quaternion.
 * All outputs are fully written here (zeros for culled Gaussians).
 * Note: dL/dscale is taken w.r.t. (mod*scale), as the published implementation does. */
void FN(orc_preprocess_bwd)(int P, int D, int M, const real* means3D, const int* radii,
                            const real* shs, const uint8_t* clamped, const real* scales,
                            const real* rotations, real mod, const real* cov3D_in,
                            int cov_precomp, int colors_precomp, const real* view,
                            const real* proj, const real* campos, int W, int H, real tanfovx,
                            real tanfovy, const real* dL_dmean2D, const real* dL_dconic,
                            const real* dL_dcolor, real* dL_dmeans3D, real* dL_dcov3D,
                            real* dL_dsh, real* dL_dscales, real* dL_drots, int i_begin,
                            int i_end) {
    const real fx = (real)W / (K(2.0) * tanfovx), fy = (real)H / (K(2.0) * tanfovy);
    (void)P;
    for (int i = i_begin; i < i_end; ++i) {
        real* gm = dL_dmeans3D + 3 * i;
        real* gc = dL_dcov3D + 6 * i;
        gm[0] = gm[1] = gm[2] = 0;
        for (int k = 0; k < 6; ++k) gc[k] = 0;
        if (dL_dsh) for (int k = 0; k < 3 * M; ++k) dL_dsh[(size_t)3 * M * i + k] = 0;
        if (dL_dscales) for (int k = 0; k < 3; ++k) dL_dscales[3 * i + k] = 0;
        if (dL_drots) for (int k = 0; k < 4; ++k) dL_drots[4 * i + k] = 0;
        if (radii[i] <= 0) continue;
        const real* p = means3D + 3 * i;
        const real* c6 = cov3D_in + 6 * i;

        /* --- EWA backward --- */
        real abc[3], Mr[6], t[3], gmul[2];
        cov2d_project(p, fx, fy, tanfovx, tanfovy, c6, view, abc, Mr, t, gmul);
        real a = abc[0], b = abc[1], c = abc[2];
        real det = a * c - b * b;
        real d2inv = K(1.0) / (det * det + K(0.0000001));
        real gA = dL_dconic[3 * i], gB = dL_dconic[3 * i + 1], gC = dL_dconic[3 * i + 2];
        real dLa = 0, dLb = 0, dLc = 0;
        if (d2inv != 0) {
            dLa = d2inv * (-c * c * gA + b * c * gB + (det - a * c) * gC);
            dLc = d2inv * (-a * a * gC + a * b * gB + (det - a * c) * gA);
            dLb = d2inv * (K(2.0) * b * c * gA - (det + K(2.0) * b * b) * gB + K(2.0) * a * b * gC);
            const real* M0 = Mr; const real* M1 = Mr + 3;
            gc[0] = M0[0] * M0[0] * dLa + M0[0] * M1[0] * dLb + M1[0] * M1[0] * dLc;
            gc[3] = M0[1] * M0[1] * dLa + M0[1] * M1[1] * dLb + M1[1] * M1[1] * dLc;
            gc[5] = M0[2] * M0[2] * dLa + M0[2] * M1[2] * dLb + M1[2] * M1[2] * dLc;
            gc[1] = K(2.0) * M0[0] * M0[1] * dLa + (M0[0] * M1[1] + M0[1] * M1[0]) * dLb + K(2.0) * M1[0] * M1[1] * dLc;
            gc[2] = K(2.0) * M0[0] * M0[2] * dLa + (M0[0] * M1[2] + M0[2] * M1[0]) * dLb + K(2.0) * M1[0] * M1[2] * dLc;
            gc[4] = K(2.0) * M0[2] * M0[1] * dLa + (M0[1] * M1[2] + M0[2] * M1[1]) * dLb + K(2.0) * M1[1] * M1[2] * dLc;
        }
        {
            const real* M0 = Mr; const real* M1 = Mr + 3;
            real S[3][3] = {{c6[0], c6[1], c6[2]}, {c6[1], c6[3], c6[4]}, {c6[2], c6[4], c6[5]}};
            real SM0[3], SM1[3]; /* Sigma * M_r^T */
            for (int k = 0; k < 3; ++k) {
                SM0[k] = S[k][0] * M0[0] + S[k][1] * M0[1] + S[k][2] * M0[2];
                SM1[k] = S[k][0] * M1[0] + S[k][1] * M1[1] + S[k][2] * M1[2];
            }
            real dM0[3], dM1[3];
            for (int k = 0; k < 3; ++k) {
                dM0[k] = K(2.0) * SM0[k] * dLa + SM1[k] * dLb;
                dM1[k] = K(2.0) * SM1[k] * dLc + SM0[k] * dLb;
            }
            /* M0 = j00*Rv[0] + j02*Rv[2], M1 = j11*Rv[1] + j12*Rv[2]; Rv[r][k] = view[4k+r] */
            real dj00 = 0, dj02 = 0, dj11 = 0, dj12 = 0;
            for (int k = 0; k < 3; ++k) {
                dj00 += view[4 * k + 0] * dM0[k];
                dj02 += view[4 * k + 2] * dM0[k];
                dj11 += view[4 * k + 1] * dM1[k];
                dj12 += view[4 * k + 2] * dM1[k];
            }
            real tz = K(1.0) / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
            real dtx = gmul[0] * -fx * tz2 * dj02;
            real dty = gmul[1] * -fy * tz2 * dj12;
            real dtz = -fx * tz2 * dj00 - fy * tz2 * dj11 + (K(2.0) * fx * t[0]) * tz3 * dj02 +
                       (K(2.0) * fy * t[1]) * tz3 * dj12;
            /* t = Rv p + tv  ->  dp = Rv^T dt */
            gm[0] = view[0] * dtx + view[1] * dty + view[2] * dtz;
            gm[1] = view[4] * dtx + view[5] * dty + view[6] * dtz;
            gm[2] = view[8] * dtx + view[9] * dty + view[10] * dtz;
        }

        /* --- mean2D -> mean3D through the perspective divide --- */
        {
            real ph[4];
            xform4x4(proj, p, ph);
            real mw = K(1.0) / (ph[3] + K(0.0000001));
            real mul1 = ph[0] * mw * mw, mul2 = ph[1] * mw * mw;
            real g0 = dL_dmean2D[2 * i], g1 = dL_dmean2D[2 * i + 1];
            gm[0] += (proj[0] * mw - proj[3] * mul1) * g0 + (proj[1] * mw - proj[3] * mul2) * g1;
            gm[1] += (proj[4] * mw - proj[7] * mul1) * g0 + (proj[5] * mw - proj[7] * mul2) * g1;
            gm[2] += (proj[8] * mw - proj[11] * mul1) * g0 + (proj[9] * mw - proj[11] * mul2) * g1;
        }

        /* --- colour -> SH coefficients and view direction --- */
        if (!colors_precomp && dL_dsh) {
            real dir0[3] = {p[0] - campos[0], p[1] - campos[1], p[2] - campos[2]};
            real len = SQRT(dir0[0] * dir0[0] + dir0[1] * dir0[1] + dir0[2] * dir0[2]);
            real x = dir0[0] / len, y = dir0[1] / len, z = dir0[2] / len;
            real gdir[3] = {0, 0, 0};
            real* gsh = dL_dsh + (size_t)3 * M * i;
            const real* sh = shs + (size_t)3 * M * i;
            for (int ch = 0; ch < 3; ++ch) {
                real g = clamped[3 * i + ch] ? K(0.0) : dL_dcolor[3 * i + ch];
#define SH(k) sh[3 * (k) + ch]
#define GSH(k) gsh[3 * (k) + ch]
                real ddx = 0, ddy = 0, ddz = 0; /* d rgb / d dir */
                GSH(0) = SH_C0 * g;
                if (D > 0) {
                    GSH(1) = -SH_C1 * y * g;
                    GSH(2) = SH_C1 * z * g;
                    GSH(3) = -SH_C1 * x * g;
                    ddx = -SH_C1 * SH(3);
                    ddy = -SH_C1 * SH(1);
                    ddz = SH_C1 * SH(2);
                    if (D > 1) {
                        real xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                        GSH(4) = SH_C2[0] * xy * g;
                        GSH(5) = SH_C2[1] * yz * g;
                        GSH(6) = SH_C2[2] * (K(2.0) * zz - xx - yy) * g;
                        GSH(7) = SH_C2[3] * xz * g;
                        GSH(8) = SH_C2[4] * (xx - yy) * g;
                        ddx += SH_C2[0] * y * SH(4) + SH_C2[2] * K(2.0) * -x * SH(6) + SH_C2[3] * z * SH(7) + SH_C2[4] * K(2.0) * x * SH(8);
                        ddy += SH_C2[0] * x * SH(4) + SH_C2[1] * z * SH(5) + SH_C2[2] * K(2.0) * -y * SH(6) + SH_C2[4] * K(2.0) * -y * SH(8);
                        ddz += SH_C2[1] * y * SH(5) + SH_C2[2] * K(2.0) * K(2.0) * z * SH(6) + SH_C2[3] * x * SH(7);
                        if (D > 2) {
                            GSH(9) = SH_C3[0] * y * (K(3.0) * xx - yy) * g;
                            GSH(10) = SH_C3[1] * xy * z * g;
                            GSH(11) = SH_C3[2] * y * (K(4.0) * zz - xx - yy) * g;
                            GSH(12) = SH_C3[3] * z * (K(2.0) * zz - K(3.0) * xx - K(3.0) * yy) * g;
                            GSH(13) = SH_C3[4] * x * (K(4.0) * zz - xx - yy) * g;
                            GSH(14) = SH_C3[5] * z * (xx - yy) * g;
                            GSH(15) = SH_C3[6] * x * (xx - K(3.0) * yy) * g;
                            ddx += SH_C3[0] * SH(9) * K(3.0) * K(2.0) * xy + SH_C3[1] * SH(10) * yz +
                                   SH_C3[2] * SH(11) * -K(2.0) * xy + SH_C3[3] * SH(12) * -K(3.0) * K(2.0) * xz +
                                   SH_C3[4] * SH(13) * (-K(3.0) * xx + K(4.0) * zz - yy) +
                                   SH_C3[5] * SH(14) * K(2.0) * xz + SH_C3[6] * SH(15) * K(3.0) * (xx - yy);
                            ddy += SH_C3[0] * SH(9) * K(3.0) * (xx - yy) + SH_C3[1] * SH(10) * xz +
                                   SH_C3[2] * SH(11) * (-K(3.0) * yy + K(4.0) * zz - xx) +
                                   SH_C3[3] * SH(12) * -K(3.0) * K(2.0) * yz + SH_C3[4] * SH(13) * -K(2.0) * xy +
                                   SH_C3[5] * SH(14) * -K(2.0) * yz + SH_C3[6] * SH(15) * -K(3.0) * K(2.0) * xy;
                            ddz += SH_C3[1] * SH(10) * xy + SH_C3[2] * SH(11) * K(4.0) * K(2.0) * yz +
                                   SH_C3[3] * SH(12) * K(3.0) * (K(2.0) * zz - xx - yy) +
                                   SH_C3[4] * SH(13) * K(4.0) * K(2.0) * xz + SH_C3[5] * SH(14) * (xx - yy);
                        }
                    }
                }
#undef SH
#undef GSH
                gdir[0] += ddx * g;
                gdir[1] += ddy * g;
                gdir[2] += ddz * g;
            }
            /* d(v/|v|) = (|v|^2 I - v v^T) / |v|^3 */
            real s2 = dir0[0] * dir0[0] + dir0[1] * dir0[1] + dir0[2] * dir0[2];
            real inv32 = K(1.0) / SQRT(s2 * s2 * s2);
            real vd = dir0[0] * gdir[0] + dir0[1] * gdir[1] + dir0[2] * gdir[2];
            for (int k = 0; k < 3; ++k) gm[k] += (s2 * gdir[k] - dir0[k] * vd) * inv32;
        }

        /* --- cov3D -> scale, quaternion --- */
        if (!cov_precomp && dL_dscales) {
            const real* q = rotations + 4 * i;
            real s[3] = {mod * scales[3 * i], mod * scales[3 * i + 1], mod * scales[3 * i + 2]};
            real r = q[0], x = q[1], y = q[2], z = q[3];
            real R[3][3] = {
                {K(1.0) - K(2.0) * (y * y + z * z), K(2.0) * (x * y - r * z), K(2.0) * (x * z + r * y)},
                {K(2.0) * (x * y + r * z), K(1.0) - K(2.0) * (x * x + z * z), K(2.0) * (y * z - r * x)},
                {K(2.0) * (x * z - r * y), K(2.0) * (y * z + r * x), K(1.0) - K(2.0) * (x * x + y * y)}};
            /* symmetric dL/dSigma with halved off-diagonals */
            real Gs[3][3] = {{gc[0], K(0.5) * gc[1], K(0.5) * gc[2]},
                             {K(0.5) * gc[1], gc[3], K(0.5) * gc[4]},
                             {K(0.5) * gc[2], K(0.5) * gc[4], gc[5]}};
            /* Sigma = R S^2 R^T: dL/dR = 2 Gs R S^2 ; dL/ds_k = 2 s_k (R^T Gs R)_kk */
            real GR[3][3];
            for (int a2 = 0; a2 < 3; ++a2)
                for (int k = 0; k < 3; ++k)
                    GR[a2][k] = Gs[a2][0] * R[0][k] + Gs[a2][1] * R[1][k] + Gs[a2][2] * R[2][k];
            real dR[3][3];
            for (int k = 0; k < 3; ++k) {
                real rgr = R[0][k] * GR[0][k] + R[1][k] * GR[1][k] + R[2][k] * GR[2][k];
                dL_dscales[3 * i + k] = K(2.0) * s[k] * rgr;
                for (int a2 = 0; a2 < 3; ++a2) dR[a2][k] = K(2.0) * GR[a2][k] * s[k] * s[k];
            }
            real* gq = dL_drots + 4 * i;
            gq[0] = K(2.0) * (z * (dR[1][0] - dR[0][1]) + y * (dR[0][2] - dR[2][0]) + x * (dR[2][1] - dR[1][2]));
            gq[1] = K(2.0) * (y * (dR[0][1] + dR[1][0]) + z * (dR[0][2] + dR[2][0]) + r * (dR[2][1] - dR[1][2])) -
                    K(4.0) * x * (dR[1][1] + dR[2][2]);
            gq[2] = K(2.0) * (x * (dR[0][1] + dR[1][0]) + r * (dR[0][2] - dR[2][0]) + z * (dR[1][2] + dR[2][1])) -
                    K(4.0) * y * (dR[0][0] + dR[2][2]);
            gq[3] = K(2.0) * (r * (dR[1][0] - dR[0][1]) + x * (dR[0][2] + dR[2][0]) + y * (dR[1][2] + dR[2][1])) -
                    K(4.0) * z * (dR[0][0] + dR[1][1]);
        }
    }
}

#ifdef ORC_RE
}  /* extern "C" */
#endif
