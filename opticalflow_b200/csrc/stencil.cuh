// ofb200 - the matrix-free 3-field operator (replaces the LIL/CSR/PETSc-AIJ assembly,
// reference source/optical_flow.py:836-1072; row formulas verified against the assembled matrix).
//
// Unknowns per pixel: ux, uy, g (net remodelling).  All coefficients of a row are evaluated at the
// row's own pixel.  The reference's 'dy' rule returns the x-derivative (:698-699), so dIdy == dIdx
// everywhere in the matrix; that is reproduced (Iy := Ix).
//
// Level l of the multigrid hierarchy uses grid spacing h = 2^l: second-order terms scale with
// hinv^2, first-order terms with hinv.  Level 0 computes the coefficients from the 3x3 neighbourhood
// of the previous frame P; coarse levels read 7 pre-scaled, full-weighting-averaged coefficient planes.
#pragma once
#include "common.cuh"

namespace ofb {

// Per-pixel coefficients, already scaled for the level.
template <typename T> struct Coef {
    T II;    // I*I * hinv^2
    T IIx;   // I*Ix * hinv          (Iy == Ix -> I*Iy identical)
    T IIxx;  // I*Ixx
    T IIyy;  // I*Iyy
    T IIxy;  // I*Ixy
    T Ih;    // I * hinv
    T Ix;    // Ix                   (zeroth-order coupling in the remodelling row)
};
constexpr int NCOEF = 7;

// level-0 coefficients from the image window (rows i-1,i,i+1; k = own index 1..V)
template <typename T>
__device__ __forceinline__ Coef<T> coef_from_image(const Row<T>& pm, const Row<T>& pc, const Row<T>& pp, int k) {
    Coef<T> c;
    const T I = pc.v[k];
    const T Ix = T(0.5) * (pp.v[k] - pm.v[k]);                       // :697  (axis 0 central difference)
    const T Ixx = pp.v[k] + pm.v[k] - T(2) * I;                      // :703
    const T Iyy = pc.v[k + 1] + pc.v[k - 1] - T(2) * I;              // :705
    const T Ixy = T(0.25) * (pp.v[k + 1] - pp.v[k - 1] - pm.v[k + 1] + pm.v[k - 1]);   // :701
    c.II = I * I; c.IIx = I * Ix; c.IIxx = I * Ixx; c.IIyy = I * Iyy; c.IIxy = I * Ixy;
    c.Ih = I; c.Ix = Ix;
    return c;
}

// 3-row windows of the three solution planes
template <typename T> struct Win {
    Row<T> m, c, p;   // rows i-1, i, i+1
};

// (S x)(i,j) for the three rows of pixel (i, j0+k-1); a = alpha*hinv^2, be = beta*hinv^2.
template <typename T>
__device__ __forceinline__ void apply_point(const Coef<T>& c, const Win<T>& ux, const Win<T>& uy, const Win<T>& g,
                                            int k, T a, T be, T& y0, T& y1, T& y2) {
    const T half = T(0.5), quarter = T(0.25);
    const T cross_uy = uy.m.v[k - 1] + uy.p.v[k + 1] - uy.m.v[k + 1] - uy.p.v[k - 1];
    const T cross_ux = ux.m.v[k - 1] + ux.p.v[k + 1] - ux.m.v[k + 1] - ux.p.v[k - 1];
    const T dux_di = ux.p.v[k] - ux.m.v[k], dux_dj = ux.c.v[k + 1] - ux.c.v[k - 1];
    const T duy_di = uy.p.v[k] - uy.m.v[k], duy_dj = uy.c.v[k + 1] - uy.c.v[k - 1];
    // ux row  (:843-887)
    y0 = (c.IIxx - T(2) * c.II - T(4) * a) * ux.c.v[k] + c.IIxy * uy.c.v[k]
       + (c.II - c.IIx + a) * ux.m.v[k] + (c.II + c.IIx + a) * ux.p.v[k]
       + a * (ux.c.v[k - 1] + ux.c.v[k + 1])
       + half * c.IIx * (duy_dj + duy_di)
       + quarter * c.II * cross_uy
       + half * c.Ih * (g.m.v[k] - g.p.v[k]);
    // uy row  (:892-936)
    y1 = (c.IIyy - T(2) * c.II - T(4) * a) * uy.c.v[k] + c.IIxy * ux.c.v[k]
       + (c.II - c.IIx + a) * uy.c.v[k - 1] + (c.II + c.IIx + a) * uy.c.v[k + 1]
       + a * (uy.m.v[k] + uy.p.v[k])
       + half * c.IIx * (dux_di + dux_dj)
       + quarter * c.II * cross_ux
       + half * c.Ih * (g.c.v[k - 1] - g.c.v[k + 1]);
    // remodelling row  (:942-960)
    y2 = (T(-1) - T(4) * be) * g.c.v[k] + c.Ix * (ux.c.v[k] + uy.c.v[k])
       + be * (g.m.v[k] + g.p.v[k] + g.c.v[k - 1] + g.c.v[k + 1])
       + half * c.Ih * (dux_di + duy_dj);
}

// Reciprocal used by the preconditioner only (smoother / Jacobi sweeps): fp32 takes the SFU approximation
// (MUFU.RCP, ~1 ulp) instead of the IEEE division sequence; fp64 stays exact.
__device__ __forceinline__ float precond_rcp(float v) { return __fdividef(1.0f, v); }
__device__ __forceinline__ double precond_rcp(double v) { return 1.0 / v; }

// z = D^-1 r for the pointwise 3x3 diagonal block [[d0,c,0],[c,d1,0],[Ix,Ix,d2]];  inv_d2 = 1/(-1-4 be).
template <typename T>
__device__ __forceinline__ void block_solve(const Coef<T>& c, T a, T inv_d2, T r0, T r1, T r2, T& z0, T& z1, T& z2) {
    const T d0 = c.IIxx - T(2) * c.II - T(4) * a;
    const T d1 = c.IIyy - T(2) * c.II - T(4) * a;
    const T inv_det = precond_rcp(d0 * d1 - c.IIxy * c.IIxy);
    z0 = (d1 * r0 - c.IIxy * r1) * inv_det;
    z1 = (d0 * r1 - c.IIxy * r0) * inv_det;
    z2 = (r2 - c.Ix * (z0 + z1)) * inv_d2;
}

}  // namespace ofb
