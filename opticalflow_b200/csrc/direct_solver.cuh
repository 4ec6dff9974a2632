// ofb200 - sparse direct solve of one frame pair on the device: replaces the reference's SuperLU branch
// (source/optical_flow.py:1146-1147, use_direct_solver=True) and is the fallback where the operator is indefinite and no
// preconditioned Krylov method converges (alpha below ~max I^2 / 1.5).
//
// Multifrontal LU on a geometric nested dissection of the interior pixels (plan: opticalflow_b200/ndplan.py).  A tree
// node owns a one-pixel separator line (a leaf: its few pixels); its front is a dense matrix over owned + border pixels,
// three unknowns per pixel, variable v = 3 * (position in the front) + field.  All nodes of a level are padded to one
// size, so a level is a batch: fronts are assembled from the stencil (nothing is read from a sparse matrix), the
// children's Schur complements are added in, F11 is factorised with partial pivoting, X = F11^-1 F12 and the Schur
// complement S = F22 - F21 X (one batched GEMM - where the ~1.4e12 fp64 flops of a 1024^2 pair are) stay in the front.
// The dense kernels are cuBLAS / cuSOLVER (batched LU for the many small fronts, one getrf per node for the few large
// ones); assembly, extend-add and the gathers/scatters of the triangular solves are the kernels below.  All in fp64.
#pragma once
#include <cublas_v2.h>
#include <cusolverDn.h>
#include "misc_kernels.cuh"

namespace ofb {

struct NdLevel {            // one tree level on the device
    int n_nodes, n_own, n_border;           // pixels per node (padded)
    const int* own;                         // [n_nodes][n_own] interior pixel ids, -1 = padding
    const int* border;                      // [n_nodes][n_border]
    const int* parent_slot;                 // [n_nodes][n_border] position of the border pixel in the parent's front
    double* F;                              // [n_nodes][nf*nf] column-major fronts, nf = 3 (n_own + n_border)
    int* piv;                               // [n_nodes][3 n_own]
    int* info;                              // [n_nodes]
    __host__ __device__ int no3() const { return 3 * n_own; }
    __host__ __device__ int nb3() const { return 3 * n_border; }
    __host__ __device__ int nf() const { return 3 * (n_own + n_border); }
};

__device__ __forceinline__ int nd_pixel(const NdLevel& L, int node, int pos) {
    return pos < L.n_own ? L.own[(long long)node * L.n_own + pos] : L.border[(long long)node * L.n_border + (pos - L.n_own)];
}

// Entry (row pixel rp, field rq; column pixel cp, field cq) of the interior operator S: the reference matrix
// (:836-1072) with its boundary rows substituted (ghost = mirror image, corner ghost = sum of its two images, :965-1070).
template <typename T>
__device__ double nd_entry(const T* __restrict__ img, const Geom& g, double a, double be, int rp, int rq, int cp, int cq) {
    const int my = g.Ny - 2;
    const int i = rp / my + 1, j = rp - (rp / my) * my + 1;          // frame coordinates
    const int ci = cp / my + 1, cj = cp - (cp / my) * my + 1;
    if (abs(ci - i) > 1 || abs(cj - j) > 1) return 0.0;
    Row<double> pm, pc, pp;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        pm.v[c] = (double)img[(long long)(i - 1) * g.pitch + j - 1 + c];
        pc.v[c] = (double)img[(long long)i * g.pitch + j - 1 + c];
        pp.v[c] = (double)img[(long long)(i + 1) * g.pitch + j - 1 + c];
    }
    const Coef<double> c = coef_from_image<double>(pm, pc, pp, 1);
    auto unit = [&](int q, int ii, int jj) -> double {       // the unit vector of the column as the stencil sees it
        if (q != cq) return 0.0;
        const bool gr = (ii == 0) || (ii == g.Nx - 1), gc = (jj == 0) || (jj == g.Ny - 1);
        if (mirror_index(ii, g.Nx) != ci || mirror_index(jj, g.Ny) != cj) return 0.0;
        return (gr && gc) ? 2.0 : 1.0;
    };
    double y[3];
    apply_at<double>(c, unit, i, j, a, be, y);
    return y[rq];
}

// Assemble the original matrix entries of a level's fronts: rows and columns of OWNED variables (entries between two
// border variables belong to fronts higher up).  grid = (nodes, owned variables, column chunks of 128).
template <typename T>
__global__ void __launch_bounds__(128) k_nd_assemble(NdLevel L, Geom g, const T* __restrict__ img, const double* __restrict__ alpha,
                                                     const double* __restrict__ beta, int pair) {
    const int node = blockIdx.x, v = blockIdx.y, o = blockIdx.z * 128 + threadIdx.x;
    const int nf = L.nf(), no3 = L.no3();
    if (o >= nf) return;
    double* F = L.F + (long long)node * nf * nf;
    const int vp = nd_pixel(L, node, v / 3), vq = v % 3;
    if (vp < 0) {                       // padded owned variable: identity row/column keeps F11 regular
        if (o == v) F[(long long)v * nf + v] = 1.0;
        return;
    }
    const int op = nd_pixel(L, node, o / 3), oq = o % 3;
    if (op < 0) return;
    const double a = alpha[pair], be = beta[pair];
    const double e = nd_entry<T>(img, g, a, be, vp, vq, op, oq);             // row v, column o
    if (e != 0.0) F[(long long)o * nf + v] = e;
    if (o >= no3) {                                                           // row o (border), column v (owned)
        const double e2 = nd_entry<T>(img, g, a, be, op, oq, vp, vq);
        if (e2 != 0.0) F[(long long)v * nf + o] = e2;
    }
}

// Add the Schur complements of the children with (index & 1) == parity into their parents' fronts (two launches, so
// that the two children of a node never write the same entry at the same time).  grid = (children / 2, nb3, chunks).
__global__ void __launch_bounds__(128) k_nd_extend_add(NdLevel C, NdLevel Pn, int parity) {
    const int child = 2 * blockIdx.x + parity, b = blockIdx.y, a = blockIdx.z * 128 + threadIdx.x;
    const int nbc = C.nb3();
    if (a >= nbc) return;
    const int sa = C.parent_slot[(long long)child * C.n_border + a / 3], sb = C.parent_slot[(long long)child * C.n_border + b / 3];
    if (sa < 0 || sb < 0) return;
    const int nfc = C.nf(), nfp = Pn.nf(), noc = C.no3();
    const double s = C.F[(long long)child * nfc * nfc + (long long)(noc + b) * nfc + (noc + a)];
    if (s != 0.0) Pn.F[(long long)(child >> 1) * nfp * nfp + (long long)(3 * sb + b % 3) * nfp + (3 * sa + a % 3)] += s;
}

// getrf's verdicts of a level -> one flag (1 + level of the first front with a zero pivot), read by the host once per pair
__global__ void k_nd_check_info(const int* __restrict__ info, int n, int level, int* __restrict__ flag) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n && info[k] != 0) atomicMax(flag, level + 1);
}

// pointer arrays of the batched cuBLAS calls: F11 (= front base), F12 block, and the per-node vectors
__global__ void k_nd_pointers(NdLevel L, double* z, double** Fptr, double** F12ptr, double** zptr) {
    const int node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= L.n_nodes) return;
    const long long nf = L.nf();
    Fptr[node] = L.F + node * nf * nf;
    F12ptr[node] = L.F + node * nf * nf + (long long)L.no3() * nf;
    zptr[node] = z + (long long)node * L.no3();
}

// global work vector w: three planes of M interior pixels, variable (pixel p, field q) at w[q * M + p]
__global__ void k_nd_gather(NdLevel L, const double* __restrict__ w, long long M, double* __restrict__ out, int border) {
    const int node = blockIdx.x, v = blockIdx.y * 128 + threadIdx.x;
    const int n3 = border ? L.nb3() : L.no3();
    if (v >= n3) return;
    const int p = border ? L.border[(long long)node * L.n_border + v / 3] : L.own[(long long)node * L.n_own + v / 3];
    out[(long long)node * n3 + v] = p < 0 ? 0.0 : w[(long long)(v % 3) * M + p];
}
__global__ void k_nd_store_own(NdLevel L, double* __restrict__ w, long long M, const double* __restrict__ z) {
    const int node = blockIdx.x, v = blockIdx.y * 128 + threadIdx.x;
    if (v >= L.no3()) return;
    const int p = L.own[(long long)node * L.n_own + v / 3];
    if (p >= 0) w[(long long)(v % 3) * M + p] = z[(long long)node * L.no3() + v];
}
// w[border] -= u   (siblings share separator pixels: atomics)
__global__ void k_nd_scatter_border(NdLevel L, double* __restrict__ w, long long M, const double* __restrict__ u) {
    const int node = blockIdx.x, v = blockIdx.y * 128 + threadIdx.x;
    if (v >= L.nb3()) return;
    const int p = L.border[(long long)node * L.n_border + v / 3];
    if (p >= 0) atomicAdd(w + (long long)(v % 3) * M + p, -u[(long long)node * L.nb3() + v]);
}

// interior of a three-plane vector of type T <-> the fp64 work vector
template <typename T>
__global__ void k_nd_load_rhs(Geom g, const T* __restrict__ b, double* __restrict__ w) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y, q = blockIdx.z;
    const int mx = g.Nx - 2, my = g.Ny - 2;
    if (j >= my || i >= mx) return;
    w[(long long)q * mx * my + (long long)i * my + j] = (double)b[q * g.plane + (long long)(i + 1) * g.pitch + j + 1];
}
template <typename T>
__global__ void k_nd_store_solution(Geom g, const double* __restrict__ w, T* __restrict__ x) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y, q = blockIdx.z;
    if (j >= g.Ny || i >= g.Nx) return;
    const int mx = g.Nx - 2, my = g.Ny - 2;
    const bool inside = i >= 1 && i <= g.Nx - 2 && j >= 1 && j <= g.Ny - 2;
    x[q * g.plane + (long long)i * g.pitch + j] = inside ? T(w[(long long)q * mx * my + (long long)(i - 1) * my + (j - 1)]) : T(0);
}
// the verdict of ofb_solve (k_verify) on the direct solution: |b - A x| <= rtol |b| + 8 u |D x|
__global__ void k_nd_relres(const PairState* __restrict__ st, double* __restrict__ relres, int* __restrict__ converged, int B, double rtol,
                            double unit_roundoff) {
    const int pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= B) return;
    const double bn = st[pair].bnorm;
    const double rr = bn > 0.0 ? sqrt(st[pair].acc[0]) / bn : 0.0;
    const double floor_rr = bn > 0.0 ? unit_roundoff * sqrt(st[pair].acc[1]) / bn : 0.0;
    if (relres != nullptr) relres[pair] = rr;
    if (converged != nullptr) converged[pair] = (isfinite(rr) && rr <= rtol + 8.0 * floor_rr) ? 1 : 0;
}

}  // namespace ofb
