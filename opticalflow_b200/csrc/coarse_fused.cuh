// ofb200 - the coarse tail of the multigrid V-cycle in ONE kernel.
//
// Levels of <= ~66x66 pixels hold 2 % of a cycle's arithmetic but, as separate launches (6 kernels per
// level per cycle, each a few microseconds of latency for microseconds of work), took 15-20 % of a solve.
// Here one CTA per frame pair runs the whole tail - pre-smoothing, residual, restriction, ... coarsest dense
// solve ..., prolongation, post-smoothing - with every vector of every tail level resident in shared memory
// (214 KB in fp32 for 66^2 + 34^2 + 18^2 + 10^2 + 6^2) and __syncthreads() between phases instead of kernel
// boundaries.  Coefficient planes are read-only and come from global memory / L1.
//
// The arithmetic is the same as the per-level kernels' (apply_point, block_solve, full weighting, bilinear
// interpolation, the LOAD_SOL ghost rule): vectors that feed the stencil keep their ghost ring materialised in
// shared memory (the writer of an interior value also writes its mirror copies; corner = 2 x value at (2,2)).
#pragma once
#include "transfer.cuh"

namespace ofb {

constexpr int FUSED_MAX_LEVELS = 8;

template <typename T> struct FusedLevel {
    int Nx, Ny, pitch;            // pitch / plane: geometry of the GLOBAL arrays of this level (coefficients, b, x)
    long long plane;
    const T* coef;                // [B][7][plane]
    double h2;                    // hinv^2
    int smem_off;                 // element offset of this level's three vectors (xa, xb, b) in shared memory
    int npix;                     // Nx*Ny (shared-memory planes are unpadded)
    int far_i, far_j;             // far-side ghost rule (Geom::far_i/far_j)
};

template <typename T> struct FusedArgs {
    int nlev;                     // tail levels, 0 = finest of the tail
    FusedLevel<T> lv[FUSED_MAX_LEVELS];
    const T* b_in;                // global right-hand side of tail level 0: [B][3][plane]
    T* x_out;                     // global result of tail level 0
    const double* alpha;
    const double* beta;
    const int* active;
    int nu_pre, nu_post;
    double w_pre[4], w_post[4];
    int dense;                    // coarsest level: 1 = dense inverse, 0 = `coarse_sweeps` Jacobi sweeps
    int dense_n;
    const T* Minv;                // [B][n*n] column-major
    int coarse_sweeps;
    double w_coarse;
};

// store an interior value and its LOAD_SOL ghost copies
template <typename T>
__device__ __forceinline__ void store_sol(T* __restrict__ f, int Nx, int Ny, int far_i, int far_j, int i, int j, T v) {
    f[i * Ny + j] = v;
    const bool ti = (i == 2), bi = (i == Nx - 1 - far_i), lj = (j == 2), rj = (j == Ny - 1 - far_j);
    if (ti) f[j] = v;
    if (bi) f[(Nx - 1) * Ny + j] = v;
    if (lj) f[i * Ny] = v;
    if (rj) f[i * Ny + Ny - 1] = v;
    const T v2 = v + v;
    if (ti && lj) f[0] = v2;
    if (ti && rj) f[Ny - 1] = v2;
    if (bi && lj) f[(Nx - 1) * Ny] = v2;
    if (bi && rj) f[(Nx - 1) * Ny + Ny - 1] = v2;
}

template <typename T>
__device__ __forceinline__ Coef<T> load_coef(const FusedLevel<T>& L, int pair, int i, int j) {
    const T* c = L.coef + (long long)pair * NCOEF * L.plane + (long long)i * L.pitch + j;
    Coef<T> cf;
    cf.II = c[0]; cf.IIx = c[L.plane]; cf.IIxx = c[2 * L.plane]; cf.IIyy = c[3 * L.plane];
    cf.IIxy = c[4 * L.plane]; cf.Ih = c[5 * L.plane]; cf.Ix = c[6 * L.plane];
    return cf;
}

// (S x)(i,j) from a shared-memory vector with materialised ghost ring
template <typename T>
__device__ __forceinline__ void stencil_smem(const T* __restrict__ x, int npix, int Ny, const Coef<T>& cf, int i, int j,
                                             T a, T be, T (&y)[3]) {
    auto acc = [&](int q, int ii, int jj) -> T { return x[q * npix + ii * Ny + jj]; };
    apply_at<T>(cf, acc, i, j, a, be, y);
}

template <typename T>
__global__ void __launch_bounds__(1024, 1) k_coarse_vcycle(FusedArgs<T> A) {
    extern __shared__ __align__(16) unsigned char fused_smem[];
    T* sm = reinterpret_cast<T*>(fused_smem);
    const int pair = blockIdx.x;
    pdl_enter();
    if (A.active != nullptr && A.active[pair] == 0) return;
    const int tid = threadIdx.x, nth = blockDim.x;
    const double alpha = A.alpha[pair], beta = A.beta[pair];
    int cur[FUSED_MAX_LEVELS];                      // which of (xa, xb) holds the current iterate of a level

    auto vec = [&](int l, int which) -> T* { return sm + A.lv[l].smem_off + which * 3 * A.lv[l].npix; };   // 0 xa, 1 xb, 2 b

    // ---- load the right-hand side of tail level 0 (ghost ring 0)
    {
        const FusedLevel<T>& L = A.lv[0];
        T* b = vec(0, 2);
        const T* src = A.b_in + (long long)pair * 3 * L.plane;
        for (int e = tid; e < 3 * L.npix; e += nth) {
            const int q = e / L.npix, p = e - q * L.npix, i = p / L.Ny, j = p - i * L.Ny;
            b[e] = src[(long long)q * L.plane + (long long)i * L.pitch + j];
        }
    }
    __syncthreads();

    // x_out = x_in + w D^-1 (b - S x_in) over one level (x_in == nullptr: zero guess, x_out = w D^-1 b)
    auto smooth_level = [&](int l, const T* xin, T* xout, const T* b, T w) {
        const FusedLevel<T>& L = A.lv[l];
        const T a = T(alpha * L.h2), be = T(beta * L.h2);
        const T inv_d2 = precond_rcp(T(-1) - T(4) * be);
        const int ni = L.Nx - 2, nj = L.Ny - 2;
        for (int p = tid; p < ni * nj; p += nth) {
            const int i = 1 + p / nj, j = 1 + p % nj;
            const Coef<T> cf = load_coef<T>(L, pair, i, j);
            const int o = i * L.Ny + j;
            T r0 = b[o], r1 = b[L.npix + o], r2 = b[2 * L.npix + o];
            T x0 = T(0), x1 = T(0), x2 = T(0);
            if (xin != nullptr) {
                T y[3];
                stencil_smem<T>(xin, L.npix, L.Ny, cf, i, j, a, be, y);
                r0 -= y[0]; r1 -= y[1]; r2 -= y[2];
                x0 = xin[o]; x1 = xin[L.npix + o]; x2 = xin[2 * L.npix + o];
            }
            T z0, z1, z2;
            block_solve<T>(cf, a, inv_d2, r0, r1, r2, z0, z1, z2);
            store_sol<T>(xout, L.Nx, L.Ny, L.far_i, L.far_j, i, j, x0 + w * z0);
            store_sol<T>(xout + L.npix, L.Nx, L.Ny, L.far_i, L.far_j, i, j, x1 + w * z1);
            store_sol<T>(xout + 2 * L.npix, L.Nx, L.Ny, L.far_i, L.far_j, i, j, x2 + w * z2);
        }
    };

    const int last = A.nlev - 1;
    // ---------------------------------------------------------------- down
    for (int l = 0; l < last; ++l) {
        const FusedLevel<T>& L = A.lv[l];
        const T* b = vec(l, 2);
        cur[l] = 0;
        smooth_level(l, nullptr, vec(l, 0), b, T(A.w_pre[0]));
        __syncthreads();
        for (int s = 1; s < A.nu_pre; ++s) {
            smooth_level(l, vec(l, cur[l]), vec(l, cur[l] ^ 1), b, T(A.w_pre[s < 4 ? s : 3]));
            cur[l] ^= 1;
            __syncthreads();
        }
        // residual into the free buffer (no ghost ring needed: the restriction mirrors indices)
        {
            const T* x = vec(l, cur[l]);
            T* r = vec(l, cur[l] ^ 1);
            const T a = T(alpha * L.h2), be = T(beta * L.h2);
            const int ni = L.Nx - 2, nj = L.Ny - 2;
            for (int p = tid; p < ni * nj; p += nth) {
                const int i = 1 + p / nj, j = 1 + p % nj;
                const Coef<T> cf = load_coef<T>(L, pair, i, j);
                T y[3];
                stencil_smem<T>(x, L.npix, L.Ny, cf, i, j, a, be, y);
                const int o = i * L.Ny + j;
                r[o] = b[o] - y[0]; r[L.npix + o] = b[L.npix + o] - y[1]; r[2 * L.npix + o] = b[2 * L.npix + o] - y[2];
            }
        }
        __syncthreads();
        // full-weighting restriction -> right-hand side of the next level (ghost ring 0)
        {
            const FusedLevel<T>& C = A.lv[l + 1];
            const T* r = vec(l, cur[l] ^ 1);
            T* bc = vec(l + 1, 2);
            const T w3[3] = {T(0.25), T(0.5), T(0.25)};
            for (int e = tid; e < 3 * C.npix; e += nth) {
                const int q = e / C.npix, p = e - q * C.npix, K = p / C.Ny, Lc = p - K * C.Ny;
                T acc = T(0);
                if (K >= 1 && K <= C.Nx - 2 && Lc >= 1 && Lc <= C.Ny - 2) {
                    const T* f = r + q * L.npix;
#pragma unroll
                    for (int a3 = 0; a3 < 3; ++a3) {
                        const int ii = reflect_clamped(2 * K - 2 + a3, L.Nx, L.far_i);
                        T rowacc = T(0);
#pragma unroll
                        for (int b3 = 0; b3 < 3; ++b3) rowacc += w3[b3] * f[ii * L.Ny + reflect_clamped(2 * Lc - 2 + b3, L.Ny, L.far_j)];
                        acc += w3[a3] * rowacc;
                    }
                }
                bc[e] = acc;
            }
        }
        __syncthreads();
    }
    // ---------------------------------------------------------------- coarsest level
    {
        const FusedLevel<T>& L = A.lv[last];
        const T* b = vec(last, 2);
        cur[last] = 0;
        if (A.dense) {
            const int n = A.dense_n;
            const T* Minv = A.Minv + (long long)pair * n * n;
            T* x = vec(last, 0);
            for (int r = tid; r < n; r += nth) {
                T acc = T(0);
                for (int c = 0; c < n; ++c) acc += Minv[(long long)c * n + r] * b[c];
                x[r] = acc;
            }
        } else {
            smooth_level(last, nullptr, vec(last, 0), b, T(A.w_coarse));
            for (int s = 1; s < A.coarse_sweeps; ++s) {
                __syncthreads();
                smooth_level(last, vec(last, cur[last]), vec(last, cur[last] ^ 1), b, T(A.w_coarse));
                cur[last] ^= 1;
            }
        }
        __syncthreads();
    }
    // ---------------------------------------------------------------- up
    for (int l = last - 1; l >= 0; --l) {
        const FusedLevel<T>& L = A.lv[l];
        const FusedLevel<T>& C = A.lv[l + 1];
        const T* b = vec(l, 2);
        {   // x += bilinear interpolation of the coarse correction (index reflection on the coarse grid)
            T* x = vec(l, cur[l]);
            const T* e = vec(l + 1, cur[l + 1]);
            const int ni = L.Nx - 2, nj = L.Ny - 2;
            for (int p = tid; p < ni * nj; p += nth) {
                const int i = 1 + p / nj, j = 1 + p % nj;
                const int i0 = reflect_clamped((i + 1) >> 1, C.Nx, C.far_i), i1 = reflect_clamped((i + 2) >> 1, C.Nx, C.far_i);
                const int j0 = reflect_clamped((j + 1) >> 1, C.Ny, C.far_j), j1 = reflect_clamped((j + 2) >> 1, C.Ny, C.far_j);
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    const T* c = e + q * C.npix;
                    const T d = T(0.25) * (c[i0 * C.Ny + j0] + c[i0 * C.Ny + j1] + c[i1 * C.Ny + j0] + c[i1 * C.Ny + j1]);
                    store_sol<T>(x + q * L.npix, L.Nx, L.Ny, L.far_i, L.far_j, i, j, x[q * L.npix + i * L.Ny + j] + d);
                }
            }
        }
        __syncthreads();
        for (int s = 0; s < A.nu_post; ++s) {
            smooth_level(l, vec(l, cur[l]), vec(l, cur[l] ^ 1), b, T(A.w_post[s < 4 ? s : 3]));
            cur[l] ^= 1;
            __syncthreads();
        }
    }
    // ---- result of tail level 0 back to global memory (interior; ghost ring 0)
    {
        const FusedLevel<T>& L = A.lv[0];
        const T* x = vec(0, cur[0]);
        T* dst = A.x_out + (long long)pair * 3 * L.plane;
        for (int e = tid; e < 3 * L.npix; e += nth) {
            const int q = e / L.npix, p = e - q * L.npix, i = p / L.Ny, j = p - i * L.Ny;
            const bool inside = (i >= 1) && (i <= L.Nx - 2) && (j >= 1) && (j <= L.Ny - 2);
            dst[(long long)q * L.plane + (long long)i * L.pitch + j] = inside ? x[e] : T(0);
        }
    }
}

}  // namespace ofb
