// ofb200 - multigrid hierarchy: coefficient coarsening, residual restriction, prolongation,
// dense coarsest-level solve.  Coarse node K sits on fine node 2K-1, so the Neumann mirror nodes
// (K=1 / i=1) of all levels coincide; coarse size = (N+3)/2.
#pragma once
#include "stencil.cuh"

namespace ofb {

__host__ __device__ __forceinline__ int coarse_size(int N) { return (N + 3) / 2; }

__device__ __forceinline__ int reflect_clamped(int i, int N, int far = 2) {
    i = min(max(i, 0), N - 1);
    return mirror_far(i, N, far);
}

// row index for grid transfers: clamp, then ghost rows to their mirror image (halo rows of a slab stay)
__device__ __forceinline__ int reflect_row(int i, const Geom& g) { return sol_row(min(max(i, 0), g.Nx - 1), g); }

// Level-1 coefficient planes straight from the image: full weighting (1/4,1/2,1/4)^2 of the
// level-0 coefficient fields (evaluated from P's 3x3 neighbourhood), pre-scaled for hinv = 1/2.
template <typename T>
__global__ void k_coef_from_image(Geom gf, const T* __restrict__ P, Geom gc, T* __restrict__ coef) {
    const int L = blockIdx.x * blockDim.x + threadIdx.x;
    const int K = blockIdx.y;
    const int pair = blockIdx.z;
    if (L >= gc.Ny) return;
    const T* img = P + (long long)pair * gf.img_pair;
    T acc[NCOEF];
#pragma unroll
    for (int q = 0; q < NCOEF; ++q) acc[q] = T(0);
    for (int di = -1; di <= 1; ++di) {
        const int i = reflect_row(2 * K - 1 + di, gf);     // slab halo rows: the image carries one more row beyond them
        const T wi = di == 0 ? T(0.5) : T(0.25);
        for (int dj = -1; dj <= 1; ++dj) {
            const int j = reflect_clamped(2 * L - 1 + dj, gf.Ny);
            const T w = wi * (dj == 0 ? T(0.5) : T(0.25));
            Row<T> pm, pc, pp;      // only v[0..2] are used (k = 1)
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                pm.v[c] = img[(long long)(i - 1) * gf.pitch + j - 1 + c];
                pc.v[c] = img[(long long)i * gf.pitch + j - 1 + c];
                pp.v[c] = img[(long long)(i + 1) * gf.pitch + j - 1 + c];
            }
            const Coef<T> c = coef_from_image<T>(pm, pc, pp, 1);
            acc[0] += w * c.II; acc[1] += w * c.IIx; acc[2] += w * c.IIxx; acc[3] += w * c.IIyy;
            acc[4] += w * c.IIxy; acc[5] += w * c.Ih; acc[6] += w * c.Ix;
        }
    }
    const T scale[NCOEF] = {T(0.25), T(0.5), T(1), T(1), T(1), T(0.5), T(1)};
    T* out = coef + (long long)pair * NCOEF * gc.plane + (long long)K * gc.pitch + L;
#pragma unroll
    for (int q = 0; q < NCOEF; ++q) out[q * gc.plane] = acc[q] * scale[q];
}

// Full-weighting restriction of `nplanes` planes, fine ghost ring by even reflection.
//   COEF = true : coefficient planes (per-plane rescale for the halved hinv, ghost ring kept)
//   COEF = false: residuals (coarse ghost ring set to 0)
//   SRC16 = true: the fine planes hold bfloat16 values (same element layout, two bytes each): the residual a float V-cycle wrote for
//                 nobody but this kernel (OpResidual::half_out)
template <typename T, bool COEF, bool SRC16 = false>
__global__ void k_restrict(Geom gf, const T* __restrict__ fine, long long fine_pair, Geom gc, T* __restrict__ coarse,
                           long long coarse_pair, int nplanes, const int* __restrict__ active) {
    pdl_enter();
    const int L = blockIdx.x * blockDim.x + threadIdx.x;
    const int K = blockIdx.y;
    const int pair = blockIdx.z;
    if (L >= gc.Ny) return;
    if (active != nullptr && !active[pair]) return;
    const bool ghost = (K == 0) || (K == gc.Nx - 1) || (L == 0) || (L == gc.Ny - 1);
    int ii[3], jj[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        ii[d] = reflect_row(2 * K - 2 + d, gf);
        jj[d] = reflect_clamped(2 * L - 2 + d, gf.Ny, gf.far_j);
    }
    const T w3[3] = {T(0.25), T(0.5), T(0.25)};
    const T scale[NCOEF] = {T(0.25), T(0.5), T(1), T(1), T(1), T(0.5), T(1)};
    for (int q = 0; q < nplanes; ++q) {
        const T* f = fine + (long long)pair * fine_pair + (long long)q * gf.plane;
        const unsigned short* f16 = reinterpret_cast<const unsigned short*>(fine) + (long long)pair * fine_pair + (long long)q * gf.plane;
        T acc = T(0);
        if (COEF || !ghost) {
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                T rowacc = T(0);
#pragma unroll
                for (int b = 0; b < 3; ++b) {
                    const long long idx = (long long)ii[a] * gf.pitch + jj[b];
                    rowacc += w3[b] * (SRC16 ? T(__uint_as_float((unsigned)f16[idx] << 16)) : f[idx]);
                }
                acc += w3[a] * rowacc;
            }
            if (COEF) acc *= scale[q];
        }
        coarse[(long long)pair * coarse_pair + (long long)q * gc.plane + (long long)K * gc.pitch + L] = acc;
    }
}

// x_fine += bilinear interpolation of the coarse correction (coarse ghost ring by even reflection).
template <typename T>
__global__ void k_prolong_add(Geom gc, const T* __restrict__ coarse, Geom gf, T* __restrict__ fine,
                              const int* __restrict__ active) {
    pdl_enter();
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    const int pair = blockIdx.z;
    if (j >= gf.Ny) return;
    if (active != nullptr && !active[pair]) return;
    if (is_ghost_row(i, gf) || j == 0 || j == gf.Ny - 1) return;      // slab halo rows are corrected like owned rows
    const int i0 = reflect_row((i + 1) >> 1, gc), i1 = reflect_row((i + 2) >> 1, gc);
    const int j0 = reflect_clamped((j + 1) >> 1, gc.Ny, gc.far_j), j1 = reflect_clamped((j + 2) >> 1, gc.Ny, gc.far_j);
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        const T* c = coarse + (long long)pair * gc.pair + (long long)q * gc.plane;
        const T e = T(0.25) * (c[(long long)i0 * gc.pitch + j0] + c[(long long)i0 * gc.pitch + j1]
                             + c[(long long)i1 * gc.pitch + j0] + c[(long long)i1 * gc.pitch + j1]);
        fine[(long long)pair * gf.pair + (long long)q * gf.plane + (long long)i * gf.pitch + j] += e;
    }
}

// Vectorised variant: one thread owns four consecutive fine columns (16-byte read-modify-write per plane); the six
// coarse values it interpolates from (columns 2m, 2m+1, 2m+2 of two coarse rows) come through the read-only path.
// Requires fp32 and 16-byte aligned fine rows with pitch % 4 == 0 (coarse levels pad their pitch: the vector that
// holds the last columns then also covers pad elements, which are written back unchanged).
__global__ void __launch_bounds__(128) k_prolong_add4(Geom gc, const float* __restrict__ coarse, Geom gf, float* __restrict__ fine,
                                                      const int* __restrict__ active) {
    pdl_enter();
    const int j4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const int i = blockIdx.y;
    const int pair = blockIdx.z;
    if (j4 >= gf.Ny) return;
    if (active != nullptr && !active[pair]) return;
    if (is_ghost_row(i, gf)) return;
    const int i0 = reflect_row((i + 1) >> 1, gc), i1 = reflect_row((i + 2) >> 1, gc);
    const int m2 = j4 >> 1;
    const int c0 = reflect_clamped(m2, gc.Ny, gc.far_j), c1 = reflect_clamped(m2 + 1, gc.Ny, gc.far_j),
              c2 = reflect_clamped(m2 + 2, gc.Ny, gc.far_j);
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        const float* c = coarse + (long long)pair * gc.pair + (long long)q * gc.plane;
        const float* r0 = c + (long long)i0 * gc.pitch;
        const float* r1 = c + (long long)i1 * gc.pitch;
        const float s0 = __ldg(r0 + c0) + __ldg(r1 + c0), s1 = __ldg(r0 + c1) + __ldg(r1 + c1), s2 = __ldg(r0 + c2) + __ldg(r1 + c2);
        float* fp = fine + (long long)pair * gf.pair + (long long)q * gf.plane + (long long)i * gf.pitch + j4;
        float4 v = *reinterpret_cast<float4*>(fp);
        const float e0 = 0.25f * (s0 + s1), e1 = 0.5f * s1, e2 = 0.25f * (s1 + s2), e3 = 0.5f * s2;
        const int last = gf.Ny - 2;                              // columns 0 and Ny-1 are ghost columns
        if (j4 > 0) v.x += e0;
        if (j4 + 1 <= last) v.y += e1;
        if (j4 + 2 <= last) v.z += e2;
        if (j4 + 3 <= last) v.w += e3;
        *reinterpret_cast<float4*>(fp) = v;
    }
}

// float32 coefficient planes [B][7][Nx][pitch] -> bfloat16 copies [B][7][Nx][pitch16] (round to nearest even; columns beyond Ny: 0).
// One thread per eight columns: two 16-byte loads, one 16-byte store (pitch % 4 == 0, pitch16 % 8 == 0, 16-byte aligned planes: the
// conditions of the streamed sweeps that read the copies).  grid = (ceil(Nx * pitch16 / 8 / 256), 1, pairs * NCOEF).
__global__ void __launch_bounds__(256) k_pack_bf16(Geom g, const float* __restrict__ coef, unsigned short* __restrict__ out, int pitch16,
                                                   long long plane16) {
    const int p8 = pitch16 >> 3;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = t / p8;
    if (i >= g.Nx) return;
    const int j = (t - i * p8) << 3;
    const int pq = blockIdx.z;                   // pair * NCOEF + plane
    const float* src = coef + (long long)pq * g.plane + (long long)i * g.pitch + j;
    float v[8];
    const float4 lo = j < g.pitch ? *reinterpret_cast<const float4*>(src) : make_float4(0.f, 0.f, 0.f, 0.f);
    const float4 hi = j + 4 < g.pitch ? *reinterpret_cast<const float4*>(src + 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    v[0] = lo.x; v[1] = lo.y; v[2] = lo.z; v[3] = lo.w; v[4] = hi.x; v[5] = hi.y; v[6] = hi.z; v[7] = hi.w;
    unsigned h[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        unsigned u = j + k < g.Ny ? __float_as_uint(v[k]) : 0u;
        u += 0x7fffu + ((u >> 16) & 1u);         // round to nearest even on the upper 16 bits (no NaN/inf in coefficients)
        h[k] = u >> 16;
    }
    const uint4 packed = make_uint4(h[0] | (h[1] << 16), h[2] | (h[3] << 16), h[4] | (h[5] << 16), h[6] | (h[7] << 16));
    *reinterpret_cast<uint4*>(out + (long long)pq * plane16 + (long long)i * pitch16 + j) = packed;
}

// ------------------------------------------------------------------ coarsest level: dense inverse
// Point evaluation of (S x)(i,j) for an arbitrary accessor, reusing apply_point (k = 1).
template <typename T, typename Acc>
__device__ __forceinline__ void apply_at(const Coef<T>& c, Acc x, int i, int j, T a, T be, T (&y)[3]) {
    Row<T> r[3][3];
#pragma unroll
    for (int q = 0; q < 3; ++q)
#pragma unroll
        for (int di = 0; di < 3; ++di)
#pragma unroll
            for (int dj = 0; dj < 3; ++dj) r[q][di].v[dj] = x(q, i + di - 1, j + dj - 1);
    Win<T> wx{r[0][0], r[0][1], r[0][2]}, wy{r[1][0], r[1][1], r[1][2]}, wg{r[2][0], r[2][1], r[2][2]};
    apply_point<T>(c, wx, wy, wg, 1, a, be, y[0], y[1], y[2]);
}

// Build the dense coarsest-level matrix (n x n, n = 3*Nx*Ny incl. ghost ring = identity rows) column by
// column from unit vectors, then invert in place by Gauss-Jordan with partial pivoting (fp64).
// One block per pair.  M is column-major scratch [B][n*n] doubles, Minv the result as T, column-major.
template <typename T>
__global__ void __launch_bounds__(1024) k_coarse_setup(Geom g, const T* __restrict__ coef, const double* __restrict__ alpha,
                                                       const double* __restrict__ beta, double h2, double* __restrict__ Mall,
                                                       T* __restrict__ Minv_all, int* __restrict__ piv_all, int use_smem) {
    const int pair = blockIdx.x;
    const int npix = g.Nx * g.Ny, n = 3 * npix;
    // the elimination is a chain of n dependent steps: run it in shared memory when the matrix fits
    // (108 x 108 doubles = 93 KB for the 6 x 6 coarsest level), in global memory otherwise
    extern __shared__ __align__(16) unsigned char coarse_smem[];
    double* M = use_smem ? reinterpret_cast<double*>(coarse_smem) : Mall + (long long)pair * n * n;
    const T* cf = coef + (long long)pair * NCOEF * g.plane;
    const double a = alpha[pair] * h2, be = beta[pair] * h2;
    // assemble: entry block for (row pixel rp, column unknown col)
    for (long long e = threadIdx.x; e < (long long)npix * n; e += blockDim.x) {
        const int rp = (int)(e % npix), col = (int)(e / npix);
        const int i = rp / g.Ny, j = rp - i * g.Ny;
        const int cq = col / npix, cp = col - cq * npix, ci = cp / g.Ny, cj = cp - ci * g.Ny;
        double y[3] = {0.0, 0.0, 0.0};
        const bool ghost = (i == 0) || (i == g.Nx - 1) || (j == 0) || (j == g.Ny - 1);
        if (ghost) {
            if (cp == rp) y[cq] = 1.0;
        } else if (abs(ci - i) <= 2) {
            Coef<double> c;
            const long long o = (long long)i * g.pitch + j;
            c.II = cf[o]; c.IIx = cf[g.plane + o]; c.IIxx = cf[2 * g.plane + o]; c.IIyy = cf[3 * g.plane + o];
            c.IIxy = cf[4 * g.plane + o]; c.Ih = cf[5 * g.plane + o]; c.Ix = cf[6 * g.plane + o];
            auto unit = [&](int q, int ii, int jj) -> double {
                // value of the unit vector e_col at (q,ii,jj) as the eliminated operator sees it
                if (q != cq) return 0.0;
                const bool gr = (ii == 0) || (ii == g.Nx - 1), gc2 = (jj == 0) || (jj == g.Ny - 1);
                const int mi = mirror_far(ii, g.Nx, g.far_i), mj = mirror_far(jj, g.Ny, g.far_j);
                if (mi != ci || mj != cj) return 0.0;
                if (ci == 0 || ci == g.Nx - 1 || cj == 0 || cj == g.Ny - 1) return 0.0;   // ghost unknowns are slaved
                return (gr && gc2) ? 2.0 : 1.0;
            };
            apply_at<double>(c, unit, i, j, a, be, y);
        }
#pragma unroll
        for (int q = 0; q < 3; ++q) M[(long long)col * n + (long long)q * npix + rp] = y[q];
    }
    __syncthreads();
    // Gauss-Jordan inversion in place with row pivoting, column-major M (element (r, c) at M[c*n + r]).
    // The n pivot steps are a dependent chain, so the cost is barriers and latency, not arithmetic: three block
    // barriers per step (pivot search | row swap | elimination with the pivot-row scaling folded in; the
    // pivot-column update overlaps the next search, which reads another column).
    int* piv = piv_all + (long long)pair * n;
    __shared__ int s_p;
    const int cgrp = threadIdx.x >> 3, r0 = threadIdx.x & 7, ngrp = blockDim.x >> 3;     // 8 threads share a column
    for (int k = 0; k < n; ++k) {
        if (threadIdx.x < 32) {   // pivot search in column k, rows k..n-1 (warp 0)
            double best = -1.0; int bi = k;
            for (int r = k + (int)threadIdx.x; r < n; r += 32) {
                const double v = fabs(M[(long long)k * n + r]);
                if (v > best) { best = v; bi = r; }
            }
            for (int o = 16; o > 0; o >>= 1) {
                const double ob = __shfl_xor_sync(FULL, best, o);
                const int oi = __shfl_xor_sync(FULL, bi, o);
                if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
            }
            if (threadIdx.x == 0) { s_p = bi; piv[k] = bi; }
        }
        __syncthreads();
        const int p = s_p;
        if (p != k) for (int c = threadIdx.x; c < n; c += blockDim.x) {   // swap rows k and p
            const double t = M[(long long)c * n + k]; M[(long long)c * n + k] = M[(long long)c * n + p]; M[(long long)c * n + p] = t;
        }
        __syncthreads();
        const double pinv = 1.0 / M[(long long)k * n + k];
        // columns c != k: scaled pivot-row entry m = M[k][c] * pinv, then M[r][c] -= M[r][k] * m for r != k
        for (int c = cgrp; c < n; c += ngrp) {
            if (c == k) continue;
            double* col = M + (long long)c * n;
            const double m = col[k] * pinv;
            const double* pc = M + (long long)k * n;
            for (int r = r0; r < n; r += 8)
                if (r != k) col[r] -= pc[r] * m;
        }
        __syncthreads();
        // pivot row: M[k][c] *= pinv (c != k); pivot column: M[r][k] = -M[r][k] * pinv (r != k), M[k][k] = pinv.
        // Nothing here is read by the next pivot search (column k+1, rows > k).
        for (int e = threadIdx.x; e < 2 * n; e += blockDim.x) {
            if (e < n) {
                const long long idx = (long long)k * n + e;                 // column k, row e
                M[idx] = (e == k) ? pinv : -M[idx] * pinv;
            } else if (e - n != k) {
                M[(long long)(e - n) * n + k] *= pinv;                      // row k, column e - n
            }
        }
    }
    __syncthreads();
    // undo the row swaps as column swaps in reverse order
    for (int k = n - 1; k >= 0; --k) {
        const int p = piv[k];
        if (p != k) for (int r = threadIdx.x; r < n; r += blockDim.x) {
            const double t = M[(long long)k * n + r]; M[(long long)k * n + r] = M[(long long)p * n + r]; M[(long long)p * n + r] = t;
        }
        __syncthreads();
    }
    T* Minv = Minv_all + (long long)pair * n * n;
    for (long long e = threadIdx.x; e < (long long)n * n; e += blockDim.x) Minv[e] = T(M[e]);
}

// x = Minv b on the coarsest level (dense, column-major, one block per pair; planes must be unpadded:
// pitch == Ny so that the flat index equals the unknown number).
template <typename T>
__global__ void k_coarse_solve(int n, const T* __restrict__ Minv_all, const T* __restrict__ b, T* __restrict__ x,
                               long long pair_stride, const int* __restrict__ active) {
    const int pair = blockIdx.x;
    if (active != nullptr && !active[pair]) return;
    extern __shared__ unsigned char smem_raw[];
    T* sb = reinterpret_cast<T*>(smem_raw);
    const T* Minv = Minv_all + (long long)pair * n * n;
    for (int c = threadIdx.x; c < n; c += blockDim.x) sb[c] = b[(long long)pair * pair_stride + c];
    __syncthreads();
    for (int r = threadIdx.x; r < n; r += blockDim.x) {
        T acc = T(0);
        for (int c = 0; c < n; ++c) acc += Minv[(long long)c * n + r] * sb[c];
        x[(long long)pair * pair_stride + r] = acc;
    }
}

}  // namespace ofb
