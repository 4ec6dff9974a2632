// ofb200 - Krylov / smoother kernels built on the row-marching sweep.
//
// BiCGStab (the reference's -ksp_type bcgs, optical_flow.py:1081) in right-preconditioned form with all
// scalars resident on the device.  Dot products are accumulated in fp64 inside the stencil sweeps
// (warp shuffle -> block -> one atomicAdd per block), never in separate reduction kernels.
#pragma once
#include "sweep.cuh"

namespace ofb {

// Per-pair solver scalars (device resident).
struct PairState {
    double rho, rho_old, alpha, omega, beta_k;
    double acc[4];          // [0] rhat.v | t.s   [1] t.t   [2] r.r   [3] rhat.r
    double bnorm, relres;   // relres: the recurrence residual |r_k| / |b| the iteration stops on
    double true_relres;     // |b - A x| / |b| recomputed from the returned x in fp64 arithmetic (k_true_residual; reference :1150-1151)
    double floor_relres;    // u |D x| / |b|: what rounding x to the storage type alone leaves in the residual (u = unit round-off)
    int active, iters, status;
    int half_pending;       // converged at the BiCGStab half step: x += alpha ph is still owed (k_update_x_r pays it)
    int restarts;           // times the recurrence was restarted from the true residual (breakdown, or recurrence and true residual parted)
};
enum Status { ST_RUNNING = 0, ST_CONVERGED = 1, ST_MAXIT = 2, ST_BREAKDOWN = 3 };

// ------------------------------------------------------------------ sweep ops
// warp-level flush of NACC fp64 accumulators: shuffle reduce, one atomicAdd per accumulator per warp, reset
template <int NACC>
__device__ __forceinline__ void warp_flush(double (&acc)[NACC], double* __restrict__ dst) {
#pragma unroll
    for (int a = 0; a < NACC; ++a) {
        const double v = warp_sum(acc[a]);
        if ((threadIdx.x & 31) == 0 && v != 0.0) atomicAdd(dst + a, v);
        acc[a] = 0.0;
    }
}

// x_out = x + w D^-1 (b - S x)
template <typename T> struct OpSmooth {
    static constexpr bool NEED_X = true;
    static constexpr int NCEN = 1;
    static constexpr int V = VecOf<T>::V;
    Geom g; const T* cen[1]; T* xo; T w;      // cen[0] = b
    template <typename Cen>
    __device__ __forceinline__ void row(int pair, int i, int j0, const Coef<T> (&cf)[V], const T (&sx)[3][V],
                                        const T (&xc)[3][V], const bool (&inside)[V], T a, T be, const Cen& cv) {
        T out[3][V], b0[V], b1[V], b2[V];
        cv.get(0, 0, b0); cv.get(0, 1, b1); cv.get(0, 2, b2);
        const T inv_d2 = precond_rcp(T(-1) - T(4) * be);
#pragma unroll
        for (int k = 0; k < V; ++k) {
            T z0, z1, z2;
            block_solve<T>(cf[k], a, inv_d2, b0[k] - sx[0][k], b1[k] - sx[1][k], b2[k] - sx[2][k], z0, z1, z2);
            out[0][k] = inside[k] ? xc[0][k] + w * z0 : T(0);
            out[1][k] = inside[k] ? xc[1][k] + w * z1 : T(0);
            out[2][k] = inside[k] ? xc[2][k] + w * z2 : T(0);
        }
        store3<T>(xo, g, pair, i, j0, out);
    }
    __device__ __forceinline__ void finish(int, bool) {}
    __device__ __forceinline__ void flush(int) {}
};

// The first post-smoothing sweep of a level with the coarse-grid correction folded in: x_out = y + w D^-1 (b - S y),
// y = x + P e.  The streamed sweep stages the two coarse rows a fine window row interpolates from and adds P e to the
// window as it is loaded (stream_sweep.cuh), so the separate prolongation pass (read x, write x: 6 words per pixel)
// disappears.  The arithmetic of a row is OpSmooth's.
template <typename T> struct OpSmoothProlong : OpSmooth<T> {
    static constexpr bool PROLONG = true;
    const T* ec;            // coarse correction [B][3][coarse plane]
    Geom gc;                // its geometry
};

// r = b - S x ; INIT: also copy to r2 and dots acc[2] += r.r, acc[3] += r.r (initial residual: rhat = r)
template <typename T, bool INIT> struct OpResidual {
    static constexpr bool NEED_X = true;
    static constexpr int NCEN = 1;
    static constexpr int V = VecOf<T>::V;
    Geom g; const T* cen[1]; T* r; T* r2; PairState* st; double acc[1];      // cen[0] = b
    int half_out;       // float, !INIT: r receives bfloat16 values (two-byte elements, same layout): a V-cycle's residual is read by the
                        // restriction only (k_restrict<.., SRC16>), and right preconditioning does not care how the preconditioner rounds
    template <typename Cen>
    __device__ __forceinline__ void row(int pair, int i, int j0, const Coef<T> (&cf)[V], const T (&sx)[3][V],
                                        const T (&xc)[3][V], const bool (&inside)[V], T a, T be, const Cen& cv) {
        T out[3][V];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            T bq[V];
            cv.get(0, q, bq);
#pragma unroll
            for (int k = 0; k < V; ++k) {
                out[q][k] = inside[k] ? bq[k] - sx[q][k] : T(0);
                if (INIT) acc[0] += (double)out[q][k] * (double)out[q][k];
            }
        }
        if constexpr (sizeof(T) == 4 && !INIT) {
            if (half_out) { store3_bf16(reinterpret_cast<unsigned short*>(r), g, pair, i, j0, out); return; }
        }
        store3<T>(r, g, pair, i, j0, out);
        if (INIT) store3<T>(r2, g, pair, i, j0, out);
    }
    __device__ __forceinline__ void finish(int pair, bool) {
        if (INIT) {
            double a2[2] = {acc[0], acc[0]};
            block_accumulate<2>(a2, st[pair].acc + 2);
        }
    }
    __device__ __forceinline__ void flush(int pair) {
        if (INIT) {
            double a2[2] = {acc[0], acc[0]};
            warp_flush<2>(a2, st[pair].acc + 2);
            acc[0] = 0.0;
        }
    }
};

// y = S x with fused dots.  NDOT == 1: acc[0] += w1.y ;  NDOT == 2: acc[0] += y.w1, acc[1] += y.y
template <typename T, int NDOT> struct OpApplyDot {
    static constexpr bool NEED_X = true;
    static constexpr int NCEN = 1;
    static constexpr int V = VecOf<T>::V;
    Geom g; T* y; const T* cen[1]; PairState* st; double acc[NDOT];          // cen[0] = w1
    template <typename Cen>
    __device__ __forceinline__ void row(int pair, int i, int j0, const Coef<T> (&cf)[V], const T (&sx)[3][V],
                                        const T (&xc)[3][V], const bool (&inside)[V], T a, T be, const Cen& cv) {
        T out[3][V];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            T wq[V];
            cv.get(0, q, wq);
#pragma unroll
            for (int k = 0; k < V; ++k) {
                out[q][k] = inside[k] ? sx[q][k] : T(0);
                acc[0] += (double)out[q][k] * (double)wq[k];
                if (NDOT == 2) acc[1] += (double)out[q][k] * (double)out[q][k];
            }
        }
        store3<T>(y, g, pair, i, j0, out);
    }
    __device__ __forceinline__ void finish(int pair, bool) { block_accumulate<NDOT>(acc, st[pair].acc); }
    __device__ __forceinline__ void flush(int pair) { warp_flush<NDOT>(acc, st[pair].acc); }
};

// y = S x, nothing else (C-ABI operator apply in eliminated form, MG residual checks)
template <typename T> struct OpApply {
    static constexpr bool NEED_X = true;
    static constexpr int NCEN = 0;
    static constexpr int V = VecOf<T>::V;
    Geom g; T* y; const T* cen[1];
    template <typename Cen>
    __device__ __forceinline__ void row(int pair, int i, int j0, const Coef<T> (&cf)[V], const T (&sx)[3][V],
                                        const T (&xc)[3][V], const bool (&inside)[V], T a, T be, const Cen& cv) {
        T out[3][V];
#pragma unroll
        for (int q = 0; q < 3; ++q)
#pragma unroll
            for (int k = 0; k < V; ++k) out[q][k] = inside[k] ? sx[q][k] : T(0);
        store3<T>(y, g, pair, i, j0, out);
    }
    __device__ __forceinline__ void finish(int, bool) {}
    __device__ __forceinline__ void flush(int) {}
};

// Pointwise ops (no solution window): direction update fused with the first Jacobi sweep of the
// preconditioner (zero initial guess: z = w D^-1 d).
//   KIND 0:  d = p <- r + beta_k (p - omega v)          (BiCGStab direction)     cen = {r, v, p}
//   KIND 1:  d = s <- r - alpha v                        (BiCGStab half step)     cen = {r, v}
//   KIND 2:  d = b (read only)                           (first sweep on a coarse level)  cen = {b}
template <typename T, int KIND> struct OpDirJacobi {
    static constexpr bool NEED_X = false;
    static constexpr int NCEN = KIND == 0 ? 3 : (KIND == 1 ? 2 : 1);
    static constexpr int V = VecOf<T>::V;
    Geom g; const T* cen[NCEN]; T* d; T* z; T w; PairState* st; double acc[1];     // acc: KIND 1 only, s.s -> st.acc[2]
    template <typename Cen>
    __device__ __forceinline__ void row(int pair, int i, int j0, const Coef<T> (&cf)[V], const T (&sx)[3][V],
                                        const T (&xc)[3][V], const bool (&inside)[V], T a, T be, const Cen& cv) {
        T dd[3][V], zz[3][V];
        if (KIND == 2) {
#pragma unroll
            for (int q = 0; q < 3; ++q) cv.get(0, q, dd[q]);
        } else {
            const T bk = T(st[pair].beta_k), om = T(st[pair].omega), al = T(st[pair].alpha);
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                T rr[V], vv[V];
                cv.get(0, q, rr);
                cv.get(1, q, vv);
                if (KIND == 0) {
                    T pp[V];
                    cv.get(NCEN - 1, q, pp);
#pragma unroll
                    for (int k = 0; k < V; ++k) dd[q][k] = inside[k] ? rr[k] + bk * (pp[k] - om * vv[k]) : T(0);
                } else {
#pragma unroll
                    for (int k = 0; k < V; ++k) {
                        dd[q][k] = inside[k] ? rr[k] - al * vv[k] : T(0);
                        acc[0] += (double)dd[q][k] * (double)dd[q][k];
                    }
                }
            }
            store3<T>(d, g, pair, i, j0, dd);
        }
        const T inv_d2 = precond_rcp(T(-1) - T(4) * be);
#pragma unroll
        for (int k = 0; k < V; ++k) {
            T z0, z1, z2;
            block_solve<T>(cf[k], a, inv_d2, dd[0][k], dd[1][k], dd[2][k], z0, z1, z2);
            zz[0][k] = inside[k] ? w * z0 : T(0);
            zz[1][k] = inside[k] ? w * z1 : T(0);
            zz[2][k] = inside[k] ? w * z2 : T(0);
        }
        store3<T>(z, g, pair, i, j0, zz);
    }
    __device__ __forceinline__ void finish(int pair, bool) { if (KIND == 1) block_accumulate<1>(acc, st[pair].acc + 2); }
    __device__ __forceinline__ void flush(int pair) { if (KIND == 1) warp_flush<1>(acc, st[pair].acc + 2); }
};

template <typename T, bool FROM_IMAGE, typename Op>
__global__ void __launch_bounds__(128) k_sweep(LevelArgs<T> L, const T* __restrict__ x, Op op) {
    pdl_enter();
    sweep_body<T, FROM_IMAGE, Op>(L, x, op);
}

// ------------------------------------------------------------------ BLAS-1 with fused dots
// omega = (t.s)/(t.t) was computed by k_scalars; x += alpha ph + omega sh ; r = s - omega t ;
// acc[2] += r.r ; acc[3] += rhat.r
template <typename T>
__global__ void __launch_bounds__(256) k_update_x_r(long long n_per_pair, long long pair_stride,
                                                    T* __restrict__ x, const T* __restrict__ ph, const T* __restrict__ sh,
                                                    T* __restrict__ r, const T* __restrict__ s, const T* __restrict__ t,
                                                    const T* __restrict__ rhat, PairState* __restrict__ st) {
    pdl_enter();
    const int pair = blockIdx.y;
    double acc[2] = {0.0, 0.0};
    if (st[pair].half_pending) {
        // the pair converged after the first half step: x += alpha ph, nothing else changes
        const T al = T(st[pair].alpha);
        const long long base = (long long)pair * pair_stride;
        for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n_per_pair; e += (long long)gridDim.x * blockDim.x)
            x[base + e] = x[base + e] + al * ph[base + e];
    } else if (st[pair].active) {
        const T al = T(st[pair].alpha), om = T(st[pair].omega);
        const long long base = (long long)pair * pair_stride;
        for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n_per_pair;
             e += (long long)gridDim.x * blockDim.x) {
            const long long k = base + e;
            x[k] = x[k] + al * ph[k] + om * sh[k];
            const T rn = s[k] - om * t[k];
            r[k] = rn;
            acc[0] += (double)rn * (double)rn;
            acc[1] += (double)rhat[k] * (double)rn;
        }
    }
    block_accumulate<2>(acc, st[pair].acc + 2);
}

// ------------------------------------------------------------------ scalar stages (one thread per pair)
// stage 0: after the initial residual   (acc[2] = r.r, acc[3] = rhat.r = r.r)
// stage 1: after v = S ph               (acc[0] = rhat.v)            -> alpha
// stage 15: after s = r - alpha v      (acc[2] = s.s)               -> half-step convergence exit
// stage 2: after t = S sh               (acc[0] = t.s, acc[1] = t.t) -> omega
// stage 3: after the x/r update         (acc[2] = r.r, acc[3] = rhat.r) -> convergence test, beta_k
__global__ void k_scalars(PairState* __restrict__ st, int* __restrict__ active_arr, int B, int stage, double rtol, int max_it,
                          int* __restrict__ n_active) {
    pdl_enter();
    const int pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= B) return;
    PairState s = st[pair];
    if (stage == 3 && s.half_pending) { s.half_pending = 0; st[pair] = s; }      // k_update_x_r has applied the half step
    if (!s.active) return;
    const double tiny = 1e-300;
    if (stage == 0) {
        s.relres = (s.bnorm > 0.0) ? sqrt(s.acc[2]) / s.bnorm : 0.0;
        s.rho = s.acc[3]; s.rho_old = 1.0; s.alpha = 1.0; s.omega = 1.0; s.beta_k = 0.0; s.iters = 0;
        if (s.relres <= rtol || s.bnorm == 0.0) { s.active = 0; s.status = ST_CONVERGED; }
    } else if (stage == 1) {
        if (!(fabs(s.acc[0]) > tiny) || !isfinite(s.acc[0])) { s.active = 0; s.status = ST_BREAKDOWN; }
        else s.alpha = s.rho / s.acc[0];
    } else if (stage == 15) {
        // half-step exit: s = r - alpha v is the residual of x + alpha ph (acc[2] = s.s)
        const double rel = sqrt(s.acc[2]) / s.bnorm;
        if (rel <= rtol) { s.iters += 1; s.relres = rel; s.active = 0; s.status = ST_CONVERGED; s.half_pending = 1; }
    } else if (stage == 2) {
        if (!(s.acc[1] > tiny) || !isfinite(s.acc[0]) || !isfinite(s.acc[1])) { s.active = 0; s.status = ST_BREAKDOWN; }
        else s.omega = s.acc[0] / s.acc[1];
    } else {
        s.iters += 1;
        s.relres = sqrt(s.acc[2]) / s.bnorm;
        if (!isfinite(s.relres)) { s.active = 0; s.status = ST_BREAKDOWN; }
        else if (s.relres <= rtol) { s.active = 0; s.status = ST_CONVERGED; }
        else if (s.iters >= max_it) { s.active = 0; s.status = ST_MAXIT; }
        else {
            const double rho_new = s.acc[3];
            if (!(fabs(s.omega) > tiny) || !(fabs(s.rho) > tiny)) { s.active = 0; s.status = ST_BREAKDOWN; }
            else { s.beta_k = (rho_new / s.rho) * (s.alpha / s.omega); s.rho_old = s.rho; s.rho = rho_new; }
        }
    }
    s.acc[0] = s.acc[1] = s.acc[2] = s.acc[3] = 0.0;
    st[pair] = s;
    active_arr[pair] = s.active;
    if (n_active != nullptr && s.active) atomicAdd(n_active, 1);
}

// After the loop: judge every pair on the TRUE residual of its x (acc[0] = |b - A x|^2, acc[1] = |D x|^2 from
// k_true_residual).  A pair passes when |b - A x| <= rtol |b| + 8 u |D x| - the normwise backward-error form of the
// reference's test (:1120, :1150-1151): x and the iterates are stored in a type with unit round-off u, which leaves
// ~2 u |A| |x| ~ 8 u |D x| in the residual whatever the solver does (float32, alpha = 1e5: a few 1e-5 |b|), so the
// recurrence residual can reach rtol while the true one cannot.  A pair that fails although its recurrence converged, or whose recurrence broke
// down, is re-activated (at most MAX_RESTARTS times): the loop restarts BiCGStab from x with r := b - A x.
constexpr int MAX_RESTARTS = 2;
__global__ void k_verify(PairState* __restrict__ st, int* __restrict__ active_arr, int B, double rtol, int max_it, double unit_roundoff,
                         int* __restrict__ n_reactivated) {
    const int pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= B) return;
    PairState s = st[pair];
    s.true_relres = (s.bnorm > 0.0) ? sqrt(s.acc[0]) / s.bnorm : 0.0;
    s.floor_relres = (s.bnorm > 0.0) ? unit_roundoff * sqrt(s.acc[1]) / s.bnorm : 0.0;
    const bool ok = isfinite(s.true_relres) && s.true_relres <= rtol + 8.0 * s.floor_relres;
    s.active = 0;
    if (ok) s.status = ST_CONVERGED;
    else if (s.status == ST_CONVERGED || s.status == ST_BREAKDOWN) {
        if (s.restarts < MAX_RESTARTS && s.iters < max_it && isfinite(s.true_relres)) {
            s.restarts += 1; s.active = 1; s.status = ST_RUNNING; s.half_pending = 0;
        } else if (s.status == ST_CONVERGED) s.status = ST_MAXIT;
    }
    s.acc[0] = s.acc[1] = s.acc[2] = s.acc[3] = 0.0;
    st[pair] = s;
    active_arr[pair] = s.active;
    if (s.active) atomicAdd(n_reactivated, 1);
}
// restart of the re-activated pairs: r = rhat = b - S x was just recomputed (acc[2] = acc[3] = r.r)
__global__ void k_restart(PairState* __restrict__ st, int B) {
    const int pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= B) return;
    PairState s = st[pair];
    if (!s.active) return;
    s.relres = sqrt(s.acc[2]) / s.bnorm;
    s.rho = s.acc[3]; s.rho_old = 1.0; s.alpha = 1.0; s.omega = 1.0; s.beta_k = 0.0;
    s.acc[0] = s.acc[1] = s.acc[2] = s.acc[3] = 0.0;
    st[pair] = s;
}

}  // namespace ofb
