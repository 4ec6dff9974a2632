// ofb200 - pre/post kernels: Gaussian blur, right-hand side, full (non-eliminated) operator apply,
// ghost-ring completion, epilogue (boundary overwrite, scaling, speed, functionals).
#pragma once
#include "transfer.cuh"
#include "solver_kernels.cuh"

namespace ofb {

// ------------------------------------------------------------------ K1 separable Gaussian
// skimage.filters.gaussian(frame, sigma, preserve_range=True) as called at optical_flow.py:303 ==
// scipy.ndimage.gaussian_filter(mode='nearest', truncate=4.0): correlate1d along axis 0, then axis 1,
// radius int(4 sigma + 0.5), clamp-to-edge.  One launch per axis; weights (double) from the host.
template <typename T, int AXIS>
__global__ void k_blur_axis(const T* __restrict__ in, T* __restrict__ out, int Nx, int Ny, long long frame_stride,
                            const double* __restrict__ weights, int radius) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    if (j >= Ny) return;
    const T* f = in + (long long)blockIdx.z * frame_stride;
    T acc = T(0);
    for (int k = -radius; k <= radius; ++k) {
        const T w = T(weights[k + radius]);
        if (AXIS == 0) acc += w * f[(long long)min(max(i + k, 0), Nx - 1) * Ny + j];
        else           acc += w * f[(long long)i * Ny + min(max(j + k, 0), Ny - 1)];
    }
    out[(long long)blockIdx.z * frame_stride + (long long)i * Ny + j] = acc;
}

// Tiled two-pass variant (the path used for radius <= 64): one CTA blurs a 128-column x TH-row tile of one frame.
// Pass 1: each warp stages one input row segment (+radius halo columns, clamp-to-edge) in shared memory with
// coalesced loads and writes its horizontally blurred row into the tile buffer H[(TH + 2 radius)][128]
// (lane-strided columns: conflict-free).  Pass 2: every thread combines 2 radius + 1 rows of H for four adjacent
// columns with 16-byte shared loads and stores 16 bytes.  The frame is read ~(TH + 2r)/TH times (L2 absorbs the
// halo rows) and written once; nothing goes to global memory between the passes.
constexpr int BLUR_BW = 128;        // tile width in columns
constexpr int BLUR_WARPS = 8;
template <typename T>
__global__ void __launch_bounds__(BLUR_WARPS * 32) k_blur_tile(const T* __restrict__ in, T* __restrict__ out, int Nx, int Ny,
                                                              long long frame_stride, const double* __restrict__ weights,
                                                              int radius, int TH, int vec_ok) {
    extern __shared__ __align__(16) unsigned char blur_smem[];
    const int taps = 2 * radius + 1;
    const int hrows = TH + 2 * radius;
    T* H = reinterpret_cast<T*>(blur_smem);                       // [hrows][BLUR_BW]
    T* rowbuf = H + (size_t)hrows * BLUR_BW;                      // [BLUR_WARPS][BLUR_BW + 2 radius]
    T* w = rowbuf + (size_t)BLUR_WARPS * (BLUR_BW + 2 * radius);  // [taps]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c0 = blockIdx.x * BLUR_BW, i0 = blockIdx.y * TH;
    const T* f = in + (long long)blockIdx.z * frame_stride;
    for (int k = threadIdx.x; k < taps; k += blockDim.x) w[k] = T(weights[k]);
    __syncthreads();
    // ---- pass 1: horizontal
    T* rb = rowbuf + (size_t)warp * (BLUR_BW + 2 * radius);
    for (int rr = warp; rr < hrows; rr += BLUR_WARPS) {
        const int src_row = min(max(i0 - radius + rr, 0), Nx - 1);
        const T* srow = f + (long long)src_row * Ny;
        for (int e = lane; e < BLUR_BW + 2 * radius; e += 32) rb[e] = srow[min(max(c0 - radius + e, 0), Ny - 1)];
        __syncwarp();
#pragma unroll
        for (int q = 0; q < BLUR_BW / 32; ++q) {
            const int col = lane + 32 * q;
            T acc = T(0);
            for (int k = 0; k < taps; ++k) acc += w[k] * rb[col + k];
            H[(size_t)rr * BLUR_BW + col] = acc;
        }
        __syncwarp();
    }
    __syncthreads();
    // ---- pass 2: vertical, four columns per thread
    constexpr int V4 = 4;
    const int cg = threadIdx.x % (BLUR_BW / V4), rg = threadIdx.x / (BLUR_BW / V4);
    const int rgs = (BLUR_WARPS * 32) / (BLUR_BW / V4);
    const int j = c0 + cg * V4;
    for (int r = rg; r < TH; r += rgs) {
        const int i = i0 + r;
        if (i >= Nx || j >= Ny) continue;
        T acc[V4] = {T(0), T(0), T(0), T(0)};
        for (int k = 0; k < taps; ++k) {
            const T* hp = H + (size_t)(r + k) * BLUR_BW + cg * V4;
            const T wk = w[k];
#pragma unroll
            for (int c = 0; c < V4; ++c) acc[c] += wk * hp[c];
        }
        T* op = out + (long long)blockIdx.z * frame_stride + (long long)i * Ny + j;
        if (vec_ok && j + V4 <= Ny) {
            if constexpr (sizeof(T) == 4) {
                *reinterpret_cast<float4*>(op) = make_float4(acc[0], acc[1], acc[2], acc[3]);
            } else {
                *reinterpret_cast<double2*>(op) = make_double2(acc[0], acc[1]);
                *reinterpret_cast<double2*>(op + 2) = make_double2(acc[2], acc[3]);
            }
        } else {
#pragma unroll
            for (int c = 0; c < V4; ++c) if (j + c < Ny) op[c] = acc[c];
        }
    }
}

// ------------------------------------------------------------------ K2 right-hand side
// b = (-I*Ixt, -I*Iyt, -It) on the interior, 0 on the ghost ring (optical_flow.py:815-823,:889,:938,:962).
template <typename T>
__global__ void k_rhs(Geom g, const T* __restrict__ P, const T* __restrict__ C, long long cur_pair, T* __restrict__ b) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    const int pair = blockIdx.z;
    if (j >= g.Ny) return;
    const T* p = P + (long long)pair * g.img_pair;
    const T* c = C + (long long)pair * cur_pair;
    T b0 = T(0), b1 = T(0), b2 = T(0);
    if (i >= 1 && i <= g.Nx - 2 && j >= 1 && j <= g.Ny - 2) {
        const long long o = (long long)i * g.pitch + j;
        const T I = p[o];
        const T Ixt = T(0.5) * (c[o + g.pitch] - c[o - g.pitch] - p[o + g.pitch] + p[o - g.pitch]);
        const T Iyt = T(0.5) * (c[o + 1] - c[o - 1] - p[o + 1] + p[o - 1]);
        b0 = -I * Ixt; b1 = -I * Iyt; b2 = -(c[o] - I);
    }
    T* out = b + (long long)pair * g.pair + (long long)i * g.pitch + j;
    out[0] = b0; out[g.plane] = b1; out[2 * g.plane] = b2;
}

// per-pair sum of squares -> st[pair].acc[0]
template <typename T>
__global__ void __launch_bounds__(256) k_norm2(long long n_per_pair, long long pair_stride, const T* __restrict__ v,
                                               PairState* __restrict__ st) {
    const int pair = blockIdx.y;
    double acc[1] = {0.0};
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n_per_pair; e += (long long)gridDim.x * blockDim.x) {
        const double t = (double)v[(long long)pair * pair_stride + e];
        acc[0] += t * t;
    }
    block_accumulate<1>(acc, st[pair].acc);
}

__global__ void k_state_init(PairState* __restrict__ st, int* __restrict__ active_arr, int B) {
    const int pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= B) return;
    PairState s;
    s.rho = s.rho_old = s.alpha = s.omega = 1.0; s.beta_k = 0.0;
    s.acc[0] = s.acc[1] = s.acc[2] = s.acc[3] = 0.0;
    s.bnorm = 0.0; s.relres = 0.0; s.true_relres = 0.0; s.floor_relres = 0.0;
    s.active = 1; s.iters = 0; s.status = ST_RUNNING; s.half_pending = 0; s.restarts = 0;
    st[pair] = s;
    active_arr[pair] = 1;
}
__global__ void k_state_bnorm(PairState* __restrict__ st, int B) {
    const int pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= B) return;
    st[pair].bnorm = sqrt(st[pair].acc[0]);
    st[pair].acc[0] = 0.0;
}
// relres = the true residual of the returned x (what the reference prints at :1150-1151), recurrence = the BiCGStab
// recurrence residual the iteration stopped on (the KSP's own norm, :1140-1142)
__global__ void k_state_export(const PairState* __restrict__ st, int B, int* __restrict__ iters, double* __restrict__ relres,
                               int* __restrict__ converged, int* __restrict__ n_unconverged, double* __restrict__ recurrence) {
    const int pair = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair >= B) return;
    if (iters) iters[pair] = st[pair].iters;
    if (relres) relres[pair] = st[pair].true_relres;
    if (recurrence) recurrence[pair] = st[pair].relres;
    const bool ok = st[pair].status == ST_CONVERGED;
    if (converged) converged[pair] = ok ? 1 : 0;
    if (n_unconverged && !ok) atomicAdd(n_unconverged, 1);
}

// Hand one counter to the host through mapped pinned memory.  A device->host memcpy of four bytes would queue on the
// copy engine behind whatever bulk transfer another stream has in flight (the previous chunk's result planes: 26 ms for
// 64 pairs of 1024^2) and stall the convergence poll of every iteration for that long.
__global__ void k_publish(const int* __restrict__ src, volatile int* __restrict__ mapped_host) {
    *mapped_host = *src;
    __threadfence_system();
}

// zero the ghost ring of a 3-plane vector (the eliminated system carries zeros there)
template <typename T>
__global__ void k_zero_ghosts(Geom g, T* __restrict__ x) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int pair = blockIdx.y;
    const int ring = 2 * g.Ny + 2 * g.Nx;
    if (t >= ring) return;
    int i, j;
    if (t < g.Ny) { i = 0; j = t; }
    else if (t < 2 * g.Ny) { i = g.Nx - 1; j = t - g.Ny; }
    else if (t < 2 * g.Ny + g.Nx) { i = t - 2 * g.Ny; j = 0; }
    else { i = t - 2 * g.Ny - g.Nx; j = g.Ny - 1; }
    T* p = x + (long long)pair * g.pair + (long long)i * g.pitch + j;
    p[0] = T(0); p[g.plane] = T(0); p[2 * g.plane] = T(0);
}

// complete the ghost ring so that x solves the reference's full system, boundary rows included:
// q(0,j)=q(2,j) ..., corner q(0,0) = q(2,0) + q(0,2) = 2 q(2,2)   (optical_flow.py:965-1070)
template <typename T>
__global__ void k_fill_ghosts(Geom g, T* __restrict__ x) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int pair = blockIdx.y;
    const int ring = 2 * g.Ny + 2 * g.Nx;
    if (t >= ring) return;
    int i, j;
    if (t < g.Ny) { i = 0; j = t; }
    else if (t < 2 * g.Ny) { i = g.Nx - 1; j = t - g.Ny; }
    else if (t < 2 * g.Ny + g.Nx) { i = t - 2 * g.Ny; j = 0; }
    else { i = t - 2 * g.Ny - g.Nx; j = g.Ny - 1; }
    if ((i == 0 && g.halo_lo) || (i == g.Nx - 1 && g.halo_hi)) return;      // slab halo rows come from the neighbour
    const bool gr = (i == 0) || (i == g.Nx - 1), gc = (j == 0) || (j == g.Ny - 1);
    const int mi = mirror_index(i, g.Nx), mj = mirror_index(j, g.Ny);
    T* base = x + (long long)pair * g.pair;
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        T v = base[q * g.plane + (long long)mi * g.pitch + mj];
        if (gr && gc) v = v + v;
        base[q * g.plane + (long long)i * g.pitch + j] = v;
    }
}

// ------------------------------------------------------------------ full operator y = A x (reference matrix,
// ghost-ring rows included; used by the C-ABI operator entry point and the parity tests)
template <typename T>
__global__ void k_apply_full(Geom g, const T* __restrict__ P, const T* __restrict__ x, T* __restrict__ y,
                             const double* __restrict__ alpha, const double* __restrict__ beta) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    const int pair = blockIdx.z;
    if (j >= g.Ny) return;
    const T* xb = x + (long long)pair * g.pair;
    const T* img = P + (long long)pair * g.img_pair;
    T out[3];
    const bool top = i == 0, bot = i == g.Nx - 1, lef = j == 0, rig = j == g.Ny - 1;
    if (top || bot || lef || rig) {
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const T* f = xb + q * g.plane;
            T v = f[(long long)i * g.pitch + j];
            if (top) v -= f[(long long)2 * g.pitch + j];
            if (bot) v -= f[(long long)(g.Nx - 3) * g.pitch + j];
            if (lef) v -= f[(long long)i * g.pitch + 2];
            if (rig) v -= f[(long long)i * g.pitch + g.Ny - 3];
            out[q] = v;
        }
    } else {
        Row<T> pm, pc, pp;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            pm.v[c] = img[(long long)(i - 1) * g.pitch + j - 1 + c];
            pc.v[c] = img[(long long)i * g.pitch + j - 1 + c];
            pp.v[c] = img[(long long)(i + 1) * g.pitch + j - 1 + c];
        }
        const Coef<T> c = coef_from_image<T>(pm, pc, pp, 1);
        auto acc = [&](int q, int ii, int jj) -> T { return xb[q * g.plane + (long long)ii * g.pitch + jj]; };
        apply_at<T>(c, acc, i, j, T(alpha[pair]), T(beta[pair]), out);
    }
    T* o = y + (long long)pair * g.pair + (long long)i * g.pitch + j;
    o[0] = out[0]; o[g.plane] = out[1]; o[2 * g.plane] = out[2];
}

// ------------------------------------------------------------------ true residual of a returned solution
// |b - A x|^2 -> st.acc[0] and |D x|^2 -> st.acc[1] (D = diagonal of A) over the interior rows, evaluated in fp64
// arithmetic whatever the storage type: the independent check the reference prints (:1150-1151).  x must have its ghost
// ring completed (k_fill_ghosts): the boundary rows of the reference system then hold exactly and contribute 0.
template <typename T>
__global__ void __launch_bounds__(128) k_true_residual(Geom g, const T* __restrict__ P, const T* __restrict__ x, const T* __restrict__ b,
                                                       const double* __restrict__ alpha, const double* __restrict__ beta,
                                                       PairState* __restrict__ st) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int pair = blockIdx.z;
    double acc[2] = {0.0, 0.0};
    if (j >= 1 && j <= g.Ny - 2) {
        const T* xb = x + (long long)pair * g.pair;
        const T* bb = b + (long long)pair * g.pair;
        const T* img = P + (long long)pair * g.img_pair;
        const double a = alpha[pair], be = beta[pair];
        auto at = [&](int q, int ii, int jj) -> double { return (double)xb[q * g.plane + (long long)ii * g.pitch + jj]; };
        for (int i = 1 + blockIdx.y; i <= g.Nx - 2; i += gridDim.y) {
            Row<double> pm, pc, pp;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                pm.v[c] = (double)img[(long long)(i - 1) * g.pitch + j - 1 + c];
                pc.v[c] = (double)img[(long long)i * g.pitch + j - 1 + c];
                pp.v[c] = (double)img[(long long)(i + 1) * g.pitch + j - 1 + c];
            }
            const Coef<double> c = coef_from_image<double>(pm, pc, pp, 1);
            double y[3];
            apply_at<double>(c, at, i, j, a, be, y);
            const long long o = (long long)i * g.pitch + j;
            const double r0 = (double)bb[o] - y[0], r1 = (double)bb[g.plane + o] - y[1], r2 = (double)bb[2 * g.plane + o] - y[2];
            acc[0] += r0 * r0 + r1 * r1 + r2 * r2;
            const double d0 = (c.IIxx - 2.0 * c.II - 4.0 * a) * at(0, i, j), d1 = (c.IIyy - 2.0 * c.II - 4.0 * a) * at(1, i, j);
            const double d2 = (-1.0 - 4.0 * be) * at(2, i, j);
            acc[1] += d0 * d0 + d1 * d1 + d2 * d2;
        }
    }
    block_accumulate<2>(acc, st[pair].acc);
}

// the exact solution of a pair with b = 0 (two identical frames) is x = 0, whatever the initial guess
template <typename T>
__global__ void k_zero_trivial(const PairState* __restrict__ st, long long n_per_pair, long long pair_stride, T* __restrict__ x) {
    const int pair = blockIdx.y;
    if (st[pair].bnorm != 0.0) return;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n_per_pair; e += (long long)gridDim.x * blockDim.x)
        x[(long long)pair * pair_stride + e] = T(0);
}

// ------------------------------------------------------------------ K6 epilogue
// apply_constant_boundary_condition (:1304-1316, even reflection, corners = f[2,2]-type value), the three
// functionals in pixel units with d/dy == d/dx (:1167-1183), scaling by delta_x/delta_t and speed (:1189-1191).
template <typename T>
__global__ void __launch_bounds__(128) k_epilogue(Geom g, const T* __restrict__ P, const T* __restrict__ C, long long cur_pair,
                                                  const T* __restrict__ x, T* __restrict__ vx, T* __restrict__ vy,
                                                  T* __restrict__ speed, T* __restrict__ gamma, double* __restrict__ functionals,
                                                  double* __restrict__ moments,
                                                  const double* __restrict__ alpha, const double* __restrict__ beta, double scale) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int pair = blockIdx.z;
    double acc[3] = {0.0, 0.0, 0.0};
    double mom[4] = {0.0, 0.0, 0.0, 0.0};
    // rows are strided over gridDim.y so that each block ends with ONE round of atomics
    if (j < g.Ny) for (int i = blockIdx.y; i < g.Nx; i += gridDim.y) {
        const T* xb = x + (long long)pair * g.pair;
        const bool halo = (i == 0 && g.halo_lo) || (i == g.Nx - 1 && g.halo_hi);
        const int mi = halo ? i : mirror_index(i, g.Nx), mj = mirror_index(j, g.Ny);
        const long long o = (long long)mi * g.pitch + mj;
        const T ux = xb[o], uy = xb[g.plane + o], gm = xb[2 * g.plane + o];
        const long long oo = (long long)pair * g.img_pair + (long long)i * g.pitch + j;
        vx[oo] = ux * T(scale); vy[oo] = uy * T(scale); gamma[oo] = gm;
        const T sp = T(sqrt((double)(ux * T(scale)) * (double)(ux * T(scale)) + (double)(uy * T(scale)) * (double)(uy * T(scale))));
        speed[oo] = sp;
        if (!halo) { mom[0] += (double)sp; mom[1] += (double)sp * (double)sp; mom[2] += (double)gm; mom[3] += (double)gm * (double)gm; }
        if (i >= 1 && i <= g.Nx - 2 && j >= 1 && j <= g.Ny - 2) {
            const int ip = sol_row(i + 1, g), im = sol_row(i - 1, g);
            const long long op = (long long)ip * g.pitch + mj, om = (long long)im * g.pitch + mj;
            const double dux = 0.5 * ((double)xb[op] - (double)xb[om]);
            const double duy = 0.5 * ((double)xb[g.plane + op] - (double)xb[g.plane + om]);
            const double dg = 0.5 * ((double)xb[2 * g.plane + op] - (double)xb[2 * g.plane + om]);
            const T* p = P + (long long)pair * g.img_pair;
            const T* c = C + (long long)pair * cur_pair;
            const long long q = (long long)i * g.pitch + j;
            const double I = (double)p[q];
            const double Ix = 0.5 * ((double)p[q + g.pitch] - (double)p[q - g.pitch]);
            const double It = (double)c[q] - I;
            const double e = It + (double)ux * Ix + (double)uy * Ix + I * dux + I * duy - (double)gm;
            acc[0] += e * e;
            acc[1] += alpha[pair] * 2.0 * (dux * dux + duy * duy);
            acc[2] += beta[pair] * 2.0 * dg * dg;
        }
    }
    block_accumulate<3>(acc, functionals + 3 * pair);
    if (moments != nullptr) {
        __syncthreads();
        block_accumulate<4>(mom, moments + 4 * pair);
    }
}

}  // namespace ofb
