// ofb200 - kernels for the callers either side of the hot path (SURVEY.md section 8f NEXT-3 / NEXT-4):
//   k_liu_shen_step     one Jacobi sweep of liu_shen_optical_flow_jit          (optical_flow.py:507-651)
//   k_liu_shen_finish   its scaling / speed epilogue                           (:665-672)
//   k_movie_max         np.max(movie)                                          (:328)
//   k_threshold_rows / k_threshold_cols   uint8 conversion, cv2.adaptiveThreshold MEAN_C   (:328-333)
//   k_clahe_hist / k_clahe_lut / k_clahe_interp   cv2 CLAHE on uint16 frames   (:362-370)
//   k_subtract_background                                                      (:195-198)
// All of it is HBM-bound element / byte work; nothing here is a contraction.
#pragma once
#include <type_traits>
#include "common.cuh"
#include "pipeline.cuh"

namespace ofb {

// -----------------------------------------------------------------------------------------------------
// Liu-Shen Jacobi sweep.  A CTA stages a (TH+2) x (TW+2) tile of vx, vy, P and C in shared memory (16-byte global
// loads for the body, the one-pixel halo mirrored about the edge pixel = apply_constant_boundary_condition on the
// reference's padded arrays, :1304-1316), then every thread updates four adjacent pixels per row from 16-byte
// shared loads and stores 16 bytes.  Algorithmic traffic: read vx, vy, P, C, write vx', vy' = 6 words per pixel.
// -----------------------------------------------------------------------------------------------------
template <typename T> struct LsCfg {
    static constexpr int TW = 128;                 // tile width (columns): one warp spans it with four columns per lane
    static constexpr int RPT = 4;                  // consecutive rows per thread (rolling three-row window)
    static constexpr int THREADS = 128;
    static constexpr int TH = RPT * THREADS / 32;  // tile height: 16 rows
    static constexpr int OFF = 4;                  // column of the first body element in a shared row (16-byte aligned)
    static constexpr int PITCH = TW + 8;
    static constexpr int MIN_CTAS = sizeof(T) == 4 ? 5 : 2;   // resident CTAs per SM the register budget is held to
    static constexpr size_t SMEM = (size_t)4 * (TH + 2) * PITCH * sizeof(T);
};

// frame row/column seen at logical index i in [-1, N]: the ghost is the mirror image about the edge pixel
__device__ __forceinline__ int ls_mirror(int i, int N) {
    i = i < 0 ? -i : i;
    i = i >= N ? 2 * N - 2 - i : i;
    return min(max(i, 0), N - 1);                  // rows/columns of a ragged tile beyond the ghost: any valid address
}

template <typename T> struct Vec4 { T v[4]; };

template <typename T>
__device__ __forceinline__ Vec4<T> ld4_global(const T* p) {
    Vec4<T> r;
    if constexpr (sizeof(T) == 4) {
        const float4 t = *reinterpret_cast<const float4*>(p);
        r.v[0] = t.x; r.v[1] = t.y; r.v[2] = t.z; r.v[3] = t.w;
    } else {
        const double2 a = *reinterpret_cast<const double2*>(p), b = *reinterpret_cast<const double2*>(p + 2);
        r.v[0] = a.x; r.v[1] = a.y; r.v[2] = b.x; r.v[3] = b.y;
    }
    return r;
}
template <typename T>
__device__ __forceinline__ void st4(T* p, const T (&v)[4]) {
    if constexpr (sizeof(T) == 4) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
        *reinterpret_cast<double2*>(p) = make_double2(v[0], v[1]);
        *reinterpret_cast<double2*>(p + 2) = make_double2(v[2], v[3]);
    }
}

// 16-byte asynchronous global -> shared copy (LDGSTS): the whole tile is in flight before anybody waits
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
}

// Six consecutive elements of a shared row starting one left of the lane's four body columns.  The neighbours come from
// the adjacent lanes by shuffle (a scalar shared load at a 16-byte lane stride would be a 4-way bank conflict); the halo
// columns are read by every lane as a broadcast (lane 0 / lane 31 read the halo word, the others one common word) and
// selected, so the row load is branch-free.
template <typename T>
__device__ __forceinline__ void ld6_shared(const T* row, int lane, T (&a)[6]) {
    constexpr int OFF = LsCfg<T>::OFF, TW = LsCfg<T>::TW;
    const Vec4<T> m = ld4_global<T>(row + OFF + 4 * lane);
    a[1] = m.v[0]; a[2] = m.v[1]; a[3] = m.v[2]; a[4] = m.v[3];
    const T left = __shfl_up_sync(FULL, m.v[3], 1), right = __shfl_down_sync(FULL, m.v[0], 1);
    const T hl = row[lane == 0 ? OFF - 1 : OFF], hr = row[lane == 31 ? OFF + TW : OFF];
    a[0] = lane == 0 ? hl : left;
    a[5] = lane == 31 ? hr : right;
}

// 1/d without the slow-path branch of the IEEE division: SFU seed + Newton steps (float: 1 step, ~1 ulp; double: the
// 20-bit seed squared twice, ~1 ulp).  The 2x2 determinants here are far from the denormal range.
__device__ __forceinline__ float ls_rcp(float d) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
    return fmaf(r, fmaf(-d, r, 1.0f), r);
}
__device__ __forceinline__ double ls_rcp(double d) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
    r = fma(r, fma(-d, r, 1.0), r);
    r = fma(r, fma(-d, r, 1.0), r);
    return fma(r, fma(-d, r, 1.0), r);
}

// One row of the Jacobi update for the lane's four pixels.  xu/xm/xd = rows i-1, i, i+1 of v_x (six columns: one halo
// each side), likewise v_y, P and C.  The derivative terms see the mirrored ghost values held in the windows; the
// 8-neighbour sums see zeros outside the frame (:522-538) through the masks m_up/m_dn (row) and ml/mr (per column);
// pa[k] = alpha * boundary_prefactor (8 inside, 5 on an edge, 3 in a corner, :619-629).
template <typename T>
__device__ __forceinline__ void ls_row_update(const T (&xu)[6], const T (&xm)[6], const T (&xd)[6], const T (&yu)[6], const T (&ym)[6],
                                              const T (&yd)[6], const T (&pu)[6], const T (&pm)[6], const T (&pd)[6], const T (&cu)[6],
                                              const T (&cm)[6], const T (&cd)[6], T m_up, T m_dn, const T (&ml)[4], const T (&mr)[4],
                                              const T (&pa)[4], T alpha, T (&ox)[4], T (&oy)[4]) {
    // Differences are kept unscaled (d = 2 * central difference, cross = 4 * mixed difference) and the factors folded
    // into the products; signs are folded into the final 2x2 solve (G = -F).
    const T quarter = T(0.25), two = T(2);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int c = k + 1;
        // image terms (:564-595); "x" is the row axis
        const T I = pm[c];
        const T dPx = pd[c] - pu[c], dPy = pm[c + 1] - pm[c - 1];                       // 2 Ix, 2 Iy
        const T Ixt2 = (cd[c] - cu[c]) - dPx, Iyt2 = (cm[c + 1] - cm[c - 1]) - dPy;     // 2 dIdx_t, 2 dIdy_t
        const T Ixx = fma(-two, I, pd[c] + pu[c]), Iyy = fma(-two, I, pm[c + 1] + pm[c - 1]);
        const T Ixy4 = (pd[c + 1] - pd[c - 1]) - (pu[c + 1] - pu[c - 1]);
        // velocity terms (:513-520, :598-605)
        const T dVxx = xd[c] - xu[c], dVxy = xm[c + 1] - xm[c - 1];                     // 2 dVx/dx, 2 dVx/dy
        const T dVyx = yd[c] - yu[c], dVyy = ym[c + 1] - ym[c - 1];
        const T cVx = (xd[c + 1] - xd[c - 1]) - (xu[c + 1] - xu[c - 1]);                // 4 d2Vx/dxdy
        const T cVy = (yd[c + 1] - yd[c - 1]) - (yu[c + 1] - yu[c - 1]);
        const T Vx_barx = xd[c] + xu[c];
        const T Vy_bary = ym[c + 1] + ym[c - 1];
        const T barx = m_up * (ml[k] * xu[c - 1] + xu[c] + mr[k] * xu[c + 1]) + ml[k] * xm[c - 1] + mr[k] * xm[c + 1]
                     + m_dn * (ml[k] * xd[c - 1] + xd[c] + mr[k] * xd[c + 1]);
        const T bary = m_up * (ml[k] * yu[c - 1] + yu[c] + mr[k] * yu[c + 1]) + ml[k] * ym[c - 1] + mr[k] * ym[c + 1]
                     + m_dn * (ml[k] * yd[c - 1] + yd[c] + mr[k] * yd[c + 1]);
        // G = -F (:636-646):  I/4 (2 Ixt2 + 2 dPx dVxx + dPy dVyx + dPx dVyy) + I^2 (Vx_barx + cVy/4) + alpha barx
        const T I2 = I * I, hI = I * quarter, I2q = I2 * quarter;
        const T ex = fma(dPx, dVxx, Ixt2), ey = fma(dPy, dVyy, Iyt2);
        const T sx = fma(dPy, dVyx, fma(dPx, dVyy, ex + ex));
        const T sy = fma(dPx, dVxy, fma(dPy, dVxx, ey + ey));
        const T G0 = fma(alpha, barx, fma(I2q, cVy, fma(I2, Vx_barx, hI * sx)));
        const T G1 = fma(alpha, bary, fma(I2q, cVx, fma(I2, Vy_bary, hI * sy)));
        // 2x2 matrix (:631-639) and its inverse applied to F = -G
        const T diag = fma(two, I2, pa[k]);
        const T a00 = fma(I, Ixx, -diag), a11 = fma(I, Iyy, -diag), a01 = hI * Ixy4;
        const T inv_det = ls_rcp(fma(a00, a11, -(a01 * a01)));
        ox[k] = fma(a01, G1, -(a11 * G0)) * inv_det;
        oy[k] = fma(a01, G0, -(a00 * G1)) * inv_det;
    }
}

template <typename T>
__global__ void __launch_bounds__(LsCfg<T>::THREADS, LsCfg<T>::MIN_CTAS)
k_liu_shen_step(const T* __restrict__ P, const T* __restrict__ C, const T* __restrict__ vx_in, const T* __restrict__ vy_in,
                T* __restrict__ vx_out, T* __restrict__ vy_out, int Nx, int Ny, long long plane, T alpha, int vec_ok) {
    using Cfg = LsCfg<T>;
    constexpr int TW = Cfg::TW, TH = Cfg::TH, OFF = Cfg::OFF, PITCH = Cfg::PITCH;
    constexpr int WARPS = Cfg::THREADS / 32;
    extern __shared__ __align__(16) unsigned char ls_smem[];
    T* s_vx = reinterpret_cast<T*>(ls_smem);
    T* s_vy = s_vx + (TH + 2) * PITCH;
    T* s_p = s_vy + (TH + 2) * PITCH;
    T* s_c = s_p + (TH + 2) * PITCH;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int c0 = blockIdx.x * TW, i0 = blockIdx.y * TH;
    const long long base = (long long)blockIdx.z * plane;
    const T* src[4] = {vx_in + base, vy_in + base, P + base, C + base};
    T* dst[4] = {s_vx, s_vy, s_p, s_c};

    // ---- stage the tile: warp per row, lane per four body columns (asynchronous 16-byte copies), then the halo columns
    {
        const int j = c0 + 4 * lane;
        const bool fast = vec_ok && j + 3 < Ny;
        for (int rr = warp; rr < TH + 2; rr += WARPS) {
            const long long grow = (long long)ls_mirror(i0 - 1 + rr, Nx) * Ny;
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                const T* g = src[a] + grow;
                T* s = dst[a] + rr * PITCH + OFF + 4 * lane;
                if (fast) {
                    cp_async16(s, g + j);
                    if constexpr (sizeof(T) == 8) cp_async16(s + 2, g + j + 2);
                } else {
                    Vec4<T> v;
#pragma unroll
                    for (int k = 0; k < 4; ++k) v.v[k] = g[ls_mirror(j + k, Ny)];
                    st4<T>(s, v.v);
                }
            }
        }
        const int jl = ls_mirror(c0 - 1, Ny), jr = ls_mirror(c0 + TW, Ny);
        for (int e = threadIdx.x; e < 2 * (TH + 2); e += Cfg::THREADS) {
            const int rr = e >> 1, side = e & 1;
            const long long g = (long long)ls_mirror(i0 - 1 + rr, Nx) * Ny + (side ? jr : jl);
            const int so = rr * PITCH + (side ? OFF + TW : OFF - 1);
#pragma unroll
            for (int a = 0; a < 4; ++a) dst[a][so] = src[a][g];
        }
        cp_async_wait_all();
    }
    __syncthreads();
    // ---- update: lane = four adjacent columns, warp = RPT consecutive rows with a rolling three-row window
    constexpr int RPT = Cfg::RPT;
    const int j = c0 + 4 * lane;
    const int r0 = warp * RPT;
    if (i0 + r0 >= Nx) return;                      // warp-uniform: the shuffles below need whole warps
    T wx[3][6], wy[3][6], wp[3][6], wc[3][6];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        ld6_shared<T>(s_vx + (r0 + q) * PITCH, lane, wx[q]);
        ld6_shared<T>(s_vy + (r0 + q) * PITCH, lane, wy[q]);
        ld6_shared<T>(s_p + (r0 + q) * PITCH, lane, wp[q]);
        ld6_shared<T>(s_c + (r0 + q) * PITCH, lane, wc[q]);
    }
    T ml[4], mr[4];
    bool edge_j[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        ml[k] = (j + k == 0) ? T(0) : T(1);
        mr[k] = (j + k == Ny - 1) ? T(0) : T(1);
        edge_j[k] = (j + k == 0) || (j + k == Ny - 1);
    }
#pragma unroll
    for (int t = 0; t < RPT; ++t) {
        const int u = t % 3, m = (t + 1) % 3, d = (t + 2) % 3;          // window slots of rows i-1, i, i+1
        ld6_shared<T>(s_vx + (r0 + t + 2) * PITCH, lane, wx[d]);
        ld6_shared<T>(s_vy + (r0 + t + 2) * PITCH, lane, wy[d]);
        ld6_shared<T>(s_p + (r0 + t + 2) * PITCH, lane, wp[d]);
        ld6_shared<T>(s_c + (r0 + t + 2) * PITCH, lane, wc[d]);
        const int i = i0 + r0 + t;
        if (i >= Nx || j >= Ny) continue;
        const bool edge_i = (i == 0) || (i == Nx - 1);
        T pa[4], ox[4], oy[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) pa[k] = alpha * ((edge_i && edge_j[k]) ? T(3) : ((edge_i || edge_j[k]) ? T(5) : T(8)));
        ls_row_update<T>(wx[u], wx[m], wx[d], wy[u], wy[m], wy[d], wp[u], wp[m], wp[d], wc[u], wc[m], wc[d], i == 0 ? T(0) : T(1),
                         i == Nx - 1 ? T(0) : T(1), ml, mr, pa, alpha, ox, oy);
        T* px = vx_out + base + (long long)i * Ny + j;
        T* py = vy_out + base + (long long)i * Ny + j;
        if (vec_ok && j + 3 < Ny) {
            st4<T>(px, ox);
            st4<T>(py, oy);
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (j + k < Ny) { px[k] = ox[k]; py[k] = oy[k]; }
        }
    }
}

// -----------------------------------------------------------------------------------------------------
// Streamed Liu-Shen sweep (frames whose rows are 16-byte multiples): the tile kernel above alternates a load phase
// and a long straight-line compute phase (instruction-fetch bound, 0.6 of the HBM peak).  Here a producer warp streams
// one row of vx, vy, P and C per stage global -> shared with bulk copies (TMA unit) into a ring, four consumer warps
// march down the rows with the same rolling register window in a three-row loop body, CTAs are persistent and every
// CTA gets the same number of rows of the (pair, column tile) columns laid end to end.  Ghost rows are the producer
// reading the mirrored source row, ghost columns a register patch in the lanes that own column 0 / Ny-1.
// -----------------------------------------------------------------------------------------------------
template <typename T> struct LsStream {
    static constexpr int CW = 4;                         // consumer warps
    static constexpr int TW = CW * WARP * 4;             // columns per tile: 512
    static constexpr int HALO = 16 / (int)sizeof(T);     // 16-byte column halo each side
    static constexpr int ROW = TW + 2 * HALO;
    static constexpr int STAGES = 4;
    static constexpr int THREADS = (CW + 1) * WARP;
    static constexpr int STAGE_ELEMS = 4 * ROW;
    static constexpr size_t RING_BYTES = (size_t)STAGES * STAGE_ELEMS * sizeof(T);
    static constexpr size_t SMEM = RING_BYTES + 2 * STAGES * sizeof(uint64_t);
    static constexpr int MIN_CTAS = sizeof(T) == 4 ? 3 : 2;
};

// six columns of a staged row (rowp[j] = column j) around the lane's four: neighbours by shuffle, the two columns
// beyond the warp's span as broadcast shared loads, mirror patch where the lane owns column 0 / Ny-1
template <typename T>
__device__ __forceinline__ void ls_load_window(const T* rowp, int cw0, int j0, int Ny, int lane, T (&a)[6]) {
    const Vec4<T> m = ld4_global<T>(rowp + j0);
    a[1] = m.v[0]; a[2] = m.v[1]; a[3] = m.v[2]; a[4] = m.v[3];
    const T left = __shfl_up_sync(FULL, m.v[3], 1), right = __shfl_down_sync(FULL, m.v[0], 1);
    const T hl = rowp[lane == 0 ? cw0 - 1 : cw0], hr = rowp[lane == 31 ? cw0 + 4 * WARP : cw0];
    a[0] = lane == 0 ? hl : left;
    a[5] = lane == 31 ? hr : right;
    if (j0 == 0) a[0] = a[2];                    // column -1 -> column 1
    if (j0 + 3 == Ny - 1) a[5] = a[3];           // column Ny -> column Ny-2
}

template <typename T>
__global__ void __launch_bounds__(LsStream<T>::THREADS, LsStream<T>::MIN_CTAS)
k_liu_shen_stream(const T* __restrict__ P, const T* __restrict__ C, const T* __restrict__ vx_in, const T* __restrict__ vy_in,
                  T* __restrict__ vx_out, T* __restrict__ vy_out, int Nx, int Ny, long long plane, T alpha, int n_pairs, int tiles_x) {
    using S = LsStream<T>;
    constexpr int STAGES = S::STAGES;
    extern __shared__ __align__(128) unsigned char lss_smem[];
    T* ring = reinterpret_cast<T*>(lss_smem);
    uint64_t* full = reinterpret_cast<uint64_t*>(lss_smem + S::RING_BYTES);
    uint64_t* empty = full + STAGES;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int e = threadIdx.x; e < STAGES * S::STAGE_ELEMS; e += blockDim.x) ring[e] = T(0);
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, S::CW); }
        mbar_fence_init();
    }
    fence_proxy_async();
    __syncthreads();

    // this CTA's share of the rows of all (pair, tile) columns laid end to end
    const long long total = (long long)n_pairs * tiles_x * Nx;
    long long cur = (long long)blockIdx.x * total / gridDim.x;
    const long long end = (long long)(blockIdx.x + 1) * total / gridDim.x;
    uint32_t n = 0;                                      // running stage counter, continues across items

    while (cur < end) {
        const long long col = cur / Nx;
        const int i_begin = (int)(cur - col * Nx);
        const int pair = (int)(col / tiles_x), tile = (int)(col - (long long)pair * tiles_x);
        const int i_end = (int)min((long long)Nx, (long long)i_begin + (end - cur));
        cur += i_end - i_begin;
        const int c0 = tile * S::TW;
        const int nk = i_end - i_begin + 2;
        const long long base = (long long)pair * plane;
        if (warp == S::CW) {
            // ---------------------------------------------------------------- producer: lane q streams plane q
            const int lo = max(c0 - S::HALO, 0), hi = min(c0 + S::TW + S::HALO, Ny);
            const uint32_t bytes = (uint32_t)(hi - lo) * (uint32_t)sizeof(T);
            const T* src = lane == 0 ? vx_in : (lane == 1 ? vy_in : (lane == 2 ? P : C));
            src += base + lo;
            const int dst_off = lane * S::ROW + (lo - (c0 - S::HALO));
            for (int k = 0; k < nk; ++k, ++n) {
                const int s = n % STAGES;
                mbar_wait_sleep(empty + s, ((n / STAGES) & 1u) ^ 1u, 128);
                if (lane == 0) mbar_arrive_expect_tx(full + s, 4u * bytes);
                __syncwarp();
                if (lane < 4) bulk_g2s(ring + s * S::STAGE_ELEMS + dst_off, src + (long long)ls_mirror(i_begin - 1 + k, Nx) * Ny, bytes, full + s);
            }
        } else {
            // ---------------------------------------------------------------- consumers
            const int cw0 = c0 + warp * (4 * WARP);
            const int j0 = cw0 + 4 * lane;
            const bool has_cols = cw0 < Ny;
            T wx[3][6], wy[3][6], wp[3][6], wc[3][6];
            const T ml[4] = {j0 == 0 ? T(0) : T(1), T(1), T(1), T(1)};
            const T mr[4] = {T(1), T(1), T(1), j0 + 3 == Ny - 1 ? T(0) : T(1)};
            const bool edge_j0 = j0 == 0, edge_j3 = j0 + 3 == Ny - 1;

            auto do_stage = [&](auto SLOT, int k) {
                constexpr int sp = decltype(SLOT)::value, su = (sp + 1) % 3, sm = (sp + 2) % 3;     // rows i+1, i-1, i
                const int s = n % STAGES;
                const uint32_t ph = (n / STAGES) & 1u;
                ++n;
                mbar_wait_sleep(full + s, ph, 32);
                if (has_cols) {
                    const T* stg = ring + s * S::STAGE_ELEMS - (c0 - S::HALO);       // stg[q*ROW + j] = column j of plane q
                    ls_load_window<T>(stg, cw0, j0, Ny, lane, wx[sp]);
                    ls_load_window<T>(stg + S::ROW, cw0, j0, Ny, lane, wy[sp]);
                    ls_load_window<T>(stg + 2 * S::ROW, cw0, j0, Ny, lane, wp[sp]);
                    ls_load_window<T>(stg + 3 * S::ROW, cw0, j0, Ny, lane, wc[sp]);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(empty + s);                               // the row now lives in registers
                if (has_cols && k >= 2 && j0 < Ny) {
                    const int i = i_begin + k - 2;
                    const bool edge_i = (i == 0) || (i == Nx - 1);
                    const T pa_in = alpha * (edge_i ? T(5) : T(8)), pa_edge = alpha * (edge_i ? T(3) : T(5));
                    const T pa[4] = {edge_j0 ? pa_edge : pa_in, pa_in, pa_in, edge_j3 ? pa_edge : pa_in};
                    T ox[4], oy[4];
                    ls_row_update<T>(wx[su], wx[sm], wx[sp], wy[su], wy[sm], wy[sp], wp[su], wp[sm], wp[sp], wc[su], wc[sm], wc[sp],
                                     i == 0 ? T(0) : T(1), i == Nx - 1 ? T(0) : T(1), ml, mr, pa, alpha, ox, oy);
                    st4<T>(vx_out + base + (long long)i * Ny + j0, ox);
                    st4<T>(vy_out + base + (long long)i * Ny + j0, oy);
                }
            };
            int k = 0;
            while (k < nk) {
                do_stage(std::integral_constant<int, 0>{}, k); if (++k >= nk) break;
                do_stage(std::integral_constant<int, 1>{}, k); if (++k >= nk) break;
                do_stage(std::integral_constant<int, 2>{}, k); ++k;
            }
        }
    }
}

// :665-672  v *= delta_x/delta_t, speed = sqrt(vx^2 + vy^2)
template <typename T>
__global__ void k_liu_shen_finish(T* __restrict__ vx, T* __restrict__ vy, T* __restrict__ speed, long long n, T scale) {
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x) {
        const T x = vx[k] * scale, y = vy[k] * scale;
        vx[k] = x;
        vy[k] = y;
        speed[k] = sqrt(x * x + y * y);
    }
}

// -----------------------------------------------------------------------------------------------------
// np.max over a movie (:328): block reduction, one atomic per block on the bit pattern of a double.
// -----------------------------------------------------------------------------------------------------
__global__ void k_fill_double(double* p, double v) { *p = v; }

__device__ __forceinline__ void atomic_max_double(double* addr, double v) {
    unsigned long long* a = reinterpret_cast<unsigned long long*>(addr);
    unsigned long long old = *a;
    while (__longlong_as_double((long long)old) < v) {
        const unsigned long long seen = atomicCAS(a, old, (unsigned long long)__double_as_longlong(v));
        if (seen == old) break;
        old = seen;
    }
}

template <typename T>
__global__ void __launch_bounds__(256) k_movie_max(const T* __restrict__ in, long long n, double* __restrict__ out) {
    __shared__ double part[8];
    double m = -INFINITY;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x)
        m = fmax(m, (double)in[k]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(FULL, m, o));
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) m = fmax(m, part[w]);
        atomic_max_double(out, m);
    }
}

// -----------------------------------------------------------------------------------------------------
// Adaptive threshold (:328-333): uint8 conversion, k x k normalised box mean with replicated borders, dst = src >
// mean - ceil(threshold).  The box sum is separable and O(1) per pixel: a warp turns a row of bytes into its prefix
// sums in shared memory and writes the k-wide window sums (16 bit: 255 * 255 < 65536); a thread per column then slides
// the k-tall window down a strip of rows.  The byte and window-sum planes of a 16-frame batch stay in the L2.
// -----------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ unsigned char to_u8_range(T v, T mx) {
    // np.array(movie/np.max(movie)*255.0, dtype='uint8'): divide, multiply (in the movie's float type), truncate
    if constexpr (sizeof(T) == 4) return (unsigned char)(int)__fmul_rn(__fdiv_rn(v, mx), 255.0f);
    else return (unsigned char)(int)__dmul_rn(__ddiv_rn(v, mx), 255.0);
}

// rows: u8 = uint8(movie / max * 255) (kept for the comparison), hs[i][j] = sum of u8[i][clamp(j-r .. j+r)]
constexpr int AT_ROW_WARPS = 8;
template <typename T>
__global__ void __launch_bounds__(AT_ROW_WARPS * 32)
k_threshold_rows(const T* __restrict__ in, const double* __restrict__ movie_max, unsigned char* __restrict__ u8,
                 unsigned short* __restrict__ hs, long long n_rows, int Ny, int r, int warps) {
    extern __shared__ __align__(16) unsigned char atr_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (warp >= warps) return;
    unsigned int* pre = reinterpret_cast<unsigned int*>(atr_smem) + (size_t)warp * (Ny + 1);     // pre[k] = sum of the first k bytes
    const T mx = (T)(*movie_max);
    for (long long row = (long long)blockIdx.x * warps + warp; row < n_rows; row += (long long)gridDim.x * warps) {
        const T* src = in + row * Ny;
        unsigned char* dst8 = u8 + row * Ny;
        unsigned int carry = 0;
        if (lane == 0) pre[0] = 0;
        for (int j0 = 0; j0 < Ny; j0 += 32) {
            const int j = j0 + lane;
            unsigned int v = 0;
            if (j < Ny) {
                const unsigned char b = to_u8_range<T>(src[j], mx);
                dst8[j] = b;
                v = b;
            }
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned int u = __shfl_up_sync(FULL, v, o);
                if (lane >= o) v += u;
            }
            if (j < Ny) pre[j + 1] = carry + v;
            carry += __shfl_sync(FULL, v, 31);
        }
        __syncwarp();
        const unsigned int first = pre[1], last = pre[Ny] - pre[Ny - 1];
        unsigned short* dsth = hs + row * Ny;
        for (int j = lane; j < Ny; j += 32) {
            const int lo = j - r, hi = j + r;                          // window [lo, hi], clamped columns repeat the edge byte
            unsigned int s = pre[min(hi, Ny - 1) + 1] - pre[max(lo, 0)];
            s += (unsigned int)max(-lo, 0) * first + (unsigned int)max(hi - (Ny - 1), 0) * last;
            dsth[j] = (unsigned short)s;
        }
        __syncwarp();
    }
}

// columns: slide the k-tall window of row sums down a strip of AT_STRIP rows; out = (src > mean - idelta) without the
// division:  mean = round(s / k^2) < src + idelta  <=>  2 s + k^2 < (src + idelta) * 2 k^2  (k odd: no ties)
constexpr int AT_STRIP = 64;
__global__ void __launch_bounds__(128)
k_threshold_cols(const unsigned char* __restrict__ u8, const unsigned short* __restrict__ hs, unsigned char* __restrict__ out, int Nx, int Ny,
                 long long plane, int window, int idelta) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= Ny) return;
    const int r = window >> 1, k2 = window * window;
    const int i0 = blockIdx.y * AT_STRIP, i1 = min(i0 + AT_STRIP, Nx);
    const long long base = (long long)blockIdx.z * plane + j;
    const unsigned short* h = hs + base;
    int s = 0;
    for (int a = i0 - r; a <= i0 + r; ++a) s += h[(long long)min(max(a, 0), Nx - 1) * Ny];
    for (int i = i0; i < i1; ++i) {
        if (i > i0) s += (int)h[(long long)min(i + r, Nx - 1) * Ny] - (int)h[(long long)max(i - 1 - r, 0) * Ny];
        const int t = min(max((int)u8[base + (long long)i * Ny] + idelta, 0), 256);
        out[base + (long long)i * Ny] = (2 * s + k2 < t * 2 * k2) ? 1 : 0;
    }
}

// -----------------------------------------------------------------------------------------------------
// CLAHE on uint16 frames (:362-370; OpenCV clahe.cpp, 16-bit path: 65536 bins).
// -----------------------------------------------------------------------------------------------------
constexpr int CLAHE_BINS = 65536;

struct ClaheGeom {
    int Nx, Ny;              // frame
    int ext_x, ext_y;        // frame extended (BORDER_REFLECT_101) to a multiple of the tile grid
    int tiles_x, tiles_y;    // tiles along the columns / the rows
    int tw, th;              // tile width (columns) / height (rows)
    int clip;                // clip limit in counts (0 = none)
    float lut_scale;         // 65535 / tile area
};

template <typename T>
__device__ __forceinline__ unsigned short to_u16(T v) { return (unsigned short)(int)v; }     // movie.astype(np.uint16)

template <typename T>
__global__ void __launch_bounds__(256) k_clahe_hist(const T* __restrict__ in, int* __restrict__ hist, ClaheGeom g, long long plane) {
    const T* f = in + (long long)blockIdx.z * plane;
    int* h = hist + (long long)blockIdx.z * g.tiles_x * g.tiles_y * CLAHE_BINS;
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= g.ext_y) return;
    const int si = i < g.Nx ? i : 2 * g.Nx - 2 - i, sj = j < g.Ny ? j : 2 * g.Ny - 2 - j;
    const unsigned short v = to_u16<T>(f[(long long)si * g.Ny + sj]);
    atomicAdd(h + (long long)((i / g.th) * g.tiles_x + j / g.tw) * CLAHE_BINS + v, 1);
}

// One CTA per (tile, frame): clip, redistribute, running sum, scale -> 16-bit look-up table.  Warp w owns the bins
// [2048 w, 2048 (w+1)) and reads them as 16-byte vectors.  Pass 1 sums min(h, clip) and the clipped excess per warp; the
// redistribution (clahe.cpp: +clipped/65536 everywhere, +1 on every `step`-th bin while the remainder lasts) is a closed
// form of the bin index, so the running sum at the start of a warp's segment follows from the per-warp sums alone and
// pass 2 (which re-reads the tile's 256 KB from the L2) is a warp scan per 128 bins with no block barrier.
__device__ __forceinline__ int clahe_extra(int b, int residual, int step, unsigned magic) {
    // 1 if bin b receives one of the `residual` leftover counts: b = q * step with q < residual
    if (residual == 0) return 0;
    if (step == 1) return b < residual ? 1 : 0;
    unsigned q = __umulhi((unsigned)b, magic);          // b / step for b < 65536 (magic = ceil(2^32 / step)), fixed up below
    if (q * (unsigned)step > (unsigned)b) --q;
    return (q * (unsigned)step == (unsigned)b && (int)q < residual) ? 1 : 0;
}

__global__ void __launch_bounds__(1024) k_clahe_lut(const int* __restrict__ hist, unsigned short* __restrict__ lut, ClaheGeom g) {
    __shared__ int s_min[32], s_clip[32];
    const long long t = (long long)blockIdx.y * gridDim.x + blockIdx.x;
    const int4* h4 = reinterpret_cast<const int4*>(hist + t * CLAHE_BINS);
    uint2* l4 = reinterpret_cast<uint2*>(lut + t * CLAHE_BINS);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int SEG = CLAHE_BINS / 32;                 // bins per warp
    const int seg = warp * SEG;
    const int clip = g.clip > 0 ? g.clip : 0x7fffffff;
    int sum_min = 0, sum_clip = 0;
#pragma unroll 4
    for (int tt = 0; tt < SEG / 128; ++tt) {
        const int4 v = h4[(seg >> 2) + tt * 32 + lane];
        const int m0 = min(v.x, clip), m1 = min(v.y, clip), m2 = min(v.z, clip), m3 = min(v.w, clip);
        sum_min += (m0 + m1) + (m2 + m3);
        sum_clip += ((v.x - m0) + (v.y - m1)) + ((v.z - m2) + (v.w - m3));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum_min += __shfl_xor_sync(FULL, sum_min, o);
        sum_clip += __shfl_xor_sync(FULL, sum_clip, o);
    }
    if (lane == 0) { s_min[warp] = sum_min; s_clip[warp] = sum_clip; }
    __syncthreads();
    int clipped = 0, before = 0;                         // before = sum of min(h, clip) over the bins below this warp's segment
    for (int w = 0; w < 32; ++w) {
        clipped += s_clip[w];
        before += w < warp ? s_min[w] : 0;
    }
    const int batch = clipped / CLAHE_BINS;
    const int residual = clipped - batch * CLAHE_BINS;
    const int step = residual != 0 ? max(CLAHE_BINS / residual, 1) : 1;
    const unsigned magic = step > 1 ? (unsigned)(((1ull << 32) + (unsigned)step - 1) / (unsigned)step) : 0u;
    // leftover counts handed to bins below `seg`: multiples of step in [0, seg), at most `residual` of them
    const int extra_before = residual != 0 ? min((seg + step - 1) / step, residual) : 0;
    int carry = before + batch * seg + extra_before;
#pragma unroll 2
    for (int tt = 0; tt < SEG / 128; ++tt) {
        const int idx = (seg >> 2) + tt * 32 + lane;
        const int4 v = h4[idx];
        const int b = idx * 4;
        const int p0 = min(v.x, clip) + batch + clahe_extra(b, residual, step, magic);
        const int p1 = p0 + min(v.y, clip) + batch + clahe_extra(b + 1, residual, step, magic);
        const int p2 = p1 + min(v.z, clip) + batch + clahe_extra(b + 2, residual, step, magic);
        const int p3 = p2 + min(v.w, clip) + batch + clahe_extra(b + 3, residual, step, magic);
        int incl = p3;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(FULL, incl, o);
            if (lane >= o) incl += u;
        }
        const int base = carry + incl - p3;
        // lut[i] = saturate_cast<ushort>(sum * lutScale): float product, round half to even
        auto q = [&](int sum) { return (unsigned)min(max(__float2int_rn(__fmul_rn(__int2float_rn(sum), g.lut_scale)), 0), 65535); };
        const unsigned q0 = q(base + p0), q1 = q(base + p1), q2 = q(base + p2), q3 = q(base + p3);
        l4[idx] = make_uint2(q0 | (q1 << 16), q2 | (q3 << 16));
        carry += __shfl_sync(FULL, incl, 31);
    }
}

template <typename T, typename TO>
__global__ void __launch_bounds__(256) k_clahe_interp(const T* __restrict__ in, const unsigned short* __restrict__ lut,
                                                      TO* __restrict__ out, ClaheGeom g, long long plane) {
    const T* f = in + (long long)blockIdx.z * plane;
    const unsigned short* l = lut + (long long)blockIdx.z * g.tiles_x * g.tiles_y * CLAHE_BINS;
    const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
    if (j >= g.Ny) return;
    const float inv_tw = __fdiv_rn(1.0f, (float)g.tw), inv_th = __fdiv_rn(1.0f, (float)g.th);
    const float txf = __fsub_rn(__fmul_rn((float)j, inv_tw), 0.5f), tyf = __fsub_rn(__fmul_rn((float)i, inv_th), 0.5f);
    int tx1 = (int)floorf(txf), ty1 = (int)floorf(tyf);
    const float xa = __fsub_rn(txf, (float)tx1), ya = __fsub_rn(tyf, (float)ty1);
    const float xa1 = __fsub_rn(1.0f, xa), ya1 = __fsub_rn(1.0f, ya);
    const int tx2 = min(tx1 + 1, g.tiles_x - 1), ty2 = min(ty1 + 1, g.tiles_y - 1);
    tx1 = max(tx1, 0);
    ty1 = max(ty1, 0);
    const int v = to_u16<T>(f[(long long)i * g.Ny + j]);
    const float l11 = l[(long long)(ty1 * g.tiles_x + tx1) * CLAHE_BINS + v], l12 = l[(long long)(ty1 * g.tiles_x + tx2) * CLAHE_BINS + v];
    const float l21 = l[(long long)(ty2 * g.tiles_x + tx1) * CLAHE_BINS + v], l22 = l[(long long)(ty2 * g.tiles_x + tx2) * CLAHE_BINS + v];
    // (l11*xa1 + l12*xa)*ya1 + (l21*xa1 + l22*xa)*ya with every product and sum rounded (no contraction)
    const float top = __fadd_rn(__fmul_rn(l11, xa1), __fmul_rn(l12, xa)), bot = __fadd_rn(__fmul_rn(l21, xa1), __fmul_rn(l22, xa));
    const float res = __fadd_rn(__fmul_rn(top, ya1), __fmul_rn(bot, ya));
    const int q = min(max(__float2int_rn(res), 0), 65535);
    out[(long long)blockIdx.z * plane + (long long)i * g.Ny + j] = (TO)q;
}

// -----------------------------------------------------------------------------------------------------
// float32 -> float64 widening of result planes on the device (the reference returns float64 arrays, :1193-1205): the
// planes then travel device -> pinned result memory as they are and the host cores do no conversion.
// -----------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_widen(const float* __restrict__ in, double* __restrict__ out, long long n) {
    const long long n4 = n >> 2;
    const float4* in4 = reinterpret_cast<const float4*>(in);
    double2* out2 = reinterpret_cast<double2*>(out);
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n4; k += (long long)gridDim.x * blockDim.x) {
        const float4 v = in4[k];
        out2[2 * k] = make_double2((double)v.x, (double)v.y);
        out2[2 * k + 1] = make_double2((double)v.z, (double)v.w);
    }
    for (long long k = 4 * n4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x)
        out[k] = (double)in[k];
}

// -----------------------------------------------------------------------------------------------------
// background subtraction (:195-198): out = movie - background where blur(movie, 10) > background, else 0
// -----------------------------------------------------------------------------------------------------
template <typename T>
__global__ void k_subtract_background(const T* __restrict__ movie, const T* __restrict__ blurred, T* __restrict__ out, long long n,
                                      T background) {
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += (long long)gridDim.x * blockDim.x)
        out[k] = blurred[k] > background ? movie[k] - background : T(0);
}

}  // namespace ofb
