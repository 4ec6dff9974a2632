// ofb200 - the row-marching stencil sweep shared by every operator-type kernel.
//
// One warp = one work item = (row strip of R rows) x (32*V columns).  The warp walks down the strip
// with a rolling 3-row register window of the three solution planes (and of the image on level 0);
// each global row is read once per strip (+2 halo rows per strip), every access is a coalesced
// 16-byte vector per lane, and left/right neighbours are exchanged with warp shuffles.
#pragma once
#include "stencil.cuh"

namespace ofb {

// Arguments common to all sweeps over one level of a batch of pairs.
template <typename T> struct LevelArgs {
    Geom g;
    const T* coef;          // level 0: previous-frame image planes [B][Nx*pitch]; level>0: [B][7] coefficient planes
    long long coef_pair;    // elements between pairs in `coef`
    const double* alpha;    // per-pair speed_alpha  (B)
    const double* beta;     // per-pair remodelling_alpha (B)
    const int* active;      // per-pair activity flag (nullptr = all active)
    double h2;              // hinv^2 of the level
    int rows_per_strip;
    int tiles_x;            // warp tiles per row
    int strips;             // row strips per pair
    int tile_fastest;       // streamed sweep: consecutive work items are neighbouring column tiles of one strip
    int n_pairs;            // streamed sweep: batch size
    int balanced;           // streamed sweep: equal row ranges per CTA instead of strips (WorkWalk in stream_sweep.cuh)
    // streamed sweeps on coarse levels (float only): the 7 coefficient planes as bfloat16, [B][7][Nx][pitch16]; nullptr = read `coef`.
    // Halves the coefficient traffic of a coarse sweep (16 -> 12.5 words per pixel); a preconditioner does not notice the 8-bit
    // mantissa (numpy study: identical iteration counts).
    const unsigned short* coef16;
    int pitch16;            // elements per row, a multiple of 8 (16-byte rows for the bulk copies)
    long long plane16, pair16;
};

template <typename T, bool FROM_IMAGE> struct CoefRows {
    Row<T> p[3];            // image rows (FROM_IMAGE)
};

// plain stores of V values into three planes of a vector at row i
template <typename T>
__device__ __forceinline__ void store3(T* __restrict__ base, const Geom& g, int pair, int i, int j0,
                                       const T (&v)[3][VecOf<T>::V]) {
    T* p = base + (long long)pair * g.pair + (long long)i * g.pitch;
#pragma unroll
    for (int q = 0; q < 3; ++q) store_own<T>(p + q * g.plane, j0, g.Ny, g.aligned != 0, v[q]);
}
// the same values as bfloat16 (round to nearest even) into planes of two-byte elements with the layout of `g`
__device__ __forceinline__ unsigned short to_bf16(float v) {
    unsigned u = __float_as_uint(v);
    u += 0x7fffu + ((u >> 16) & 1u);
    return (unsigned short)(u >> 16);
}
__device__ __forceinline__ void store3_bf16(unsigned short* __restrict__ base, const Geom& g, int pair, int i, int j0, const float (&v)[3][4]) {
    unsigned short* p = base + (long long)pair * g.pair + (long long)i * g.pitch;
#pragma unroll
    for (int q = 0; q < 3; ++q) {
        unsigned short* row = p + q * g.plane;
        if (g.aligned != 0 && j0 + 4 <= g.Ny) {
            uint2 t;        // one cvt per pair of values: the upper half of the result is the first operand
            asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(t.x) : "f"(v[q][1]), "f"(v[q][0]));
            asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(t.y) : "f"(v[q][3]), "f"(v[q][2]));
            *reinterpret_cast<uint2*>(row + j0) = t;
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (j0 + k < g.Ny) row[j0 + k] = to_bf16(v[q][k]);
        }
    }
}

template <typename T>
__device__ __forceinline__ void load3(const T* __restrict__ base, const Geom& g, int pair, int i, int j0,
                                      T (&v)[3][VecOf<T>::V]) {
    const T* p = base + (long long)pair * g.pair + (long long)i * g.pitch;
#pragma unroll
    for (int q = 0; q < 3; ++q) load_own<T>(p + q * g.plane, j0, g.Ny, g.aligned != 0, v[q]);
}

// Centre-row inputs of an op (NCEN three-plane vectors read at the row's own pixels only) are fetched
// through an accessor at the point of use, so that they never sit in registers across the stencil
// arithmetic: from global memory in the register sweep, from the staged row in the streamed sweep.
template <typename T> struct CenGlobal {
    const T* const* cen; const Geom* g; int pair, i, j0;
    __device__ __forceinline__ void get(int n, int q, T (&v)[VecOf<T>::V]) const {
        load_own<T>(cen[n] + (long long)pair * g->pair + (long long)q * g->plane + (long long)i * g->pitch, j0, g->Ny,
                    g->aligned != 0, v);
    }
};
template <typename T, int ROW> struct CenShared {
    const T* base;          // staged centre row 0 at this lane's first column
    __device__ __forceinline__ void get(int n, int q, T (&v)[VecOf<T>::V]) const {
        using VT = typename VecOf<T>::type;
        const VT t = *reinterpret_cast<const VT*>(base + (3 * n + q) * ROW);
        const T* p = reinterpret_cast<const T*>(&t);
#pragma unroll
        for (int k = 0; k < VecOf<T>::V; ++k) v[k] = p[k];
    }
};

// Op interface (all members __device__):
//   static constexpr bool NEED_X;          // needs the solution window (false: pointwise ops)
//   static constexpr int NCEN;             // number of centre-only vector inputs
//   const T* cen[max(NCEN,1)];             // their base pointers ([B][3][plane] layout)
//   void row(int pair, int i, int j0, const Coef<T> (&cf)[V], const T (&sx)[3][V],
//            const T (&xc)[3][V], const bool (&inside)[V], T a, T be, const Cen& cv);   // cv.get(n, q, v)
//   void finish(int pair, bool has_work);  // k_sweep: block-level reductions (called by all threads)
//   void flush(int pair);                  // k_stream: warp-level reductions + reset (called by whole warps)
template <typename T, bool FROM_IMAGE, typename Op>
__device__ __forceinline__ void sweep_body(const LevelArgs<T>& L, const T* __restrict__ x, Op& op) {
    constexpr int V = VecOf<T>::V;
    const Geom& g = L.g;
    const int lane = threadIdx.x & 31;
    const int warp_in_block = threadIdx.x >> 5;
    const int warps_per_block = blockDim.x >> 5;
    const int pair = blockIdx.y;
    const bool pair_active = (L.active == nullptr) || (L.active[pair] != 0);
    const int item = blockIdx.x * warps_per_block + warp_in_block;
    const bool has_work = pair_active && item < L.tiles_x * L.strips;
    if (has_work) {
        const int strip = item / L.tiles_x, tile = item - strip * L.tiles_x;
        const int c0 = tile * (WARP * V);
        const int j0 = c0 + lane * V;
        const int i_begin = strip * L.rows_per_strip;
        const int i_end = min(i_begin + L.rows_per_strip, g.Nx);
        const bool edge_warp = (c0 == 0) || (c0 + WARP * V + 1 >= g.Ny - 1);
        const T a = T(L.alpha[pair] * L.h2), be = T(L.beta[pair] * L.h2);
        const T* xb = Op::NEED_X ? x + (long long)pair * g.pair : nullptr;
        const T* cb = L.coef + (long long)pair * L.coef_pair;

        Row<T> ux[3], uy[3], gg[3], pw[3];
        auto load_x = [&](int slot, int i) {
            if constexpr (Op::NEED_X) {
                load_row<T, LOAD_SOL>(ux[slot], xb, i, j0, g, lane, edge_warp);
                load_row<T, LOAD_SOL>(uy[slot], xb + g.plane, i, j0, g, lane, edge_warp);
                load_row<T, LOAD_SOL>(gg[slot], xb + 2 * g.plane, i, j0, g, lane, edge_warp);
            }
            if constexpr (FROM_IMAGE) load_row<T, LOAD_RAW>(pw[slot], cb, i, j0, g, lane, false);
        };
        auto step = [&](int sm, int sc, int sp, int i) {
            // rows i-1,i,i+1 are in slots sm,sc,sp
            Coef<T> cf[V];
            bool inside[V];
            T sx[3][V], xc[3][V];
            const bool row_inside = (i >= 1) && (i <= g.Nx - 2);
            if constexpr (!FROM_IMAGE) {
                T cr[NCOEF][V];
                const T* rowp = cb + (long long)i * g.pitch;
#pragma unroll
                for (int q = 0; q < NCOEF; ++q) load_own<T>(rowp + q * g.plane, j0, g.Ny, g.aligned != 0, cr[q]);
#pragma unroll
                for (int k = 0; k < V; ++k) {
                    cf[k].II = cr[0][k]; cf[k].IIx = cr[1][k]; cf[k].IIxx = cr[2][k]; cf[k].IIyy = cr[3][k];
                    cf[k].IIxy = cr[4][k]; cf[k].Ih = cr[5][k]; cf[k].Ix = cr[6][k];
                }
            }
#pragma unroll
            for (int k = 0; k < V; ++k) {
                const int j = j0 + k;
                inside[k] = row_inside && (j >= 1) && (j <= g.Ny - 2);
                if constexpr (FROM_IMAGE) cf[k] = coef_from_image<T>(pw[sm], pw[sc], pw[sp], k + 1);
                if constexpr (Op::NEED_X) {
                    Win<T> wx{ux[sm], ux[sc], ux[sp]}, wy{uy[sm], uy[sc], uy[sp]}, wg{gg[sm], gg[sc], gg[sp]};
                    apply_point<T>(cf[k], wx, wy, wg, k + 1, a, be, sx[0][k], sx[1][k], sx[2][k]);
                    xc[0][k] = ux[sc].v[k + 1]; xc[1][k] = uy[sc].v[k + 1]; xc[2][k] = gg[sc].v[k + 1];
                } else {
                    sx[0][k] = sx[1][k] = sx[2][k] = T(0);
                    xc[0][k] = xc[1][k] = xc[2][k] = T(0);
                }
            }
            const CenGlobal<T> cv{op.cen, &g, pair, i, j0};
            op.row(pair, i, j0, cf, sx, xc, inside, a, be, cv);
        };
        load_x(0, i_begin - 1);
        load_x(1, i_begin);
        int i = i_begin;
        while (i < i_end) {
            load_x(2, i + 1); step(0, 1, 2, i); if (++i >= i_end) break;
            load_x(0, i + 1); step(1, 2, 0, i); if (++i >= i_end) break;
            load_x(1, i + 1); step(2, 0, 1, i); ++i;
        }
    }
    op.finish(pair, has_work);
}

}  // namespace ofb
