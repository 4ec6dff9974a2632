// ofb200 - streamed stencil sweep: the large-level variant of the row-marching sweep (sweep.cuh).
//
// One CTA = 4 consumer warps + 1 producer warp, persistent over work items (column tile x row strip x pair).
// The producer warp streams every input row of the item (solution planes, image or coefficient planes,
// centre-only vectors) global -> shared with 1-D bulk copies (cp.async.bulk, TMA unit; SASS UBLKCP) into a
// STAGES-deep ring; completion is signalled on a "full" mbarrier by transaction bytes.  Consumer warps pull a
// staged row into their rolling 3-row register window, release the stage on an "empty" mbarrier, and compute.
// The copy engine therefore runs several rows ahead of the arithmetic independent of register pressure,
// which is what the register-only sweep could not do (16 warps/SM x 7 loads in flight: latency bound).
// The ring continues across work items, so there is no pipeline drain between items.
//
// Requirements (checked on the host, otherwise the register sweep is used): 16-byte aligned plane bases,
// pitch % V == 0.
#pragma once
#include <type_traits>
#include "pipeline.cuh"
#include "solver_kernels.cuh"
#include "transfer.cuh"

namespace ofb {

template <typename T> struct StreamCfg {
    static constexpr int V = VecOf<T>::V;
    static constexpr int CW = 4;                         // consumer warps per CTA
    static constexpr int TW = CW * WARP * V;             // columns per tile: 512 (f32) / 256 (f64)
    static constexpr int HALO = 16 / (int)sizeof(T);     // 16-byte aligned column halo
    static constexpr int ROW = TW + 2 * HALO;            // elements per staged row (2080 bytes)
    static constexpr int STAGES = 4;
    static constexpr int THREADS = (CW + 1) * WARP;
};

// Ops with `static constexpr bool PROLONG = true` (OpSmoothProlong) get the coarse-grid correction added to the solution
// window as it is loaded: 3 planes x 2 coarse rows are staged with every window row.
template <typename Op, typename = void> struct OpProlong : std::false_type {};
template <typename Op> struct OpProlong<Op, std::void_t<decltype(Op::PROLONG)>> : std::integral_constant<bool, Op::PROLONG> {};

#ifndef OFB_PROLONG_STAGES
#define OFB_PROLONG_STAGES 3
#endif

template <typename T, bool FROM_IMAGE, typename Op> struct StreamLayout {
    using C = StreamCfg<T>;
    static constexpr int NWIN = (Op::NEED_X ? 3 : 0) + (FROM_IMAGE ? 1 : 0);      // rows staged one row ahead
    static constexpr int NCOR = OpProlong<Op>::value ? 6 : 0;                     // coarse correction rows (staged with them)
    static constexpr int COR_ELEMS = C::TW / 2 + 4;                               // coarse columns a tile interpolates from
    static constexpr int NCORSLOTS = (NCOR * COR_ELEMS + C::ROW - 1) / C::ROW;    // the six half-width rows are packed: 3 row slots
    static constexpr int NAHEAD = NWIN + NCORSLOTS;                               // row slots staged one row ahead
    static constexpr int NCENROWS = 3 * Op::NCEN + (FROM_IMAGE ? 0 : NCOEF);      // rows staged at the centre row
    static constexpr int NROWS = NAHEAD + NCENROWS;                               // row slots per stage
    static constexpr int NLANES = NWIN + NCOR + NCENROWS;                         // producer lanes (one per copied row)
    static_assert(NLANES <= WARP, "one producer lane per staged row");
    // three CTAs per SM: the correction rows bring a stage to 10 row slots, which fit three deep
    static constexpr int STAGES = NCOR > 0 ? OFB_PROLONG_STAGES : C::STAGES;
    static constexpr int STAGE_ELEMS = NROWS * C::ROW;
    static constexpr size_t RING_BYTES = (size_t)STAGES * STAGE_ELEMS * sizeof(T);
    static constexpr size_t SMEM_BYTES = RING_BYTES + 2 * STAGES * sizeof(uint64_t);
};

template <int N> using IC = std::integral_constant<int, N>;

// Work distribution over the persistent CTAs (identical walk in the producer and the consumer warps).
//   strips   : work items = (pair, column tile, strip of rows_per_strip rows); CTA c takes items c, c+G, c+2G, ...
//   balanced : the (pair, tile) columns are laid end to end, rows fastest, and cut into G equal row ranges; a CTA
//              walks its range and restarts its window where the range runs into the next column.  Every CTA gets
//              the same number of rows (+- 1), so there is no partly filled last round: 16 pairs of 1024^2 on 444
//              CTAs cost 74 + ~3 rows per CTA instead of two rounds of 38 + 2.5.
template <typename T> struct WorkWalk {
    const LevelArgs<T>& L;
    long long cur, end;
    int item, n_items;
    int scan_pair, scan_seen;        // balanced: pairs examined so far / active pairs among them
    __device__ __forceinline__ WorkWalk(const LevelArgs<T>& L_, int n_items_, int n_pairs)
        : L(L_), cur(0), end(0), item(blockIdx.x), n_items(n_items_), scan_pair(0), scan_seen(0) {
        if (L.balanced) {
            // only the pairs that are still iterating are laid end to end: late iterations with a few active pairs
            // are spread over all CTAs like the early ones
            int n_act = n_pairs;
            if (L.active != nullptr) {
                n_act = 0;
                for (int p = 0; p < n_pairs; ++p) n_act += (L.active[p] != 0) ? 1 : 0;
            }
            const long long total = (long long)n_act * L.tiles_x * L.g.Nx;
            cur = (long long)blockIdx.x * total / gridDim.x;
            end = (long long)(blockIdx.x + 1) * total / gridDim.x;
        }
    }
    __device__ __forceinline__ bool next(int& pair, int& tile, int& i_begin, int& i_end) {
        if (L.balanced) {
            if (cur >= end) return false;
            const long long col = cur / L.g.Nx;
            i_begin = (int)(cur - col * L.g.Nx);
            const int logical = (int)(col / L.tiles_x);                    // index among the active pairs
            tile = (int)(col - (long long)logical * L.tiles_x);
            if (L.active != nullptr) {
                while (true) {                                             // advance to the logical-th active pair
                    if (L.active[scan_pair] != 0) { if (scan_seen == logical) break; ++scan_seen; }
                    ++scan_pair;
                }
                pair = scan_pair;
            } else {
                pair = logical;
            }
            i_end = (int)min((long long)L.g.Nx, (long long)i_begin + (end - cur));
            cur += i_end - i_begin;
            return true;
        }
        if (item >= n_items) return false;
        const int items_per_pair = L.tiles_x * L.strips;
        pair = item / items_per_pair;
        const int rem = item - pair * items_per_pair;
        tile = L.tile_fastest ? rem % L.tiles_x : rem / L.strips;
        const int strip = L.tile_fastest ? rem / L.tiles_x : rem - tile * L.strips;
        i_begin = strip * L.rows_per_strip;
        i_end = min(i_begin + L.rows_per_strip, L.g.Nx);
        item += gridDim.x;
        return true;
    }
};

template <typename T, bool FROM_IMAGE, typename Op>
__global__ void __launch_bounds__(StreamCfg<T>::THREADS, 3) k_stream(LevelArgs<T> L, const T* __restrict__ x, Op op, int n_items) {
    using C = StreamCfg<T>;
    using Lay = StreamLayout<T, FROM_IMAGE, Op>;
    constexpr int V = C::V, STAGES = Lay::STAGES, NWIN = Lay::NWIN, NCOR = Lay::NCOR, NAHEAD = Lay::NAHEAD, NLANES = Lay::NLANES;
    constexpr int NCENROWS = Lay::NCENROWS;
    constexpr bool PROLONG = OpProlong<Op>::value;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    T* ring = reinterpret_cast<T*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + Lay::RING_BYTES);
    uint64_t* empty = full + STAGES;
    const Geom& g = L.g;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    for (int e = threadIdx.x; e < STAGES * Lay::STAGE_ELEMS; e += blockDim.x) ring[e] = T(0);
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, C::CW); }
        mbar_fence_init();
    }
    fence_proxy_async();
    __syncthreads();
    pdl_enter();                                        // everything above overlaps the tail of the previous kernel

    uint32_t n = 0;                                     // running stage counter (continues across items)

    if (warp == C::CW) {
        // ------------------------------------------------------------------ producer warp
        // lane q owns staged row q: [0,3) solution planes (NEED_X), then the image plane (FROM_IMAGE),
        // then the centre rows: NCEN vectors x 3 planes, then 7 coefficient planes (!FROM_IMAGE).
        const bool is_win = lane < NWIN;
        const bool is_sol = Op::NEED_X && lane < 3;
        const bool is_cor = PROLONG && lane >= NWIN && lane < NWIN + NCOR;      // plane (lane-NWIN)/2, coarse row a / b
        WorkWalk<T> walk(L, n_items, L.n_pairs);
        int pair, tile, i_begin, i_end;
        while (walk.next(pair, tile, i_begin, i_end)) {
            if (L.active != nullptr && L.active[pair] == 0) continue;
            const int c0 = tile * C::TW;
            const int lo = max(c0 - C::HALO, 0), hi = min(c0 + C::TW + C::HALO, g.pitch);
            const uint32_t bytes = (uint32_t)(hi - lo) * (uint32_t)sizeof(T);
            const int nk = i_end - i_begin + 2;
            const T* base = nullptr;
            // coarse levels, float: the coefficient rows come from the bfloat16 planes (half the bytes; no halo columns needed)
            bool coef_half = false;
            const unsigned short* base16 = nullptr;
            uint32_t bytes16 = 0;
            if constexpr (!FROM_IMAGE && sizeof(T) == 4) {
                if (L.coef16 != nullptr) {
                    bytes16 = (uint32_t)(min(c0 + C::TW, L.pitch16) - c0) * 2u;
                    coef_half = lane >= NWIN + NCOR + 3 * Op::NCEN && lane < NLANES;
                    if (coef_half)
                        base16 = L.coef16 + (long long)pair * L.pair16 + (long long)(lane - NWIN - NCOR - 3 * Op::NCEN) * L.plane16 + c0;
                }
            }
            uint32_t cbytes = 0;                    // bytes of one coarse correction row of this tile
            int cpitch = 0;
            if constexpr (PROLONG) {
                const int lo_c = c0 >> 1, hi_c = min(lo_c + Lay::COR_ELEMS, op.gc.pitch);
                cbytes = (uint32_t)(hi_c - lo_c) * (uint32_t)sizeof(T);
                cpitch = op.gc.pitch;
                if (is_cor) base = op.ec + (long long)pair * op.gc.pair + (long long)((lane - NWIN) >> 1) * op.gc.plane + lo_c;
            }
            if (lane < NLANES && !is_cor) {
                if (is_sol) base = x + (long long)pair * g.pair + (long long)lane * g.plane;
                else if (is_win) base = L.coef + (long long)pair * L.coef_pair;               // image plane
                else {
                    const int c = lane - NWIN - NCOR;
                    if (c < 3 * Op::NCEN) base = op.cen[c / 3] + (long long)pair * g.pair + (long long)(c % 3) * g.plane;
                    else base = L.coef + (long long)pair * L.coef_pair + (long long)(c - 3 * Op::NCEN) * g.plane;
                }
                base += lo;
            }
            // row slot of the lane's copy: window rows first, then the packed correction rows, then the centre rows
            const int dst_off = is_cor ? NWIN * C::ROW + (lane - NWIN) * Lay::COR_ELEMS
                                       : (is_win ? lane : lane - NCOR + Lay::NCORSLOTS) * C::ROW + (lo - (c0 - C::HALO));
            for (int k = 0; k < nk; ++k, ++n) {
                const int s = n % STAGES;
                const uint32_t ph = (n / STAGES) & 1u;
                mbar_wait_sleep(empty + s, ph ^ 1u, 256);
                const bool cen_valid = k >= 2;
                if (lane == 0) {
                    uint32_t tx = (uint32_t)(cen_valid ? NWIN + NCENROWS : NWIN) * bytes + (uint32_t)NCOR * cbytes;
                    if (bytes16 != 0 && cen_valid) tx = (uint32_t)(NWIN + 3 * Op::NCEN) * bytes + (uint32_t)NCOEF * bytes16;
                    mbar_arrive_expect_tx(full + s, tx);
                }
                __syncwarp();
                bool cor_done = false;
                if constexpr (PROLONG) {
                    if (is_cor) {
                        // the two coarse rows the fine window row (after its ghost redirect) interpolates from
                        const int fine = sol_row(i_begin - 1 + k, g);
                        const int crow = reflect_row((fine + 1 + ((lane - NWIN) & 1)) >> 1, op.gc);
                        bulk_g2s(ring + s * Lay::STAGE_ELEMS + dst_off, base + (long long)crow * cpitch, cbytes, full + s);
                        cor_done = true;
                    }
                }
                if (coef_half) {
                    // the row slot of the plane, from its first element: 1024 bytes of bfloat16 in a 2080-byte slot
                    if (cen_valid)
                        bulk_g2s(ring + s * Lay::STAGE_ELEMS + (lane - NCOR + Lay::NCORSLOTS) * C::ROW,
                                 base16 + (long long)(i_begin - 2 + k) * L.pitch16, bytes16, full + s);
                } else if (!cor_done && lane < NLANES && (is_win || cen_valid)) {
                    int row;
                    if (is_win) {
                        row = i_begin - 1 + k;
                        row = is_sol ? sol_row(row, g) : min(max(row, 0), g.Nx - 1);
                    } else {
                        row = i_begin - 2 + k;
                    }
                    bulk_g2s(ring + s * Lay::STAGE_ELEMS + dst_off, base + (long long)row * g.pitch, bytes, full + s);
                }
            }
        }
    } else {
        // ------------------------------------------------------------------ consumer warps
        Row<T> ux[3], uy[3], gg[3], pw[3];
        WorkWalk<T> walk(L, n_items, L.n_pairs);
        int pair, tile, i_begin, i_end;
        while (walk.next(pair, tile, i_begin, i_end)) {
            if (L.active != nullptr && L.active[pair] == 0) continue;
            const int c0 = tile * C::TW;
            const int cw0 = c0 + warp * (WARP * V);
            const int j0 = cw0 + lane * V;
            const bool warp_has_cols = cw0 < g.Ny;
            const bool edge_warp = (cw0 == 0) || (cw0 + WARP * V + 1 >= g.Ny - 1);
            const int nk = i_end - i_begin + 2;
            const T a = T(L.alpha[pair] * L.h2), be = T(L.beta[pair] * L.h2);

            auto do_stage = [&](auto SLOT, int k) {
                constexpr int sp = decltype(SLOT)::value, sm = (sp + 1) % 3, sc = (sp + 2) % 3;
                const int s = n % STAGES;
                const uint32_t ph = (n / STAGES) & 1u;
                ++n;
                mbar_wait_sleep(full + s, ph, 32);
                const T* stg = ring + s * Lay::STAGE_ELEMS - (c0 - C::HALO);      // stg[q*ROW + j] = column j of staged row q
                const int i = i_begin + k - 2;                                    // centre row of this stage
                const bool compute = warp_has_cols && k >= 2;
                if (warp_has_cols) {
                    const int iw = i + 1;                                         // window row delivered by this stage
                    const bool ghost_row = is_ghost_row(iw, g);
                    if constexpr (Op::NEED_X) {
                        load_row_ptr<T, LOAD_SOL, true>(ux[sp], stg, ghost_row, j0, g, lane, edge_warp);
                        load_row_ptr<T, LOAD_SOL, true>(uy[sp], stg + C::ROW, ghost_row, j0, g, lane, edge_warp);
                        load_row_ptr<T, LOAD_SOL, true>(gg[sp], stg + 2 * C::ROW, ghost_row, j0, g, lane, edge_warp);
                        if constexpr (PROLONG) {
                            // y = x + P e on the window row: bilinear interpolation from the two staged coarse rows.  Window
                            // column t (fine column j0-1+t, j0 a multiple of 4) takes the coarse columns m+(t>>1), m+((t+1)>>1),
                            // m = j0/2; coarse ghost columns are read at their mirror image (edge warps only); the ghost-column
                            // patch of the window is linear, so the correction is patched the same way and added.
                            const T* cor = ring + s * Lay::STAGE_ELEMS + NWIN * C::ROW - (c0 >> 1);     // cor[r*COR_ELEMS + m] = coarse column m
                            const int m0 = j0 >> 1;
                            const bool one_row = (sol_row(iw, g) & 1) != 0;       // odd fine rows sit on a coarse row: both staged rows are it
                            Row<T>* dst3[3] = {&ux[sp], &uy[sp], &gg[sp]};
#pragma unroll
                            for (int q = 0; q < 3; ++q) {
                                const T* ca = cor + (2 * q) * Lay::COR_ELEMS;
                                const T* cb = ca + Lay::COR_ELEMS;
                                T sm_[4];
                                if (!edge_warp) {
                                    // four consecutive coarse columns from an even index: two 8-byte loads per coarse row
                                    const float2 a0 = *reinterpret_cast<const float2*>(ca + m0), a1 = *reinterpret_cast<const float2*>(ca + m0 + 2);
                                    float2 b0 = a0, b1 = a1;
                                    if (!one_row) {
                                        b0 = *reinterpret_cast<const float2*>(cb + m0);
                                        b1 = *reinterpret_cast<const float2*>(cb + m0 + 2);
                                    }
                                    sm_[0] = a0.x + b0.x; sm_[1] = a0.y + b0.y; sm_[2] = a1.x + b1.x; sm_[3] = a1.y + b1.y;
                                } else {
#pragma unroll
                                    for (int t = 0; t < 4; ++t) {
                                        const int mc = reflect_clamped(m0 + t, op.gc.Ny, op.gc.far_j);
                                        sm_[t] = ca[mc] + cb[mc];
                                    }
                                }
                                Row<T> er;
                                er.v[0] = T(0.5) * sm_[0];
                                er.v[1] = T(0.25) * (sm_[0] + sm_[1]);
                                er.v[2] = T(0.5) * sm_[1];
                                er.v[3] = T(0.25) * (sm_[1] + sm_[2]);
                                er.v[4] = T(0.5) * sm_[2];
                                er.v[5] = T(0.25) * (sm_[2] + sm_[3]);
                                if (edge_warp) patch_ghost_columns<T, LOAD_SOL>(er, ghost_row, j0, g);
#pragma unroll
                                for (int t = 0; t < V + 2; ++t) dst3[q]->v[t] += er.v[t];
                            }
                        }
                    }
                    if constexpr (FROM_IMAGE) load_row_ptr<T, LOAD_RAW, true>(pw[sp], stg + (NWIN - 1) * C::ROW, false, j0, g, lane, false);
                }
                if (compute) {
                    // centre rows (vectors, coefficient planes) are read from the stage at the point of use;
                    // the stage is released after the row has been computed
                    Coef<T> cf[V];
                    bool inside[V];
                    T sx[3][V], xc[3][V];
                    const bool row_inside = (i >= 1) && (i <= g.Nx - 2);
                    if constexpr (!FROM_IMAGE) {
                        using VT = typename VecOf<T>::type;
                        T cr[NCOEF][V];
                        bool half_rows = false;
                        if constexpr (sizeof(T) == 4) half_rows = L.coef16 != nullptr;
                        if (half_rows) {
                            if constexpr (sizeof(T) == 4) {
                                // four bfloat16 of this lane's columns: 8 bytes at (j0 - c0) elements into the plane's row slot
                                const unsigned short* slot0 = reinterpret_cast<const unsigned short*>(ring + s * Lay::STAGE_ELEMS +
                                                                                                      (NAHEAD + 3 * Op::NCEN) * C::ROW) + (j0 - c0);
#pragma unroll
                                for (int q = 0; q < NCOEF; ++q) {
                                    const uint2 raw = *reinterpret_cast<const uint2*>(slot0 + (long long)q * C::ROW * 2);
                                    cr[q][0] = __uint_as_float(raw.x << 16); cr[q][1] = __uint_as_float(raw.x & 0xffff0000u);
                                    cr[q][2] = __uint_as_float(raw.y << 16); cr[q][3] = __uint_as_float(raw.y & 0xffff0000u);
                                }
                            }
                        } else {
#pragma unroll
                        for (int q = 0; q < NCOEF; ++q) {
                            const VT t = *reinterpret_cast<const VT*>(stg + (NAHEAD + 3 * Op::NCEN + q) * C::ROW + j0);
                            const T* p = reinterpret_cast<const T*>(&t);
#pragma unroll
                            for (int kk = 0; kk < V; ++kk) cr[q][kk] = p[kk];
                        }
                        }
#pragma unroll
                        for (int kk = 0; kk < V; ++kk) {
                            cf[kk].II = cr[0][kk]; cf[kk].IIx = cr[1][kk]; cf[kk].IIxx = cr[2][kk]; cf[kk].IIyy = cr[3][kk];
                            cf[kk].IIxy = cr[4][kk]; cf[kk].Ih = cr[5][kk]; cf[kk].Ix = cr[6][kk];
                        }
                    }
#pragma unroll
                    for (int kk = 0; kk < V; ++kk) {
                        const int j = j0 + kk;
                        inside[kk] = row_inside && (j >= 1) && (j <= g.Ny - 2);
                        if constexpr (FROM_IMAGE) cf[kk] = coef_from_image<T>(pw[sm], pw[sc], pw[sp], kk + 1);
                        if constexpr (Op::NEED_X) {
                            Win<T> wx{ux[sm], ux[sc], ux[sp]}, wy{uy[sm], uy[sc], uy[sp]}, wg{gg[sm], gg[sc], gg[sp]};
                            apply_point<T>(cf[kk], wx, wy, wg, kk + 1, a, be, sx[0][kk], sx[1][kk], sx[2][kk]);
                            xc[0][kk] = ux[sc].v[kk + 1]; xc[1][kk] = uy[sc].v[kk + 1]; xc[2][kk] = gg[sc].v[kk + 1];
                        } else {
                            sx[0][kk] = sx[1][kk] = sx[2][kk] = T(0);
                            xc[0][kk] = xc[1][kk] = xc[2][kk] = T(0);
                        }
                    }
                    const CenShared<T, C::ROW> cv{stg + NAHEAD * C::ROW + j0};
                    op.row(pair, i, j0, cf, sx, xc, inside, a, be, cv);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(empty + s);
            };
            int k = 0;
            while (k < nk) {
                do_stage(IC<0>{}, k); if (++k >= nk) break;
                do_stage(IC<1>{}, k); if (++k >= nk) break;
                do_stage(IC<2>{}, k); ++k;
            }
            op.flush(pair);
        }
    }
}

}  // namespace ofb
