// ofb200 - common device helpers (sm_100a).
//
// Data layout used by every kernel: SoA planes, row-major, fastest axis = image axis 2 ("y", j).
//   plane(q)[i*pitch + j],  q-th plane at  base + q*plane_stride,  pair b at  base + b*pair_stride.
// A warp owns 32*V consecutive columns (V = 16 bytes / sizeof(T): float4 / double2 accesses) and
// marches down a strip of rows keeping a 3-row register window; horizontal neighbours come from
// warp shuffles, so there is no shared-memory staging and no block barrier in the stencil sweeps.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ofb {

constexpr int WARP = 32;
constexpr unsigned FULL = 0xffffffffu;

template <typename T> struct VecOf;
template <> struct VecOf<float>  { static constexpr int V = 4; using type = float4; };
template <> struct VecOf<double> { static constexpr int V = 2; using type = double2; };

// geometry of one level (all planes of a level share it)
struct Geom {
    int Nx, Ny, pitch;           // rows, columns, row pitch in elements
    long long plane;             // elements between field planes
    long long pair;              // elements between batch entries (vectors: 3*plane or more)
    long long img_pair;          // elements between batch entries for single-plane images
    int aligned;                 // 1 if every row start is 16-byte aligned (pitch % V == 0 and base aligned)
    int far_i, far_j;            // far-side ghost rule per axis: ghost(N-1) = q(N-1-far); 2 = mirror about the last
                                 // interior node (the reference's rule, level 0), 1 = mirror about the half point
                                 // beyond it (coarse levels whose fine parent has an even number of interior nodes)
    int halo_lo, halo_hi;        // slab decomposition: row 0 / row Nx-1 hold the neighbouring slab's boundary row (real
                                 // data, refreshed by the halo exchange) instead of being a mirror ghost row
};

// ghost-ring handling when a plane is read
enum LoadMode { LOAD_RAW = 0,    // values as stored (images, right-hand sides)
                LOAD_SOL = 1,    // solution-type: ghost ring from the reference's boundary rows
                                 // (optical_flow.py:965-1070): q(0,j)=q(2,j) ..., corners = q(2,0)+q(0,2)
                LOAD_EVEN = 2 }; // even reflection (corner = q(2,2)); used by grid transfers

__device__ __forceinline__ int mirror_index(int i, int N) {
    return i == 0 ? 2 : (i == N - 1 ? N - 3 : i);
}
__device__ __forceinline__ int mirror_far(int i, int N, int far) {
    return i == 0 ? 2 : (i == N - 1 ? N - 1 - far : i);
}

// Row a solution-type plane is read from for logical row i: mirror ghost rows are redirected to their interior
// image, halo rows of a slab are read as stored.
__device__ __forceinline__ int sol_row(int i, const Geom& g) {
    if (i <= 0) return g.halo_lo ? 0 : 2;
    if (i >= g.Nx - 1) return g.halo_hi ? g.Nx - 1 : g.Nx - 1 - g.far_i;
    return i;
}
__device__ __forceinline__ bool is_ghost_row(int i, const Geom& g) {
    return (i == 0 && !g.halo_lo) || (i == g.Nx - 1 && !g.halo_hi);
}

// One row of one plane as seen by a lane: v[0] = left halo (column j0-1), v[1..V] own columns,
// v[V+1] = right halo (column j0+V).
template <typename T> struct Row {
    T v[VecOf<T>::V + 2];
};

template <typename T>
__device__ __forceinline__ void load_own(const T* __restrict__ rowptr, int j0, int Ny, bool aligned,
                                         T (&out)[VecOf<T>::V]) {
    constexpr int V = VecOf<T>::V;
    if (aligned && j0 + V <= Ny) {
        typename VecOf<T>::type t = *reinterpret_cast<const typename VecOf<T>::type*>(rowptr + j0);
        const T* p = reinterpret_cast<const T*>(&t);
#pragma unroll
        for (int k = 0; k < V; ++k) out[k] = p[k];
    } else {
#pragma unroll
        for (int k = 0; k < V; ++k) {
            int j = j0 + k;
            out[k] = (j < Ny) ? rowptr[j] : T(0);
        }
    }
}

template <typename T>
__device__ __forceinline__ void store_own(T* __restrict__ rowptr, int j0, int Ny, bool aligned,
                                          const T (&val)[VecOf<T>::V]) {
    constexpr int V = VecOf<T>::V;
    if (aligned && j0 + V <= Ny) {
        typename VecOf<T>::type t;
        T* p = reinterpret_cast<T*>(&t);
#pragma unroll
        for (int k = 0; k < V; ++k) p[k] = val[k];
        *reinterpret_cast<typename VecOf<T>::type*>(rowptr + j0) = t;
    } else {
#pragma unroll
        for (int k = 0; k < V; ++k) {
            int j = j0 + k;
            if (j < Ny) rowptr[j] = val[k];
        }
    }
}

// Ghost columns of a window row are patched in registers from the lane's own window: column 0 <- column 2,
// column Ny-1 <- column Ny-1-far_j (window index kg = Ny - j0, mirror at kg - far_j); doubled on ghost rows of
// solution-type planes (the reference's corner rows).  Linear in the row, so a correction added to a row can be
// patched separately (fused prolongation in the streamed smoother).
template <typename T, int MODE>
__device__ __forceinline__ void patch_ghost_columns(Row<T>& r, bool ghost_row, int j0, const Geom& g) {
    constexpr int V = VecOf<T>::V;
    if (j0 == 0) r.v[1] = (MODE == LOAD_SOL && ghost_row) ? r.v[3] + r.v[3] : r.v[3];
    if (j0 == 1) r.v[0] = (MODE == LOAD_SOL && ghost_row) ? r.v[2] + r.v[2] : r.v[2];
    const int kg = g.Ny - j0;
#pragma unroll
    for (int k = 2; k <= V + 1; ++k)
        if (kg == k) {
            const T t = (g.far_j == 2) ? r.v[k - 2] : r.v[k - 1];
            r.v[k] = (MODE == LOAD_SOL && ghost_row) ? t + t : t;
        }
}

// Load one row with halos from `rowptr` (rowptr[j] = column j; global memory or a shared-memory row
// staged by the streamed sweeps).  `edge_warp` is warp-uniform: true when the warp's column range
// (with halo) touches column 0 or Ny-1, the only place the ghost remap costs anything.
// FULLVEC: the row is 16-byte aligned and readable past Ny (zero-filled staging) - always one vector load.
template <typename T, int MODE, bool FULLVEC>
__device__ __forceinline__ void load_row_ptr(Row<T>& r, const T* __restrict__ rowptr, bool ghost_row, int j0,
                                             const Geom& g, int lane, bool edge_warp) {
    constexpr int V = VecOf<T>::V;
    T own[V];
    if (FULLVEC) {
        typename VecOf<T>::type t = *reinterpret_cast<const typename VecOf<T>::type*>(rowptr + j0);
        const T* p = reinterpret_cast<const T*>(&t);
#pragma unroll
        for (int k = 0; k < V; ++k) own[k] = p[k];
    } else {
        load_own<T>(rowptr, j0, g.Ny, g.aligned != 0, own);
    }
#pragma unroll
    for (int k = 0; k < V; ++k) r.v[k + 1] = own[k];
    if (FULLVEC) {
        // staged row: the halo columns are in shared memory next to the own vector
        r.v[0] = rowptr[j0 - 1];
        r.v[V + 1] = rowptr[j0 + V];
    } else {
        T left = __shfl_up_sync(FULL, own[V - 1], 1);
        T right = __shfl_down_sync(FULL, own[0], 1);
        if (lane == 0) left = (j0 > 0) ? rowptr[j0 - 1] : T(0);
        if (lane == WARP - 1) right = (j0 + V < g.Ny) ? rowptr[j0 + V] : T(0);
        r.v[0] = left;
        r.v[V + 1] = right;
    }
    if (MODE != LOAD_RAW && edge_warp) patch_ghost_columns<T, MODE>(r, ghost_row, j0, g);
}

// Load row i of a plane in global memory with halos.
template <typename T, int MODE>
__device__ __forceinline__ void load_row(Row<T>& r, const T* __restrict__ plane, int i, int j0,
                                         const Geom& g, int lane, bool edge_warp) {
    int row = min(max(i, 0), g.Nx - 1);
    if (MODE != LOAD_RAW) row = sol_row(i, g);
    load_row_ptr<T, MODE, false>(r, plane + (long long)row * g.pitch, is_ghost_row(i, g), j0, g, lane, edge_warp);
}

// ---------------------------------------------------------------- programmatic dependent launch
// With OFB_FLAG_PDL the kernels of the solver loop are launched with the programmatic-stream-serialisation
// attribute: a kernel lets its successor be scheduled right away (its CTAs take SM slots as this kernel's CTAs
// retire and run their prologue) and waits here for its predecessor's results before it touches global memory.
// Without the attribute both instructions are no-ops.  Measured on the 16-pair step: neutral to -0.3 % (the
// ~400 dependent launches are already queued back to back and the prologues are short), so it is off by default.
__device__ __forceinline__ void pdl_enter() {
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

// ---------------------------------------------------------------- reductions
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    return v;
}

// Block reduction of NACC doubles, one atomicAdd per accumulator per block.
template <int NACC>
__device__ __forceinline__ void block_accumulate(double (&acc)[NACC], double* __restrict__ dst) {
    __shared__ double smem[NACC][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nwarps = (blockDim.x * blockDim.y + 31) >> 5;
#pragma unroll
    for (int a = 0; a < NACC; ++a) {
        double v = warp_sum(acc[a]);
        if (lane == 0) smem[a][warp] = v;
    }
    __syncthreads();
    if (warp == 0) {
#pragma unroll
        for (int a = 0; a < NACC; ++a) {
            double v = (lane < nwarps) ? smem[a][lane] : 0.0;
            v = warp_sum(v);
            if (lane == 0) atomicAdd(dst + a, v);
        }
    }
}

}  // namespace ofb
