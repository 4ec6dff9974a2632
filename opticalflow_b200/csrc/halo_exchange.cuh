// ofb200 - halo-row exchange between neighbouring slabs by peer stores over NVLink (config d-5).
//
// One launch per exchange, two CTAs: CTA 0 talks to rank-1, CTA 1 to rank+1.  Each CTA
//   1. writes this rank's boundary row of every plane straight into the neighbour's mailbox (16-byte peer stores),
//   2. fences (system scope) and publishes the exchange number in the neighbour's arrival flag,
//   3. waits for its own arrival flag (set by the neighbour's step 2) and copies the mailbox rows into the halo rows.
// Step 1/2 never depend on the neighbour, so two ranks entering the same exchange cannot dead-lock; the wait has a
// wall-clock bound and raises an error word instead of hanging.  Compared with a grouped ncclSend/ncclRecv per
// plane this is one small launch (~5 us) instead of ~20 us, which matters because a V-cycle exchanges after every
// sweep on every distributed level.
#pragma once
#include <stdint.h>
#include "common.cuh"

namespace ofb {

template <typename T> struct HaloArgs {
    T* base;                         // first plane of pair 0
    long long plane, pair_stride;    // elements between planes / pairs
    int nplanes, B, Nx, pitch;       // planes per pair, pairs, local rows, row length in elements
    unsigned char* local;            // this rank's mailbox
    unsigned char* peer[2];          // neighbours' mailboxes (nullptr: no neighbour on that side)
    unsigned long long seq;          // exchange number (>= 1), identical on both sides
    size_t header, slot_bytes;
    unsigned long long timeout_ns;
};

__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void store_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long load_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

template <typename T>
__global__ void __launch_bounds__(1024) k_halo_exchange(HaloArgs<T> A) {
    const int dir = blockIdx.x;                       // 0: neighbour rank-1 (my row 1 <-> my row 0), 1: rank+1
    if (A.peer[dir] == nullptr) return;
    const int parity = (int)(A.seq & 1ull);
    const int send_row = dir == 0 ? 1 : A.Nx - 2, halo_row = dir == 0 ? 0 : A.Nx - 1;
    const int rows = A.B * A.nplanes;
    constexpr int V = VecOf<T>::V;
    using VT = typename VecOf<T>::type;
    const bool vec = (A.pitch % V == 0) && ((reinterpret_cast<uintptr_t>(A.base) & 15) == 0) && (A.plane % V == 0) &&
                     (A.pair_stride % V == 0);
    // my rows arrive at the neighbour from the opposite side: its slot [1 - dir]
    T* out = reinterpret_cast<T*>(A.peer[dir] + A.header + (size_t)((1 - dir) * 2 + parity) * A.slot_bytes);
    for (int r = 0; r < rows; ++r) {
        const int pair = r / A.nplanes, q = r - pair * A.nplanes;
        const T* src = A.base + (long long)pair * A.pair_stride + (long long)q * A.plane + (long long)send_row * A.pitch;
        T* dst = out + (long long)r * A.pitch;
        if (vec) {
            for (int e = threadIdx.x; e < A.pitch / V; e += blockDim.x)
                reinterpret_cast<VT*>(dst)[e] = reinterpret_cast<const VT*>(src)[e];
        } else {
            for (int e = threadIdx.x; e < A.pitch; e += blockDim.x) dst[e] = src[e];
        }
    }
    __threadfence_system();
    __syncthreads();
    unsigned long long* my_flags = reinterpret_cast<unsigned long long*>(A.local);
    if (threadIdx.x == 0) {
        unsigned long long* peer_flags = reinterpret_cast<unsigned long long*>(A.peer[dir]);
        store_release_sys(peer_flags + (1 - dir) * 2 + parity, A.seq);
        const unsigned long long t0 = global_timer_ns();
        while (load_acquire_sys(my_flags + dir * 2 + parity) < A.seq) {
            if (global_timer_ns() - t0 > A.timeout_ns) {
                *reinterpret_cast<int*>(A.local + 4 * sizeof(unsigned long long)) = 1;      // error word
                break;
            }
            __nanosleep(64);
        }
    }
    __syncthreads();
    const T* in = reinterpret_cast<const T*>(A.local + A.header + (size_t)(dir * 2 + parity) * A.slot_bytes);
    for (int r = 0; r < rows; ++r) {
        const int pair = r / A.nplanes, q = r - pair * A.nplanes;
        T* dst = A.base + (long long)pair * A.pair_stride + (long long)q * A.plane + (long long)halo_row * A.pitch;
        const T* src = in + (long long)r * A.pitch;
        if (vec) {
            for (int e = threadIdx.x; e < A.pitch / V; e += blockDim.x)
                reinterpret_cast<VT*>(dst)[e] = __ldcg(reinterpret_cast<const VT*>(src) + e);
        } else {
            for (int e = threadIdx.x; e < A.pitch; e += blockDim.x) dst[e] = __ldcg(src + e);
        }
    }
}

}  // namespace ofb
