// i7_xchg.cuh — best-trajectory selection fused with the multi-GPU exchange over peer memory (device code shared by the
// stand-alone exchange kernel and the epilogue of the fused solve kernel).
#pragma once
#include <math_constants.h>

#include "i7_rbd.cuh"

namespace i7 {

// -----------------------------------------------------------------------------------------------------------
// Multi-GPU exchange fused with the selection: ONE kernel per rank and tick packs the local record, stores it into
// slot `rank` of EVERY peer's exchange buffer over NVLink (peer-mapped memory, cudaIpc), raises a per-slot flag on
// every peer (release), waits until the flags of all ranks have reached this tick on its own buffer (acquire) and
// picks the winner — the all-gather and both selection kernels of the NCCL path in a single launch, no host-side
// collective call.  Buffers are double-buffered by tick parity: a rank can be at most one tick ahead of a peer (it
// needs the peer's flag of the current tick to finish), so the data of tick t is never overwritten before every peer
// has consumed it.  Each rank runs this on its own GPU; a rank that does not arrive within `timeout_ns` makes the
// others give up with *err = 1 instead of spinning forever.
//   buffer of a rank (bytes): [parity 0: world x stride x T | parity 1: ... | flags: 2 x world x u64]
struct P2PLayout { size_t data_bytes, flags_off, total; };
__host__ __device__ inline P2PLayout p2p_layout(int world, int stride, int esize) {
    P2PLayout L;
    L.data_bytes = (((size_t)world * stride * esize) + 255) & ~size_t(255);
    L.flags_off = 2 * L.data_bytes;
    L.total = L.flags_off + 2 * (size_t)world * sizeof(unsigned long long);
    return L;
}
I7_DEV unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
I7_DEV void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
I7_DEV unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Runs on ONE CTA (any size that is a multiple of 32).  stats / xu may have been written by other CTAs of the same
// launch (the epilogue form): they are read with ld.global.cg.  wait_ns (optional): time spent waiting for the peers.
template <typename T>
__device__ __noinline__ void pack_exchange_reduce(int batch, int nz, const T* stats, int stats_stride, int cost_col,
                                                  const T* xu, int index_offset, int rank, int world,
                                                  unsigned char* const* peers, unsigned long long seq,
                                                  unsigned long long timeout_ns, T* out, int* err, long long* wait_ns) {
    __shared__ T s_val[32];
    __shared__ int s_idx[32];
    __shared__ int s_win, s_timed_out;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_timed_out = 0;
    const int stride = batch + 2 + nz;
    const P2PLayout L = p2p_layout(world, stride, (int)sizeof(T));
    const int par = (int)(seq & 1);
    unsigned char* mine = peers[rank];
    T* packed = reinterpret_cast<T*>(mine + par * L.data_bytes) + (size_t)rank * stride;
    // ---- pack (same selection as pack_best_kernel: lowest cost, lowest index on ties, NaN never wins) ----
    T best = T(0); int bi = -1;
#pragma unroll 4
    for (int b = tid; b < batch; b += blockDim.x) {
        const T c = __ldcg(stats + (size_t)b * stats_stride + cost_col);
        packed[b] = c;
        if (c == c && (bi < 0 || c < best)) { best = c; bi = b; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        const T ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (oi >= 0 && (bi < 0 || ov < best || (ov == best && oi < bi))) { best = ov; bi = oi; }
    }
    if (lane == 0) { s_val[warp] = best; s_idx[warp] = bi; }
    __syncthreads();
    if (warp == 0) {
        const int nw = blockDim.x >> 5;
        best = lane < nw ? s_val[lane] : T(0); bi = lane < nw ? s_idx[lane] : -1;
        for (int o = 16; o > 0; o >>= 1) {
            const T ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (oi >= 0 && (bi < 0 || ov < best || (ov == best && oi < bi))) { best = ov; bi = oi; }
        }
        if (lane == 0) { s_val[0] = best; s_idx[0] = bi; }
    }
    __syncthreads();
    bi = s_idx[0];
    T* tail = packed + batch;
    if (tid == 0) { tail[0] = s_val[0]; tail[1] = T(bi < 0 ? -1 : bi + index_offset); }
    for (int t = tid; t < nz; t += blockDim.x) tail[2 + t] = bi < 0 ? T(0) : __ldcg(xu + (size_t)bi * nz + t);
    __syncthreads();
    // ---- scatter the record into slot `rank` of every peer (posted stores over NVLink) ----
    for (int r = 0; r < world; ++r) {
        if (r == rank) continue;
        T* dst = reinterpret_cast<T*>(peers[r] + par * L.data_bytes) + (size_t)rank * stride;
        for (int t = tid; t < stride; t += blockDim.x) dst[t] = packed[t];
    }
    __threadfence_system();
    __syncthreads();
    if (tid < world) {
        unsigned long long* f = reinterpret_cast<unsigned long long*>(peers[tid] + L.flags_off) + (size_t)par * world + rank;
        st_release_sys(f, seq);
    }
    // ---- wait for every rank's record of this tick in the own buffer ----
    const unsigned long long t_wait0 = (wait_ns && tid == 0) ? global_ns() : 0ull;
    if (tid < world) {
        const unsigned long long* f = reinterpret_cast<const unsigned long long*>(mine + L.flags_off) + (size_t)par * world + tid;
        const unsigned long long t0 = global_ns();
        while (ld_acquire_sys(f) < seq) {
            if (global_ns() - t0 > timeout_ns) { *err = 1; s_timed_out = 1; break; }
            __nanosleep(200);
        }
    }
    __threadfence_system();
    __syncthreads();
    if (wait_ns && tid == 0) *wait_ns += (long long)(global_ns() - t_wait0);     // summed over the ticks; the host divides
    if (s_timed_out) {
        // a rank did not arrive: no winner can be named.  The caller gets a record that cannot be mistaken for one
        // (cost NaN, index -1, zero trajectory) and i7_p2p_status reports the time-out until i7_p2p_reset
        for (int t = tid; t < 2 + nz; t += blockDim.x) out[t] = t == 0 ? T(CUDART_NAN) : (t == 1 ? T(-1) : T(0));
        return;
    }
    // ---- winner over the ranks (same rule as reduce_best_kernel: lowest cost, lowest rank on ties) ----
    const T* gathered = reinterpret_cast<const T*>(mine + par * L.data_bytes);
    if (tid == 0) {
        int win = -1; T bc = T(0);
        for (int r = 0; r < world; ++r) {
            const T* tl = gathered + (size_t)r * stride + batch;
            const T c = __ldcg(tl), ix = __ldcg(tl + 1);
            if (ix >= T(0) && (win < 0 || c < bc)) { bc = c; win = r; }
        }
        s_win = win < 0 ? 0 : win;
    }
    __syncthreads();
    const T* tl = gathered + (size_t)s_win * stride + batch;
    for (int t = tid; t < 2 + nz; t += blockDim.x) out[t] = __ldcg(tl + t);
}

}  // namespace i7
