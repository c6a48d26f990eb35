// i7_api.cu — C ABI of the B200-native SQP-MPC solve loop (see include/indy7_b200.h for the boundary).
#include "../../include/indy7_b200.h"
#undef I7_NJ
#include "i7_launch.h"
#include "i7_aux.cuh"
#include "i7_admm.cuh"

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

namespace {

thread_local std::string g_create_error;

struct Arena {           // a device buffer with a pinned host mirror
    void* dev = nullptr;
    void* host = nullptr;
    size_t bytes = 0;
};

}  // namespace

struct i7_handle {
    i7_config cfg;
    int esize = 8;
    size_t nz = 0;
    i7::Robot<double> rob64;
    i7::Robot<float> rob32;
    Arena in, out;               // packed inputs / outputs for host-pointer calls
    void* d_fext = nullptr;      // batch x 6, Featherstone order, handle dtype
    bool has_fext = false;
    long long* d_prof = nullptr; // optional phase-clock buffer (i7_enable_profile)
    void* d_spill = nullptr;     // gains/z region when it does not fit shared memory
    bool spill = false;
    int smem_bytes = 0;
    int threads = 0;
    int max_smem_optin = 0;
    bool use_cluster = false;    // i7_solve runs the cluster-per-problem (latency) kernel
    int cluster_size = 0;        // CTAs per problem of that kernel: 8, 4 or 2
    int cluster_smem_bytes = 0;
    void* d_xbuf = nullptr;      // exchange buffer of the cluster kernel (batch x cluster_xbuf_len(N))
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // the pinned input mirror (in.host) is read asynchronously by a copy / by the kernel itself when a call with host
    // inputs and device outputs returns without waiting: the next call that stages through it waits for this event
    cudaEvent_t mirror_ev = nullptr;
    bool mirror_pending = false;
    // host-buffer calls on big batches run as a software pipeline of chunks over these streams (copy-in of chunk
    // i+1 and copy-out of chunk i-1 overlap the kernel of chunk i); pipe_chunk = problems per chunk, 0 = off
    static constexpr int kPipeStreams = 6, kPipeMaxChunks = 64;    // streams in use: pipe_nstreams (I7_PIPE_STREAMS)
    cudaStream_t pipe_stream[kPipeStreams] = {};
    cudaEvent_t pipe_fork = nullptr, pipe_join[kPipeStreams] = {};
    int pipe_nstreams = 3;
    cudaEvent_t pipe_done[kPipeMaxChunks] = {};
    // mode 2 (default): all copies-in on ONE stream back to back, kernels round-robin on the compute streams (each
    // waits for its chunk's inputs), all copies-out on ONE stream — a compute stream never waits for a copy-out
    cudaStream_t pipe_in = nullptr, pipe_out = nullptr;
    cudaEvent_t pipe_in_done[kPipeMaxChunks] = {}, pipe_k_done[kPipeMaxChunks] = {};
    int pipe_mode = 2;
    int pipe_chunk = 0;
    // peer-memory exchange (i7_p2p_*): own buffer, the peers' mapped buffers, device copy of the pointer table
    int p2p_world = 0, p2p_rank = 0;
    void* p2p_mine = nullptr;
    void* p2p_peer[16] = {};
    unsigned char** p2p_table = nullptr;
    int* p2p_err = nullptr;
    unsigned* d_sel_counter = nullptr;   // i7_find_best_idx: ticket counter + error scratch (lazy)
    void* d_sel_errs = nullptr;
    unsigned long long p2p_seq = 0;
    unsigned* d_xchg_ticket = nullptr;   // i7_solve_select: problems finished in the current launch
    long long* d_xchg_wait = nullptr;    // ... and the time its epilogue spent waiting for the peers (ns, summed)
    // qp_mode "osqp": per-problem ADMM warm start (x, y, y of the limit rows), QP blocks, factor workspace, QP solution,
    // QP record, "SQP loop ended" flags, statistics scratch
    void *d_wx = nullptr, *d_wy = nullptr, *d_wyl = nullptr, *d_blk = nullptr, *d_fac = nullptr, *d_zbuf = nullptr,
         *d_qpinfo = nullptr, *d_stats = nullptr;
    int* d_done = nullptr;
    int admm_smem_bytes = 0;
    i7::AdmmSettings admm{};
    int opt_x0_into_xu = 0;      // i7_set_option(I7_OPT_X0_INTO_WARM_START)
    int zero_copy = 1;           // I7_ZERO_COPY=0: page-locked user buffers go through copies / the chunked pipeline
    int no_events = 0;           // I7_TIMING_EVENTS=0: no timing events around mapped-memory solves (3 us of host time each)
    int direct_max_batch = 0;    // host-buffer solves up to this batch use mapped host memory instead of copies
    bool timed = false;
    int64_t launches = 0;
    std::string err;
};

namespace {

#define I7_CUDA(h, call)                                                                         \
    do {                                                                                         \
        cudaError_t e_ = (call);                                                                 \
        if (e_ != cudaSuccess) {                                                                 \
            (h)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                       \
            return I7_ERR_CUDA;                                                                  \
        }                                                                                        \
    } while (0)

template <typename T>
void fill_robot(const i7_model& m, i7::Robot<T>& r) {
    for (int i = 0; i < 6; ++i) {
        for (int a = 0; a < 9; ++a) r.R[i][a] = T(m.R[i][a]);
        for (int a = 0; a < 3; ++a) { r.p[i][a] = T(m.p[i][a]); r.h[i][a] = T(m.mass[i] * m.com[i][a]); }
        r.m[i] = T(m.mass[i]);
        // Ib = Ic + m (c.c I - c c')
        const double* c = m.com[i];
        const double cc = c[0] * c[0] + c[1] * c[1] + c[2] * c[2];
        double Ib[3][3];
        for (int a = 0; a < 3; ++a)
            for (int b = 0; b < 3; ++b)
                Ib[a][b] = m.Ic[i][3 * a + b] + m.mass[i] * ((a == b ? cc : 0.0) - c[a] * c[b]);
        r.Ib[i][0] = T(Ib[0][0]); r.Ib[i][1] = T(Ib[0][1]); r.Ib[i][2] = T(Ib[0][2]);
        r.Ib[i][3] = T(Ib[1][1]); r.Ib[i][4] = T(Ib[1][2]); r.Ib[i][5] = T(Ib[2][2]);
    }
    for (int a = 0; a < 3; ++a) r.g[a] = T(m.gravity[a]);
}

// 0 = pageable host memory, 1 = device / managed memory, 2 = page-locked (pinned / registered) host memory
int ptr_kind(const void* p) {
    if (!p) return 0;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return 0; }
    if (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged) return 1;
    if (at.type == cudaMemoryTypeHost) return 2;
    return 0;
}

i7_status arena_reserve(i7_handle* h, Arena& a, size_t bytes) {
    if (bytes <= a.bytes) return I7_OK;
    if (h->mirror_pending) { cudaEventSynchronize(h->mirror_ev); h->mirror_pending = false; }
    if (a.dev) cudaFree(a.dev);
    if (a.host) cudaFreeHost(a.host);
    a.dev = a.host = nullptr; a.bytes = 0;
    I7_CUDA(h, cudaMalloc(&a.dev, bytes));
    I7_CUDA(h, cudaMallocHost(&a.host, bytes));
    a.bytes = bytes;
    return I7_OK;
}

// Staging of the buffers of one call.  Device pointers are used as they are.  Page-locked host buffers are copied
// straight to / from their slice of the device arena with cudaMemcpyAsync (no intermediate copy).  Pageable host
// buffers go through the pinned mirror of the arena (one memcpy + one bulk transfer per direction).
struct Stage {
    size_t in_off = 0, out_off = 0;
    i7_handle* h;
    // direct: the kernel reads its inputs from and writes its results to the page-locked mirror itself (mapped host
    // memory, same address on the device under unified addressing) — no copy operations in the stream.  For a handful
    // of problems the two copies cost more stream latency (~10 us) than the few KB cost over PCIe inside the kernel.
    bool direct = false;
    // zero_copy: page-locked USER buffers are handed to the kernel as they are (cudaHostAlloc / registered memory is
    // device-accessible under unified addressing): every CTA pulls its problem over PCIe when it starts and pushes its
    // result when it ends, so the transfers of a big batch are spread over the whole launch with no copy operation,
    // no staging and a single launch.
    bool zero_copy = false, zc_pending = false;
    explicit Stage(i7_handle* h_, bool direct_ = false) : h(h_), direct(direct_) {}
    static void* mapped(const void* p) {
        void* d = nullptr;
        if (cudaHostGetDevicePointer(&d, const_cast<void*>(p), 0) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        return d;
    }
    struct Xfer { void* user; size_t off; size_t bytes; bool pinned; };
    Xfer ins[8], outs[12];
    int nin = 0, nout = 0;
    bool used_mirror = false;
    const void* in(const void* p, size_t bytes) {
        if (!p) return nullptr;
        const int kind = ptr_kind(p);
        if (kind == 1) return p;
        if (kind == 2 && zero_copy)
            if (const void* d = mapped(p)) return d;
        if (kind == 0 || direct) {
            if (h->mirror_pending) { cudaEventSynchronize(h->mirror_ev); h->mirror_pending = false; }
            std::memcpy((char*)h->in.host + in_off, p, bytes);
            used_mirror = true;
        }
        if (!direct) ins[nin++] = Xfer{const_cast<void*>(p), in_off, bytes, kind == 2};
        const void* d = (char*)(direct ? h->in.host : h->in.dev) + in_off;
        in_off += (bytes + 15) & ~size_t(15);
        return d;
    }
    void* out(void* p, size_t bytes) {
        if (!p) return nullptr;
        const int kind = ptr_kind(p);
        if (kind == 1) return p;
        if (kind == 2 && zero_copy)
            if (void* d = mapped(p)) { zc_pending = true; return d; }
        outs[nout++] = Xfer{p, out_off, bytes, kind == 2 && !direct};
        void* d = (char*)(direct ? h->out.host : h->out.dev) + out_off;
        out_off += (bytes + 15) & ~size_t(15);
        return d;
    }
    // pageable buffers sit back to back in the pinned mirror: runs of them travel as ONE copy (a single-problem solve
    // is 1 transfer in and 1 out instead of 3 + 2; each call costs microseconds of host time on the latency path)
    static size_t al(size_t b) { return (b + 15) & ~size_t(15); }
    i7_status flush_in(cudaStream_t s) {
        for (int i = 0; i < nin;) {
            if (ins[i].pinned) {
                I7_CUDA(h, cudaMemcpyAsync((char*)h->in.dev + ins[i].off, ins[i].user, ins[i].bytes, cudaMemcpyHostToDevice, s));
                ++i;
                continue;
            }
            int j = i;
            while (j + 1 < nin && !ins[j + 1].pinned && ins[j + 1].off == ins[j].off + al(ins[j].bytes)) ++j;
            const size_t span = ins[j].off + ins[j].bytes - ins[i].off;
            I7_CUDA(h, cudaMemcpyAsync((char*)h->in.dev + ins[i].off, (char*)h->in.host + ins[i].off, span, cudaMemcpyHostToDevice, s));
            i = j + 1;
        }
        return I7_OK;
    }
    i7_status flush_out(cudaStream_t s) {
        if (!nout && zc_pending) I7_CUDA(h, cudaStreamSynchronize(s));      // results land in the caller's memory
        else if (!nout && used_mirror) {                                      // asynchronous return: the mirror stays in use
            I7_CUDA(h, cudaEventRecord(h->mirror_ev, s));
            h->mirror_pending = true;
        }
        if (!nout) return I7_OK;
        for (int i = 0; i < nout && !direct;) {
            if (outs[i].pinned) {
                I7_CUDA(h, cudaMemcpyAsync(outs[i].user, (char*)h->out.dev + outs[i].off, outs[i].bytes, cudaMemcpyDeviceToHost, s));
                ++i;
                continue;
            }
            int j = i;
            while (j + 1 < nout && !outs[j + 1].pinned && outs[j + 1].off == outs[j].off + al(outs[j].bytes)) ++j;
            const size_t span = outs[j].off + outs[j].bytes - outs[i].off;
            I7_CUDA(h, cudaMemcpyAsync((char*)h->out.host + outs[i].off, (char*)h->out.dev + outs[i].off, span, cudaMemcpyDeviceToHost, s));
            i = j + 1;
        }
        I7_CUDA(h, cudaStreamSynchronize(s));
        for (int i = 0; i < nout; ++i)
            if (!outs[i].pinned) std::memcpy(outs[i].user, (char*)h->out.host + outs[i].off, outs[i].bytes);
        return I7_OK;
    }
};

template <typename T> const i7::Robot<T>& get_rob(const i7_handle* h);
template <> const i7::Robot<double>& get_rob<double>(const i7_handle* h) { return h->rob64; }
template <> const i7::Robot<float>& get_rob<float>(const i7_handle* h) { return h->rob32; }

template <typename T>
i7_status launch_sqp(i7_handle* h, const i7::Robot<T>& rob, i7::SolveParams<T>& P, cudaStream_t s, size_t first_problem = 0) {
    P.spill_ws = h->spill ? (T*)h->d_spill + first_problem * (size_t)i7::make_layout(h->cfg.N, true).reg_size : nullptr;
    cudaError_t e;
    if (h->use_cluster && P.mode == i7::MODE_SOLVE) {
        if (h->cluster_size == 8) e = i7::launch_sqp_cluster<T, 8>(rob, P, (T*)h->d_xbuf, h->cluster_smem_bytes, s);
        else if (h->cluster_size == 4) e = i7::launch_sqp_cluster<T, 4>(rob, P, (T*)h->d_xbuf, h->cluster_smem_bytes, s);
        else e = i7::launch_sqp_cluster<T, 2>(rob, P, (T*)h->d_xbuf, h->cluster_smem_bytes, s);
    }
    else if (h->threads == 384) e = h->spill ? i7::launch_sqp_cta<T, 384, true>(rob, P, h->smem_bytes, s) : i7::launch_sqp_cta<T, 384, false>(rob, P, h->smem_bytes, s);
    else if (h->threads == 256) e = h->spill ? i7::launch_sqp_cta<T, 256, true>(rob, P, h->smem_bytes, s) : i7::launch_sqp_cta<T, 256, false>(rob, P, h->smem_bytes, s);
    else e = h->spill ? i7::launch_sqp_cta<T, 128, true>(rob, P, h->smem_bytes, s) : i7::launch_sqp_cta<T, 128, false>(rob, P, h->smem_bytes, s);
    ++h->launches;
    I7_CUDA(h, e);
    return I7_OK;
}

template <typename T>
cudaError_t allow_big_smem(int threads, int optin_bytes, int* cap) {
    // the spill / no-spill instances of one CTA size have the same static footprint; raise the cap of both
    cudaError_t e;
    if (threads == 384) { e = i7::smem_cap_sqp_cta<T, 384, true>(optin_bytes, cap); if (e == cudaSuccess) e = i7::smem_cap_sqp_cta<T, 384, false>(optin_bytes, cap); }
    else if (threads == 256) { e = i7::smem_cap_sqp_cta<T, 256, true>(optin_bytes, cap); if (e == cudaSuccess) e = i7::smem_cap_sqp_cta<T, 256, false>(optin_bytes, cap); }
    else { e = i7::smem_cap_sqp_cta<T, 128, true>(optin_bytes, cap); if (e == cudaSuccess) e = i7::smem_cap_sqp_cta<T, 128, false>(optin_bytes, cap); }
    return e;
}

template <typename T>
void base_params(const i7_handle* h, i7::SolveParams<T>& P, int batch, int goal_stride, int mode) {
    std::memset(&P, 0, sizeof(P));
    const i7_config& c = h->cfg;
    P.N = c.N; P.batch = batch; P.goal_stride = goal_stride; P.mode = mode;
    P.max_qp = c.max_qp; P.regularize = c.regularize;
    P.dt = T(c.dt); P.dQ = T(c.dQ_cost); P.R = T(c.R_cost); P.QN = T(c.QN_cost); P.eps = T(c.eps);
    P.mu = T(c.mu); P.step_tol = T(c.step_tol);
    P.fext = h->has_fext ? (const T*)h->d_fext : nullptr;
    P.prof = (mode == i7::MODE_SOLVE) ? h->d_prof : nullptr;
    P.x0_into_xu = (mode == i7::MODE_SOLVE) ? h->opt_x0_into_xu : 0;
}

i7_status check_batch(i7_handle* h, int batch, int goal_stride) {
    if (!h) return I7_ERR_INVALID;
    if (batch < 1 || batch > h->cfg.batch) { h->err = "batch out of range for this handle"; return I7_ERR_INVALID; }
    if (goal_stride != 3 && goal_stride != 6) { h->err = "goal_stride must be 3 or 6"; return I7_ERR_INVALID; }
    return I7_OK;
}

// generic driver for every mode of the fused kernel
struct SqpCall {
    int mode, batch, goal_stride;
    const void *xu, *x0, *goals, *xu_full;
    void *xu_out, *stats, *z_out, *y_out, *Adata, *Pdata, *g, *l, *alpha, *merits;
};

template <typename T>
i7_status run_sqp_t(i7_handle* h, const SqpCall& c, cudaStream_t s, bool zero_copy = false) {
    const int N = h->cfg.N;
    const size_t nz = h->nz, B = c.batch, es = sizeof(T);
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    Stage st(h, c.mode == i7::MODE_SOLVE && c.batch <= h->direct_max_batch);
    st.zero_copy = zero_copy;
    i7::SolveParams<T> P;
    base_params<T>(h, P, c.batch, c.goal_stride, c.mode);
    P.xu_in = (const T*)st.in(c.xu, B * nz * es);
    P.x0 = (const T*)st.in(c.x0, B * 12 * es);
    P.goals = (const T*)st.in(c.goals, B * N * c.goal_stride * es);
    P.xu_full = (const T*)st.in(c.xu_full, B * nz * es);
    P.xu_out = (T*)st.out(c.xu_out, B * nz * es);
    P.stats = (T*)st.out(c.stats, B * I7_STATS_LEN * es);
    P.z_out = (T*)st.out(c.z_out, B * nz * es);
    P.y_out = (T*)st.out(c.y_out, B * 12 * N * es);
    P.Adata = (T*)st.out(c.Adata, B * (360 * (N - 1) + 144) * es);
    P.Pdata = (T*)st.out(c.Pdata, B * (27 * N + 6 * (N - 1)) * es);
    P.g = (T*)st.out(c.g, B * nz * es);
    P.l = (T*)st.out(c.l, B * 12 * N * es);
    P.alpha_out = (T*)st.out(c.alpha, B * es);
    P.merits_out = (T*)st.out(c.merits, B * (c.mode == i7::MODE_MERIT ? 6 : 9) * es);
    i7_status rc = st.flush_in(s);
    if (rc) return rc;
    const bool ev = c.mode == i7::MODE_SOLVE && !(h->no_events && st.direct);
    if (ev) I7_CUDA(h, cudaEventRecord(h->ev0, s));
    rc = launch_sqp<T>(h, get_rob<T>(h), P, s);
    if (rc) return rc;
    if (ev) { I7_CUDA(h, cudaEventRecord(h->ev1, s)); h->timed = true; }
    else if (c.mode == i7::MODE_SOLVE) h->timed = false;
    return st.flush_out(s);
}

// i7_solve with every buffer in HOST memory and a batch of several chunks: a software pipeline.  Chunk i's inputs go
// host -> device on pipe stream i mod 3, its kernel follows on the same stream, then its outputs device -> host; the
// three streams run concurrently, so the copy engines (both directions) and the SMs are busy at the same time and the
// tail wave of one chunk's kernel overlaps the head of the next.  Pageable buffers pass through the pinned mirror of
// the arena chunk by chunk (the host memcpy of chunk i+1 overlaps the GPU work on chunk i).  Synchronous, like every
// host-buffer call.
// problems per chunk.  Smaller chunks for smaller batches were measured and lose: every chunk costs ~25 us of host
// time to issue (copies, launch, events), which for 64-problem chunks is more than the exposed copies they save.
int pipe_chunk_for(const i7_handle* h, int /*batch*/) { return h->pipe_chunk; }

template <typename T>
i7_status run_solve_pipelined(i7_handle* h, const SqpCall& c, cudaStream_t s) {
    const int N = h->cfg.N;
    const size_t nz = h->nz, B = c.batch, es = sizeof(T);
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    const size_t per_in[3] = {nz * es, 12 * es, (size_t)N * c.goal_stride * es};
    const size_t per_out[2] = {nz * es, (size_t)I7_STATS_LEN * es};
    const void* uin[3] = {c.xu, c.x0, c.goals};
    void* uout[2] = {c.xu_out, c.stats};
    size_t off_in[3], off_out[2], o = 0;
    for (int a = 0; a < 3; ++a) { off_in[a] = o; o += (B * per_in[a] + 15) & ~size_t(15); }
    o = 0;
    for (int a = 0; a < 2; ++a) { off_out[a] = o; o += (B * per_out[a] + 15) & ~size_t(15); }
    bool pin_in[3], pin_out[2];
    for (int a = 0; a < 3; ++a) pin_in[a] = ptr_kind(uin[a]) == 2;
    for (int a = 0; a < 2; ++a) pin_out[a] = uout[a] && ptr_kind(uout[a]) == 2;
    const size_t chunk = (size_t)pipe_chunk_for(h, c.batch);
    const bool m2 = h->pipe_mode == 2;
    if (h->mirror_pending) { I7_CUDA(h, cudaEventSynchronize(h->mirror_ev)); h->mirror_pending = false; }
    const int nchunks = (int)((B + chunk - 1) / chunk);
    I7_CUDA(h, cudaEventRecord(h->ev0, s));
    I7_CUDA(h, cudaEventRecord(h->pipe_fork, s));
    for (int k = 0; k < h->pipe_nstreams; ++k) I7_CUDA(h, cudaStreamWaitEvent(h->pipe_stream[k], h->pipe_fork, 0));
    if (m2) {
        I7_CUDA(h, cudaStreamWaitEvent(h->pipe_in, h->pipe_fork, 0));
        I7_CUDA(h, cudaStreamWaitEvent(h->pipe_out, h->pipe_fork, 0));
    }
    for (int i = 0; i < nchunks; ++i) {
        cudaStream_t ps = h->pipe_stream[i % h->pipe_nstreams];
        cudaStream_t s_in = m2 ? h->pipe_in : ps, s_out = m2 ? h->pipe_out : ps;
        const size_t b0 = (size_t)i * chunk, nb = B - b0 < chunk ? B - b0 : chunk;
        for (int a = 0; a < 3; ++a) {
            char* dev = (char*)h->in.dev + off_in[a] + b0 * per_in[a];
            const char* src = (const char*)uin[a] + b0 * per_in[a];
            if (!pin_in[a]) {
                char* mir = (char*)h->in.host + off_in[a] + b0 * per_in[a];
                std::memcpy(mir, src, nb * per_in[a]);
                src = mir;
            }
            I7_CUDA(h, cudaMemcpyAsync(dev, src, nb * per_in[a], cudaMemcpyHostToDevice, s_in));
        }
        if (m2) {
            I7_CUDA(h, cudaEventRecord(h->pipe_in_done[i], s_in));
            I7_CUDA(h, cudaStreamWaitEvent(ps, h->pipe_in_done[i], 0));
        }
        i7::SolveParams<T> P;
        base_params<T>(h, P, (int)nb, c.goal_stride, c.mode);
        if (P.fext) P.fext += b0 * 6;
        if (P.prof) P.prof += b0 * 32;
        P.xu_in = (const T*)((char*)h->in.dev + off_in[0] + b0 * per_in[0]);
        P.x0 = (const T*)((char*)h->in.dev + off_in[1] + b0 * per_in[1]);
        P.goals = (const T*)((char*)h->in.dev + off_in[2] + b0 * per_in[2]);
        P.xu_out = (T*)((char*)h->out.dev + off_out[0] + b0 * per_out[0]);
        P.stats = uout[1] ? (T*)((char*)h->out.dev + off_out[1] + b0 * per_out[1]) : nullptr;
        i7_status rc = launch_sqp<T>(h, get_rob<T>(h), P, ps, b0);
        if (rc) return rc;
        if (m2) {
            I7_CUDA(h, cudaEventRecord(h->pipe_k_done[i], ps));
            I7_CUDA(h, cudaStreamWaitEvent(s_out, h->pipe_k_done[i], 0));
        }
        for (int a = 0; a < 2; ++a) {
            if (!uout[a]) continue;
            const char* dev = (const char*)h->out.dev + off_out[a] + b0 * per_out[a];
            char* dst = pin_out[a] ? (char*)uout[a] + b0 * per_out[a] : (char*)h->out.host + off_out[a] + b0 * per_out[a];
            I7_CUDA(h, cudaMemcpyAsync(dst, dev, nb * per_out[a], cudaMemcpyDeviceToHost, s_out));
        }
        I7_CUDA(h, cudaEventRecord(h->pipe_done[i], s_out));
    }
    for (int k = 0; k < h->pipe_nstreams; ++k) {
        I7_CUDA(h, cudaEventRecord(h->pipe_join[k], h->pipe_stream[k]));
        I7_CUDA(h, cudaStreamWaitEvent(s, h->pipe_join[k], 0));
    }
    if (m2) {       // every kernel is awaited by the copy-out stream; the copy-in stream by the kernels
        I7_CUDA(h, cudaStreamWaitEvent(s, h->pipe_done[nchunks - 1], 0));
    }
    I7_CUDA(h, cudaEventRecord(h->ev1, s));
    h->timed = true;
    // pageable outputs: unpack chunk by chunk as they land
    for (int i = 0; i < nchunks; ++i) {
        const size_t b0 = (size_t)i * chunk, nb = B - b0 < chunk ? B - b0 : chunk;
        bool any = false;
        for (int a = 0; a < 2; ++a) any = any || (uout[a] && !pin_out[a]);
        if (!any) break;
        I7_CUDA(h, cudaEventSynchronize(h->pipe_done[i]));
        for (int a = 0; a < 2; ++a)
            if (uout[a] && !pin_out[a])
                std::memcpy((char*)uout[a] + b0 * per_out[a], (char*)h->out.host + off_out[a] + b0 * per_out[a], nb * per_out[a]);
    }
    I7_CUDA(h, cudaStreamSynchronize(s));
    return I7_OK;
}

bool pipeline_eligible(const i7_handle* h, const SqpCall& c) {
    if (c.mode != i7::MODE_SOLVE || h->pipe_chunk <= 0 || h->use_cluster) return false;
    const int chunk = pipe_chunk_for(h, c.batch);
    if (c.batch < 2 * chunk || (c.batch + chunk - 1) / chunk > i7_handle::kPipeMaxChunks) return false;
    if (c.xu_full || c.z_out || c.y_out || c.Adata || c.Pdata || c.g || c.l || c.alpha || c.merits) return false;
    const void* ptrs[5] = {c.xu, c.x0, c.goals, c.xu_out, c.stats};
    for (int a = 0; a < 5; ++a)
        if (ptrs[a] && ptr_kind(ptrs[a]) == 1) return false;      // device-resident buffers: one launch, no copies
    return true;
}

// qp_mode "osqp": an SQP pass is three launches — linearise + export the QP blocks (MODE_EXPORT of the fused kernel),
// ADMM (i7_admm.cuh), line search + update (MODE_STEP) — all on the caller's stream, the per-problem "loop ended" flags
// stay on the device.  Host buffers go through the arena copies.
i7_status run_osqp(i7_handle* h, const SqpCall& c, cudaStream_t s) {
    using T = double;
    const int N = h->cfg.N;
    const size_t nz = h->nz, B = c.batch, es = sizeof(T);
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    Stage st(h);
    const T* d_xu = (const T*)st.in(c.xu, B * nz * es);
    const T* d_x0 = (const T*)st.in(c.x0, B * 12 * es);
    const T* d_goals = (const T*)st.in(c.goals, B * N * c.goal_stride * es);
    T* d_out = (T*)st.out(c.xu_out, B * nz * es);
    T* d_stats = (T*)st.out(c.stats, B * I7_STATS_LEN * es);
    T* d_z = (T*)st.out(c.z_out, B * nz * es);
    T* d_y = (T*)st.out(c.y_out, B * 12 * N * es);
    i7_status rc = st.flush_in(s);
    if (rc) return rc;
    auto admm = [&](const T* xu_lin, T* z_out, T* y_out, const int* done) -> i7_status {
        i7::AdmmParams<T> A;
        std::memset(&A, 0, sizeof(A));
        A.blk = (const T*)h->d_blk; A.xu = xu_lin; A.x0 = d_x0;
        A.wx = (T*)h->d_wx; A.wy = (T*)h->d_wy; A.wyl = (T*)h->d_wyl;
        A.z_out = z_out; A.y_out = y_out; A.info = (T*)h->d_qpinfo; A.fac = (T*)h->d_fac; A.done = done;
        A.N = N; A.batch = c.batch; A.dt = h->cfg.dt; A.S = h->admm; A.prof = h->d_prof;
        I7_CUDA(h, i7::launch_admm(A, h->admm_smem_bytes, s));
        ++h->launches;
        return I7_OK;
    };
    auto fused = [&](int mode, const T* xu_in, const T* xu_full, T* xu_out, T* stats, int* done, int qp_index) -> i7_status {
        i7::SolveParams<T> P;
        base_params<T>(h, P, c.batch, c.goal_stride, mode);
        P.xu_in = xu_in; P.x0 = d_x0; P.goals = d_goals; P.xu_full = xu_full; P.xu_out = xu_out; P.stats = stats;
        P.blk_out = (T*)h->d_blk; P.qp_info = (const T*)h->d_qpinfo; P.done = done; P.qp_index = qp_index;
        return launch_sqp<T>(h, get_rob<T>(h), P, s);
    };
    if (c.mode == i7::MODE_QP) {
        if ((rc = fused(i7::MODE_EXPORT, d_xu, nullptr, nullptr, nullptr, nullptr, 0))) return rc;
        if ((rc = admm(d_xu, d_z, d_y, nullptr))) return rc;
        return st.flush_out(s);
    }
    T* stats = d_stats ? d_stats : (T*)h->d_stats;
    I7_CUDA(h, cudaEventRecord(h->ev0, s));
    if (d_out != d_xu) I7_CUDA(h, cudaMemcpyAsync(d_out, d_xu, B * nz * es, cudaMemcpyDeviceToDevice, s));
    if (h->opt_x0_into_xu)       // I7_OPT_X0_INTO_WARM_START: XU[:12] = xcur (src/osqp_mpc.py:69)
        I7_CUDA(h, cudaMemcpy2DAsync(d_out, nz * es, d_x0, 12 * es, 12 * es, B, cudaMemcpyDeviceToDevice, s));
    I7_CUDA(h, cudaMemsetAsync(h->d_done, 0, B * sizeof(int), s));
    for (int qp = 0; qp < h->cfg.max_qp; ++qp) {
        if ((rc = fused(i7::MODE_EXPORT, d_out, nullptr, nullptr, nullptr, h->d_done, qp))) return rc;
        if ((rc = admm(d_out, (T*)h->d_zbuf, nullptr, h->d_done))) return rc;
        if ((rc = fused(i7::MODE_STEP, d_out, (const T*)h->d_zbuf, d_out, stats, h->d_done, qp))) return rc;
    }
    I7_CUDA(h, cudaEventRecord(h->ev1, s));
    h->timed = true;
    return st.flush_out(s);
}

i7_status run_sqp(i7_handle* h, const SqpCall& c, void* stream) {
    i7_status rc = check_batch(h, c.batch, c.goal_stride);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (h->cfg.qp_mode == I7_QP_OSQP && (c.mode == i7::MODE_SOLVE || c.mode == i7::MODE_QP)) return run_osqp(h, c, s);
    if (h->zero_copy && c.mode == i7::MODE_SOLVE && c.batch > h->direct_max_batch) {
        // every host buffer of the call page-locked: one launch straight on the caller's memory
        const void* ptrs[5] = {c.xu, c.x0, c.goals, c.xu_out, c.stats};
        bool all = true, any = false;       // no pageable buffer, at least one page-locked (the rest on the device)
        for (int a = 0; a < 5; ++a) {
            const int k = ptrs[a] ? ptr_kind(ptrs[a]) : 1;
            all = all && k != 0;
            any = any || k == 2;
        }
        if (all && any) return h->cfg.dtype == I7_F64 ? run_sqp_t<double>(h, c, s, true) : run_sqp_t<float>(h, c, s, true);
    }
    if (pipeline_eligible(h, c))
        return h->cfg.dtype == I7_F64 ? run_solve_pipelined<double>(h, c, s) : run_solve_pipelined<float>(h, c, s);
    return h->cfg.dtype == I7_F64 ? run_sqp_t<double>(h, c, s) : run_sqp_t<float>(h, c, s);
}

}  // namespace

extern "C" {

const char* i7_version(void) { return "indy7_b200 0.1.0 (sm_100a)"; }

void i7_default_config(i7_config* cfg) {
    std::memset(cfg, 0, sizeof(*cfg));
    cfg->N = 32; cfg->batch = 1; cfg->dtype = I7_F64; cfg->device = 0;
    cfg->dt = 0.01; cfg->dQ_cost = 0.01; cfg->R_cost = 1e-5; cfg->QN_cost = 100.0; cfg->eps = 1.0;
    cfg->regularize = 1; cfg->max_qp = 2; cfg->mu = 10.0; cfg->step_tol = 1e-3; cfg->threads = 0;
    cfg->qp_mode = I7_QP_EXACT; cfg->limits = 0;
    cfg->admm_rho = 0.1; cfg->admm_rho_u = 1e-3; cfg->admm_sigma = 1e-6; cfg->admm_alpha = 1.6;
    cfg->admm_eps_abs = 1e-3; cfg->admm_eps_rel = 1e-3; cfg->admm_max_iter = 4000; cfg->admm_check = 25;
    cfg->model.gravity[2] = -9.81;
    for (int i = 0; i < I7_NJ; ++i) { cfg->model.q_min[i] = -1e30; cfg->model.q_max[i] = 1e30; cfg->model.v_max[i] = 1e30; cfg->model.u_max[i] = 1e30; }
}

const char* i7_last_error(const i7_handle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

void i7_destroy(i7_handle* h) {
    if (!h) return;
    cudaSetDevice(h->cfg.device);
    if (h->in.dev) cudaFree(h->in.dev);
    if (h->in.host) cudaFreeHost(h->in.host);
    if (h->out.dev) cudaFree(h->out.dev);
    if (h->out.host) cudaFreeHost(h->out.host);
    if (h->d_fext) cudaFree(h->d_fext);
    if (h->d_spill) cudaFree(h->d_spill);
    if (h->d_xbuf) cudaFree(h->d_xbuf);
    if (h->d_prof) cudaFree(h->d_prof);
    for (void* p : {h->d_wx, h->d_wy, h->d_wyl, h->d_blk, h->d_fac, h->d_zbuf, h->d_qpinfo, h->d_stats, (void*)h->d_done})
        if (p) cudaFree(p);
    for (int r = 0; r < h->p2p_world; ++r)
        if (h->p2p_peer[r] && r != h->p2p_rank) cudaIpcCloseMemHandle(h->p2p_peer[r]);
    if (h->p2p_mine) cudaFree(h->p2p_mine);
    if (h->p2p_table) cudaFree(h->p2p_table);
    if (h->p2p_err) cudaFree(h->p2p_err);
    if (h->d_xchg_ticket) cudaFree(h->d_xchg_ticket);
    if (h->d_xchg_wait) cudaFree(h->d_xchg_wait);
    if (h->d_sel_counter) cudaFree(h->d_sel_counter);
    if (h->d_sel_errs) cudaFree(h->d_sel_errs);
    for (int k = 0; k < i7_handle::kPipeStreams; ++k) {
        if (h->pipe_stream[k]) cudaStreamDestroy(h->pipe_stream[k]);
        if (h->pipe_join[k]) cudaEventDestroy(h->pipe_join[k]);
    }
    for (int k = 0; k < i7_handle::kPipeMaxChunks; ++k) if (h->pipe_done[k]) cudaEventDestroy(h->pipe_done[k]);
    if (h->pipe_fork) cudaEventDestroy(h->pipe_fork);
    if (h->pipe_in) cudaStreamDestroy(h->pipe_in);
    if (h->pipe_out) cudaStreamDestroy(h->pipe_out);
    for (int k = 0; k < i7_handle::kPipeMaxChunks; ++k) {
        if (h->pipe_in_done[k]) cudaEventDestroy(h->pipe_in_done[k]);
        if (h->pipe_k_done[k]) cudaEventDestroy(h->pipe_k_done[k]);
    }
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->mirror_ev) cudaEventDestroy(h->mirror_ev);
    delete h;
}

i7_status i7_create(const i7_config* cfg, i7_handle** out) {
    if (!cfg || !out) { g_create_error = "null argument"; return I7_ERR_INVALID; }
    *out = nullptr;
    if (cfg->N < 2 || cfg->N > 512 || cfg->batch < 1 || (cfg->dtype != I7_F64 && cfg->dtype != I7_F32) ||
        cfg->max_qp < 1 || cfg->max_qp > I7_MAX_QP || !(cfg->dt > 0)) {
        g_create_error = "invalid configuration (N in [2,512], batch >= 1, dtype f64|f32, max_qp in [1,2], dt > 0)";
        return I7_ERR_INVALID;
    }
    if ((cfg->qp_mode != I7_QP_EXACT && cfg->qp_mode != I7_QP_OSQP) || (cfg->limits && cfg->qp_mode != I7_QP_OSQP)) {
        g_create_error = "qp_mode must be I7_QP_EXACT or I7_QP_OSQP; limits need I7_QP_OSQP (the exact solve is equality-only)";
        return I7_ERR_INVALID;
    }
    if (cfg->qp_mode == I7_QP_OSQP) {
        if (cfg->dtype != I7_F64) { g_create_error = "I7_QP_OSQP runs in f64 only (sigma = 1e-6 against rho = 100)"; return I7_ERR_UNSUPPORTED; }
        if (!(cfg->admm_rho > 0) || !(cfg->admm_rho_u > 0) || !(cfg->admm_sigma > 0) || !(cfg->admm_alpha > 0 && cfg->admm_alpha < 2) ||
            cfg->admm_max_iter < 1 || cfg->admm_check < 1) {
            g_create_error = "invalid ADMM settings (rho, rho_u, sigma > 0; 0 < alpha < 2; max_iter, check >= 1)";
            return I7_ERR_INVALID;
        }
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) +
                         " (indy7_b200 has no CPU fallback)";
        cudaGetLastError();
        return I7_ERR_CUDA;
    }
    if (cfg->device < 0 || cfg->device >= ndev) { g_create_error = "device ordinal out of range"; return I7_ERR_INVALID; }
    i7_handle* h = new i7_handle();
    h->cfg = *cfg;
    h->esize = cfg->dtype == I7_F64 ? 8 : 4;
    h->nz = 18 * (size_t)cfg->N - 6;
    fill_robot<double>(cfg->model, h->rob64);
    fill_robot<float>(cfg->model, h->rob32);
    auto fail = [&](i7_status rc) { g_create_error = h->err; i7_destroy(h); return rc; };
#define I7_TRY(call) do { i7_status rc_ = (call); if (rc_) return fail(rc_); } while (0)
#define I7_CUDA_C(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { h->err = std::string(#call) + ": " + cudaGetErrorString(e_); return fail(I7_ERR_CUDA); } } while (0)
    I7_CUDA_C(cudaSetDevice(cfg->device));
    I7_CUDA_C(cudaDeviceGetAttribute(&h->max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, cfg->device));
    bool want_spill = false;
    // threads per CTA of the throughput form: 384 while every problem has an SM to itself; for bigger batches 128 when
    // shared memory lets two or more CTAs share an SM (more independent problems per SM beats wider CTAs: measured
    // 1.60 -> 1.27 ms for 4096 fp32 problems, four CTAs per SM), 256 when only one fits (long horizons)
    {
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, cfg->device);
        const size_t per_sm = (size_t)h->max_smem_optin + 1024;
        const size_t need_n = (size_t)i7::make_layout(cfg->N, false).total * h->esize + 1024;
        const size_t need_s = (size_t)i7::make_layout(cfg->N, true).total * h->esize + 1024;
        const int fit_n = (int)(per_sm / need_n), fit_s = (int)(per_sm / need_s);
        // records in shared memory (fit_n CTAs per SM) or in the global workspace with the shared-memory tile (fit_s):
        // the tile form when the records do not fit at all, and for throughput batches when the resident form would leave
        // ONE CTA per SM while the tile form runs three or more (fp32 N=128, four per SM: 1.30 -> 0.90 ms for 512 problems;
        // with only two — fp64 N=64 — the resident 256-thread form stays ahead: 0.81 vs 0.85 ms)
        want_spill = fit_n == 0 || (cfg->batch > sms && fit_n < 2 && fit_s >= 3);
        if (const char* e7 = std::getenv("I7_FORCE_SPILL")) want_spill = (std::atoi(e7) != 0 || fit_n == 0) && fit_s >= 1;
        const int fit = want_spill ? fit_s : fit_n;
        h->threads = cfg->threads ? cfg->threads : (cfg->batch <= sms ? 384 : (fit >= 2 ? 128 : 256));
    }
    if (h->threads != 128 && h->threads != 256 && h->threads != 384) { h->err = "threads must be 0, 128, 256 or 384"; return fail(I7_ERR_INVALID); }
    int cap = 0;
    {
        cudaError_t ea;
        const int o = h->max_smem_optin;
        ea = cfg->dtype == I7_F64 ? allow_big_smem<double>(h->threads, o, &cap) : allow_big_smem<float>(h->threads, o, &cap);
        if (ea != cudaSuccess) { h->err = std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(ea); return fail(I7_ERR_CUDA); }
    }
    i7::Layout L = i7::make_layout(cfg->N, false);
    h->spill = false;
    if (want_spill || (size_t)L.total * h->esize > (size_t)cap) {
        L = i7::make_layout(cfg->N, true);
        h->spill = true;
        if ((size_t)L.total * h->esize > (size_t)cap) {
            h->err = "horizon too long for the shared-memory layout of this dtype";
            return fail(I7_ERR_UNSUPPORTED);
        }
        I7_CUDA_C(cudaMalloc(&h->d_spill, (size_t)cfg->batch * L.reg_size * h->esize));
    }
    h->smem_bytes = L.total * h->esize;
    // latency regime: a cluster of 8, 4 or 2 CTAs per problem — the largest cluster size for which every cluster of the
    // batch is resident at once (clusters are placed per GPC, so that is fewer than SMs / size; a second wave of
    // clusters is slower than the CTA form).  cfg->cluster: 0 automatic, 1 CTA form, 2 | 4 | 8 forced.
    {
        if (cfg->cluster != 0 && cfg->cluster != 1 && cfg->cluster != 2 && cfg->cluster != 4 && cfg->cluster != 8) {
            h->err = "cluster must be 0, 1, 2, 4 or 8"; return fail(I7_ERR_INVALID);
        }
        const bool f64 = cfg->dtype == I7_F64;
        const int sizes[3] = {8, 4, 2};
        for (int si = 0; si < 3 && !h->use_cluster && cfg->cluster != 1 && cfg->qp_mode == I7_QP_EXACT; ++si) {
            const int cl = sizes[si];
            if (cfg->cluster != 0 && cfg->cluster != cl) continue;
            int ccap = 0, nact = 0;
            cudaError_t ec = cl == 8 ? (f64 ? i7::smem_cap_sqp_cluster<double, 8>(h->max_smem_optin, &ccap) : i7::smem_cap_sqp_cluster<float, 8>(h->max_smem_optin, &ccap))
                           : cl == 4 ? (f64 ? i7::smem_cap_sqp_cluster<double, 4>(h->max_smem_optin, &ccap) : i7::smem_cap_sqp_cluster<float, 4>(h->max_smem_optin, &ccap))
                                     : (f64 ? i7::smem_cap_sqp_cluster<double, 2>(h->max_smem_optin, &ccap) : i7::smem_cap_sqp_cluster<float, 2>(h->max_smem_optin, &ccap));
            if (ec != cudaSuccess) { h->err = std::string("cudaFuncSetAttribute(cluster): ") + cudaGetErrorString(ec); return fail(I7_ERR_CUDA); }
            const int bytes = i7::cluster_smem_len(cfg->N, cl) * h->esize;
            if (bytes > ccap) {
                if (cfg->cluster == cl) { h->err = "horizon too long for the cluster kernel"; return fail(I7_ERR_UNSUPPORTED); }
                continue;
            }
            if (cfg->cluster == 0) {
                cudaError_t eo = cl == 8 ? (f64 ? i7::max_active_sqp_clusters<double, 8>(bytes, &nact) : i7::max_active_sqp_clusters<float, 8>(bytes, &nact))
                               : cl == 4 ? (f64 ? i7::max_active_sqp_clusters<double, 4>(bytes, &nact) : i7::max_active_sqp_clusters<float, 4>(bytes, &nact))
                                         : (f64 ? i7::max_active_sqp_clusters<double, 2>(bytes, &nact) : i7::max_active_sqp_clusters<float, 2>(bytes, &nact));
                if (eo != cudaSuccess) { (void)cudaGetLastError(); nact = 0; }
                if (cfg->batch > nact) continue;
            }
            h->use_cluster = true; h->cluster_size = cl; h->cluster_smem_bytes = bytes;
        }
        if (h->use_cluster) I7_CUDA_C(cudaMalloc(&h->d_xbuf, (size_t)cfg->batch * i7::cluster_xbuf_len(cfg->N) * h->esize));
    }
    if (cfg->qp_mode == I7_QP_OSQP) {
        const size_t B = cfg->batch, N = cfg->N;
        int acap = 0;
        I7_CUDA_C(i7::smem_cap_admm(h->max_smem_optin, &acap));
        h->admm_smem_bytes = i7::admm_smem_elems(cfg->N) * (int)sizeof(double);
        if (h->admm_smem_bytes > acap) { h->err = "horizon too long for the ADMM kernel's shared-memory vectors"; return fail(I7_ERR_UNSUPPORTED); }
        i7::AdmmSettings& S = h->admm;
        S.rho_eq = 1e3 * cfg->admm_rho; S.rho_x = cfg->admm_rho; S.rho_u = cfg->admm_rho_u; S.sigma = cfg->admm_sigma;
        S.alpha = cfg->admm_alpha; S.eps_abs = cfg->admm_eps_abs; S.eps_rel = cfg->admm_eps_rel;
        S.max_iter = cfg->admm_max_iter; S.check = cfg->admm_check; S.limits = cfg->limits ? 1 : 0;
        for (int i = 0; i < 6; ++i) { S.qmin[i] = cfg->model.q_min[i]; S.qmax[i] = cfg->model.q_max[i]; S.vmax[i] = cfg->model.v_max[i]; S.umax[i] = cfg->model.u_max[i]; }
        I7_CUDA_C(cudaMalloc(&h->d_wx, B * 18 * N * 8));
        I7_CUDA_C(cudaMalloc(&h->d_wy, B * 12 * N * 8));
        I7_CUDA_C(cudaMalloc(&h->d_wyl, B * 18 * N * 8));
        I7_CUDA_C(cudaMemset(h->d_wx, 0, B * 18 * N * 8));
        I7_CUDA_C(cudaMemset(h->d_wy, 0, B * 12 * N * 8));
        I7_CUDA_C(cudaMemset(h->d_wyl, 0, B * 18 * N * 8));
        I7_CUDA_C(cudaMalloc(&h->d_blk, B * N * i7::QB_STRIDE * 8));
        I7_CUDA_C(cudaMalloc(&h->d_fac, B * N * (size_t)i7::admm_fac_stride() * 8));
        I7_CUDA_C(cudaMalloc(&h->d_zbuf, B * h->nz * 8));
        I7_CUDA_C(cudaMalloc(&h->d_qpinfo, B * 4 * 8));
        I7_CUDA_C(cudaMemset(h->d_qpinfo, 0, B * 4 * 8));
        I7_CUDA_C(cudaMalloc(&h->d_stats, B * I7_STATS_LEN * 8));
        I7_CUDA_C(cudaMalloc((void**)&h->d_done, B * sizeof(int)));
    }
    const size_t B = cfg->batch, N = cfg->N, es = h->esize;
    // worst case over all entry points: inputs xu, x0, goals(6), xu_full; outputs of the dump mode
    const size_t in_bytes = B * (2 * h->nz + 12 + 6 * N) * es + 256;
    const size_t out_bytes = B * ((360 * (N - 1) + 144) + (33 * N) + h->nz + 12 * N + h->nz + 64) * es + 256;
    I7_TRY(arena_reserve(h, h->in, in_bytes));
    I7_TRY(arena_reserve(h, h->out, out_bytes));
    I7_CUDA_C(cudaMalloc(&h->d_fext, B * 6 * es));
    I7_CUDA_C(cudaMemset(h->d_fext, 0, B * 6 * es));
    I7_CUDA_C(cudaEventCreate(&h->ev0));
    I7_CUDA_C(cudaEventCreate(&h->ev1));
    I7_CUDA_C(cudaEventCreateWithFlags(&h->mirror_ev, cudaEventDisableTiming));
    {
        const char* env = std::getenv("I7_DIRECT_MAX_BATCH");
        int can_map = 0;
        cudaDeviceGetAttribute(&can_map, cudaDevAttrCanMapHostMemory, cfg->device);
        h->direct_max_batch = can_map ? (env ? std::atoi(env) : 8) : 0;
        if (const char* e3 = std::getenv("I7_TIMING_EVENTS")) h->no_events = std::atoi(e3) == 0;
        if (const char* e6 = std::getenv("I7_ZERO_COPY")) h->zero_copy = std::atoi(e6) != 0;
        if (!can_map) h->zero_copy = 0;
    }
    // chunked host pipeline for big batches (I7_PIPE_CHUNK problems per chunk; 0 disables)
    {
        const char* env = std::getenv("I7_PIPE_CHUNK");
        h->pipe_chunk = env ? std::atoi(env) : 256;
        if (const char* e5 = std::getenv("I7_PIPE_MODE")) h->pipe_mode = std::atoi(e5) == 1 ? 1 : 2;
        if (const char* e4 = std::getenv("I7_PIPE_STREAMS")) {
            const int n = std::atoi(e4);
            h->pipe_nstreams = n < 1 ? 1 : (n > i7_handle::kPipeStreams ? i7_handle::kPipeStreams : n);
        }
        if (h->pipe_chunk > 0 && cfg->batch >= 2 * h->pipe_chunk) {
            I7_CUDA_C(cudaStreamCreateWithFlags(&h->pipe_in, cudaStreamNonBlocking));
            I7_CUDA_C(cudaStreamCreateWithFlags(&h->pipe_out, cudaStreamNonBlocking));
            for (int k = 0; k < i7_handle::kPipeMaxChunks; ++k) {
                I7_CUDA_C(cudaEventCreateWithFlags(&h->pipe_in_done[k], cudaEventDisableTiming));
                I7_CUDA_C(cudaEventCreateWithFlags(&h->pipe_k_done[k], cudaEventDisableTiming));
            }
            for (int k = 0; k < i7_handle::kPipeStreams; ++k) {
                I7_CUDA_C(cudaStreamCreateWithFlags(&h->pipe_stream[k], cudaStreamNonBlocking));
                I7_CUDA_C(cudaEventCreateWithFlags(&h->pipe_join[k], cudaEventDisableTiming));
            }
            I7_CUDA_C(cudaEventCreateWithFlags(&h->pipe_fork, cudaEventDisableTiming));
            for (int k = 0; k < i7_handle::kPipeMaxChunks; ++k) I7_CUDA_C(cudaEventCreateWithFlags(&h->pipe_done[k], cudaEventDisableTiming));
        } else {
            h->pipe_chunk = 0;
        }
    }
#undef I7_TRY
#undef I7_CUDA_C
    *out = h;
    return I7_OK;
}

i7_status i7_set_weights(i7_handle* h, double dQ_cost, double R_cost, double QN_cost, int32_t regularize, double eps) {
    if (!h) return I7_ERR_INVALID;
    h->cfg.dQ_cost = dQ_cost; h->cfg.R_cost = R_cost; h->cfg.QN_cost = QN_cost;
    h->cfg.regularize = regularize; h->cfg.eps = eps;
    return I7_OK;
}

i7_status i7_set_option(i7_handle* h, int32_t option, int32_t value) {
    if (!h) return I7_ERR_INVALID;
    switch (option) {
        case I7_OPT_X0_INTO_WARM_START: h->opt_x0_into_xu = value != 0; return I7_OK;
        default: h->err = "unknown option"; return I7_ERR_INVALID;
    }
}

i7_status i7_set_external_wrench(i7_handle* h, const double* f_ext, int32_t batch) {
    if (!h) return I7_ERR_INVALID;
    if (!f_ext) { h->has_fext = false; return I7_OK; }
    if (batch < 1 || batch > h->cfg.batch) { h->err = "batch out of range for this handle"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    // Pinocchio [linear; angular] -> Featherstone [angular; linear], converted to the handle dtype
    const size_t n = (size_t)batch * 6;
    std::string buf(n * h->esize, '\0');
    for (int b = 0; b < batch; ++b)
        for (int a = 0; a < 6; ++a) {
            const double v = f_ext[6 * b + (a < 3 ? a + 3 : a - 3)];
            if (h->esize == 8) ((double*)buf.data())[6 * b + a] = v; else ((float*)buf.data())[6 * b + a] = (float)v;
        }
    I7_CUDA(h, cudaMemcpy(h->d_fext, buf.data(), buf.size(), cudaMemcpyHostToDevice));
    h->has_fext = true;
    return I7_OK;
}

i7_status i7_reset(i7_handle* h) {
    if (!h) return I7_ERR_INVALID;
    h->timed = false;
    if (h->cfg.qp_mode == I7_QP_OSQP) {      // OSQP object re-created: primal and dual warm start gone
        const size_t B = h->cfg.batch, N = h->cfg.N;
        I7_CUDA(h, cudaSetDevice(h->cfg.device));
        I7_CUDA(h, cudaDeviceSynchronize());
        I7_CUDA(h, cudaMemset(h->d_wx, 0, B * 18 * N * 8));
        I7_CUDA(h, cudaMemset(h->d_wy, 0, B * 12 * N * 8));
        I7_CUDA(h, cudaMemset(h->d_wyl, 0, B * 18 * N * 8));
    }
    return I7_OK;
}
i7_status i7_reset_rho(i7_handle* h) {
    if (!h) return I7_ERR_INVALID;
    if (h->cfg.qp_mode == I7_QP_OSQP) { h->admm.rho_eq = 1e3 * h->cfg.admm_rho; h->admm.rho_x = h->cfg.admm_rho; h->admm.rho_u = h->cfg.admm_rho_u; }
    return I7_OK;
}
i7_status i7_reset_lambda(i7_handle* h) {
    if (!h) return I7_ERR_INVALID;
    if (h->cfg.qp_mode == I7_QP_OSQP) {
        const size_t B = h->cfg.batch, N = h->cfg.N;
        I7_CUDA(h, cudaSetDevice(h->cfg.device));
        I7_CUDA(h, cudaDeviceSynchronize());
        I7_CUDA(h, cudaMemset(h->d_wy, 0, B * 12 * N * 8));
        I7_CUDA(h, cudaMemset(h->d_wyl, 0, B * 18 * N * 8));
    }
    return I7_OK;
}

i7_status i7_solve(i7_handle* h, int32_t batch, const void* xu_in, const void* x0, const void* goals,
                   int32_t goal_stride, void* xu_out, void* stats, void* stream) {
    if (!h || !xu_in || !x0 || !goals || !xu_out) { if (h) h->err = "null argument"; return I7_ERR_INVALID; }
    SqpCall c{};
    c.mode = i7::MODE_SOLVE; c.batch = batch; c.goal_stride = goal_stride;
    c.xu = xu_in; c.x0 = x0; c.goals = goals; c.xu_out = xu_out; c.stats = stats;
    return run_sqp(h, c, stream);
}

i7_status i7_last_solve_time_us(i7_handle* h, double* us) {
    if (!h || !us) return I7_ERR_INVALID;
    if (!h->timed) { h->err = "no solve has been timed yet"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaEventSynchronize(h->ev1));
    float ms = 0.f;
    I7_CUDA(h, cudaEventElapsedTime(&ms, h->ev0, h->ev1));
    *us = 1e3 * ms;
    return I7_OK;
}

int64_t i7_launch_count(const i7_handle* h) { return h ? h->launches : 0; }

i7_status i7_kernel_form(const i7_handle* h, int32_t* threads, int32_t* ctas_per_problem, int32_t* smem_bytes) {
    if (!h) return I7_ERR_INVALID;
    if (threads) *threads = h->use_cluster ? i7::kClusterThreads : h->threads;
    if (ctas_per_problem) *ctas_per_problem = h->use_cluster ? h->cluster_size : 1;
    if (smem_bytes) *smem_bytes = h->use_cluster ? h->cluster_smem_bytes : h->smem_bytes;
    return I7_OK;
}

i7_status i7_pack_best(i7_handle* h, int32_t batch, const void* stats, const void* xu, int32_t index_offset,
                       void* packed, void* stream) {
    if (!h || !stats || !xu || !packed || batch < 1 || batch > h->cfg.batch) { if (h) h->err = "bad argument"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    const size_t es = h->esize, nz = h->nz;
    Stage st(h);
    const void* d_st = st.in(stats, (size_t)batch * I7_STATS_LEN * es);
    const void* d_xu = st.in(xu, (size_t)batch * nz * es);
    void* d_out = st.out(packed, ((size_t)batch + 2 + nz) * es);
    i7_status rc = st.flush_in(s);
    if (rc) return rc;
    if (h->esize == 8) i7::pack_best_kernel<double><<<1, 256, 0, s>>>(batch, (int)nz, (const double*)d_st, I7_STATS_LEN, 8, (const double*)d_xu, index_offset, (double*)d_out);
    else i7::pack_best_kernel<float><<<1, 256, 0, s>>>(batch, (int)nz, (const float*)d_st, I7_STATS_LEN, 8, (const float*)d_xu, index_offset, (float*)d_out);
    ++h->launches;
    I7_CUDA(h, cudaGetLastError());
    return st.flush_out(s);
}

i7_status i7_reduce_best(i7_handle* h, int32_t nranks, int32_t stride, const void* gathered, void* out, void* stream) {
    if (!h || !gathered || !out || nranks < 1 || stride < (int)h->nz + 2) { if (h) h->err = "bad argument"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    const size_t es = h->esize, nz = h->nz;
    i7_status rc;
    if ((rc = arena_reserve(h, h->in, (size_t)nranks * stride * es + 256))) return rc;
    Stage st(h);
    const void* d_in = st.in(gathered, (size_t)nranks * stride * es);
    void* d_out = st.out(out, (2 + nz) * es);
    if ((rc = st.flush_in(s))) return rc;
    if (h->esize == 8) i7::reduce_best_kernel<double><<<1, 256, 0, s>>>(nranks, stride, (int)nz, (const double*)d_in, (double*)d_out);
    else i7::reduce_best_kernel<float><<<1, 256, 0, s>>>(nranks, stride, (int)nz, (const float*)d_in, (float*)d_out);
    ++h->launches;
    I7_CUDA(h, cudaGetLastError());
    return st.flush_out(s);
}

i7_status i7_profile_phases(i7_handle* h, int32_t enable, int64_t* clocks16, int32_t batch) {
    if (!h) return I7_ERR_INVALID;
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    if (enable && !h->d_prof) {
        I7_CUDA(h, cudaMalloc((void**)&h->d_prof, (size_t)h->cfg.batch * 32 * sizeof(long long)));
        I7_CUDA(h, cudaMemset(h->d_prof, 0, (size_t)h->cfg.batch * 32 * sizeof(long long)));
    }
    if (clocks16 && h->d_prof && batch > 0 && batch <= h->cfg.batch) {
        I7_CUDA(h, cudaDeviceSynchronize());
        I7_CUDA(h, cudaMemcpy(clocks16, h->d_prof, (size_t)batch * 32 * sizeof(long long), cudaMemcpyDeviceToHost));
    }
    if (!enable && h->d_prof) { cudaFree(h->d_prof); h->d_prof = nullptr; }
    return I7_OK;
}

i7_status i7_p2p_create(i7_handle* h, int32_t world, int32_t rank, unsigned char* handle_out) {
    if (!h || !handle_out || world < 1 || world > 16 || rank < 0 || rank >= world) { if (h) h->err = "bad argument"; return I7_ERR_INVALID; }
    if (h->p2p_mine) { h->err = "exchange buffer already created"; return I7_ERR_INVALID; }
    static_assert(sizeof(cudaIpcMemHandle_t) == I7_IPC_HANDLE_BYTES, "IPC handle size");
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    const int stride = h->cfg.batch + 2 + (int)h->nz;
    const i7::P2PLayout L = i7::p2p_layout(world, stride, h->esize);
    I7_CUDA(h, cudaMalloc(&h->p2p_mine, L.total));
    I7_CUDA(h, cudaMemset(h->p2p_mine, 0, L.total));
    I7_CUDA(h, cudaMalloc(&h->p2p_err, sizeof(int)));
    I7_CUDA(h, cudaMemset(h->p2p_err, 0, sizeof(int)));
    I7_CUDA(h, cudaDeviceSynchronize());
    cudaIpcMemHandle_t ih;
    I7_CUDA(h, cudaIpcGetMemHandle(&ih, h->p2p_mine));
    std::memcpy(handle_out, &ih, sizeof(ih));
    h->p2p_world = world; h->p2p_rank = rank;
    return I7_OK;
}

i7_status i7_p2p_connect(i7_handle* h, const unsigned char* handles) {
    if (!h || !handles || !h->p2p_mine) { if (h) h->err = "i7_p2p_create first"; return I7_ERR_INVALID; }
    if (h->p2p_table) { h->err = "peers already connected"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    for (int r = 0; r < h->p2p_world; ++r) {
        if (r == h->p2p_rank) { h->p2p_peer[r] = h->p2p_mine; continue; }
        cudaIpcMemHandle_t ih;
        std::memcpy(&ih, handles + (size_t)r * I7_IPC_HANDLE_BYTES, sizeof(ih));
        I7_CUDA(h, cudaIpcOpenMemHandle(&h->p2p_peer[r], ih, cudaIpcMemLazyEnablePeerAccess));
    }
    I7_CUDA(h, cudaMalloc(&h->p2p_table, 16 * sizeof(unsigned char*)));
    I7_CUDA(h, cudaMemcpy(h->p2p_table, h->p2p_peer, 16 * sizeof(unsigned char*), cudaMemcpyHostToDevice));
    return I7_OK;
}

i7_status i7_p2p_exchange(i7_handle* h, int32_t batch, const void* stats, const void* xu, int32_t index_offset, void* out,
                          void* stream) {
    if (!h || !stats || !xu || !out || !h->p2p_table) { if (h) h->err = "i7_p2p_connect first"; return I7_ERR_INVALID; }
    if (batch != h->cfg.batch) { h->err = "the exchange is laid out for the handle's batch on every rank"; return I7_ERR_INVALID; }
    if (ptr_kind(stats) != 1 || ptr_kind(xu) != 1 || ptr_kind(out) != 1) { h->err = "i7_p2p_exchange takes device pointers"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned long long seq = ++h->p2p_seq;
    const unsigned long long timeout_ns = 4000000000ull;
    if (h->esize == 8)
        i7::pack_exchange_reduce_kernel<double><<<1, 256, 0, s>>>(batch, (int)h->nz, (const double*)stats, I7_STATS_LEN, 8, (const double*)xu,
                                                                 index_offset, h->p2p_rank, h->p2p_world, h->p2p_table, seq, timeout_ns, (double*)out, h->p2p_err);
    else
        i7::pack_exchange_reduce_kernel<float><<<1, 256, 0, s>>>(batch, (int)h->nz, (const float*)stats, I7_STATS_LEN, 8, (const float*)xu,
                                                                index_offset, h->p2p_rank, h->p2p_world, h->p2p_table, seq, timeout_ns, (float*)out, h->p2p_err);
    ++h->launches;
    I7_CUDA(h, cudaGetLastError());
    return I7_OK;
}

i7_status i7_solve_select(i7_handle* h, int32_t batch, const void* xu_in, const void* x0, const void* goals, int32_t goal_stride,
                          void* xu_out, void* stats, int32_t index_offset, void* best_out, void* stream) {
    if (!h || !xu_in || !x0 || !goals || !xu_out || !stats || !best_out) { if (h) h->err = "null argument"; return I7_ERR_INVALID; }
    if (!h->p2p_table) { h->err = "i7_p2p_connect first (a world of one rank is fine)"; return I7_ERR_INVALID; }
    if (batch != h->cfg.batch) { h->err = "the exchange is laid out for the handle's batch on every rank"; return I7_ERR_INVALID; }
    if (h->use_cluster || h->cfg.qp_mode != I7_QP_EXACT) { h->err = "i7_solve_select needs the one-CTA-per-problem form (cluster = 1) of the exact QP mode"; return I7_ERR_UNSUPPORTED; }
    const void* ptrs[6] = {xu_in, x0, goals, xu_out, stats, best_out};
    for (const void* p : ptrs) if (ptr_kind(p) != 1) { h->err = "i7_solve_select takes device pointers"; return I7_ERR_INVALID; }
    i7_status rc = check_batch(h, batch, goal_stride);
    if (rc) return rc;
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    if (!h->d_xchg_ticket) {
        I7_CUDA(h, cudaMalloc((void**)&h->d_xchg_ticket, sizeof(unsigned)));
        I7_CUDA(h, cudaMemset(h->d_xchg_ticket, 0, sizeof(unsigned)));
        I7_CUDA(h, cudaMalloc((void**)&h->d_xchg_wait, sizeof(long long)));
        I7_CUDA(h, cudaMemset(h->d_xchg_wait, 0, sizeof(long long)));
    }
    i7::XchgArgs X{};
    X.peers = h->p2p_table; X.seq = ++h->p2p_seq; X.timeout_ns = 4000000000ull; X.out = best_out; X.err = h->p2p_err;
    X.ticket = h->d_xchg_ticket; X.wait_ns = h->d_xchg_wait; X.rank = h->p2p_rank; X.world = h->p2p_world;
    X.index_offset = index_offset; X.enabled = 1;
    I7_CUDA(h, cudaEventRecord(h->ev0, s));
    if (h->cfg.dtype == I7_F64) {
        i7::SolveParams<double> P;
        base_params<double>(h, P, batch, goal_stride, i7::MODE_SOLVE);
        P.xu_in = (const double*)xu_in; P.x0 = (const double*)x0; P.goals = (const double*)goals; P.xu_out = (double*)xu_out;
        P.stats = (double*)stats; P.xchg = X;
        rc = launch_sqp<double>(h, h->rob64, P, s);
    } else {
        i7::SolveParams<float> P;
        base_params<float>(h, P, batch, goal_stride, i7::MODE_SOLVE);
        P.xu_in = (const float*)xu_in; P.x0 = (const float*)x0; P.goals = (const float*)goals; P.xu_out = (float*)xu_out;
        P.stats = (float*)stats; P.xchg = X;
        rc = launch_sqp<float>(h, h->rob32, P, s);
    }
    if (rc) return rc;
    I7_CUDA(h, cudaEventRecord(h->ev1, s));
    h->timed = true;
    return I7_OK;
}

i7_status i7_exchange_wait_ns(i7_handle* h, int64_t* total_ns, int32_t reset) {
    if (!h || !total_ns) return I7_ERR_INVALID;
    *total_ns = 0;
    if (!h->d_xchg_wait) return I7_OK;
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    I7_CUDA(h, cudaDeviceSynchronize());
    long long v = 0;
    I7_CUDA(h, cudaMemcpy(&v, h->d_xchg_wait, sizeof(v), cudaMemcpyDeviceToHost));
    *total_ns = v;
    if (reset) I7_CUDA(h, cudaMemset(h->d_xchg_wait, 0, sizeof(long long)));
    return I7_OK;
}

i7_status i7_p2p_gathered(i7_handle* h, const void** gathered) {
    if (!h || !gathered || !h->p2p_mine || h->p2p_seq == 0) { if (h) h->err = "no exchange yet"; return I7_ERR_INVALID; }
    const i7::P2PLayout L = i7::p2p_layout(h->p2p_world, h->cfg.batch + 2 + (int)h->nz, h->esize);
    *gathered = (const char*)h->p2p_mine + (h->p2p_seq & 1) * L.data_bytes;
    return I7_OK;
}

i7_status i7_p2p_status(i7_handle* h, void* stream) {
    if (!h || !h->p2p_err) { if (h) h->err = "no exchange buffer"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaStreamSynchronize((cudaStream_t)stream));
    int e = 0;
    I7_CUDA(h, cudaMemcpy(&e, h->p2p_err, sizeof(int), cudaMemcpyDeviceToHost));
    if (e) { h->err = "peer-memory exchange timed out waiting for a rank"; return I7_ERR_CUDA; }
    return I7_OK;
}

i7_status i7_p2p_reset(i7_handle* h) {
    if (!h || !h->p2p_mine) { if (h) h->err = "no exchange buffer"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    I7_CUDA(h, cudaDeviceSynchronize());
    const i7::P2PLayout L = i7::p2p_layout(h->p2p_world, h->cfg.batch + 2 + (int)h->nz, h->esize);
    I7_CUDA(h, cudaMemset((char*)h->p2p_mine + L.flags_off, 0, L.total - L.flags_off));
    I7_CUDA(h, cudaMemset(h->p2p_err, 0, sizeof(int)));
    I7_CUDA(h, cudaDeviceSynchronize());
    h->p2p_seq = 0;
    return I7_OK;
}

i7_status i7_measure_fma_peak(i7_handle* h, double* tflops) {
    if (!h || !tflops) return I7_ERR_INVALID;
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    int sms = 0;
    I7_CUDA(h, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->cfg.device));
    const int iters = 1 << 15, threads = 256, blocks = sms * 8;
    void* sink = nullptr;
    I7_CUDA(h, cudaMalloc(&sink, 64));
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    double best = 0;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(a, 0);
        if (h->esize == 8) i7::fma_peak_kernel<double><<<blocks, threads>>>(iters, 1.0 + rep, (double*)sink);
        else i7::fma_peak_kernel<float><<<blocks, threads>>>(iters, 1.0f + rep, (float*)sink);
        cudaEventRecord(b, 0);
        cudaEventSynchronize(b);
        float ms = 0;
        cudaEventElapsedTime(&ms, a, b);
        const double tf = 2.0 * 8.0 * iters * (double)threads * blocks / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    h->launches += 5;
    cudaEventDestroy(a); cudaEventDestroy(b);
    cudaFree(sink);
    I7_CUDA(h, cudaGetLastError());
    *tflops = best;
    return I7_OK;
}

i7_status i7_qp_state(i7_handle* h, int32_t batch, void* x, void* y, void* y_lim, void* info, void* stream) {
    if (!h) return I7_ERR_INVALID;
    if (h->cfg.qp_mode != I7_QP_OSQP) { h->err = "i7_qp_state: the exact QP solve keeps no state (qp_mode I7_QP_OSQP only)"; return I7_ERR_INVALID; }
    if (batch < 1 || batch > h->cfg.batch) { h->err = "batch out of range for this handle"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    const size_t N = h->cfg.N, nz = h->nz;
    // the warm start is kept padded to 18 numbers per knot; rows of the caller's arrays are 18N-6 / 18N-18 long
    if (x) I7_CUDA(h, cudaMemcpy2DAsync(x, nz * 8, h->d_wx, 18 * N * 8, nz * 8, batch, cudaMemcpyDefault, s));
    if (y) I7_CUDA(h, cudaMemcpyAsync(y, h->d_wy, (size_t)batch * 12 * N * 8, cudaMemcpyDefault, s));
    if (y_lim) I7_CUDA(h, cudaMemcpy2DAsync(y_lim, (nz - 12) * 8, (char*)h->d_wyl + 12 * 8, 18 * N * 8, (nz - 12) * 8, batch, cudaMemcpyDefault, s));
    if (info) I7_CUDA(h, cudaMemcpyAsync(info, h->d_qpinfo, (size_t)batch * 4 * 8, cudaMemcpyDefault, s));
    I7_CUDA(h, cudaStreamSynchronize(s));
    return I7_OK;
}

i7_status i7_solve_qp(i7_handle* h, int32_t batch, const void* xu, const void* x0, const void* goals,
                      int32_t goal_stride, void* z_out, void* y_out, void* stream) {
    if (!h || !xu || !x0 || !goals || !z_out) { if (h) h->err = "null argument"; return I7_ERR_INVALID; }
    SqpCall c{};
    c.mode = i7::MODE_QP; c.batch = batch; c.goal_stride = goal_stride;
    c.xu = xu; c.x0 = x0; c.goals = goals; c.z_out = z_out; c.y_out = y_out;
    return run_sqp(h, c, stream);
}

i7_status i7_dump_qp(i7_handle* h, int32_t batch, const void* xu, const void* x0, const void* goals,
                     int32_t goal_stride, void* Adata, void* Pdata, void* g, void* l, void* stream) {
    if (!h || !xu || !x0 || !goals) { if (h) h->err = "null argument"; return I7_ERR_INVALID; }
    SqpCall c{};
    c.mode = i7::MODE_DUMP; c.batch = batch; c.goal_stride = goal_stride;
    c.xu = xu; c.x0 = x0; c.goals = goals; c.Adata = Adata; c.Pdata = Pdata; c.g = g; c.l = l;
    return run_sqp(h, c, stream);
}

i7_status i7_merit(i7_handle* h, int32_t batch, const void* xu, const void* xu_ref, const void* goals,
                   int32_t goal_stride, void* out6, void* stream) {
    if (!h || !xu || !xu_ref || !goals || !out6) { if (h) h->err = "null argument"; return I7_ERR_INVALID; }
    SqpCall c{};
    c.mode = i7::MODE_MERIT; c.batch = batch; c.goal_stride = goal_stride;
    c.xu = xu; c.xu_full = xu_ref; c.goals = goals; c.merits = out6;
    return run_sqp(h, c, stream);
}

i7_status i7_linesearch(i7_handle* h, int32_t batch, const void* xu, const void* xu_full, const void* goals,
                        int32_t goal_stride, void* alpha, void* merits9, void* stream) {
    if (!h || !xu || !xu_full || !goals || !alpha || !merits9) { if (h) h->err = "null argument"; return I7_ERR_INVALID; }
    SqpCall c{};
    c.mode = i7::MODE_LINESEARCH; c.batch = batch; c.goal_stride = goal_stride;
    c.xu = xu; c.xu_full = xu_full; c.goals = goals; c.alpha = alpha; c.merits = merits9;
    return run_sqp(h, c, stream);
}

}  // extern "C"

namespace {
template <typename T>
i7_status dynamics_t(i7_handle* h, int n, const void* q, const void* v, const void* tau, const void* f_ext, void* qdd,
                     void* dq, void* dv, void* minv, cudaStream_t s) {
    const size_t es = sizeof(T);
    const size_t need_in = (size_t)n * 24 * es + 256, need_out = (size_t)n * (6 + 108) * es + 256;
    i7_status rc;
    if ((rc = arena_reserve(h, h->in, need_in))) return rc;
    if ((rc = arena_reserve(h, h->out, need_out))) return rc;
    Stage st(h);
    const T* dq_ = (const T*)st.in(q, n * 6 * es);
    const T* dv_ = (const T*)st.in(v, n * 6 * es);
    const T* dt_ = (const T*)st.in(tau, n * 6 * es);
    const T* df_ = (const T*)st.in(f_ext, n * 6 * es);
    T* o_qdd = (T*)st.out(qdd, n * 6 * es);
    T* o_dq = (T*)st.out(dq, n * 36 * es);
    T* o_dv = (T*)st.out(dv, n * 36 * es);
    T* o_mi = (T*)st.out(minv, n * 36 * es);
    if ((rc = st.flush_in(s))) return rc;
    const auto& rob = get_rob<T>(h);
    i7::dynamics_kernel<T><<<(n + 63) / 64, 64, 0, s>>>(rob, n, dq_, dv_, dt_, df_, o_qdd, o_dq, o_dv, o_mi);
    ++h->launches;
    I7_CUDA(h, cudaGetLastError());
    return st.flush_out(s);
}

template <typename T>
i7_status eepos_t(i7_handle* h, int n, const void* q, void* p, void* J, cudaStream_t s) {
    const size_t es = sizeof(T);
    i7_status rc;
    if ((rc = arena_reserve(h, h->in, (size_t)n * 6 * es + 256))) return rc;
    if ((rc = arena_reserve(h, h->out, (size_t)n * 21 * es + 256))) return rc;
    Stage st(h);
    const T* dq_ = (const T*)st.in(q, n * 6 * es);
    T* o_p = (T*)st.out(p, n * 3 * es);
    T* o_J = (T*)st.out(J, n * 18 * es);
    if ((rc = st.flush_in(s))) return rc;
    const auto& rob = get_rob<T>(h);
    i7::eepos_kernel<T><<<(n + 127) / 128, 128, 0, s>>>(rob, n, dq_, o_p, o_J);
    ++h->launches;
    I7_CUDA(h, cudaGetLastError());
    return st.flush_out(s);
}

template <typename T>
i7_status sim_forward_t(i7_handle* h, int batch, const void* x, const void* u, double dt, int bcast, void* x_next,
                        cudaStream_t s) {
    const size_t es = sizeof(T);
    Stage st(h);
    const size_t rows = bcast ? 1 : batch;
    const T* dx = (const T*)st.in(x, rows * 12 * es);
    const T* du = (const T*)st.in(u, rows * 6 * es);
    T* o = (T*)st.out(x_next, (size_t)batch * 12 * es);
    i7_status rc;
    if ((rc = st.flush_in(s))) return rc;
    const auto& rob = get_rob<T>(h);
    // a warp per lane while that still fills the device only a few times over, a thread per lane beyond
    if (batch <= 512)
        i7::sim_forward_warp_kernel<T><<<batch, 32, 0, s>>>(rob, batch, dx, du, bcast,
                                                           h->has_fext ? (const T*)h->d_fext : nullptr, T(dt), o,
                                                           i7::SelectArgs<T>{nullptr, nullptr, nullptr, nullptr, 0});
    else
        i7::sim_forward_kernel<T><<<(batch + 63) / 64, 64, 0, s>>>(rob, batch, dx, du, bcast,
                                                                   h->has_fext ? (const T*)h->d_fext : nullptr, T(dt), o);
    ++h->launches;
    I7_CUDA(h, cudaGetLastError());
    return st.flush_out(s);
}

template <typename T>
i7_status find_best_idx_t(i7_handle* h, int batch, const void* x_last, const void* u_last, const void* xs, double dt,
                          int tie_last, void* x_next, void* errors, void* result, cudaStream_t s) {
    const size_t es = sizeof(T);
    if (!h->d_sel_counter) {
        I7_CUDA(h, cudaMalloc((void**)&h->d_sel_counter, sizeof(unsigned)));
        I7_CUDA(h, cudaMemset(h->d_sel_counter, 0, sizeof(unsigned)));
        I7_CUDA(h, cudaMalloc(&h->d_sel_errs, (size_t)h->cfg.batch * es));
    }
    Stage st(h);
    st.zero_copy = h->zero_copy != 0;       // page-locked buffers (30 numbers in, 2 out) are read and written in place
    const T* dx = (const T*)st.in(x_last, 12 * es);
    const T* du = (const T*)st.in(u_last, 6 * es);
    const T* dxs = (const T*)st.in(xs, 12 * es);
    T* o = x_next ? (T*)st.out(x_next, (size_t)batch * 12 * es) : nullptr;
    T* de = errors ? (T*)st.out(errors, (size_t)batch * es) : (T*)h->d_sel_errs;
    T* dr = (T*)st.out(result, 2 * es);
    i7_status rc;
    if ((rc = st.flush_in(s))) return rc;
    i7::sim_forward_warp_kernel<T><<<batch, 32, 0, s>>>(get_rob<T>(h), batch, dx, du, 1,
                                                       h->has_fext ? (const T*)h->d_fext : nullptr, T(dt), o,
                                                       i7::SelectArgs<T>{dxs, de, dr, h->d_sel_counter, tie_last});
    ++h->launches;
    I7_CUDA(h, cudaGetLastError());
    return st.flush_out(s);
}
template <typename T>
i7_status shift_t(i7_handle* h, int batch, int steps, int tail, const void* xu_new, const void* xu_prev, const void* x_cur,
                  void* xu_out, void* x0_out, cudaStream_t s) {
    const size_t es = sizeof(T), nz = h->nz;
    Stage st(h);
    st.zero_copy = h->zero_copy != 0;       // page-locked buffers (the measured states of a tick) are read in place
    const T* d_new = (const T*)st.in(xu_new, (size_t)batch * nz * es);
    const T* d_prev = (const T*)st.in(xu_prev, (size_t)batch * nz * es);
    const T* d_cur = (const T*)st.in(x_cur, (size_t)batch * 12 * es);
    T* d_out = (T*)st.out(xu_out, (size_t)batch * nz * es);
    T* d_x0 = (T*)st.out(x0_out, (size_t)batch * 12 * es);
    i7_status rc;
    if ((rc = st.flush_in(s))) return rc;
    const size_t total = (size_t)batch * nz;
    const int blocks = (int)((total + 255) / 256 < 1184 ? (total + 255) / 256 : 1184);
    i7::shift_kernel<T><<<blocks, 256, 0, s>>>(batch, h->cfg.N, steps, tail, d_new, d_prev, d_cur, d_out, d_x0);
    ++h->launches;
    I7_CUDA(h, cudaGetLastError());
    return st.flush_out(s);
}

template <typename T>
i7_status goal_window_t(i7_handle* h, int batch, const void* traj, int traj_len, int traj_stride, const int32_t* offsets,
                        int tick, int wrap, int out_stride, void* goals, cudaStream_t s) {
    const size_t es = sizeof(T);
    i7_status rc;
    const size_t need = (size_t)traj_len * traj_stride * es + (size_t)batch * sizeof(int32_t) + 256;
    if (ptr_kind(traj) != 1 && (rc = arena_reserve(h, h->in, need))) return rc;
    Stage st(h);
    const T* d_traj = (const T*)st.in(traj, (size_t)traj_len * traj_stride * es);
    const int* d_off = (const int*)st.in(offsets, (size_t)batch * sizeof(int32_t));
    T* d_goals = (T*)st.out(goals, (size_t)batch * h->cfg.N * out_stride * es);
    if ((rc = st.flush_in(s))) return rc;
    const size_t total = (size_t)batch * h->cfg.N * out_stride;
    const int blocks = (int)((total + 255) / 256 < 1184 ? (total + 255) / 256 : 1184);
    i7::goal_window_kernel<T><<<blocks, 256, 0, s>>>(batch, h->cfg.N, d_traj, traj_len, traj_stride, d_off, tick, wrap, out_stride, d_goals);
    ++h->launches;
    I7_CUDA(h, cudaGetLastError());
    return st.flush_out(s);
}
}  // namespace

extern "C" {

i7_status i7_shift(i7_handle* h, int32_t batch, int32_t steps, int32_t tail, const void* xu_new, const void* xu_prev,
                   const void* x_cur, void* xu_out, void* x0_out, void* stream) {
    if (!h || !xu_new || !xu_out) { if (h) h->err = "null argument"; return I7_ERR_INVALID; }
    if (batch < 1 || batch > h->cfg.batch) { h->err = "batch out of range for this handle"; return I7_ERR_INVALID; }
    if (steps < 0 || steps >= h->cfg.N || tail < 0 || tail > 2) { h->err = "steps must be in [0, N), tail in {0, 1, 2}"; return I7_ERR_INVALID; }
    if (tail == 0 && !xu_prev) { h->err = "tail 0 (src/osqp_mpc.py:65-70) keeps the end of the previous warm start: xu_prev required"; return I7_ERR_INVALID; }
    if (xu_out == xu_new && steps > 0) { h->err = "i7_shift cannot run in place"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    return h->cfg.dtype == I7_F64 ? shift_t<double>(h, batch, steps, tail, xu_new, xu_prev, x_cur, xu_out, x0_out, s)
                                  : shift_t<float>(h, batch, steps, tail, xu_new, xu_prev, x_cur, xu_out, x0_out, s);
}

i7_status i7_goal_window(i7_handle* h, int32_t batch, const void* traj, int32_t traj_len, int32_t traj_stride,
                         const int32_t* offsets, int32_t tick, int32_t wrap, int32_t out_stride, void* goals, void* stream) {
    if (!h || !traj || !goals) { if (h) h->err = "null argument"; return I7_ERR_INVALID; }
    if (batch < 1 || batch > h->cfg.batch) { h->err = "batch out of range for this handle"; return I7_ERR_INVALID; }
    if (traj_len < 1 || traj_stride < 3 || (out_stride != 3 && out_stride != 6)) { h->err = "traj_len >= 1, traj_stride >= 3, out_stride 3 or 6"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    return h->cfg.dtype == I7_F64 ? goal_window_t<double>(h, batch, traj, traj_len, traj_stride, offsets, tick, wrap, out_stride, goals, s)
                                  : goal_window_t<float>(h, batch, traj, traj_len, traj_stride, offsets, tick, wrap, out_stride, goals, s);
}

i7_status i7_find_best_idx(i7_handle* h, int32_t batch, const void* x_last, const void* u_last, const void* xs, double dt,
                           int32_t tie_last, void* x_next, void* errors, void* result, void* stream) {
    if (!h || !x_last || !u_last || !xs || !result) { if (h) h->err = "null argument"; return I7_ERR_INVALID; }
    if (batch < 1 || batch > h->cfg.batch) { h->err = "batch out of range for this handle"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    return h->cfg.dtype == I7_F64
               ? find_best_idx_t<double>(h, batch, x_last, u_last, xs, dt, tie_last, x_next, errors, result, s)
               : find_best_idx_t<float>(h, batch, x_last, u_last, xs, dt, tie_last, x_next, errors, result, s);
}

i7_status i7_dynamics(i7_handle* h, int32_t n, const void* q, const void* v, const void* tau, const void* f_ext,
                      void* qdd, void* dqdd_dq, void* dqdd_dv, void* Minv, void* stream) {
    if (!h || n < 1 || !q || !v || !tau) { if (h) h->err = "bad argument"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    return h->cfg.dtype == I7_F64 ? dynamics_t<double>(h, n, q, v, tau, f_ext, qdd, dqdd_dq, dqdd_dv, Minv, s)
                                  : dynamics_t<float>(h, n, q, v, tau, f_ext, qdd, dqdd_dq, dqdd_dv, Minv, s);
}

i7_status i7_eepos(i7_handle* h, int32_t n, const void* q, void* p, void* J, void* stream) {
    if (!h || n < 1 || !q || !p) { if (h) h->err = "bad argument"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    return h->cfg.dtype == I7_F64 ? eepos_t<double>(h, n, q, p, J, s) : eepos_t<float>(h, n, q, p, J, s);
}

i7_status i7_sim_forward(i7_handle* h, int32_t batch, const void* x, const void* u, double dt, int32_t bcast,
                         void* x_next, void* stream) {
    if (!h || !x || !u || !x_next) { if (h) h->err = "null argument"; return I7_ERR_INVALID; }
    if (batch < 1 || batch > h->cfg.batch) { h->err = "batch out of range for this handle"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    return h->cfg.dtype == I7_F64 ? sim_forward_t<double>(h, batch, x, u, dt, bcast, x_next, s)
                                  : sim_forward_t<float>(h, batch, x, u, dt, bcast, x_next, s);
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// i7_find_best_idx_static: the hypothesis selection of i7_find_best_idx with its scratch (ticket counter + one error per
// lane) in a file-scope device array instead of the lazily cudaMalloc'ed buffers of the handle — for callers that must
// not allocate (CUDA graph capture, tight serving loops).  No counterpart in the reference.
// Every GPU that loads the library holds the 256 + 8 * I7_STATIC_MAX_BATCH bytes until exit.
// In host code the name of the array is a host-side stand-in, NOT a device address: the address handed to the kernel
// comes from cudaGetSymbolAddress, taken once per device with that device current.
// ---------------------------------------------------------------------------------------------------------------------
namespace {
constexpr int kStaticScratchHeader = 256;       // the ticket counter lives in the first 4 bytes
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchHeader + 8 * I7_STATIC_MAX_BATCH];

constexpr int kMaxDevices = 64;
std::atomic<void*> g_static_scratch_addr[kMaxDevices];

// device address of g_static_scratch on `device` (which must be current)
i7_status static_scratch_ptr(i7_handle* h, int device, unsigned char** out) {
    if (device < 0 || device >= kMaxDevices) { h->err = "device ordinal beyond the static scratch table"; return I7_ERR_INVALID; }
    void* p = g_static_scratch_addr[device].load(std::memory_order_acquire);
    if (!p) {
        I7_CUDA(h, cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) { h->err = "cudaGetSymbolAddress returned a null address for the static scratch"; return I7_ERR_CUDA; }
        g_static_scratch_addr[device].store(p, std::memory_order_release);     // same value from every thread
    }
    *out = (unsigned char*)p;
    return I7_OK;
}

template <typename T>
i7_status find_best_idx_static_t(i7_handle* h, int batch, const void* x_last, const void* u_last, const void* xs, double dt,
                                 int tie_last, void* x_next, void* errors, void* result, cudaStream_t s) {
    const size_t es = sizeof(T);
    unsigned char* scratch = nullptr;
    i7_status rc;
    if ((rc = static_scratch_ptr(h, h->cfg.device, &scratch))) return rc;
    unsigned* counter = (unsigned*)scratch;
    T* errs_scratch = (T*)(scratch + kStaticScratchHeader);
    // the array keeps the previous call's contents: the ticket counter must start at zero (a launch that did not run
    // to its end would leave it elsewhere); the errors are all rewritten before they are read
    I7_CUDA(h, cudaMemsetAsync(counter, 0, kStaticScratchHeader, s));
    Stage st(h);                                // staging arenas of the handle, sized at i7_create: nothing is allocated
    const T* dx = (const T*)st.in(x_last, 12 * es);
    const T* du = (const T*)st.in(u_last, 6 * es);
    const T* dxs = (const T*)st.in(xs, 12 * es);
    T* o = x_next ? (T*)st.out(x_next, (size_t)batch * 12 * es) : nullptr;
    T* dr = (T*)st.out(result, 2 * es);
    if ((rc = st.flush_in(s))) return rc;
    i7::sim_forward_warp_kernel<T><<<batch, 32, 0, s>>>(get_rob<T>(h), batch, dx, du, 1,
                                                       h->has_fext ? (const T*)h->d_fext : nullptr, T(dt), o,
                                                       i7::SelectArgs<T>{dxs, errs_scratch, dr, counter, tie_last});
    ++h->launches;
    I7_CUDA(h, cudaGetLastError());
    if (errors)                                 // optional copy of the per-lane errors out of the scratch (host or device)
        I7_CUDA(h, cudaMemcpyAsync(errors, errs_scratch, (size_t)batch * es, cudaMemcpyDefault, s));
    if ((rc = st.flush_out(s))) return rc;
    if (errors && ptr_kind(errors) != 1) I7_CUDA(h, cudaStreamSynchronize(s));
    return I7_OK;
}
}  // namespace

extern "C" {

i7_status i7_find_best_idx_static(i7_handle* h, int32_t batch, const void* x_last, const void* u_last, const void* xs,
                                  double dt, int32_t tie_last, void* x_next, void* errors, void* result, void* stream) {
    if (!h || !x_last || !u_last || !xs || !result) { if (h) h->err = "null argument"; return I7_ERR_INVALID; }
    if (batch < 1 || batch > h->cfg.batch) { h->err = "batch out of range for this handle"; return I7_ERR_INVALID; }
    if (batch > I7_STATIC_MAX_BATCH) {
        h->err = "i7_find_best_idx_static: batch " + std::to_string(batch) + " exceeds the static scratch (I7_STATIC_MAX_BATCH = " +
                 std::to_string(I7_STATIC_MAX_BATCH) + " lanes); use i7_find_best_idx";
        return I7_ERR_INVALID;
    }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    return h->cfg.dtype == I7_F64
               ? find_best_idx_static_t<double>(h, batch, x_last, u_last, xs, dt, tie_last, x_next, errors, result, s)
               : find_best_idx_static_t<float>(h, batch, x_last, u_last, xs, dt, tie_last, x_next, errors, result, s);
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// i7_resample_wrench: gato_controller.py:120-129 resample_f_ext_batch (and the notebook / batch-sample variants) on the
// device, chained to i7_find_best_idx without a host round trip (best index read from its device result).
// ---------------------------------------------------------------------------------------------------------------------
namespace {
template <typename T>
i7_status resample_wrench_t(i7_handle* h, int batch, int best_idx, const void* best_result, const double* f_ext,
                            const double* noise, int keep_lane, int flags, double decay, double* f_ext_out, cudaStream_t s) {
    Stage st(h);
    st.zero_copy = h->zero_copy != 0;       // page-locked draws are read in place
    const T* d_res = best_idx < 0 ? (const T*)st.in(best_result, 2 * sizeof(T)) : nullptr;
    const double* d_f = (const double*)st.in(f_ext, (size_t)batch * 6 * sizeof(double));
    const double* d_n = (const double*)st.in(noise, (size_t)batch * 6 * sizeof(double));
    // in place on device memory: the kernel reads the selected lane before it writes any (one CTA)
    double* d_o = (f_ext_out == f_ext && ptr_kind(f_ext) == 1) ? const_cast<double*>(d_f)
                                                               : (double*)st.out(f_ext_out, (size_t)batch * 6 * sizeof(double));
    i7_status rc;
    if ((rc = st.flush_in(s))) return rc;
    const bool apply = (flags & I7_RESAMPLE_APPLY) != 0;
    i7::resample_wrench_kernel<T><<<1, 256, 0, s>>>(batch, best_idx, d_res, d_f, d_n, keep_lane,
                                                    (flags & I7_RESAMPLE_KEEP_TORQUE) ? 1 : 0, decay, d_o,
                                                    apply ? (T*)h->d_fext : nullptr);
    ++h->launches;
    I7_CUDA(h, cudaGetLastError());
    if (apply) h->has_fext = true;
    return st.flush_out(s);
}
}  // namespace

extern "C" {

i7_status i7_resample_wrench(i7_handle* h, int32_t batch, int32_t best_idx, const void* best_result, const double* f_ext,
                             const double* noise, int32_t keep_lane, int32_t flags, double decay, double* f_ext_out,
                             void* stream) {
    if (!h || !f_ext || !noise || !f_ext_out) { if (h) h->err = "null argument"; return I7_ERR_INVALID; }
    if (batch < 1 || batch > h->cfg.batch) { h->err = "batch out of range for this handle"; return I7_ERR_INVALID; }
    if (best_idx >= batch || (best_idx < 0 && !best_result)) {
        h->err = "best_idx must be a lane of the batch, or negative with best_result = the result of i7_find_best_idx";
        return I7_ERR_INVALID;
    }
    if (keep_lane < I7_KEEP_NONE) { h->err = "keep_lane: a lane index, I7_KEEP_BEST or I7_KEEP_NONE"; return I7_ERR_INVALID; }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    return h->cfg.dtype == I7_F64
               ? resample_wrench_t<double>(h, batch, best_idx, best_result, f_ext, noise, keep_lane, flags, decay, f_ext_out, s)
               : resample_wrench_t<float>(h, batch, best_idx, best_result, f_ext, noise, keep_lane, flags, decay, f_ext_out, s);
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// i7_broadcast_best: the hand-over at the end of a controller tick (gato_controller.py:226,249).
// ---------------------------------------------------------------------------------------------------------------------
namespace {
template <typename T>
i7_status broadcast_best_t(i7_handle* h, int batch, const void* xu_new, int best_idx, const void* best_result, void* xu_out,
                           void* best_out, cudaStream_t s) {
    const size_t es = sizeof(T), nz = h->nz;
    Stage st(h);
    const T* d_new = (const T*)st.in(xu_new, (size_t)batch * nz * es);
    const T* d_res = best_idx < 0 ? (const T*)st.in(best_result, 2 * es) : nullptr;
    T* d_out = !xu_out ? nullptr
                       : ((xu_out == xu_new && ptr_kind(xu_new) == 1) ? const_cast<T*>(d_new) : (T*)st.out(xu_out, (size_t)batch * nz * es));
    T* d_best = (T*)st.out(best_out, nz * es);
    i7_status rc;
    if ((rc = st.flush_in(s))) return rc;
    const size_t total = (size_t)batch * nz;
    const int blocks = (int)((total + 255) / 256 < 1184 ? (total + 255) / 256 : 1184);
    i7::broadcast_best_kernel<T><<<blocks, 256, 0, s>>>(batch, (int)nz, d_new, best_idx, d_res, d_out, d_best);
    ++h->launches;
    I7_CUDA(h, cudaGetLastError());
    return st.flush_out(s);
}
}  // namespace

extern "C" {

i7_status i7_broadcast_best(i7_handle* h, int32_t batch, const void* xu_new, int32_t best_idx, const void* best_result,
                            void* xu_out, void* best_out, void* stream) {
    if (!h || !xu_new || (!xu_out && !best_out)) { if (h) h->err = "null argument"; return I7_ERR_INVALID; }
    if (batch < 1 || batch > h->cfg.batch) { h->err = "batch out of range for this handle"; return I7_ERR_INVALID; }
    if (best_idx >= batch || (best_idx < 0 && !best_result)) {
        h->err = "best_idx must be a lane of the batch, or negative with best_result = the result of i7_find_best_idx";
        return I7_ERR_INVALID;
    }
    I7_CUDA(h, cudaSetDevice(h->cfg.device));
    cudaStream_t s = (cudaStream_t)stream;
    return h->cfg.dtype == I7_F64 ? broadcast_best_t<double>(h, batch, xu_new, best_idx, best_result, xu_out, best_out, s)
                                  : broadcast_best_t<float>(h, batch, xu_new, best_idx, best_result, xu_out, best_out, s);
}

}  // extern "C"
