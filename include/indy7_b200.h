/* indy7_b200.h — C ABI of the B200-native SQP-MPC solve loop (drop-in for the hot path of A2R-Lab/indy7-mpc).
 *
 * Boundary being replaced (reference file:line, relative to the reference checkout):
 *   src/osqp_solver.py:7-46    OSQPSolver.__init__            -> i7_create / i7_set_weights
 *   src/osqp_solver.py:70-101  compute_dynamics_jacobians /
 *                              update_constraint_matrix        -> i7_dump_qp (Adata, l), i7_dynamics
 *   src/osqp_solver.py:103-135 update_cost_matrix              -> i7_dump_qp (Pdata, g)
 *   src/osqp_solver.py:137-143 setup_and_solve_qp (osqp.solve) -> i7_solve_qp
 *   src/osqp_solver.py:146-155 eepos / d_eepos                 -> i7_eepos
 *   src/osqp_sqp.py:13-47      eepos_cost / integrator_err     -> i7_merit
 *   src/osqp_sqp.py:49-74      linesearch                      -> i7_linesearch
 *   src/osqp_sqp.py:76-93      sqp                             -> i7_solve
 *   src/utils.py:3-18          rk4 (plant step)                -> i7_sim_forward
 *   bindings.batch_sqp.SQPSolverfloat_<B> (binary absent; call sites gato_controller.py:53-138,
 *   src/gato_mpc_batch.py:14-55): solve / reset / resetRho / resetLambda / set_external_wrench_batch /
 *   sim_forward                                                -> i7_solve / i7_reset* / i7_set_external_wrench /
 *                                                                 i7_sim_forward
 *   native handle precedent: indy7-sdk/src/trajopt/include/trajopt_solver.cuh:8-226 (TrajoptSolver owns the
 *   device workspaces; caller owns inputs/outputs)            -> i7_handle
 *
 * Conventions
 *   - plain C, no torch / C++ types; every function returns an i7_status (0 = OK) and never throws.
 *   - a handle is bound to one CUDA device, one arithmetic type (dtype) and a maximum batch; it is not
 *     re-entrant.  Distinct handles may be used from distinct threads / streams.
 *   - every data pointer may be HOST or DEVICE memory (queried with cudaPointerGetAttributes).  Element
 *     type = the handle's dtype (double for I7_F64, float for I7_F32), C-contiguous, row = one problem.
 *     With device pointers a call is asynchronous on `stream`.  With host pointers the call returns after the
 *     results are in the caller's memory: page-locked buffers (cudaHostAlloc / cudaHostRegister) are read and
 *     written by the kernels themselves (zero-copy), pageable ones pass through the handle's page-locked mirror
 *     (details at i7_solve).
 *   - trajectory layout (src/osqp_solver.py:22,88-91): XU = [x_0 u_0 x_1 u_1 ... x_{N-1}], x = [q(6) v(6)],
 *     length 18N-6.  Goals: N knots of `goal_stride` numbers (3 = [x y z]; 6 = GATO format [x y z 0 0 0],
 *     src/gato_mpc_batch.py:80-88), only the first three of each knot are read.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).
 *   - CUDA graphs: with device pointers only, i7_solve (exact QP mode), i7_pack_best, i7_shift, i7_goal_window,
 *     i7_sim_forward and i7_broadcast_best issue nothing but kernel launches and device-to-device copies on `stream` and
 *     may be recorded under stream capture after one ordinary call of the same kind on the handle (first-use set-up);
 *     the replayed graph is bit-identical to the eager calls (tests/test_gpu_parity_r2.py::test_device_tick_under_cuda_graph).
 */
#ifndef INDY7_B200_H
#define INDY7_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define I7_NJ 6            /* joints (nq = nv = nu) */
#define I7_NX 12
#define I7_NXU 18
#define I7_MAX_QP 2        /* src/osqp_sqp.py:77  `for qp in range(2)` */
#define I7_NALPHA 8        /* src/osqp_sqp.py:58  alphas 1, 1/2, ..., 1/128 */
#define I7_STATS_LEN 12

typedef enum {
    I7_OK = 0,
    I7_ERR_INVALID = 1,     /* bad argument / shape */
    I7_ERR_CUDA = 2,        /* CUDA runtime error; text in i7_last_error */
    I7_ERR_NOMEM = 3,
    I7_ERR_UNSUPPORTED = 4
} i7_status;

typedef enum { I7_F64 = 0, I7_F32 = 1 } i7_dtype;

/* Robot constants of a serial chain of I7_NJ revolute-z joints (what pin.buildModelsFromUrdf yields for
 * description/indy7.urdf:64-245).  Always doubles, converted to the handle dtype at create time. */
typedef struct {
    double R[I7_NJ][9];     /* joint placement rotation, row-major, parent <- joint frame (Rz(yaw)Ry(pitch)Rx(roll)) */
    double p[I7_NJ][3];     /* joint placement translation in the parent frame */
    double mass[I7_NJ];
    double com[I7_NJ][3];   /* centre of mass in the link frame */
    double Ic[I7_NJ][9];    /* rotational inertia about the COM, link frame, row-major */
    double gravity[3];      /* world gravity, default (0, 0, -9.81) */
    /* joint limits (description/indy7.urdf:203-238 <limit lower upper effort velocity>); read only when
     * i7_config.limits != 0.  The reference QP has no inequality rows (src/osqp_solver.py:41 passes l == u). */
    double q_min[I7_NJ], q_max[I7_NJ], v_max[I7_NJ], u_max[I7_NJ];
} i7_model;

/* How the QP of an SQP pass is solved (src/osqp_solver.py:137-143):
 *   I7_QP_EXACT  block recursion over the knots: the exact KKT point OSQP converges to; no warm-start state; one
 *                launch per i7_solve (default — the throughput / latency path)
 *   I7_QP_OSQP   the reference's own regime: OSQP's ADMM (rho = 0.1, x 1e3 on equality rows; sigma = 1e-6; alpha = 1.6;
 *                eps_abs = eps_rel = 1e-3 checked every 25 iterations; no Ruiz scaling, no adaptive rho), warm-started
 *                per problem from the previous call on the same handle slot — (x, y) live in the handle like the OSQP
 *                object lives in the reference solver (src/osqp_solver.py:39-41).  f64 only; three launches per SQP pass
 *                (linearise, ADMM, line search + update).  Required for `limits`. */
typedef enum { I7_QP_EXACT = 0, I7_QP_OSQP = 1 } i7_qp_mode;

typedef struct {
    int32_t N;              /* knot points (32 in the drivers, 64 in gato_controller.py:314, 128 in trajopt_node.cu:178) */
    int32_t batch;          /* maximum number of independent problems per call */
    int32_t dtype;          /* i7_dtype */
    int32_t device;         /* CUDA device ordinal */
    double dt;              /* src/osqp_solver.py:7 defaults: 0.01 */
    double dQ_cost;         /* 0.01  */
    double R_cost;          /* 1e-5  */
    double QN_cost;         /* 100   */
    double eps;             /* 1     */
    int32_t regularize;     /* 1     */
    int32_t max_qp;         /* outer SQP iterations, <= I7_MAX_QP (reference: 2) */
    double mu;              /* merit penalty, src/osqp_sqp.py:50: 10 */
    double step_tol;        /* src/osqp_sqp.py:90: 1e-3 */
    int32_t threads;        /* CTA size of the one-CTA-per-problem kernel: 0 = choose from batch (384 up to 148 problems, else 256 with two CTAs per SM); 128 | 256 | 384 */
    int32_t cluster;        /* i7_solve kernel form: 0 = auto (the largest cluster of 8, 4 or 2 CTAs per problem for which
                               the whole batch is co-resident, else one CTA per problem), 1 = always one CTA per
                               problem, 2 | 4 | 8 = always that cluster form */
    int32_t qp_mode;        /* i7_qp_mode; default I7_QP_EXACT */
    int32_t limits;         /* != 0: box rows q_min <= q_k <= q_max, |v_k| <= v_max (k >= 1), |u_k| <= u_max in every QP
                               (needs I7_QP_OSQP) */
    double admm_rho;        /* 0.1   rho of inequality rows on states; equality rows carry 1e3 x admm_rho */
    double admm_rho_u;      /* 1e-3  rho of the effort-limit rows (unscaled ADMM: 0.1 stalls against R_cost = 1e-5) */
    double admm_sigma;      /* 1e-6 */
    double admm_alpha;      /* 1.6  */
    double admm_eps_abs;    /* 1e-3 */
    double admm_eps_rel;    /* 1e-3 */
    int32_t admm_max_iter;  /* 4000 */
    int32_t admm_check;     /* 25   */
    i7_model model;
} i7_config;

/* Per-problem statistics record written by i7_solve: I7_STATS_LEN numbers of the handle dtype.
 *   [0] qp_iters        what src/osqp_sqp.py:92 appends (qp+1)
 *   [1] n_linesearch    number of line searches run (= alphas appended, src/osqp_sqp.py:73)
 *   [2] n_steps         number of accepted steps (= stepsizes appended, src/osqp_sqp.py:88)
 *   [3] status          0 ok, 1 = a line search failed (alpha 0, not an error: src/osqp_sqp.py:72,81-82),
 *                       2 = non-finite value met
 *   [4],[5] alpha of line search 0 / 1 (0 when failed or not run)
 *   [6],[7] 2-norm of accepted step 0 / 1 (0 when absent)
 *   [8] merit of the returned trajectory as evaluated by the last line search (the accepted candidate, or the
 *       base merit when that line search failed) — the per-problem cost used for best-trajectory selection
 *   [9] base merit of the first line search (merit of the input trajectory)
 *   [10],[11] I7_QP_OSQP: ADMM iterations of QP 0 / QP 1 (what OSQP reports as info.iter); I7_QP_EXACT: 0      */

typedef struct i7_handle i7_handle;

const char* i7_version(void);
void i7_default_config(i7_config* cfg);         /* reference defaults, N=32, batch=1, f64, Indy7 model zeroed */
i7_status i7_create(const i7_config* cfg, i7_handle** out);
void i7_destroy(i7_handle* h);
const char* i7_last_error(const i7_handle* h);  /* h may be NULL: last create error */

i7_status i7_set_weights(i7_handle* h, double dQ_cost, double R_cost, double QN_cost, int32_t regularize, double eps);
/* Run-time switches of a handle.
 *   I7_OPT_X0_INTO_WARM_START (default 0): i7_solve / i7_solve_select overwrite the first state of every warm start with x0
 *   while loading it — the `XU[:12] = xcur` of the reference loop (src/osqp_mpc.py:69) done inside the solve, so the
 *   shift of a tick (i7_shift with x_cur == NULL) can run at the END of the previous tick, off the critical path. */
#define I7_OPT_X0_INTO_WARM_START 1
i7_status i7_set_option(i7_handle* h, int32_t option, int32_t value);
/* f_ext: batch x 6 DOUBLES on the host, Pinocchio order [fx fy fz tx ty tz], expressed in the local frame of
 * the last joint (src/gato_mpc_batch_sample.py:97-101,158); NULL clears. */
i7_status i7_set_external_wrench(i7_handle* h, const double* f_ext, int32_t batch);
/* bindings.batch_sqp compatibility (reset / resetRho / resetLambda, gato_controller.py:70-78).  I7_QP_OSQP: i7_reset
 * clears the whole warm start (x and y of every problem slot), i7_reset_lambda the duals only, i7_reset_rho restores
 * the configured rho (rho is never adapted, so it has nothing to undo).  I7_QP_EXACT carries no solver state: the
 * three calls only clear the timing state. */
i7_status i7_reset(i7_handle* h);
i7_status i7_reset_rho(i7_handle* h);
i7_status i7_reset_lambda(i7_handle* h);

/* One SQP solve per problem (src/osqp_sqp.py:76-93): xu_out[b] = sqp(x0[b], goals[b], xu_in[b]).
 * xu_in/xu_out: batch x (18N-6); x0: batch x 12; goals: batch x (N*goal_stride); stats: batch x I7_STATS_LEN
 * (may be NULL).  xu_out may alias xu_in.
 * Host buffers: the call returns when the results are in the caller's memory.  Page-locked host buffers
 * (cudaHostAlloc / cudaHostRegister) are read and written by the kernel itself (zero-copy; I7_ZERO_COPY=0 at
 * i7_create selects copy operations instead); pageable ones pass through the handle's page-locked mirror, chunk by
 * chunk for 512 problems and more.  Host inputs with DEVICE outputs: the call is asynchronous on `stream`, and
 * page-locked inputs must stay untouched until the stream is synchronised. */
i7_status i7_solve(i7_handle* h, int32_t batch, const void* xu_in, const void* x0, const void* goals,
                   int32_t goal_stride, void* xu_out, void* stats, void* stream);

/* Device-time of the latest i7_solve kernel in microseconds (CUDA events on the launch stream; blocks).  With the
 * environment variable I7_TIMING_EVENTS=0 at i7_create, small host-buffer solves (the mapped-memory path) are not
 * timed (the two event records cost 3 us of host time per solve) and this call then returns I7_ERR_INVALID. */
i7_status i7_last_solve_time_us(i7_handle* h, double* us);
/* Number of kernel launches issued by this handle so far. */
int64_t i7_launch_count(const i7_handle* h);
/* The kernel form i7_solve uses for this handle: threads per CTA, CTAs per problem (1 = CTA form; 2, 4, 8 = cluster
 * forms) and dynamic shared memory per CTA in bytes.  (The reference picks its kernel at compile time through the
 * SQPSolverfloat_<B> template instances, gato_controller.py:53-68; here it is a run-time choice, reported for logs.) */
i7_status i7_kernel_form(const i7_handle* h, int32_t* threads, int32_t* ctas_per_problem, int32_t* smem_bytes);

/* One rk4 plant step of length dt per lane under that lane's external wrench (src/utils.py:3-18):
 * x, u: batch x 12 / batch x 6 (or a single row broadcast to every lane when bcast != 0); out: batch x 12. */
i7_status i7_sim_forward(i7_handle* h, int32_t batch, const void* x, const void* u, double dt, int32_t bcast,
                         void* x_next, void* stream);

/* MPC tick plumbing on the device (reference: src/osqp_mpc.py:65-70; device precedent just_shift<T>,
 * indy7-sdk/src/trajopt/include/trajopt_solver.cuh:153-172): the warm start of the next tick = the new solution moved
 * `steps` knots towards the present.  xu_out[b] = xu_new[b][18 steps:], the last `steps` knots filled per `tail`:
 *   0  the reference loop: they keep the values of the previous warm start xu_prev, and the last state becomes
 *      [1 x 6, 0 x 6] (src/osqp_mpc.py:66,70);  1  trajopt_solver.cuh: x_{N-1} repeated, controls zeroed;
 *   2  the last [u, x] pair repeated (batch drivers, bench.py).
 * x_cur (batch x 12, may be NULL) overwrites x_0 (src/osqp_mpc.py:69); NULL: x_0 = x_steps of the solution.
 * x0_out (batch x 12, may be NULL) receives the new x_0.  xu_prev may be NULL unless tail == 0.  Not in place. */
i7_status i7_shift(i7_handle* h, int32_t batch, int32_t steps, int32_t tail, const void* xu_new, const void* xu_prev,
                   const void* x_cur, void* xu_out, void* x0_out, void* stream);
/* Goal window of every lane from ONE reference trajectory (gato_controller.py:210-218 slices
 * ref_traj[offset : offset + 6N] of the padded figure-8; trajopt_solver.cuh:97-146 keeps the same window on the
 * device): goals[b][k] = traj[offsets[b] + tick + k] (first three numbers of each sample of `traj_stride`), the index
 * wrapped (wrap != 0) or clamped to the last sample.  offsets: batch int32 (host or device; NULL = 0).  out_stride 3, or
 * 6 for the GATO format [x y z 0 0 0]. */
i7_status i7_goal_window(i7_handle* h, int32_t batch, const void* traj, int32_t traj_len, int32_t traj_stride,
                         const int32_t* offsets, int32_t tick, int32_t wrap, int32_t out_stride, void* goals, void* stream);

/* Wrench-hypothesis selection of the batch drivers in ONE launch (gato_controller.py:109-118 find_best_idx;
 * src/gato_mpc_batch_sample.py:254-300 find_best_trajectory_id after the wrench conversion): one rk4 plant step of
 * length dt from the single row (x_last, u_last) per lane under that lane's external wrench, error_i =
 * ||x_next_i - xs||_2, then the argmin in index order: tie_last != 0 compares with `<=` (last of equal minima,
 * gato_controller.py:115), tie_last == 0 with `<` (first, gato_mpc_batch_sample.py:283); NaN errors never win.
 * result: 2 numbers [best index (as a number; -1 when no lane qualified), its error].  x_next (batch x 12) and
 * errors (batch) are optional outputs (NULL to skip).  Pointers host or device, handle dtype. */
i7_status i7_find_best_idx(i7_handle* h, int32_t batch, const void* x_last, const void* u_last, const void* xs, double dt,
                           int32_t tie_last, void* x_next, void* errors, void* result, void* stream);

/* i7_find_best_idx for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same
 * arguments, same kernel and same results as i7_find_best_idx, but the scratch memory of the selection (ticket counter
 * + one error per lane, which i7_find_best_idx keeps in buffers it allocates on first use) is a file-scope
 * `static __device__` array of the library: 256 + 8 * I7_STATIC_MAX_BATCH bytes (32.25 KB), held by every GPU that
 * loads the library until the process exits.  The call performs no cudaMalloc / cudaFree and calls no allocating
 * routine (host buffers travel through the staging arenas the handle sized at i7_create).  batch >
 * I7_STATIC_MAX_BATCH is rejected with I7_ERR_INVALID (text in i7_last_error).  ALL calls on a device — from any
 * handle and any stream — share the one array, which keeps the last call's contents: the call zeroes the ticket
 * counter on `stream` before its launch (every error is rewritten before it is read), and two calls that overlap on
 * the GPU (different streams, or different handles without ordering between them) RACE on it — order them. */
#define I7_STATIC_MAX_BATCH 4096
i7_status i7_find_best_idx_static(i7_handle* h, int32_t batch, const void* x_last, const void* u_last, const void* xs,
                                  double dt, int32_t tie_last, void* x_next, void* errors, void* result, void* stream);

/* resample_f_ext_batch (gato_controller.py:120-129) on the device, and its two variants in the drivers (notebook
 * run_mpc_fig8; src/gato_mpc_batch_sample.py:286-298): around the selected lane `best`,
 *     f[b] = f[best] + noise[b];  f[keep] = f[best];  f[:, 3:] = 0;  f[0] = 0;  f *= decay
 * in exactly that order in double precision (bit-identical to the numpy statements).  `noise` (batch x 6) are the
 * caller's normal draws: the reference seeds np.random on the host (gato_controller.py:15) and its stream is what parity
 * means here, so the draws stay with the caller.  best_idx >= 0: the lane; best_idx < 0: the lane is read from
 * `best_result` (the 2-number result of i7_find_best_idx, host or device, handle dtype) — with device buffers the two
 * calls chain on a stream without a host round trip; a result of -1 (no lane qualified) leaves the batch unchanged.
 * keep_lane: a lane index (the notebook keeps lane 1), I7_KEEP_BEST (gato_controller.py:125) or I7_KEEP_NONE.
 * flags: I7_RESAMPLE_KEEP_TORQUE — the kept lane is restored AFTER the torque part is cleared (notebook order);
 * I7_RESAMPLE_APPLY — the new batch also becomes the handle's wrench batch, as by i7_set_external_wrench.
 * f_ext / noise / f_ext_out: batch x 6 doubles [linear; angular], host or device; f_ext_out may equal f_ext. */
#define I7_KEEP_BEST (-1)
#define I7_KEEP_NONE (-2)
#define I7_RESAMPLE_KEEP_TORQUE 1
#define I7_RESAMPLE_APPLY 2
i7_status i7_resample_wrench(i7_handle* h, int32_t batch, int32_t best_idx, const void* best_result, const double* f_ext,
                             const double* noise, int32_t keep_lane, int32_t flags, double decay, double* f_ext_out,
                             void* stream);

/* Hand-over at the end of a controller tick (gato_controller.py:226,249): XU_best = XU_new[best] -> best_out (18N-6
 * numbers), and XU_batch[:, :] = XU_best -> xu_out (batch x (18N-6); may equal xu_new; NULL to skip).  best_idx >= 0: the
 * lane; < 0: read from `best_result`, the result of i7_find_best_idx (host or device, handle dtype); a result of -1 (no
 * lane qualified — the reference would index with None) takes lane 0, the zero-wrench hypothesis.  Pointers host or
 * device.  With i7_goal_window, i7_solve, i7_find_best_idx and i7_resample_wrench this makes the whole
 * GATO_Controller.joint_callback tick (gato_controller.py:201-256) a chain of launches on one stream. */
i7_status i7_broadcast_best(i7_handle* h, int32_t batch, const void* xu_new, int32_t best_idx, const void* best_result,
                            void* xu_out, void* best_out, void* stream);

/* Best-trajectory selection over a batch (north_star: "a single NCCL gather of per-problem costs and best
 * trajectory"; reference analogue gato_controller.py:105-118 argmin over hypotheses).  Writes, on the device,
 * packed = [cost_0 .. cost_{batch-1} | best_cost, best_index (as a number), XU_best (18N-6)] where cost_b =
 * stats[b][8]; `index_offset` is added to best_index (rank * shard size).  Ties -> lowest index.  All pointers
 * host or device; packed has batch + 2 + 18N-6 elements of the handle dtype. */
i7_status i7_pack_best(i7_handle* h, int32_t batch, const void* stats, const void* xu, int32_t index_offset,
                       void* packed, void* stream);
/* Final reduction after the gather: `gathered` holds `nranks` records of `stride` elements, each ending with
 * [best_cost, best_index, XU_best]; writes out = [best_cost, best_index, XU_best] of the winning rank. */
i7_status i7_reduce_best(i7_handle* h, int32_t nranks, int32_t stride, const void* gathered, void* out, void* stream);

/* Multi-GPU exchange over peer memory (one process per GPU on one node), the fused form of
 * i7_pack_best -> all-gather -> i7_reduce_best (the reference analogue is the host loop find_best_trajectory_id,
 * src/gato_mpc_batch_sample.py:254-300; the transport has no reference counterpart):
 *   i7_p2p_create   allocates this rank's exchange buffer (sized for the handle's batch) and returns its 64-byte
 *                   inter-process handle; the caller exchanges the handles of all ranks (any host-side all-gather);
 *   i7_p2p_connect  maps every rank's buffer (handles = world x 64 bytes, in rank order);
 *   i7_p2p_exchange ONE kernel: packs [local costs | best_cost, best_index, XU_best], stores it into every peer's
 *                   buffer over NVLink, waits for all ranks and writes the overall winner [best_cost, global index,
 *                   XU_best] to `out` (device pointer).  `batch` must be the same on every rank.  Every rank must call
 *                   it once per tick; a rank missing for longer than the timeout raises I7_ERR_CUDA at i7_p2p_status.
 *   i7_p2p_gathered device pointer to the world x (batch + 2 + nz) records of the latest exchange
 *   i7_p2p_status   0, or an error once an exchange has timed out (synchronises the stream given).  A timed-out
 *                   exchange writes [NaN, -1, 0...] to `out`, never a stale winner.
 *   i7_p2p_reset    clears the time-out state, the arrival flags and the tick counter of this rank.  COLLECTIVE: every
 *                   rank calls it, with a host-side barrier before and after (no exchange may be in flight). */
#define I7_IPC_HANDLE_BYTES 64
i7_status i7_p2p_create(i7_handle* h, int32_t world, int32_t rank, unsigned char* handle_out);
i7_status i7_p2p_connect(i7_handle* h, const unsigned char* handles);
i7_status i7_p2p_exchange(i7_handle* h, int32_t batch, const void* stats, const void* xu, int32_t index_offset, void* out,
                          void* stream);
i7_status i7_p2p_gathered(i7_handle* h, const void** gathered);
/* i7_solve and i7_p2p_exchange in ONE launch: the CTA that finishes the last problem of the batch runs the selection
 * and the exchange as the kernel's epilogue (ranks that finish early post their record while the others still compute;
 * no second launch).  Same results as the two calls, bit for bit.  Device pointers only; needs i7_p2p_connect (a world
 * of one rank gives the single-GPU selection), the one-CTA-per-problem kernel form (i7_config.cluster = 1 or a batch
 * beyond the latency forms) and I7_QP_EXACT.  best_out: [best_cost, global index, XU_best] (2 + 18N-6 numbers). */
i7_status i7_solve_select(i7_handle* h, int32_t batch, const void* xu_in, const void* x0, const void* goals, int32_t goal_stride,
                          void* xu_out, void* stats, int32_t index_offset, void* best_out, void* stream);
/* Time the epilogues of i7_solve_select have spent waiting for the other ranks, in nanoseconds, summed since the last
 * reset (synchronises the device). */
i7_status i7_exchange_wait_ns(i7_handle* h, int64_t* total_ns, int32_t reset);
i7_status i7_p2p_status(i7_handle* h, void* stream);
i7_status i7_p2p_reset(i7_handle* h);

/* Measurement utility for roofline denominators: sustained FMA throughput of the CUDA-core pipe of the
 * handle's dtype (dependent-chain-free FMA loop on every SM), in TFLOP/s. */
i7_status i7_measure_fma_peak(i7_handle* h, double* tflops);

/* Debugging aid: with enable != 0 every later i7_solve records clock64() at its phase boundaries; clocks16
 * (batch x 32, may be NULL) receives the stamps of the latest solve: per SQP iteration i (0,1) entries
 * 5i+0 start, +1 after setup, +2 after the QP solve, +3 after the merit evaluation, +4 after the update;
 * [10..13] cycles of the first QP solve spent in the 6x6 solves / value updates / roll-out / control recovery;
 * [16..24] stamps at the sub-phase boundaries of the first setup. */
i7_status i7_profile_phases(i7_handle* h, int32_t enable, int64_t* clocks32, int32_t batch);

/* ---- stage-level entry points (parity tests and debugging; same device code as i7_solve) ---- */
/* State of the I7_QP_OSQP solver after the latest i7_solve / i7_solve_qp (host or device pointers, any may be NULL):
 * x: batch x (18N-6) primal iterate; y: batch x 12N duals of the equality rows; y_lim: batch x (18N-18) duals of the
 * limit rows (z order from x_1 on; zeros without limits); info: batch x 4 = [iterations, primal residual, dual
 * residual, status (0 solved, 1 iteration limit)] of the latest QP. */
i7_status i7_qp_state(i7_handle* h, int32_t batch, void* x, void* y, void* y_lim, void* info, void* stream);

/* The full-step QP solution of `setup_and_solve_qp(xu, x0, goals)` (src/osqp_solver.py:137-143): z_out = `.x`
 * (batch x (18N-6)); y_out (batch x 12N, may be NULL) = `.y`, the duals of the equality rows in the row order of the
 * reference's A (block 0: -x_0 = -xs; block k+1: A_k x_k + B_k u_k - x_{k+1} = -c_k), sign convention of OSQP
 * (P z + g + A'y = 0). */
i7_status i7_solve_qp(i7_handle* h, int32_t batch, const void* xu, const void* x0, const void* goals,
                      int32_t goal_stride, void* z_out, void* y_out, void* stream);
/* The QP data in the reference's CSC value order, per problem:
 * Adata: 360(N-1)+144, Pdata: 33N-6... (N*27 + (N-1)*6), g: 18N-6, l: 12N.  Any pointer may be NULL. */
i7_status i7_dump_qp(i7_handle* h, int32_t batch, const void* xu, const void* x0, const void* goals,
                     int32_t goal_stride, void* Adata, void* Pdata, void* g, void* l, void* stream);
/* merit terms of one trajectory against a reference iterate: out[b] = {qcost, vcost, ucost, integrator_err,
 * ||xu[:12]-xu_ref[:12]||, merit} (src/osqp_sqp.py:13-47,52-56). */
i7_status i7_merit(i7_handle* h, int32_t batch, const void* xu, const void* xu_ref, const void* goals,
                   int32_t goal_stride, void* out6, void* stream);
/* line search (src/osqp_sqp.py:49-74): alpha[b] and merits[b] = {base, merit(alpha_1..alpha_8)} (all nine
 * are evaluated; the accepted alpha is the first with merit <= base). */
i7_status i7_linesearch(i7_handle* h, int32_t batch, const void* xu, const void* xu_full, const void* goals,
                        int32_t goal_stride, void* alpha, void* merits9, void* stream);
/* Forward dynamics and its analytic derivatives at n samples (pin.computeABADerivatives,
 * src/osqp_solver.py:71): q, v, tau: n x 6; outputs (any may be NULL) qdd: n x 6; dqdd_dq, dqdd_dv, Minv:
 * n x 36 row-major.  f_ext (n x 6, Pinocchio order, handle dtype) may be NULL. */
i7_status i7_dynamics(i7_handle* h, int32_t n, const void* q, const void* v, const void* tau, const void* f_ext,
                      void* qdd, void* dqdd_dq, void* dqdd_dv, void* Minv, void* stream);
/* End-effector position and translational Jacobian (src/osqp_solver.py:146-155): q: n x 6 -> p: n x 3,
 * J: n x 18 row-major (3 x 6).  J may be NULL. */
i7_status i7_eepos(i7_handle* h, int32_t n, const void* q, void* p, void* J, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* INDY7_B200_H */
