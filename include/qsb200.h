/*
 * qsb200.h - C ABI of the B200-native spin-ensemble time-evolution engine (libqsb200.so).
 *
 * The reference (TimHarrelson/QuantumSimulations) has no FFI: its hot path is the Python call
 *     qt.sesolve(H, psi0, t, e_ops=[6 ops], options)        dipolar_ensemble_with_rare.py:653-666
 * on the operator assembled by build_hamiltonian_rare()     dipolar_ensemble_with_rare.py:453-588.
 * These entry points are what a ctypes binding of that call would bind; each one cites the reference
 * lines it replaces.  Plain pointers and sizes only; no torch / Python types.  All arithmetic is
 * IEEE float64 / complex128 (interleaved re,im).  Every function returns a qsb_status; on failure
 * qsb_last_error() holds a message (thread-local).  There is NO CPU fallback: without a CUDA device
 * every compute entry point fails with QSB_ERR_NO_DEVICE.
 *
 * Index convention (reference: qt.tensor ordering, dipolar_ensemble_with_rare.py:43-45): site k of
 * the reference is BIT n-1-k of the amplitude index; bit value 0 is Iz=+1/2.  Everything in this ABI
 * is indexed by bit.
 */
#ifndef QSB200_H
#define QSB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QSB_VERSION 200
#define QSB_MAX_QUBITS 34
#define QSB_SMALL_MAX_QUBITS 10 /* whole trajectory on-chip (one CTA) up to here */
#define QSB_HBM_PROPAGATOR_MAX_QUBITS 14 /* dense propagator resident in HBM up to here (1 GB at 13, 4 GB at 14) */
#define QSB_NOBS 8
#define QSB_MAX_TAYLOR_ORDER 30

typedef enum {
  QSB_OK = 0,
  QSB_ERR_BAD_ARG = 1,
  QSB_ERR_NSTEPS = 2,         /* > nsteps internal steps inside one output interval (QuTiP: "Excess work") */
  QSB_ERR_STEP_UNDERFLOW = 3, /* step size fell below 16 ulp of t                                         */
  QSB_ERR_CUDA = 4,
  QSB_ERR_OOM = 5,
  QSB_ERR_NO_DEVICE = 6,
  QSB_ERR_UNSUPPORTED = 7
} qsb_status;

/* Integrators.  Both control the local error in the zvode sense: weighted RMS over complex components,
 * weight 1/(atol + rtol*|y_i|), accept when <= 1 (reference knobs: options["atol"/"rtol"],
 * dipolar_ensemble_with_rare.py:644-647). */
typedef enum {
  QSB_METHOD_DP5 = 0,   /* Dormand-Prince 5(4), FSAL, 6 H-applies per step attempt                          */
  QSB_METHOD_TAYLOR = 1 /* order-p Taylor pair p(p-1) for the constant-H case, p H-applies per step attempt */
} qsb_method;

/* Coupling table: H = sum_b h[a][b] I^a_b + sum_{b<c} J[a][b][c] I^a_b I^a_c, a in {x,y,z}, I = sigma/2.
 * Replaces the 2^N x 2^N CSR operator of build_hamiltonian_rare (dipolar_ensemble_with_rare.py:504-570). */
typedef struct {
  int32_t n_qubits;   /* 1 .. QSB_MAX_QUBITS                                                     */
  const double* h;    /* [3][n]      fields (rad/s)                                              */
  const double* J;    /* [3][n][n]   symmetric pair couplings (rad/s); diagonal ignored          */
  uint64_t sea_mask;  /* bits summed into Ix_sea, Iy_sea, Iz_sea  (dipolar_ensemble_with_rare.py:573-575) */
  uint64_t rare_mask; /* single bit of Iz_R, Ix_R, Iy_R           (dipolar_ensemble_with_rare.py:576-578) */
} qsb_table;

/* Solver knobs = the reference's options dict (dipolar_ensemble_with_rare.py:639-651). */
typedef struct {
  double atol;      /* > 0                                                                          */
  double rtol;      /* >= 0                                                                         */
  double max_step;  /* hard cap on h; 0 = none                                                      */
  int64_t nsteps;   /* max internal steps per output interval (zvode MXSTEP semantics); 0 = no cap  */
  int32_t method;   /* qsb_method                                                                   */
  int32_t order;    /* Taylor order p (4..QSB_MAX_TAYLOR_ORDER); ignored for DP5; 0 = default 20    */
  double first_step;/* 0 = automatic                                                                */
} qsb_solver_opts;

typedef struct {
  int64_t n_accepted;
  int64_t n_rejected;
  int64_t n_rhs;     /* H-applies                                              */
  double last_h;
  double device_ms;  /* CUDA-event time of the whole evolve on its stream      */
  int32_t path;      /* 0 = on-chip small-N kernel, 1 = tiled HBM path, 2 = on-chip propagator (counters summed
                        over the 2^n column trajectories that built U), 3 = HBM-resident propagator, 4 = on-chip
                        propagator + Pauli-string observables (qsb_propagate_expect) */
  int32_t status;    /* qsb_status of this trajectory (batch calls)            */
} qsb_stats;

typedef struct qsb_plan qsb_plan;

/* ---- library / device ------------------------------------------------------------------------ */
int qsb_version(void);
const char* qsb_last_error(void);
int qsb_device_count(int* count);

/* ---- plans: one compiled table on one device --------------------------------------------------- */
/* Replaces build_hamiltonian_rare's operator assembly (dipolar_ensemble_with_rare.py:453-588).       */
int qsb_plan_create(const qsb_table* table, int device, qsb_plan** plan);
int qsb_plan_destroy(qsb_plan* plan);
/* Device bytes the tiled path needs for `method` (state, stages, transform scratch). 0 for n <= 10.  */
int qsb_plan_workspace_bytes(const qsb_plan* plan, int method, int order, size_t* bytes);
/* Optional: hand the plan a caller-owned device buffer (e.g. a torch allocation) instead of letting it
 * cudaMalloc.  Must be >= qsb_plan_workspace_bytes and 256-byte aligned; caller keeps ownership.      */
int qsb_plan_set_workspace(qsb_plan* plan, void* device_ptr, size_t bytes);

/* ---- the hot path ------------------------------------------------------------------------------ */
/* Replaces qt.sesolve(H, psi0, t, e_ops, options) + the norm loop (dipolar_ensemble_with_rare.py:653-669).
 * psi0: computational-basis index (psi0_host == NULL; initial_state_rare, :591-606) or a full host vector
 * of 2^n complex128.  t_out[0] is the start time.  out_obs is HOST memory, [n_out][QSB_NOBS], columns
 * Ix_sea, Iy_sea, Iz_sea, Iz_R, Ix_R, Iy_R (all on the output-normalised ket, as QuTiP reports them),
 * state_norm (norm after that normalisation), raw_norm (the integrator's own norm, which QuTiP drops). */
int qsb_evolve(qsb_plan* plan, uint64_t psi0_index, const double* psi0_host, int64_t n_out,
               const double* t_out, const qsb_solver_opts* opts, double* out_obs, qsb_stats* stats);

/* Many independent small trajectories (n <= QSB_SMALL_MAX_QUBITS) in ONE launch, one CTA each: the
 * sweep's serial loop over detunings x variants (sweep_sea_detuning.py:611, :694-702).  tables[i] may
 * differ in every field incl. n_qubits.  t_out is [n_traj][n_out] when t_out_per_traj != 0, else
 * [n_out] shared.  out_obs: [n_traj][n_out][QSB_NOBS] host.  opts/stats: [n_traj].                   */
int qsb_evolve_batch(int32_t n_traj, const qsb_table* tables, const uint64_t* psi0_index, int64_t n_out,
                     const double* t_out, int32_t t_out_per_traj, const qsb_solver_opts* opts,
                     double* out_obs, qsb_stats* stats, int device);

/* Constant-H propagator path for UNIFORM output grids (n <= QSB_SMALL_MAX_QUBITS, same n in every table of a call).
 * The reference's grid is np.linspace(0, t_final, steps) (dipolar_ensemble_with_rare.py:626) and its Hamiltonian has
 * no time dependence (:514, :570), so all output intervals share ONE propagator U = exp(-i H dt):
 *   1. the 2^n columns of exp(-i H dt / 2^s) are integrated by the adaptive integrator opts[i] (same kernel, same
 *      error control and nsteps / max_step semantics as qsb_evolve_batch; one CTA per column, all trajectories of the
 *      call in one launch), and the matrix is squared s times on chip (scaling and squaring).  s is chosen per
 *      trajectory so that ||H|| dt / 2^s <= ~8 (qsb_debug_squarings), capped by max_squarings (< 0: no cap; 0: the
 *      columns are integrated over the whole interval);
 *   2. psi_k = U psi_{k-1} for k = 1 .. n_out-1 with the observables of every psi_k reduced exactly as
 *      qsb_evolve_batch reports them (U stays in shared memory / registers / L2).  Chains of >= 256 outputs are cut
 *      into ~128 segments of 2^k outputs, each its own CTA, seeded with (U^(2^k))^c psi_0; k depends on n_out only.
 * A 20000-output, 30-second trajectory (sweep_sea_detuning.py:1201-1252) then costs 2^n integrations over a few
 * integrator steps, ~20 dense squarings and 2 * 10^4 dense 2^n x 2^n matvecs instead of ~10^9 integrator steps.
 * Output k is the state at t0 + k dt for any t0; dt is [n_traj] when dt_per_traj != 0, else one shared value.  out_obs (host): [n_traj][n_out][QSB_NOBS], or
 * [n_traj][QSB_NOBS][n_out] when obs_by_column != 0 (one contiguous trace per observable, the shape simulate_rare
 * hands back, dipolar_ensemble_with_rare.py:671-679).                                                              */
int qsb_propagate_batch(int32_t n_traj, const qsb_table* tables, const uint64_t* psi0_index, int64_t n_out,
                        const double* dt, int32_t dt_per_traj, const qsb_solver_opts* opts, int32_t max_squarings,
                        int32_t obs_by_column, double* out_obs, qsb_stats* stats, int device);
/* The same constant-H propagator path one size class up, QSB_SMALL_MAX_QUBITS < n <= QSB_HBM_PROPAGATOR_MAX_QUBITS, ONE
 * trajectory per call, U = exp(-i H dt) resident in HBM (1 GB at n = 13, 4 GB at 14).  The reference's sweep function defaults to
 * n_sea = 12, i.e. n = 13 (sweep_sea_detuning.py:356-376), where a full-size trajectory (30 s, 20000 outputs, ||H|| dt ~
 * 10^3) is hours of stepping.  All 2^n columns are integrated at once over dt / 2^s by the tiled path's adaptive
 * integrator (the matrix is the state of a 2n-qubit plan whose upper n bits are decoupled; opts as in
 * qsb_propagate_batch), s squarings, U^seg_len for the time-split seeds and the chain X <- U X over all segments are
 * cuBLAS Zgemm calls, the observables of every state come from one kernel per output.  out_obs / obs_by_column / the
 * n_rhs counter (column H-applies) as in qsb_propagate_batch; stats->path = 3.  Needs 5 x 16 x 4^n bytes of HBM.     */
int qsb_propagate_hbm(const qsb_table* table, uint64_t psi0_index, int64_t n_out, double dt, const qsb_solver_opts* opts,
                      int32_t max_squarings, int32_t obs_by_column, double* out_obs, qsb_stats* stats, int device);
/* qsb_propagate_batch keeps its stream and device buffers (up to a few GiB) per device between calls; this frees
 * them (device < 0: on every device).  Safe to call at any time; the next call re-allocates.                        */
int qsb_release_caches(int device);
/* Step 1 alone, for parity tests: u_host[(j * 2^n + x) * 2 + {0,1}] = re, im of <x| exp(-i H dt) |j>.              */
int qsb_build_propagator(const qsb_table* table, double dt, const qsb_solver_opts* opts, int32_t max_squarings,
                         double* u_host, qsb_stats* stats, int device);
/* Host-only: the number of squarings step 1 uses for this table and interval when max_squarings < 0.               */
int qsb_debug_squarings(const qsb_table* table, double dt, int32_t* n_squarings);
/* Host-only: launch shape of step 2 for n qubits, out = {threads, column split, columns of U cached in shared
 * memory, dynamic shared-memory bytes, slots of the state ring between two rounds of observable reductions}.   */
int qsb_debug_propagate_shape(int32_t n_qubits, int32_t* out /* [5] */);

/* ---- pieces of the path, exposed for parity tests and the benchmark ---------------------------- */
/* out = H psi on host vectors of 2^n complex128 (one H-apply through the same kernels evolve uses).  */
int qsb_apply_h(qsb_plan* plan, const double* psi_host, double* out_host);
/* qsb_apply_h WITHOUT any device allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same inputs, outputs and arithmetic as qsb_apply_h on an n_qubits <= QSB_STATIC_MAX_QUBITS plan (host
 * vectors of 2^n complex128), but the descriptor / input / output device buffers that qsb_apply_h cudaMallocs per call
 * live in ONE file-scope `static __device__` array of QSB_STATIC_SCRATCH_BYTES (36 KB, held by every GPU that loads the
 * library until exit).  It calls no cudaMalloc, cudaFree or allocating library routine.  Larger plans are rejected with
 * QSB_ERR_UNSUPPORTED.  Every call on a device shares that array, which keeps the previous call's contents (the
 * output slot is cleared at the start of each call): the call blocks until its result is on the host, but calls that
 * overlap on one GPU (several host threads, several plans) RACE on it - serialise them.                          */
#define QSB_STATIC_MAX_QUBITS 10
#define QSB_STATIC_SCRATCH_BYTES (4096 + 2 * 16 * 1024)
int qsb_apply_h_static(qsb_plan* plan, const double* psi_host, double* out_host);
/* The same H-apply at sizes no host vector reaches (BASELINE configs C4 / C5): the device fills the state with the
 * PRODUCT state psi[x] = prod_b phi[b][bit_b(x)] (phi: [n][2] complex128, interleaved re,im), applies H through the
 * kernels qsb_evolve uses, and returns rows `rows[i]` (indices into THIS plan's slice; collective when sharded) of
 * H psi in out_host [n_rows] complex128.  On a product state every row of H psi has an O(n^2) closed form
 * (oracle/product.py), which is what pins the operator of dipolar_ensemble_with_rare.py:549-568 at N = 28 .. 33.  */
int qsb_apply_h_product(qsb_plan* plan, const double* phi, const uint64_t* rows, int64_t n_rows, double* out_host);
/* Stepping session with the state resident in HBM (tiled path, n > QSB_SMALL_MAX_QUBITS).            */
int qsb_session_begin(qsb_plan* plan, uint64_t psi0_index, const double* psi0_host, double t0,
                      const qsb_solver_opts* opts);
/* Advance exactly n_attempts step attempts (accepted or rejected) - the benchmark's unit of work.
 * device_ms = CUDA-event time on the plan's stream; happly_ms = summed event time of the H-apply
 * kernels only (0 when timing of sub-ranges is off).                                                 */
int qsb_session_step(qsb_plan* plan, int64_t n_attempts, double* t_now, qsb_stats* stats, double* device_ms);
/* Advance the open session to exactly t_target (steps land on it; zvode's nsteps cap applies to this interval).
 * With qsb_session_state / qsb_session_begin(psi0_host, t0, opts.first_step = stats.last_h) this is what trajectory
 * checkpoint / resume is made of: the reference restarts a 10^9-step trajectory from zero
 * (sweep_sea_detuning.py:483-485 names a fresh directory per run and keeps no solver state).                     */
int qsb_session_advance(qsb_plan* plan, double t_target, double* t_now, qsb_stats* stats);
/* Observables of the current state, [QSB_NOBS] host.                                                 */
int qsb_session_observe(qsb_plan* plan, double* obs);
/* Copy the current state to host (2^n complex128).                                                   */
int qsb_session_state(qsb_plan* plan, double* psi_host);
int qsb_session_end(qsb_plan* plan);

/* ---- local dimension 4: the spin-3/2 rare spin on two index bits ----------------------------------- */
/* The reference declares a J = 3/2 rare spin (4x4 _Jx, _Jy, _Jz, dipolar_ensemble_with_rare.py:22-25, used at
 * :499-501, :523-528, :563-568, :576-578; `is_spin_three_half=True` is the dataclass default, :377), although its own
 * builder cannot run it (:486 forces all local dimensions to 2).  Here the four-level site is the two LOWEST index bits,
 * level m = 2 a + b with m = 0 <-> J_z = +3/2 (qt.tensor order: the rare site is the last factor):
 *     J_z = 2 I^z_a + I^z_b,    J_x = sqrt(3) I^x_b + 2 (I^x_a I^x_b + I^y_a I^y_b),
 *     J_y = sqrt(3) I^y_b + 2 (I^y_a I^x_b - I^x_a I^y_b).
 * In the frame rotating about J_z by the drive phase phi_R the rare drive w1R (cos phi_R J_x + sin phi_R J_y) becomes
 * w1R J_x, so the whole Hamiltonian fits qsb_table's same-axis form and the time evolution runs through the kernels of
 * the spin-1/2 path.  The observables do not: <J_x>, <J_y> are two-bit (one of them mixed-axis) Pauli strings.  They are
 * given as a list of weighted strings; output column `column` (0 .. 5 = Ix_sea, Iy_sea, Iz_sea, Iz_R, Ix_R, Iy_R)
 * receives sum_t weight_t <psi|P_t|psi> / <psi|psi>,  P = i^popcount(x_mask & z_mask) X^x_mask Z^z_mask  (a bit in
 * both masks carries Y = i X Z); columns 6 / 7 are state_norm / raw_norm as everywhere else.                        */
typedef struct {
  uint64_t x_mask;  /* bits carrying X or Y                                       */
  uint64_t z_mask;  /* bits carrying Z or Y                                       */
  double weight;
  int32_t column;   /* 0 .. 5                                                     */
  int32_t reserved; /* 0                                                          */
} qsb_pauli_term;
#define QSB_MAX_PAULI_TERMS 512
/* qsb_propagate_batch for ONE trajectory with such observables, n <= QSB_SMALL_MAX_QUBITS (n_sea <= 8 with a spin-3/2
 * rare spin): the same column integrations and on-chip squarings build U = exp(-i H dt); the matvec chain keeps every
 * psi_k in device memory (n_out * 2^n * 16 bytes, at most 16 GiB) and one more kernel reduces the strings on all of
 * them.  out_obs / obs_by_column / stats as in qsb_propagate_batch; stats->path = 4.                               */
int qsb_propagate_expect(const qsb_table* table, uint64_t psi0_index, int64_t n_out, double dt, const qsb_solver_opts* opts,
                         int32_t max_squarings, const qsb_pauli_term* terms, int32_t n_terms, int32_t obs_by_column,
                         double* out_obs, qsb_stats* stats, int device);
/* The same for n_traj trajectories of equal n_qubits that SHARE the list of strings (a sweep: same n_sea and phi_R,
 * sweep_sea_detuning.py:611-702) in one pass: arguments as qsb_propagate_batch, out_obs [n_traj][n_out][QSB_NOBS] or
 * [n_traj][QSB_NOBS][n_out].  Trajectories are taken in chunks whose kept states stay below ~8 GiB.                 */
int qsb_propagate_expect_batch(int32_t n_traj, const qsb_table* tables, const uint64_t* psi0_index, int64_t n_out,
                               const double* dt, int32_t dt_per_traj, const qsb_solver_opts* opts, int32_t max_squarings,
                               const qsb_pauli_term* terms, int32_t n_terms, int32_t obs_by_column, double* out_obs,
                               qsb_stats* stats, int device);
/* qsb_propagate_hbm with such observables, QSB_SMALL_MAX_QUBITS < n <= QSB_HBM_PROPAGATOR_MAX_QUBITS (9 <= n_sea <= 12 with a
 * spin-3/2 rare spin; n_sea = 12 is the reference's dataclass default, dipolar_ensemble_with_rare.py:332-384, and its sweep
 * function's default, sweep_sea_detuning.py:363): same columns, cuBLAS squarings and chain; the strings are reduced on
 * the chain's state buffer after every matvec.  stats->path = 3.                                                    */
int qsb_propagate_hbm_expect(const qsb_table* table, uint64_t psi0_index, int64_t n_out, double dt, const qsb_solver_opts* opts,
                             int32_t max_squarings, const qsb_pauli_term* terms, int32_t n_terms, int32_t obs_by_column,
                             double* out_obs, qsb_stats* stats, int device);
/* The same observables on the current state of an open session (tiled path, unsharded plans), [QSB_NOBS] host:
 * qsb_session_advance to every output time + this call is the spin-3/2 trajectory for n > QSB_SMALL_MAX_QUBITS.    */
int qsb_session_expect(qsb_plan* plan, const qsb_pauli_term* terms, int32_t n_terms, double* obs);


/* ---- sharded state: one trajectory whose 2^n amplitudes do not fit one GPU ------------------------ */
/* (BASELINE config C5; no counterpart in the reference, whose sesolve call - dipolar_ensemble_with_rare.py:653 -
 * holds the whole ket in one process.)  The amplitude index's top log2(world) bits are the rank: rank r holds
 * amplitudes [r * 2^(n-log2 world), (r+1) * 2^(n-log2 world)).  One qsb_plan per rank (normally one process per
 * GPU); all ranks then make the SAME sequence of calls (qsb_session_*, qsb_evolve, qsb_apply_h) collectively, with
 * psi0_host / psi_host / out_host holding only the rank's own slice.  Every pass of the H-apply is local except the
 * centre pass, which reads and writes the peers' transform scratch directly over NVLink (CUDA IPC / peer access);
 * scalars (error norm, observable sums) are combined through the caller's all-gather hook in rank order, so every
 * rank takes bit-identical step decisions.                                                                        */

/* Gather `count` doubles from every rank: recv[r*count .. (r+1)*count) = rank r's send.  count may be 0 (a pure
 * barrier).  Return 0 on success.  In one process per GPU this is a torch.distributed / NCCL / MPI all-gather; for
 * ranks living in one process (threads) use qsb_local_group_*.                                                  */
typedef int (*qsb_allgather_fn)(void* ctx, const double* send, int32_t count, double* recv);

#define QSB_SHARD_BLOB_BYTES 96
int qsb_shard_plan_create(const qsb_table* table, int device, int32_t rank, int32_t world, qsb_plan** plan);
/* Describe this rank's exchange scratch (pid, device, address, CUDA IPC handle) in QSB_SHARD_BLOB_BYTES bytes. */
int qsb_shard_export(qsb_plan* plan, void* blob);
/* blobs: [world][QSB_SHARD_BLOB_BYTES], rank-ordered, as gathered from every rank's qsb_shard_export.          */
int qsb_shard_connect(qsb_plan* plan, const void* blobs, qsb_allgather_fn allgather, void* ctx);
/* Optional, one process per GPU only: a rank barrier that is QUEUED ON A CUDA STREAM instead of blocking the host -
 * `stream` is a cudaStream_t; work queued on it after the call must not start before every rank's work queued on its
 * counterpart stream before ITS call is complete (e.g. a one-element NCCL all-reduce issued on that stream).  With it
 * the H-apply runs its two branches as a pipeline: the NVLink-bound centre pass of one branch on a second,
 * low-priority stream under the HBM-bound passes of the other; every rank queues the barriers in the same order.
 * Without it (fn = NULL, the default) the passes are serialised around host-side barriers through the all-gather
 * hook, which is what ranks that share a GPU (threads, tests) must use: kernels that wait on one another as separate
 * launches on one GPU may never be co-resident.                                                                 */
typedef int (*qsb_stream_barrier_fn)(void* ctx, void* stream);
int qsb_shard_set_stream_barrier(qsb_plan* plan, qsb_stream_barrier_fn fn, void* ctx);
/* The plan's own all-gather hook, for host-side agreement between the ranks (checkpoint / resume: every rank must
 * pick the trajectory up at the same output time).  recv: [world][count]; count = 0 is a barrier.              */
int qsb_shard_allgather(qsb_plan* plan, const double* send, int32_t count, double* recv);
/* Host-only (no device needed): the tile layouts the tiled path uses for n index bits split over `world` ranks.
 * out[k] = {lk, gs, ge} for k < *n_layouts (<= 6): tile-local bits [0,lk) are index bits [0,lk), tile-local bits
 * [lk,12) are index bits [gs,ge).  Layout 0 is the contiguous one; the last one is the centre and must contain all
 * rank bits [n - log2 world, n).                                                                               */
int qsb_debug_layouts(int32_t n_qubits, int32_t world, int32_t* n_layouts, int32_t* out /* [6][3] */);
/* Host-only: how the strided passes (mid .. centre .. mid) of an unsharded n-qubit plan are fused into one L2-resident
 * persistent launch on a device with n_sms SMs.  out = {on, f, m, fbits, tiles_lb, lag, vmask, fixed_vmask[6]}:
 * layouts f .. f+m (the centre) are fused, a unit fixes fbits index bits and has 2^tiles_lb tiles per phase, phases of
 * one unit are `lag` units apart in the ticket order, and tile v of fused layout k belongs to the unit whose id is the
 * bits of v selected by fixed_vmask[k].  on = 0: every pass is its own launch.                                  */
int qsb_debug_fusion(int32_t n_qubits, int32_t n_sms, int64_t* out /* [13] */);
/* Amplitudes held by this plan (2^n for an ordinary plan).                                                     */
int qsb_plan_local_dim(const qsb_plan* plan, uint64_t* dim);

/* All-gather for `world` ranks that are threads of this process (tests; one process driving several GPUs).      */
typedef struct qsb_local_group qsb_local_group;
int qsb_local_group_create(int32_t world, qsb_local_group** group);
int qsb_local_group_destroy(qsb_local_group* group);
void* qsb_local_group_member(qsb_local_group* group, int32_t rank); /* the ctx to pass with qsb_local_allgather */
int qsb_local_allgather(void* member_ctx, const double* send, int32_t count, double* recv);
/* A qsb_stream_barrier_fn for such ranks: drains the stream and joins the thread barrier ON THE HOST (no kernel waits
 * on another, so the ranks may share a GPU).  It serialises what an NCCL barrier overlaps; it exists so that the
 * pipelined H-apply of qsb_shard_set_stream_barrier can be tested on a one-GPU box.                              */
int qsb_local_stream_barrier(void* member_ctx, void* stream);

/* Benchmark aid: when enabled, every launch of the tiled path is bracketed by CUDA events on its stream (this
 * serialises host and device; never enable it inside a timed throughput run).  Classes: 0 pass_first, 1 pass_mid,
 * 2 pass_centre, 3 pass_last, 4 pass_first with a fused stage combination, 5 error_sums, 6 strided_fused (mid .. centre
 * .. mid as one L2-resident persistent launch), 7 pass_last_first (a DP5 stage's pass_last fused with the next stage's pass_first).  ms / count / bytes: summed
 * milliseconds, launches and bytes moved by design (vectors read + written) per class since the last enable.        */
/* Benchmark aid: measured FP64 FMA rate of `device` in TFLOP/s (DFMA-chain microbenchmark over all SMs) - the
 * denominator of the FP64 co-bound bench.py reports next to the HBM roofline.                                       */
int qsb_debug_fp64_peak(int device, double* tflops);
#define QSB_N_PROFILE 8
int qsb_profile(int enable);
int qsb_profile_read(double* ms, int64_t* count, double* bytes, int32_t n);

#ifdef __cplusplus
}
#endif
#endif /* QSB200_H */
