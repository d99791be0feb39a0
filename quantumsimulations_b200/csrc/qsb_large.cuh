// Tiled HBM path (n >= 12): matrix-free H-apply as three diagonal forms in three bases,
//     H psi = Dz psi + W Dx W psi + S W Dy W S^dagger psi,      W = Hadamard^(x n), S = diag(1,i)^(x n),
// because every term of the table is a product of same-axis spin operators (include/qsb200.h).  The
// Walsh-Hadamard transforms are done 12 bits at a time on 4096-amplitude tiles staged in shared memory;
// bits above 12 are reached by further passes whose tiles gather strided 2^lk-amplitude runs.
#pragma once
#include <cuda.h>

#include <vector>

#include "qsb_common.cuh"

namespace qsb {

constexpr int TB = 12;                 // tile bits
constexpr int TILE = 1 << TB;          // amplitudes per tile
constexpr int TPT = 256;               // threads per tile (16 amplitudes each)
constexpr int TILE_PHYS = TILE + TILE / 16;  // padded length in cplx (bank-conflict-free 16-strided access)
constexpr int MAX_LAYOUTS = 6;
constexpr int MAX_VEC_IN = 8;
constexpr int PSTRIDE = 10;             // doubles per tile in the partial-sum buffer

// One tile layout: tile-local bits [0,lk) = global bits [0,lk); local [lk,12) = global [gs, gs+12-lk).
struct Layout {
  int lk, gs, ge;
  int m8, m4, m0;        // Hadamard stage masks for local-bit windows [8,12), [4,8), [0,4)
  uint64_t tile_mask;    // global bits covered by the tile
};

struct DiagTables {      // device pointers
  const double* h;       // [3][n]
  const double* J;       // [3][n][n]
  // In-tile part of a diagonal form, factored for the register window the pass evaluates it in (InTileDiag):
  const double* Ix;      // [5][256] centre layout, x basis (2^-n folded in)
  const double* Iy;      // [5][256] centre layout, y basis
  const double* Iz;      // [5][256] layout 0, z basis, window 8
  double Wx[16], Wy[16], Wz[16];  // window-window pair sums; live in the kernel parameter (constant) bank
  const double* Tx;      // [n_tiles][40] per-tile rows (window-folded lin/cst) of the centre layout, x basis
  const double* Ty;      //               ... y basis
  const double* Tz;      //               ... layout 0, z basis
};

// Fused epilogue of the last pass: k = -i H v (or H v), then stage algebra on k.
struct Epilogue {
  const cplx* vin[MAX_VEC_IN];
  double c[MAX_VEC_IN];  // o2 = sum c_j vin_j + ck k + cs v     (v = the vector the apply acted on)
  double e[MAX_VEC_IN];  // err = sum e_j vin_j + ek k
  double ck, ek, alpha;  // o1 = alpha k
  double cs;             // weight of v in o2: v is in registers anyway (the Dz term), so adding it costs no traffic
  cplx* o1;
  cplx* o2;
  int n_in;
  int want_err;          // accumulate |err|^2 / sc^2 into partials
  int sc_i0, sc_i1;      // sc = atol + rtol max(|vin[sc_i0]|, sc_i1 >= 0 ? |vin[sc_i1]| : sc_i1 == -1 ? |o2| : |v|)
  int raw_h;             // 1: k = H v (no -i), for qsb_apply_h
  int pf;                // prefetch policy of pass_last (experiment knob QSB200_LAST_PF): bit 0 = tX to L1 instead of L2,
                         // bit 1 = do not prefetch the epilogue's input vectors
  double atol, rtol;
};

constexpr int MAX_WORLD = 8;
constexpr int MAX_PIPE = 16;            // chunk groups of the overlapped sharded pipeline

// Stage combination fused into the load phase of pass_first: the vector the H-apply acts on is FORMED there,
//     v = sum_q c_q vin_q      (e.g. the DP5 stage input y + h sum_j a_sj k_j),
// written once to `out` (pass_last needs it for the Dz term) and transformed in the same pass.  n_in = 0: no
// combination, pass_first just reads its src.
// Optionally the same loads also form a second combination, eout = sum_q e_q vin_q: the embedded pair's error
// estimate over the stages that exist so far (DP5: h sum_{j<=6} e_j k_j while pass_first of the last stage holds
// k_1..k_6 anyway); the last stage's pass_last adds its own k and takes the norm (Epilogue, sc_i1 = -2).
struct Combine {
  const cplx* vin[MAX_VEC_IN];
  double c[MAX_VEC_IN];
  double e[MAX_VEC_IN];
  cplx* out;
  cplx* eout;
  int n_in;
};

// Weighted error sums of an embedded pair, as its own streaming sweep (one CTA per 4096 amplitudes):
//     partials[tile][7] = sum |sum_q e_q vin_q|^2 / sc^2,  sc = atol + rtol max(|vin[sc_i0]|, |vin[sc_i1]|)
struct ErrSpec {
  const cplx* vin[MAX_VEC_IN];
  double e[MAX_VEC_IN];
  int n_in, sc_i0, sc_i1;
  double atol, rtol;
};

struct LargeGeom {
  int n;                 // index bits of the whole state (>= 12)
  int n_loc;             // index bits held by this rank: n - log2(world); the top n - n_loc bits are the rank
  int rank, world;       // world = 1: unsharded
  uint64_t rank_bits;    // rank << n_loc
  int n_layouts;         // 1..MAX_LAYOUTS; layout 0 is the contiguous one, the last one is the centre
  Layout lay[MAX_LAYOUTS];
  uint64_t sea_mask, rare_mask;
  int has_x, has_y;      // basis has any non-zero coefficient
};

// Sharded state: how a rank reaches its peers.  The centre layout's group ends with the rank bits, so the centre
// pass is the ONLY pass that touches other ranks' memory: it loads its tiles straight from the peers' transform
// scratch (P2P loads over NVLink), applies W_g D W_g, and stores them back in place (P2P stores).  `sync` orders
// that pass against the peers' neighbouring passes (drain own stream + a rank barrier); see DESIGN.md.
struct ShardLink {
  cplx* tx[MAX_WORLD];
  cplx* ty[MAX_WORLD];
  // Push-only exchange (QSB200_SHARD_PUSH=1 at plan creation): receive buffers of every rank, [n_tiles_loc][4096] in
  // centre-tile order.  The last forward mid pass stores each amplitude straight into the buffer of the rank that owns
  // its centre tile (posted NVLink writes), so the centre pass loads locally and only its stores are remote.
  cplx* rx[MAX_WORLD];
  cplx* ry[MAX_WORLD];
  bool push;
  int (*sync)(void* ctx);
  void* ctx;
  // Overlapped mode (one process per GPU): `stream_barrier` enqueues a rank barrier ON a stream (an NCCL collective:
  // work queued on that stream afterwards starts only when every rank's earlier work on its counterpart stream is
  // complete) without blocking the host.  The two branches then run as a pipeline: the NVLink-bound centre pass of
  // one branch on `centre_stream` (low priority) under the HBM-bound mid passes of the other on the plan's stream.
  // Null: the serialised path with the host-side `sync` (ranks that are threads sharing GPUs; nothing may spin there).
  int (*stream_barrier)(void* ctx, cudaStream_t stream);
  void* barrier_ctx;
  cudaStream_t centre_stream;
  int centre_ctas;                      // CTAs of the persistent centre launch (QSB200_CENTRE_CTAS; default = SM count)
  cudaEvent_t ev_fwd[MAX_PIPE], ev_centre[MAX_PIPE];  // per chunk group: mids forward + barrier done (plan stream) / centre + barrier done (centre stream)
  int pipe_lb;                          // log2 of the chunk groups the strided passes are pipelined over (QSB200_SHARD_PIPE; default 3)
  // TMA descriptors of the peers' scratch vectors for the ring version of the centre pass, [branch][peer]; built on
  // first use after the link is connected (host cache)
  CUtensorMap ring_maps[2][MAX_WORLD];
  bool ring_ready;
};

// L2-resident fusion of the strided passes f .. centre .. f into one persistent launch (qsb_large_fused.inl).
constexpr int MAX_PHASES = 2 * (MAX_LAYOUTS - 2) + 1;
struct FusedSpec {
  int f, m;                 // first fused layout, centre - f; phases P = 2 m + 1
  int fbits;                // number of index bits a unit fixes
  int tiles_lb;             // log2 T, tiles of one phase of one unit
  int lag;                  // pipeline lag in units between consecutive phases
  unsigned fixed_vmask[MAX_LAYOUTS];  // per layout: the bits of a tile number v that carry the unit id
  unsigned vmask;           // n_tiles - 1
};
struct StridedFusion {
  bool on = false;
  FusedSpec spec = {};
  int grid = 0;             // persistent CTAs (1 per SM)
  int* ctrl = nullptr;      // device: ticket + per (unit, phase) counters; owned by the plan
  size_t ctrl_bytes = 0;
  // TMA descriptors of the scratch vectors, [branch][layout], rebuilt when the vectors change (host cache)
  CUtensorMap maps[2][MAX_LAYOUTS];
  const void* maps_tx = nullptr;
  const void* maps_ty = nullptr;
};
// Decides whether (and from which layout on) the strided passes of geometry g are fused: the smallest f whose unit
// is at most 2^unit_lb amplitudes.  Unsharded plans with >= 3 layouts only.  Off unless QSB200_FUSE_LB sets unit_lb
// (20 = 16 MB units); QSB200_FUSE_LAG overrides the pipeline lag.  Does not allocate: the caller provides ctrl.
StridedFusion plan_fusion(const LargeGeom& g, int n_sms);

// Layouts over n index bits of which the top log2(world) are rank bits (they end up inside the centre group).
LargeGeom make_geometry(int n, int rank = 0, int world = 1);
size_t tile_diag_doubles(int n_loc);
// d_h/d_J: device [n] / [n][n] of ONE basis.  Fills d_out[n_tiles_loc][40] for tiles v0 .. v0 + n_tiles_loc - 1 of
// layout L; extra_bits are index bits outside the address space of this rank that belong to every such tile.
// W: the register window of the pass that will consume the rows (centre_window(L) for the centre, 8 for layout 0).
void launch_build_tile_diag(const double* d_h, const double* d_J, int n, const Layout& L, int W, double scale,
                            double* d_out, uint64_t n_tiles_loc, uint64_t v0, uint64_t extra_bits, cudaStream_t stream);
// Per-device kernel attributes (dynamic shared memory opt-in); call once per device before launching.
void prepare_device();
// In-tile diagonal D_in(u) = sum_b h_b s_b/2 + sum_{b<c} J_bc s_b s_c/4 over the 12 tile bits, factored for a thread
// that holds the 16 amplitudes u = i0(tt) | j << W (window bits [W, W+4), the other 8 bits fixed by tt):
//   D_in = A[tt] + sum_k s_k(j) Lk[tt] + Ww[j],   tab = {A[256], L0[256], L1[256], L2[256], L3[256]}, Ww[16].
// 5 coalesced loads per thread instead of 16 gathers from a 4096-entry table, and far fewer live registers.
struct InTileDiag {
  double tab[5 * 256];
  double ww[16];
};
InTileDiag host_in_tile_diag(const double* h, const double* J, int n, int basis, const Layout& L, int W, double scale);
// Register window in which pass_centre evaluates the diagonal for layout L (8, 4 or 0).
int centre_window(const Layout& L);

// One H-apply: src -> epilogue.  tx/ty: scratch vectors of 2^n cplx.  partials: device doubles,
// [n_tiles][PSTRIDE] = (nrm2, zsea, zrare, xsea, xrare, ysea, yrare, err, aux).  collect_obs: fill the observable slots.
// link: null when unsharded.  Returns a qsb_status (only the rank barrier can fail).
// comb: optional stage combination; when comb->n_in > 0 the apply acts on sum_q c_q vin_q, which is also written
// to comb->out, and `src` must equal comb->out.  Epilogues take at most 2 input vectors.
// next: when given, the apply ends in pass_last_first instead of pass_last - it writes k = -i H v to ep.o1 AND begins
// the next apply (forms next->cb + c_fresh k into next->cb.out, which must be `src`, and its tX / tY); the next call
// then passes skip_first.  Plain k-producing epilogues only (no o2, no error sums, alpha = 1).
struct FuseNext {
  Combine cb;      // the next stage's combination WITHOUT the fresh k
  double c_fresh;  // weight of the fresh k in it
};
int launch_happly(const LargeGeom& g, const DiagTables& tab, const cplx* src, cplx* tx, cplx* ty,
                  const Epilogue& ep, int collect_obs, double* partials, cudaStream_t stream,
                  ShardLink* link = nullptr, const Combine* comb = nullptr, StridedFusion* fu = nullptr,
                  const FuseNext* next = nullptr, bool skip_first = false);
void launch_error_sums(const LargeGeom& g, const ErrSpec& es, double* partials, cudaStream_t stream);
// Observables only (pass_first + mids + pass_centre), filling partial slots 0..6.
int launch_observe(const LargeGeom& g, const DiagTables& tab, const cplx* src, cplx* tx, cplx* ty,
                   double* partials, cudaStream_t stream, ShardLink* link = nullptr,
                   StridedFusion* fu = nullptr);
// out[PSTRIDE] = column sums of partials[n_tiles][PSTRIDE] in a fixed order.
void launch_reduce_partials(const double* partials, int64_t n_tiles, double* out8, cudaStream_t stream);
// psi <- basis state (index < dim) or zero (index >= dim)
void launch_set_basis_state(cplx* psi, uint64_t dim, uint64_t index, cudaStream_t stream);
// Product state psi[x] = prod_b phi.v[b][bit_b(x)] over this rank's 2^n_loc amplitudes (bits >= n_loc from rank_bits),
// and a row gather; together with one H-apply they are the closed-form parity check at sizes the oracle cannot hold.
struct ProductPhi { cplx v[QSB_MAX_QUBITS][2]; };
void launch_product_state(cplx* psi, const ProductPhi& phi, int n, int n_loc, uint64_t rank_bits, cudaStream_t stream);
void launch_gather_rows(const cplx* src, const uint64_t* d_rows, int64_t n_rows, cplx* d_out, cudaStream_t stream);
// Dense output of a Taylor step: out = y + sum_{m < n} c[m] term[m]  (c[m] = s^(m+1), s = fraction of the step).
struct DenseSpec {
  const cplx* y;
  const cplx* term[QSB_MAX_TAYLOR_ORDER];
  double c[QSB_MAX_TAYLOR_ORDER];
  cplx* out;
  int n;
};
void launch_dense_eval(const DenseSpec& d, uint64_t dim, cudaStream_t stream);
// ---- HBM-resident propagator (11 <= n <= 14 qubits; qsb_propagate_hbm) -------------------------------------------------
// psi[col * 2^n + row] = (row == col): the identity as 2^n basis-state columns, i.e. the start of the integration that
// yields ALL columns of exp(-i H t) in one go (a 2n-qubit plan whose upper n bits are decoupled spectators).
void launch_set_identity(cplx* psi, int n, cudaStream_t stream);
// Observables of n_cols states of 2^n amplitudes stored as columns (x[col * 2^n + row]); column c is the state at output
// index o = c * seg_len + k and is skipped when o >= n_out.  out: [n_out][QSB_NOBS], or [QSB_NOBS][n_out] (by_column).
void launch_observe_columns(const cplx* x, int n, int n_cols, int64_t seg_len, int64_t k, int64_t n_out, uint64_t sea_mask,
                            uint64_t rare_mask, double* out, int by_column, cudaStream_t stream);
// Measured FP64 FMA rate of the current device in TFLOP/s (8 independent DFMA chains per thread, all SMs, best of 3).
double measure_fp64_peak_tflops(cudaStream_t stream);
size_t happly_smem_bytes();

// Per-kernel-class CUDA-event timing (bench.py's roofline leg).  Classes: 0 pass_first, 1 pass_mid, 2 pass_centre,
// 3 pass_last, 4 pass_first with a fused stage combination, 5 error_sums, 6 the fused strided passes, 7 pass_last_first.  bytes = what the launch moves by design
// (vectors read + written, in bytes), so ms and bytes of a class give its achieved GB/s.
constexpr int N_PROFILE = 8;
struct Profile {
  bool on = false;
  double ms[N_PROFILE] = {};
  long long count[N_PROFILE] = {};
  double bytes[N_PROFILE] = {};
};
Profile& profile();

}  // namespace qsb
