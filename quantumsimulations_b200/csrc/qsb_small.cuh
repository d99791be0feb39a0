// On-chip path (n <= QSB_SMALL_MAX_QUBITS): descriptor shared by host compiler and kernel.
#pragma once
#include "qsb_common.cuh"

namespace qsb {

constexpr int SMALL_MAXQ = QSB_SMALL_MAX_QUBITS;
constexpr int SMALL_MAXPAIR = SMALL_MAXQ * (SMALL_MAXQ - 1) / 2;

struct SmallFlip {  // out[x] += (hxh - i s(x) hyh) psi[x ^ mask],  s = +1 when the bit is 0
  double hxh, hyh;
  uint32_t mask, pad;
};
struct SmallPair {  // out[x] += (equal bits ? c_same : c_diff) psi[x ^ mask]
  double c_same, c_diff;
  uint32_t mask, pad;
};

struct SmallDesc {
  int32_t n, n_flip, n_pair, method;
  int32_t order, t_per_traj, pad0, pad1;
  uint32_t sea_mask, rare_mask, psi0, pad2;
  int64_t nsteps;
  double atol, rtol, max_step, first_step;
  double hz_half[SMALL_MAXQ];      // hz_b / 2
  double jzz_quarter[SMALL_MAXPAIR];  // Jzz_bc / 4, (b<c) row-major
  SmallFlip flip[SMALL_MAXQ];
  SmallPair pair[SMALL_MAXPAIR];
  // dense per-bit / per-bit-pair copies for the register-resident kernel (row-major b<c like jzz_quarter)
  double hx_half[SMALL_MAXQ], hy_half[SMALL_MAXQ];
  double pair_same[SMALL_MAXPAIR], pair_diff[SMALL_MAXPAIR];
};

struct SmallStats {
  int64_t n_accepted, n_rejected, n_rhs;
  double last_h;
  int32_t status, pad;
};

// Launch: one CTA per trajectory, blockDim.x = max(32, 2^max_n) over the batch.  uniform_n != 0 promises
// that every trajectory has exactly max_n qubits, which enables the register-resident row product (n <= 8).
void launch_small_evolve(const SmallDesc* d_desc, int n_traj, int max_n, int uniform_n, const double* d_t_out,
                         int64_t n_out, double* d_obs, SmallStats* d_stats, cudaStream_t stream);

// Propagator build: CTA b integrates basis state b % 2^n of trajectory b / 2^n (all n_traj tables have exactly n qubits)
// over [t_pairs[2 traj], t_pairs[2 traj + 1]] and stores its final state in columns[b][0 .. 2^n), so that
// columns[traj] is U = exp(-i H dt) column-major.  d_stats: [n_traj << n], one per column.  Descriptors must have
// t_per_traj = 1.
void launch_small_columns(const SmallDesc* d_desc, int n_traj, int n, const double* d_t_pairs, SmallStats* d_stats,
                          cplx* d_columns, cudaStream_t stream);

// columns[traj] <- columns[traj]^(2^n_sq[traj]) (dense complex matrix squarings, one CTA per trajectory);
// d_tmp: scratch of the same size as d_columns.
void launch_small_square(cplx* d_columns, cplx* d_tmp, const int32_t* d_n_sq, int n_traj, int n, cudaStream_t stream);

// psi_k = U psi_{k-1} for k < n_out with the observables of every psi_k (one CTA per trajectory, uniform n).
// by_column != 0: d_obs is [n_traj][QSB_NOBS][n_out] instead of [n_traj][n_out][QSB_NOBS].
// Time-split chains: with d_seeds (psi at the first output of each of the n_seg segments of seg_len outputs,
// [n_traj][n_seg][2^n]) the launch has n_traj * n_seg CTAs.  d_states_out != nullptr: store the states instead of the
// observables ([n_traj][n_out][2^n]); that is how the seeds are produced, from U^seg_len and n_out = n_seg.
int launch_small_propagate(const SmallDesc* d_desc, int n_traj, int n, const cplx* d_columns, int64_t n_out,
                           double* d_obs, int by_column, cudaStream_t stream, const cplx* d_seeds = nullptr,
                           int n_seg = 1, int64_t seg_len = 0, cplx* d_states_out = nullptr);
// Launch shape of the propagate kernel for n qubits (host-only; exposed for tests).
void small_propagate_shape(int n, int* threads, int* split, int* n_cache, int* ring, size_t* smem);

// out = H in for one n <= 10 table (test hook: same RHS routine the trajectory kernel uses).
void launch_small_apply(const SmallDesc* d_desc, int n, const cplx* d_in, cplx* d_out, cudaStream_t stream);

// The same H-apply without any device allocation (host vectors in and out, scratch in a file-scope device array of
// small_apply_static_bytes() bytes; see qsb_apply_h_static).  The target device must be current.
cudaError_t launch_small_apply_static(const SmallDesc* h_desc, int n, const cplx* h_in, cplx* h_out, cudaStream_t stream);
size_t small_apply_static_bytes();

}  // namespace qsb
