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

// out = H in for one n <= 10 table (test hook: same RHS routine the trajectory kernel uses).
void launch_small_apply(const SmallDesc* d_desc, int n, const cplx* d_in, cplx* d_out, cudaStream_t stream);

}  // namespace qsb
