// On-chip trajectory kernel for n <= 10 qubits (configs C1-C3: the reference's real workload is n = 7).
//
// One CTA integrates one whole trajectory: thread x owns amplitude x, its RK stages live in registers,
// the current stage vector lives in shared memory (double-buffered) so neighbours x^mask can be read,
// the adaptive controller runs redundantly in every thread, and per output time eight reduced doubles
// leave the SM.  The state never touches HBM.  Replaces the zvode step loop + Python RHS callback +
// QuTiP expectation evaluation behind qt.sesolve (dipolar_ensemble_with_rare.py:653-669).
//
// The loop is instruction-issue / FP64-latency bound (one warp per SM sub-partition at n = 7), so the
// row product is specialised: for n <= 8 every thread keeps the coefficients of ITS row in registers
// (selected once by its own bit pattern) and the neighbour offsets are compile-time constants, which
// leaves one LDS.128 + 2-4 DFMA per term.  Larger n falls back to term lists in shared memory.
#include "qsb_small.cuh"

#include <cstdlib>
#include <mutex>

namespace qsb {
namespace {

struct SmallShared {
  SmallDesc d;
  double red[7 * 32 + 8 + 64];  // block_reduce<7> scratch, then two rows for weighted_rms
};

__device__ __forceinline__ double row_diagonal(const SmallDesc& d, uint32_t x) {
  double diag = 0.0;
  int pi = 0;
  for (int b = 0; b < d.n; ++b) {
    double sb = (x >> b & 1) ? -1.0 : 1.0;
    diag = fma(sb, d.hz_half[b], diag);
    for (int c = b + 1; c < d.n; ++c, ++pi) diag = fma((x >> c & 1) ? -sb : sb, d.jzz_quarter[pi], diag);
  }
  return diag;
}

// ---- row product, generic: term lists in shared memory ------------------------------------------------
struct RhsGeneric {
  const SmallDesc* d;
  double diag;
  uint32_t x;
  __device__ __forceinline__ void setup(const SmallDesc& desc, uint32_t xs) {
    d = &desc;
    x = xs;
    diag = row_diagonal(desc, xs);
  }
  __device__ __forceinline__ cplx operator()(const cplx* __restrict__ v) const {
    cplx acc0 = diag * v[x], acc1 = mk(0.0, 0.0);
    for (int f = 0; f < d->n_flip; ++f) {
      const SmallFlip& e = d->flip[f];
      cplx u = v[x ^ e.mask];
      double im = (x & e.mask) ? e.hyh : -e.hyh;
      caxpy(acc1, mk(e.hxh, im), u);
    }
#pragma unroll 2
    for (int p = 0; p < d->n_pair; p += 2) {
      const SmallPair& e0 = d->pair[p];
      cplx u0 = v[x ^ e0.mask];
      axpy(acc0, (__popc(x & e0.mask) & 1) ? e0.c_diff : e0.c_same, u0);
      if (p + 1 < d->n_pair) {
        const SmallPair& e1 = d->pair[p + 1];
        cplx u1 = v[x ^ e1.mask];
        axpy(acc1, (__popc(x & e1.mask) & 1) ? e1.c_diff : e1.c_same, u1);
      }
    }
    return mul_mi(acc0 + acc1);  // d(psi)/dt = -i H psi
  }
};

// ---- row product, register-resident coefficients for exactly NB qubits ------------------------------------
template <int NB>
struct RhsReg {
  static constexpr int NP = NB * (NB - 1) / 2;
  double diag;
  uint32_t x;
  double fre[NB], fim[NB];
  double pc[NP > 0 ? NP : 1];
  __device__ __forceinline__ void setup(const SmallDesc& d, uint32_t xs) {
    x = xs;
    diag = row_diagonal(d, xs);
    int pi = 0;
#pragma unroll
    for (int b = 0; b < NB; ++b) {
      fre[b] = d.hx_half[b];
      fim[b] = (xs >> b & 1) ? d.hy_half[b] : -d.hy_half[b];
#pragma unroll
      for (int c = b + 1; c < NB; ++c, ++pi)
        pc[pi] = ((xs >> b ^ xs >> c) & 1) ? d.pair_diff[pi] : d.pair_same[pi];
    }
  }
  __device__ __forceinline__ cplx operator()(const cplx* __restrict__ v) const {
    cplx acc[4] = {diag * v[x], mk(0.0, 0.0), mk(0.0, 0.0), mk(0.0, 0.0)};
    int pi = 0;
#pragma unroll
    for (int b = 0; b < NB; ++b) {
      caxpy(acc[(b + 1) & 3], mk(fre[b], fim[b]), v[x ^ (1u << b)]);
#pragma unroll
      for (int c = b + 1; c < NB; ++c, ++pi) axpy(acc[pi & 3], pc[pi], v[x ^ ((1u << b) | (1u << c))]);
    }
    return mul_mi((acc[0] + acc[1]) + (acc[2] + acc[3]));
  }
};

// Eight observable sums of the vector currently in `v` (must be visible to the whole CTA).
__device__ void small_observe(const cplx* __restrict__ v, uint32_t x, bool live, const SmallDesc& d,
                              double* red, double* __restrict__ out) {
  double s[7] = {0, 0, 0, 0, 0, 0, 0};  // nrm2, zsea, zrare, csea.re, csea.im, crare.re, crare.im
  if (live) {
    cplx me = v[x];
    double p = norm2(me);
    s[0] = p;
    int msea = __popc(d.sea_mask) - 2 * __popc(x & d.sea_mask);
    s[1] = p * (double)msea;
    s[2] = (x & d.rare_mask) ? -p : p;
    for (int b = 0; b < d.n; ++b) {
      uint32_t m = 1u << b;
      if (x & m) continue;
      bool is_sea = d.sea_mask & m, is_rare = d.rare_mask & m;
      if (!is_sea && !is_rare) continue;
      cplx u = v[x | m];  // conj(me) * u
      double re = fma(me.x, u.x, me.y * u.y), im = fma(me.x, u.y, -me.y * u.x);
      if (is_sea) { s[3] += re; s[4] += im; }
      if (is_rare) { s[5] += re; s[6] += im; }
    }
  }
  block_reduce<7>(s, red);
  if (threadIdx.x == 0) {
    double inv = 1.0 / s[0];
    out[0] = s[3] * inv;        // Ix_sea
    out[1] = s[4] * inv;        // Iy_sea
    out[2] = 0.5 * s[1] * inv;  // Iz_sea
    out[3] = 0.5 * s[2] * inv;  // Iz_R
    out[4] = s[5] * inv;        // Ix_R
    out[5] = s[6] * inv;        // Iy_R
    out[6] = 1.0;               // norm of the output-normalised ket
    out[7] = sqrt(s[0]);        // raw integrator norm
  }
}

// Weighted RMS with ONE barrier: shuffle tree per warp, per-warp partials to shared memory (two scratch rows
// used alternately, so the next call cannot overwrite values still being read), then every thread adds the
// partials in the same order.  `local` must already be zero on lanes that do not own an amplitude.
__device__ __forceinline__ double weighted_rms(double local, double* red, int dim, int& flip) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) local += __shfl_xor_sync(0xffffffffu, local, off);
  if (nwarp > 1) {
    double* row = red + 32 * flip;
    flip ^= 1;
    if (lane == 0) row[warp] = local;
    __syncthreads();
    local = 0.0;
    for (int w = 0; w < nwarp; ++w) local += row[w];
  }
  return sqrt(local / (double)dim);
}

}  // namespace
}  // namespace qsb

#include "qsb_small_kernel.inl"
