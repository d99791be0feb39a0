// Weighted Pauli-string expectation values of device-resident state vectors (qsb_propagate_expect, qsb_session_expect).
//
// The reference's spin-3/2 rare spin (dipolar_ensemble_with_rare.py:22-25, :499-501, :576-578) occupies two index bits
// here, and its J_x / J_y are sums of one- and two-bit Pauli strings (sqrt(3) I^x_b + 2 (I^x_a I^x_b + I^y_a I^y_b) and
// the mixed-axis counterpart), which the mask-based observable sums fused into the integrator kernels do not cover.
// These two kernels evaluate sum_t w_t <psi|P_t|psi> / <psi|psi> per output column for arbitrary strings P_t on
// states that already sit in device memory (the state-storing matvec chain of the propagator path, or a session's
// vector); they are not on the spin-1/2 hot path.
#include "qsb_expect.cuh"

namespace qsb {
namespace {

struct ColumnRanges {
  int begin[7];  // terms [begin[c], begin[c + 1]) feed output column c (the host sorts the terms by column)
};

// grid.x = n_states * nb: CTA (s, b) reduces amplitudes b * 256 + tid, + nb * 256, ... of state s.
__global__ void __launch_bounds__(EXPECT_THREADS)
expect_partial_kernel(const cplx* __restrict__ states, uint64_t dim, int nb, const ExpectTerm* __restrict__ terms,
                      int n_terms, ColumnRanges cr, double* __restrict__ partials) {
  __shared__ ExpectTerm st[EXPECT_MAX_TERMS];
  __shared__ double red[7 * 32 + 7];
  for (int i = threadIdx.x; i < n_terms; i += blockDim.x) st[i] = terms[i];
  __syncthreads();
  const uint64_t s = blockIdx.x / (unsigned)nb;
  const unsigned b = blockIdx.x - (unsigned)(s * (unsigned)nb);
  const cplx* __restrict__ psi = states + s * dim;
  double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  for (uint64_t x = (uint64_t)b * EXPECT_THREADS + threadIdx.x; x < dim; x += (uint64_t)nb * EXPECT_THREADS) {
    const cplx me = psi[x];
    acc[6] += norm2(me);
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      double sum = 0.0;
      for (int t = cr.begin[c]; t < cr.begin[c + 1]; ++t) {
        const ExpectTerm e = st[t];
        const cplx u = psi[x ^ e.x_mask];
        // conj(u) * me, times i^ny, real part
        const double re = fma(u.x, me.x, u.y * me.y), im = fma(u.x, me.y, -u.y * me.x);
        double v = e.ny == 0 ? re : e.ny == 1 ? -im : e.ny == 2 ? -re : im;
        if (__popcll(x & e.z_mask) & 1) v = -v;
        sum = fma(e.weight, v, sum);
      }
      acc[c] += sum;
    }
  }
  block_reduce<7>(acc, red);
  if (threadIdx.x == 0) {
    double* dst = partials + (size_t)blockIdx.x * 8;
#pragma unroll
    for (int k = 0; k < 7; ++k) dst[k] = acc[k];
    dst[7] = 0.0;
  }
}

// One thread per state: the nb partial rows are added in block order, then normalised like every other path reports
// its observables (on the output-normalised ket, QuTiP's normalize_output; column 7 keeps the raw norm).
__global__ void expect_finish_kernel(const double* __restrict__ partials, int nb, int64_t n_states, double* __restrict__ out,
                                     int64_t stride_state, int64_t stride_col) {
  const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_states) return;
  double acc[7] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  for (int b = 0; b < nb; ++b) {
    const double* src = partials + ((size_t)s * nb + b) * 8;
#pragma unroll
    for (int k = 0; k < 7; ++k) acc[k] += src[k];
  }
  const double inv = 1.0 / acc[6];
  double* dst = out + s * stride_state;
#pragma unroll
  for (int k = 0; k < 6; ++k) dst[k * stride_col] = acc[k] * inv;
  dst[6 * stride_col] = 1.0;
  dst[7 * stride_col] = sqrt(acc[6]);
}

}  // namespace

int expect_blocks(uint64_t dim) {
  uint64_t nb = dim / (4 * EXPECT_THREADS);
  if (nb < 1) nb = 1;
  if (nb > EXPECT_MAX_BLOCKS) nb = EXPECT_MAX_BLOCKS;
  return (int)nb;
}

void launch_expect(const cplx* states, int64_t n_states, uint64_t dim, const ExpectTerm* terms, int n_terms,
                   const int* column_begin, double* partials, double* out, int64_t stride_state, int64_t stride_col,
                   cudaStream_t stream) {
  const int nb = expect_blocks(dim);
  ColumnRanges cr;
  for (int c = 0; c < 7; ++c) cr.begin[c] = column_begin[c];
  expect_partial_kernel<<<(unsigned)(n_states * nb), EXPECT_THREADS, 0, stream>>>(states, dim, nb, terms, n_terms, cr, partials);
  expect_finish_kernel<<<(unsigned)((n_states + 127) / 128), 128, 0, stream>>>(partials, nb, n_states, out, stride_state,
                                                                              stride_col);
}

}  // namespace qsb
