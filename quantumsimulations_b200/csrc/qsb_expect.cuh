// Weighted Pauli-string expectation values of state vectors resident in device memory: the observables of the
// spin-3/2 rare spin (two index bits), which the six mask-based sums fused into the integrator kernels cannot express.
#pragma once
#include "qsb_common.cuh"

namespace qsb {

constexpr int EXPECT_MAX_TERMS = 512;
constexpr int EXPECT_THREADS = 256;
constexpr int EXPECT_MAX_BLOCKS = 1024;  // CTAs per state (grid.x)

// Device copy of one qsb_pauli_term: P = i^popc(x_mask & z_mask) X^x_mask Z^z_mask (Y = i X Z), so
// <psi|P|psi> = sum_x conj(psi[x ^ x_mask]) i^ny (-1)^popc(x & z_mask) psi[x].  `ny` = popc(x_mask & z_mask) & 3.
struct ExpectTerm {
  uint64_t x_mask, z_mask;
  double weight;
  int32_t column, ny;
};

// CTAs per state for vectors of `dim` amplitudes.
int expect_blocks(uint64_t dim);

// states: [n_states][dim] complex128 (device), terms: [n_terms] (device, sorted by column), column_begin: [7] (host;
// terms [column_begin[c], column_begin[c + 1]) feed column c), partials: [n_states][expect_blocks(dim)][8] (device
// scratch); n_states * expect_blocks(dim) must fit a grid dimension.  Writes, per state s, the QSB_NOBS columns of the C ABI: out[s * stride_state + k * stride_col],
// k = 0..5 the weighted sums of column k divided by <psi|psi>, 6 = 1.0, 7 = sqrt(<psi|psi>).  Two launches, fixed
// summation order (bitwise reproducible), no atomics.
void launch_expect(const cplx* states, int64_t n_states, uint64_t dim, const ExpectTerm* terms, int n_terms,
                   const int* column_begin, double* partials, double* out, int64_t stride_state, int64_t stride_col, cudaStream_t stream);

}  // namespace qsb
