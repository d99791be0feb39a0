// Shared device/host helpers for libqsb200: complex128 arithmetic on double2, the Dormand-Prince
// tableau, deterministic block reductions, error plumbing.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>

#include "../../include/qsb200.h"

namespace qsb {

// ---- error plumbing -------------------------------------------------------------------------------
void set_error(const std::string& msg);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);
#define QSB_CUDA(call)                                                        \
  do {                                                                        \
    cudaError_t e__ = (call);                                                 \
    if (e__ != cudaSuccess) return qsb::cuda_fail(e__, #call, __FILE__, __LINE__); \
  } while (0)

// ---- complex128 as double2 -------------------------------------------------------------------------
typedef double2 cplx;
__host__ __device__ __forceinline__ cplx mk(double r, double i) { return make_double2(r, i); }
__host__ __device__ __forceinline__ cplx operator+(cplx a, cplx b) { return mk(a.x + b.x, a.y + b.y); }
__host__ __device__ __forceinline__ cplx operator-(cplx a, cplx b) { return mk(a.x - b.x, a.y - b.y); }
__host__ __device__ __forceinline__ cplx operator*(double s, cplx a) { return mk(s * a.x, s * a.y); }
__host__ __device__ __forceinline__ cplx cmul(cplx a, cplx b) {
  return mk(fma(a.x, b.x, -a.y * b.y), fma(a.x, b.y, a.y * b.x));
}
// acc += s * a   (real s)
__host__ __device__ __forceinline__ void axpy(cplx& acc, double s, cplx a) {
  acc.x = fma(s, a.x, acc.x);
  acc.y = fma(s, a.y, acc.y);
}
// acc += c * a   (complex c)
__host__ __device__ __forceinline__ void caxpy(cplx& acc, cplx c, cplx a) {
  acc.x = fma(c.x, a.x, fma(-c.y, a.y, acc.x));
  acc.y = fma(c.x, a.y, fma(c.y, a.x, acc.y));
}
__host__ __device__ __forceinline__ double norm2(cplx a) { return fma(a.x, a.x, a.y * a.y); }
// -i * a
__host__ __device__ __forceinline__ cplx mul_mi(cplx a) { return mk(a.y, -a.x); }
// a * (-i)^p  and  a * (+i)^p,  p mod 4
__host__ __device__ __forceinline__ cplx mul_mi_pow(cplx a, int p) {
  p &= 3;
  return p == 0 ? a : p == 1 ? mk(a.y, -a.x) : p == 2 ? mk(-a.x, -a.y) : mk(-a.y, a.x);
}
__host__ __device__ __forceinline__ cplx mul_pi_pow(cplx a, int p) { return mul_mi_pow(a, 4 - (p & 3)); }

// ---- Dormand-Prince 5(4) ---------------------------------------------------------------------------
// a[s][j]: stage s+2 input = y + h * sum_j a[s][j] k_{j+1};  row 5 is the 5th-order solution (FSAL).
struct DP5 {
  static constexpr int S = 7;
  double a[6][6];
  double e[7];  // error weights: err = h * sum e_j k_j
};
__host__ __device__ inline DP5 dp5_tableau() {
  DP5 t = {};
  t.a[0][0] = 1.0 / 5.0;
  t.a[1][0] = 3.0 / 40.0;        t.a[1][1] = 9.0 / 40.0;
  t.a[2][0] = 44.0 / 45.0;       t.a[2][1] = -56.0 / 15.0;      t.a[2][2] = 32.0 / 9.0;
  t.a[3][0] = 19372.0 / 6561.0;  t.a[3][1] = -25360.0 / 2187.0; t.a[3][2] = 64448.0 / 6561.0;  t.a[3][3] = -212.0 / 729.0;
  t.a[4][0] = 9017.0 / 3168.0;   t.a[4][1] = -355.0 / 33.0;     t.a[4][2] = 46732.0 / 5247.0;  t.a[4][3] = 49.0 / 176.0;   t.a[4][4] = -5103.0 / 18656.0;
  t.a[5][0] = 35.0 / 384.0;      t.a[5][1] = 0.0;               t.a[5][2] = 500.0 / 1113.0;    t.a[5][3] = 125.0 / 192.0;  t.a[5][4] = -2187.0 / 6784.0;  t.a[5][5] = 11.0 / 84.0;
  t.e[0] = 71.0 / 57600.0; t.e[1] = 0.0; t.e[2] = -71.0 / 16695.0; t.e[3] = 71.0 / 1920.0;
  t.e[4] = -17253.0 / 339200.0; t.e[5] = 22.0 / 525.0; t.e[6] = -1.0 / 40.0;
  return t;
}

// ---- step-size controller (shared by every path so that CPU port and GPU make the same decisions) --
// err = weighted RMS error (accept when <= 1); q = order of the error estimate + 1.
struct Controller {
  static constexpr double safety = 0.9, fac_min = 0.2, fac_max = 5.0;
  __host__ __device__ static double factor(double err, double inv_q, bool after_reject) {
    // single precision is plenty for a step-size factor and ~30x cheaper than double pow on the device
#ifdef __CUDA_ARCH__
    double f = (err > 0.0) ? safety * (double)__powf((float)err, -(float)inv_q) : fac_max;
#else
    double f = (err > 0.0) ? safety * (double)powf((float)err, -(float)inv_q) : fac_max;
#endif
    f = f < fac_min ? fac_min : f;
    double cap = after_reject ? 1.0 : fac_max;
    return f > cap ? cap : f;
  }
};

// ---- deterministic block reduction of K doubles ----------------------------------------------------
// Fixed shuffle tree inside each warp, then warp 0 sums the per-warp partials in warp order.
// All threads must call; result valid in every thread (broadcast through shared memory).
template <int K>
__device__ __forceinline__ void block_reduce(double (&v)[K], double* scratch /* [K*32 + K] */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int k = 0; k < K; ++k) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], off);
  }
  if (nwarp == 1) return;  // single warp: every lane already holds the total
  __syncthreads();
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < K; ++k) scratch[k * 32 + warp] = v[k];
  }
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int k = 0; k < K; ++k) {
      double x = lane < nwarp ? scratch[k * 32 + lane] : 0.0;
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
      if (lane == 0) scratch[K * 32 + k] = x;
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < K; ++k) v[k] = scratch[K * 32 + k];
}

}  // namespace qsb
