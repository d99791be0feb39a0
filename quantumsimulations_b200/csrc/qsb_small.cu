// On-chip trajectory kernel for n <= 10 qubits (configs C1-C3: the reference's real workload is n = 7).
//
// One CTA integrates one whole trajectory: thread x owns amplitude x, its RK stages live in registers,
// the current stage vector lives in shared memory (double-buffered) so neighbours x^mask can be read,
// the adaptive controller runs redundantly in every thread, and per output time eight reduced doubles
// leave the SM.  The state never touches HBM.  Replaces the zvode step loop + Python RHS callback +
// QuTiP expectation evaluation behind qt.sesolve (dipolar_ensemble_with_rare.py:653-669).
#include "qsb_small.cuh"

namespace qsb {
namespace {

struct SmallShared {
  SmallDesc d;
  double red[8 * 32 + 8];
};

__device__ __forceinline__ cplx small_rhs(const cplx* __restrict__ v, uint32_t x, double diag,
                                          const SmallDesc& d) {
  cplx acc = diag * v[x];
  for (int f = 0; f < d.n_flip; ++f) {
    const SmallFlip& e = d.flip[f];
    cplx u = v[x ^ e.mask];
    double im = (x & e.mask) ? e.hyh : -e.hyh;
    caxpy(acc, mk(e.hxh, im), u);
  }
  for (int p = 0; p < d.n_pair; ++p) {
    const SmallPair& e = d.pair[p];
    cplx u = v[x ^ e.mask];
    double c = (__popc(x & e.mask) & 1) ? e.c_diff : e.c_same;
    axpy(acc, c, u);
  }
  return mul_mi(acc);  // d(psi)/dt = -i H psi
}

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

__device__ __forceinline__ double weighted_rms(double local, double* red, int dim) {
  double s[1] = {local};
  block_reduce<1>(s, red);
  return sqrt(s[0] / (double)dim);
}

__global__ void small_evolve_kernel(const SmallDesc* __restrict__ descs, const double* __restrict__ t_out_all,
                                    int64_t n_out, double* __restrict__ obs_all, SmallStats* __restrict__ stats) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SmallShared& sh = *reinterpret_cast<SmallShared*>(smem_raw);
  cplx* buf0 = reinterpret_cast<cplx*>(smem_raw + ((sizeof(SmallShared) + 15) & ~size_t(15)));
  const int traj = blockIdx.x;
  {  // descriptor -> shared
    const int* src = reinterpret_cast<const int*>(descs + traj);
    int* dst = reinterpret_cast<int*>(&sh.d);
    for (int i = threadIdx.x; i < (int)(sizeof(SmallDesc) / 4); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const SmallDesc& d = sh.d;
  const int dim = 1 << d.n;
  cplx* buf1 = buf0 + blockDim.x;
  const uint32_t x = threadIdx.x;
  const bool live = x < (uint32_t)dim;
  const double* t_out = t_out_all + (d.t_per_traj ? (size_t)traj * n_out : 0);
  double* obs = obs_all + (size_t)traj * n_out * QSB_NOBS;

  double diag = 0.0;
  if (live) {
    int pi = 0;
    for (int b = 0; b < d.n; ++b) {
      double sb = (x >> b & 1) ? -1.0 : 1.0;
      diag = fma(sb, d.hz_half[b], diag);
      for (int c = b + 1; c < d.n; ++c, ++pi) {
        double sc = (x >> c & 1) ? -sb : sb;
        diag = fma(sc, d.jzz_quarter[pi], diag);
      }
    }
  }

  const DP5 T = dp5_tableau();
  cplx y = mk(live && x == d.psi0 ? 1.0 : 0.0, 0.0);
  cplx k[7];
  double t = t_out[0];
  int64_t n_acc = 0, n_rej = 0, n_rhs = 0;
  int status = QSB_OK;
  const bool taylor = d.method == QSB_METHOD_TAYLOR;
  const int order = taylor ? d.order : 5;
  const double inv_q = taylor ? 1.0 / (double)order : 0.2;

  // output 0 and k1 = f(y0)
  buf0[x] = y;
  __syncthreads();
  small_observe(buf0, x, live, d, sh.red, obs);
  k[0] = live ? small_rhs(buf0, x, diag, d) : mk(0, 0);
  ++n_rhs;
  double h;
  {
    double w = 1.0 / (d.atol + d.rtol * sqrt(norm2(y)));
    double d0 = weighted_rms(live ? norm2(y) * w * w : 0.0, sh.red, dim);
    double d1 = weighted_rms(live ? norm2(k[0]) * w * w : 0.0, sh.red, dim);
    h = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
    if (d.first_step > 0.0) h = d.first_step;
    if (d.max_step > 0.0 && h > d.max_step) h = d.max_step;
  }
  bool after_reject = false;
  int cur = 1;  // buffer the NEXT stage vector is written to (buf0 holds y0, already consumed after sync below)
  __syncthreads();

  for (int64_t io = 1; io < n_out && status == QSB_OK; ++io) {
    const double t_target = t_out[io];
    int64_t in_interval = 0;
    while (t < t_target) {
      if (d.nsteps > 0 && in_interval >= d.nsteps) { status = QSB_ERR_NSTEPS; break; }
      double hmax = d.max_step > 0.0 ? fmin(h, d.max_step) : h;
      double remaining = t_target - t;
      double hu = hmax;
      bool hit = false;
      if (hu >= remaining * (1.0 - 1e-12)) { hu = remaining; hit = true; }
      else if (2.0 * hu > remaining) hu = 0.5 * remaining;
      if (hu <= 16.0 * 2.220446049250313e-16 * fabs(t)) { status = QSB_ERR_STEP_UNDERFLOW; break; }

      cplx ynew;
      double err_local;
      if (!taylor) {
#pragma unroll
        for (int s = 0; s < 6; ++s) {
          cplx acc = mk(0, 0);
#pragma unroll
          for (int j = 0; j <= s; ++j) axpy(acc, T.a[s][j], k[j]);
          ynew = y;
          axpy(ynew, hu, acc);
          cplx* b = cur ? buf1 : buf0;
          b[x] = ynew;
          __syncthreads();
          k[s + 1] = live ? small_rhs(b, x, diag, d) : mk(0, 0);
          cur ^= 1;
        }
        n_rhs += 6;
        cplx e = mk(0, 0);
#pragma unroll
        for (int j = 0; j < 7; ++j) axpy(e, T.e[j], k[j]);
        double sc = d.atol + d.rtol * sqrt(fmax(norm2(y), norm2(ynew)));
        err_local = hu * hu * norm2(e) / (sc * sc);
      } else {
        cplx term = y;
        ynew = y;
        for (int m = 0; m < order; ++m) {
          cplx f;
          if (m == 0) {
            f = k[0];  // f(y) carried over from the previous accepted step or the start
          } else {
            cplx* b = cur ? buf1 : buf0;
            b[x] = term;
            __syncthreads();
            f = live ? small_rhs(b, x, diag, d) : mk(0, 0);
            cur ^= 1;
          }
          term = (hu / (double)(m + 1)) * f;
          ynew = ynew + term;
        }
        n_rhs += order - 1;
        double sc = d.atol + d.rtol * sqrt(fmax(norm2(y), norm2(ynew)));
        err_local = norm2(term) / (sc * sc);
      }
      double err = weighted_rms(live ? err_local : 0.0, sh.red, dim);
      ++in_interval;
      if (err <= 1.0) {
        y = ynew;
        t = hit ? t_target : t + hu;
        ++n_acc;
        if (!taylor) {
          k[0] = k[6];
        } else {  // Taylor has no FSAL: evaluate f(y_new) now
          cplx* b = cur ? buf1 : buf0;
          b[x] = y;
          __syncthreads();
          k[0] = live ? small_rhs(b, x, diag, d) : mk(0, 0);
          cur ^= 1;
          ++n_rhs;
        }
        double hn = hu * Controller::factor(err, inv_q, after_reject);
        h = (hu < hmax) ? fmax(hn, h) : hn;  // a step clipped to land on t_target must not shrink h
        after_reject = false;
      } else {
        ++n_rej;
        h = hu * Controller::factor(err, inv_q, true);
        after_reject = true;
      }
    }
    if (status != QSB_OK) break;
    cplx* b = cur ? buf1 : buf0;
    b[x] = y;
    __syncthreads();
    small_observe(b, x, live, d, sh.red, obs + io * QSB_NOBS);
    cur ^= 1;
  }
  if (threadIdx.x == 0) {
    SmallStats st;
    st.n_accepted = n_acc; st.n_rejected = n_rej; st.n_rhs = n_rhs; st.last_h = h; st.status = status; st.pad = 0;
    stats[traj] = st;
  }
}

__global__ void small_apply_kernel(const SmallDesc* __restrict__ desc, const cplx* __restrict__ in, cplx* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SmallShared& sh = *reinterpret_cast<SmallShared*>(smem_raw);
  cplx* buf = reinterpret_cast<cplx*>(smem_raw + ((sizeof(SmallShared) + 15) & ~size_t(15)));
  {
    const int* src = reinterpret_cast<const int*>(desc);
    int* dst = reinterpret_cast<int*>(&sh.d);
    for (int i = threadIdx.x; i < (int)(sizeof(SmallDesc) / 4); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const SmallDesc& d = sh.d;
  const uint32_t x = threadIdx.x;
  const bool live = x < (1u << d.n);
  double diag = 0.0;
  if (live) {
    int pi = 0;
    for (int b = 0; b < d.n; ++b) {
      double sb = (x >> b & 1) ? -1.0 : 1.0;
      diag = fma(sb, d.hz_half[b], diag);
      for (int c = b + 1; c < d.n; ++c, ++pi) diag = fma((x >> c & 1) ? -sb : sb, d.jzz_quarter[pi], diag);
    }
  }
  buf[x] = live ? in[x] : mk(0, 0);
  __syncthreads();
  if (live) {
    cplx k = small_rhs(buf, x, diag, d);  // -i H psi
    out[x] = mk(-k.y, k.x);               // i * k = H psi
  }
}

}  // namespace

void launch_small_apply(const SmallDesc* d_desc, int n, const cplx* d_in, cplx* d_out, cudaStream_t stream) {
  int threads = 1 << n;
  if (threads < 32) threads = 32;
  size_t smem = ((sizeof(SmallShared) + 15) & ~size_t(15)) + sizeof(cplx) * threads;
  small_apply_kernel<<<1, threads, smem, stream>>>(d_desc, d_in, d_out);
}

void launch_small_evolve(const SmallDesc* d_desc, int n_traj, int max_n, const double* d_t_out, int64_t n_out,
                         double* d_obs, SmallStats* d_stats, cudaStream_t stream) {
  int threads = 1 << max_n;
  if (threads < 32) threads = 32;
  size_t smem = ((sizeof(SmallShared) + 15) & ~size_t(15)) + 2 * sizeof(cplx) * threads;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(small_evolve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
    attr_set = true;
  }
  small_evolve_kernel<<<n_traj, threads, smem, stream>>>(d_desc, d_t_out, n_out, d_obs, d_stats);
}

}  // namespace qsb
