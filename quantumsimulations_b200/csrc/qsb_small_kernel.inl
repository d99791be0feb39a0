// Trajectory kernel body + launchers; included from qsb_small.cu.
namespace qsb {
namespace {

template <class RHS, int MAXT>
__global__ void __launch_bounds__(MAXT, 1)
small_evolve_kernel(const SmallDesc* __restrict__ descs, const double* __restrict__ t_out_all, int64_t n_out,
                    double* __restrict__ obs_all, SmallStats* __restrict__ stats) {
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
  const uint32_t xs = live ? x : 0;  // idle lanes compute a harmless copy of row 0 and never store it
  double* rms_red = sh.red + 7 * 32 + 8;
  int flip = 0;
  const double* t_out = t_out_all + (d.t_per_traj ? (size_t)traj * n_out : 0);
  double* obs = obs_all + (size_t)traj * n_out * QSB_NOBS;

  RHS rhs;
  rhs.setup(d, xs);

  const DP5 T = dp5_tableau();
  cplx y = mk(live && x == d.psi0 ? 1.0 : 0.0, 0.0);
  cplx k[7];
  double t = t_out[0];
  int64_t n_acc = 0, n_rej = 0, n_rhs = 0;
  int status = QSB_OK;
  const bool taylor = d.method == QSB_METHOD_TAYLOR;
  const int order = taylor ? d.order : 5;
  const double inv_q = taylor ? 1.0 / (double)order : 0.2;
  const double atol = d.atol, rtol = d.rtol, max_step = d.max_step;
  const int64_t nsteps = d.nsteps;

  // output 0 and k1 = f(y0)
  buf0[x] = y;
  __syncthreads();
  small_observe(buf0, x, live, d, sh.red, obs);
  k[0] = rhs(buf0);
  ++n_rhs;
  double h;
  {
    double w = 1.0 / (atol + rtol * sqrt(norm2(y)));
    double d0 = weighted_rms(live ? norm2(y) * w * w : 0.0, rms_red, dim, flip);
    double d1 = weighted_rms(live ? norm2(k[0]) * w * w : 0.0, rms_red, dim, flip);
    h = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
    if (d.first_step > 0.0) h = d.first_step;
    if (max_step > 0.0 && h > max_step) h = max_step;
  }
  bool after_reject = false;
  int cur = 1;  // buffer the NEXT stage vector goes to
  __syncthreads();

  for (int64_t io = 1; io < n_out && status == QSB_OK; ++io) {
    const double t_target = t_out[io];
    int64_t in_interval = 0;
    while (t < t_target) {
      if (nsteps > 0 && in_interval >= nsteps) { status = QSB_ERR_NSTEPS; break; }
      double hmax = max_step > 0.0 ? fmin(h, max_step) : h;
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
          cplx a0 = mk(0, 0), a1 = mk(0, 0);  // two chains for the stage combination
#pragma unroll
          for (int j = 0; j <= s; ++j) axpy((j & 1) ? a1 : a0, T.a[s][j], k[j]);
          ynew = y;
          axpy(ynew, hu, a0 + a1);
          cplx* b = cur ? buf1 : buf0;
          b[x] = ynew;
          __syncthreads();
          k[s + 1] = rhs(b);
          cur ^= 1;
        }
        n_rhs += 6;
        cplx e0 = mk(0, 0), e1 = mk(0, 0);
#pragma unroll
        for (int j = 0; j < 7; ++j) axpy((j & 1) ? e1 : e0, T.e[j], k[j]);
        double sc = atol + rtol * sqrt(fmax(norm2(y), norm2(ynew)));
        err_local = hu * hu * norm2(e0 + e1) / (sc * sc);
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
            f = rhs(b);
            cur ^= 1;
          }
          term = (hu / (double)(m + 1)) * f;
          ynew = ynew + term;
        }
        n_rhs += order - 1;
        double sc = atol + rtol * sqrt(fmax(norm2(y), norm2(ynew)));
        err_local = norm2(term) / (sc * sc);
      }
      double err = weighted_rms(live ? err_local : 0.0, rms_red, dim, flip);
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
          k[0] = rhs(b);
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

__global__ void small_apply_kernel(const SmallDesc* __restrict__ desc, const cplx* __restrict__ in,
                                   cplx* __restrict__ out, int use_reg) {
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
  const uint32_t xs = live ? x : 0;
  buf[x] = live ? in[x] : mk(0, 0);
  __syncthreads();
  cplx k;  // -i H psi through whichever row product the trajectory kernel would use for this n
  if (use_reg && d.n == 7) { RhsReg<7> r; r.setup(d, xs); k = r(buf); }
  else if (use_reg && d.n == 8) { RhsReg<8> r; r.setup(d, xs); k = r(buf); }
  else if (use_reg && d.n == 5) { RhsReg<5> r; r.setup(d, xs); k = r(buf); }
  else if (use_reg && d.n == 3) { RhsReg<3> r; r.setup(d, xs); k = r(buf); }
  else { RhsGeneric r; r.setup(d, xs); k = r(buf); }
  if (live) out[x] = mk(-k.y, k.x);  // i * k = H psi
}

template <class RHS>
static void launch_evolve_t(const SmallDesc* d_desc, int n_traj, int threads, const double* d_t_out, int64_t n_out,
                            double* d_obs, SmallStats* d_stats, cudaStream_t stream) {
  size_t smem = ((sizeof(SmallShared) + 15) & ~size_t(15)) + 2 * sizeof(cplx) * threads;
  if (threads > 512)  // 1024-thread CTAs are capped at 64 registers per thread (n = 10 only)
    small_evolve_kernel<RHS, 1024><<<n_traj, threads, smem, stream>>>(d_desc, d_t_out, n_out, d_obs, d_stats);
  else if (threads > 256)
    small_evolve_kernel<RHS, 512><<<n_traj, threads, smem, stream>>>(d_desc, d_t_out, n_out, d_obs, d_stats);
  else
    small_evolve_kernel<RHS, 256><<<n_traj, threads, smem, stream>>>(d_desc, d_t_out, n_out, d_obs, d_stats);
}

}  // namespace

void launch_small_apply(const SmallDesc* d_desc, int n, const cplx* d_in, cplx* d_out, cudaStream_t stream) {
  int threads = 1 << n;
  if (threads < 32) threads = 32;
  size_t smem = ((sizeof(SmallShared) + 15) & ~size_t(15)) + sizeof(cplx) * threads;
  const int use_reg = getenv("QSB200_SMALL_GENERIC") ? 0 : 1;
  small_apply_kernel<<<1, threads, smem, stream>>>(d_desc, d_in, d_out, use_reg);
}

// `uniform_n`: every trajectory of the batch has exactly max_n qubits (required by the register-resident rows).
void launch_small_evolve(const SmallDesc* d_desc, int n_traj, int max_n, int uniform_n, const double* d_t_out,
                         int64_t n_out, double* d_obs, SmallStats* d_stats, cudaStream_t stream) {
  int threads = 1 << max_n;
  if (threads < 32) threads = 32;
  const bool reg = uniform_n && !getenv("QSB200_SMALL_GENERIC");
#define QSB_REG_CASE(NB)                                                                                     \
  case NB: launch_evolve_t<RhsReg<NB>>(d_desc, n_traj, threads, d_t_out, n_out, d_obs, d_stats, stream); return;
  if (reg) {
    switch (max_n) {
      QSB_REG_CASE(2) QSB_REG_CASE(3) QSB_REG_CASE(4) QSB_REG_CASE(5) QSB_REG_CASE(6) QSB_REG_CASE(7) QSB_REG_CASE(8)
      default: break;
    }
  }
#undef QSB_REG_CASE
  launch_evolve_t<RhsGeneric>(d_desc, n_traj, threads, d_t_out, n_out, d_obs, d_stats, stream);
}

}  // namespace qsb
