// Trajectory kernel body + launchers; included from qsb_small.cu.
namespace qsb {
namespace {

// COL = false: CTA b integrates trajectory b of the batch and reports its observables.
// COL = true (propagator build): CTA b integrates column b % dim of trajectory b / dim's propagator, i.e. it starts
// from basis state b % dim, reports no observables and leaves its final state in columns[b][0..dim).
template <class RHS, int MAXT, bool COL>
__global__ void __launch_bounds__(MAXT, 1)
small_evolve_kernel(const SmallDesc* __restrict__ descs, const double* __restrict__ t_out_all, int64_t n_out,
                    double* __restrict__ obs_all, SmallStats* __restrict__ stats, cplx* __restrict__ columns,
                    int col_dim) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SmallShared& sh = *reinterpret_cast<SmallShared*>(smem_raw);
  cplx* buf0 = reinterpret_cast<cplx*>(smem_raw + ((sizeof(SmallShared) + 15) & ~size_t(15)));
  const int traj = COL ? (int)(blockIdx.x / (unsigned)col_dim) : (int)blockIdx.x;
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
  double* obs = COL ? nullptr : obs_all + (size_t)traj * n_out * QSB_NOBS;
  const uint32_t psi0 = COL ? blockIdx.x - (unsigned)traj * (unsigned)col_dim : d.psi0;

  RHS rhs;
  rhs.setup(d, xs);

  const DP5 T = dp5_tableau();
  cplx y = mk(live && x == psi0 ? 1.0 : 0.0, 0.0);
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
  if (!COL) small_observe(buf0, x, live, d, sh.red, obs);
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
    if (COL) continue;
    cplx* b = cur ? buf1 : buf0;
    b[x] = y;
    __syncthreads();
    small_observe(b, x, live, d, sh.red, obs + io * QSB_NOBS);
    cur ^= 1;
  }
  if (COL && live) columns[(size_t)blockIdx.x * dim + x] = y;
  if (threadIdx.x == 0) {
    SmallStats st;
    st.n_accepted = n_acc; st.n_rejected = n_rej; st.n_rhs = n_rhs; st.last_h = h; st.status = status; st.pad = 0;
    stats[blockIdx.x] = st;
  }
}

// Observable sums of one state by ONE warp (shuffle tree only, fixed order): lane l owns amplitudes l, l + 32, ...
// Same sums and output columns as small_observe.
__device__ __forceinline__ void warp_observe(const cplx* __restrict__ v, int dim, const SmallDesc& d,
                                             double* __restrict__ out, int64_t ks) {  // column k goes to out[k * ks]
  const int lane = threadIdx.x & 31;
  double s[7] = {0, 0, 0, 0, 0, 0, 0};  // nrm2, zsea, zrare, csea.re, csea.im, crare.re, crare.im
  for (uint32_t x = lane; x < (uint32_t)dim; x += 32) {
    cplx me = v[x];
    double p = norm2(me);
    s[0] += p;
    s[1] += p * (double)(__popc(d.sea_mask) - 2 * __popc(x & d.sea_mask));
    s[2] += (x & d.rare_mask) ? -p : p;
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
#pragma unroll
  for (int k = 0; k < 7; ++k) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s[k] += __shfl_xor_sync(0xffffffffu, s[k], off);
  }
  if (lane == 0) {
    double inv = 1.0 / s[0];
    out[0] = s[3] * inv;             // Ix_sea
    out[ks] = s[4] * inv;            // Iy_sea
    out[2 * ks] = 0.5 * s[1] * inv;  // Iz_sea
    out[3 * ks] = 0.5 * s[2] * inv;  // Iz_R
    out[4 * ks] = s[5] * inv;        // Ix_R
    out[5 * ks] = s[6] * inv;        // Iy_R
    out[6 * ks] = 1.0;               // norm of the output-normalised ket
    out[7 * ks] = sqrt(s[0]);        // raw norm
  }
}

// psi_k = U psi_{k-1}, k = 1 .. n_out-1, with the six observables of every psi_k: the constant-H fast path for
// uniform output grids (the reference's grid is np.linspace(0, t_final, steps), dipolar_ensemble_with_rare.py:626,
// and its H has no time dependence, :514 / :570).  One CTA per trajectory.  U arrives column-major
// (ut[j * dim + x] = <x|U|j>, the layout small_evolve_kernel<COL> leaves behind) and stays on the SM: the first
// n_cache columns in shared memory, each thread's next NREG columns in registers (at n = 7 that is all of U), and
// only what is left streams from L2 every step.
// Lane l of warp w is (xl = l & 7, ql = l >> 3): it accumulates the columns j = q (mod split) of row x with
// q = ql & (split - 1), x = 8 (w 4/split + ql / split) + xl, so a quarter warp reads 128 contiguous bytes of a
// column (conflict-free LDS.128) and the `split` partial sums of a row meet in one warp: they are added by a fixed
// xor-shuffle tree (bitwise reproducible) and psi_k costs ONE barrier.  The states go into a ring of `ring` slots;
// whenever it is full, warp r reduces the observables of slot r (warp_observe), so the reductions need no barriers
// of their own.
template <int MAXT, int NREG>
__global__ void __launch_bounds__(MAXT, 1)
small_propagate_kernel(const SmallDesc* __restrict__ descs, const cplx* __restrict__ ut_all, int64_t n_out,
                       double* __restrict__ obs_all, int split, int n_cache, int ring, int by_column,
                       const cplx* __restrict__ seeds, int n_seg, int64_t seg_len, cplx* __restrict__ states_out) {
  // Time-split chains: CTA b works on segment c = b % n_seg of trajectory b / n_seg, i.e. on the outputs
  // [c seg_len, min((c + 1) seg_len, n_out)), starting from seeds[b] = psi_{c seg_len} (seeds == nullptr, n_seg == 1: the
  // whole chain from the basis state).  states_out != nullptr: the states themselves are stored instead of their
  // observables (states_out[traj][output][x]) - how the seeds are made, with U^seg_len in place of U.
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SmallShared& sh = *reinterpret_cast<SmallShared*>(smem_raw);
  const int traj = blockIdx.x / n_seg, seg = blockIdx.x - traj * n_seg;
  {
    const int* src = reinterpret_cast<const int*>(descs + traj);
    int* dst = reinterpret_cast<int*>(&sh.d);
    for (int i = threadIdx.x; i < (int)(sizeof(SmallDesc) / 4); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const SmallDesc& d = sh.d;
  const int n = d.n, dim = 1 << n;
  cplx* states = reinterpret_cast<cplx*>(smem_raw + ((sizeof(SmallShared) + 15) & ~size_t(15)));  // [ring][dim]
  cplx* us = states + (size_t)ring * dim;                                                       // [n_cache][dim]
  const cplx* __restrict__ ut = ut_all + ((size_t)traj << (2 * n));
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  const int xl = lane & 7, ql = lane >> 3;
  const int q = ql & (split - 1);
  const int x = 8 * (warp * (4 / split) + ql / split) + xl;
  const bool worker = x < dim;
  const int64_t o0 = (int64_t)seg * seg_len;                                  // first output of this CTA
  const int64_t n_mine = (n_out - o0 < seg_len) ? n_out - o0 : seg_len;       // outputs of this CTA
  double* obs = obs_all + (size_t)traj * n_out * QSB_NOBS;

  for (int i = threadIdx.x; i < n_cache * dim; i += blockDim.x) us[i] = ut[i];
  for (int i = threadIdx.x; i < dim; i += blockDim.x)
    states[i] = seeds ? seeds[(size_t)blockIdx.x * dim + i] : mk((uint32_t)i == d.psi0 ? 1.0 : 0.0, 0.0);
  // first column of this thread that is not in shared memory, and the register-resident columns from there on
  int j_tail = q;
  while (j_tail < n_cache) j_tail += split;
  cplx ureg[NREG > 0 ? NREG : 1];
#pragma unroll
  for (int r = 0; r < NREG; ++r) {
    const int j = j_tail + r * split;
    ureg[r] = (worker && j < dim) ? ut[(size_t)j * dim + x] : mk(0, 0);
  }
  __syncthreads();

  for (int64_t base = 0; base < n_mine; base += ring) {  // slot r of this round holds psi_{o0 + base + r}
    const int count = (int)((n_mine - base < ring) ? n_mine - base : ring);
    for (int r = (base == 0 ? 1 : 0); r < count; ++r) {
      const cplx* cur = states + (size_t)(r == 0 ? ring - 1 : r - 1) * dim;
      cplx a0 = mk(0, 0), a1 = mk(0, 0);
      if (worker) {
        int j = q;
        for (; j + split < n_cache; j += 2 * split) {
          caxpy(a0, us[(size_t)j * dim + x], cur[j]);
          caxpy(a1, us[(size_t)(j + split) * dim + x], cur[j + split]);
        }
        if (j < n_cache) { caxpy(a0, us[(size_t)j * dim + x], cur[j]); j += split; }
#pragma unroll
        for (int rr = 0; rr < NREG; ++rr) {
          if (j < dim) caxpy((rr & 1) ? a1 : a0, ureg[rr], cur[j]);
          j += split;
        }
        for (; j + 3 * split < dim; j += 4 * split) {  // four independent L2 loads in flight per thread
          cplx u0 = __ldg(ut + (size_t)j * dim + x), u1 = __ldg(ut + (size_t)(j + split) * dim + x);
          cplx u2 = __ldg(ut + (size_t)(j + 2 * split) * dim + x), u3 = __ldg(ut + (size_t)(j + 3 * split) * dim + x);
          caxpy(a0, u0, cur[j]);
          caxpy(a1, u1, cur[j + split]);
          caxpy(a0, u2, cur[j + 2 * split]);
          caxpy(a1, u3, cur[j + 3 * split]);
        }
        for (; j < dim; j += split) caxpy(a0, __ldg(ut + (size_t)j * dim + x), cur[j]);
      }
      cplx sum = a0 + a1;
      for (int off = 8; off < 8 * split; off <<= 1) {  // the split partial sums of row x: lanes differing in q
        sum.x += __shfl_xor_sync(0xffffffffu, sum.x, off);
        sum.y += __shfl_xor_sync(0xffffffffu, sum.y, off);
      }
      if (worker && q == 0) states[(size_t)r * dim + x] = sum;
      __syncthreads();
    }
    // by_column: obs[traj][k][output] (what a caller that wants one contiguous trace per observable asks for)
    if (states_out) {
      cplx* dst = states_out + ((size_t)traj * n_out + o0 + base) * dim;
      for (int i = threadIdx.x; i < count * dim; i += blockDim.x) dst[i] = states[i];
    } else {
      for (int r = warp; r < count; r += nwarp)
        warp_observe(states + (size_t)r * dim, dim, d, obs + (o0 + base + r) * (by_column ? 1 : QSB_NOBS),
                     by_column ? n_out : 1);
    }
    __syncthreads();
  }
}

// n = 7 (the reference's sweep geometry) variant of small_propagate_kernel with TWO rows per thread, which halves the
// psi_j operand loads (45 % of the shared-memory wavefronts of the one-row kernel).  512 threads; lane l of warp w is
// (xl = l & 3, q = l >> 2): rows x = 4 w + xl and x + 64, columns j = q (mod 8).  A quarter warp reads two 64-byte
// pieces of neighbouring columns; the cached columns are stored with a stride of 128 + 4 elements so that those two
// pieces fall into different halves of the 32 banks.  Everything else (ring of states, per-warp observables, time-split
// segments, seeds, bitwise-fixed reduction order) is as in the general kernel.
constexpr int P2_DIM = 128, P2_CS = P2_DIM + 4, P2_SPLIT = 8, P2_NREG = 4, P2_THREADS = 512;

__global__ void __launch_bounds__(P2_THREADS, 1)
small_propagate_kernel_n7(const SmallDesc* __restrict__ descs, const cplx* __restrict__ ut_all, int64_t n_out,
                          double* __restrict__ obs_all, int n_cache, int ring, int by_column,
                          const cplx* __restrict__ seeds, int n_seg, int64_t seg_len, cplx* __restrict__ states_out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SmallShared& sh = *reinterpret_cast<SmallShared*>(smem_raw);
  const int traj = blockIdx.x / n_seg, seg = blockIdx.x - traj * n_seg;
  {
    const int* src = reinterpret_cast<const int*>(descs + traj);
    int* dst = reinterpret_cast<int*>(&sh.d);
    for (int i = threadIdx.x; i < (int)(sizeof(SmallDesc) / 4); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const SmallDesc& d = sh.d;
  constexpr int dim = P2_DIM;
  cplx* states = reinterpret_cast<cplx*>(smem_raw + ((sizeof(SmallShared) + 15) & ~size_t(15)));  // [ring][dim]
  cplx* us = states + (size_t)ring * dim;                                                       // [n_cache][P2_CS]
  const cplx* __restrict__ ut = ut_all + (size_t)traj * dim * dim;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = P2_THREADS / 32;
  const int q = lane >> 2;
  const int x = 4 * warp + (lane & 3);  // second row: x + 64
  const int64_t o0 = (int64_t)seg * seg_len;
  const int64_t n_mine = (n_out - o0 < seg_len) ? n_out - o0 : seg_len;
  double* obs = obs_all + (size_t)traj * n_out * QSB_NOBS;

  for (int i = threadIdx.x; i < n_cache * dim; i += blockDim.x) us[(i >> 7) * P2_CS + (i & (dim - 1))] = ut[i];
  for (int i = threadIdx.x; i < dim; i += blockDim.x)
    states[i] = seeds ? seeds[(size_t)blockIdx.x * dim + i] : mk((uint32_t)i == d.psi0 ? 1.0 : 0.0, 0.0);
  int j_tail = q;
  while (j_tail < n_cache) j_tail += P2_SPLIT;
  cplx ur0[P2_NREG], ur1[P2_NREG];  // register-resident columns j_tail + 8 r of both rows
#pragma unroll
  for (int r = 0; r < P2_NREG; ++r) {
    const int j = j_tail + r * P2_SPLIT;
    ur0[r] = j < dim ? ut[(size_t)j * dim + x] : mk(0, 0);
    ur1[r] = j < dim ? ut[(size_t)j * dim + x + 64] : mk(0, 0);
  }
  __syncthreads();

  for (int64_t base = 0; base < n_mine; base += ring) {
    const int count = (int)((n_mine - base < ring) ? n_mine - base : ring);
    for (int r = (base == 0 ? 1 : 0); r < count; ++r) {
      const cplx* cur = states + (size_t)(r == 0 ? ring - 1 : r - 1) * dim;
      cplx a0 = mk(0, 0), a1 = mk(0, 0), b0 = mk(0, 0), b1 = mk(0, 0);
      int j = q;
      for (; j + P2_SPLIT < n_cache; j += 2 * P2_SPLIT) {
        const cplx p0 = cur[j], p1 = cur[j + P2_SPLIT];
        caxpy(a0, us[j * P2_CS + x], p0);
        caxpy(b0, us[j * P2_CS + x + 64], p0);
        caxpy(a1, us[(j + P2_SPLIT) * P2_CS + x], p1);
        caxpy(b1, us[(j + P2_SPLIT) * P2_CS + x + 64], p1);
      }
      if (j < n_cache) {
        const cplx p0 = cur[j];
        caxpy(a0, us[j * P2_CS + x], p0);
        caxpy(b0, us[j * P2_CS + x + 64], p0);
        j += P2_SPLIT;
      }
#pragma unroll
      for (int rr = 0; rr < P2_NREG; ++rr) {
        if (j < dim) {
          const cplx p0 = cur[j];
          caxpy((rr & 1) ? a1 : a0, ur0[rr], p0);
          caxpy((rr & 1) ? b1 : b0, ur1[rr], p0);
        }
        j += P2_SPLIT;
      }
      for (; j < dim; j += P2_SPLIT) {  // only if fewer columns are cached than the launcher normally arranges
        const cplx p0 = cur[j];
        caxpy(a0, __ldg(ut + (size_t)j * dim + x), p0);
        caxpy(b0, __ldg(ut + (size_t)j * dim + x + 64), p0);
      }
      cplx sa = a0 + a1, sb = b0 + b1;
#pragma unroll
      for (int off = 4; off < 32; off <<= 1) {  // the 8 partial sums of a row: lanes differing in q
        sa.x += __shfl_xor_sync(0xffffffffu, sa.x, off);
        sa.y += __shfl_xor_sync(0xffffffffu, sa.y, off);
        sb.x += __shfl_xor_sync(0xffffffffu, sb.x, off);
        sb.y += __shfl_xor_sync(0xffffffffu, sb.y, off);
      }
      if (q == 0) {
        states[(size_t)r * dim + x] = sa;
        states[(size_t)r * dim + x + 64] = sb;
      }
      __syncthreads();
    }
    if (states_out) {
      cplx* dst = states_out + ((size_t)traj * n_out + o0 + base) * dim;
      for (int i = threadIdx.x; i < count * dim; i += blockDim.x) dst[i] = states[i];
    } else {
      for (int r = warp; r < count; r += nwarp)
        warp_observe(states + (size_t)r * dim, dim, d, obs + (o0 + base + r) * (by_column ? 1 : QSB_NOBS),
                     by_column ? n_out : 1);
    }
    __syncthreads();
  }
}

// M segments of ONE trajectory per CTA (time-split chains only): the M states advance together, so every element of U
// that a thread loads from shared memory (or holds in registers) feeds M complex multiply-adds instead of one - the
// one-segment kernel spends 1630 of its 2970 shared-memory wavefronts per output re-reading U
// (profiles/r1zz_ncu_propagator_summary.json).  CTA b works on segments [M g, M g + M) of trajectory b / n_grp,
// g = b % n_grp, n_grp = ceil(n_seg / M).  Thread mapping, column cache and register columns are those of
// small_propagate_kernel_n7; each (row, segment) has ONE accumulator (2 rows x M segments = 2 M independent chains),
// summed over j = q, q + 8, .. in order and then over q by the fixed xor-shuffle tree, so a segment's numbers do not
// depend on which other segments share its CTA.  The ring holds `ring` states per segment.
template <int M>
__global__ void __launch_bounds__(P2_THREADS, 1)
small_propagate_kernel_n7m(const SmallDesc* __restrict__ descs, const cplx* __restrict__ ut_all, int64_t n_out,
                           double* __restrict__ obs_all, int n_cache, int ring, int by_column,
                           const cplx* __restrict__ seeds, int n_seg, int64_t seg_len) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SmallShared& sh = *reinterpret_cast<SmallShared*>(smem_raw);
  const int n_grp = (n_seg + M - 1) / M;
  const int traj = blockIdx.x / n_grp, seg0 = (blockIdx.x - traj * n_grp) * M;
  {
    const int* src = reinterpret_cast<const int*>(descs + traj);
    int* dst = reinterpret_cast<int*>(&sh.d);
    for (int i = threadIdx.x; i < (int)(sizeof(SmallDesc) / 4); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const SmallDesc& d = sh.d;
  constexpr int dim = P2_DIM;
  cplx* states = reinterpret_cast<cplx*>(smem_raw + ((sizeof(SmallShared) + 15) & ~size_t(15)));  // [M][ring][dim]
  cplx* us = states + (size_t)M * ring * dim;                                                     // [n_cache][P2_CS]
  const cplx* __restrict__ ut = ut_all + (size_t)traj * dim * dim;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = P2_THREADS / 32;
  const int q = lane >> 2;
  const int x = 4 * warp + (lane & 3);  // second row: x + 64
  double* obs = obs_all + (size_t)traj * n_out * QSB_NOBS;
  // outputs of segment seg0 + s: [o0, o0 + n_mine); segments past n_seg are idle copies (never observed)
  auto first_output = [&](int s) { return (int64_t)(seg0 + s) * seg_len; };
  auto outputs_of = [&](int s) -> int64_t {
    if (seg0 + s >= n_seg) return 0;
    const int64_t left = n_out - first_output(s);
    return left < seg_len ? (left > 0 ? left : 0) : seg_len;
  };
  const int64_t n_max = outputs_of(0);  // segments are full except the trajectory's last one

  for (int i = threadIdx.x; i < n_cache * dim; i += blockDim.x) us[(i >> 7) * P2_CS + (i & (dim - 1))] = ut[i];
  for (int i = threadIdx.x; i < M * dim; i += blockDim.x) {
    const int s = i >> 7, xi = i & (dim - 1);
    states[(size_t)s * ring * dim + xi] =
        seg0 + s < n_seg ? seeds[((size_t)traj * n_seg + seg0 + s) * dim + xi] : mk(0.0, 0.0);
  }
  int j_tail = q;
  while (j_tail < n_cache) j_tail += P2_SPLIT;
  cplx ur0[P2_NREG], ur1[P2_NREG];
#pragma unroll
  for (int r = 0; r < P2_NREG; ++r) {
    const int j = j_tail + r * P2_SPLIT;
    ur0[r] = j < dim ? ut[(size_t)j * dim + x] : mk(0, 0);
    ur1[r] = j < dim ? ut[(size_t)j * dim + x + 64] : mk(0, 0);
  }
  __syncthreads();

  for (int64_t base = 0; base < n_max; base += ring) {
    const int count = (int)((n_max - base < ring) ? n_max - base : ring);
    for (int r = (base == 0 ? 1 : 0); r < count; ++r) {
      const cplx* cur = states + (size_t)(r == 0 ? ring - 1 : r - 1) * dim;  // segment s: + s * ring * dim
      cplx a[M], b[M];
#pragma unroll
      for (int s = 0; s < M; ++s) { a[s] = mk(0, 0); b[s] = mk(0, 0); }
      int j = q;
      for (; j < n_cache; j += P2_SPLIT) {
        const cplx u0 = us[j * P2_CS + x], u1 = us[j * P2_CS + x + 64];
#pragma unroll
        for (int s = 0; s < M; ++s) {
          const cplx p0 = cur[(size_t)s * ring * dim + j];
          caxpy(a[s], u0, p0);
          caxpy(b[s], u1, p0);
        }
      }
#pragma unroll
      for (int rr = 0; rr < P2_NREG; ++rr) {
        if (j < dim) {
#pragma unroll
          for (int s = 0; s < M; ++s) {
            const cplx p0 = cur[(size_t)s * ring * dim + j];
            caxpy(a[s], ur0[rr], p0);
            caxpy(b[s], ur1[rr], p0);
          }
        }
        j += P2_SPLIT;
      }
      for (; j < dim; j += P2_SPLIT) {  // only if fewer columns are cached than the launcher normally arranges
        const cplx u0 = __ldg(ut + (size_t)j * dim + x), u1 = __ldg(ut + (size_t)j * dim + x + 64);
#pragma unroll
        for (int s = 0; s < M; ++s) {
          const cplx p0 = cur[(size_t)s * ring * dim + j];
          caxpy(a[s], u0, p0);
          caxpy(b[s], u1, p0);
        }
      }
#pragma unroll
      for (int s = 0; s < M; ++s) {
#pragma unroll
        for (int off = 4; off < 32; off <<= 1) {  // the 8 partial sums of a row: lanes differing in q
          a[s].x += __shfl_xor_sync(0xffffffffu, a[s].x, off);
          a[s].y += __shfl_xor_sync(0xffffffffu, a[s].y, off);
          b[s].x += __shfl_xor_sync(0xffffffffu, b[s].x, off);
          b[s].y += __shfl_xor_sync(0xffffffffu, b[s].y, off);
        }
        if (q == 0) {
          states[((size_t)s * ring + r) * dim + x] = a[s];
          states[((size_t)s * ring + r) * dim + x + 64] = b[s];
        }
      }
      __syncthreads();
    }
    for (int idx = warp; idx < M * count; idx += nwarp) {
      const int s = idx / count, r = idx - s * count;
      if (base + r < outputs_of(s))
        warp_observe(states + ((size_t)s * ring + r) * dim, dim, d,
                     obs + (first_output(s) + base + r) * (by_column ? 1 : QSB_NOBS), by_column ? n_out : 1);
    }
    __syncthreads();
  }
}

// U <- U^(2^s): s = n_sq[traj] in-place squarings of one column-major propagator per CTA (scaling and squaring: the
// columns were integrated over dt / 2^s only).  C[x][j] = sum_k A[x][k] A[k][j]; thread (x, q) owns row x of the
// columns j = 4 (q + split i) .. +3, so the A[x][k] load is coalesced over x and shared by four accumulators while
// A[k][j] is a warp-uniform load; both operands stay in L1/L2 (the matrix is 256 KB at n = 7).  The two buffers
// swap roles after every squaring; results land back in `u_all`.  No __restrict__ / __ldg: the kernel reads what its
// own CTA wrote before the last barrier.
template <int MAXT>
__global__ void __launch_bounds__(MAXT, 1)
small_square_kernel(cplx* u_all, cplx* tmp_all, const int32_t* __restrict__ n_sq, int n, int split) {
  const int dim = 1 << n;
  const size_t off = (size_t)blockIdx.x << (2 * n);
  cplx* a = u_all + off;
  cplx* b = tmp_all + off;
  const int x = threadIdx.x & (dim - 1);
  const int q = threadIdx.x >> n;
  const int s_total = n_sq[blockIdx.x];
  for (int s = 0; s < s_total; ++s) {
    if (q < split) {
      for (int j0 = 4 * q; j0 < dim; j0 += 4 * split) {
        cplx acc[4] = {mk(0, 0), mk(0, 0), mk(0, 0), mk(0, 0)};
#pragma unroll 4
        for (int k = 0; k < dim; ++k) {
          const cplx ax = a[(size_t)k * dim + x];
#pragma unroll
          for (int r = 0; r < 4; ++r)
            if (j0 + r < dim) caxpy(acc[r], ax, a[(size_t)(j0 + r) * dim + k]);
        }
#pragma unroll
        for (int r = 0; r < 4; ++r)
          if (j0 + r < dim) b[(size_t)(j0 + r) * dim + x] = acc[r];
      }
    }
    __syncthreads();
    cplx* t = a; a = b; b = t;
  }
  if (s_total & 1) {  // the result sits in tmp: copy it home
    for (int i = threadIdx.x; i < dim * dim; i += blockDim.x) b[i] = a[i];
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

// n_ctas = trajectories (COL = false) or trajectories x 2^n columns (COL = true, col_dim = 2^n).
template <class RHS, bool COL, bool WIDE>
static void launch_evolve_t(const SmallDesc* d_desc, int n_ctas, int threads, const double* d_t_out, int64_t n_out,
                            double* d_obs, SmallStats* d_stats, cplx* d_columns, int col_dim, cudaStream_t stream) {
  size_t smem = ((sizeof(SmallShared) + 15) & ~size_t(15)) + 2 * sizeof(cplx) * threads;
  if (WIDE && threads > 512)  // 1024-thread CTAs are capped at 64 registers per thread (n = 10 only)
    small_evolve_kernel<RHS, WIDE ? 1024 : 256, COL><<<n_ctas, threads, smem, stream>>>(d_desc, d_t_out, n_out, d_obs, d_stats, d_columns, col_dim);
  else if (WIDE && threads > 256)
    small_evolve_kernel<RHS, WIDE ? 512 : 256, COL><<<n_ctas, threads, smem, stream>>>(d_desc, d_t_out, n_out, d_obs, d_stats, d_columns, col_dim);
  else
    small_evolve_kernel<RHS, 256, COL><<<n_ctas, threads, smem, stream>>>(d_desc, d_t_out, n_out, d_obs, d_stats, d_columns, col_dim);
}

}  // namespace

void launch_small_apply(const SmallDesc* d_desc, int n, const cplx* d_in, cplx* d_out, cudaStream_t stream) {
  int threads = 1 << n;
  if (threads < 32) threads = 32;
  size_t smem = ((sizeof(SmallShared) + 15) & ~size_t(15)) + sizeof(cplx) * threads;
  const int use_reg = getenv("QSB200_SMALL_GENERIC") ? 0 : 1;
  small_apply_kernel<<<1, threads, smem, stream>>>(d_desc, d_in, d_out, use_reg);
}

// ---- allocation-free H-apply (qsb_apply_h_static) ------------------------------------------------------------------
// Scratch of the variant below: [0, 4096) the compiled descriptor, then 2^10 input amplitudes, then 2^10 output
// amplitudes.  File-scope device memory: every GPU that loads the library holds these 36 KB until exit, every call
// shares them (calls that overlap on the GPU race), and they keep the last call's contents.
constexpr size_t STATIC_DESC_BYTES = 4096;
constexpr size_t STATIC_VEC_BYTES = sizeof(cplx) << SMALL_MAXQ;
constexpr size_t STATIC_SCRATCH_BYTES = STATIC_DESC_BYTES + 2 * STATIC_VEC_BYTES;
static_assert(sizeof(SmallDesc) <= STATIC_DESC_BYTES, "descriptor slot of the static scratch is too small");
static_assert(STATIC_SCRATCH_BYTES == QSB_STATIC_SCRATCH_BYTES && SMALL_MAXQ == QSB_STATIC_MAX_QUBITS, "header and scratch disagree");
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_SCRATCH_BYTES];

namespace {
// small_apply_kernel with its three device buffers inside ONE scratch block (the address comes from the host, which
// obtained it with cudaGetSymbolAddress: in host code the array's name is not a device address).
__global__ void small_apply_static_kernel(unsigned char* __restrict__ scratch, int use_reg) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SmallShared& sh = *reinterpret_cast<SmallShared*>(smem_raw);
  cplx* buf = reinterpret_cast<cplx*>(smem_raw + ((sizeof(SmallShared) + 15) & ~size_t(15)));
  const cplx* in = reinterpret_cast<const cplx*>(scratch + STATIC_DESC_BYTES);
  cplx* out = reinterpret_cast<cplx*>(scratch + STATIC_DESC_BYTES + STATIC_VEC_BYTES);
  {
    const int* src = reinterpret_cast<const int*>(scratch);
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
  cplx k;
  if (use_reg && d.n == 7) { RhsReg<7> r; r.setup(d, xs); k = r(buf); }
  else if (use_reg && d.n == 8) { RhsReg<8> r; r.setup(d, xs); k = r(buf); }
  else if (use_reg && d.n == 5) { RhsReg<5> r; r.setup(d, xs); k = r(buf); }
  else if (use_reg && d.n == 3) { RhsReg<3> r; r.setup(d, xs); k = r(buf); }
  else { RhsGeneric r; r.setup(d, xs); k = r(buf); }
  if (live) out[x] = mk(-k.y, k.x);  // i * k = H psi
}
}  // namespace

// Device address of the scratch on the CURRENT device, resolved once per device (cudaGetSymbolAddress answers for the
// device that is current when it is called).
static cudaError_t static_scratch_ptr(unsigned char** out) {
  constexpr int MAX_DEV = 64;
  static std::mutex mu;
  static unsigned char* ptrs[MAX_DEV] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= MAX_DEV) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lk(mu);
  if (!ptrs[dev]) {
    unsigned char* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_static_scratch);
    if (e != cudaSuccess) return e;
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return cudaSuccess;
}

size_t small_apply_static_bytes() { return STATIC_SCRATCH_BYTES; }

// out_host = H in_host for one n <= SMALL_MAXQ table, with NO device allocation: descriptor, input and output live in
// g_static_scratch.  The device must be current; blocks until the result is on the host.
cudaError_t launch_small_apply_static(const SmallDesc* h_desc, int n, const cplx* h_in, cplx* h_out, cudaStream_t stream) {
  unsigned char* scratch = nullptr;
  cudaError_t e = static_scratch_ptr(&scratch);
  if (e != cudaSuccess) return e;
  const size_t bytes = sizeof(cplx) << n;
  e = cudaMemcpyAsync(scratch, h_desc, sizeof(SmallDesc), cudaMemcpyHostToDevice, stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(scratch + STATIC_DESC_BYTES, h_in, bytes, cudaMemcpyHostToDevice, stream);
  // the array keeps the previous call's output: start from zero
  if (e == cudaSuccess) e = cudaMemsetAsync(scratch + STATIC_DESC_BYTES + STATIC_VEC_BYTES, 0, STATIC_VEC_BYTES, stream);
  if (e != cudaSuccess) return e;
  int threads = 1 << n;
  if (threads < 32) threads = 32;
  const size_t smem = ((sizeof(SmallShared) + 15) & ~size_t(15)) + sizeof(cplx) * threads;
  const int use_reg = getenv("QSB200_SMALL_GENERIC") ? 0 : 1;
  small_apply_static_kernel<<<1, threads, smem, stream>>>(scratch, use_reg);
  e = cudaGetLastError();
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(h_out, scratch + STATIC_DESC_BYTES + STATIC_VEC_BYTES, bytes, cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  return e;
}

template <bool COL>
static void launch_evolve_any(const SmallDesc* d_desc, int n_ctas, int max_n, int uniform_n, const double* d_t_out,
                              int64_t n_out, double* d_obs, SmallStats* d_stats, cplx* d_columns, cudaStream_t stream) {
  int threads = 1 << max_n;
  if (threads < 32) threads = 32;
  const bool reg = uniform_n && !getenv("QSB200_SMALL_GENERIC");
  const int col_dim = 1 << max_n;
#define QSB_REG_CASE(NB)                                                                                         \
  case NB:                                                                                                       \
    launch_evolve_t<RhsReg<NB>, COL, false>(d_desc, n_ctas, threads, d_t_out, n_out, d_obs, d_stats, d_columns,  \
                                            col_dim, stream);                                                    \
    return;
  if (reg) {
    switch (max_n) {
      QSB_REG_CASE(2) QSB_REG_CASE(3) QSB_REG_CASE(4) QSB_REG_CASE(5) QSB_REG_CASE(6) QSB_REG_CASE(7) QSB_REG_CASE(8)
      default: break;
    }
  }
#undef QSB_REG_CASE
  launch_evolve_t<RhsGeneric, COL, true>(d_desc, n_ctas, threads, d_t_out, n_out, d_obs, d_stats, d_columns, col_dim,
                                         stream);
}

// `uniform_n`: every trajectory of the batch has exactly max_n qubits (required by the register-resident rows).
void launch_small_evolve(const SmallDesc* d_desc, int n_traj, int max_n, int uniform_n, const double* d_t_out,
                         int64_t n_out, double* d_obs, SmallStats* d_stats, cudaStream_t stream) {
  launch_evolve_any<false>(d_desc, n_traj, max_n, uniform_n, d_t_out, n_out, d_obs, d_stats, nullptr, stream);
}

void launch_small_columns(const SmallDesc* d_desc, int n_traj, int n, const double* d_t_pairs, SmallStats* d_stats,
                          cplx* d_columns, cudaStream_t stream) {
  launch_evolve_any<true>(d_desc, n_traj << n, n, 1, d_t_pairs, 2, nullptr, d_stats, d_columns, stream);
}

static constexpr size_t PROPAGATE_SMEM_MAX = 226 * 1024;  // of the 227 KB a CTA may opt in to
static constexpr int PROPAGATE_NREG = 8;                   // columns of U per thread kept in registers (<= 512 threads)

void small_propagate_shape(int n, int* threads, int* split, int* n_cache, int* ring, size_t* smem) {
  const int dim = 1 << n;
  int sp = dim >= 512 ? 1 : (dim > 128 ? 2 : 4);  // partial sums of a row live in one warp: 1, 2 or 4
  if (sp > dim) sp = dim;
  int th = dim * sp;
  if (th < 32) th = 32;
  // ring of states between two rounds of observable reductions: one slot per warp, at most 16 KB, at least 2
  int rg = th / 32;
  const int fit = (int)(16384 / ((size_t)dim * sizeof(cplx)));
  if (rg > fit) rg = fit;
  if (rg < 2) rg = 2;
  const size_t fixed = ((sizeof(SmallShared) + 15) & ~size_t(15)) + (size_t)rg * dim * sizeof(cplx);
  size_t room = (PROPAGATE_SMEM_MAX - fixed) / ((size_t)dim * sizeof(cplx));
  int nc = room > (size_t)dim ? dim : (int)room;
  *threads = th; *split = sp; *n_cache = nc; *ring = rg;
  *smem = fixed + (size_t)nc * dim * sizeof(cplx);
}

void launch_small_square(cplx* d_columns, cplx* d_tmp, const int32_t* d_n_sq, int n_traj, int n, cudaStream_t stream) {
  int threads, split, n_cache, ring;
  size_t smem;
  small_propagate_shape(n, &threads, &split, &n_cache, &ring, &smem);
  if (threads > 512)
    small_square_kernel<1024><<<n_traj, threads, 0, stream>>>(d_columns, d_tmp, d_n_sq, n, split);
  else
    small_square_kernel<512><<<n_traj, threads, 0, stream>>>(d_columns, d_tmp, d_n_sq, n, split);
}

int launch_small_propagate(const SmallDesc* d_desc, int n_traj, int n, const cplx* d_columns, int64_t n_out,
                           double* d_obs, int by_column, cudaStream_t stream, const cplx* d_seeds, int n_seg,
                           int64_t seg_len, cplx* d_states_out) {
  int threads, split, n_cache, ring;
  size_t smem;
  small_propagate_shape(n, &threads, &split, &n_cache, &ring, &smem);
  if (n_seg < 1) n_seg = 1;
  if (seg_len < 1) seg_len = n_out;
  const unsigned grid = (unsigned)n_traj * (unsigned)n_seg;
  // n = 7 time-split chains: PROP_M segments of a trajectory per CTA share every U element they load.  (QSB200_PROP_M=1:
  // the one-segment kernel; seeds / state-storing chains always use it.)
  constexpr int PROP_M = 4;
  const char* pm = getenv("QSB200_PROP_M");
  if (n == 7 && d_seeds && !d_states_out && n_seg >= 2 && !(pm && atoi(pm) == 1) && !getenv("QSB200_PROP_ONE_ROW")) {
    const int ring_m = 2;  // states per segment between two rounds of observable reductions (M x 2 x 2 KB = 16 KB)
    const size_t fixed = ((sizeof(SmallShared) + 15) & ~size_t(15)) + (size_t)PROP_M * ring_m * P2_DIM * sizeof(cplx);
    int nc = (int)((PROPAGATE_SMEM_MAX - fixed) / ((size_t)P2_CS * sizeof(cplx)));
    if (nc > P2_DIM) nc = P2_DIM;
    const size_t smem2 = fixed + (size_t)nc * P2_CS * sizeof(cplx);
    const unsigned grid_m = (unsigned)n_traj * (unsigned)((n_seg + PROP_M - 1) / PROP_M);
    QSB_CUDA(cudaFuncSetAttribute(small_propagate_kernel_n7m<PROP_M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
    small_propagate_kernel_n7m<PROP_M><<<grid_m, P2_THREADS, smem2, stream>>>(d_desc, d_columns, n_out, d_obs, nc, ring_m,
                                                                               by_column, d_seeds, n_seg, seg_len);
    QSB_CUDA(cudaGetLastError());
    return QSB_OK;
  }
  if (n == 7 && !getenv("QSB200_PROP_ONE_ROW")) {  // two rows per thread, padded column stride
    const size_t fixed = ((sizeof(SmallShared) + 15) & ~size_t(15)) + (size_t)ring * P2_DIM * sizeof(cplx);
    int nc = (int)((PROPAGATE_SMEM_MAX - fixed) / ((size_t)P2_CS * sizeof(cplx)));
    if (nc > P2_DIM) nc = P2_DIM;
    const size_t smem2 = fixed + (size_t)nc * P2_CS * sizeof(cplx);
    QSB_CUDA(cudaFuncSetAttribute(small_propagate_kernel_n7, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
    small_propagate_kernel_n7<<<grid, P2_THREADS, smem2, stream>>>(d_desc, d_columns, n_out, d_obs, nc, ring, by_column,
                                                                   d_seeds, n_seg, seg_len, d_states_out);
    QSB_CUDA(cudaGetLastError());
    return QSB_OK;
  }
  if (threads > 512) {
    QSB_CUDA(cudaFuncSetAttribute(small_propagate_kernel<1024, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    small_propagate_kernel<1024, 0><<<grid, threads, smem, stream>>>(d_desc, d_columns, n_out, d_obs, split, n_cache, ring,
                                                                     by_column, d_seeds, n_seg, seg_len, d_states_out);
  } else {
    QSB_CUDA(cudaFuncSetAttribute(small_propagate_kernel<512, PROPAGATE_NREG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    small_propagate_kernel<512, PROPAGATE_NREG><<<grid, threads, smem, stream>>>(
        d_desc, d_columns, n_out, d_obs, split, n_cache, ring, by_column, d_seeds, n_seg, seg_len, d_states_out);
  }
  QSB_CUDA(cudaGetLastError());
  return QSB_OK;
}

}  // namespace qsb
