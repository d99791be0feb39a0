// Tiled-path stepping session (host-side adaptive controller around launch_happly) and qsb_evolve.
using namespace qsb;

#include "qsb_api_shard.inl"

struct Session {
  qsb_solver_opts o;
  int order = 5;
  double inv_q = 0.2;
  double t = 0.0, h = 0.0;
  bool after_reject = false;
  bool have_k1 = false;
  int64_t n_acc = 0, n_rej = 0, n_rhs = 0;
  uint64_t dim = 0;        // amplitudes of the padded vector
  double norm_dim = 0.0;   // 2^n of the user's problem: divisor of the weighted RMS norms
  cplx *y = nullptr, *ys = nullptr, *k[7] = {}, *tx = nullptr, *ty = nullptr;
  double* partials = nullptr;
  int64_t n_tiles = 0;
  double obs_sums[PSTRIDE] = {};  // sums belonging to the current y (when valid)
  bool obs_valid = false;
  // Dense output (Taylor, constant H): when the output grid is finer than the controller's step, ONE natural step is
  // taken with every term kept (dense[m] = term_m = (hu^m / m!) (-iH)^m y, m = 1 .. order), and every output time
  // inside it is read off the same polynomial, psi(t + s hu) = y + sum_m s^m term_m - one combination + one
  // observation per output instead of `order` H-applies per output.  Owned by the plan (qsb_plan::dense_buf).
  cplx* dense[QSB_MAX_TAYLOR_ORDER + 1] = {};
  bool dense_on = false;
};

static void carve(qsb_plan* p, Session* s) {
  char* base = (char*)p->ws;
  size_t vec = sizeof(cplx) << p->n_loc;
  int i = 0;
  auto next = [&]() { return (cplx*)(base + vec * (i++)); };
  s->y = next(); s->ys = next();
  if (p->world > 1) { s->tx = (cplx*)p->xbuf; s->ty = (cplx*)((char*)p->xbuf + vec); }
  else { s->tx = next(); s->ty = next(); }
  if (s->o.method == QSB_METHOD_DP5) {
    for (int j = 0; j < 7; ++j) s->k[j] = next();
  } else {
    s->k[0] = next();  // running Taylor term
  }
  s->partials = (double*)(base + vec * i);
}

static int fetch_sums(qsb_plan* p, Session* s, double* out) {
  launch_reduce_partials(s->partials, s->n_tiles, p->d_red, p->stream);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaMemcpyAsync(p->h_red, p->d_red, PSTRIDE * sizeof(double), cudaMemcpyDeviceToHost, p->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
  if (p->world == 1) {
    if (e != cudaSuccess) return cuda_fail(e, "fetch_sums", __FILE__, __LINE__);
    memcpy(out, p->h_red, PSTRIDE * sizeof(double));
    return QSB_OK;
  }
  // sharded: the last slot (unused by the kernels) carries this rank's failure flag through the rank sum, so a
  // failing rank takes its peers down with it instead of leaving them at the all-gather
  memcpy(out, p->h_red, PSTRIDE * sizeof(double));
  if (e != cudaSuccess) for (int k = 0; k < PSTRIDE; ++k) out[k] = 0.0;
  out[PSTRIDE - 1] = (e == cudaSuccess) ? 0.0 : 1.0;
  int rc = shard_sum(p, out, PSTRIDE);
  if (e != cudaSuccess) return cuda_fail(e, "fetch_sums", __FILE__, __LINE__);
  if (rc) return rc;
  if (out[PSTRIDE - 1] != 0.0) return fail(QSB_ERR_CUDA, "a peer rank reported a CUDA failure");
  return QSB_OK;
}

static ShardLink* link_of(qsb_plan* p) { return p->world > 1 && p->connected ? &p->link : nullptr; }

static Epilogue blank_epilogue(const Session* s) {
  Epilogue e = {};
  e.alpha = 1.0;
  e.sc_i0 = 0; e.sc_i1 = 0;
  e.atol = s->o.atol; e.rtol = s->o.rtol;
  return e;
}

static void obs_from_sums(const qsb_plan* p, const double* s, double* obs) {
  const double nrm2 = s[0], sx = std::ldexp(0.5, -p->n_eff) / nrm2;
  obs[0] = s[3] * sx;            // Ix_sea
  obs[1] = s[5] * sx;            // Iy_sea
  obs[2] = 0.5 * s[1] / nrm2;    // Iz_sea
  obs[3] = 0.5 * s[2] / nrm2;    // Iz_R
  obs[4] = s[4] * sx;            // Ix_R
  obs[5] = s[6] * sx;            // Iy_R
  obs[6] = 1.0;
  obs[7] = std::sqrt(nrm2);
}

// k1 = f(y) with observables of y and the start-up step size.
static int session_prime(qsb_plan* p, Session* s) {
  Epilogue e = blank_epilogue(s);
  e.o1 = s->k[0];
  e.vin[0] = s->y; e.n_in = 1; e.c[0] = 0.0; e.e[0] = 0.0; e.ek = 1.0; e.want_err = 1;
  if (s->o.method == QSB_METHOD_TAYLOR) e.o1 = nullptr;  // Taylor recomputes its first term with h folded in
  int rc = launch_happly(p->g, p->tab, s->y, s->tx, s->ty, e, 1, s->partials, p->stream, link_of(p), nullptr, &p->fu);
  if (rc) return rc;
  QSB_CUDA(cudaGetLastError());
  ++s->n_rhs;
  double sums[PSTRIDE];
  rc = fetch_sums(p, s, sums);
  if (rc) return rc;
  memcpy(s->obs_sums, sums, sizeof sums);
  s->obs_valid = true;
  s->have_k1 = (s->o.method == QSB_METHOD_DP5);
  if (s->h == 0.0) {
    double d1 = std::sqrt(sums[7] / s->norm_dim), d0 = std::sqrt(sums[8] / s->norm_dim);
    s->h = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
    if (s->o.first_step > 0.0) s->h = s->o.first_step;
    if (s->o.max_step > 0.0 && s->h > s->o.max_step) s->h = s->o.max_step;
  }
  return QSB_OK;
}

// One step attempt of size hu.  Returns the weighted RMS error in *err; obs sums of the candidate in cand_sums.
// Every stage input y + hu sum_j a_sj k_j is formed inside pass_first of the apply that consumes it (Combine), the
// applies end in the plain k-producing pass_last, and the embedded error estimate rides on the last stage (below).
static bool dp5_error_sweep() {  // A/B knob: the embedded error estimate as its own 8 S streaming sweep (the round-1 form)
  static int on = -1;
  if (on < 0) { const char* e = getenv("QSB200_DP5_ERR_SWEEP"); on = (e && *e == '1') ? 1 : 0; }
  return on == 1;
}

static int attempt_dp5(qsb_plan* p, Session* s, double hu, double* err, double* cand_sums) {
  static const DP5 T = dp5_tableau();
  const bool sweep = dp5_error_sweep();
  // Stages 0 .. 3 end in pass_last_first: the kernel that finishes apply #st also forms the input of apply #st + 1 from
  // the k it has just produced (still in registers) and the older stage vectors, and starts its transform; the next
  // apply then skips its pass_first.  (QSB200_DP5_FUSE=0: every stage is pass_first .. pass_last, the round-1 form.)
  static const bool fuse_stages = [] { const char* e = getenv("QSB200_DP5_FUSE"); return !(e && *e == '0'); }();
  auto combination = [&](int st, bool skip_fresh) {  // y + hu sum_j a[st][j] k_j, optionally without the newest k
    Combine cb = {};
    cb.vin[0] = s->y; cb.c[0] = 1.0; cb.e[0] = 0.0; cb.n_in = 1;
    for (int j = 0; j <= st; ++j) {
      if (skip_fresh && j == st) continue;
      if (T.a[st][j] != 0.0) { cb.vin[cb.n_in] = s->k[j]; cb.c[cb.n_in] = hu * T.a[st][j]; cb.e[cb.n_in] = hu * T.e[j]; ++cb.n_in; }
    }
    cb.out = s->ys;
    return cb;
  };
  bool fused_in = false;  // the previous apply's last kernel has already begun this one
  for (int st = 0; st <= 5; ++st) {  // apply #st: src = ys = y + hu sum_j a[st][j] k_j, produces k_{st+1} (0-based k[st+1])
    Combine cb = combination(st, false);
    Epilogue e = blank_epilogue(s);
    e.o1 = s->k[st + 1];
    if (st == 5 && !sweep) {
      // The embedded error estimate rides on the last stage: its pass_first reads y, k_1, k_3 .. k_6 anyway (the
      // tableau's zero pattern, a_72 = e_2 = 0, is the same for both rows) and also forms Q = hu sum_{j<=6} e_j k_j,
      // parked in k_7's own storage; its pass_last then reads Q (in place) and y and reduces
      // |Q + hu e_7 k_7|^2 / sc^2 with sc = atol + rtol max(|y|, |y_new|): 3 S of traffic instead of an 8 S sweep.
      cb.eout = s->k[6];
      e.vin[0] = s->k[6]; e.c[0] = 0.0; e.e[0] = 1.0;
      e.vin[1] = s->y;    e.c[1] = 0.0; e.e[1] = 0.0;
      e.n_in = 2;
      e.want_err = 1; e.ek = hu * T.e[6]; e.sc_i0 = 1; e.sc_i1 = -2;
    }
    FuseNext fn = {};
    const bool fuse_out = fuse_stages && st <= 3 && T.a[st + 1][st + 1] != 0.0;
    if (fuse_out) {
      fn.cb = combination(st + 1, true);
      fn.c_fresh = hu * T.a[st + 1][st + 1];
    }
    // the last apply acts on y_new (FSAL): its pass_first / pass_centre also collect y_new's observable sums
    int rc = launch_happly(p->g, p->tab, s->ys, s->tx, s->ty, e, st == 5 ? 1 : 0, s->partials, p->stream, link_of(p), &cb, &p->fu,
                           fuse_out ? &fn : nullptr, fused_in);
    if (rc) return rc;
    fused_in = fuse_out;
  }
  if (sweep) {
    ErrSpec es = {};
    es.vin[0] = s->y; es.vin[1] = s->ys; es.n_in = 2;
    for (int j = 0; j < 7; ++j)
      if (T.e[j] != 0.0) { es.vin[es.n_in] = s->k[j]; es.e[es.n_in] = hu * T.e[j]; ++es.n_in; }
    es.sc_i0 = 0; es.sc_i1 = 1;
    es.atol = s->o.atol; es.rtol = s->o.rtol;
    launch_error_sums(p->g, es, s->partials, p->stream);
  }
  QSB_CUDA(cudaGetLastError());
  s->n_rhs += 6;
  int rc = fetch_sums(p, s, cand_sums);
  if (rc) return rc;
  *err = std::sqrt(cand_sums[7] / s->norm_dim);
  return QSB_OK;
}

static int attempt_taylor(qsb_plan* p, Session* s, double hu, double* err, double* cand_sums, bool keep_terms = false) {
  // k[0] holds the running term, term_{m+1} = hu/(m+1) * f(term_m), term_0 = y; ys becomes y + sum of terms.
  // The vector an apply acts on (term_m) is in pass_last's registers anyway, so ys is only touched by every second
  // apply, which adds BOTH term_m and term_{m+1}:  m = 1: ys = y + term_1 + term_2;  m = 3, 5, ..: ys += term_m +
  // term_{m+1};  an odd order's last apply adds its single term.  The other applies end in the plain term-producing
  // pass_last (4 S instead of 6 S), and no initial copy of y is needed.
  for (int m = 0; m < s->order; ++m) {
    const bool last = m == s->order - 1;
    Epilogue e = blank_epilogue(s);
    e.alpha = hu / (double)(m + 1);
    e.o1 = keep_terms ? s->dense[m + 1] : s->k[0];
    if ((m & 1) || last) {
      e.o2 = s->ys;
      e.vin[0] = m <= 1 ? s->y : s->ys; e.c[0] = 1.0; e.n_in = 1;
      e.ck = e.alpha;
      e.cs = (m & 1) ? 1.0 : 0.0;
    }
    if (last) {  // last term = error estimate; y is read here for the error weights
      e.vin[e.n_in] = s->y; e.c[e.n_in] = 0.0;
      e.want_err = 1; e.ek = e.alpha; e.sc_i0 = e.n_in; e.sc_i1 = -1;
      ++e.n_in;
    }
    const cplx* src = m == 0 ? s->y : (keep_terms ? s->dense[m] : s->k[0]);
    int rc = launch_happly(p->g, p->tab, src, s->tx, s->ty, e, 0, s->partials, p->stream, link_of(p), nullptr, &p->fu);
    if (rc) return rc;
  }
  QSB_CUDA(cudaGetLastError());
  s->n_rhs += s->order;
  int rc = fetch_sums(p, s, cand_sums);
  if (rc) return rc;
  *err = std::sqrt(cand_sums[7] / s->norm_dim);
  return QSB_OK;
}

// Advance until t_target (if finite) or for max_attempts attempts, whichever comes first.
static int session_advance(qsb_plan* p, Session* s, double t_target, int64_t max_attempts, int64_t nsteps_cap,
                           int64_t* attempts_done) {
  int64_t in_interval = 0;
  *attempts_done = 0;
  const bool bounded = std::isfinite(t_target);
  while ((!bounded || s->t < t_target) && *attempts_done < max_attempts) {
    if (nsteps_cap > 0 && in_interval >= nsteps_cap)
      return fail(QSB_ERR_NSTEPS, "more than nsteps internal steps inside one output interval (excess work)");
    if (s->o.method == QSB_METHOD_DP5 && !s->have_k1) {
      int rc = session_prime(p, s);
      if (rc) return rc;
    }
    double hmax = s->o.max_step > 0.0 ? std::min(s->h, s->o.max_step) : s->h;
    double hu = hmax;
    bool hit = false;
    if (bounded) {
      double remaining = t_target - s->t;
      if (hu >= remaining * (1.0 - 1e-12)) { hu = remaining; hit = true; }
      else if (2.0 * hu > remaining) hu = 0.5 * remaining;
    }
    if (hu <= 16.0 * 2.220446049250313e-16 * std::fabs(s->t)) return fail(QSB_ERR_STEP_UNDERFLOW, "step size underflow");
    double err = 0.0, cand[PSTRIDE];
    int rc = (s->o.method == QSB_METHOD_DP5) ? attempt_dp5(p, s, hu, &err, cand) : attempt_taylor(p, s, hu, &err, cand);
    if (rc) return rc;
    ++in_interval;
    ++*attempts_done;
    if (!(err == err)) return fail(QSB_ERR_CUDA, "non-finite error estimate (state blew up)");
    if (err <= 1.0) {
      std::swap(s->y, s->ys);
      if (s->o.method == QSB_METHOD_DP5) {
        std::swap(s->k[0], s->k[6]);
        memcpy(s->obs_sums, cand, sizeof cand);
        s->obs_valid = true;
      } else {
        s->obs_valid = false;
      }
      s->t = hit ? t_target : s->t + hu;
      ++s->n_acc;
      double hn = hu * Controller::factor(err, s->inv_q, s->after_reject);
      s->h = (hu < hmax) ? std::max(hn, s->h) : hn;
      s->after_reject = false;
    } else {
      ++s->n_rej;
      s->h = hu * Controller::factor(err, s->inv_q, true);
      s->after_reject = true;
    }
  }
  return QSB_OK;
}

static int session_observe(qsb_plan* p, Session* s, double* obs) {
  if (!s->obs_valid) {
    int rc = launch_observe(p->g, p->tab, s->y, s->tx, s->ty, s->partials, p->stream, link_of(p), &p->fu);
    if (rc) return rc;
    QSB_CUDA(cudaGetLastError());
    rc = fetch_sums(p, s, s->obs_sums);
    if (rc) return rc;
    s->obs_valid = true;
  }
  obs_from_sums(p, s->obs_sums, obs);
  return QSB_OK;
}

// Observables of an arbitrary local vector (tx / ty are clobbered).
static int observe_vec(qsb_plan* p, Session* s, const cplx* vec, double* obs) {
  int rc = launch_observe(p->g, p->tab, vec, s->tx, s->ty, s->partials, p->stream, link_of(p), &p->fu);
  if (rc) return rc;
  QSB_CUDA(cudaGetLastError());
  double sums[PSTRIDE];
  rc = fetch_sums(p, s, sums);
  if (rc) return rc;
  obs_from_sums(p, sums, obs);
  return QSB_OK;
}

// Dense output is offered when the kept terms fit a modest extra allocation: Taylor sessions of plans with at most
// 2^22 local amplitudes (64 MB per vector; QSB200_DENSE_OUTPUT=0 turns it off, =1 lifts the size limit).
static bool dense_setup(qsb_plan* p, Session* s) {
  s->dense_on = false;
  if (s->o.method != QSB_METHOD_TAYLOR) return false;
  const char* e = getenv("QSB200_DENSE_OUTPUT");
  if (e && *e == '0') return false;
  if (p->n_loc > 22 && !(e && *e == '1')) return false;
  const size_t vec = sizeof(cplx) << p->n_loc, need = vec * (size_t)s->order;
  if (p->dense_bytes < need) {
    if (p->dense_buf) cudaFree(p->dense_buf);
    p->dense_buf = nullptr; p->dense_bytes = 0;
    if (cudaMalloc(&p->dense_buf, need) != cudaSuccess) { cudaGetLastError(); p->dense_buf = nullptr; return false; }
    p->dense_bytes = need;
  }
  for (int m = 1; m <= s->order; ++m) s->dense[m] = (cplx*)((char*)p->dense_buf + vec * (size_t)(m - 1));
  s->dense_on = true;
  return true;
}

// One natural step from s->t that is NOT clipped to the next output time; every output time it covers is evaluated
// from the step's own Taylor polynomial.  *k: next output index (advanced past the outputs served).  Returns with the
// step accepted (state advanced) or rejected (h reduced); the caller loops.  *in_interval counts attempts since the
// last served output (zvode's nsteps cap applies per output interval).
static int dense_step(qsb_plan* p, Session* s, const double* t_out, int64_t n_out, int64_t* k, double* out_obs,
                      int64_t* in_interval) {
  if (s->o.nsteps > 0 && *in_interval >= s->o.nsteps)
    return fail(QSB_ERR_NSTEPS, "more than nsteps internal steps inside one output interval (excess work)");
  const double t_final = t_out[n_out - 1];
  const double hmax = s->o.max_step > 0.0 ? std::min(s->h, s->o.max_step) : s->h;
  double hu = hmax;
  bool hit = false;
  const double remaining = t_final - s->t;
  if (hu >= remaining * (1.0 - 1e-12)) { hu = remaining; hit = true; }
  if (hu <= 16.0 * 2.220446049250313e-16 * std::fabs(s->t)) return fail(QSB_ERR_STEP_UNDERFLOW, "step size underflow");
  double err = 0.0, cand[PSTRIDE];
  int rc = attempt_taylor(p, s, hu, &err, cand, true);
  if (rc) return rc;
  ++*in_interval;
  if (!(err == err)) return fail(QSB_ERR_CUDA, "non-finite error estimate (state blew up)");
  if (err > 1.0) {
    ++s->n_rej;
    s->h = hu * Controller::factor(err, s->inv_q, true);
    s->after_reject = true;
    return QSB_OK;
  }
  const double t_new = hit ? t_final : s->t + hu;
  while (*k < n_out && (t_out[*k] <= t_new || (hit && *k == n_out - 1))) {
    const double frac = (hit && *k == n_out - 1) ? 1.0 : (t_out[*k] - s->t) / hu;
    const cplx* vec = s->ys;                       // the full step
    if (frac < 1.0 - 1e-13) {
      DenseSpec d = {};
      d.y = s->y; d.out = s->k[0]; d.n = s->order;
      double pw = 1.0;
      for (int m = 0; m < s->order; ++m) { pw *= frac; d.term[m] = s->dense[m + 1]; d.c[m] = pw; }
      launch_dense_eval(d, 1ull << p->n_loc, p->stream);
      vec = s->k[0];
    }
    rc = observe_vec(p, s, vec, out_obs + *k * QSB_NOBS);
    if (rc) return rc;
    ++*k;
    *in_interval = 0;
  }
  std::swap(s->y, s->ys);
  s->obs_valid = false;
  s->t = t_new;
  ++s->n_acc;
  const double hn = hu * Controller::factor(err, s->inv_q, s->after_reject);
  s->h = (hu < hmax) ? std::max(hn, s->h) : hn;
  s->after_reject = false;
  return QSB_OK;
}

#include "qsb_api_entry.inl"
