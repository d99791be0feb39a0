// extern "C" entry points of the tiled path + qsb_evolve + qsb_apply_h; included from qsb_api_session.inl.

static int upload_state(qsb_plan* p, cplx* dst, uint64_t psi0_index, const double* psi0_host) {
  const uint64_t dim = 1ull << p->n_loc, udim = p->world > 1 ? dim : (1ull << p->n);
  if (!psi0_host && p->identity_bits) {  // the propagator build: all basis states at once (qsb_propagate_hbm)
    launch_set_identity(dst, p->identity_bits, p->stream);
    QSB_CUDA(cudaGetLastError());
  } else if (!psi0_host) {
    if (psi0_index >> p->n) return fail(QSB_ERR_BAD_ARG, "psi0_index out of range");
    // sharded: only the owner of the index sets a 1 (local index); everybody else zero-fills
    uint64_t local = psi0_index;
    if (p->world > 1) local = (psi0_index >> p->n_loc) == (uint64_t)p->rank ? (psi0_index & (dim - 1)) : ~0ull;
    launch_set_basis_state(dst, dim, local, p->stream);
    QSB_CUDA(cudaGetLastError());
  } else {
    if (dim != udim) QSB_CUDA(cudaMemsetAsync(dst, 0, sizeof(cplx) * dim, p->stream));
    QSB_CUDA(cudaMemcpyAsync(dst, psi0_host, sizeof(cplx) * udim, cudaMemcpyHostToDevice, p->stream));
  }
  return QSB_OK;
}

extern "C" {

int qsb_session_begin(qsb_plan* p, uint64_t psi0_index, const double* psi0_host, double t0,
                      const qsb_solver_opts* opts) {
  if (!p) return fail(QSB_ERR_BAD_ARG, "null plan");
  if (p->small) return fail(QSB_ERR_UNSUPPORTED, "sessions are for the tiled path (n_qubits > 10)");
  if (p->world > 1 && !p->connected) return fail(QSB_ERR_BAD_ARG, "sharded plan is not connected to its peers (qsb_shard_connect)");
  int rc = check_opts(opts);
  if (rc) return rc;
  QSB_CUDA(cudaSetDevice(p->device));
  qsb_session_end(p);
  rc = ensure_workspace(p, opts->method);
  if (rc) return rc;
  Session* s = new Session;
  s->o = *opts;
  s->order = (opts->method == QSB_METHOD_TAYLOR) ? (opts->order ? opts->order : 20) : 5;
  s->inv_q = (opts->method == QSB_METHOD_TAYLOR) ? 1.0 / s->order : 0.2;
  s->t = t0;
  s->dim = 1ull << p->n_loc;
  s->norm_dim = std::ldexp(1.0, p->n);
  s->n_tiles = 1ll << (p->n_loc - TB);
  carve(p, s);
  p->ses = s;
  rc = upload_state(p, s->y, psi0_index, psi0_host);
  if (!rc) rc = session_prime(p, s);
  if (rc) { delete s; p->ses = nullptr; }
  return rc;
}

int qsb_session_step(qsb_plan* p, int64_t n_attempts, double* t_now, qsb_stats* stats, double* device_ms) {
  if (!p || !p->ses) return fail(QSB_ERR_BAD_ARG, "no open session");
  if (n_attempts < 0) return fail(QSB_ERR_BAD_ARG, "n_attempts < 0");
  Session* s = p->ses;
  QSB_CUDA(cudaSetDevice(p->device));
  cudaEvent_t e0, e1;
  QSB_CUDA(cudaEventCreate(&e0));
  QSB_CUDA(cudaEventCreate(&e1));
  QSB_CUDA(cudaEventRecord(e0, p->stream));
  int64_t done = 0;
  int rc = session_advance(p, s, INFINITY, n_attempts, 0, &done);
  cudaEventRecord(e1, p->stream);
  cudaEventSynchronize(e1);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (t_now) *t_now = s->t;
  if (device_ms) *device_ms = ms;
  if (stats) {
    stats->n_accepted = s->n_acc; stats->n_rejected = s->n_rej; stats->n_rhs = s->n_rhs;
    stats->last_h = s->h; stats->device_ms = ms; stats->path = 1; stats->status = rc;
  }
  return rc;
}

int qsb_session_advance(qsb_plan* p, double t_target, double* t_now, qsb_stats* stats) {
  if (!p || !p->ses) return fail(QSB_ERR_BAD_ARG, "no open session");
  Session* s = p->ses;
  if (!(t_target >= s->t)) return fail(QSB_ERR_BAD_ARG, "t_target lies before the session's current time");
  QSB_CUDA(cudaSetDevice(p->device));
  int64_t done = 0;
  int rc = session_advance(p, s, t_target, INT64_MAX, s->o.nsteps, &done);
  if (t_now) *t_now = s->t;
  if (stats) {
    stats->n_accepted = s->n_acc; stats->n_rejected = s->n_rej; stats->n_rhs = s->n_rhs;
    stats->last_h = s->h; stats->device_ms = 0.0; stats->path = 1; stats->status = rc;
  }
  return rc;
}

int qsb_session_observe(qsb_plan* p, double* obs) {
  if (!p || !p->ses || !obs) return fail(QSB_ERR_BAD_ARG, "no open session or null obs");
  QSB_CUDA(cudaSetDevice(p->device));
  return session_observe(p, p->ses, obs);
}

int qsb_session_state(qsb_plan* p, double* psi_host) {
  if (!p || !p->ses || !psi_host) return fail(QSB_ERR_BAD_ARG, "no open session or null buffer");
  QSB_CUDA(cudaSetDevice(p->device));
  QSB_CUDA(cudaMemcpyAsync(psi_host, p->ses->y, sizeof(cplx) << (p->world > 1 ? p->n_loc : p->n), cudaMemcpyDeviceToHost,
                           p->stream));
  QSB_CUDA(cudaStreamSynchronize(p->stream));
  return QSB_OK;
}

int qsb_session_end(qsb_plan* p) {
  if (p && p->ses) {
    cudaStreamSynchronize(p->stream);
    delete p->ses;
    p->ses = nullptr;
  }
  return QSB_OK;
}

int qsb_evolve(qsb_plan* p, uint64_t psi0_index, const double* psi0_host, int64_t n_out, const double* t_out,
               const qsb_solver_opts* opts, double* out_obs, qsb_stats* stats) {
  if (!p || !out_obs) return fail(QSB_ERR_BAD_ARG, "null plan or output buffer");
  int rc = check_opts(opts);
  if (rc) return rc;
  rc = check_grid(n_out, t_out);
  if (rc) return rc;
  if (p->small) {
    if (psi0_host) return fail(QSB_ERR_UNSUPPORTED, "the on-chip path starts from a basis state (psi0_index)");
    qsb_table t;
    t.n_qubits = p->n; t.h = p->h.data(); t.J = p->J.data(); t.sea_mask = p->sea_mask; t.rare_mask = p->rare_mask;
    return qsb_evolve_batch(1, &t, &psi0_index, n_out, t_out, 0, opts, out_obs, stats, p->device);
  }
  QSB_CUDA(cudaSetDevice(p->device));
  cudaEvent_t e0, e1;
  QSB_CUDA(cudaEventCreate(&e0));
  QSB_CUDA(cudaEventCreate(&e1));
  QSB_CUDA(cudaEventRecord(e0, p->stream));
  rc = qsb_session_begin(p, psi0_index, psi0_host, t_out[0], opts);
  if (!rc) {
    Session* s = p->ses;
    rc = session_observe(p, s, out_obs);
    const bool dense = dense_setup(p, s);
    int64_t in_interval = 0;
    for (int64_t k = 1; k < n_out && !rc;) {
      const double hmax = s->o.max_step > 0.0 ? std::min(s->h, s->o.max_step) : s->h;
      if (dense && hmax > (t_out[k] - s->t) * (1.0 + 1e-9)) {
        // the grid is finer than the step the controller wants: one unclipped step serves all outputs inside it
        rc = dense_step(p, s, t_out, n_out, &k, out_obs, &in_interval);
        continue;
      }
      int64_t done = 0;
      rc = session_advance(p, s, t_out[k], INT64_MAX, opts->nsteps, &done);
      if (!rc) rc = session_observe(p, s, out_obs + k * QSB_NOBS);
      ++k;
      in_interval = 0;
    }
    cudaEventRecord(e1, p->stream);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (stats) {
      stats->n_accepted = s->n_acc; stats->n_rejected = s->n_rej; stats->n_rhs = s->n_rhs;
      stats->last_h = s->h; stats->device_ms = ms; stats->path = 1; stats->status = rc;
    }
    std::string keep = g_err;
    qsb_session_end(p);
    g_err = keep;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  return rc;
}

int qsb_debug_fp64_peak(int device, double* tflops) {
  if (!tflops) return fail(QSB_ERR_BAD_ARG, "null argument");
  int rc = pick_device(device);
  if (rc) return rc;
  *tflops = measure_fp64_peak_tflops(nullptr);
  QSB_CUDA(cudaGetLastError());
  return QSB_OK;
}

int qsb_profile(int enable) {
  Profile& pr = profile();
  pr.on = enable != 0;
  for (int i = 0; i < N_PROFILE; ++i) { pr.ms[i] = 0.0; pr.count[i] = 0; pr.bytes[i] = 0.0; }
  return QSB_OK;
}

int qsb_profile_read(double* ms, int64_t* count, double* bytes, int32_t n) {
  if (!ms || !count || !bytes || n < 0) return fail(QSB_ERR_BAD_ARG, "null argument");
  for (int i = 0; i < n; ++i) {
    const bool have = i < N_PROFILE;
    ms[i] = have ? profile().ms[i] : 0.0;
    count[i] = have ? profile().count[i] : 0;
    bytes[i] = have ? profile().bytes[i] : 0.0;
  }
  return QSB_OK;
}

int qsb_apply_h(qsb_plan* p, const double* psi_host, double* out_host) {
  if (!p || !psi_host || !out_host) return fail(QSB_ERR_BAD_ARG, "null argument");
  QSB_CUDA(cudaSetDevice(p->device));
  const size_t ubytes = sizeof(cplx) << (p->world > 1 ? p->n_loc : p->n);
  if (p->small) {
    qsb_table t;
    t.n_qubits = p->n; t.h = p->h.data(); t.J = p->J.data(); t.sea_mask = p->sea_mask; t.rare_mask = p->rare_mask;
    qsb_solver_opts o = {1e-8, 1e-6, 0.0, 0, QSB_METHOD_DP5, 0, 0.0};
    SmallDesc d;
    compile_small(&t, 0, &o, 0, &d);
    SmallDesc* dd = nullptr;
    cplx *din = nullptr, *dout = nullptr;
    struct Guard {  // early returns free what has been allocated
      SmallDesc*& a; cplx*& b; cplx*& c; cudaStream_t s;
      ~Guard() { cudaStreamSynchronize(s); cudaFree(a); cudaFree(b); cudaFree(c); }
    } guard{dd, din, dout, p->stream};
    QSB_CUDA(cudaMalloc(&dd, sizeof d));
    QSB_CUDA(cudaMalloc(&din, ubytes));
    QSB_CUDA(cudaMalloc(&dout, ubytes));
    QSB_CUDA(cudaMemcpyAsync(dd, &d, sizeof d, cudaMemcpyHostToDevice, p->stream));
    QSB_CUDA(cudaMemcpyAsync(din, psi_host, ubytes, cudaMemcpyHostToDevice, p->stream));
    launch_small_apply(dd, p->n, din, dout, p->stream);
    QSB_CUDA(cudaGetLastError());
    QSB_CUDA(cudaMemcpyAsync(out_host, dout, ubytes, cudaMemcpyDeviceToHost, p->stream));
    QSB_CUDA(cudaStreamSynchronize(p->stream));
    return QSB_OK;
  }
  if (p->ses) return fail(QSB_ERR_BAD_ARG, "qsb_apply_h cannot run while a session is open (shared workspace)");
  if (p->world > 1 && !p->connected) return fail(QSB_ERR_BAD_ARG, "sharded plan is not connected to its peers (qsb_shard_connect)");
  int rc = ensure_workspace(p, QSB_METHOD_TAYLOR);
  if (rc) return rc;
  Session s;
  s.o = {1e-8, 1e-6, 0.0, 0, QSB_METHOD_TAYLOR, 12, 0.0};
  carve(p, &s);
  rc = upload_state(p, s.y, 0, psi_host);
  if (rc) return rc;
  Epilogue e = blank_epilogue(&s);
  e.raw_h = 1;
  e.o1 = s.ys;
  rc = launch_happly(p->g, p->tab, s.y, s.tx, s.ty, e, 0, s.partials, p->stream, link_of(p), nullptr, &p->fu);
  if (rc) return rc;
  QSB_CUDA(cudaGetLastError());
  QSB_CUDA(cudaMemcpyAsync(out_host, s.ys, ubytes, cudaMemcpyDeviceToHost, p->stream));
  QSB_CUDA(cudaStreamSynchronize(p->stream));
  return QSB_OK;
}


int qsb_apply_h_product(qsb_plan* p, const double* phi, const uint64_t* rows, int64_t n_rows, double* out_host) {
  if (!p || !phi || !rows || !out_host || n_rows < 1) return fail(QSB_ERR_BAD_ARG, "null argument or n_rows < 1");
  if (p->small) return fail(QSB_ERR_UNSUPPORTED, "qsb_apply_h_product is a hook of the tiled path (n_qubits > 10)");
  if (p->ses) return fail(QSB_ERR_BAD_ARG, "qsb_apply_h_product cannot run while a session is open (shared workspace)");
  if (p->world > 1 && !p->connected) return fail(QSB_ERR_BAD_ARG, "sharded plan is not connected to its peers (qsb_shard_connect)");
  const uint64_t dim = 1ull << p->n_loc;
  for (int64_t i = 0; i < n_rows; ++i)
    if (rows[i] >= dim) return fail(QSB_ERR_BAD_ARG, "row index outside this plan's local slice");
  QSB_CUDA(cudaSetDevice(p->device));
  int rc = ensure_workspace(p, QSB_METHOD_TAYLOR);
  if (rc) return rc;
  Session s;
  s.o = {1e-8, 1e-6, 0.0, 0, QSB_METHOD_TAYLOR, 12, 0.0};
  carve(p, &s);
  ProductPhi pp;
  for (int b = 0; b < QSB_MAX_QUBITS; ++b) {  // padded spectator bits (b >= n) stay in |0>
    pp.v[b][0] = b < p->n ? mk(phi[4 * b], phi[4 * b + 1]) : mk(1.0, 0.0);
    pp.v[b][1] = b < p->n ? mk(phi[4 * b + 2], phi[4 * b + 3]) : mk(0.0, 0.0);
  }
  launch_product_state(s.y, pp, p->n_eff, p->n_loc, p->g.rank_bits, p->stream);
  QSB_CUDA(cudaGetLastError());
  Epilogue e = blank_epilogue(&s);
  e.raw_h = 1;
  e.o1 = s.ys;
  rc = launch_happly(p->g, p->tab, s.y, s.tx, s.ty, e, 0, s.partials, p->stream, link_of(p), nullptr, &p->fu);
  if (rc) return rc;
  QSB_CUDA(cudaGetLastError());
  // the running Taylor term's slot is free here: row indices and gathered values live in it
  uint64_t* d_rows = (uint64_t*)s.k[0];
  cplx* d_out = (cplx*)(d_rows + ((n_rows + 1) & ~(int64_t)1));  // 16-byte aligned
  if ((size_t)(n_rows + 1) * (sizeof(uint64_t) + sizeof(cplx)) > (sizeof(cplx) << p->n_loc))
    return fail(QSB_ERR_BAD_ARG, "too many rows for the scratch vector");
  QSB_CUDA(cudaMemcpyAsync(d_rows, rows, sizeof(uint64_t) * n_rows, cudaMemcpyHostToDevice, p->stream));
  launch_gather_rows(s.ys, d_rows, n_rows, d_out, p->stream);
  QSB_CUDA(cudaGetLastError());
  QSB_CUDA(cudaMemcpyAsync(out_host, d_out, sizeof(cplx) * n_rows, cudaMemcpyDeviceToHost, p->stream));
  QSB_CUDA(cudaStreamSynchronize(p->stream));
  return QSB_OK;
}


int qsb_apply_h_static(qsb_plan* p, const double* psi_host, double* out_host) {
  if (!p || !psi_host || !out_host) return fail(QSB_ERR_BAD_ARG, "null argument");
  if (!p->small || p->n > QSB_STATIC_MAX_QUBITS)
    return fail(QSB_ERR_UNSUPPORTED, "qsb_apply_h_static: the static scratch holds at most 2^" + std::to_string(QSB_STATIC_MAX_QUBITS) +
                                         " amplitudes (n_qubits <= " + std::to_string(QSB_STATIC_MAX_QUBITS) + "); this plan has n_qubits = " +
                                         std::to_string(p->n) + " - use qsb_apply_h");
  QSB_CUDA(cudaSetDevice(p->device));
  qsb_table t;
  t.n_qubits = p->n; t.h = p->h.data(); t.J = p->J.data(); t.sea_mask = p->sea_mask; t.rare_mask = p->rare_mask;
  qsb_solver_opts o = {1e-8, 1e-6, 0.0, 0, QSB_METHOD_DP5, 0, 0.0};
  SmallDesc d;
  compile_small(&t, 0, &o, 0, &d);
  // no cudaMalloc / cudaFree on this path: descriptor, input and output go through the library's static device scratch
  QSB_CUDA(launch_small_apply_static(&d, p->n, (const cplx*)psi_host, (cplx*)out_host, p->stream));
  return QSB_OK;
}

}  // extern "C"

#include "qsb_api_prophbm.inl"
