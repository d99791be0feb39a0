// Local dimension 4 (the reference's spin-3/2 rare spin, dipolar_ensemble_with_rare.py:22-25 / :499-501) on two index
// bits: trajectories whose observables are weighted sums of Pauli strings.  Included from qsb_api.cu.
//
// With m = 2 a + b the level index of the four-level site (m = 0 is J_z = +3/2), J_z = 2 I^z_a + I^z_b and
// J_x = sqrt(3) I^x_b + 2 (I^x_a I^x_b + I^y_a I^y_b): in the frame rotating about J_z by the drive phase the whole
// Hamiltonian stays inside the same-axis coupling table the integrator kernels take, so the time evolution runs through
// the existing kernels unchanged.  What they cannot produce is <J_x>, <J_y> (two-bit strings, one of them mixed-axis):
// the state-storing matvec chain of the propagator path (n <= 10) or a session's vector (n > 10) is handed to
// launch_expect (qsb_expect.cu) instead of the fused mask-based observable sums.

static int stage_terms(const qsb_pauli_term* terms, int32_t n_terms, int n_qubits, std::vector<ExpectTerm>* out,
                       int* column_begin /* [7] */) {
  if (!terms || n_terms < 1 || n_terms > EXPECT_MAX_TERMS)
    return fail(QSB_ERR_BAD_ARG, "need 1 .. " + std::to_string(EXPECT_MAX_TERMS) + " Pauli terms");
  const uint64_t all = (n_qubits >= 64) ? ~0ull : ((1ull << n_qubits) - 1);
  for (int c = 0; c < 7; ++c) column_begin[c] = 0;
  for (int t = 0; t < n_terms; ++t) {
    const qsb_pauli_term& e = terms[t];
    if ((e.x_mask | e.z_mask) & ~all) return fail(QSB_ERR_BAD_ARG, "Pauli term " + std::to_string(t) + " touches bits above n_qubits");
    if (e.column < 0 || e.column > 5) return fail(QSB_ERR_BAD_ARG, "Pauli term " + std::to_string(t) + ": column must be 0 .. 5");
    if (!std::isfinite(e.weight)) return fail(QSB_ERR_BAD_ARG, "Pauli term " + std::to_string(t) + ": non-finite weight");
    ++column_begin[e.column + 1];
  }
  for (int c = 0; c < 6; ++c) column_begin[c + 1] += column_begin[c];
  out->resize((size_t)n_terms);
  int fill[6];
  for (int c = 0; c < 6; ++c) fill[c] = column_begin[c];
  for (int t = 0; t < n_terms; ++t) {  // counting sort by column: terms of one column keep the caller's order
    const qsb_pauli_term& e = terms[t];
    ExpectTerm& d = (*out)[(size_t)fill[e.column]++];
    d.x_mask = e.x_mask; d.z_mask = e.z_mask; d.weight = e.weight; d.column = e.column;
    d.ny = __builtin_popcountll(e.x_mask & e.z_mask) & 3;
  }
  return QSB_OK;
}

namespace {
struct ExpectScratch {  // device buffers, events and the stream of one qsb_propagate_expect call
  cudaStream_t st = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  std::vector<void*> bufs;
  int get(size_t bytes, void** out) {
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes ? bytes : 16);
    if (e != cudaSuccess) {
      cudaGetLastError();
      return fail(QSB_ERR_OOM, "cudaMalloc of " + std::to_string(bytes) + " bytes failed: " + cudaGetErrorString(e));
    }
    bufs.push_back(p);
    *out = p;
    return QSB_OK;
  }
  ~ExpectScratch() {
    if (st) cudaStreamSynchronize(st);
    for (void* p : bufs) cudaFree(p);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (st) cudaStreamDestroy(st);
  }
};
}  // namespace

// Shared body of qsb_propagate_expect / qsb_propagate_expect_batch: n_traj trajectories of equal n_qubits that share one
// list of Pauli strings (a sweep: same n_sea, same phi_R), chunked so that the kept states of one pass stay below ~8 GiB.
static int propagate_expect_impl(int32_t n_traj, const qsb_table* tables, const uint64_t* psi0_index, int64_t n_out,
                                 const double* dt, int32_t dt_per_traj, const qsb_solver_opts* opts, int32_t max_squarings,
                                 const qsb_pauli_term* terms, int32_t n_terms, int32_t obs_by_column, double* out_obs,
                                 qsb_stats* stats, int device) {
  if (n_traj < 1 || !tables || !psi0_index || !opts || !out_obs || !dt || n_out < 1)
    return fail(QSB_ERR_BAD_ARG, "qsb_propagate_expect: null argument, n_traj < 1 or n_out < 1");
  const int n = tables[0].n_qubits;
  int rc;
  for (int i = 0; i < n_traj; ++i) {
    if ((rc = check_table(&tables[i]))) return rc;
    if (tables[i].n_qubits > QSB_SMALL_MAX_QUBITS)
      return fail(QSB_ERR_UNSUPPORTED, "qsb_propagate_expect handles n_qubits <= " + std::to_string(QSB_SMALL_MAX_QUBITS) +
                                           " (11 .. 14: qsb_propagate_hbm_expect; any size: a session with qsb_session_expect)");
    if (tables[i].n_qubits != n) return fail(QSB_ERR_BAD_ARG, "qsb_propagate_expect_batch needs the same n_qubits in every table");
    if ((rc = check_opts(&opts[i]))) return rc;
    if (psi0_index[i] >> n) return fail(QSB_ERR_BAD_ARG, "psi0_index out of range");
    const double h = dt[dt_per_traj ? i : 0];
    if (!(h > 0.0) || !std::isfinite(h)) return fail(QSB_ERR_BAD_ARG, "dt must be finite and > 0");
  }
  std::vector<ExpectTerm> staged;
  int column_begin[7];
  if ((rc = stage_terms(terms, n_terms, n, &staged, column_begin))) return rc;
  const size_t dim = (size_t)1 << n;
  const int nb = expect_blocks(dim);
  const double state_bytes = (double)n_out * (double)dim * sizeof(cplx);
  if (state_bytes > 16.0 * 1073741824.0 || (double)n_out * nb > 2.0e9)
    return fail(QSB_ERR_UNSUPPORTED, "qsb_propagate_expect keeps every output state in device memory: n_out * 2^n * 16 bytes must stay below 16 GiB");
  if ((rc = pick_device(device))) return rc;
  const size_t u_bytes = sizeof(cplx) * dim * dim;
  // trajectories per pass: kept states within ~8 GiB (one trajectory may take up to 16), propagators within ~2 GiB
  int chunk = (int)std::max(1.0, std::min((double)n_traj, 8.0 * 1073741824.0 / state_bytes));
  chunk = (int)std::max<size_t>(1, std::min<size_t>((size_t)chunk, ((size_t)2 << 30) / u_bytes));
  if (const char* e = getenv("QSB200_PROP_CHUNK")) chunk = std::max(1, std::min(chunk, atoi(e)));  // test knob

  // Time-split chain, exactly as in propagate_impl: ~128 segments of 2^seg_lb outputs (a function of n_out only), each
  // its own CTA, seeded with (U^(2^seg_lb))^c psi_0 from seg_lb more squarings of a copy of U.
  int seg_lb = 0, n_seg = 1;
  int64_t seg_len = n_out;
  if (n_out >= 256) {
    seg_lb = 6;
    while (((int64_t)1 << seg_lb) * 128 < n_out) ++seg_lb;
    if (const char* e = getenv("QSB200_PROP_SEGMENT")) seg_lb = std::max(0, std::min(40, atoi(e)));  // 0: no split
    if (seg_lb > 0 && ((int64_t)1 << seg_lb) < n_out) {
      seg_len = (int64_t)1 << seg_lb;
      n_seg = (int)((n_out + seg_len - 1) / seg_len);
    } else {
      seg_lb = 0;
    }
  }
  std::vector<int32_t> n_sq(n_traj, 0);
  bool any_sq = false;
  for (int i = 0; i < n_traj; ++i) {
    int s = max_squarings == 0 ? 0 : auto_squarings(&tables[i], dt[dt_per_traj ? i : 0]);
    if (max_squarings > 0 && s > max_squarings) s = max_squarings;
    n_sq[i] = s;
    any_sq = any_sq || s > 0;
  }

  std::vector<SmallStats> hs((size_t)chunk * dim);  // declared before `sc`: its destructor drains the stream that copies into it
  std::vector<SmallDesc> descs(chunk);
  std::vector<double> tp(2 * (size_t)chunk);
  std::vector<int32_t> seg_sq(chunk, seg_lb);
  ExpectScratch sc;
  QSB_CUDA(cudaStreamCreateWithFlags(&sc.st, cudaStreamNonBlocking));
  QSB_CUDA(cudaEventCreate(&sc.e0));
  QSB_CUDA(cudaEventCreate(&sc.e1));
  cudaStream_t st = sc.st;
  SmallDesc* d_desc = nullptr;
  double *d_t = nullptr, *d_partials = nullptr, *d_obs = nullptr;
  SmallStats* d_stats = nullptr;
  cplx *d_u = nullptr, *d_tmp = nullptr, *d_states = nullptr, *d_v = nullptr, *d_seeds = nullptr;
  int32_t *d_sq = nullptr, *d_seg_sq = nullptr;
  ExpectTerm* d_terms = nullptr;
  if ((rc = sc.get(sizeof(SmallDesc) * chunk, (void**)&d_desc))) return rc;
  if ((rc = sc.get(sizeof(double) * 2 * chunk, (void**)&d_t))) return rc;
  if ((rc = sc.get(sizeof(SmallStats) * dim * chunk, (void**)&d_stats))) return rc;
  if ((rc = sc.get(u_bytes * chunk, (void**)&d_u))) return rc;
  if (any_sq || n_seg > 1) {
    if ((rc = sc.get(u_bytes * chunk, (void**)&d_tmp))) return rc;
    if ((rc = sc.get(sizeof(int32_t) * chunk, (void**)&d_sq))) return rc;
  }
  if ((rc = sc.get(sizeof(cplx) * dim * (size_t)n_out * chunk, (void**)&d_states))) return rc;
  if ((rc = sc.get(sizeof(ExpectTerm) * staged.size(), (void**)&d_terms))) return rc;
  if ((rc = sc.get(sizeof(double) * 8 * (size_t)nb * (size_t)n_out * chunk, (void**)&d_partials))) return rc;
  if ((rc = sc.get(sizeof(double) * QSB_NOBS * (size_t)n_out * chunk, (void**)&d_obs))) return rc;
  if (n_seg > 1) {
    if ((rc = sc.get(u_bytes * chunk, (void**)&d_v))) return rc;
    if ((rc = sc.get(sizeof(cplx) * dim * (size_t)n_seg * chunk, (void**)&d_seeds))) return rc;
    if ((rc = sc.get(sizeof(int32_t) * chunk, (void**)&d_seg_sq))) return rc;
    QSB_CUDA(cudaMemcpyAsync(d_seg_sq, seg_sq.data(), sizeof(int32_t) * chunk, cudaMemcpyHostToDevice, st));
  }
  QSB_CUDA(cudaMemcpyAsync(d_terms, staged.data(), sizeof(ExpectTerm) * staged.size(), cudaMemcpyHostToDevice, st));

  int worst = QSB_OK;
  for (int base = 0; base < n_traj; base += chunk) {
    const int m = std::min(chunk, n_traj - base);
    for (int i = 0; i < m; ++i) {
      compile_small(&tables[base + i], psi0_index[base + i], &opts[base + i], 1, &descs[i]);
      tp[2 * i] = 0.0;
      tp[2 * i + 1] = std::ldexp(dt[dt_per_traj ? base + i : 0], -n_sq[base + i]);
    }
    if (any_sq) QSB_CUDA(cudaMemcpyAsync(d_sq, n_sq.data() + base, sizeof(int32_t) * m, cudaMemcpyHostToDevice, st));
    QSB_CUDA(cudaMemcpyAsync(d_desc, descs.data(), sizeof(SmallDesc) * m, cudaMemcpyHostToDevice, st));
    QSB_CUDA(cudaMemcpyAsync(d_t, tp.data(), sizeof(double) * 2 * m, cudaMemcpyHostToDevice, st));
    QSB_CUDA(cudaEventRecord(sc.e0, st));
    launch_small_columns(d_desc, m, n, d_t, d_stats, d_u, st);  // U^(1 / 2^s), one CTA per column
    QSB_CUDA(cudaGetLastError());
    if (any_sq) {
      launch_small_square(d_u, d_tmp, d_sq, m, n, st);
      QSB_CUDA(cudaGetLastError());
    }
    if (n_seg > 1) {
      QSB_CUDA(cudaMemcpyAsync(d_v, d_u, u_bytes * m, cudaMemcpyDeviceToDevice, st));
      launch_small_square(d_v, d_tmp, d_seg_sq, m, n, st);                                               // V = U^seg_len
      QSB_CUDA(cudaGetLastError());
      rc = launch_small_propagate(d_desc, m, n, d_v, n_seg, nullptr, 0, st, nullptr, 1, n_seg, d_seeds);  // seeds
      if (rc) return rc;
      rc = launch_small_propagate(d_desc, m, n, d_u, n_out, nullptr, 0, st, d_seeds, n_seg, seg_len, d_states);  // every psi_k, kept
    } else {
      rc = launch_small_propagate(d_desc, m, n, d_u, n_out, nullptr, 0, st, nullptr, 1, n_out, d_states);  // psi_k = U psi_{k-1}, kept
    }
    if (rc) return rc;
    if (obs_by_column) {  // [traj][QSB_NOBS][n_out]: one reduction per trajectory
      for (int i = 0; i < m; ++i)
        launch_expect(d_states + (size_t)i * n_out * dim, n_out, dim, d_terms, (int)staged.size(), column_begin,
                      d_partials + (size_t)i * n_out * nb * 8, d_obs + (size_t)i * n_out * QSB_NOBS, 1, n_out, st);
    } else {              // [traj][n_out][QSB_NOBS] is one run of states
      launch_expect(d_states, (int64_t)m * n_out, dim, d_terms, (int)staged.size(), column_begin, d_partials, d_obs, QSB_NOBS, 1, st);
    }
    QSB_CUDA(cudaGetLastError());
    QSB_CUDA(cudaEventRecord(sc.e1, st));
    QSB_CUDA(cudaMemcpyAsync(hs.data(), d_stats, sizeof(SmallStats) * dim * m, cudaMemcpyDeviceToHost, st));
    QSB_CUDA(cudaMemcpyAsync(out_obs + (size_t)base * n_out * QSB_NOBS, d_obs, sizeof(double) * QSB_NOBS * (size_t)n_out * m,
                             cudaMemcpyDeviceToHost, st));
    QSB_CUDA(cudaStreamSynchronize(st));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, sc.e0, sc.e1);
    for (int i = 0; i < m; ++i) {
      qsb_stats acc = {0, 0, 0, 0.0, (double)ms, 4, QSB_OK};
      for (size_t c = 0; c < dim; ++c) {
        const SmallStats& s = hs[(size_t)i * dim + c];
        acc.n_accepted += s.n_accepted; acc.n_rejected += s.n_rejected; acc.n_rhs += s.n_rhs;
        acc.last_h = s.last_h;
        if (s.status != QSB_OK && acc.status == QSB_OK) acc.status = s.status;
      }
      if (stats) stats[base + i] = acc;
      if (acc.status != QSB_OK && worst == QSB_OK) {
        worst = acc.status;
        g_err = "propagator of trajectory " + std::to_string(base + i) + (acc.status == QSB_ERR_NSTEPS
                    ? ": more than nsteps internal steps inside one output interval (excess work)"
                    : ": step size underflow");
      }
    }
  }
  return worst;
}

extern "C" {

int qsb_propagate_expect(const qsb_table* table, uint64_t psi0_index, int64_t n_out, double dt, const qsb_solver_opts* opts,
                         int32_t max_squarings, const qsb_pauli_term* terms, int32_t n_terms, int32_t obs_by_column,
                         double* out_obs, qsb_stats* stats, int device) {
  return propagate_expect_impl(1, table, &psi0_index, n_out, &dt, 0, opts, max_squarings, terms, n_terms, obs_by_column, out_obs,
                               stats, device);
}

int qsb_propagate_expect_batch(int32_t n_traj, const qsb_table* tables, const uint64_t* psi0_index, int64_t n_out,
                               const double* dt, int32_t dt_per_traj, const qsb_solver_opts* opts, int32_t max_squarings,
                               const qsb_pauli_term* terms, int32_t n_terms, int32_t obs_by_column, double* out_obs,
                               qsb_stats* stats, int device) {
  return propagate_expect_impl(n_traj, tables, psi0_index, n_out, dt, dt_per_traj, opts, max_squarings, terms, n_terms,
                               obs_by_column, out_obs, stats, device);
}

int qsb_session_expect(qsb_plan* p, const qsb_pauli_term* terms, int32_t n_terms, double* obs) {
  if (!p || !p->ses || !obs) return fail(QSB_ERR_BAD_ARG, "no open session or null obs");
  if (p->world > 1) return fail(QSB_ERR_UNSUPPORTED, "qsb_session_expect: not available on sharded plans");
  std::vector<ExpectTerm> staged;
  int column_begin[7];
  int rc = stage_terms(terms, n_terms, p->n, &staged, column_begin);
  if (rc) return rc;
  QSB_CUDA(cudaSetDevice(p->device));
  const uint64_t dim = 1ull << p->n;  // amplitudes above 2^n (spectator padding of plans below 12 qubits) are zero
  const int nb = expect_blocks(dim);
  const size_t terms_bytes = (sizeof(ExpectTerm) * staged.size() + 255) & ~size_t(255);
  const size_t part_bytes = sizeof(double) * 8 * (size_t)nb;
  char* d_all = nullptr;
  cudaError_t e = cudaMalloc((void**)&d_all, terms_bytes + part_bytes + sizeof(double) * QSB_NOBS);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(QSB_ERR_OOM, std::string("qsb_session_expect: cudaMalloc failed: ") + cudaGetErrorString(e));
  }
  struct Free { char* p; cudaStream_t s; ~Free() { cudaStreamSynchronize(s); cudaFree(p); } } guard{d_all, p->stream};
  ExpectTerm* d_terms = (ExpectTerm*)d_all;
  double* d_partials = (double*)(d_all + terms_bytes);
  double* d_out = (double*)(d_all + terms_bytes + part_bytes);
  QSB_CUDA(cudaMemcpyAsync(d_terms, staged.data(), sizeof(ExpectTerm) * staged.size(), cudaMemcpyHostToDevice, p->stream));
  launch_expect(p->ses->y, 1, dim, d_terms, (int)staged.size(), column_begin, d_partials, d_out, QSB_NOBS, 1, p->stream);
  QSB_CUDA(cudaGetLastError());
  QSB_CUDA(cudaMemcpyAsync(obs, d_out, sizeof(double) * QSB_NOBS, cudaMemcpyDeviceToHost, p->stream));
  QSB_CUDA(cudaStreamSynchronize(p->stream));
  return QSB_OK;
}

}  // extern "C"
