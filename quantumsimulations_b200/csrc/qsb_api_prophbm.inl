// HBM-resident constant-H propagator for 11 <= n <= 14 qubits (qsb_propagate_hbm); included from qsb_api_entry.inl.
//
// Same idea as the on-chip propagator path (qsb_propagate_batch) one size class up.  U = exp(-i H dt) no longer fits an
// SM (1 GB at n = 13), but it fits HBM comfortably, and the pieces are already there:
//   * ALL 2^n columns are integrated at once by the tiled path: the 2^n x 2^n matrix is the state vector of a 2n-qubit
//     plan whose upper n bits are decoupled spectators (zero coefficients), index = column * 2^n + row, started from
//     the identity - the same adaptive Taylor / DP5 session, error control and nsteps / max_step semantics;
//   * scaling and squaring, U^seg_len for the time-split seeds and the chain X <- U X over all segments at once are
//     plain dense complex GEMMs: cuBLAS Zgemm (a library GEMM, nothing fused into it);
//   * the observables of every state of every segment come from observe_columns_kernel.
// A 20000-output trajectory at the sweep's tolerances then costs seconds instead of the hours the stepped path needs
// (||H|| dt_out ~ 10^3 .. 10^4 H-applies per output interval).
// cuBLAS is loaded on first use (dlopen), not linked: only this path needs it, and a box without libcublas on the loader
// path must still be able to load the engine.
typedef cublasStatus_t (*CublasCreateFn)(cublasHandle_t*);
typedef cublasStatus_t (*CublasSetStreamFn)(cublasHandle_t, cudaStream_t);
typedef cublasStatus_t (*CublasZgemmFn)(cublasHandle_t, cublasOperation_t, cublasOperation_t, int, int, int, const cuDoubleComplex*,
                                        const cuDoubleComplex*, int, const cuDoubleComplex*, int, const cuDoubleComplex*,
                                        cuDoubleComplex*, int);
typedef cublasStatus_t (*CublasZgemvFn)(cublasHandle_t, cublasOperation_t, int, int, const cuDoubleComplex*, const cuDoubleComplex*, int,
                                        const cuDoubleComplex*, int, const cuDoubleComplex*, cuDoubleComplex*, int);
struct CublasCtx {
  std::mutex mu;
  std::mutex busy[64];   // one qsb_propagate_hbm at a time per device: the handle and its stream binding are shared
  cublasHandle_t h[64] = {};
  void* lib = nullptr;
  bool tried = false;
  CublasCreateFn create = nullptr;
  CublasSetStreamFn set_stream = nullptr;
  CublasZgemmFn zgemm = nullptr;
  CublasZgemmFn zgemm3m = nullptr;  // Gauss's three-multiplication complex GEMM (25 % fewer flops); optional
  CublasZgemvFn zgemv = nullptr;
};
static CublasCtx& cublas_ctx() {
  static CublasCtx c;
  return c;
}
static int cublas_for(int device, cublasHandle_t* out) {
  CublasCtx& c = cublas_ctx();
  std::lock_guard<std::mutex> lk(c.mu);
  if (!c.tried) {
    c.tried = true;
    for (const char* name : {"libcublas.so.12", "libcublas.so"}) {
      c.lib = dlopen(name, RTLD_NOW | RTLD_LOCAL);
      if (c.lib) break;
    }
    if (c.lib) {
      c.create = (CublasCreateFn)dlsym(c.lib, "cublasCreate_v2");
      c.set_stream = (CublasSetStreamFn)dlsym(c.lib, "cublasSetStream_v2");
      c.zgemm = (CublasZgemmFn)dlsym(c.lib, "cublasZgemm_v2");
      c.zgemv = (CublasZgemvFn)dlsym(c.lib, "cublasZgemv_v2");
      const char* e3 = getenv("QSB200_HBM_GEMM3M");
      if (!(e3 && *e3 == '0')) c.zgemm3m = (CublasZgemmFn)dlsym(c.lib, "cublasZgemm3m");
    }
  }
  if (!c.lib || !c.create || !c.set_stream || !c.zgemm || !c.zgemv)
    return fail(QSB_ERR_UNSUPPORTED, "qsb_propagate_hbm needs cuBLAS (libcublas.so.12 could not be loaded)");
  if (!c.h[device & 63] && c.create(&c.h[device & 63]) != CUBLAS_STATUS_SUCCESS) {
    c.h[device & 63] = nullptr;
    return fail(QSB_ERR_CUDA, "cublasCreate failed");
  }
  *out = c.h[device & 63];
  return QSB_OK;
}

static int stage_terms(const qsb_pauli_term* terms, int32_t n_terms, int n_qubits, std::vector<ExpectTerm>* out,
                       int* column_begin);  // qsb_api_expect.inl

// terms == nullptr: the six mask-based observables (qsb_propagate_hbm); else weighted Pauli strings reduced by
// launch_expect on the same chain buffer (qsb_propagate_hbm_expect: a local-dimension-4 table, qsb_api_expect.inl).
static int propagate_hbm_impl(const qsb_table* table, uint64_t psi0_index, int64_t n_out, double dt,
                              const qsb_solver_opts* opts, int32_t max_squarings, int32_t obs_by_column, double* out_obs,
                              qsb_stats* stats, int device, const qsb_pauli_term* terms, int32_t n_terms) {
  if (!table || !opts || !out_obs || n_out < 1) return fail(QSB_ERR_BAD_ARG, "qsb_propagate_hbm: null argument or n_out < 1");
  int rc = check_table(table);
  if (rc) return rc;
  rc = check_opts(opts);
  if (rc) return rc;
  const int n = table->n_qubits;
  std::vector<ExpectTerm> staged;
  int column_begin[7] = {};
  if (terms && (rc = stage_terms(terms, n_terms, n, &staged, column_begin))) return rc;
  if (n <= QSB_SMALL_MAX_QUBITS || n > QSB_HBM_PROPAGATOR_MAX_QUBITS)
    return fail(QSB_ERR_UNSUPPORTED, "qsb_propagate_hbm handles " + std::to_string(QSB_SMALL_MAX_QUBITS + 1) + " <= n_qubits <= " +
                                         std::to_string(QSB_HBM_PROPAGATOR_MAX_QUBITS) + " (smaller: qsb_propagate_batch)");
  if (psi0_index >> n) return fail(QSB_ERR_BAD_ARG, "psi0_index out of range");
  if (!(dt > 0.0) || !std::isfinite(dt)) return fail(QSB_ERR_BAD_ARG, "dt must be finite and > 0");
  const size_t dim = (size_t)1 << n;

  // squarings: integrate over dt / 2^s with ||H|| dt / 2^s <= ~8 (table_norm_bound overestimates ||H||).  Measured at
  // n = 13: a squaring is a 2^(3n+3)-flop GEMM, 0.13 s; integrating over dt / 256 takes 0.88 s and halves with every
  // further squaring, so the total is smallest two squarings beyond the <= 32 rule used first (1.5 s against 1.9 s)
  int s_sq = 0;
  if (max_squarings != 0) {
    const double x = table_norm_bound(table) * dt / 8.0;
    while (s_sq < 40 && std::ldexp(1.0, s_sq) < x) ++s_sq;
    if (max_squarings > 0 && s_sq > max_squarings) s_sq = max_squarings;
  }
  // the 2n-qubit table: the upper n bits carry no term at all
  const int n2 = 2 * n;
  std::vector<double> h2(3 * (size_t)n2, 0.0), J2(3 * (size_t)n2 * n2, 0.0);
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < n; ++b) {
      h2[(size_t)a * n2 + b] = table->h[a * n + b];
      for (int c = 0; c < n; ++c)
        if (b != c) J2[((size_t)a * n2 + b) * n2 + c] = table->J[((size_t)a * n + b) * n + c];
    }
  qsb_table t2 = {n2, h2.data(), J2.data(), table->sea_mask, table->rare_mask};
  qsb_plan* bp = nullptr;
  rc = plan_create(&t2, device, 0, 1, &bp);
  if (rc) return rc;
  struct PlanGuard { qsb_plan* p; ~PlanGuard() { qsb_plan_destroy(p); } } guard{bp};
  bp->identity_bits = n;
  std::lock_guard<std::mutex> one_at_a_time(cublas_ctx().busy[device & 63]);
  cublasHandle_t cb = nullptr;
  rc = cublas_for(device, &cb);
  if (rc) return rc;
  if (cublas_ctx().set_stream(cb, bp->stream) != CUBLAS_STATUS_SUCCESS) return fail(QSB_ERR_CUDA, "cublasSetStream failed");
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  QSB_CUDA(cudaEventCreate(&e0));
  QSB_CUDA(cudaEventCreate(&e1));
  struct EvGuard { cudaEvent_t &a, &b; ~EvGuard() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); } } evg{e0, e1};
  QSB_CUDA(cudaEventRecord(e0, bp->stream));

  // ---- columns: one session over the whole matrix ----------------------------------------------------------------
  rc = qsb_session_begin(bp, 0, nullptr, 0.0, opts);
  if (rc) return rc;
  Session* s = bp->ses;
  int64_t done = 0;
  rc = session_advance(bp, s, std::ldexp(dt, -s_sq), INT64_MAX, opts->nsteps, &done);
  if (rc) return rc;
  // the session's five vectors double as scratch from here on (the plan owns them)
  cuDoubleComplex* U = (cuDoubleComplex*)s->y;
  cuDoubleComplex* T1 = (cuDoubleComplex*)s->ys;
  cuDoubleComplex* V = (cuDoubleComplex*)s->k[0];
  cuDoubleComplex* T2 = (cuDoubleComplex*)s->tx;
  char* misc = (char*)s->ty;
  const cuDoubleComplex one = make_cuDoubleComplex(1.0, 0.0), zero = make_cuDoubleComplex(0.0, 0.0);
  auto gemm = [&](const cuDoubleComplex* A, const cuDoubleComplex* B, cuDoubleComplex* C, int cols) -> int {
    // square products (squarings) take the three-multiplication form when cuBLAS has it; the skinny chain GEMMs do not
    CublasZgemmFn f = (cols == (int)dim && cublas_ctx().zgemm3m) ? cublas_ctx().zgemm3m : cublas_ctx().zgemm;
    if (f(cb, CUBLAS_OP_N, CUBLAS_OP_N, (int)dim, cols, (int)dim, &one, A, (int)dim, B, (int)dim, &zero, C, (int)dim) != CUBLAS_STATUS_SUCCESS)
      return fail(QSB_ERR_CUDA, "cublasZgemm failed");
    return QSB_OK;
  };
  for (int i = 0; i < s_sq; ++i) {
    if ((rc = gemm(U, U, T1, (int)dim))) return rc;
    std::swap(U, T1);
  }
  // ---- time-split chain: segments of seg_len outputs, all advanced together ---------------------------------------
  int seg_lb = 0, n_seg = 1;
  int64_t seg_len = n_out;
  if (n_out >= 256) {
    seg_lb = 6;
    while (((int64_t)1 << seg_lb) * 128 < n_out) ++seg_lb;
    seg_len = (int64_t)1 << seg_lb;
    n_seg = (int)((n_out + seg_len - 1) / seg_len);
  }
  const size_t xbytes = sizeof(cplx) * dim * (size_t)n_seg;
  cuDoubleComplex* X0 = (cuDoubleComplex*)misc;
  cuDoubleComplex* X1 = (cuDoubleComplex*)(misc + ((xbytes + 255) & ~(size_t)255));
  double* d_obs = (double*)(misc + 2 * ((xbytes + 255) & ~(size_t)255));
  const size_t obs_bytes = sizeof(double) * (size_t)n_out * QSB_NOBS;
  if (2 * ((xbytes + 255) & ~(size_t)255) + obs_bytes > (sizeof(cplx) << n2))
    return fail(QSB_ERR_UNSUPPORTED, "too many outputs for the scratch vector of the HBM propagator path");
  QSB_CUDA(cudaMemsetAsync(X0, 0, xbytes, bp->stream));
  launch_set_basis_state((cplx*)X0, dim, psi0_index, bp->stream);  // column 0 = the initial basis state
  if (n_seg > 1) {
    QSB_CUDA(cudaMemcpyAsync(V, U, sizeof(cplx) << n2, cudaMemcpyDeviceToDevice, bp->stream));
    cuDoubleComplex* Vt = T2;
    for (int i = 0; i < seg_lb; ++i) {  // V = U^seg_len
      if ((rc = gemm(V, V, Vt, (int)dim))) return rc;
      std::swap(V, Vt);
    }
    for (int c = 1; c < n_seg; ++c)  // seeds: column c = V column c-1
      if (cublas_ctx().zgemv(cb, CUBLAS_OP_N, (int)dim, (int)dim, &one, V, (int)dim, X0 + (size_t)(c - 1) * dim, 1, &zero,
                      X0 + (size_t)c * dim, 1) != CUBLAS_STATUS_SUCCESS)
        return fail(QSB_ERR_CUDA, "cublasZgemv failed");
  }
  QSB_CUDA(cudaGetLastError());
  // Pauli-string observables: the terms and the per-CTA partial sums live in one small allocation of their own
  char* d_expect = nullptr;
  struct ExpectFree { char*& p; cudaStream_t s; ~ExpectFree() { if (p) { cudaStreamSynchronize(s); cudaFree(p); } } } xfree{d_expect, bp->stream};
  ExpectTerm* d_terms = nullptr;
  double* d_partials = nullptr;
  if (terms) {
    const size_t tb = (sizeof(ExpectTerm) * staged.size() + 255) & ~size_t(255);
    cudaError_t e = cudaMalloc((void**)&d_expect, tb + sizeof(double) * 8 * (size_t)expect_blocks(dim) * (size_t)n_seg);
    if (e != cudaSuccess) {
      cudaGetLastError();
      d_expect = nullptr;
      return fail(QSB_ERR_OOM, std::string("qsb_propagate_hbm_expect: cudaMalloc failed: ") + cudaGetErrorString(e));
    }
    d_terms = (ExpectTerm*)d_expect;
    d_partials = (double*)(d_expect + tb);
    QSB_CUDA(cudaMemcpyAsync(d_terms, staged.data(), sizeof(ExpectTerm) * staged.size(), cudaMemcpyHostToDevice, bp->stream));
  }
  for (int64_t k = 0; k < seg_len; ++k) {
    if (terms) {  // column c of X0 is output c seg_len + k; the last segment may be shorter
      const int64_t valid = k < n_out ? (n_out - k + seg_len - 1) / seg_len : 0;
      if (valid > 0) {
        if (obs_by_column) launch_expect((const cplx*)X0, valid, dim, d_terms, (int)staged.size(), column_begin, d_partials,
                                         d_obs + k, seg_len, n_out, bp->stream);
        else launch_expect((const cplx*)X0, valid, dim, d_terms, (int)staged.size(), column_begin, d_partials,
                           d_obs + k * QSB_NOBS, seg_len * QSB_NOBS, 1, bp->stream);
      }
    } else {
      launch_observe_columns((const cplx*)X0, n, n_seg, seg_len, k, n_out, table->sea_mask, table->rare_mask, d_obs,
                             obs_by_column != 0, bp->stream);
    }
    if (k + 1 < seg_len) {
      if ((rc = gemm(U, X0, X1, n_seg))) return rc;
      std::swap(X0, X1);
    }
  }
  QSB_CUDA(cudaGetLastError());
  QSB_CUDA(cudaEventRecord(e1, bp->stream));
  QSB_CUDA(cudaMemcpyAsync(out_obs, d_obs, obs_bytes, cudaMemcpyDeviceToHost, bp->stream));
  QSB_CUDA(cudaStreamSynchronize(bp->stream));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  if (stats) {
    stats->n_accepted = s->n_acc; stats->n_rejected = s->n_rej;
    stats->n_rhs = s->n_rhs * (int64_t)dim;   // column H-applies, comparable with qsb_propagate_batch's counter
    stats->last_h = s->h; stats->device_ms = ms; stats->path = 3; stats->status = QSB_OK;
  }
  return QSB_OK;
}

extern "C" int qsb_propagate_hbm(const qsb_table* table, uint64_t psi0_index, int64_t n_out, double dt,
                                 const qsb_solver_opts* opts, int32_t max_squarings, int32_t obs_by_column, double* out_obs,
                                 qsb_stats* stats, int device) {
  return propagate_hbm_impl(table, psi0_index, n_out, dt, opts, max_squarings, obs_by_column, out_obs, stats, device, nullptr, 0);
}

extern "C" int qsb_propagate_hbm_expect(const qsb_table* table, uint64_t psi0_index, int64_t n_out, double dt,
                                        const qsb_solver_opts* opts, int32_t max_squarings, const qsb_pauli_term* terms,
                                        int32_t n_terms, int32_t obs_by_column, double* out_obs, qsb_stats* stats, int device) {
  if (!terms) return fail(QSB_ERR_BAD_ARG, "qsb_propagate_hbm_expect: terms is null");
  return propagate_hbm_impl(table, psi0_index, n_out, dt, opts, max_squarings, obs_by_column, out_obs, stats, device, terms, n_terms);
}
