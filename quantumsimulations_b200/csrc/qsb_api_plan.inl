// Plans, the batch path and library-level entry points; included from qsb_api.cu.
using namespace qsb;

struct Session;

struct qsb_plan {
  int n = 0;        // user qubits
  int n_eff = 0;    // qubits of the tiled path (n padded to >= 12 with decoupled spectator bits)
  int n_loc = 0;    // index bits of the local shard (n_eff when unsharded)
  int rank = 0, world = 1;
  int device = 0;
  bool small = false;
  std::vector<double> h, J;  // [3][n], [3][n][n] as given
  uint64_t sea_mask = 0, rare_mask = 0;
  cudaStream_t stream = nullptr;
  // tiled path
  LargeGeom g = {};
  DiagTables tab = {};
  double* d_tables = nullptr;   // h, J and the factored in-tile diagonal tables in one allocation
  double* d_tiles = nullptr;    // per-tile diagonal rows Tx, Ty, Tz
  void* ws = nullptr;           // workspace (vectors + partials)
  size_t ws_bytes = 0;
  bool ws_owned = false;
  double* d_red = nullptr;      // [PSTRIDE] reduced sums
  double* h_red = nullptr;      // pinned mirror
  Session* ses = nullptr;
  int identity_bits = 0;        // > 0: a session's default initial state is the 2^bits x 2^bits identity (qsb_propagate_hbm)
  void* dense_buf = nullptr;    // `order` vectors of kept Taylor terms (dense output, qsb_evolve); grown on demand
  size_t dense_bytes = 0;
  StridedFusion fu;             // L2-resident fusion of the strided passes (unsharded plans); fu.ctrl is owned here
  // sharded state (world > 1): exchange scratch (tx, ty) the peers map, and the peers' mappings
  void* xbuf = nullptr;
  int xvecs = 2;                // vectors in xbuf: tx, ty [, rx, ry]
  ShardLink link = {};
  bool connected = false;
  void* ipc_opened[MAX_WORLD] = {};
  qsb_allgather_fn allgather = nullptr;
  void* allgather_ctx = nullptr;
};

static int n_vectors(int method) { return method == QSB_METHOD_TAYLOR ? 5 : 11; }

static size_t workspace_need(const qsb_plan* p, int method) {
  if (p->small) return 0;
  size_t vec = sizeof(cplx) << p->n_loc;
  size_t partials = sizeof(double) * PSTRIDE * ((size_t)1 << (p->n_loc - TB));
  // a sharded plan keeps tx/ty in its own exchange allocation (the peers map it), not in the workspace
  return (n_vectors(method) - (p->world > 1 ? 2 : 0)) * vec + ((partials + 255) & ~size_t(255));
}

static int ensure_workspace(qsb_plan* p, int method) {
  size_t need = workspace_need(p, method);
  if (p->ws && p->ws_bytes >= need) return QSB_OK;
  if (p->ws && !p->ws_owned)
    return fail(QSB_ERR_BAD_ARG, "caller-provided workspace is too small: need " + std::to_string(need) + " bytes");
  if (p->ws) { cudaFree(p->ws); p->ws = nullptr; }
  cudaError_t e = cudaMalloc(&p->ws, need);
  if (e != cudaSuccess) {
    p->ws = nullptr; p->ws_bytes = 0;
    cudaGetLastError();
    return fail(QSB_ERR_OOM, "cudaMalloc of " + std::to_string(need) + " workspace bytes failed: " + cudaGetErrorString(e));
  }
  p->ws_bytes = need;
  p->ws_owned = true;
  return QSB_OK;
}

extern "C" {

int qsb_version(void) { return QSB_VERSION; }
const char* qsb_last_error(void) { return g_err.c_str(); }

int qsb_device_count(int* count) {
  if (!count) return fail(QSB_ERR_BAD_ARG, "count is null");
  int c = 0;
  cudaError_t e = cudaGetDeviceCount(&c);
  if (e != cudaSuccess) { cudaGetLastError(); c = 0; }
  *count = c;
  return QSB_OK;
}

}  // extern "C"

static int plan_create(const qsb_table* table, int device, int rank, int world, qsb_plan** out) {
  if (!out) return fail(QSB_ERR_BAD_ARG, "plan out-pointer is null");
  *out = nullptr;
  int rc = check_table(table);
  if (rc) return rc;
  int wl = 0;
  while ((1 << wl) < world) ++wl;
  if (world < 1 || world > MAX_WORLD || (1 << wl) != world || rank < 0 || rank >= world)
    return fail(QSB_ERR_BAD_ARG, "world must be 1, 2, 4 or 8 and 0 <= rank < world");
  const int n = table->n_qubits;
  if (world > 1 && n - wl < TB)
    return fail(QSB_ERR_UNSUPPORTED, "a sharded plan needs n_qubits - log2(world) >= 12 (one tile per rank)");
  rc = pick_device(device);
  if (rc) return rc;
  std::unique_ptr<qsb_plan> p(new qsb_plan);
  p->n = n;
  p->device = device;
  p->rank = rank;
  p->world = world;
  p->small = world == 1 && n <= QSB_SMALL_MAX_QUBITS;
  p->h.assign(table->h, table->h + 3 * (size_t)n);
  p->J.assign(table->J, table->J + 3 * (size_t)n * n);
  p->sea_mask = table->sea_mask;
  p->rare_mask = table->rare_mask;
  if (world > 1) {  // sharded: the plan's stream outranks the centre stream of the overlapped mode (qsb_shard_set_stream_barrier)
    int least = 0, greatest = 0;
    QSB_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    QSB_CUDA(cudaStreamCreateWithPriority(&p->stream, cudaStreamNonBlocking, greatest));
  } else {
    QSB_CUDA(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
  }
  if (!p->small) {
    const int ne = std::max(n, TB);
    p->n_eff = ne;
    p->n_loc = ne - wl;
    prepare_device();
    p->g = make_geometry(ne, rank, world);
    p->g.sea_mask = table->sea_mask;
    p->g.rare_mask = table->rare_mask;
    // pad the table with decoupled spectator bits (zero coefficients) up to 12 qubits
    std::vector<double> h(3 * (size_t)ne, 0.0), J(3 * (size_t)ne * ne, 0.0);
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < n; ++b) {
        h[(size_t)a * ne + b] = table->h[a * n + b];
        for (int c = 0; c < n; ++c)
          if (b != c) J[((size_t)a * ne + b) * ne + c] = table->J[((size_t)a * n + b) * n + c];
      }
    const Layout& Lc = p->g.lay[p->g.n_layouts - 1];
    const int wc = centre_window(Lc);
    const InTileDiag qx = host_in_tile_diag(h.data(), J.data(), ne, 0, Lc, wc, std::ldexp(1.0, -ne));
    const InTileDiag qy = host_in_tile_diag(h.data(), J.data(), ne, 1, Lc, wc, std::ldexp(1.0, -ne));
    const InTileDiag qz = host_in_tile_diag(h.data(), J.data(), ne, 2, p->g.lay[0], 8, 1.0);
    std::vector<double> blob;
    blob.insert(blob.end(), h.begin(), h.end());
    blob.insert(blob.end(), J.begin(), J.end());
    size_t off_q = blob.size();
    blob.insert(blob.end(), qx.tab, qx.tab + 5 * 256);
    blob.insert(blob.end(), qy.tab, qy.tab + 5 * 256);
    blob.insert(blob.end(), qz.tab, qz.tab + 5 * 256);
    memcpy(p->tab.Wx, qx.ww, sizeof qx.ww);
    memcpy(p->tab.Wy, qy.ww, sizeof qy.ww);
    memcpy(p->tab.Wz, qz.ww, sizeof qz.ww);
    QSB_CUDA(cudaMalloc(&p->d_tables, blob.size() * sizeof(double)));
    QSB_CUDA(cudaMemcpy(p->d_tables, blob.data(), blob.size() * sizeof(double), cudaMemcpyHostToDevice));
    p->tab.h = p->d_tables;
    p->tab.J = p->d_tables + h.size();
    p->tab.Ix = p->d_tables + off_q;
    p->tab.Iy = p->tab.Ix + 5 * 256;
    p->tab.Iz = p->tab.Iy + 5 * 256;
    const size_t td = tile_diag_doubles(p->n_loc);
    QSB_CUDA(cudaMalloc(&p->d_tiles, 3 * td * sizeof(double)));
    const double sxy = std::ldexp(1.0, -ne);
    // this rank's slice of the centre tiles (the rank bits are inside those tiles); its own layout-0 tiles
    const uint64_t ntl = 1ull << (p->n_loc - TB), v0 = (uint64_t)rank * ntl;
    launch_build_tile_diag(p->tab.h, p->tab.J, ne, Lc, wc, sxy, p->d_tiles, ntl, v0, 0, p->stream);
    launch_build_tile_diag(p->tab.h + ne, p->tab.J + (size_t)ne * ne, ne, Lc, wc, sxy, p->d_tiles + td, ntl, v0, 0, p->stream);
    launch_build_tile_diag(p->tab.h + 2 * ne, p->tab.J + 2 * (size_t)ne * ne, ne, p->g.lay[0], 8, 1.0, p->d_tiles + 2 * td,
                           ntl, 0, p->g.rank_bits, p->stream);
    QSB_CUDA(cudaGetLastError());
    QSB_CUDA(cudaStreamSynchronize(p->stream));
    p->tab.Tx = p->d_tiles;
    p->tab.Ty = p->d_tiles + td;
    p->tab.Tz = p->d_tiles + 2 * td;
    QSB_CUDA(cudaMalloc(&p->d_red, PSTRIDE * sizeof(double)));
    QSB_CUDA(cudaMallocHost(&p->h_red, PSTRIDE * sizeof(double)));
    {
      int n_sms = 0;
      QSB_CUDA(cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, device));
      p->fu = plan_fusion(p->g, n_sms);
      if (p->fu.on) QSB_CUDA(cudaMalloc(&p->fu.ctrl, p->fu.ctrl_bytes));
    }
    if (world > 1) {
      // exchange allocation the peers map: tx, ty and - for the push-only exchange - the receive buffers rx, ry
      const char* pe = getenv("QSB200_SHARD_PUSH");
      p->xvecs = (pe && *pe == '1') ? 4 : 2;
      cudaError_t e = cudaMalloc(&p->xbuf, p->xvecs * (sizeof(cplx) << p->n_loc));
      if (e != cudaSuccess) {
        cudaGetLastError();
        p->xbuf = nullptr;
        qsb_plan* raw = p.release();
        qsb_plan_destroy(raw);
        return fail(QSB_ERR_OOM, std::string("cudaMalloc of the shard exchange scratch failed: ") + cudaGetErrorString(e));
      }
    }
  }
  *out = p.release();
  return QSB_OK;
}

extern "C" {

int qsb_plan_create(const qsb_table* table, int device, qsb_plan** out) { return plan_create(table, device, 0, 1, out); }

int qsb_session_end(qsb_plan* plan);

int qsb_plan_destroy(qsb_plan* p) {
  if (!p) return QSB_OK;
  cudaSetDevice(p->device);
  qsb_session_end(p);
  if (p->stream) cudaStreamSynchronize(p->stream);
  for (int r = 0; r < MAX_WORLD; ++r)
    if (p->ipc_opened[r]) cudaIpcCloseMemHandle(p->ipc_opened[r]);
  if (p->link.centre_stream) {
    cudaStreamSynchronize(p->link.centre_stream);
    for (int b = 0; b < MAX_PIPE; ++b) { cudaEventDestroy(p->link.ev_fwd[b]); cudaEventDestroy(p->link.ev_centre[b]); }
    cudaStreamDestroy(p->link.centre_stream);
  }
  if (p->dense_buf) cudaFree(p->dense_buf);
  if (p->xbuf) cudaFree(p->xbuf);
  if (p->ws && p->ws_owned) cudaFree(p->ws);
  if (p->d_tables) cudaFree(p->d_tables);
  if (p->d_tiles) cudaFree(p->d_tiles);
  if (p->d_red) cudaFree(p->d_red);
  if (p->fu.ctrl) cudaFree(p->fu.ctrl);
  if (p->h_red) cudaFreeHost(p->h_red);
  if (p->stream) cudaStreamDestroy(p->stream);
  delete p;
  return QSB_OK;
}

int qsb_plan_workspace_bytes(const qsb_plan* p, int method, int order, size_t* bytes) {
  (void)order;
  if (!p || !bytes) return fail(QSB_ERR_BAD_ARG, "null argument");
  *bytes = workspace_need(p, method);
  return QSB_OK;
}

int qsb_plan_set_workspace(qsb_plan* p, void* device_ptr, size_t bytes) {
  if (!p) return fail(QSB_ERR_BAD_ARG, "null plan");
  if (p->ses) return fail(QSB_ERR_BAD_ARG, "cannot change the workspace while a session is open");
  if (((uintptr_t)device_ptr) & 255) return fail(QSB_ERR_BAD_ARG, "workspace must be 256-byte aligned");
  if (p->ws && p->ws_owned) cudaFree(p->ws);
  p->ws = device_ptr;
  p->ws_bytes = device_ptr ? bytes : 0;
  p->ws_owned = false;
  return QSB_OK;
}

// ---- on-chip batch path ---------------------------------------------------------------------------
int qsb_evolve_batch(int32_t n_traj, const qsb_table* tables, const uint64_t* psi0_index, int64_t n_out,
                     const double* t_out, int32_t t_out_per_traj, const qsb_solver_opts* opts, double* out_obs,
                     qsb_stats* stats, int device) {
  if (n_traj < 1 || !tables || !psi0_index || !opts || !out_obs)
    return fail(QSB_ERR_BAD_ARG, "qsb_evolve_batch: null argument or n_traj < 1");
  int max_n = 0, min_n = QSB_MAX_QUBITS;
  for (int i = 0; i < n_traj; ++i) {
    int rc = check_table(&tables[i]);
    if (rc) return rc;
    if (tables[i].n_qubits > QSB_SMALL_MAX_QUBITS)
      return fail(QSB_ERR_UNSUPPORTED, "qsb_evolve_batch handles n_qubits <= " + std::to_string(QSB_SMALL_MAX_QUBITS) +
                                           "; use one plan + qsb_evolve per larger trajectory");
    rc = check_opts(&opts[i]);
    if (rc) return rc;
    if (psi0_index[i] >> tables[i].n_qubits) return fail(QSB_ERR_BAD_ARG, "psi0_index out of range");
    rc = check_grid(n_out, t_out + (t_out_per_traj ? (size_t)i * n_out : 0));
    if (rc) return rc;
    max_n = std::max(max_n, tables[i].n_qubits);
    min_n = std::min(min_n, tables[i].n_qubits);
  }
  int rc = pick_device(device);
  if (rc) return rc;
  std::vector<SmallDesc> descs(n_traj);
  for (int i = 0; i < n_traj; ++i) compile_small(&tables[i], psi0_index[i], &opts[i], t_out_per_traj, &descs[i]);

  SmallDesc* d_desc = nullptr;
  double *d_t = nullptr, *d_obs = nullptr;
  SmallStats* d_stats = nullptr;
  const size_t nt = (size_t)n_out * (t_out_per_traj ? n_traj : 1);
  const size_t nobs = (size_t)n_traj * n_out * QSB_NOBS;
  cudaStream_t st = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  std::vector<SmallStats> hs(n_traj);  // declared before the guard: it must outlive the guard's stream drain
  struct Guard {  // every early return (QSB_CUDA) releases what has been created so far
    SmallDesc*& a; double*& b; double*& c; SmallStats*& d; cudaStream_t& s; cudaEvent_t &x, &y;
    ~Guard() {
      if (s) cudaStreamSynchronize(s);
      cudaFree(a); cudaFree(b); cudaFree(c); cudaFree(d);
      if (x) cudaEventDestroy(x);
      if (y) cudaEventDestroy(y);
      if (s) cudaStreamDestroy(s);
    }
  } guard{d_desc, d_t, d_obs, d_stats, st, e0, e1};
  QSB_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  QSB_CUDA(cudaEventCreate(&e0));
  QSB_CUDA(cudaEventCreate(&e1));
  QSB_CUDA(cudaMalloc(&d_desc, sizeof(SmallDesc) * n_traj));
  QSB_CUDA(cudaMalloc(&d_t, sizeof(double) * nt));
  QSB_CUDA(cudaMalloc(&d_obs, sizeof(double) * nobs));
  QSB_CUDA(cudaMalloc(&d_stats, sizeof(SmallStats) * n_traj));
  QSB_CUDA(cudaMemcpyAsync(d_desc, descs.data(), sizeof(SmallDesc) * n_traj, cudaMemcpyHostToDevice, st));
  QSB_CUDA(cudaMemcpyAsync(d_t, t_out, sizeof(double) * nt, cudaMemcpyHostToDevice, st));
  QSB_CUDA(cudaEventRecord(e0, st));
  launch_small_evolve(d_desc, n_traj, max_n, min_n == max_n, d_t, n_out, d_obs, d_stats, st);
  QSB_CUDA(cudaGetLastError());
  QSB_CUDA(cudaEventRecord(e1, st));
  QSB_CUDA(cudaMemcpyAsync(out_obs, d_obs, sizeof(double) * nobs, cudaMemcpyDeviceToHost, st));
  QSB_CUDA(cudaMemcpyAsync(hs.data(), d_stats, sizeof(SmallStats) * n_traj, cudaMemcpyDeviceToHost, st));
  QSB_CUDA(cudaStreamSynchronize(st));
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  int worst = QSB_OK;  // (buffers, events and the stream are released by `guard`)
  for (int i = 0; i < n_traj; ++i) {
    if (stats) {
      stats[i].n_accepted = hs[i].n_accepted; stats[i].n_rejected = hs[i].n_rejected; stats[i].n_rhs = hs[i].n_rhs;
      stats[i].last_h = hs[i].last_h; stats[i].device_ms = ms; stats[i].path = 0; stats[i].status = hs[i].status;
    }
    if (hs[i].status != QSB_OK && worst == QSB_OK) {
      worst = hs[i].status;
      g_err = "trajectory " + std::to_string(i) + (hs[i].status == QSB_ERR_NSTEPS
                  ? ": more than nsteps internal steps inside one output interval (excess work)"
                  : ": step size underflow");
    }
  }
  return worst;
}

}  // extern "C"

// Shared body of qsb_propagate_batch / qsb_build_propagator: columns of U = exp(-i H dt) through the on-chip
// trajectory kernel (one CTA per column), then - when n_out > 0 - the matvec chain with observables.
// Upper bound on ||H||: every one-site operator has norm 1/2, every pair product 1/4.
static double table_norm_bound(const qsb_table* t) {
  const int n = t->n_qubits;
  double s = 0.0;
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < n; ++b) {
      s += 0.5 * std::fabs(t->h[a * n + b]);
      for (int c = b + 1; c < n; ++c) s += 0.25 * std::fabs(t->J[((size_t)a * n + b) * n + c]);
    }
  return s;
}

// Squarings for one propagator: the columns are integrated over dt / 2^s with ||H|| dt / 2^s <= ~8 (a handful of
// Taylor(20) steps, enough for the controller to settle), then U is squared s times.
static int auto_squarings(const qsb_table* t, double dt) {
  const double x = table_norm_bound(t) * dt / 8.0;
  int s = 0;
  while (s < 40 && std::ldexp(1.0, s) < x) ++s;
  return s;
}

// Per-device context of the propagator path, kept between calls: a sweep driver calls qsb_propagate_batch once per
// sweep (or once per grid), and stream / event creation plus 7-10 cudaMalloc + cudaFree pairs per call were a third
// of the full-size C2 sweep's wall time (40 ms for 17 ms of kernels).  Buffers only grow; qsb_release_caches() frees
// them.  One call at a time per device (the mutex is held for the whole call).
struct PropCtx {
  static constexpr int NBUF = 10;
  std::mutex mu;
  cudaStream_t st = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  void* buf[NBUF] = {};
  size_t cap[NBUF] = {};
  int ensure_stream() {
    if (st) return QSB_OK;
    QSB_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    QSB_CUDA(cudaEventCreate(&e0));
    QSB_CUDA(cudaEventCreate(&e1));
    return QSB_OK;
  }
  int get(int slot, size_t bytes, void** out) {
    if (cap[slot] < bytes) {
      if (buf[slot]) { cudaFree(buf[slot]); buf[slot] = nullptr; cap[slot] = 0; }
      cudaError_t e = cudaMalloc(&buf[slot], bytes);
      if (e != cudaSuccess) {
        buf[slot] = nullptr;
        cudaGetLastError();
        return fail(QSB_ERR_OOM, "cudaMalloc of " + std::to_string(bytes) + " bytes for the propagator path failed: " + cudaGetErrorString(e));
      }
      cap[slot] = bytes;
    }
    *out = buf[slot];
    return QSB_OK;
  }
  void release() {
    if (st) cudaStreamSynchronize(st);
    for (int i = 0; i < NBUF; ++i) { if (buf[i]) cudaFree(buf[i]); buf[i] = nullptr; cap[i] = 0; }
  }
};
static PropCtx& prop_ctx(int device) {
  static PropCtx ctxs[64];
  return ctxs[device & 63];
}

// First touch of a large pageable host buffer costs ~0.35 s per GB on one thread (page faults); done here by a few
// threads WHILE the kernels that will fill it are running, it costs nothing.
static void prefault_async(std::vector<std::thread>& pool, void* ptr, size_t bytes) {
  if (bytes < ((size_t)8 << 20)) return;
  unsigned nt = std::min(8u, std::max(1u, std::thread::hardware_concurrency() / 2));
  const size_t page = 4096, per = ((bytes / nt) + page - 1) & ~(page - 1);
  for (unsigned t = 0; t < nt; ++t) {
    const size_t lo = (size_t)t * per, hi = std::min(bytes, lo + per);
    if (lo >= hi) break;
    pool.emplace_back([=] {
      volatile char* p = (volatile char*)ptr;
      for (size_t o = lo; o < hi; o += page) p[o] = 0;
    });
  }
}

static int propagate_impl(int32_t n_traj, const qsb_table* tables, const uint64_t* psi0_index, int64_t n_out,
                          const double* dt, int32_t dt_per_traj, const qsb_solver_opts* opts, int32_t max_squarings,
                          int32_t obs_by_column, double* out_obs, qsb_stats* stats, double* u_host, int device) {
  const int n = tables[0].n_qubits;
  for (int i = 0; i < n_traj; ++i) {
    int rc = check_table(&tables[i]);
    if (rc) return rc;
    if (tables[i].n_qubits > QSB_SMALL_MAX_QUBITS)
      return fail(QSB_ERR_UNSUPPORTED, "the propagator path handles n_qubits <= " + std::to_string(QSB_SMALL_MAX_QUBITS));
    if (tables[i].n_qubits != n) return fail(QSB_ERR_BAD_ARG, "the propagator path needs the same n_qubits in every table of a call");
    rc = check_opts(&opts[i]);
    if (rc) return rc;
    if (psi0_index && (psi0_index[i] >> n)) return fail(QSB_ERR_BAD_ARG, "psi0_index out of range");
    const double h = dt[dt_per_traj ? i : 0];
    if (!(h > 0.0) || !std::isfinite(h)) return fail(QSB_ERR_BAD_ARG, "dt must be finite and > 0");
  }
  int rc = pick_device(device);
  if (rc) return rc;
  const size_t dim = (size_t)1 << n;
  const size_t u_bytes = sizeof(cplx) * dim * dim;
  // trajectories per pass: keep the propagators of one pass within ~4 GiB
  int chunk = (int)std::max<size_t>(1, std::min<size_t>((size_t)n_traj, ((size_t)4 << 30) / u_bytes));
  if (const char* e = getenv("QSB200_PROP_CHUNK")) chunk = std::max(1, std::min(chunk, atoi(e)));  // test knob
  if (n_out > 0) {  // ... and their observable traces within ~2 GiB
    const size_t per_traj = sizeof(double) * (size_t)n_out * QSB_NOBS;
    chunk = (int)std::max<size_t>(1, std::min<size_t>((size_t)chunk, ((size_t)2 << 30) / per_traj));
  }

  PropCtx& ctx = prop_ctx(device);
  std::lock_guard<std::mutex> ctx_lock(ctx.mu);
  rc = ctx.ensure_stream();
  if (rc) return rc;
  cudaStream_t st = ctx.st;
  cudaEvent_t e0 = ctx.e0, e1 = ctx.e1;
  SmallDesc* d_desc = nullptr;
  double *d_t = nullptr, *d_obs = nullptr;
  SmallStats* d_stats = nullptr;
  cplx *d_u = nullptr, *d_tmp = nullptr;
  int32_t* d_sq = nullptr;
  struct Drain {  // an early return must not leave work of this call running on the shared stream
    cudaStream_t s;
    std::vector<std::thread> pool;
    ~Drain() { cudaStreamSynchronize(s); for (auto& t : pool) t.join(); }
  } drain{st, {}};
  std::vector<int32_t> n_sq(n_traj, 0);
  bool any_sq = false;
  for (int i = 0; i < n_traj; ++i) {
    int s = max_squarings == 0 ? 0 : auto_squarings(&tables[i], dt[dt_per_traj ? i : 0]);
    if (max_squarings > 0 && s > max_squarings) s = max_squarings;
    n_sq[i] = s;
    any_sq = any_sq || s > 0;
  }
  if ((rc = ctx.get(0, sizeof(SmallDesc) * chunk, (void**)&d_desc))) return rc;
  if ((rc = ctx.get(1, sizeof(double) * 2 * chunk, (void**)&d_t))) return rc;
  if ((rc = ctx.get(2, sizeof(SmallStats) * chunk * dim, (void**)&d_stats))) return rc;
  if ((rc = ctx.get(3, u_bytes * chunk, (void**)&d_u))) return rc;
  if (any_sq) {
    if ((rc = ctx.get(4, u_bytes * chunk, (void**)&d_tmp))) return rc;
    if ((rc = ctx.get(5, sizeof(int32_t) * chunk, (void**)&d_sq))) return rc;
  }
  if (n_out > 0 && (rc = ctx.get(6, sizeof(double) * chunk * n_out * QSB_NOBS, (void**)&d_obs))) return rc;

  // Time-split chains: a trajectory's outputs are cut into segments of 2^seg_lb outputs that run as separate CTAs, each
  // from its own seed psi_{c 2^seg_lb} = (U^(2^seg_lb))^c psi_0 (seg_lb more squarings of a copy of U, then a chain of
  // n_seg matvecs that stores states).  The segment length depends on n_out ONLY (~128 segments, at least 64 outputs
  // each), so a trajectory's traces do not depend on what else is in the batch.
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
  cplx *d_v = nullptr, *d_seeds = nullptr;
  int32_t* d_seg_sq = nullptr;
  std::vector<int32_t> seg_sq(chunk, seg_lb);
  if (n_seg > 1) {
    if ((rc = ctx.get(7, u_bytes * chunk, (void**)&d_v))) return rc;
    if ((rc = ctx.get(8, sizeof(cplx) * dim * (size_t)n_seg * chunk, (void**)&d_seeds))) return rc;
    if ((rc = ctx.get(9, sizeof(int32_t) * chunk, (void**)&d_seg_sq))) return rc;
    if (!d_tmp && (rc = ctx.get(4, u_bytes * chunk, (void**)&d_tmp))) return rc;
    QSB_CUDA(cudaMemcpyAsync(d_seg_sq, seg_sq.data(), sizeof(int32_t) * chunk, cudaMemcpyHostToDevice, st));
  }

  std::vector<SmallDesc> descs(chunk);
  std::vector<double> tp(2 * (size_t)chunk);
  std::vector<SmallStats> hs((size_t)chunk * dim);
  int worst = QSB_OK;
  for (int base = 0; base < n_traj; base += chunk) {
    const int m = std::min(chunk, n_traj - base);
    for (int i = 0; i < m; ++i) {
      compile_small(&tables[base + i], psi0_index ? psi0_index[base + i] : 0, &opts[base + i], 1, &descs[i]);
      tp[2 * i] = 0.0;
      tp[2 * i + 1] = std::ldexp(dt[dt_per_traj ? base + i : 0], -n_sq[base + i]);
    }
    if (any_sq)
      QSB_CUDA(cudaMemcpyAsync(d_sq, n_sq.data() + base, sizeof(int32_t) * m, cudaMemcpyHostToDevice, st));
    QSB_CUDA(cudaMemcpyAsync(d_desc, descs.data(), sizeof(SmallDesc) * m, cudaMemcpyHostToDevice, st));
    QSB_CUDA(cudaMemcpyAsync(d_t, tp.data(), sizeof(double) * 2 * m, cudaMemcpyHostToDevice, st));
    QSB_CUDA(cudaEventRecord(e0, st));
    launch_small_columns(d_desc, m, n, d_t, d_stats, d_u, st);
    QSB_CUDA(cudaGetLastError());
    if (any_sq) {
      launch_small_square(d_u, d_tmp, d_sq, m, n, st);
      QSB_CUDA(cudaGetLastError());
    }
    if (n_out > 0 && n_seg > 1) {
      QSB_CUDA(cudaMemcpyAsync(d_v, d_u, u_bytes * m, cudaMemcpyDeviceToDevice, st));
      launch_small_square(d_v, d_tmp, d_seg_sq, m, n, st);                                        // V = U^seg_len
      QSB_CUDA(cudaGetLastError());
      rc = launch_small_propagate(d_desc, m, n, d_v, n_seg, nullptr, 0, st, nullptr, 1, n_seg, d_seeds);  // seeds
      if (rc) return rc;
      rc = launch_small_propagate(d_desc, m, n, d_u, n_out, d_obs, obs_by_column != 0, st, d_seeds, n_seg, seg_len, nullptr);
      if (rc) return rc;
    } else if (n_out > 0) {
      rc = launch_small_propagate(d_desc, m, n, d_u, n_out, d_obs, obs_by_column != 0, st);
      if (rc) return rc;
    }
    QSB_CUDA(cudaEventRecord(e1, st));
    if (n_out > 0 && base == 0)  // the kernels are running: fault the caller's output pages in meanwhile
      prefault_async(drain.pool, out_obs, sizeof(double) * (size_t)n_traj * n_out * QSB_NOBS);
    if (base == 0) { for (auto& t : drain.pool) t.join(); drain.pool.clear(); }
    QSB_CUDA(cudaMemcpyAsync(hs.data(), d_stats, sizeof(SmallStats) * m * dim, cudaMemcpyDeviceToHost, st));
    if (n_out > 0)
      QSB_CUDA(cudaMemcpyAsync(out_obs + (size_t)base * n_out * QSB_NOBS, d_obs, sizeof(double) * m * n_out * QSB_NOBS,
                               cudaMemcpyDeviceToHost, st));
    if (u_host)
      QSB_CUDA(cudaMemcpyAsync(u_host + 2 * dim * dim * (size_t)base, d_u, u_bytes * m, cudaMemcpyDeviceToHost, st));
    QSB_CUDA(cudaStreamSynchronize(st));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    for (int i = 0; i < m; ++i) {
      qsb_stats acc = {0, 0, 0, 0.0, (double)ms, 2, QSB_OK};
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

int qsb_propagate_batch(int32_t n_traj, const qsb_table* tables, const uint64_t* psi0_index, int64_t n_out,
                        const double* dt, int32_t dt_per_traj, const qsb_solver_opts* opts, int32_t max_squarings,
                        int32_t obs_by_column, double* out_obs, qsb_stats* stats, int device) {
  if (n_traj < 1 || !tables || !psi0_index || !opts || !out_obs || !dt || n_out < 1)
    return fail(QSB_ERR_BAD_ARG, "qsb_propagate_batch: null argument, n_traj < 1 or n_out < 1");
  return propagate_impl(n_traj, tables, psi0_index, n_out, dt, dt_per_traj, opts, max_squarings, obs_by_column, out_obs,
                        stats, nullptr, device);
}

int qsb_build_propagator(const qsb_table* table, double dt, const qsb_solver_opts* opts, int32_t max_squarings,
                         double* u_host, qsb_stats* stats, int device) {
  if (!table || !opts || !u_host) return fail(QSB_ERR_BAD_ARG, "qsb_build_propagator: null argument");
  return propagate_impl(1, table, nullptr, 0, &dt, 0, opts, max_squarings, 0, nullptr, stats, u_host, device);
}

int qsb_release_caches(int device) {
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess) { cudaGetLastError(); return QSB_OK; }
  for (int dev = 0; dev < count && dev < 64; ++dev) {
    if (device >= 0 && dev != device) continue;
    PropCtx& c = prop_ctx(dev);
    std::lock_guard<std::mutex> lk(c.mu);
    if (!c.st) continue;
    if (cudaSetDevice(dev) != cudaSuccess) { cudaGetLastError(); continue; }
    c.release();
  }
  return QSB_OK;
}

int qsb_debug_squarings(const qsb_table* table, double dt, int32_t* n_squarings) {
  if (!n_squarings) return fail(QSB_ERR_BAD_ARG, "null argument");
  int rc = check_table(table);
  if (rc) return rc;
  *n_squarings = auto_squarings(table, dt);
  return QSB_OK;
}

int qsb_debug_propagate_shape(int32_t n_qubits, int32_t* out) {
  if (!out || n_qubits < 1 || n_qubits > QSB_SMALL_MAX_QUBITS) return fail(QSB_ERR_BAD_ARG, "need 1 <= n_qubits <= 10 and out");
  int th, sp, nc, rg;
  size_t smem;
  small_propagate_shape(n_qubits, &th, &sp, &nc, &rg, &smem);
  out[0] = th; out[1] = sp; out[2] = nc; out[3] = (int32_t)smem; out[4] = rg;
  return QSB_OK;
}

}  // extern "C"

#include "qsb_api_session.inl"
