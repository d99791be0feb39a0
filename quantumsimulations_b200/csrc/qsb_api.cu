// C ABI of libqsb200 (include/qsb200.h): plans, the on-chip batch path, the tiled HBM path with its
// host-side step controller, and the test/benchmark hooks.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <memory>
#include <mutex>
#include <thread>
#include <string>
#include <vector>

#include <cublas_v2.h>  // types only: the library itself is dlopen-ed on first use (qsb_api_prophbm.inl)
#include <dlfcn.h>

#include "qsb_expect.cuh"
#include "qsb_large.cuh"
#include "qsb_small.cuh"

namespace qsb {

static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }
int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  char buf[512];
  snprintf(buf, sizeof buf, "CUDA error %d (%s) at %s:%d in %s", (int)e, cudaGetErrorString(e), file, line, what);
  g_err = buf;
  cudaGetLastError();
  if (e == cudaErrorMemoryAllocation) return QSB_ERR_OOM;
  if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) return QSB_ERR_NO_DEVICE;
  return QSB_ERR_CUDA;
}
static int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}

// ---- table validation / compilation ------------------------------------------------------------------
static int check_table(const qsb_table* t) {
  if (!t || !t->h || !t->J) return fail(QSB_ERR_BAD_ARG, "table, table->h and table->J must be non-null");
  if (t->n_qubits < 1 || t->n_qubits > QSB_MAX_QUBITS)
    return fail(QSB_ERR_BAD_ARG, "n_qubits must be in 1.." + std::to_string(QSB_MAX_QUBITS));
  const int n = t->n_qubits;
  const uint64_t all = (n >= 64) ? ~0ull : ((1ull << n) - 1);
  if (t->sea_mask & ~all) return fail(QSB_ERR_BAD_ARG, "sea_mask has bits above n_qubits");
  if (!t->rare_mask || (t->rare_mask & (t->rare_mask - 1)) || (t->rare_mask & ~all))
    return fail(QSB_ERR_BAD_ARG, "rare_mask must be exactly one bit below n_qubits");
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < n; ++b) {
      if (!std::isfinite(t->h[a * n + b])) return fail(QSB_ERR_BAD_ARG, "non-finite field coefficient");
      for (int c = 0; c < n; ++c) {
        double v = t->J[((size_t)a * n + b) * n + c];
        if (!std::isfinite(v)) return fail(QSB_ERR_BAD_ARG, "non-finite pair coefficient");
        if (b != c && v != t->J[((size_t)a * n + c) * n + b]) return fail(QSB_ERR_BAD_ARG, "J must be symmetric");
      }
    }
  return QSB_OK;
}

static int check_opts(const qsb_solver_opts* o) {
  if (!o) return fail(QSB_ERR_BAD_ARG, "opts must be non-null");
  if (!(o->atol > 0.0) || !(o->rtol >= 0.0) || !std::isfinite(o->atol) || !std::isfinite(o->rtol))
    return fail(QSB_ERR_BAD_ARG, "need atol > 0 and rtol >= 0");
  if (o->max_step < 0.0 || o->first_step < 0.0 || o->nsteps < 0)
    return fail(QSB_ERR_BAD_ARG, "max_step, first_step and nsteps must be >= 0");
  if (o->method != QSB_METHOD_DP5 && o->method != QSB_METHOD_TAYLOR) return fail(QSB_ERR_BAD_ARG, "unknown method");
  if (o->method == QSB_METHOD_TAYLOR && o->order != 0 && (o->order < 4 || o->order > QSB_MAX_TAYLOR_ORDER))
    return fail(QSB_ERR_BAD_ARG, "Taylor order must be 0 (default) or 4..30");
  return QSB_OK;
}

static int check_grid(int64_t n_out, const double* t) {
  if (n_out < 1 || !t) return fail(QSB_ERR_BAD_ARG, "need n_out >= 1 and a time grid");
  for (int64_t k = 1; k < n_out; ++k)
    if (!(t[k] > t[k - 1])) return fail(QSB_ERR_BAD_ARG, "t_out must be strictly increasing");
  return QSB_OK;
}

static void compile_small(const qsb_table* t, uint64_t psi0, const qsb_solver_opts* o, int t_per_traj, SmallDesc* d) {
  memset(d, 0, sizeof *d);
  const int n = t->n_qubits;
  d->n = n;
  d->method = o->method;
  d->order = (o->method == QSB_METHOD_TAYLOR) ? (o->order ? o->order : 20) : 5;
  d->t_per_traj = t_per_traj;
  d->sea_mask = (uint32_t)t->sea_mask;
  d->rare_mask = (uint32_t)t->rare_mask;
  d->psi0 = (uint32_t)psi0;
  d->nsteps = o->nsteps;
  d->atol = o->atol; d->rtol = o->rtol; d->max_step = o->max_step; d->first_step = o->first_step;
  const double *hx = t->h, *hy = t->h + n, *hz = t->h + 2 * n;
  const double *jx = t->J, *jy = t->J + (size_t)n * n, *jz = t->J + 2 * (size_t)n * n;
  int pi = 0;
  for (int b = 0; b < n; ++b) {
    d->hz_half[b] = 0.5 * hz[b];
    d->hx_half[b] = 0.5 * hx[b];
    d->hy_half[b] = 0.5 * hy[b];
    if (hx[b] != 0.0 || hy[b] != 0.0) {
      SmallFlip& f = d->flip[d->n_flip++];
      f.hxh = 0.5 * hx[b]; f.hyh = 0.5 * hy[b]; f.mask = 1u << b;
    }
    for (int c = b + 1; c < n; ++c, ++pi) {
      d->jzz_quarter[pi] = 0.25 * jz[b * n + c];
      double xx = jx[b * n + c], yy = jy[b * n + c];
      d->pair_same[pi] = 0.25 * (xx - yy);
      d->pair_diff[pi] = 0.25 * (xx + yy);
      if (xx != 0.0 || yy != 0.0) {
        SmallPair& p = d->pair[d->n_pair++];
        p.c_same = 0.25 * (xx - yy); p.c_diff = 0.25 * (xx + yy); p.mask = (1u << b) | (1u << c);
      }
    }
  }
}

static int pick_device(int device) {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    cudaGetLastError();
    return fail(QSB_ERR_NO_DEVICE, "no CUDA device available (this library has no CPU fallback)");
  }
  if (device < 0 || device >= count) return fail(QSB_ERR_BAD_ARG, "device index out of range");
  QSB_CUDA(cudaSetDevice(device));
  return QSB_OK;
}

}  // namespace qsb

#include "qsb_api_plan.inl"
#include "qsb_api_expect.inl"
