// Sharded-state plumbing: exchange-scratch export/connect (CUDA IPC or same-process pointers), the rank barrier the
// centre pass needs, cross-rank scalar sums, and an in-process all-gather for ranks that are threads.
// Included from qsb_api_entry.inl.
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <unistd.h>

namespace {

// QSB200_TRACE=1: one stderr line per rank-level event of the sharded path (debugging aid)
bool shard_trace_on() {
  static int on = -1;
  if (on < 0) { const char* e = getenv("QSB200_TRACE"); on = (e && *e && *e != '0') ? 1 : 0; }
  return on == 1;
}
#define SHARD_TRACE(p, ...)                                          \
  do {                                                               \
    if (shard_trace_on()) {                                          \
      fprintf(stderr, "[qsb rank %d/%d] ", (p)->rank, (p)->world);   \
      fprintf(stderr, __VA_ARGS__);                                  \
      fprintf(stderr, "\n");                                         \
      fflush(stderr);                                                \
    }                                                                \
  } while (0)

struct ShardBlob {
  int64_t pid;
  int32_t device;
  int32_t rank;
  uint64_t ptr;
  uint64_t bytes;
  cudaIpcMemHandle_t handle;  // 64 bytes
};
static_assert(sizeof(ShardBlob) == QSB_SHARD_BLOB_BYTES, "blob layout");

// ShardLink::sync: everything this rank has queued must be complete before the peers are told so.
int shard_sync(void* ctx) {
  qsb_plan* p = (qsb_plan*)ctx;
  SHARD_TRACE(p, "sync: draining stream");
  // A rank whose stream failed still joins the barrier and tells the others, so that errors are collective
  // (a rank that simply returned would leave its peers waiting).
  const cudaError_t e = cudaStreamSynchronize(p->stream);
  double mine = (e == cudaSuccess) ? 0.0 : 1.0, all[MAX_WORLD];
  SHARD_TRACE(p, "sync: barrier");
  if (p->allgather(p->allgather_ctx, &mine, 1, all)) return fail(QSB_ERR_CUDA, "rank barrier (all-gather hook) failed");
  if (e != cudaSuccess) return cuda_fail(e, "cudaStreamSynchronize (sharded pass)", __FILE__, __LINE__);
  for (int r = 0; r < p->world; ++r)
    if (all[r] != 0.0) return fail(QSB_ERR_CUDA, "rank " + std::to_string(r) + " reported a CUDA failure in a sharded pass");
  return QSB_OK;
}

}  // namespace

// Replace the local sums by the rank-ordered sum over all ranks (identical bits on every rank).
static int shard_sum(qsb_plan* p, double* sums, int count) {
  if (p->world == 1) return QSB_OK;
  if (!p->connected) return fail(QSB_ERR_BAD_ARG, "sharded plan is not connected to its peers (qsb_shard_connect)");
  double all[MAX_WORLD * PSTRIDE];
  if (count > PSTRIDE) return fail(QSB_ERR_BAD_ARG, "shard_sum: too many values");
  SHARD_TRACE(p, "sum of %d scalars", count);
  if (p->allgather(p->allgather_ctx, sums, count, all)) return fail(QSB_ERR_CUDA, "all-gather hook failed");
  for (int k = 0; k < count; ++k) {
    double acc = 0.0;
    for (int r = 0; r < p->world; ++r) acc += all[r * count + k];
    sums[k] = acc;
  }
  return QSB_OK;
}

struct qsb_local_group {
  int world = 0;
  std::mutex m;
  std::condition_variable cv;
  int arrived = 0;
  uint64_t generation = 0;
  bool broken = false;      // a barrier timed out: every later call fails at once instead of waiting again
  double timeout_s = 120.0; // QSB200_BARRIER_TIMEOUT_S
  std::vector<double> buf;  // [world][count] of the current round
  struct Member { qsb_local_group* g; int rank; };
  std::vector<Member> members;
  bool barrier() {  // false: a peer never arrived (it failed outside the collective sequence)
    std::unique_lock<std::mutex> lk(m);
    if (broken) return false;
    const uint64_t gen = generation;
    if (++arrived == world) {
      arrived = 0;
      ++generation;
      cv.notify_all();
      return true;
    }
    if (!cv.wait_for(lk, std::chrono::duration<double>(timeout_s), [&] { return generation != gen || broken; }) || broken) {
      broken = true;
      cv.notify_all();
      return false;
    }
    return true;
  }
};

extern "C" {

int qsb_shard_plan_create(const qsb_table* table, int device, int32_t rank, int32_t world, qsb_plan** plan) {
  return plan_create(table, device, rank, world, plan);
}

int qsb_debug_layouts(int32_t n_qubits, int32_t world, int32_t* n_layouts, int32_t* out) {
  if (!n_layouts || !out) return fail(QSB_ERR_BAD_ARG, "null argument");
  int wl = 0;
  while ((1 << wl) < world) ++wl;
  if (world < 1 || world > MAX_WORLD || (1 << wl) != world || n_qubits - wl < TB || n_qubits > QSB_MAX_QUBITS)
    return fail(QSB_ERR_BAD_ARG, "unsupported (n_qubits, world)");
  LargeGeom g = make_geometry(n_qubits, 0, world);
  *n_layouts = g.n_layouts;
  for (int k = 0; k < g.n_layouts; ++k) { out[3 * k] = g.lay[k].lk; out[3 * k + 1] = g.lay[k].gs; out[3 * k + 2] = g.lay[k].ge; }
  return QSB_OK;
}

int qsb_debug_fusion(int32_t n_qubits, int32_t n_sms, int64_t* out) {
  if (!out) return fail(QSB_ERR_BAD_ARG, "null argument");
  if (n_qubits < TB || n_qubits > QSB_MAX_QUBITS) return fail(QSB_ERR_BAD_ARG, "unsupported n_qubits");
  const LargeGeom g = make_geometry(n_qubits, 0, 1);
  const StridedFusion fu = plan_fusion(g, n_sms);
  const FusedSpec& s = fu.spec;
  const int64_t head[7] = {fu.on ? 1 : 0, s.f, s.m, s.fbits, s.tiles_lb, s.lag, (int64_t)s.vmask};
  for (int i = 0; i < 7; ++i) out[i] = head[i];
  for (int k = 0; k < MAX_LAYOUTS; ++k) out[7 + k] = s.fixed_vmask[k];
  return QSB_OK;
}

int qsb_plan_local_dim(const qsb_plan* p, uint64_t* dim) {
  if (!p || !dim) return fail(QSB_ERR_BAD_ARG, "null argument");
  *dim = p->small ? (1ull << p->n) : (p->world > 1 ? (1ull << p->n_loc) : (1ull << p->n));
  return QSB_OK;
}

int qsb_shard_export(qsb_plan* p, void* blob) {
  if (!p || !blob) return fail(QSB_ERR_BAD_ARG, "null argument");
  if (p->world < 2 || !p->xbuf) return fail(QSB_ERR_BAD_ARG, "not a sharded plan");
  QSB_CUDA(cudaSetDevice(p->device));
  ShardBlob b;
  memset(&b, 0, sizeof b);
  b.pid = (int64_t)getpid();
  b.device = p->device;
  b.rank = p->rank;
  b.ptr = (uint64_t)(uintptr_t)p->xbuf;
  b.bytes = (uint64_t)p->xvecs * (sizeof(cplx) << p->n_loc);
  QSB_CUDA(cudaIpcGetMemHandle(&b.handle, p->xbuf));
  memcpy(blob, &b, sizeof b);
  return QSB_OK;
}

int qsb_shard_connect(qsb_plan* p, const void* blobs, qsb_allgather_fn allgather, void* ctx) {
  if (!p || !blobs || !allgather) return fail(QSB_ERR_BAD_ARG, "null argument");
  if (p->world < 2) return fail(QSB_ERR_BAD_ARG, "not a sharded plan");
  if (p->connected) return fail(QSB_ERR_BAD_ARG, "plan is already connected");
  QSB_CUDA(cudaSetDevice(p->device));
  const size_t vec = sizeof(cplx) << p->n_loc;
  for (int r = 0; r < p->world; ++r) {
    ShardBlob b;
    memcpy(&b, (const char*)blobs + (size_t)r * sizeof b, sizeof b);
    if (b.rank != r || b.bytes != (uint64_t)p->xvecs * vec)
      return fail(QSB_ERR_BAD_ARG, "blob " + std::to_string(r) + " does not describe rank " + std::to_string(r) +
                                       " of a plan with this shard size");
    char* base = nullptr;
    if (r == p->rank) {
      base = (char*)p->xbuf;
    } else if (b.pid == (int64_t)getpid()) {
      // the peer is a thread of this process: its pointer is valid here; other devices need peer access
      if (b.device != p->device) {
        int can = 0;
        QSB_CUDA(cudaDeviceCanAccessPeer(&can, p->device, b.device));
        if (!can) return fail(QSB_ERR_UNSUPPORTED, "no peer access between devices " + std::to_string(p->device) + " and " + std::to_string(b.device));
        cudaError_t e = cudaDeviceEnablePeerAccess(b.device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(e, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__);
        cudaGetLastError();
      }
      base = (char*)(uintptr_t)b.ptr;
    } else {
      void* mapped = nullptr;
      QSB_CUDA(cudaIpcOpenMemHandle(&mapped, b.handle, cudaIpcMemLazyEnablePeerAccess));
      p->ipc_opened[r] = mapped;
      base = (char*)mapped;
    }
    p->link.tx[r] = (cplx*)base;
    p->link.ty[r] = (cplx*)(base + vec);
    p->link.rx[r] = p->xvecs == 4 ? (cplx*)(base + 2 * vec) : nullptr;
    p->link.ry[r] = p->xvecs == 4 ? (cplx*)(base + 3 * vec) : nullptr;
  }
  p->allgather = allgather;
  p->allgather_ctx = ctx;
  p->link.push = p->xvecs == 4;
  p->link.ring_ready = false;  // the peers' pointers changed: the TMA descriptors are rebuilt on first use
  p->link.sync = shard_sync;
  p->link.ctx = p;
  p->connected = true;
  SHARD_TRACE(p, "connected; barrier");
  // nobody may start a collective call before every rank has mapped every peer
  if (allgather(ctx, nullptr, 0, nullptr)) return fail(QSB_ERR_CUDA, "rank barrier (all-gather hook) failed");
  return QSB_OK;
}

int qsb_shard_set_stream_barrier(qsb_plan* p, qsb_stream_barrier_fn fn, void* ctx) {
  if (!p) return fail(QSB_ERR_BAD_ARG, "null plan");
  if (p->world < 2) return fail(QSB_ERR_BAD_ARG, "not a sharded plan");
  if (p->ses) return fail(QSB_ERR_BAD_ARG, "cannot change the barrier while a session is open");
  QSB_CUDA(cudaSetDevice(p->device));
  if (fn && !p->link.centre_stream) {
    int least = 0, greatest = 0;
    QSB_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    QSB_CUDA(cudaStreamCreateWithPriority(&p->link.centre_stream, cudaStreamNonBlocking, least));
    for (int b = 0; b < MAX_PIPE; ++b) {
      QSB_CUDA(cudaEventCreateWithFlags(&p->link.ev_fwd[b], cudaEventDisableTiming));
      QSB_CUDA(cudaEventCreateWithFlags(&p->link.ev_centre[b], cudaEventDisableTiming));
    }
  }
  {
    int n_sms = 0;
    QSB_CUDA(cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, p->device));
    p->link.centre_ctas = std::max(1, n_sms);
    if (const char* e = getenv("QSB200_CENTRE_CTAS")) p->link.centre_ctas = std::max(1, atoi(e));
    p->link.pipe_lb = 3;  // measured on 2 x B200 at N = 31: 118 / 107 / 102 / 98 ms per H-apply for 1 / 2 / 4 / 8 groups
    if (const char* e = getenv("QSB200_SHARD_PIPE")) p->link.pipe_lb = std::max(0, std::min(4, atoi(e)));
  }
  p->link.stream_barrier = (int (*)(void*, cudaStream_t))fn;
  p->link.barrier_ctx = ctx;
  return QSB_OK;
}

int qsb_shard_allgather(qsb_plan* p, const double* send, int32_t count, double* recv) {
  if (!p || count < 0 || (count > 0 && (!send || !recv))) return fail(QSB_ERR_BAD_ARG, "null argument");
  if (p->world < 2 || !p->connected) return fail(QSB_ERR_BAD_ARG, "not a connected sharded plan");
  if (p->allgather(p->allgather_ctx, send, count, recv)) return fail(QSB_ERR_CUDA, "all-gather hook failed");
  return QSB_OK;
}

int qsb_local_group_create(int32_t world, qsb_local_group** group) {
  if (!group || world < 1 || world > MAX_WORLD) return fail(QSB_ERR_BAD_ARG, "bad local group arguments");
  qsb_local_group* g = new qsb_local_group;
  g->world = world;
  g->members.resize(world);
  for (int r = 0; r < world; ++r) g->members[r] = {g, r};
  g->buf.assign((size_t)world * PSTRIDE * 2, 0.0);
  if (const char* e = getenv("QSB200_BARRIER_TIMEOUT_S")) { double t = atof(e); if (t > 0.0) g->timeout_s = t; }
  *group = g;
  return QSB_OK;
}

int qsb_local_group_destroy(qsb_local_group* g) {
  delete g;
  return QSB_OK;
}

void* qsb_local_group_member(qsb_local_group* g, int32_t rank) {
  if (!g || rank < 0 || rank >= g->world) return nullptr;
  return &g->members[rank];
}

// A qsb_stream_barrier_fn for ranks that are threads of one process: it BLOCKS the host (drain the stream, then the
// thread barrier), which satisfies the hook's contract - work queued after the call starts only when every rank's
// earlier work on its counterpart stream is complete - without any kernel waiting on another (ranks may share a GPU).
// It serialises what the NCCL version overlaps; its purpose is to run the pipelined code path (chunk maps, persistent
// centre pass, event wiring) on a one-GPU box.
int qsb_local_stream_barrier(void* member_ctx, void* stream) {
  auto* m = (qsb_local_group::Member*)member_ctx;
  if (!m) return 1;
  const cudaError_t e = cudaStreamSynchronize((cudaStream_t)stream);
  const bool ok = m->g->barrier();
  if (e != cudaSuccess) { cudaGetLastError(); return 1; }
  return ok ? 0 : 1;
}

int qsb_local_allgather(void* member_ctx, const double* send, int32_t count, double* recv) {
  auto* m = (qsb_local_group::Member*)member_ctx;
  if (!m) return 1;
  qsb_local_group* g = m->g;
  if (count < 0 || (size_t)count * g->world > g->buf.size()) return 1;
  for (int k = 0; k < count; ++k) g->buf[(size_t)m->rank * count + k] = send[k];
  if (!g->barrier()) return 1;
  for (int k = 0; k < count * g->world; ++k) recv[k] = g->buf[k];
  if (!g->barrier()) return 1;  // nobody overwrites buf before everyone has read it
  return 0;
}

}  // extern "C"
