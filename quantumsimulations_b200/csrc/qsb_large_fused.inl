// L2-resident fusion of the strided passes (mid ... centre ... mid) of one H-apply; included from
// qsb_large_kernels.inl.
//
// The strided layouts f .. centre all leave the same index bits untouched: the low pad bits [lkmax, 12) of their
// tiles and the groups of the layouts before f.  Fixing those bits selects a *unit*: 2^(n - fbits) amplitudes (times
// two branches) that the 2 m + 1 passes f, f+1, .., centre, .., f+1, f transform among themselves.  A unit is sized to
// sit in L2 (16 MB at N = 28 against 126 MB), so only the first pass of a unit reads HBM and only the last one's
// stores have to reach it: 4 S of DRAM traffic instead of 4 S per pass (measured: 2.0 S read, 2.35 S written).
//
// One persistent launch, ONE CTA per SM, walks a software pipeline of tile tasks in ticket order.  Slot s holds, in
// this order, the tiles of phase P-1 of unit s - (P-1) lag, .., phase 1 of unit s - lag, phase 0 of unit s; phase p of
// a unit starts when all T tiles of its phase p - 1 have been published (a counter per unit and phase, release /
// acquire at gpu scope).  Every dependency has a smaller ticket than its dependant and the thread that waits for a
// dependency never stops storing and publishing the CTA's own finished tiles, so the pipeline cannot deadlock.
//
// Inside the CTA the tile moves asynchronously: a producer thread draws tickets and issues ONE TMA tensor load per
// tile (a strided tile is a 5-d box of the state vector: pad runs x group) into a ring of three 64 KB shared-memory
// buffers, NG groups of 256 threads transform landed tiles in place (the butterflies and diagonal of pass_mid /
// pass_centre, register windows [8,12) and [4,8) only, which need no bank-conflict padding), and the producer sends a
// finished buffer back with ONE TMA tensor store.  Global latency is therefore never exposed to the arithmetic, which
// is what capped the register-staged passes at 2 CTAs x 8 warps per SM.
//
// The arithmetic per tile is exactly that of pass_mid / pass_centre, so results are bitwise those of the unfused path.
namespace qsb {
namespace {

constexpr int NBUF = 3;          // tile buffers per CTA
#ifndef QSB_FUSED_NG
#define QSB_FUSED_NG 2
#endif
constexpr int FUSED_NG = QSB_FUSED_NG;  // compute groups of TPT threads (2: 17 warps -> 96 registers per thread)
constexpr int TILE_BYTES = TILE * (int)sizeof(cplx);
constexpr int ROW_BYTES = TD_STRIDE * (int)sizeof(double);

constexpr int MAX_MAPS = MAX_WORLD > MAX_LAYOUTS ? MAX_WORLD : MAX_LAYOUTS;
struct alignas(64) TmaMaps {
  // fused strided pass: [branch][layout], the branch's scratch vector as a 5-d tensor of doubles whose box is a tile;
  // sharded centre ring: [branch][peer], the PEER's scratch vector, box = the slab of a centre tile that peer holds
  CUtensorMap m[2][MAX_MAPS];
};

struct TaskDesc {
  int kind;      // 0 mid, 1 centre, 2 exit
  int li;        // layout
  int branch;
  unsigned v;    // tile number (index arithmetic)
  unsigned slot; // this rank's number for the tile (diagonal rows, partial sums); = v when unsharded
  int c2, c4;    // its TMA box coordinates (dims 2 and 4; the others are 0)
  int ctr;       // counter to bump when the tile's store has completed (-1: nobody waits)
  int map0, n_parts;  // the tile is moved as n_parts boxes through maps m[branch][map0 .. map0 + n_parts)
};

struct FusedSmem {
  cplx buf[NBUF][TILE];
  double intile[2][5 * 256];      // factored in-tile diagonal tables of the centre layout, x and y basis
  double row[NBUF][TD_STRIDE];    // per-tile diagonal row of the centre tile in the buffer
  double red[FUSED_NG][2 * 32 + 2];
  unsigned long long full[NBUF];      // tile (and row) landed
  unsigned long long computed[NBUF];  // all TPT threads of the computing group are done with the buffer
  unsigned long long empty[NBUF];     // the buffer's TMA store has read it: free for the next load
  int final_count;                    // tiles this CTA loaded in total, once the tickets have run out (-1 before)
  TaskDesc desc[NBUF];
};

__device__ __forceinline__ unsigned deposit_bits(unsigned x, unsigned mask) {  // software pdep
  unsigned out = 0;
  while (mask) {
    const unsigned low = mask & (0u - mask);
    if (x & 1u) out |= low;
    x >>= 1;
    mask ^= low;
  }
  return out;
}

__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(void* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(void* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(void* bar, int bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_test(void* bar, int parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
// Every wait of the fused pass is bounded: a wait that lasts longer than WATCHDOG_NS is a broken pipeline, and a trap
// (the launch fails with a CUDA error) is better than a GPU that never comes back.
constexpr unsigned long long WATCHDOG_NS = 20ull * 1000 * 1000 * 1000;
__device__ __forceinline__ unsigned long long now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
struct Watchdog {
  unsigned long long t0 = 0;
  unsigned spins = 0;
  __device__ __forceinline__ void tick() {
    if ((++spins & 1023u) == 0) {
      const unsigned long long t = now_ns();
      if (!t0) t0 = t;
      else if (t - t0 > WATCHDOG_NS) __trap();
    }
  }
};
__device__ __forceinline__ void mbar_wait(void* bar, int parity) {
  uint32_t ok;
  Watchdog wd;
  do {
    wd.tick();
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void tma_load_5d(void* dst, const CUtensorMap* map, void* bar, int c2, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %3, %4, %3, %5}], [%2];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(0), "r"(c2), "r"(c4) : "memory");
}
__device__ __forceinline__ void tma_store_5d(const CUtensorMap* map, const void* src, int c2, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.global.shared::cta.bulk_group [%0, {%2, %2, %3, %2, %4}], [%1];"
      ::"l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(src)), "r"(0), "r"(c2), "r"(c4) : "memory");
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, int bytes, void* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void group_bar(int g) { asm volatile("bar.sync %0, %1;" ::"r"(1 + g), "r"(TPT) : "memory"); }

// block_reduce<2> for one compute group: the same shuffle trees, so the sums are bitwise those of pass_centre
__device__ __forceinline__ void group_reduce2(double (&v)[2], double* scratch, int tt, int g) {
  const int lane = tt & 31, warp = tt >> 5;
#pragma unroll
  for (int k = 0; k < 2; ++k) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], off);
  }
  group_bar(g);
  if (lane == 0) { scratch[warp] = v[0]; scratch[32 + warp] = v[1]; }
  group_bar(g);
  if (warp == 0) {
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      double x = lane < TPT / 32 ? scratch[k * 32 + lane] : 0.0;
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
      v[k] = x;
    }
  }
}

// in-place Hadamard on one group's bits of a tile that sits (unpadded) in shared memory
__device__ __forceinline__ void smem_mid_tile(const Layout& L, cplx* sm, int tt, int g) {
  cplx r[16];
  tile_load16<8>(sm, tt, r);
  bfly16<0>(r, L.m8);
  tile_store16<8>(sm, tt, r);
  if (L.m4) {
    group_bar(g);
    tile_load16<4>(sm, tt, r);
    bfly16<0>(r, L.m4);
    tile_store16<4>(sm, tt, r);
  }
}

// t <- W_g D W_g t on one centre tile in shared memory (W: the register window the diagonal is evaluated in)
template <int W, bool COLLECT>
__device__ __forceinline__ void smem_centre_tile(const PassArgs& a, const Layout& L, uint64_t v, unsigned slot, int branch,
                                                 cplx* sm, const double* row, const double* intile, double* red, int tt,
                                                 int g, double* __restrict__ partials) {
  const double (&ww)[16] = branch ? a.tab.Wy : a.tab.Wx;
  TileDiag<W> td;
  td.load_shared(row, intile, tt);
  const uint64_t base = tile_base(L, v);
  cplx r[16];
  double s[2] = {0.0, 0.0};
  tile_load16<8>(sm, tt, r);
  if constexpr (W == 8) {
    centre_core<8, COLLECT>(r, tt, L.m8, td, ww, L, base, a.g, s);
  } else {
    bfly16<0>(r, L.m8);
    tile_store16<8>(sm, tt, r);
    group_bar(g);
    tile_load16<4>(sm, tt, r);
    centre_core<4, COLLECT>(r, tt, L.m4, td, ww, L, base, a.g, s);
    tile_store16<4>(sm, tt, r);
    group_bar(g);
    tile_load16<8>(sm, tt, r);
    bfly16<0>(r, L.m8);
  }
  tile_store16<8>(sm, tt, r);
  if (COLLECT) {
    group_reduce2(s, red, tt, g);
    if (tt == 0) {
      double* p = partials + (size_t)slot * PSTRIDE + 3 + 2 * branch;
      p[0] = s[0]; p[1] = s[1];
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// The three roles of a ring CTA.  Block: FUSED_NG * TPT compute threads + a loader warp + a storer warp (lane 0 of
// each works).
//   loader : (next task) -> buffer free (empty[b]) -> descriptor, TMA load(s) (full[b])
//   compute: full[b] -> transform in place -> computed[b]
//   storer : computed[b] -> TMA store(s) -> empty[b] once the store has read the buffer -> publishes the PREVIOUS
//            tile once its store has completed (so a completion wait never sits between two stores)
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void ring_init(FusedSmem& S, const PassArgs& a, int tid) {
  for (int i = tid; i < 5 * 256; i += blockDim.x) {
    S.intile[0][i] = a.tab.Ix[i];
    S.intile[1][i] = a.tab.Iy[i];
  }
  if (tid == 0) {
    for (int b = 0; b < NBUF; ++b) { mbar_init(&S.full[b], 1); mbar_init(&S.computed[b], TPT); mbar_init(&S.empty[b], 1); }
    S.final_count = -1;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async;" ::: "memory");
  }
  __syncthreads();
}

// Loader side of one task: wait for the buffer, publish the descriptor, issue the loads.
__device__ __forceinline__ void ring_issue(FusedSmem& S, const PassArgs& a, const TmaMaps& maps, int seq, const TaskDesc& task) {
  const int b = seq % NBUF, round = seq / NBUF;
  if (round > 0) mbar_wait(&S.empty[b], (round - 1) & 1);
  S.desc[b] = task;
  const bool is_centre = task.kind == 1;
  mbar_arrive_expect_tx(&S.full[b], TILE_BYTES + (is_centre ? ROW_BYTES : 0));
  const int part = TILE / task.n_parts;
  for (int r = 0; r < task.n_parts; ++r)
    tma_load_5d(S.buf[b] + r * part, &maps.m[task.branch][task.map0 + r], &S.full[b], task.c2, task.c4);
  if (is_centre)
    bulk_load(S.row[b], (task.branch ? a.tab.Ty : a.tab.Tx) + (size_t)task.slot * TD_STRIDE, ROW_BYTES, &S.full[b]);
}

__device__ __forceinline__ void ring_finish_loads(FusedSmem& S, int seq) {
  *(volatile int*)&S.final_count = seq;
  for (int g = 0; g < FUSED_NG; ++g) {  // one exit marker per compute group
    const int b = seq % NBUF, round = seq / NBUF;
    if (round > 0) mbar_wait(&S.empty[b], (round - 1) & 1);
    S.desc[b].kind = 2;
    mbar_arrive(&S.full[b]);
    ++seq;
  }
}

// ctrl == nullptr: nothing is published (no other CTA depends on this one's tiles).
__device__ __forceinline__ void ring_storer(FusedSmem& S, const TmaMaps& maps, int* ctrl) {
  // prev_ctr: the counter of the last stored tile, still to be bumped once that store has completed.  While stores
  // follow each other the completion wait rides behind the next store; when the storer idles it publishes at once
  // (the tile some loader is waiting for may depend on exactly this publication).
  int prev_ctr = -1;
  Watchdog wd;
  for (int seq = 0;; ++seq) {
    const int b = seq % NBUF, par = (seq / NBUF) & 1;
    bool done = false;
    while (!mbar_test(&S.computed[b], par)) {
      if (prev_ctr >= 0) {
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        asm volatile("fence.proxy.async;" ::: "memory");
        __threadfence();
        atomicAdd(ctrl + prev_ctr, 1);
        prev_ctr = -1;
      }
      const int fc = *(volatile int*)&S.final_count;
      if (fc >= 0 && seq >= fc) { done = true; break; }
      wd.tick();
    }
    if (done) break;
    const TaskDesc d = S.desc[b];
    const int part = TILE / d.n_parts;
    for (int r = 0; r < d.n_parts; ++r) tma_store_5d(&maps.m[d.branch][d.map0 + r], S.buf[b] + r * part, d.c2, d.c4);
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    mbar_arrive(&S.empty[b]);
    if (seq > 0 && prev_ctr >= 0) {
      asm volatile("cp.async.bulk.wait_group 1;" ::: "memory");
      asm volatile("fence.proxy.async;" ::: "memory");
      __threadfence();
      atomicAdd(ctrl + prev_ctr, 1);
    }
    prev_ctr = ctrl ? d.ctr : -1;
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // the kernel must not end with stores in flight
  if (prev_ctr >= 0) {
    asm volatile("fence.proxy.async;" ::: "memory");
    __threadfence();
    atomicAdd(ctrl + prev_ctr, 1);
  }
}

template <int W, bool COLLECT>
__device__ __forceinline__ void ring_compute(FusedSmem& S, const PassArgs& a, int tid, double* __restrict__ partials) {
  const int g = tid / TPT, tt = tid - g * TPT;
  for (int seq = g;; seq += FUSED_NG) {
    const int b = seq % NBUF;
    mbar_wait(&S.full[b], (seq / NBUF) & 1);
    const TaskDesc d = S.desc[b];
    if (d.kind == 2) break;
    if (d.kind == 1)
      smem_centre_tile<W, COLLECT>(a, a.g.lay[d.li], d.v, d.slot, d.branch, S.buf[b], S.row[b], S.intile[d.branch], S.red[g],
                                   tt, g, partials);
    else
      smem_mid_tile(a.g.lay[d.li], S.buf[b], tt, g);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // the tile is about to be read by the TMA store
    mbar_arrive(&S.computed[b]);
  }
}

// ---------------------------------------------------------------------------------------------------
// pass_strided_fused: the L2-resident mid .. centre .. mid pipeline of an unsharded plan.
// ctrl[0]: ticket counter; ctrl[1 + unit * P + phase]: published tiles.  Zeroed by the launcher before every launch.
// ---------------------------------------------------------------------------------------------------
template <int W, bool COLLECT>
__global__ void __launch_bounds__(FUSED_NG * TPT + 64, 1)
pass_strided_fused(PassArgs a, FusedSpec fs, const __grid_constant__ TmaMaps maps, double* __restrict__ partials,
                   int* ctrl) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FusedSmem& S = *reinterpret_cast<FusedSmem*>(smem_raw);
  const int tid = threadIdx.x;
  ring_init(S, a, tid);
  if (tid == FUSED_NG * TPT) {
    const int centre = fs.f + fs.m;
    const int P = 2 * fs.m + 1;
    const int T = 1 << fs.tiles_lb;
    const int U = 2 << fs.fbits;  // unit = (fixed bits, branch); branch is the lowest bit of the unit number
    const int per_slot = P << fs.tiles_lb;
    const int total = (U + (P - 1) * fs.lag) * per_slot;
    int seq = 0;                          // tiles loaded by this CTA; buffer = seq % NBUF
    int next_k = atomicAdd(&ctrl[0], 1);  // always one ticket ahead: the draw's latency hides behind the waits below
    while (true) {
      int k, p = 0, u = 0, q = 0;
      bool valid = false;
      while (true) {
        k = next_k;
        if (k >= total) break;
        next_k = atomicAdd(&ctrl[0], 1);
        const int slot = k / per_slot, rr = k - slot * per_slot;
        p = P - 1 - (rr >> fs.tiles_lb);
        q = rr & (T - 1);
        u = slot - p * fs.lag;
        if (u >= 0 && u < U) { valid = true; break; }
      }
      if (!valid) break;
      TaskDesc d;
      d.branch = u & 1;
      d.li = fs.f + (p <= fs.m ? p : 2 * fs.m - p);
      const unsigned fm = fs.fixed_vmask[d.li];
      d.v = deposit_bits((unsigned)(u >> 1), fm) | deposit_bits((unsigned)q, fs.vmask & ~fm);
      d.slot = d.v;
      if (p > 0) {  // every tile of the unit's previous phase must have been stored and published
        const int* flag = ctrl + 1 + u * P + (p - 1);
        Watchdog wd;
        while (ld_acquire(flag) < T) { __nanosleep(32); wd.tick(); }
        asm volatile("fence.proxy.async;" ::: "memory");
      }
      const Layout& L = a.g.lay[d.li];
      const int midbits = L.gs - L.lk;
      d.kind = d.li == centre ? 1 : 0;
      d.c2 = (int)(d.v & ((1u << midbits) - 1));
      d.c4 = (int)(d.v >> midbits);
      d.ctr = p < P - 1 ? 1 + u * P + p : -1;
      d.map0 = d.li;
      d.n_parts = 1;
      ring_issue(S, a, maps, seq, d);
      ++seq;
    }
    ring_finish_loads(S, seq);
  } else if (tid == FUSED_NG * TPT + 32) {
    ring_storer(S, maps, ctrl);
  } else if (tid < FUSED_NG * TPT) {
    ring_compute<W, COLLECT>(S, a, tid, partials);
  }
}

// ---------------------------------------------------------------------------------------------------
// pass_centre_ring: the sharded centre pass (the one exchange step of a sharded H-apply) on the same ring.
// A centre tile has one slab of TILE / world amplitudes on every rank (the rank bits are the top bits of the tile);
// rank r owns tiles v0 .. v0 + n_tiles_loc of the centre layout and both branches.  The loader pulls the world slabs
// of a tile straight out of the peers' scratch vectors with one TMA tensor load each (NVLink reads issued by the copy
// engine, never waited for by arithmetic), the compute groups apply W_g D W_g in shared memory, and the storer
// pushes the slabs back in place with one TMA store each: exchange and compute are one kernel, and up to three tiles
// per SM are in flight on the links at any time.  Tiles are dealt round-robin to the CTAs; nothing is published
// (the rank barriers around the pass order it against the peers' neighbouring passes).
// ---------------------------------------------------------------------------------------------------
template <int W, bool COLLECT>
__global__ void __launch_bounds__(FUSED_NG * TPT + 64, 1)
pass_centre_ring(PassArgs a, int li, unsigned v0, unsigned n_tiles_loc, int world, const __grid_constant__ TmaMaps maps,
                 double* __restrict__ partials) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FusedSmem& S = *reinterpret_cast<FusedSmem*>(smem_raw);
  const int tid = threadIdx.x;
  ring_init(S, a, tid);
  if (tid == FUSED_NG * TPT) {
    int seq = 0;
    for (unsigned i = blockIdx.x; i < 2u * n_tiles_loc; i += gridDim.x, ++seq) {
      TaskDesc d;
      d.kind = 1;
      d.li = li;
      d.branch = (int)(i & 1u);   // x and y tiles alternate: both branches' slabs share the peers' pages
      d.slot = i >> 1;
      d.v = v0 + d.slot;
      d.c2 = (int)d.v;            // the centre group reaches the top index bit: every outer bit of the tile is a "mid" bit
      d.c4 = 0;
      d.ctr = -1;
      d.map0 = 0;
      d.n_parts = world;
      ring_issue(S, a, maps, seq, d);
    }
    ring_finish_loads(S, seq);
  } else if (tid == FUSED_NG * TPT + 32) {
    ring_storer(S, maps, nullptr);
  } else if (tid < FUSED_NG * TPT) {
    ring_compute<W, COLLECT>(S, a, tid, partials);
  }
}

}  // namespace
}  // namespace qsb
