// Tiled HBM path kernels.  See qsb_large.cuh for the algebra; DESIGN.md for traffic accounting.
#include "qsb_large.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>

namespace qsb {

// ---------------------------------------------------------------------------------------------------
// Host: layouts.  Layout 0 = bits [0,12).  Remaining bits are split into groups of at most 8 (so that a
// group's Hadamard needs only the two coalescing-friendly windows), balanced; each group's tile is
// padded with the lowest 12-|g| bits so every tile is 4096 amplitudes made of >= 256-byte runs.
// ---------------------------------------------------------------------------------------------------
static void set_masks(Layout& L) {
  L.m8 = L.m4 = L.m0 = 0;
  int first = (L.ge > L.gs) ? L.lk : 0;  // layout 0 transforms all 12 bits
  if (L.ge == L.gs && L.lk == TB) first = 0;
  for (int b = first; b < TB; ++b) {
    int* m = b >= 8 ? &L.m8 : b >= 4 ? &L.m4 : &L.m0;
    *m |= 1 << (b & 3);
  }
}

LargeGeom make_geometry(int n) {
  LargeGeom g = {};
  g.n = n;
  Layout& L0 = g.lay[0];
  L0.lk = TB; L0.gs = TB; L0.ge = TB;
  L0.tile_mask = (1ull << TB) - 1;
  set_masks(L0);
  g.n_layouts = 1;
  int rest = n - TB;
  if (rest > 0) {
    int gmax = 8;  // test knob: smaller groups exercise the multi-group passes at small n
    if (const char* e = std::getenv("QSB200_MAX_GROUP")) gmax = std::max(1, std::min(8, atoi(e)));
    int ng = (rest + gmax - 1) / gmax;
    if (ng > MAX_LAYOUTS - 1) ng = MAX_LAYOUTS - 1;
    int start = TB;
    for (int k = 0; k < ng; ++k) {
      int sz = rest / ng + (k < rest % ng ? 1 : 0);
      Layout& L = g.lay[g.n_layouts++];
      L.lk = TB - sz; L.gs = start; L.ge = start + sz;
      L.tile_mask = ((1ull << L.lk) - 1) | (((1ull << sz) - 1) << start);
      set_masks(L);
      start += sz;
    }
  } else {
    // n == 12: the centre pass is layout 0 with nothing left to transform (diagonal multiply only)
    Layout& L = g.lay[g.n_layouts++];
    L = L0;
    L.m8 = L.m4 = L.m0 = 0;
  }
  return g;
}

static inline int global_bit(const Layout& L, int b) { return b < L.lk ? b : L.gs + (b - L.lk); }

std::vector<double> host_q_table(const double* h, const double* J, int n, int basis, const Layout& L, double scale) {
  std::vector<double> q(TILE);
  const double* hb = h + (size_t)basis * n;
  const double* Jb = J + (size_t)basis * n * n;
  for (int u = 0; u < TILE; ++u) {
    double acc = 0.0;
    for (int b = 0; b < TB; ++b) {
      int gb = global_bit(L, b);
      double sb = (u >> b & 1) ? -1.0 : 1.0;
      acc += 0.5 * hb[gb] * sb;
      for (int c = b + 1; c < TB; ++c) {
        int gc = global_bit(L, c);
        double sc = (u >> c & 1) ? -sb : sb;
        acc += 0.25 * Jb[(size_t)gb * n + gc] * sc;
      }
    }
    q[u] = acc * scale;
  }
  return q;
}

namespace {

// ---------------------------------------------------------------------------------------------------
// Device helpers
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ int phys(int i) { return i + (i >> 4); }

template <int W>
__device__ __forceinline__ int elem(int tt, int j) {
  return (tt & ((1 << W) - 1)) | (j << W) | ((tt >> W) << (W + 4));
}

__device__ __forceinline__ uint64_t tile_base(const Layout& L, uint64_t v) {
  int midbits = L.gs - L.lk;
  uint64_t mid = v & ((1ull << midbits) - 1), hi = v >> midbits;
  return (mid << L.lk) | (hi << L.ge);
}
__device__ __forceinline__ uint64_t gidx(const Layout& L, uint64_t base, int i) {
  return (uint64_t)(i & ((1 << L.lk) - 1)) | ((uint64_t)(i >> L.lk) << L.gs) | base;
}

// 16-point butterflies on the stages selected by mask (all unnormalised):
//   MODE 0  Hadamard           (a,b) -> (a+b, a-b)
//   MODE 1  M  = H S H         (a,b) -> (p+iq, p-iq),  p = a+b, q = a-b      [W^-1 S W = (M/2)^(x n)]
//   MODE 2  M' = H S^dagger H  (a,b) -> (p-iq, p+iq)
template <int MODE>
__device__ __forceinline__ void bfly16(cplx (&r)[16], int mask) {
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    if (mask >> s & 1) {
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        if (!(j >> s & 1)) {
          cplx a = r[j], b = r[j | (1 << s)];
          cplx p = a + b, q = a - b;
          if (MODE == 0) {
            r[j] = p;
            r[j | (1 << s)] = q;
          } else if (MODE == 1) {
            r[j] = mk(p.x - q.y, p.y + q.x);
            r[j | (1 << s)] = mk(p.x + q.y, p.y - q.x);
          } else {
            r[j] = mk(p.x + q.y, p.y - q.x);
            r[j | (1 << s)] = mk(p.x - q.y, p.y + q.x);
          }
        }
      }
    }
  }
}

template <int W>
__device__ __forceinline__ void smem_load16(const cplx* sm, int tt, cplx (&r)[16]) {
#pragma unroll
  for (int j = 0; j < 16; ++j) r[j] = sm[phys(elem<W>(tt, j))];
}
template <int W>
__device__ __forceinline__ void smem_store16(cplx* sm, int tt, const cplx (&r)[16]) {
#pragma unroll
  for (int j = 0; j < 16; ++j) sm[phys(elem<W>(tt, j))] = r[j];
}
template <int W, int MODE>
__device__ __forceinline__ void smem_round(cplx* sm, int tt, int mask) {
  cplx r[16];
  smem_load16<W>(sm, tt, r);
  bfly16<MODE>(r, mask);
  smem_store16<W>(sm, tt, r);
}

// Per-CTA part of a diagonal form D(u; v) = Q[u] + cst + LinLo[u & 63] + LinHi[u >> 6]  (see DESIGN.md):
// lin_b(v) = scale/4 * sum_{c outer} J[gb(b)][c] s_c(v),  cst(v) = scale * (sum_c h_c s_c/2 + sum_{c<c'} J s s'/4)
// over outer bits only.  Needs >= 128 threads; ends with __syncthreads().
struct DiagShared {
  double lin[TB];
  double lo[64], hi[64];
  double cst;
};
__device__ void diag_prologue(DiagShared& ds, const double* __restrict__ h, const double* __restrict__ J, int n,
                              const Layout& L, uint64_t base, double scale, int tt) {
  const uint64_t outer = ~L.tile_mask & ((n >= 64) ? ~0ull : ((1ull << n) - 1));
  const int lane = tt & 31, warp = tt >> 5;
  if (warp == 0) {
    if (lane < TB) {
      int gb = lane < L.lk ? lane : L.gs + (lane - L.lk);
      double acc = 0.0;
      for (int c = 0; c < n; ++c)
        if (outer >> c & 1) acc = fma((base >> c & 1) ? -0.25 : 0.25, J[(size_t)gb * n + c], acc);
      ds.lin[lane] = acc * scale;
    }
  } else if (warp == 1) {
    double part = 0.0;
    for (int c = lane; c < n; c += 32) {
      if (!(outer >> c & 1)) continue;
      double sc = (base >> c & 1) ? -1.0 : 1.0;
      double acc = 0.5 * h[c];
      for (int c2 = c + 1; c2 < n; ++c2)
        if (outer >> c2 & 1) acc = fma((base >> c2 & 1) ? -0.25 : 0.25, J[(size_t)c * n + c2], acc);
      part = fma(sc, acc, part);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) part += __shfl_xor_sync(0xffffffffu, part, off);
    if (lane == 0) ds.cst = part * scale;
  }
  __syncthreads();
  if (tt < 128) {
    int a = tt & 63, half = tt >> 6;
    double acc = 0.0;
#pragma unroll
    for (int b = 0; b < 6; ++b) acc = fma((a >> b & 1) ? -1.0 : 1.0, ds.lin[half * 6 + b], acc);
    (half ? ds.hi : ds.lo)[a] = acc;
  }
  __syncthreads();
}
__device__ __forceinline__ double diag_at(const DiagShared& ds, const double* __restrict__ Q, int u) {
  return __ldg(Q + u) + ds.cst + ds.lo[u & 63] + ds.hi[u >> 6];
}

__device__ __forceinline__ int sea_weight(uint64_t idx, uint64_t sea_mask, int n_sea) {
  return n_sea - 2 * __popcll(idx & sea_mask);
}

}  // namespace
}  // namespace qsb

#include "qsb_large_kernels.inl"
