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

LargeGeom make_geometry(int n, int rank, int world) {
  LargeGeom g = {};
  int wl = 0;
  while ((1 << wl) < world) ++wl;
  g.n = n;
  g.n_loc = n - wl;
  g.rank = rank;
  g.world = world;
  g.rank_bits = (uint64_t)rank << g.n_loc;
  Layout& L0 = g.lay[0];
  L0.lk = TB; L0.gs = TB; L0.ge = TB;
  L0.tile_mask = (1ull << TB) - 1;
  set_masks(L0);
  g.n_layouts = 1;
  int rest = n - TB;
  if (rest > 0) {
    int gmax = 8;  // test knob: smaller groups exercise the multi-group passes at small n
    if (const char* e = std::getenv("QSB200_MAX_GROUP")) gmax = std::max(std::max(1, wl), std::min(8, atoi(e)));
    // group sizes, ascending bit order; the LAST group must contain all wl rank bits (it becomes the centre, the
    // only pass that reaches the peers), so it is sized first and the other bits are balanced over the rest
    int sizes[MAX_LAYOUTS] = {0};
    // Unsharded plans take 9-bit groups (tiles of 128-byte runs: pass_mid 6.1 instead of 6.7 TB/s,
    // profiles/r2_short_run_pass.json) where that saves a whole layout, i.e. two more HBM passes: N = 29 and 30
    // become 12 | 9 | 8 and 12 | 9 | 9 (19 S per H-apply) instead of four layouts (27 S).
    // (Not with the opt-in L2-fused strided pass, QSB200_FUSE_LB, which is written for groups of at most 8 bits.)
    if (world == 1 && gmax == 8 && !std::getenv("QSB200_MAX_GROUP") && !std::getenv("QSB200_FUSE_LB") &&
        (rest + 8) / 9 < (rest + 7) / 8)
      gmax = 9;
    int ng = (rest + gmax - 1) / gmax;
    if (ng > MAX_LAYOUTS - 1) ng = MAX_LAYOUTS - 1;
    int last = std::max(wl, (rest + ng - 1) / ng);
    if (world > 1 && ng * gmax >= rest) {  // (not when MAX_LAYOUTS already forces oversized groups: test knob only)
      // Sharded: the centre pass is NVLink-bound and its P2P throughput falls with the number of far-apart rows a
      // tile gathers (2^last rows, 2^(n_loc ...) bytes apart): measured at 2^30 amplitudes per GPU on 2 B200s, 3-5
      // group bits reach 91 % of the peer-copy peak, 6 bits 82 %, 7 bits 65 %, 8 bits 49 %
      // (profiles/r1zz_centre_group_2gpu.log).  So the centre group is capped at 5 bits wherever that does not add
      // a layout (= one more pair of HBM passes).
      const int need = rest - (ng - 1) * gmax;  // fewest bits the last group can take with ng groups
      last = std::max(std::max(wl, need), std::min(last, 5));
    }
    // experiment knob: size of the centre group (rows per centre tile = 2^last, run length 2^(12-last) amplitudes)
    if (const char* e = std::getenv("QSB200_CENTRE_GROUP")) last = std::max(std::max(1, wl), std::min(gmax, atoi(e)));
    if (last > rest) last = rest;
    int others = rest - last;
    ng = 1 + (others + gmax - 1) / gmax;
    if (ng > MAX_LAYOUTS - 1) ng = MAX_LAYOUTS - 1;
    for (int k = 0; k + 1 < ng; ++k) sizes[k] = others / (ng - 1) + (k < others % (ng - 1) ? 1 : 0);
    sizes[ng - 1] = last;
    if (world == 1 && ng > 1) {  // the centre (last) is the heavier kernel: give it the smallest group, the mids the wide ones
      int lo = 0;
      for (int k = 1; k < ng; ++k)
        if (sizes[k] < sizes[lo]) lo = k;
      std::swap(sizes[lo], sizes[ng - 1]);
    }
    // experiment knob: explicit group sizes in ascending bit order, e.g. QSB200_GROUPS=10,6 (each 1..12, summing to
    // n - 12, the last one holding all rank bits).  Groups wider than 8 bits give tiles made of runs shorter than 256
    // bytes (lk = 12 - size pad bits); that is how the short-run passes of DESIGN.md section 10 were timed.
    if (const char* e = std::getenv("QSB200_GROUPS")) {
      int cand[MAX_LAYOUTS] = {0}, nc = 0, sum = 0;
      bool ok = true;
      for (const char* q = e; *q && ok;) {
        char* end = nullptr;
        long v = strtol(q, &end, 10);
        if (end == q || v < 1 || v > TB || nc >= MAX_LAYOUTS - 1) { ok = false; break; }
        cand[nc++] = (int)v;
        sum += (int)v;
        q = (*end == ',') ? end + 1 : end;
        if (*end && *end != ',') ok = false;
      }
      if (ok && nc > 0 && sum == rest && cand[nc - 1] >= wl) {
        ng = nc;
        for (int k = 0; k < ng; ++k) sizes[k] = cand[k];
      }
    }
    int start = TB;
    for (int k = 0; k < ng; ++k) {
      int sz = sizes[k];
      Layout& L = g.lay[g.n_layouts++];
      L.lk = TB - sz; L.gs = start; L.ge = start + sz;
      L.tile_mask = ((1ull << L.lk) - 1) | (((1ull << sz) - 1) << start);
      set_masks(L);
      start += sz;
    }
    // Unsharded plans may apply the diagonal on ANY group (only the rank bits tie a sharded centre to the top one).
    // QSB200_CENTRE_LOW=1 makes the lowest strided group the centre, so that the far-strided top group is served by
    // the lighter pass_mid (experiment knob).
    if (world == 1 && g.n_layouts > 2) {
      const char* e = std::getenv("QSB200_CENTRE_LOW");
      if (e && *e == '1') {
        const Layout low = g.lay[1];
        for (int k = 1; k + 1 < g.n_layouts; ++k) g.lay[k] = g.lay[k + 1];
        g.lay[g.n_layouts - 1] = low;
      }
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

int centre_window(const Layout& L) { return (!L.m4 && !L.m0) ? 8 : (!L.m0 ? 4 : 0); }

InTileDiag host_in_tile_diag(const double* h, const double* J, int n, int basis, const Layout& L, int W, double scale) {
  InTileDiag d;
  const double* hb = h + (size_t)basis * n;
  const double* Jb = J + (size_t)basis * n * n;
  auto Jg = [&](int b, int c) { return Jb[(size_t)global_bit(L, b) * n + global_bit(L, c)]; };
  for (int tt = 0; tt < 256; ++tt) {
    // elem<W>(tt, 0): tt's low W bits are tile bits [0, W), the rest tile bits [W + 4, 12)
    const int i0 = (tt & ((1 << W) - 1)) | ((tt >> W) << (W + 4));
    auto sgn = [&](int b) { return (i0 >> b & 1) ? -1.0 : 1.0; };
    auto in_window = [&](int b) { return b >= W && b < W + 4; };
    double A = 0.0;
    for (int b = 0; b < TB; ++b) {
      if (in_window(b)) continue;
      A += 0.5 * hb[global_bit(L, b)] * sgn(b);
      for (int c = b + 1; c < TB; ++c)
        if (!in_window(c)) A += 0.25 * Jg(b, c) * sgn(b) * sgn(c);
    }
    d.tab[tt] = A * scale;
    for (int k = 0; k < 4; ++k) {
      const int wb = W + k;
      double l = 0.5 * hb[global_bit(L, wb)];
      for (int c = 0; c < TB; ++c)
        if (!in_window(c)) l += 0.25 * Jg(wb, c) * sgn(c);
      d.tab[(1 + k) * 256 + tt] = l * scale;
    }
  }
  for (int j = 0; j < 16; ++j) {
    double acc = 0.0;
    for (int k = 0; k < 4; ++k)
      for (int k2 = k + 1; k2 < 4; ++k2)
        acc += 0.25 * Jg(W + k, W + k2) * ((j >> k & 1) ? -1.0 : 1.0) * ((j >> k2 & 1) ? -1.0 : 1.0);
    d.ww[j] = acc * scale;
  }
  return d;
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
// Hides a value's provenance from the optimiser.  The store addresses of a pass equal its load addresses; without
// this the compiler keeps all sixteen 64-bit addresses live across the whole kernel (32 registers -> spills).
// Index offsets of the four window bits [W, W+4) of a tile in layout L: the 16 amplitudes a thread holds for window
// W sit at idx0 + (sum of off[k] over the set bits of j), which costs 8 registers instead of 16 addresses.
struct WinOff {
  uint64_t off[4];
  // late0: 0, optionally a "late zero" (see pass_mid / pass_centre) that keeps the offsets for a pass's stores from
  // being shared with (and kept alive since) the ones for its loads
  template <int W>
  __device__ __forceinline__ void init(const Layout& L, unsigned late0) {
    const int lk = L.lk + (int)late0, gs = L.gs;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int b = W + k;
      off[k] = 1ull << (b < lk ? b : gs + (b - lk));
    }
  }
  __device__ __forceinline__ uint64_t at(int j) const {  // j: compile-time constant after unrolling
    uint64_t o = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (j >> k & 1) o += off[k];
    return o;
  }
};

__device__ __forceinline__ uint64_t gidx(const Layout& L, uint64_t base, int i) {
  return (uint64_t)(i & ((1 << L.lk) - 1)) | ((uint64_t)(i >> L.lk) << L.gs) | base;
}

// 16-point butterflies on the stages selected by mask (all unnormalised):
//   MODE 0  Hadamard            (a,b) -> (a+b, a-b)
//   MODE 1  K  = M  / (1+i)     (a,b) -> (a - i b, b - i a)      M  = H S H         [W^-1 S W = (M/2)^(x n)]
//   MODE 2  K' = M' / (1-i)     (a,b) -> (a + i b, b + i a)      M' = H S^dagger H
// M = [[1+i, 1-i], [1-i, 1+i]] = (1+i) [[1, -i], [-i, 1]], so the scalar is pulled out of every stage: a K stage costs
// the four additions of a Hadamard stage (M done literally costs eight), and over the twelve bits of a tile the
// scalars multiply to the REAL number (1+i)^12 = (1-i)^12 = -64, which the callers fold into their 2^-12 (M_SCALE).
template <int MODE>
__device__ __forceinline__ void bfly16(cplx (&r)[16], int mask) {
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    if (mask >> s & 1) {
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        if (!(j >> s & 1)) {
          cplx a = r[j], b = r[j | (1 << s)];
          if (MODE == 0) {
            r[j] = a + b;
            r[j | (1 << s)] = a - b;
          } else if (MODE == 1) {
            r[j] = mk(a.x + b.y, a.y - b.x);
            r[j | (1 << s)] = mk(b.x + a.y, b.y - a.x);
          } else {
            r[j] = mk(a.x - b.y, a.y + b.x);
            r[j | (1 << s)] = mk(b.x - a.y, b.y + a.x);
          }
        }
      }
    }
  }
}

// Padded shared-memory position of elem<W>(tt, j): phys(elem) is linear in j because the pad (one slot per 16)
// never splits a window, so a thread addresses its 16 slots as ONE base + compile-time offsets j * PS<W>
// (no per-element address registers).
template <int W>
struct PhysStride { static constexpr int value = (W == 8) ? 256 + 16 : (W == 4) ? 16 + 1 : 1; };
template <int W>
__device__ __forceinline__ int phys_base(int tt) { return phys(elem<W>(tt, 0)); }

template <int W>
__device__ __forceinline__ void smem_load16(const cplx* sm, int tt, cplx (&r)[16]) {
  const cplx* p = sm + phys_base<W>(tt);
#pragma unroll
  for (int j = 0; j < 16; ++j) r[j] = p[j * PhysStride<W>::value];
}
template <int W>
__device__ __forceinline__ void smem_store16(cplx* sm, int tt, const cplx (&r)[16]) {
  cplx* p = sm + phys_base<W>(tt);
#pragma unroll
  for (int j = 0; j < 16; ++j) p[j * PhysStride<W>::value] = r[j];
}
// The same accesses on an UNPADDED tile (element i at slot i).  Windows [8,12) and [4,8) are conflict-free without
// the pad (a quarter-warp reads 8 consecutive 16-byte slots); only window [0,4) needs it.  Passes that never touch
// window [0,4) - every strided pass, whose low >= 4 tile bits are pad bits - use these and 64 KB of shared memory
// instead of 69.6 KB: two CTAs then fit the 132 KB carve-out step instead of 164 KB, which leaves 32 KB more L1 for
// loads in flight.
template <int W>
__device__ __forceinline__ void tile_load16(const cplx* sm, int tt, cplx (&r)[16]) {
  const cplx* p = sm + elem<W>(tt, 0);
#pragma unroll
  for (int j = 0; j < 16; ++j) r[j] = p[j << W];
}
template <int W>
__device__ __forceinline__ void tile_store16(cplx* sm, int tt, const cplx (&r)[16]) {
  cplx* p = sm + elem<W>(tt, 0);
#pragma unroll
  for (int j = 0; j < 16; ++j) p[j << W] = r[j];
}
// PAD selects between the two at compile time
template <int W, bool PAD>
__device__ __forceinline__ void any_load16(const cplx* sm, int tt, cplx (&r)[16]) {
  if constexpr (PAD) smem_load16<W>(sm, tt, r); else tile_load16<W>(sm, tt, r);
}
template <int W, bool PAD>
__device__ __forceinline__ void any_store16(cplx* sm, int tt, const cplx (&r)[16]) {
  if constexpr (PAD) smem_store16<W>(sm, tt, r); else tile_store16<W>(sm, tt, r);
}

template <int W, int MODE>
__device__ __forceinline__ void smem_round(cplx* sm, int tt, int mask) {
  cplx r[16];
  smem_load16<W>(sm, tt, r);
  bfly16<MODE>(r, mask);
  smem_store16<W>(sm, tt, r);
}

// Diagonal form on a tile:  D(u; v) = D_in(u) + cst(v) + sum_b s_b(u) lin_b(v)   (u = tile-local index, v = tile).
// The tile-dependent part is stored per tile, already folded for the register window W the consuming pass uses
// (one 320-byte row per tile, built once per plan by build_tile_diag below):
//   row[0..15]  F_lo[tt & 15]  = cst + sum over the tile bits fed by tt's low nibble  of +-lin_b
//   row[16..31] F_hi[tt >> 4]  =       sum over the tile bits fed by tt's high nibble of +-lin_b
//   row[32..35] lin of the four window bits
// so a thread needs 2 + 4 row values and its 5 in-tile table values: 11 loads, ~22 transient registers.
constexpr int TD_STRIDE = 40;

// Everything of D that a thread can hoist out of its 16-element loop for window W: the per-tile cst, the eight
// non-window bits (fixed by the thread index) and this thread's row of the factored in-tile table; what is left per
// element is four sign-folded adds and one constant-bank operand.
template <int W>
struct TileDiag {
  double fixed;
  double w[4];
  __device__ __forceinline__ void load(const double* __restrict__ row, const double* __restrict__ intile, int tt) {
    fixed = (__ldg(row + (tt & 15)) + __ldg(row + 16 + (tt >> 4))) + __ldg(intile + tt);
#pragma unroll
    for (int k = 0; k < 4; ++k) w[k] = __ldg(row + 32 + k) + __ldg(intile + (1 + k) * 256 + tt);
  }
  // the same sums from shared-memory copies of the row and the table (the TMA-fed fused pass)
  __device__ __forceinline__ void load_shared(const double* row, const double* intile, int tt) {
    fixed = (row[tt & 15] + row[16 + (tt >> 4)]) + intile[tt];
#pragma unroll
    for (int k = 0; k < 4; ++k) w[k] = row[32 + k] + intile[(1 + k) * 256 + tt];
  }
  // j must be a compile-time constant after unrolling: the sign pattern folds into add/sub and ww[j] is read
  // straight from the kernel-parameter bank
  __device__ __forceinline__ double at(const double (&ww)[16], int j) const {
    double a = fixed + ww[j];
#pragma unroll
    for (int k = 0; k < 4; ++k) a += (j >> k & 1) ? -w[k] : w[k];
    return a;
  }
};

__device__ __forceinline__ int sea_weight(uint64_t idx, uint64_t sea_mask, int n_sea) {
  return n_sea - 2 * __popcll(idx & sea_mask);
}

// One thread per tile: lin_b(v) = scale/4 * sum_{c outer} J[gb(b)][c] s_c(v),
// cst(v) = scale * (sum_{c outer} h_c s_c/2 + sum_{c<c' outer} J[c][c'] s_c s_c'/4), folded for window W.
__global__ void build_tile_diag_kernel(const double* __restrict__ h, const double* __restrict__ J, int n, Layout L, int W,
                                       double scale, double* __restrict__ out, uint64_t n_tiles, uint64_t v0,
                                       uint64_t extra_bits) {
  uint64_t v = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= n_tiles) return;
  const uint64_t base = tile_base(L, v + v0) | extra_bits;
  const uint64_t outer = ~L.tile_mask & ((1ull << n) - 1);
  double* row = out + v * TD_STRIDE;
  double lin[TB];
  for (int b = 0; b < TB; ++b) {
    int gb = b < L.lk ? b : L.gs + (b - L.lk);
    double acc = 0.0;
    for (int c = 0; c < n; ++c)
      if (outer >> c & 1) acc = fma((base >> c & 1) ? -0.25 : 0.25, J[(size_t)gb * n + c], acc);
    lin[b] = acc * scale;
  }
  double cst = 0.0;
  for (int c = 0; c < n; ++c) {
    if (!(outer >> c & 1)) continue;
    double acc = 0.5 * h[c];
    for (int c2 = c + 1; c2 < n; ++c2)
      if (outer >> c2 & 1) acc = fma((base >> c2 & 1) ? -0.25 : 0.25, J[(size_t)c * n + c2], acc);
    cst = fma((base >> c & 1) ? -1.0 : 1.0, acc, cst);
  }
  cst *= scale;
  // thread bits in ascending order: tt bit q feeds tile bit (q < W ? q : q + 4)
  for (int half = 0; half < 2; ++half)
    for (int m = 0; m < 16; ++m) {
      double acc = half ? 0.0 : cst;
      for (int q = 0; q < 4; ++q) {
        const int tq = 4 * half + q, b = tq < W ? tq : tq + 4;
        acc += (m >> q & 1) ? -lin[b] : lin[b];
      }
      row[16 * half + m] = acc;
    }
  for (int k = 0; k < 4; ++k) row[32 + k] = lin[W + k];
  row[36] = row[37] = row[38] = row[39] = 0.0;
}

}  // namespace

size_t tile_diag_doubles(int n_loc) { return (size_t)TD_STRIDE << (n_loc - TB); }

void launch_build_tile_diag(const double* d_h, const double* d_J, int n, const Layout& L, int W, double scale,
                            double* d_out, uint64_t n_tiles, uint64_t v0, uint64_t extra_bits, cudaStream_t stream) {
  const unsigned blocks = (unsigned)((n_tiles + 127) / 128);
  build_tile_diag_kernel<<<blocks, 128, 0, stream>>>(d_h, d_J, n, L, W, scale, d_out, n_tiles, v0, extra_bits);
}

}  // namespace qsb

#include "qsb_large_kernels.inl"
