// Kernels of the tiled path; included at the end of qsb_large.cu.
//
// Every pass is one 256-thread CTA per 4096-amplitude tile with a single padded shared-memory buffer
// (2 CTAs per SM).  The y-branch never needs its own copy of the layout-0 transform:
//     W0 S0^dagger = (M'/2)^(x12) W0      and      S0 W0 = W0 (M/2)^(x12),
// so pass_first derives tY from tX in place, and pass_last folds tY into tX before ONE inverse transform.
namespace qsb {
namespace {

struct PassArgs {
  LargeGeom g;
  DiagTables tab;
};

// (M/2)^(x12) = M_SCALE K^(x12): the 2^-12 of the twelve halvings times (1 +- i)^12 = -64 (see bfly16)
constexpr double M_SCALE = -64.0 / (double)TILE;

// Which tiles a launch of a strided-layout pass covers.  Tile v of a layout is numbered by its outer bits in
// ascending order, so index bit b in [lk, 12) - a pad bit that NO strided layout transforms once b >= the largest lk -
// is bit (b - lk) of v.  Those bits select a "chunk": all strided passes of one chunk touch the same amplitudes and
// nothing of any other chunk, which is what lets the sharded mode pipeline chunk c's NVLink-bound centre pass under
// the HBM-bound passes of chunks c - 1 and c + 1.  A launch covers chunks c0 .. c0 + 2^lb - 1 of a cw-bit chunk field
// that starts at bit `shift` of v:
//     bx = [low: shift bits][chunk - c0: lb bits][high]   ->   v = vbase + low + (chunk << shift) + (high << (shift + cw)).
// {vbase, 0, 0, 0, 0} is v = vbase + bx.  vbase: first tile of this rank's slice of the centre layout (sharded), else 0.
struct TileMap {
  unsigned vbase;
  int shift, cw, lb;
  unsigned c0;
  unsigned zero;  // always 0; see opaque()
  int branch0;    // branch of blockIdx.y == 0 (a launch with gridDim.y == 1 covers ONE branch: sharded overlap)
  unsigned count; // pass_centre<PERSIST> only: the gridDim.x CTAs walk bx = 0 .. count-1
  __device__ __forceinline__ uint64_t tile(unsigned bx) const {
    const unsigned low = bx & ((1u << shift) - 1), cl = (bx >> shift) & ((1u << lb) - 1), hi = bx >> (shift + lb);
    return (uint64_t)vbase + low + ((uint64_t)(c0 + cl) << shift) + ((uint64_t)hi << (shift + cw));
  }
};

// ---------------------------------------------------------------------------------------------------
// pass_first: layout 0.  tX = W0 psi;  tY = W0 S^dagger psi = c^dagger(v) (M'/2)^(x12) tX.
// Optionally the z-basis observable partial sums.
// ---------------------------------------------------------------------------------------------------
// COMB: psi is formed here as sum_q c_q vin_q (Combine) and written to cb.out on the way; the loads go in two
// half-bursts of 8 per input vector so that accumulators + in-flight values fit the register file.
// ERR (with COMB): the loads also accumulate cb.eout = sum_q e_q vin_q, in quarter-bursts of 4 so that the second
// accumulator fits the register file.
template <bool COMB, bool ERR = false>
__global__ void __launch_bounds__(TPT, 2)
pass_first(PassArgs a, const cplx* __restrict__ psi, cplx* __restrict__ tx, cplx* __restrict__ ty, Combine cb,
           int collect_obs, double* __restrict__ partials) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* sm = reinterpret_cast<cplx*>(smem_raw);
  __shared__ double red[3 * 32 + 3];
  const int tt = threadIdx.x;
  const uint64_t base = (uint64_t)blockIdx.x << TB;
  cplx r[16];
  if (!COMB) {
#pragma unroll
    for (int j = 0; j < 16; ++j) r[j] = psi[base | (uint64_t)elem<8>(tt, j)];
  } else if (ERR) {
    const uint64_t i0 = base | (uint64_t)tt;  // elem<8>(tt, j) = tt + 256 j
#pragma unroll
    for (int q4 = 0; q4 < 16; q4 += 4) {
      cplx er[4];
      {
        const cplx* v0 = cb.vin[0];
        const double c0 = cb.c[0], e0 = cb.e[0];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const cplx w = v0[i0 + 256u * (q4 + j)];
          r[q4 + j] = c0 * w;
          er[j] = e0 * w;
        }
      }
      for (int q = 1; q < cb.n_in; ++q) {
        const cplx* vq = cb.vin[q];
        const double cq = cb.c[q], eq = cb.e[q];
        cplx w[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) w[j] = vq[i0 + 256u * (q4 + j)];
#pragma unroll
        for (int j = 0; j < 4; ++j) { axpy(r[q4 + j], cq, w[j]); axpy(er[j], eq, w[j]); }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        cb.out[i0 + 256u * (q4 + j)] = r[q4 + j];
        cb.eout[i0 + 256u * (q4 + j)] = er[j];
      }
    }
  } else {
    const uint64_t i0 = base | (uint64_t)tt;  // elem<8>(tt, j) = tt + 256 j
#pragma unroll
    for (int half = 0; half < 16; half += 8) {
      {
        const cplx* v0 = cb.vin[0];
        const double c0 = cb.c[0];
#pragma unroll
        for (int j = 0; j < 8; ++j) r[half + j] = c0 * v0[i0 + 256u * (half + j)];
      }
      for (int q = 1; q < cb.n_in; ++q) {
        const cplx* vq = cb.vin[q];
        const double cq = cb.c[q];
        cplx w[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) w[j] = vq[i0 + 256u * (half + j)];
#pragma unroll
        for (int j = 0; j < 8; ++j) axpy(r[half + j], cq, w[j]);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) cb.out[i0 + 256u * (half + j)] = r[half + j];
    }
  }
  double s[3] = {0.0, 0.0, 0.0};
  if (collect_obs) {
    const int n_sea = __popcll(a.g.sea_mask);
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      uint64_t idx = base | a.g.rank_bits | (uint64_t)elem<8>(tt, j);
      double p = norm2(r[j]);
      s[0] += p;
      s[1] = fma(p, (double)sea_weight(idx, a.g.sea_mask, n_sea), s[1]);
      s[2] += (idx & a.g.rare_mask) ? -p : p;
    }
  }
  bfly16<0>(r, 0xF);
  smem_store16<8>(sm, tt, r);
  __syncthreads();
  smem_round<0, 0>(sm, tt, 0xF);
  __syncthreads();
  smem_load16<4>(sm, tt, r);
  bfly16<0>(r, 0xF);
#pragma unroll
  for (int j = 0; j < 16; ++j) tx[base | (uint64_t)elem<4>(tt, j)] = r[j];
  // y branch, continuing from the registers (each thread owns its window-4 elements: no barrier needed here)
  bfly16<2>(r, 0xF);
  smem_store16<4>(sm, tt, r);
  __syncthreads();
  smem_round<0, 2>(sm, tt, 0xF);
  __syncthreads();
  smem_load16<8>(sm, tt, r);
  bfly16<2>(r, 0xF);
  const int pv = __popcll(base | a.g.rank_bits);  // S_rest^dagger = (-i)^popcount(outer bits)
#pragma unroll
  for (int j = 0; j < 16; ++j) ty[base | (uint64_t)elem<8>(tt, j)] = mul_mi_pow(M_SCALE * r[j], pv);
  if (collect_obs) {
    block_reduce<3>(s, red);
    if (threadIdx.x == 0) {
      double* p = partials + (size_t)blockIdx.x * PSTRIDE;
      p[0] = s[0]; p[1] = s[1]; p[2] = s[2];
    }
  }
}

// Peer bases of both branches for the sharded centre pass: element idx lives on rank idx >> n_loc.
struct PeerPtrs {
  cplx* t[2][MAX_WORLD];
};

// Push-only exchange: where the last forward mid pass delivers its output.  The amplitude at local index i of this
// rank belongs to centre tile v = index bits [lk_c, gs_c) (owner = the top bits of v, v >> tl), tile-local position
// u = pad bits [0, lk_c) | centre-group bits (local ones [gs_c, n_loc), then the rank) << lk_c.
struct PushSpec {
  cplx* r[2][MAX_WORLD];  // receive buffers [branch][owner]
  int lk_c, gs_c, tl;     // centre layout's pad bits / group start; log2 of the centre tiles per rank
  uint64_t rank_hi;       // rank << (n_loc - gs_c): this rank's bits in the centre group
  int on;
  __device__ __forceinline__ cplx* dest(int branch, uint64_t i) const {
    const uint64_t v = (i & ((1ull << gs_c) - 1)) >> lk_c;
    const uint64_t u = (i & ((1ull << lk_c) - 1)) | (((i >> gs_c) | rank_hi) << lk_c);
    return r[branch][v >> tl] + (((v & ((1ull << tl) - 1)) << TB) | u);
  }
};

// ---------------------------------------------------------------------------------------------------
// pass_mid: in-place Hadamard on one group's bits (layout k, 0 < k < centre).  blockIdx.y = branch.
// ---------------------------------------------------------------------------------------------------
// PAD: the tile buffer has the bank-conflict pad (needed only if the layout transforms window [0,4), which no layout
// built by make_geometry does: groups have at most 8 bits)
// PUSH: the stores go to the owners' receive buffers (PushSpec) instead of back in place.
template <bool PAD, bool PUSH = false>
__global__ void __launch_bounds__(TPT, 2)
pass_mid(Layout L, TileMap tm, cplx* __restrict__ tx, cplx* __restrict__ ty, PushSpec ps) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* sm = reinterpret_cast<cplx*>(smem_raw);
  const int branch = tm.branch0 + blockIdx.y;
  cplx* t = branch ? ty : tx;
  const int tt = threadIdx.x;
  const uint64_t base = tile_base(L, tm.tile(blockIdx.x));
  cplx r[16];
  {
    WinOff wo;
    wo.init<8>(L, 0);
    const cplx* p = t + gidx(L, base, elem<8>(tt, 0));
#pragma unroll
    for (int j = 0; j < 16; ++j) r[j] = p[wo.at(j)];
  }
  bfly16<0>(r, L.m8);
  any_store16<8, PAD>(sm, tt, r);
  __syncthreads();
  if constexpr (PAD) {
    if (L.m0) {
      smem_round<0, 0>(sm, tt, L.m0);
      __syncthreads();
    }
  }
  any_load16<4, PAD>(sm, tt, r);
  bfly16<0>(r, L.m4);
  {
    const unsigned late0 = tm.zero & (unsigned)__double2loint(r[0].x);  // 0, but only known once the data is final
    WinOff wo;
    wo.init<4>(L, late0);
    if constexpr (PUSH) {
      const uint64_t i0 = gidx(L, base | late0, elem<4>(tt, 0));
#pragma unroll
      for (int j = 0; j < 16; ++j) *ps.dest(branch, i0 + wo.at(j)) = r[j];
    } else {
      cplx* p = t + gidx(L, base | late0, elem<4>(tt, 0));
#pragma unroll
      for (int j = 0; j < 16; ++j) p[wo.at(j)] = r[j];
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// pass_centre: t <- W_g D W_g t on the centre layout; D is the x- or y-basis diagonal form evaluated on
// the full index.  Optionally that basis' observable partial sums (taken on the fully transformed vector).
// ---------------------------------------------------------------------------------------------------
template <int W, bool COLLECT>
__device__ __forceinline__ void centre_core(cplx (&r)[16], int tt, int mask, const TileDiag<W>& td, const double (&ww)[16],
                                            const Layout& L, uint64_t base, const LargeGeom& g, double (&s)[2]) {
  bfly16<0>(r, mask);
  if (COLLECT) {
    const int n_sea = __popcll(g.sea_mask);
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      uint64_t idx = gidx(L, base, elem<W>(tt, j));
      double p = norm2(r[j]);
      s[0] = fma(p, (double)sea_weight(idx, g.sea_mask, n_sea), s[0]);
      s[1] += (idx & g.rare_mask) ? -p : p;
    }
  }
#pragma unroll
  for (int j = 0; j < 16; ++j) r[j] = td.at(ww, j) * r[j];
  bfly16<0>(r, mask);
}

// W = centre_window(L): the register window the diagonal is evaluated in (8: the group fits window [8,12) and never
// touches shared memory; 4: N = 28; 0: groups wider than 8 bits do not exist, but layout 0 as centre (n = 12) and
// short groups with lk > 8 land here).  The diagonal coefficients (10 registers) are fetched FIRST so that their
// latency hides behind the tile loads and nothing but the data is live across the butterflies.
// SHARDED: the group's top bits are the rank, so tile rows come from (and go back to) the peers' scratch vectors
// through P2P loads/stores; this rank owns tiles tm.vbase .. of the centre layout and the diagonal rows / partial
// sums are numbered locally (blockIdx.x).
// (Three CTAs per SM - 80 registers, which the W = 4 body fits without spilling, and the large carve-out - were
// measured at N = 28: 3.99 ms against 3.28 ms, profiles/r2_centre_occ3.json; L1 capacity, not a third tile, is what
// the loads in flight need.)
// PUSHED (sharded): the tile was delivered by the peers' last forward mid pass into this rank's receive buffer
// (rbuf[branch], [slot][4096] in tile-local order): the loads are local and contiguous, only the stores are remote.
template <bool SHARDED, bool COLLECT, int W, bool PERSIST = false, bool PUSHED = false>
__global__ void __launch_bounds__(TPT, 2)
pass_centre(PassArgs a, Layout L, TileMap tm, cplx* __restrict__ tx, cplx* __restrict__ ty, PeerPtrs peers,
            double* __restrict__ partials, const cplx* __restrict__ rbuf_x = nullptr, const cplx* __restrict__ rbuf_y = nullptr) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* sm = reinterpret_cast<cplx*>(smem_raw);
  __shared__ double red[2 * 32 + 2];
  const int branch = tm.branch0 + blockIdx.y;  // 0: x, 1: y
  cplx* t = branch ? ty : tx;
  const double (&ww)[16] = branch ? a.tab.Wy : a.tab.Wx;
  const int tt = threadIdx.x;
  const uint64_t lmask = (1ull << a.g.n_loc) - 1;
  // PERSIST: gridDim.x CTAs walk the tm.count tiles of this rank.  The overlapped sharded mode launches one CTA per
  // SM, so that the HBM-bound passes of the other branch keep the SMs' second CTA slot while this NVLink-bound pass
  // is in flight (two full-occupancy kernels would only take turns).  (A compile-time switch: the loop costs the
  // one-tile-per-CTA kernels of the unsharded hot path registers they do not have.)
  unsigned bx = blockIdx.x;
  do {
    const uint64_t v = tm.tile(bx);
    const size_t slot = (size_t)(v - tm.vbase);  // this rank's number for the tile (diagonal rows, partial sums)
    TileDiag<W> td;
    td.load((branch ? a.tab.Ty : a.tab.Tx) + slot * TD_STRIDE, branch ? a.tab.Iy : a.tab.Ix, tt);
    const uint64_t base = tile_base(L, v);
    cplx r[16];
    double s[2] = {0.0, 0.0};
    if constexpr (PUSHED) {
      const cplx* rb = (branch ? rbuf_y : rbuf_x) + (slot << TB) + elem<8>(tt, 0);
#pragma unroll
      for (int j = 0; j < 16; ++j) r[j] = rb[j << 8];
    } else {
      WinOff wo;
      wo.init<8>(L, 0);
      const uint64_t idx0 = gidx(L, base, elem<8>(tt, 0));
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const uint64_t idx = idx0 + wo.at(j);
        r[j] = SHARDED ? peers.t[branch][idx >> a.g.n_loc][idx & lmask] : t[idx];
      }
    }
    if constexpr (W == 8) {
      centre_core<8, COLLECT>(r, tt, L.m8, td, ww, L, base, a.g, s);
    } else {
      constexpr bool PAD = W == 0;  // only window [0,4) needs the padded buffer
      bfly16<0>(r, L.m8);
      any_store16<8, PAD>(sm, tt, r);
      __syncthreads();
      if constexpr (W == 4) {
        tile_load16<4>(sm, tt, r);
        centre_core<4, COLLECT>(r, tt, L.m4, td, ww, L, base, a.g, s);
        tile_store16<4>(sm, tt, r);
      } else {
        smem_round<4, 0>(sm, tt, L.m4);
        __syncthreads();
        smem_load16<0>(sm, tt, r);
        centre_core<0, COLLECT>(r, tt, L.m0, td, ww, L, base, a.g, s);
        smem_store16<0>(sm, tt, r);
        __syncthreads();
        smem_round<4, 0>(sm, tt, L.m4);
      }
      __syncthreads();
      any_load16<8, PAD>(sm, tt, r);
      bfly16<0>(r, L.m8);
    }
    {
      const unsigned late0 = tm.zero & (unsigned)__double2loint(r[0].x);  // 0, but only known once the data is final
      WinOff wo;
      wo.init<8>(L, late0);
      const uint64_t idx0 = gidx(L, base | late0, elem<8>(tt, 0));
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const uint64_t idx = idx0 + wo.at(j);
        if (SHARDED) peers.t[branch][idx >> a.g.n_loc][idx & lmask] = r[j];
        else t[idx] = r[j];
      }
    }
    if (COLLECT) {
      block_reduce<2>(s, red);
      if (threadIdx.x == 0) {
        double* p = partials + slot * PSTRIDE + 3 + 2 * branch;
        p[0] = s[0]; p[1] = s[1];
      }
    }
    if constexpr (!PERSIST) break;
    __syncthreads();  // the next tile's first store must not overtake this tile's last shared-memory reads
    bx += gridDim.x;
  } while (bx < tm.count);
}

// ---------------------------------------------------------------------------------------------------
// pass_centre_pipe: the sharded centre pass as a persistent, double-buffered kernel (one CTA per SM).  The NVLink
// round trip of a tile is several microseconds; pass_centre exposes it once per tile (load -> transform -> store with
// nothing in flight during the transform).  Here tile i + 1 is already on its way - cp.async copies straight from the
// peers' scratch into the OTHER 64 KB shared-memory buffer - while tile i is transformed in place in this one, so the
// links always have 64 KB per SM in flight.  Same arithmetic per tile as pass_centre (bitwise).  W = 8 or 4 only
// (window [0,4) would need the padded buffer).
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}

template <bool COLLECT, int W>
__global__ void __launch_bounds__(TPT, 2)
pass_centre_pipe(PassArgs a, Layout L, TileMap tm, PeerPtrs peers, double* __restrict__ partials) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* bufs = reinterpret_cast<cplx*>(smem_raw);  // [2][TILE], unpadded
  __shared__ double red[2 * 32 + 2];
  const int branch = tm.branch0 + blockIdx.y;
  const double (&ww)[16] = branch ? a.tab.Wy : a.tab.Wx;
  const int tt = threadIdx.x;
  const uint64_t lmask = (1ull << a.g.n_loc) - 1;
  auto issue = [&](unsigned bx, cplx* dst) {
    const uint64_t base = tile_base(L, tm.tile(bx));
    WinOff wo;
    wo.init<8>(L, 0);
    const uint64_t idx0 = gidx(L, base, elem<8>(tt, 0));
    cplx* d = dst + elem<8>(tt, 0);
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const uint64_t idx = idx0 + wo.at(j);
      cp_async16(d + (j << 8), &peers.t[branch][idx >> a.g.n_loc][idx & lmask]);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  unsigned bx = blockIdx.x;
  int cur = 0;
  if (bx < tm.count) issue(bx, bufs);
  while (bx < tm.count) {
    const unsigned next = bx + gridDim.x;
    cplx* sm = bufs + (size_t)cur * TILE;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();  // tile bx has landed for every thread, and nobody still reads the other buffer (tile bx - stride)
    if (next < tm.count) issue(next, bufs + (size_t)(cur ^ 1) * TILE);
    const uint64_t v = tm.tile(bx);
    const size_t slot = (size_t)(v - tm.vbase);
    TileDiag<W> td;
    td.load((branch ? a.tab.Ty : a.tab.Tx) + slot * TD_STRIDE, branch ? a.tab.Iy : a.tab.Ix, tt);
    const uint64_t base = tile_base(L, v);
    cplx r[16];
    double s[2] = {0.0, 0.0};
    tile_load16<8>(sm, tt, r);
    if constexpr (W == 8) {
      centre_core<8, COLLECT>(r, tt, L.m8, td, ww, L, base, a.g, s);
    } else {
      bfly16<0>(r, L.m8);
      tile_store16<8>(sm, tt, r);
      __syncthreads();
      tile_load16<4>(sm, tt, r);
      centre_core<4, COLLECT>(r, tt, L.m4, td, ww, L, base, a.g, s);
      tile_store16<4>(sm, tt, r);
      __syncthreads();
      tile_load16<8>(sm, tt, r);
      bfly16<0>(r, L.m8);
    }
    {
      const unsigned late0 = tm.zero & (unsigned)__double2loint(r[0].x);
      WinOff wo;
      wo.init<8>(L, late0);
      const uint64_t idx0 = gidx(L, base | late0, elem<8>(tt, 0));
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const uint64_t idx = idx0 + wo.at(j);
        peers.t[branch][idx >> a.g.n_loc][idx & lmask] = r[j];
      }
    }
    if (COLLECT) {
      block_reduce<2>(s, red);
      if (threadIdx.x == 0) {
        double* p = partials + slot * PSTRIDE + 3 + 2 * branch;
        p[0] = s[0]; p[1] = s[1];
      }
    }
    bx = next;
    cur ^= 1;
  }
}

// ---------------------------------------------------------------------------------------------------
// error_sums: the embedded-pair error norm as one streaming sweep; CTA b covers amplitudes [4096 b, 4096 (b+1)),
// so its partial sums land in the same per-tile slots (7: error, 8: |vin[sc_i0]|^2 / sc^2) the fused epilogues use.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPT, 4)
error_sums_kernel(ErrSpec es, double* __restrict__ partials) {
  __shared__ double red[2 * 32 + 2];
  const uint64_t i0 = ((uint64_t)blockIdx.x << TB) | threadIdx.x;
  double err_acc = 0.0, aux_acc = 0.0;
#pragma unroll 1
  for (int j0 = 0; j0 < 16; j0 += 4) {
    cplx er[4];
    double m0[4], m1[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) { er[j] = mk(0.0, 0.0); m0[j] = 0.0; m1[j] = 0.0; }
    for (int q = 0; q < es.n_in; ++q) {
      const cplx* vq = es.vin[q];
      const double eq = es.e[q];
      cplx w[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) w[j] = vq[i0 + 256u * (j0 + j)];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        axpy(er[j], eq, w[j]);
        if (q == es.sc_i0) m0[j] = norm2(w[j]);
        if (q == es.sc_i1) m1[j] = norm2(w[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const double sc = es.atol + es.rtol * sqrt(fmax(m0[j], m1[j]));
      err_acc += norm2(er[j]) / (sc * sc);
      aux_acc += m0[j] / (sc * sc);
    }
  }
  double s[2] = {err_acc, aux_acc};
  block_reduce<2>(s, red);
  if (threadIdx.x == 0) {
    partials[(size_t)blockIdx.x * PSTRIDE + 7] = s[0];
    partials[(size_t)blockIdx.x * PSTRIDE + 8] = s[1];
  }
}

}  // namespace
}  // namespace qsb

#include "qsb_large_fused.inl"
#include "qsb_large_last.inl"
