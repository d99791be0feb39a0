// pass_last + small utility kernels + launchers; included from qsb_large_kernels.inl.
namespace qsb {
namespace {

// ---------------------------------------------------------------------------------------------------
// pass_last: layout 0.  k = -i (Dz v + W0 [tX + c(v) (M/2)^(x12) tY]), then the fused stage algebra
// (Epilogue): o1 = alpha k, o2 = sum c_j vin_j + ck k, err += |sum e_j vin_j + ek k|^2 / sc^2.
// The last butterfly round leaves thread t with amplitudes t + 256 j, so the epilogue streams the other
// vectors straight from registers with fully coalesced accesses.
// ---------------------------------------------------------------------------------------------------
// EPI bit 0: o2 is produced; bit 1: the error sums are produced.  The epilogue works on batches of B amplitudes:
// all loads of a batch are issued before anything of that batch is stored, because o1/o2 may alias inputs at
// the SAME index (in-place stage vectors) and the compiler must therefore keep loads behind earlier stores.
// B x NQ: elements per batch x input vectors the batch can take (n_in <= NQ): <4, 2> for short epilogues (Taylor
// terms, first DP5 stage), <2, 8> for the long DP5 stage combinations; EPI == 0 has no inputs (<8, 0>).
template <int EPI, int B, int NQ>
__global__ void __launch_bounds__(TPT, 2)
pass_last(PassArgs a, const cplx* src, const cplx* __restrict__ tx, const cplx* __restrict__ ty,
          Epilogue ep, double* __restrict__ partials) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* sm = reinterpret_cast<cplx*>(smem_raw);
  __shared__ double red[2 * 32 + 2];
  const int tt = threadIdx.x;
  const uint64_t base = (uint64_t)blockIdx.x << TB;
  cplx r[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) r[j] = ty[base | (uint64_t)elem<8>(tt, j)];
  {  // pull the tiles needed later (tX in three rounds, the epilogue inputs after six) towards L2
    const size_t off = (size_t)tt * 256;
    const char* p = reinterpret_cast<const char*>(tx + base) + off;
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p + 128));
    p = reinterpret_cast<const char*>(src + base) + off;
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p + 128));
    if (EPI) {
      for (int q = 0; q < ep.n_in; ++q) {
        p = reinterpret_cast<const char*>(ep.vin[q] + base) + off;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p + 128));
      }
    }
  }
  bfly16<1>(r, 0xF);
  smem_store16<8>(sm, tt, r);
  __syncthreads();
  smem_round<0, 1>(sm, tt, 0xF);
  __syncthreads();
  smem_load16<4>(sm, tt, r);
  bfly16<1>(r, 0xF);
  const int pv = __popcll(base | a.g.rank_bits);  // S_rest = (+i)^popcount(outer bits)
#pragma unroll
  for (int j = 0; j < 16; ++j) r[j] = mul_pi_pow(INV_TILE * r[j], pv) + tx[base | (uint64_t)elem<4>(tt, j)];
  bfly16<0>(r, 0xF);
  smem_store16<4>(sm, tt, r);
  __syncthreads();
  smem_round<0, 0>(sm, tt, 0xF);
  __syncthreads();
  smem_load16<8>(sm, tt, r);
  bfly16<0>(r, 0xF);

  // ---- epilogue ----------------------------------------------------------------------------------------------
  // The transformed tile is parked in this thread's own (now free) shared-memory slots, so the registers can hold a
  // whole batch of outstanding loads: B elements x (src + up to NQ input vectors), all issued before the first use.
  // The groups of two input vectors sit behind uniform branches on n_in, which resolve at once and therefore do not
  // break up the load burst.
  {
    cplx* park = sm + phys_base<8>(tt);
#pragma unroll
    for (int j = 0; j < 16; ++j) park[j * PhysStride<8>::value] = r[j];
  }
  const cplx* park = sm + phys_base<8>(tt);
  TileDiag<8> td;
  td.load(a.tab.Tz + (size_t)blockIdx.x * TD_STRIDE, a.tab.Iz, tt);
  const uint64_t idx0 = base | (uint64_t)tt;  // element j of this thread is idx0 + 256 j
  double err_acc = 0.0, aux_acc = 0.0;
  const int n_in = EPI ? ep.n_in : 0;
#pragma unroll
  for (int j0 = 0; j0 < 16; j0 += B) {
    cplx v[B], w[NQ > 0 ? NQ : 1][B];
#pragma unroll
    for (int jj = 0; jj < B; ++jj) v[jj] = src[idx0 + 256u * (j0 + jj)];
#pragma unroll
    for (int q0 = 0; q0 < NQ; q0 += 2) {
      if (q0 < n_in) {
#pragma unroll
        for (int jj = 0; jj < B; ++jj) w[q0][jj] = ep.vin[q0][idx0 + 256u * (j0 + jj)];
      }
      if (q0 + 1 < n_in) {
#pragma unroll
        for (int jj = 0; jj < B; ++jj) w[q0 + 1][jj] = ep.vin[q0 + 1][idx0 + 256u * (j0 + jj)];
      }
    }
    cplx k[B];
#pragma unroll
    for (int jj = 0; jj < B; ++jj) {  // k = -i (Dz v + r)
      cplx acc = td.at(a.tab.Wz, j0 + jj) * v[jj] + park[(j0 + jj) * PhysStride<8>::value];
      k[jj] = ep.raw_h ? acc : mul_mi(acc);
    }
    if (EPI == 0) {
      if (ep.o1) {
#pragma unroll
        for (int jj = 0; jj < B; ++jj) ep.o1[idx0 + 256u * (j0 + jj)] = ep.alpha * k[jj];
      }
    } else {
      cplx o2[(EPI & 1) ? B : 1], er[(EPI & 2) ? B : 1];
      double m0[(EPI & 2) ? B : 1], m1[(EPI & 2) ? B : 1];
#pragma unroll
      for (int jj = 0; jj < B; ++jj) {
        if (EPI & 1) o2[jj] = ep.ck * k[jj];
        if (EPI & 2) { er[jj] = ep.ek * k[jj]; m0[jj] = 0.0; m1[jj] = 0.0; }
      }
#pragma unroll
      for (int q = 0; q < NQ; ++q) {
        if (q < n_in) {
          const double cq = ep.c[q], eq = ep.e[q];
#pragma unroll
          for (int jj = 0; jj < B; ++jj) {
            if (EPI & 1) axpy(o2[jj], cq, w[q][jj]);
            if (EPI & 2) {
              axpy(er[jj], eq, w[q][jj]);
              if (q == ep.sc_i0) m0[jj] = norm2(w[q][jj]);
              if (q == ep.sc_i1) m1[jj] = norm2(w[q][jj]);
            }
          }
        }
      }
      if (ep.o1) {
#pragma unroll
        for (int jj = 0; jj < B; ++jj) ep.o1[idx0 + 256u * (j0 + jj)] = ep.alpha * k[jj];
      }
#pragma unroll
      for (int jj = 0; jj < B; ++jj) {
        if (EPI & 1) ep.o2[idx0 + 256u * (j0 + jj)] = o2[jj];
        if (EPI & 2) {
          if (ep.sc_i1 < 0) m1[jj] = (EPI & 1) ? norm2(o2[jj]) : 0.0;
          double sc = ep.atol + ep.rtol * sqrt(fmax(m0[jj], m1[jj]));
          err_acc += norm2(er[jj]) / (sc * sc);
          aux_acc += m0[jj] / (sc * sc);
        }
      }
    }
  }
  if (EPI & 2) {
    double s[2] = {err_acc, aux_acc};
    block_reduce<2>(s, red);
    if (threadIdx.x == 0) {
      partials[(size_t)blockIdx.x * PSTRIDE + 7] = s[0];
      partials[(size_t)blockIdx.x * PSTRIDE + 8] = s[1];
    }
  }
}

__global__ void reduce_partials_kernel(const double* __restrict__ partials, int64_t n_tiles, double* __restrict__ out8) {
  __shared__ double red[PSTRIDE * 32 + PSTRIDE];
  double s[PSTRIDE];
#pragma unroll
  for (int k = 0; k < PSTRIDE; ++k) s[k] = 0.0;
  for (int64_t t = threadIdx.x; t < n_tiles; t += blockDim.x) {
#pragma unroll
    for (int k = 0; k < PSTRIDE; ++k) s[k] += partials[t * PSTRIDE + k];
  }
  block_reduce<PSTRIDE>(s, red);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < PSTRIDE; ++k) out8[k] = s[k];
  }
}

__global__ void set_basis_kernel(cplx* psi, uint64_t dim, uint64_t index) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (; i < dim; i += stride) psi[i] = mk(i == index ? 1.0 : 0.0, 0.0);  // index >= dim: all zero
}

}  // namespace

size_t happly_smem_bytes() { return (size_t)TILE_PHYS * sizeof(cplx); }

Profile& profile() {
  static Profile p;
  return p;
}

namespace {
// Times one launch with CUDA events on its own stream when profiling is on (serialises; bench-only).
struct Timed {
  int cls;
  cudaStream_t st;
  double bytes;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  Timed(int c, cudaStream_t s, double b = 0.0) : cls(c), st(s), bytes(b) {
    if (profile().on) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, st); }
  }
  ~Timed() {
    if (e0) {
      cudaEventRecord(e1, st);
      cudaEventSynchronize(e1);
      float ms = 0.f;
      cudaEventElapsedTime(&ms, e0, e1);
      profile().ms[cls] += ms;
      profile().count[cls] += 1;
      profile().bytes[cls] += bytes;
      cudaEventDestroy(e0); cudaEventDestroy(e1);
    }
  }
};
}  // namespace


template <int W>
static void set_centre_attrs(int sm) {
  cudaFuncSetAttribute(pass_centre<false, false, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
  cudaFuncSetAttribute(pass_centre<false, true, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
  cudaFuncSetAttribute(pass_centre<true, false, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
  cudaFuncSetAttribute(pass_centre<true, true, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
}

template <bool SHARDED, int W>
static void launch_centre(const PassArgs& a, const Layout& L, const TileMap& tm, cplx* tx, cplx* ty, const PeerPtrs& pp,
                          int collect_obs, double* partials, unsigned n_tiles, size_t sm, cudaStream_t stream) {
  if (collect_obs) pass_centre<SHARDED, true, W><<<dim3(n_tiles, 2), TPT, sm, stream>>>(a, L, tm, tx, ty, pp, partials);
  else pass_centre<SHARDED, false, W><<<dim3(n_tiles, 2), TPT, sm, stream>>>(a, L, tm, tx, ty, pp, partials);
}
template <bool SHARDED>
static void launch_centre(const PassArgs& a, const Layout& L, const TileMap& tm, cplx* tx, cplx* ty, const PeerPtrs& pp,
                          int collect_obs, double* partials, unsigned n_tiles, size_t sm, cudaStream_t stream) {
  switch (centre_window(L)) {
    case 8: launch_centre<SHARDED, 8>(a, L, tm, tx, ty, pp, collect_obs, partials, n_tiles, sm, stream); break;
    case 4: launch_centre<SHARDED, 4>(a, L, tm, tx, ty, pp, collect_obs, partials, n_tiles, sm, stream); break;
    default: launch_centre<SHARDED, 0>(a, L, tm, tx, ty, pp, collect_obs, partials, n_tiles, sm, stream); break;
  }
}

void prepare_device() {
  const int sm = (int)((size_t)TILE_PHYS * sizeof(cplx));
  cudaFuncSetAttribute(pass_first<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
  cudaFuncSetAttribute(pass_first<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
  cudaFuncSetAttribute(pass_last<0, 8, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
  cudaFuncSetAttribute(pass_last<1, 4, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
  cudaFuncSetAttribute(pass_last<2, 4, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
  cudaFuncSetAttribute(pass_last<3, 4, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
  cudaFuncSetAttribute(pass_mid, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
  set_centre_attrs<8>(sm);
  set_centre_attrs<4>(sm);
  set_centre_attrs<0>(sm);
}

// The strided-layout passes of one H-apply (mids forward, centre, mids back).
// (Issuing mid -> centre -> mid chunk batch by chunk batch through TileMap so that the 2nd and 3rd pass of a
// batch would hit L2 was measured and is slower than three full-grid launches: profiles/r1g_l2_chunk_experiment.log.)
// Sharded: every rank runs the centre pass on its 1/world slice of the centre tiles, reading and writing the
// peers' tx/ty in place; the rank barrier before it makes the peers' forward passes visible, the one after it
// keeps the peers' backward passes from starting on rows that are still being written.
static int launch_strided_passes(const LargeGeom& g, const PassArgs& a, cplx* tx, cplx* ty, int collect_obs,
                                 double* partials, cudaStream_t stream, const ShardLink* link) {
  const size_t sm1 = (size_t)TILE_PHYS * sizeof(cplx);
  const unsigned n_tiles = 1u << (g.n_loc - TB);
  const int centre = g.n_layouts - 1;
  const TileMap all = {0u, 0, 0, 0u};
  const double S = (double)(sizeof(cplx) << g.n_loc);  // bytes of one local vector
  for (int k = 1; k < centre; ++k) { Timed t(1, stream, 4 * S); pass_mid<<<dim3(n_tiles, 2), TPT, sm1, stream>>>(g.lay[k], all, tx, ty); }
  if (g.world == 1) {
    Timed t(2, stream, 4 * S);
    PeerPtrs none = {};
    launch_centre<false>(a, g.lay[centre], all, tx, ty, none, collect_obs, partials, n_tiles, sm1, stream);
  } else {
    if (!link) { set_error("sharded plan is not connected to its peers (qsb_shard_connect)"); return QSB_ERR_BAD_ARG; }
    int rc = link->sync(link->ctx);
    if (rc) return rc;
    {
      Timed t(2, stream, 4 * S);
      PeerPtrs pp = {};
      for (int r = 0; r < g.world; ++r) { pp.t[0][r] = link->tx[r]; pp.t[1][r] = link->ty[r]; }
      const TileMap mine = {(unsigned)g.rank * n_tiles, 0, 0, 0u};
      launch_centre<true>(a, g.lay[centre], mine, tx, ty, pp, collect_obs, partials, n_tiles, sm1, stream);
    }
    rc = link->sync(link->ctx);
    if (rc) return rc;
  }
  for (int k = centre - 1; k >= 1; --k) { Timed t(1, stream, 4 * S); pass_mid<<<dim3(n_tiles, 2), TPT, sm1, stream>>>(g.lay[k], all, tx, ty); }
  return QSB_OK;
}

int launch_happly(const LargeGeom& g, const DiagTables& tab, const cplx* src, cplx* tx, cplx* ty,
                  const Epilogue& ep, int collect_obs, double* partials, cudaStream_t stream, const ShardLink* link,
                  const Combine* comb) {
  const size_t sm2 = (size_t)TILE_PHYS * sizeof(cplx);
  PassArgs a;
  a.g = g;
  a.tab = tab;
  const unsigned n_tiles = 1u << (g.n_loc - TB);
  if (ep.n_in > 2) { set_error("internal: fused epilogues take at most 2 input vectors"); return QSB_ERR_BAD_ARG; }
  const double S = (double)(sizeof(cplx) << g.n_loc);
  {
    const bool cmb = comb && comb->n_in > 0;
    Timed t(cmb ? 4 : 0, stream, cmb ? (comb->n_in + 3) * S : 3 * S);
    if (comb && comb->n_in > 0) {
      if (comb->out != src) { set_error("internal: Combine::out must be the vector the apply acts on"); return QSB_ERR_BAD_ARG; }
      pass_first<true><<<n_tiles, TPT, sm2, stream>>>(a, src, tx, ty, *comb, collect_obs, partials);
    } else {
      Combine none = {};
      pass_first<false><<<n_tiles, TPT, sm2, stream>>>(a, src, tx, ty, none, collect_obs, partials);
    }
  }
  int rc = launch_strided_passes(g, a, tx, ty, collect_obs, partials, stream, link);
  if (rc) return rc;
  {
    const int epi = (ep.o2 ? 1 : 0) | (ep.want_err ? 2 : 0);
    Timed t(3, stream, (3 + ep.n_in + (ep.o1 ? 1 : 0) + (ep.o2 ? 1 : 0)) * S);
    switch (epi) {
      case 0: pass_last<0, 8, 0><<<n_tiles, TPT, sm2, stream>>>(a, src, tx, ty, ep, partials); break;
      case 1: pass_last<1, 4, 2><<<n_tiles, TPT, sm2, stream>>>(a, src, tx, ty, ep, partials); break;
      case 2: pass_last<2, 4, 2><<<n_tiles, TPT, sm2, stream>>>(a, src, tx, ty, ep, partials); break;
      default: pass_last<3, 4, 2><<<n_tiles, TPT, sm2, stream>>>(a, src, tx, ty, ep, partials); break;
    }
  }
  return QSB_OK;
}

int launch_observe(const LargeGeom& g, const DiagTables& tab, const cplx* src, cplx* tx, cplx* ty,
                   double* partials, cudaStream_t stream, const ShardLink* link) {
  // the transform half of an H-apply: z sums in pass_first, x/y sums in pass_centre (tx/ty are clobbered)
  const size_t sm2 = (size_t)TILE_PHYS * sizeof(cplx);
  PassArgs a;
  a.g = g;
  a.tab = tab;
  const unsigned n_tiles = 1u << (g.n_loc - TB);
  Combine none = {};
  pass_first<false><<<n_tiles, TPT, sm2, stream>>>(a, src, tx, ty, none, 1, partials);
  return launch_strided_passes(g, a, tx, ty, 1, partials, stream, link);
}

void launch_error_sums(const LargeGeom& g, const ErrSpec& es, double* partials, cudaStream_t stream) {
  Timed t(5, stream, es.n_in * (double)(sizeof(cplx) << g.n_loc));
  error_sums_kernel<<<1u << (g.n_loc - TB), TPT, 0, stream>>>(es, partials);
}

void launch_reduce_partials(const double* partials, int64_t n_tiles, double* out8, cudaStream_t stream) {
  reduce_partials_kernel<<<1, 1024, 0, stream>>>(partials, n_tiles, out8);
}

void launch_set_basis_state(cplx* psi, uint64_t dim, uint64_t index, cudaStream_t stream) {
  int blocks = (int)std::min<uint64_t>((dim + 255) / 256, 148 * 16);
  set_basis_kernel<<<blocks, 256, 0, stream>>>(psi, dim, index);
}

}  // namespace qsb
