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
    if (ep.pf & 1) {
      asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(p + 128));
    } else {
      asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
      asm volatile("prefetch.global.L2 [%0];" ::"l"(p + 128));
    }
    p = reinterpret_cast<const char*>(src + base) + off;
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p + 128));
    if (EPI && !(ep.pf & 2)) {
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
  for (int j = 0; j < 16; ++j) r[j] = mul_pi_pow(M_SCALE * r[j], pv) + tx[base | (uint64_t)elem<4>(tt, j)];
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
  // (EPI == 0 has no input vectors: src alone fits next to the tile, which then stays in registers.)
  if (EPI != 0) {
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
      cplx acc = td.at(a.tab.Wz, j0 + jj) * v[jj] + (EPI != 0 ? park[(j0 + jj) * PhysStride<8>::value] : r[j0 + jj]);
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
        if (EPI & 1) { o2[jj] = ep.ck * k[jj]; axpy(o2[jj], ep.cs, v[jj]); }
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
          if (ep.sc_i1 == -2) m1[jj] = norm2(v[jj]);
          else if (ep.sc_i1 < 0) m1[jj] = (EPI & 1) ? norm2(o2[jj]) : 0.0;
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

// ---------------------------------------------------------------------------------------------------
// pass_last_first: pass_last of one DP5 stage fused with pass_first<COMB> of the NEXT stage.  Both work on the same
// layout-0 tile, so the fresh k never leaves the registers between them: the next stage's combination
//     v' = sum_q c_q vin_q + c_fresh k        (vin: y and the OLDER stage vectors)
// takes it from there instead of re-reading it (1 S less traffic per stage, one launch less), and v' replaces v in
// `ys` - the same addresses this thread has just read for the Dz term.  tX / tY are rewritten in place, tile by tile,
// with the mappings they were read with.  The arithmetic of the two halves is that of pass_last<0, 8, 0> and
// pass_first<true>; only the order of the combination's summands differs (the fresh k first).
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPT, 2)
pass_last_first(PassArgs a, cplx* ys, cplx* tx, cplx* ty, cplx* __restrict__ kout, Combine cb, double c_fresh) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* sm = reinterpret_cast<cplx*>(smem_raw);
  const int tt = threadIdx.x;
  const uint64_t base = (uint64_t)blockIdx.x << TB;
  cplx r[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) r[j] = ty[base | (uint64_t)elem<8>(tt, j)];
  {
    const size_t off = (size_t)tt * 256;
    const char* p = reinterpret_cast<const char*>(tx + base) + off;
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p + 128));
    p = reinterpret_cast<const char*>(ys + base) + off;
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p + 128));
  }
  bfly16<1>(r, 0xF);
  smem_store16<8>(sm, tt, r);
  __syncthreads();
  smem_round<0, 1>(sm, tt, 0xF);
  __syncthreads();
  smem_load16<4>(sm, tt, r);
  bfly16<1>(r, 0xF);
  const int pv = __popcll(base | a.g.rank_bits);
#pragma unroll
  for (int j = 0; j < 16; ++j) r[j] = mul_pi_pow(M_SCALE * r[j], pv) + tx[base | (uint64_t)elem<4>(tt, j)];
  bfly16<0>(r, 0xF);
  smem_store16<4>(sm, tt, r);
  __syncthreads();
  smem_round<0, 0>(sm, tt, 0xF);
  __syncthreads();
  smem_load16<8>(sm, tt, r);
  bfly16<0>(r, 0xF);
  const uint64_t idx0 = base | (uint64_t)tt;  // element j of this thread is idx0 + 256 j
  {
    TileDiag<8> td;
    td.load(a.tab.Tz + (size_t)blockIdx.x * TD_STRIDE, a.tab.Iz, tt);
#pragma unroll
    for (int j0 = 0; j0 < 16; j0 += 8) {
      cplx v[8];
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) v[jj] = ys[idx0 + 256u * (j0 + jj)];
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) r[j0 + jj] = mul_mi(td.at(a.tab.Wz, j0 + jj) * v[jj] + r[j0 + jj]);  // k = -i (Dz v + ..)
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) kout[idx0 + 256u * (j0 + jj)] = r[j0 + jj];
    }
  }
  // ---- the next stage's pass_first<COMB>, with k already in r ------------------------------------------------------
#pragma unroll
  for (int half = 0; half < 16; half += 8) {
#pragma unroll
    for (int j = 0; j < 8; ++j) r[half + j] = c_fresh * r[half + j];
    for (int q = 0; q < cb.n_in; ++q) {
      const cplx* vq = cb.vin[q];
      const double cq = cb.c[q];
      cplx w[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) w[j] = vq[idx0 + 256u * (half + j)];
#pragma unroll
      for (int j = 0; j < 8; ++j) axpy(r[half + j], cq, w[j]);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) ys[idx0 + 256u * (half + j)] = r[half + j];
  }
  bfly16<0>(r, 0xF);
  __syncthreads();  // every thread is past its smem_load16<8> above before the tile buffer is written again
  smem_store16<8>(sm, tt, r);
  __syncthreads();
  smem_round<0, 0>(sm, tt, 0xF);
  __syncthreads();
  smem_load16<4>(sm, tt, r);
  bfly16<0>(r, 0xF);
#pragma unroll
  for (int j = 0; j < 16; ++j) tx[base | (uint64_t)elem<4>(tt, j)] = r[j];
  bfly16<2>(r, 0xF);
  smem_store16<4>(sm, tt, r);
  __syncthreads();
  smem_round<0, 2>(sm, tt, 0xF);
  __syncthreads();
  smem_load16<8>(sm, tt, r);
  bfly16<2>(r, 0xF);
#pragma unroll
  for (int j = 0; j < 16; ++j) ty[base | (uint64_t)elem<8>(tt, j)] = mul_mi_pow(M_SCALE * r[j], pv);
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

// psi[x] = prod_b phi[b][bit_b(x | rank_bits)]: a state whose H-image has a closed form on every row (parity hook
// for sizes no host vector reaches).  Thread i owns x = i mod 2^12 .. in strides; the product over the bits above 12
// is shared by a whole 4096-amplitude block, so it is formed once per block.
__global__ void product_state_kernel(cplx* __restrict__ psi, ProductPhi phi, int n, uint64_t rank_bits) {
  const uint64_t base = (uint64_t)blockIdx.x << TB;
  cplx hi = mk(1.0, 0.0);
  for (int b = TB; b < n; ++b) hi = cmul(hi, phi.v[b][((base | rank_bits) >> b) & 1]);
  for (int i = threadIdx.x; i < TILE; i += blockDim.x) {
    cplx a = hi;
#pragma unroll
    for (int b = 0; b < TB; ++b) a = cmul(a, phi.v[b][(i >> b) & 1]);
    psi[base | (uint64_t)i] = a;
  }
}

__global__ void gather_rows_kernel(const cplx* __restrict__ src, const uint64_t* __restrict__ rows, int64_t n_rows,
                                   cplx* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_rows) out[i] = src[rows[i]];
}

// out[i] = y[i] + sum_m c[m] term[m][i]: the Taylor polynomial of one step read off at an interior time.
__global__ void __launch_bounds__(256)
dense_eval_kernel(DenseSpec d, uint64_t dim) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < dim; i += stride) {
    cplx acc = d.y[i];
    for (int m = 0; m < d.n; ++m) axpy(acc, d.c[m], d.term[m][i]);
    d.out[i] = acc;
  }
}

// FP64 issue-rate microbenchmark (the denominator of bench.py's FP64 co-bound): 8 independent DFMA chains per thread.
__global__ void __launch_bounds__(256)
fp64_peak_kernel(double* out, int iters, double a, double b) {
  double x[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) x[k] = (double)(threadIdx.x + k);
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 8; ++k) x[k] = fma(x[k], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += x[k];
  if (s == 123.456) out[0] = s;  // keeps the chains alive without a store in practice
}

__global__ void set_identity_kernel(cplx* psi, int n) {
  const uint64_t dim2 = 1ull << (2 * n), stride = (uint64_t)gridDim.x * blockDim.x, mask = (1ull << n) - 1;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < dim2; i += stride)
    psi[i] = mk((i & mask) == (i >> n) ? 1.0 : 0.0, 0.0);
}

// One CTA per column: the eight numbers qsb_evolve reports for a state (same sums as small_observe / warp_observe),
// reduced in a fixed order.
__global__ void __launch_bounds__(256)
observe_columns_kernel(const cplx* __restrict__ x, int n, int64_t seg_len, int64_t k, int64_t n_out, uint64_t sea_mask,
                       uint64_t rare_mask, double* __restrict__ out, int by_column) {
  __shared__ double red[7 * 32 + 7];
  const int64_t o = (int64_t)blockIdx.x * seg_len + k;
  if (o >= n_out) return;  // uniform per CTA
  const uint32_t dim = 1u << n;
  const cplx* v = x + ((size_t)blockIdx.x << n);
  const int n_sea = __popcll(sea_mask);
  double s[7] = {0, 0, 0, 0, 0, 0, 0};  // nrm2, zsea, zrare, csea.re, csea.im, crare.re, crare.im
  for (uint32_t i = threadIdx.x; i < dim; i += blockDim.x) {
    const cplx me = v[i];
    const double p = norm2(me);
    s[0] += p;
    s[1] = fma(p, (double)(n_sea - 2 * __popcll((uint64_t)i & sea_mask)), s[1]);
    s[2] += ((uint64_t)i & rare_mask) ? -p : p;
    for (int b = 0; b < n; ++b) {
      const uint32_t m = 1u << b;
      if (i & m) continue;
      const bool is_sea = sea_mask >> b & 1, is_rare = rare_mask >> b & 1;
      if (!is_sea && !is_rare) continue;
      const cplx u = v[i | m];  // conj(me) * u
      const double re = fma(me.x, u.x, me.y * u.y), im = fma(me.x, u.y, -me.y * u.x);
      if (is_sea) { s[3] += re; s[4] += im; }
      if (is_rare) { s[5] += re; s[6] += im; }
    }
  }
  block_reduce<7>(s, red);
  if (threadIdx.x == 0) {
    const double inv = 1.0 / s[0];
    const double val[QSB_NOBS] = {s[3] * inv, s[4] * inv, 0.5 * s[1] * inv, 0.5 * s[2] * inv, s[5] * inv, s[6] * inv, 1.0, sqrt(s[0])};
    for (int j = 0; j < QSB_NOBS; ++j) out[by_column ? (size_t)j * n_out + o : (size_t)o * QSB_NOBS + j] = val[j];
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
  tile_kernel_attrs(pass_centre<false, false, W>, sm);
  tile_kernel_attrs(pass_centre<false, true, W>, sm);
  tile_kernel_attrs(pass_centre<true, false, W>, sm);
  tile_kernel_attrs(pass_centre<true, true, W>, sm);
  tile_kernel_attrs(pass_centre<true, false, W, true>, sm);
  tile_kernel_attrs(pass_centre<true, true, W, true>, sm);
  tile_kernel_attrs(pass_centre<true, false, W, true, true>, sm);
  tile_kernel_attrs(pass_centre<true, true, W, true, true>, sm);
  tile_kernel_attrs(pass_centre<true, false, W, false, true>, sm);
  tile_kernel_attrs(pass_centre<true, true, W, false, true>, sm);
  if constexpr (W != 0) {
    const int sm2 = 2 * TILE * (int)sizeof(cplx);
    tile_kernel_attrs(pass_centre_pipe<false, W>, sm2);
    tile_kernel_attrs(pass_centre_pipe<true, W>, sm2);
  }
}

// Shared memory of a strided pass: the padded buffer only if the layout transforms window [0,4).
static size_t strided_smem(bool pad) { return (size_t)(pad ? TILE_PHYS : TILE) * sizeof(cplx); }

// nb: branches per launch - 2 (both, blockIdx.y) or 1 (only tm.branch0)
// push: non-null for the LAST forward mid pass of a push-only sharded exchange (its output goes to the owners' buffers)
static void launch_mid(const Layout& L, const TileMap& tm, cplx* tx, cplx* ty, unsigned n_tiles, cudaStream_t stream, int nb = 2,
                       const PushSpec* push = nullptr) {
  PushSpec none = {};
  if (push) {
    if (L.m0) pass_mid<true, true><<<dim3(n_tiles, nb), TPT, strided_smem(true), stream>>>(L, tm, tx, ty, *push);
    else pass_mid<false, true><<<dim3(n_tiles, nb), TPT, strided_smem(false), stream>>>(L, tm, tx, ty, *push);
    return;
  }
  if (L.m0) pass_mid<true><<<dim3(n_tiles, nb), TPT, strided_smem(true), stream>>>(L, tm, tx, ty, none);
  else pass_mid<false><<<dim3(n_tiles, nb), TPT, strided_smem(false), stream>>>(L, tm, tx, ty, none);
}

// Opt-in.  Measured on 2 x B200 at N = 31 (profiles/r2k_centre_pipe_2gpu.jsonl): the centre pass alone gains 3 %
// (49.9 vs 51.5 ms per H-apply) but the overlapped H-apply LOSES, 115.7 vs 98.2 ms - its 128 KB of shared memory shrink
// the L1 of the co-resident HBM passes, whose loads in flight are bounded by L1 capacity.
static int centre_pipe_mode() {  // read per launch: tests switch it inside one process
  const char* e = std::getenv("QSB200_CENTRE_PIPE");
  return e ? atoi(e) : 0;
}

template <bool SHARDED, int W>
static void launch_centre(const PassArgs& a, const Layout& L, const TileMap& tm, cplx* tx, cplx* ty, const PeerPtrs& pp,
                          int collect_obs, double* partials, unsigned n_tiles, size_t sm, cudaStream_t stream, int nb,
                          const cplx* rbx = nullptr, const cplx* rby = nullptr) {
  sm = W == 8 ? 0 : strided_smem(W == 0);  // W == 8: the whole group sits in one register window, no staging at all
  if constexpr (SHARDED) {
    if (rbx) {  // push-only exchange: local loads from the receive buffers
      if (tm.count) {
        if (collect_obs) pass_centre<true, true, W, true, true><<<dim3(n_tiles, nb), TPT, sm, stream>>>(a, L, tm, tx, ty, pp, partials, rbx, rby);
        else pass_centre<true, false, W, true, true><<<dim3(n_tiles, nb), TPT, sm, stream>>>(a, L, tm, tx, ty, pp, partials, rbx, rby);
      } else {
        if (collect_obs) pass_centre<true, true, W, false, true><<<dim3(n_tiles, nb), TPT, sm, stream>>>(a, L, tm, tx, ty, pp, partials, rbx, rby);
        else pass_centre<true, false, W, false, true><<<dim3(n_tiles, nb), TPT, sm, stream>>>(a, L, tm, tx, ty, pp, partials, rbx, rby);
      }
      return;
    }
  }
  if constexpr (SHARDED && W != 0) {
    // persistent + double-buffered (cp.async prefetch of the next tile from the peers); QSB200_CENTRE_PIPE=1: on,
    // =2: also in the serialised sharded path (how a one-GPU box tests the kernel)
    if (tm.count && centre_pipe_mode() != 0) {
      const size_t sm2 = 2 * (size_t)TILE * sizeof(cplx);
      if (collect_obs) pass_centre_pipe<true, W><<<dim3(n_tiles, nb), TPT, sm2, stream>>>(a, L, tm, pp, partials);
      else pass_centre_pipe<false, W><<<dim3(n_tiles, nb), TPT, sm2, stream>>>(a, L, tm, pp, partials);
      return;
    }
  }
  if constexpr (SHARDED) {
    if (tm.count) {  // persistent walk (overlapped sharded mode)
      if (collect_obs) pass_centre<true, true, W, true><<<dim3(n_tiles, nb), TPT, sm, stream>>>(a, L, tm, tx, ty, pp, partials);
      else pass_centre<true, false, W, true><<<dim3(n_tiles, nb), TPT, sm, stream>>>(a, L, tm, tx, ty, pp, partials);
      return;
    }
  }
  if (collect_obs) pass_centre<SHARDED, true, W><<<dim3(n_tiles, nb), TPT, sm, stream>>>(a, L, tm, tx, ty, pp, partials);
  else pass_centre<SHARDED, false, W><<<dim3(n_tiles, nb), TPT, sm, stream>>>(a, L, tm, tx, ty, pp, partials);
}
template <bool SHARDED>
static void launch_centre(const PassArgs& a, const Layout& L, const TileMap& tm, cplx* tx, cplx* ty, const PeerPtrs& pp,
                          int collect_obs, double* partials, unsigned n_tiles, size_t sm, cudaStream_t stream, int nb = 2,
                          const cplx* rbx = nullptr, const cplx* rby = nullptr) {
  switch (centre_window(L)) {
    case 8: launch_centre<SHARDED, 8>(a, L, tm, tx, ty, pp, collect_obs, partials, n_tiles, sm, stream, nb, rbx, rby); break;
    case 4: launch_centre<SHARDED, 4>(a, L, tm, tx, ty, pp, collect_obs, partials, n_tiles, sm, stream, nb, rbx, rby); break;
    default: launch_centre<SHARDED, 0>(a, L, tm, tx, ty, pp, collect_obs, partials, n_tiles, sm, stream, nb, rbx, rby); break;
  }
}

StridedFusion plan_fusion(const LargeGeom& g, int n_sms) {
  StridedFusion fu;
  // Opt-in (QSB200_FUSE_LB=20: units of 2^20 amplitudes = 16 MB): measured on B200 the fused launch does cut the
  // DRAM traffic of the strided passes to 2.0 S read + 2.35 S written, but it is bound by tile turnaround inside the
  // SM (shared-memory bandwidth and a three-buffer pipeline), 9.05 ms against 8.76 ms for the three separate
  // HBM-bound launches at N = 28 (profiles/r1m_fused_strided_experiment.md).
  int unit_lb = 0;
  if (const char* e = std::getenv("QSB200_FUSE_LB")) unit_lb = atoi(e);
  const int centre = g.n_layouts - 1;
  if (unit_lb <= 0 || g.world != 1 || centre < 2 || n_sms <= 0) return fu;
  for (int f = 1; f < centre; ++f) {
    int lkmax = 0;
    for (int k = f; k <= centre; ++k) lkmax = std::max(lkmax, g.lay[k].lk);
    // bits a unit fixes: tile pad bits no fused layout transforms or reads across, and the groups of layouts < f
    uint64_t fixed = ((1ull << TB) - 1) & ~((1ull << lkmax) - 1);
    for (int k = 1; k < f; ++k) fixed |= ((1ull << (g.lay[k].ge - g.lay[k].gs)) - 1) << g.lay[k].gs;
    const int fbits = __builtin_popcountll(fixed);
    if (g.n_loc - fbits > unit_lb) continue;
    FusedSpec& s = fu.spec;
    s.f = f;
    s.m = centre - f;
    s.fbits = fbits;
    s.tiles_lb = g.n_loc - TB - fbits;
    s.vmask = (1u << (g.n_loc - TB)) - 1;
    for (int k = f; k <= centre; ++k) {
      unsigned vm = 0;
      int pos = 0;
      for (int b = 0; b < g.n_loc; ++b) {
        if (g.lay[k].tile_mask >> b & 1) continue;  // outer bits of the layout, ascending = bits of v
        if (fixed >> b & 1) vm |= 1u << pos;
        ++pos;
      }
      s.fixed_vmask[k] = vm;
    }
    // A dependant is drawn lag P T - 2 T tickets after the last tile of its dependency: make that at least one wave
    // of CTAs where the units in flight ((P - 1) lag + 1 of them) still fit comfortably in L2.
    const int P = 2 * s.m + 1, grid = n_sms;
    const long long PT = (long long)P << s.tiles_lb, T = 1ll << s.tiles_lb;
    const double unit_mb = (double)(sizeof(cplx) << (g.n_loc - fbits)) / (1 << 20);
    s.lag = 1;
    while (s.lag * PT - 2 * T < NBUF * grid && ((P - 1) * (s.lag + 1) + 1) * unit_mb <= 64.0) ++s.lag;
    if (const char* e = std::getenv("QSB200_FUSE_LAG")) s.lag = std::max(1, atoi(e));
    fu.grid = grid;
    fu.ctrl_bytes = sizeof(int) * (1 + ((size_t)2 << fbits) * P);
    fu.on = true;
    return fu;
  }
  return fu;
}

// (No carve-out preference is set: the driver's default leaves the rest of the 256 KB to L1, and L1 capacity is what
// bounds the loads in flight - with the maximum carve-out pass_mid drops from 6.7 to 5.2 TB/s.)
template <typename K>
static void tile_kernel_attrs(K kernel, int sm) {
  cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, sm);
}

void prepare_device() {
  const int sm = (int)((size_t)TILE_PHYS * sizeof(cplx));
  tile_kernel_attrs(pass_first<false>, sm);
  tile_kernel_attrs(pass_first<true>, sm);
  tile_kernel_attrs(pass_first<true, true>, sm);
  tile_kernel_attrs(pass_last<0, 8, 0>, sm);
  tile_kernel_attrs(pass_last_first, sm);
  tile_kernel_attrs(pass_last<1, 4, 2>, sm);
  tile_kernel_attrs(pass_last<2, 4, 2>, sm);
  tile_kernel_attrs(pass_last<3, 4, 2>, sm);
  tile_kernel_attrs(pass_mid<true>, sm);
  tile_kernel_attrs(pass_mid<false>, sm);
  tile_kernel_attrs(pass_mid<true, true>, sm);
  tile_kernel_attrs(pass_mid<false, true>, sm);
  set_centre_attrs<8>(sm);
  set_centre_attrs<4>(sm);
  set_centre_attrs<0>(sm);
  const int fsm = (int)sizeof(FusedSmem);
  cudaFuncSetAttribute(pass_strided_fused<8, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fsm);
  cudaFuncSetAttribute(pass_strided_fused<8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, fsm);
  cudaFuncSetAttribute(pass_strided_fused<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fsm);
  cudaFuncSetAttribute(pass_strided_fused<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, fsm);
  cudaFuncSetAttribute(pass_centre_ring<8, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fsm);
  cudaFuncSetAttribute(pass_centre_ring<8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, fsm);
  cudaFuncSetAttribute(pass_centre_ring<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fsm);
  cudaFuncSetAttribute(pass_centre_ring<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, fsm);
}

// The strided-layout passes of one H-apply (mids forward, centre, mids back).
// (Issuing mid -> centre -> mid chunk batch by chunk batch through TileMap so that the 2nd and 3rd pass of a
// batch would hit L2 was measured and is slower than three full-grid launches: profiles/r1g_l2_chunk_experiment.log.)
// Sharded: every rank runs the centre pass on its 1/world slice of the centre tiles, reading and writing the
// peers' tx/ty in place; the rank barrier before it makes the peers' forward passes visible, the one after it
// keeps the peers' backward passes from starting on rows that are still being written.
// cuTensorMapEncodeTiled through the runtime's driver entry point query (libqsb200 does not link libcuda).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) != cudaSuccess ||
        qr != cudaDriverEntryPointSuccess)
      p = nullptr;
    cudaGetLastError();
    return (EncodeTiledFn)p;
  }();
  return fn;
}

// The vector `base` (2^n_loc cplx) as a 5-d tensor of doubles whose box {2^(a+1), 2^(lk-a), 1, 2^(ge-gs), 1} is one
// tile of layout L in tile-local order: dims = pad bits [0,a) x 2 doubles, pad bits [a,lk), bits [lk,gs), the group
// [gs,ge), bits [ge,n_loc).  a = min(lk, 7) keeps the innermost box extent within TMA's 256 elements.
// A sharded centre group ends above n_loc (its top bits are the rank): the box then covers the LOCAL group bits only,
// i.e. the slab of the tile that this vector's owner holds.
static bool encode_tile_map(CUtensorMap* map, const Layout& L, int n_loc, void* base) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc) return false;
  const int a = std::min(L.lk, 7);
  const int ge = std::min(L.ge, n_loc);
  const cuuint64_t dims[5] = {2ull << a, 1ull << (L.lk - a), 1ull << (L.gs - L.lk), 1ull << (ge - L.gs), 1ull << (n_loc - ge)};
  const cuuint64_t strides[4] = {sizeof(cplx) << a, sizeof(cplx) << L.lk, sizeof(cplx) << L.gs, sizeof(cplx) << ge};
  const cuuint32_t box[5] = {2u << a, 1u << (L.lk - a), 1u, 1u << (ge - L.gs), 1u};
  const cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 5, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static bool ensure_tile_maps(const LargeGeom& g, StridedFusion& fu, cplx* tx, cplx* ty) {
  if (fu.maps_tx == tx && fu.maps_ty == ty) return true;
  for (int k = fu.spec.f; k < g.n_layouts; ++k)
    if (!encode_tile_map(&fu.maps[0][k], g.lay[k], g.n_loc, tx) || !encode_tile_map(&fu.maps[1][k], g.lay[k], g.n_loc, ty))
      return false;
  fu.maps_tx = tx;
  fu.maps_ty = ty;
  return true;
}

static bool ensure_ring_maps(const LargeGeom& g, ShardLink& link) {
  if (link.ring_ready) return true;
  const Layout& L = g.lay[g.n_layouts - 1];
  for (int r = 0; r < g.world; ++r)
    if (!encode_tile_map(&link.ring_maps[0][r], L, g.n_loc, link.tx[r]) || !encode_tile_map(&link.ring_maps[1][r], L, g.n_loc, link.ty[r]))
      return false;
  link.ring_ready = true;
  return true;
}

// Opt-in (QSB200_SHARD_RING=1).  Measured on 2 x B200 at N = 28 the ring and the register-staged P2P centre pass are
// both NVLink-bound (691 vs 703 GB/s per direction, 91 % of the peer-copy peak), so the ring buys nothing there; with
// one process per GPU (CUDA-IPC mappings) and >= 8 GB vectors its TMA loads from the peers did not complete within
// the watchdog, while ranks that are threads of one process work at every size (profiles/r1s_shard_ring_experiment.md).
static bool shard_ring_enabled() {
  static int on = -1;
  if (on < 0) { const char* e = std::getenv("QSB200_SHARD_RING"); on = (e && *e == '1') ? 1 : 0; }
  return on == 1;
}

template <int W>
static void launch_centre_ring(const PassArgs& a, const LargeGeom& g, const ShardLink& link, int collect_obs, double* partials,
                               unsigned n_tiles, cudaStream_t stream) {
  TmaMaps maps;
  for (int br = 0; br < 2; ++br) memcpy(maps.m[br], link.ring_maps[br], sizeof(CUtensorMap) * MAX_WORLD);
  int dev = 0, n_sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
  const int grid = (int)std::min<unsigned>(2u * n_tiles, (unsigned)std::max(1, n_sms));
  const int threads = FUSED_NG * TPT + 64, centre = g.n_layouts - 1;
  const unsigned v0 = (unsigned)g.rank * n_tiles;
  if (collect_obs) pass_centre_ring<W, true><<<grid, threads, sizeof(FusedSmem), stream>>>(a, centre, v0, n_tiles, g.world, maps, partials);
  else pass_centre_ring<W, false><<<grid, threads, sizeof(FusedSmem), stream>>>(a, centre, v0, n_tiles, g.world, maps, partials);
}

template <int W>
static void launch_fused(const PassArgs& a, const StridedFusion& fu, int collect_obs, double* partials, cudaStream_t stream) {
  TmaMaps maps;
  for (int br = 0; br < 2; ++br) memcpy(maps.m[br], fu.maps[br], sizeof(CUtensorMap) * MAX_LAYOUTS);
  const int threads = FUSED_NG * TPT + 64;
  if (collect_obs) pass_strided_fused<W, true><<<fu.grid, threads, sizeof(FusedSmem), stream>>>(a, fu.spec, maps, partials, fu.ctrl);
  else pass_strided_fused<W, false><<<fu.grid, threads, sizeof(FusedSmem), stream>>>(a, fu.spec, maps, partials, fu.ctrl);
}

static int launch_strided_passes(const LargeGeom& g, const PassArgs& a, cplx* tx, cplx* ty, int collect_obs,
                                 double* partials, cudaStream_t stream, ShardLink* link, StridedFusion* fu) {
  const size_t sm1 = (size_t)TILE_PHYS * sizeof(cplx);
  const unsigned n_tiles = 1u << (g.n_loc - TB);
  const int centre = g.n_layouts - 1;
  const TileMap all = {0u, 0, 0, 0, 0u, 0u, 0, 0u};
  const double S = (double)(sizeof(cplx) << g.n_loc);  // bytes of one local vector
  if (fu && fu->on && g.world == 1 && centre_window(g.lay[centre]) != 0 && ensure_tile_maps(g, *fu, tx, ty)) {
    // layouts before f stay full-grid launches; f .. centre .. f is one persistent launch
    const int f = fu->spec.f;
    for (int k = 1; k < f; ++k) { Timed t(1, stream, 4 * S); launch_mid(g.lay[k], all, tx, ty, n_tiles, stream); }
    {
      Timed t(6, stream, 4 * S);
      cudaMemsetAsync(fu->ctrl, 0, fu->ctrl_bytes, stream);
      if (centre_window(g.lay[centre]) == 8) launch_fused<8>(a, *fu, collect_obs, partials, stream);
      else launch_fused<4>(a, *fu, collect_obs, partials, stream);
    }
    for (int k = f - 1; k >= 1; --k) { Timed t(1, stream, 4 * S); launch_mid(g.lay[k], all, tx, ty, n_tiles, stream); }
    return QSB_OK;
  }
  const bool overlapped = g.world > 1 && link && link->stream_barrier &&
                          !(shard_ring_enabled() && centre_window(g.lay[centre]) != 0);
  const bool push_serial = !overlapped && g.world > 1 && link && link->push && centre >= 2 &&
                           !(shard_ring_enabled() && centre_window(g.lay[centre]) != 0);
  PushSpec push_s = {};
  if (push_serial) {
    for (int r = 0; r < g.world; ++r) { push_s.r[0][r] = link->rx[r]; push_s.r[1][r] = link->ry[r]; }
    push_s.lk_c = g.lay[centre].lk;
    push_s.gs_c = g.lay[centre].gs;
    push_s.tl = g.n_loc - TB;
    push_s.rank_hi = (uint64_t)g.rank << (g.n_loc - g.lay[centre].gs);
    push_s.on = 1;
  }
  if (!overlapped)
    for (int k = 1; k < centre; ++k) {
      Timed t(1, stream, 4 * S);
      launch_mid(g.lay[k], all, tx, ty, n_tiles, stream, 2, (push_serial && k == centre - 1) ? &push_s : nullptr);
    }
  if (g.world == 1) {
    Timed t(2, stream, 4 * S);
    // experiment knob: the TMA-ring centre kernel (one persistent CTA per SM, three tile buffers) on an unsharded plan
    static const bool ring1 = [] { const char* e = std::getenv("QSB200_CENTRE_RING"); return e && *e == '1'; }();
    const int w = centre_window(g.lay[centre]);
    const bool top = g.lay[centre].ge == g.n_loc;  // the ring's box coordinates assume the group reaches the top bit
    if (ring1 && w != 0 && top && fu && encode_tile_map(&fu->maps[0][0], g.lay[centre], g.n_loc, tx) &&
        encode_tile_map(&fu->maps[1][0], g.lay[centre], g.n_loc, ty)) {
      ShardLink self = {};
      self.ring_maps[0][0] = fu->maps[0][0];
      self.ring_maps[1][0] = fu->maps[1][0];
      LargeGeom g1 = g;
      g1.rank = 0;
      if (w == 8) launch_centre_ring<8>(a, g1, self, collect_obs, partials, n_tiles, stream);
      else launch_centre_ring<4>(a, g1, self, collect_obs, partials, n_tiles, stream);
    } else {
      PeerPtrs none = {};
      launch_centre<false>(a, g.lay[centre], all, tx, ty, none, collect_obs, partials, n_tiles, sm1, stream);
    }
  } else {
    if (!link) { set_error("sharded plan is not connected to its peers (qsb_shard_connect)"); return QSB_ERR_BAD_ARG; }
    PeerPtrs pp = {};
    for (int r = 0; r < g.world; ++r) { pp.t[0][r] = link->tx[r]; pp.t[1][r] = link->ty[r]; }
    const int w = centre_window(g.lay[centre]);
    // push-only exchange: needs a mid pass to deliver the tiles (centre >= 2) and the peers' receive buffers
    PushSpec push = {};
    const bool pushing = link->push && centre >= 2 && !(shard_ring_enabled() && w != 0);
    if (pushing) {
      for (int r = 0; r < g.world; ++r) { push.r[0][r] = link->rx[r]; push.r[1][r] = link->ry[r]; }
      push.lk_c = g.lay[centre].lk;
      push.gs_c = g.lay[centre].gs;
      push.tl = g.n_loc - TB;
      push.rank_hi = (uint64_t)g.rank << (g.n_loc - g.lay[centre].gs);
      push.on = 1;
    }
    const cplx* rbx = pushing ? link->rx[g.rank] : nullptr;
    const cplx* rby = pushing ? link->ry[g.rank] : nullptr;
    if (overlapped) {
      // Overlapped: a software pipeline over chunk groups.  Index bits [lmax, 12) - pad bits of every strided tile, lmax
      // the largest lk - are transformed by no strided pass, so they cut the vectors into 2^(12 - lmax) chunks whose
      // strided passes are independent; K = 2^pipe_lb groups of chunks are pipelined:
      //   plan stream S (high priority):  fwd 0, B, fwd 1, B, [cen 0] back 0, fwd 2, B, [cen 1] back 1, ... [cen K-1] back K-1
      //   centre stream C (low priority): [fwd c] centre x c, centre y c, B       for c = 0 .. K-1
      // (fwd / back: the mid passes of both branches on one group; B: a rank barrier queued on the stream - an NCCL
      // collective, the host never waits; [..]: an event wait).  The centre pass is NVLink-bound and persistent, one
      // CTA per SM, so the HBM-bound launches of S keep the second CTA slot of every SM; with K >= 3 the links stay
      // busy from the end of fwd 0 to the start of back K-1.  Every rank queues the same barriers in the same order
      // on each stream (one communicator per stream).
      cudaStream_t C = link->centre_stream;
      int lmax = 0;
      for (int k = 1; k <= centre; ++k) lmax = std::max(lmax, g.lay[k].lk);
      // chunk field = index bits [lmax, lmax + cw): below 12, and below the bits of the centre layout's tile number
      // that say which rank owns the tile (its top log2(world) bits = index bits [gs - wl, gs)); tiny shards only
      int wl = 0;
      while ((1 << wl) < g.world) ++wl;
      const int cw = std::max(0, std::min(TB, g.lay[centre].gs - wl) - lmax);
      const int plb = std::max(0, std::min(std::min(link->pipe_lb, cw), 4));
      const int K = 1 << plb, lb = cw - plb;                     // groups, log2 chunks per group
      const unsigned per_group = n_tiles >> plb;                 // tiles of one layout in one group
      auto group_map = [&](const Layout& L, int c, unsigned vbase, int br, unsigned count) {
        TileMap m = {vbase, lmax - L.lk, cw, lb, (unsigned)c << lb, 0u, br, count};
        return m;
      };
      auto fwd = [&](int c) -> int {
        for (int k = 1; k < centre; ++k) {
          Timed t(1, stream, 4 * S / K);
          launch_mid(g.lay[k], group_map(g.lay[k], c, 0u, 0, 0u), tx, ty, per_group, stream, 2,
                     (pushing && k == centre - 1) ? &push : nullptr);
        }
        if (link->stream_barrier(link->barrier_ctx, stream)) { set_error("rank barrier (stream hook) failed"); return QSB_ERR_CUDA; }
        cudaEventRecord(link->ev_fwd[c], stream);
        cudaStreamWaitEvent(C, link->ev_fwd[c], 0);
        for (int br = 0; br < 2; ++br) {
          Timed t(2, C, 2 * S / K);
          launch_centre<true>(a, g.lay[centre], group_map(g.lay[centre], c, (unsigned)g.rank * n_tiles, br, per_group), tx, ty,
                              pp, collect_obs, partials, std::min<unsigned>(per_group, (unsigned)link->centre_ctas), sm1, C, 1,
                              rbx, rby);
        }
        if (link->stream_barrier(link->barrier_ctx, C)) { set_error("rank barrier (stream hook) failed"); return QSB_ERR_CUDA; }
        cudaEventRecord(link->ev_centre[c], C);
        return QSB_OK;
      };
      auto back = [&](int c) {
        cudaStreamWaitEvent(stream, link->ev_centre[c], 0);
        for (int k = centre - 1; k >= 1; --k) {
          Timed t(1, stream, 4 * S / K);
          launch_mid(g.lay[k], group_map(g.lay[k], c, 0u, 0, 0u), tx, ty, per_group, stream);
        }
      };
      int rc = fwd(0);
      for (int c = 1; c < K && !rc; ++c) {
        rc = fwd(c);
        back(c - 1);
      }
      if (rc) return rc;
      back(K - 1);
      return QSB_OK;
    }
    int rc = link->sync(link->ctx);
    if (rc) return rc;
    {
      Timed t(2, stream, 4 * S);
      if (shard_ring_enabled() && w != 0 && ensure_ring_maps(g, *link)) {
        if (w == 8) launch_centre_ring<8>(a, g, *link, collect_obs, partials, n_tiles, stream);
        else launch_centre_ring<4>(a, g, *link, collect_obs, partials, n_tiles, stream);
      } else {
        if (centre_pipe_mode() == 2 && w != 0) {
          int dev = 0, n_sms = 148;
          cudaGetDevice(&dev);
          cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
          const TileMap mine = {(unsigned)g.rank * n_tiles, 0, 0, 0, 0u, 0u, 0, n_tiles};
          launch_centre<true>(a, g.lay[centre], mine, tx, ty, pp, collect_obs, partials,
                              std::min<unsigned>(n_tiles, (unsigned)n_sms), sm1, stream);
        } else {
          const TileMap mine = {(unsigned)g.rank * n_tiles, 0, 0, 0, 0u, 0u, 0, 0u};
          launch_centre<true>(a, g.lay[centre], mine, tx, ty, pp, collect_obs, partials, n_tiles, sm1, stream, 2, rbx, rby);
        }
      }
    }
    rc = link->sync(link->ctx);
    if (rc) return rc;
  }
  for (int k = centre - 1; k >= 1; --k) { Timed t(1, stream, 4 * S); launch_mid(g.lay[k], all, tx, ty, n_tiles, stream); }
  return QSB_OK;
}

int launch_happly(const LargeGeom& g, const DiagTables& tab, const cplx* src, cplx* tx, cplx* ty,
                  const Epilogue& ep, int collect_obs, double* partials, cudaStream_t stream, ShardLink* link,
                  const Combine* comb, StridedFusion* fu, const FuseNext* next, bool skip_first) {
  const size_t sm2 = (size_t)TILE_PHYS * sizeof(cplx);
  PassArgs a;
  a.g = g;
  a.tab = tab;
  const unsigned n_tiles = 1u << (g.n_loc - TB);
  if (ep.n_in > 2) { set_error("internal: fused epilogues take at most 2 input vectors"); return QSB_ERR_BAD_ARG; }
  const double S = (double)(sizeof(cplx) << g.n_loc);
  if (!skip_first) {
    const bool cmb = comb && comb->n_in > 0;
    Timed t(cmb ? 4 : 0, stream, cmb ? (comb->n_in + 3 + (comb->eout ? 1 : 0)) * S : 3 * S);
    if (comb && comb->n_in > 0) {
      if (comb->out != src) { set_error("internal: Combine::out must be the vector the apply acts on"); return QSB_ERR_BAD_ARG; }
      if (comb->eout) pass_first<true, true><<<n_tiles, TPT, sm2, stream>>>(a, src, tx, ty, *comb, collect_obs, partials);
      else pass_first<true><<<n_tiles, TPT, sm2, stream>>>(a, src, tx, ty, *comb, collect_obs, partials);
    } else {
      Combine none = {};
      pass_first<false><<<n_tiles, TPT, sm2, stream>>>(a, src, tx, ty, none, collect_obs, partials);
    }
  }
  int rc = launch_strided_passes(g, a, tx, ty, collect_obs, partials, stream, link, fu);
  if (rc) return rc;
  if (next) {
    if (ep.o2 || ep.want_err || ep.raw_h || ep.alpha != 1.0 || !ep.o1 || next->cb.out != src) {
      set_error("internal: pass_last_first needs a plain k-producing epilogue and next->cb.out == src");
      return QSB_ERR_BAD_ARG;
    }
    Timed t(7, stream, (3 + 1 + next->cb.n_in + 3) * S);
    pass_last_first<<<n_tiles, TPT, sm2, stream>>>(a, const_cast<cplx*>(src), tx, ty, ep.o1, next->cb, next->c_fresh);
    return QSB_OK;
  }
  {
    static const int last_pf = [] { const char* e = std::getenv("QSB200_LAST_PF"); return e ? atoi(e) : 0; }();
    Epilogue epx = ep;
    epx.pf = last_pf;
    const Epilogue& ep = epx;
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
                   double* partials, cudaStream_t stream, ShardLink* link, StridedFusion* fu) {
  // the transform half of an H-apply: z sums in pass_first, x/y sums in pass_centre (tx/ty are clobbered)
  const size_t sm2 = (size_t)TILE_PHYS * sizeof(cplx);
  PassArgs a;
  a.g = g;
  a.tab = tab;
  const unsigned n_tiles = 1u << (g.n_loc - TB);
  Combine none = {};
  pass_first<false><<<n_tiles, TPT, sm2, stream>>>(a, src, tx, ty, none, 1, partials);
  return launch_strided_passes(g, a, tx, ty, 1, partials, stream, link, fu);
}

void launch_error_sums(const LargeGeom& g, const ErrSpec& es, double* partials, cudaStream_t stream) {
  Timed t(5, stream, es.n_in * (double)(sizeof(cplx) << g.n_loc));
  error_sums_kernel<<<1u << (g.n_loc - TB), TPT, 0, stream>>>(es, partials);
}

void launch_reduce_partials(const double* partials, int64_t n_tiles, double* out8, cudaStream_t stream) {
  reduce_partials_kernel<<<1, 1024, 0, stream>>>(partials, n_tiles, out8);
}

void launch_product_state(cplx* psi, const ProductPhi& phi, int n, int n_loc, uint64_t rank_bits, cudaStream_t stream) {
  product_state_kernel<<<1u << (n_loc - TB), 256, 0, stream>>>(psi, phi, n, rank_bits);
}

void launch_gather_rows(const cplx* src, const uint64_t* d_rows, int64_t n_rows, cplx* d_out, cudaStream_t stream) {
  gather_rows_kernel<<<(unsigned)((n_rows + 255) / 256), 256, 0, stream>>>(src, d_rows, n_rows, d_out);
}

void launch_dense_eval(const DenseSpec& d, uint64_t dim, cudaStream_t stream) {
  const int blocks = (int)std::min<uint64_t>((dim + 255) / 256, 148 * 8);
  dense_eval_kernel<<<blocks, 256, 0, stream>>>(d, dim);
}

double measure_fp64_peak_tflops(cudaStream_t stream) {
  double* d = nullptr;
  if (cudaMalloc(&d, sizeof(double)) != cudaSuccess) { cudaGetLastError(); return 0.0; }
  int dev = 0, n_sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev);
  const int blocks = n_sms * 8, iters = 1 << 15;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  fp64_peak_kernel<<<blocks, 256, 0, stream>>>(d, 1 << 10, 0.999999, 1e-9);  // warm-up
  double best = 0.0;
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0, stream);
    fp64_peak_kernel<<<blocks, 256, 0, stream>>>(d, iters, 0.999999, 1e-9);
    cudaEventRecord(e1, stream);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double tf = 2.0 * 8.0 * 256.0 * blocks * (double)iters / (ms * 1e-3) / 1e12;
    if (tf > best) best = tf;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(d);
  return best;
}

void launch_set_identity(cplx* psi, int n, cudaStream_t stream) {
  set_identity_kernel<<<148 * 8, 256, 0, stream>>>(psi, n);
}

void launch_observe_columns(const cplx* x, int n, int n_cols, int64_t seg_len, int64_t k, int64_t n_out, uint64_t sea_mask,
                            uint64_t rare_mask, double* out, int by_column, cudaStream_t stream) {
  observe_columns_kernel<<<n_cols, 256, 0, stream>>>(x, n, seg_len, k, n_out, sea_mask, rare_mask, out, by_column);
}

void launch_set_basis_state(cplx* psi, uint64_t dim, uint64_t index, cudaStream_t stream) {
  int blocks = (int)std::min<uint64_t>((dim + 255) / 256, 148 * 16);
  set_basis_kernel<<<blocks, 256, 0, stream>>>(psi, dim, index);
}

}  // namespace qsb
