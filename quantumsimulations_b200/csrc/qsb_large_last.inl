// pass_last + small utility kernels + launchers; included from qsb_large_kernels.inl.
namespace qsb {
namespace {

// ---------------------------------------------------------------------------------------------------
// pass_last: layout 0.  k = -i (Dz v + W0 [tX + c(v) (M/2)^(x12) tY]), then the fused stage algebra
// (Epilogue): o1 = alpha k, o2 = sum c_j vin_j + ck k, err += |sum e_j vin_j + ek k|^2 / sc^2.
// The last butterfly round leaves thread t with amplitudes t + 256 j, so the epilogue streams the other
// vectors straight from registers with fully coalesced accesses.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TPT, 2)
pass_last(PassArgs a, const cplx* src, const cplx* __restrict__ tx, const cplx* __restrict__ ty,
          Epilogue ep, double* __restrict__ partials) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cplx* sm = reinterpret_cast<cplx*>(smem_raw);
  __shared__ DiagShared ds;
  __shared__ double red[2 * 32 + 2];
  const int tt = threadIdx.x;
  const uint64_t base = (uint64_t)blockIdx.x << TB;
  const int n = a.g.n;
  cplx r[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) r[j] = ty[base | (uint64_t)elem<8>(tt, j)];
  {  // pull the tX tile (needed three rounds from now) towards L2 while the tY loads are in flight
    const char* p = reinterpret_cast<const char*>(tx + base) + (size_t)tt * 256;
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p + 128));
  }
  diag_prologue(ds, a.tab.h + 2 * (size_t)n, a.tab.J + 2 * (size_t)n * n, n, a.g.lay[0], base, 1.0, tt);
  bfly16<1>(r, 0xF);
  smem_store16<8>(sm, tt, r);
  __syncthreads();
  smem_round<0, 1>(sm, tt, 0xF);
  __syncthreads();
  smem_load16<4>(sm, tt, r);
  bfly16<1>(r, 0xF);
  const int pv = __popcll(base);  // S_rest = (+i)^popcount(outer bits)
#pragma unroll
  for (int j = 0; j < 16; ++j) r[j] = mul_pi_pow(INV_TILE * r[j], pv) + tx[base | (uint64_t)elem<4>(tt, j)];
  bfly16<0>(r, 0xF);
  smem_store16<4>(sm, tt, r);
  __syncthreads();
  smem_round<0, 0>(sm, tt, 0xF);
  __syncthreads();
  smem_load16<8>(sm, tt, r);
  bfly16<0>(r, 0xF);

  double err_acc = 0.0, aux_acc = 0.0;
#pragma unroll 4
  for (int j = 0; j < 16; ++j) {
    const int i = elem<8>(tt, j);
    const uint64_t idx = base | (uint64_t)i;
    cplx v = src[idx];
    cplx acc = diag_at(ds, a.tab.Qz, i) * v + r[j];
    cplx k = ep.raw_h ? acc : mul_mi(acc);
    if (ep.o1) ep.o1[idx] = ep.alpha * k;
    if (ep.o2 || ep.want_err) {
      cplx o2 = ep.ck * k, er = ep.ek * k;
      double m0 = 0.0, m1 = 0.0;
#pragma unroll
      for (int q = 0; q < MAX_VEC_IN; ++q) {
        if (q < ep.n_in) {
          cplx w = ep.vin[q][idx];
          axpy(o2, ep.c[q], w);
          axpy(er, ep.e[q], w);
          if (q == ep.sc_i0) m0 = norm2(w);
          if (q == ep.sc_i1) m1 = norm2(w);
        }
      }
      if (ep.o2) ep.o2[idx] = o2;
      if (ep.want_err) {
        if (ep.sc_i1 < 0) m1 = norm2(o2);
        double sc = ep.atol + ep.rtol * sqrt(fmax(m0, m1));
        err_acc += norm2(er) / (sc * sc);
        aux_acc += m0 / (sc * sc);
      }
    }
  }
  if (ep.want_err) {
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
  for (; i < dim; i += stride) psi[i] = mk(i == index ? 1.0 : 0.0, 0.0);
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
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  Timed(int c, cudaStream_t s) : cls(c), st(s) {
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
      cudaEventDestroy(e0); cudaEventDestroy(e1);
    }
  }
};
}  // namespace

void launch_happly(const LargeGeom& g, const DiagTables& tab, const cplx* src, cplx* tx, cplx* ty,
                   const Epilogue& ep, int collect_obs, double* partials, cudaStream_t stream) {
  static bool attr = false;
  const size_t sm1 = (size_t)TILE_PHYS * sizeof(cplx), sm2 = sm1;
  if (!attr) {
    cudaFuncSetAttribute(pass_first, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2);
    cudaFuncSetAttribute(pass_last, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2);
    cudaFuncSetAttribute(pass_mid, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1);
    cudaFuncSetAttribute(pass_centre, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1);
    attr = true;
  }
  PassArgs a;
  a.g = g;
  a.tab = tab;
  const unsigned n_tiles = 1u << (g.n - TB);
  const int centre = g.n_layouts - 1;
  { Timed t(0, stream); pass_first<<<n_tiles, TPT, sm2, stream>>>(a, src, tx, ty, collect_obs, partials); }
  for (int k = 1; k < centre; ++k) { Timed t(1, stream); pass_mid<<<dim3(n_tiles, 2), TPT, sm1, stream>>>(g.lay[k], tx, ty); }
  { Timed t(2, stream); pass_centre<<<dim3(n_tiles, 2), TPT, sm1, stream>>>(a, g.lay[centre], tx, ty, collect_obs, partials); }
  for (int k = centre - 1; k >= 1; --k) { Timed t(1, stream); pass_mid<<<dim3(n_tiles, 2), TPT, sm1, stream>>>(g.lay[k], tx, ty); }
  { Timed t(3, stream); pass_last<<<n_tiles, TPT, sm2, stream>>>(a, src, tx, ty, ep, partials); }
}

void launch_observe(const LargeGeom& g, const DiagTables& tab, const cplx* src, cplx* tx, cplx* ty,
                    double* partials, cudaStream_t stream) {
  // the transform half of an H-apply: z sums in pass_first, x/y sums in pass_centre (tx/ty are clobbered)
  Epilogue none = {};
  (void)none;
  const size_t sm1 = (size_t)TILE_PHYS * sizeof(cplx), sm2 = sm1;
  PassArgs a;
  a.g = g;
  a.tab = tab;
  const unsigned n_tiles = 1u << (g.n - TB);
  const int centre = g.n_layouts - 1;
  cudaFuncSetAttribute(pass_first, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm2);
  cudaFuncSetAttribute(pass_mid, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1);
  cudaFuncSetAttribute(pass_centre, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm1);
  pass_first<<<n_tiles, TPT, sm2, stream>>>(a, src, tx, ty, 1, partials);
  for (int k = 1; k < centre; ++k) pass_mid<<<dim3(n_tiles, 2), TPT, sm1, stream>>>(g.lay[k], tx, ty);
  pass_centre<<<dim3(n_tiles, 2), TPT, sm1, stream>>>(a, g.lay[centre], tx, ty, 1, partials);
}

void launch_reduce_partials(const double* partials, int64_t n_tiles, double* out8, cudaStream_t stream) {
  reduce_partials_kernel<<<1, 1024, 0, stream>>>(partials, n_tiles, out8);
}

void launch_set_basis_state(cplx* psi, uint64_t dim, uint64_t index, cudaStream_t stream) {
  int blocks = (int)std::min<uint64_t>((dim + 255) / 256, 148 * 16);
  set_basis_kernel<<<blocks, 256, 0, stream>>>(psi, dim, index);
}

}  // namespace qsb
