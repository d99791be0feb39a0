// Microbenchmark for the cluster-tile plan of DESIGN.md section 10: the all-to-all a 2^c-CTA cluster would run once per
// pass to exchange the cross-CTA bits of a (12 + c)-bit tile.  Every CTA holds a 64 KB tile (4096 complex128) in
// shared memory; per iteration it sends slab p (4096 / C amplitudes) to CTA p of its cluster through distributed shared
// memory (16-byte st.shared::cluster, or 16-byte ld.shared::cluster pulls with --pull), then the cluster synchronises.
// Prints bytes per cycle per SM and GB/s per SM for cluster sizes 2 .. 16.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/dsmem_alltoall.bin scripts/dsmem_alltoall.cu
//   scripts/dsmem_alltoall.bin [--pull]
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>

namespace cg = cooperative_groups;

constexpr int TILE = 4096;  // complex128 amplitudes per CTA
constexpr int THREADS = 256;

template <bool PULL>
__global__ void __launch_bounds__(THREADS, 1) alltoall_kernel(int iters, double2* sink, long long* cycles) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double2* a = reinterpret_cast<double2*>(smem_raw);  // source tile
  double2* b = a + TILE;                              // destination tile
  cg::cluster_group cluster = cg::this_cluster();
  const int C = cluster.num_blocks(), r = cluster.block_rank();
  const int slab = TILE / C;
  for (int i = threadIdx.x; i < TILE; i += THREADS) a[i] = make_double2(r, i);
  cluster.sync();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    for (int pp = 1; pp < C; ++pp) {  // start at the next rank so that the clusters' traffic is spread over the peers
      const int p = (r + pp) % C;
      if (PULL) {
        const double2* remote = cluster.map_shared_rank(a, p);
        for (int i = threadIdx.x; i < slab; i += THREADS) b[p * slab + i] = remote[r * slab + i];
      } else {
        double2* remote = cluster.map_shared_rank(b, p);
        for (int i = threadIdx.x; i < slab; i += THREADS) remote[r * slab + i] = a[p * slab + i];
      }
    }
    for (int i = threadIdx.x; i < slab; i += THREADS) b[r * slab + i] = a[r * slab + i];
    cluster.sync();
    double2* t = a; a = b; b = t;
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) {
    cycles[blockIdx.x] = t1 - t0;
    sink[blockIdx.x] = a[(blockIdx.x * 7) % TILE];
  }
}

template <bool PULL>
static int run(int C, int iters) {
  auto kern = alltoall_kernel<PULL>;
  const size_t smem = 2 * TILE * sizeof(double2);
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (C > 8) cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(THREADS);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = C; attr.val.clusterDim.y = 1; attr.val.clusterDim.z = 1;
  cfg.attrs = &attr; cfg.numAttrs = 1;
  int n_clusters = 0;
  cfg.gridDim = dim3(C);
  cudaError_t e = cudaOccupancyMaxActiveClusters(&n_clusters, kern, &cfg);
  if (e != cudaSuccess || n_clusters < 1) {
    printf("{\"cluster\": %d, \"error\": \"%s\"}\n", C, cudaGetErrorString(e));
    cudaGetLastError();
    return 1;
  }
  const int grid = n_clusters * C;
  cfg.gridDim = dim3(grid);
  double2* sink; long long* cyc;
  cudaMalloc(&sink, grid * sizeof(double2));
  cudaMalloc(&cyc, grid * sizeof(long long));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaLaunchKernelEx(&cfg, kern, 10, sink, cyc);  // warm-up
  cudaEventRecord(e0);
  e = cudaLaunchKernelEx(&cfg, kern, iters, sink, cyc);
  cudaEventRecord(e1);
  cudaError_t e2 = cudaDeviceSynchronize();
  if (e != cudaSuccess || e2 != cudaSuccess) {
    printf("{\"cluster\": %d, \"error\": \"%s / %s\"}\n", C, cudaGetErrorString(e), cudaGetErrorString(e2));
    return 1;
  }
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  long long* h = new long long[grid];
  cudaMemcpy(h, cyc, grid * sizeof(long long), cudaMemcpyDeviceToHost);
  double mean = 0.0; long long worst = 0;
  for (int i = 0; i < grid; ++i) { mean += (double)h[i]; if (h[i] > worst) worst = h[i]; }
  mean /= grid;
  const double sent = (double)TILE * sizeof(double2) * (C - 1) / C;  // bytes a CTA sends (and receives) per iteration
  printf("{\"cluster\": %d, \"mode\": \"%s\", \"ctas\": %d, \"iters\": %d, \"bytes_sent_per_cta_per_iter\": %.0f, "
         "\"cycles_per_iter_mean\": %.0f, \"cycles_per_iter_max\": %.0f, \"bytes_per_cycle_per_sm_out\": %.2f, "
         "\"GBps_per_sm_out\": %.1f, \"ms\": %.3f}\n",
         C, PULL ? "pull (ld.shared::cluster)" : "push (st.shared::cluster)", grid, iters, sent, mean / iters,
         (double)worst / iters, sent / (mean / iters), sent * iters / (ms * 1e-3) / 1e9, ms);
  delete[] h;
  cudaFree(sink); cudaFree(cyc);
  return 0;
}

int main(int argc, char** argv) {
  const bool pull = argc > 1 && !strcmp(argv[1], "--pull");
  for (int C = 2; C <= 16; C *= 2) {
    if (pull) run<true>(C, 2000); else run<false>(C, 2000);
  }
  return 0;
}
