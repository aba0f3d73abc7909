// mma_bench.cu — measures tcgen05.mma throughput per SM for several shapes,
// with operands resident in smem (no TMA) and optionally with a TMA-like
// bulk-copy stream running into other smem at the same time.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I csrc -o /tmp/mma_bench tools/mma_bench.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../csrc/sm100_ptx.cuh"
using namespace sm100;

struct Res { long long cycles; };

template <int CG>
__global__ void __launch_bounds__(128, 1)
bench(int n, int iters, int k_per_iter, long long *out, const uint8_t *gsrc, int do_copy) {
  extern __shared__ uint8_t raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar, cbar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // fill operand area with small bf16 values
  for (int i = threadIdx.x; i < (64 * 1024) / 4; i += blockDim.x) ((uint32_t *)smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_init(&cbar, 1); fence_mbar_init(); }
  fence_proxy_async();
  if (CG == 2) cluster_sync_all();
  if (warp == 0) { if (CG == 1) tmem_alloc<512>(&slot); else tmem_alloc_2sm<512>(&slot); }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();
  tc_fence_after();
  const uint32_t tm = slot;
  const bool leader = (CG == 1) || cluster_ctarank() == 0;
  long long t0 = 0, t1 = 0;
  if (warp == 1 && lane == 0 && leader) {
    const uint32_t sa = smem_u32(smem);
    const uint32_t sb = sa + 16384;
    const uint32_t idesc = umma_idesc_bf16(CG == 1 ? 128 : 256, n);
    t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      for (int k = 0; k < k_per_iter; ++k) {
        uint64_t da = umma_desc_sw128(sa) + 2 * (k & 3);
        uint64_t db = umma_desc_sw128(sb) + 2 * (k & 3);
        if (CG == 1) umma_bf16(tm, da, db, idesc, 1u);
        else umma_bf16_2sm(tm, da, db, idesc, 1u);
      }
    }
    if (CG == 1) umma_commit(&bar); else umma_commit_2sm(&bar, 1);
    mbar_wait(&bar, 0);
    t1 = clock64();
    out[blockIdx.x] = (t1 - t0);
  }
  if (warp == 2 && lane == 0 && do_copy) {
    // stream bulk copies global -> smem while the MMAs run: 4 in flight, each
    // into its own 32 KB slot that overlaps the operand area like a real ring
    __shared__ uint64_t cb[4];
    for (int i = 0; i < 4; ++i) mbar_init(&cb[i], 1);
    fence_mbar_init();
    const int total = iters * k_per_iter / 4;
    long long c0 = clock64();
    for (int it = 0; it < total + 4; ++it) {
      const int sl = it & 3;
      if (it >= 4) mbar_wait(&cb[sl], ((it >> 2) - 1) & 1);
      if (it < total) {
        mbar_expect_tx(&cb[sl], do_copy);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     :: "r"(smem_u32(smem + sl * 40960)), "l"(gsrc + (size_t)((blockIdx.x * 7 + it) % 64) * 65536), "r"(do_copy), "r"(smem_u32(&cb[sl])) : "memory");
      }
    }
    out[148 + blockIdx.x] = clock64() - c0;
  }
  tc_fence_before();
  __syncthreads();
  if (CG == 2) cluster_sync_all();
  if (warp == 0) { tc_fence_after(); if (CG == 1) tmem_dealloc<512>(tm); else tmem_dealloc_2sm<512>(tm); }
}

template <int CG>
void run(int n, int do_copy, const uint8_t *gsrc, long long *dout) {
  const int iters = 200, kpi = 16;
  const int smem = 4 * 40960 + 2048;
  cudaFuncSetAttribute(bench<CG>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(148); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = CG; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  cudaMemset(dout, 0, 296 * 8);
  cudaError_t e = cudaLaunchKernelEx(&cfg, bench<CG>, n, iters, kpi, dout, gsrc, do_copy);
  e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("CG=%d N=%d: %s\n", CG, n, cudaGetErrorString(e)); exit(1); }
  std::vector<long long> h(296);
  cudaMemcpy(h.data(), dout, 296 * 8, cudaMemcpyDeviceToHost);
  double sum = 0; int cnt = 0;
  for (int i = 0; i < 148; ++i) if (h[i] > 0) { sum += h[i]; cnt++; }
  double csum = 0; int ccnt = 0;
  for (int i = 148; i < 296; ++i) if (h[i] > 0) { csum += h[i]; ccnt++; }
  if (ccnt) printf("   copy stream: %.1f B/clk/SM  ", (double)do_copy * (iters * kpi / 4) / (csum / ccnt));
  double per = sum / cnt / (iters * kpi);
  double floor_c = (CG == 1 ? 128.0 : 256.0) * n / (256.0 * CG);
  printf("cta_group=%d M=%d N=%3d copy=%5d B/4mma : %.1f cyc/MMA (floor %.0f, %.0f%%)  [%d issuers]\n", CG, CG * 128, n, do_copy, per, floor_c, 100 * floor_c / per, cnt);
}

int main() {
  uint8_t *gsrc; long long *dout;
  cudaMalloc(&gsrc, 64 * 65536); cudaMemset(gsrc, 0, 64 * 65536);
  cudaMalloc(&dout, 296 * 8);
  for (int copy : {0, 28672, 40960}) {
    for (int n : {128, 192, 256}) run<1>(n, copy, gsrc, dout);
    for (int n : {128, 192, 256}) run<2>(n, copy, gsrc, dout);
  }
  return 0;
}
