// Micro-benchmark (test tooling): per-SM throughput of cp.async.bulk global->shared and shared->global
// as a function of the copy size and of how many warps issue.  Answers: is a bulk copy a fixed
// cost per operation or a cost per byte?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_probe tma_probe.cu && ./tma_probe
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void bulk_store(void* dst, unsigned src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}

// Each CTA: `warps` issuing warps; per round every issuing warp's lane 0 issues `per` copies of `bytes`
// (loads if mode 0, stores if mode 1, per/2 of each if mode 2); rounds are `depth` deep in flight.
__global__ void __launch_bounds__(512, 1) probe(unsigned char* g, size_t g_stride, int bytes, int per, int warps, int rounds, int depth, int mode, long long* cyc) {
  extern __shared__ __align__(128) unsigned char sm[];
  __shared__ unsigned long long bars[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned sb = smem_u32(sm), b0 = smem_u32(bars);
  if (threadIdx.x == 0) { for (int i = 0; i < 8; ++i) mbar_init(b0 + 8 * i, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();
  unsigned char* base = g + (size_t)blockIdx.x * g_stride;
  const int S = depth + 1;
  const int n_ops = per * warps;  // per round
  const long long t0 = clock64();
  unsigned phase = 0;
  // slot s of stage st: smem offset ((st * n_ops) + s) * bytes ; global offset ((r * n_ops) + s) * bytes (streaming)
  auto issue = [&](int r) {
    const int st = r % S;
    if (warp == 0 && lane == 0 && mode != 1) mbar_expect_tx(b0 + 8 * st, (unsigned)(mode == 2 ? (per / 2) * warps : n_ops) * bytes);
    __syncwarp();
    if (warp < warps && lane == 0) {
      for (int i = 0; i < per; ++i) {
        const int s = warp * per + i;
        const bool store = mode == 1 || (mode == 2 && (i & 1));
        unsigned char* gp = base + ((size_t)(r % 64) * n_ops + s) * bytes;
        const unsigned sp = sb + ((st * n_ops) + s) * bytes;
        if (store) bulk_store(gp, sp, bytes); else bulk_load(sp, gp, bytes, b0 + 8 * st);
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  };
  for (int r = 0; r < depth && r < rounds; ++r) issue(r);
  for (int r = 0; r < rounds; ++r) {
    const int st = r % S;
    if (r + depth < rounds) issue(r + depth);
    if (mode != 1) { mbar_wait(b0 + 8 * st, (phase >> st) & 1u); phase ^= 1u << st; }
    if (mode != 0 && warp < warps && lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
    __syncthreads();
  }
  if (warp < warps && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) cyc[blockIdx.x] = clock64() - t0;
}

int main() {
  int dev = 0; cudaSetDevice(dev);
  cudaDeviceProp prop; cudaGetDeviceProperties(&prop, dev);
  const int nsm = prop.multiProcessorCount;
  const size_t stride = 64ull * 64 * 16384;  // per CTA: 64 rounds x up to 64 ops x 16 KB
  unsigned char* g; cudaMalloc(&g, stride * nsm); cudaMemset(g, 1, stride * nsm);
  long long* cyc; cudaMalloc(&cyc, sizeof(long long) * nsm);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int rounds = 400;
  printf("mode bytes per warps depth | us  cycles/round  cycles/op  B/clk/SM  TB/s(all SMs)\n");
  for (int mode = 0; mode < 3; ++mode)
    for (int bytes : {512, 1024, 2048, 4096, 8192})
      for (int warps : {1, 2, 4, 8})
        for (int per : {1, 2, 4, 8})
          for (int depth : {1, 2, 3}) {
            const int n_ops = per * warps;
            if ((size_t)(depth + 1) * n_ops * bytes > 190 * 1024 || n_ops > 64 || n_ops < 4 || n_ops > 16) continue;
            if (mode == 2 && per < 2) continue;
            cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
            probe<<<nsm, 512, 200 * 1024>>>(g, stride, bytes, per, warps, 20, depth, mode, cyc);
            cudaEventRecord(e0);
            probe<<<nsm, 512, 200 * 1024>>>(g, stride, bytes, per, warps, rounds, depth, mode, cyc);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            long long c0; cudaMemcpy(&c0, cyc, sizeof c0, cudaMemcpyDeviceToHost);
            const double cpr = (double)c0 / rounds;
            printf("%d %5d %d %d %d | %8.1f %8.0f %7.0f %7.1f %6.2f\n", mode, bytes, per, warps, depth, ms * 1e3, cpr, cpr / n_ops,
                   (double)n_ops * bytes / cpr, (double)n_ops * bytes * rounds * nsm / (ms * 1e-3) / 1e12);
            if (cudaGetLastError() != cudaSuccess) { printf("error\n"); return 1; }
          }
  return 0;
}
