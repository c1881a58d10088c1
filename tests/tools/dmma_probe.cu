// fp64 tensor-core probe (sm_100a): throughput of mma.sync m8n8k4 against m16n8k8 / m16n8k16
// from registers only, and a check of the larger shapes' fragment layouts against a host product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_probe tests/tools/dmma_probe.cu && ./dmma_probe
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ void mma884(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c[0]), "+d"(c[1]) : "d"(a), "d"(b));
}
__device__ __forceinline__ void mma1688(double (&c)[4], const double (&a)[4], const double (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}
__device__ __forceinline__ void mma16816(double (&c)[4], const double (&a)[8], const double (&b)[4]) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
               : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
               : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]), "d"(b[0]), "d"(b[1]),
                 "d"(b[2]), "d"(b[3]));
}

// ---- throughput: every warp runs `iters` rounds of CH independent accumulation chains
template <int SHAPE, int CH>
__global__ void __launch_bounds__(256) rate_kernel(int iters, double* out) {
  const int lane = threadIdx.x & 31;
  double acc = 0.0;
  if (SHAPE == 0) {
    double c[CH][2];
    for (int i = 0; i < CH; ++i) c[i][0] = c[i][1] = 0.0;
    const double a = 1.0 + lane * 1e-3, b = 1.0 - lane * 1e-3;
    for (int it = 0; it < iters; ++it)
#pragma unroll
      for (int i = 0; i < CH; ++i) mma884(c[i], a, b);
    for (int i = 0; i < CH; ++i) acc += c[i][0] + c[i][1];
  } else if (SHAPE == 1) {
    double c[CH][4];
    for (int i = 0; i < CH; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.0;
    double a[4], b[2];
    for (int i = 0; i < 4; ++i) a[i] = 1.0 + (lane + i) * 1e-3;
    for (int i = 0; i < 2; ++i) b[i] = 1.0 - (lane + i) * 1e-3;
    for (int it = 0; it < iters; ++it)
#pragma unroll
      for (int i = 0; i < CH; ++i) mma1688(c[i], a, b);
    for (int i = 0; i < CH; ++i) acc += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  } else {
    double c[CH][4];
    for (int i = 0; i < CH; ++i) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0.0;
    double a[8], b[4];
    for (int i = 0; i < 8; ++i) a[i] = 1.0 + (lane + i) * 1e-3;
    for (int i = 0; i < 4; ++i) b[i] = 1.0 - (lane + i) * 1e-3;
    for (int it = 0; it < iters; ++it)
#pragma unroll
      for (int i = 0; i < CH; ++i) mma16816(c[i], a, b);
    for (int i = 0; i < CH; ++i) acc += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  }
  if (acc == 1.2345) out[0] = acc;
}

template <int SHAPE, int CH>
static void rate(const char* name, int n_sms, int ctas_per_sm, double fma_per_mma) {
  double* d;
  CK(cudaMalloc(&d, 8));
  const int iters = 4000;
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  rate_kernel<SHAPE, CH><<<n_sms * ctas_per_sm, 256>>>(100, d);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  rate_kernel<SHAPE, CH><<<n_sms * ctas_per_sm, 256>>>(iters, d);
  CK(cudaEventRecord(e1));
  CK(cudaDeviceSynchronize());
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double mmas = (double)n_sms * ctas_per_sm * 8 * CH * iters;
  printf("%-10s chains/warp %2d, %d x 8 warps/SM: %7.2f TFLOP/s  (%.2f ns per mma per SM)\n", name, CH, ctas_per_sm,
         2.0 * fma_per_mma * mmas / (ms * 1e-3) / 1e12, ms * 1e6 / (mmas / n_sms));
  CK(cudaFree(d));
}

// ---- layout check: C[16][8] = A[16][16] B[16][8] with the fragment maps this repo assumes
__global__ void layout_kernel(const double* A, const double* B, double* C16, double* C8) {
  const int lane = threadIdx.x, g = lane >> 2, t = lane & 3;
  double a[8], b[4], c[4] = {0, 0, 0, 0};
  for (int i = 0; i < 8; ++i) a[i] = A[(g + 8 * (i & 1)) * 16 + 4 * (i >> 1) + t];  // a_i: row g + 8 (i odd), column 4 (i / 2) + t
  for (int i = 0; i < 4; ++i) b[i] = B[(4 * i + t) * 8 + g];                        // b_i: row 4 i + t, column g
  mma16816(c, a, b);
  for (int i = 0; i < 4; ++i) C16[(g + 8 * (i >> 1)) * 8 + 2 * t + (i & 1)] = c[i];   // c_i: row g + 8 (i >= 2), column 2 t + (i & 1)
  // m16n8k8 twice (k = 0..7, 8..15)
  double c2[4] = {0, 0, 0, 0};
  for (int h = 0; h < 2; ++h) {
    double a4[4], b2[2];
    for (int i = 0; i < 4; ++i) a4[i] = A[(g + 8 * (i & 1)) * 16 + 8 * h + 4 * (i >> 1) + t];
    for (int i = 0; i < 2; ++i) b2[i] = B[(8 * h + 4 * i + t) * 8 + g];
    mma1688(c2, a4, b2);
  }
  for (int i = 0; i < 4; ++i) C8[(g + 8 * (i >> 1)) * 8 + 2 * t + (i & 1)] = c2[i];
}

int main() {
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  printf("%s, %d SMs, %d MHz\n", prop.name, prop.multiProcessorCount, prop.clockRate / 1000);
  std::vector<double> A(256), B(128), C(128, 0.0), C16(128), C8(128);
  srand(1);
  for (auto& x : A) x = rand() / (double)RAND_MAX - 0.5;
  for (auto& x : B) x = rand() / (double)RAND_MAX - 0.5;
  for (int i = 0; i < 16; ++i)
    for (int j = 0; j < 8; ++j)
      for (int k = 0; k < 16; ++k) C[i * 8 + j] += A[i * 16 + k] * B[k * 8 + j];
  double *dA, *dB, *dC, *dC8;
  CK(cudaMalloc(&dA, 2048)); CK(cudaMalloc(&dB, 1024)); CK(cudaMalloc(&dC, 1024)); CK(cudaMalloc(&dC8, 1024));
  CK(cudaMemcpy(dA, A.data(), 2048, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, B.data(), 1024, cudaMemcpyHostToDevice));
  layout_kernel<<<1, 32>>>(dA, dB, dC, dC8);
  CK(cudaMemcpy(C16.data(), dC, 1024, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(C8.data(), dC8, 1024, cudaMemcpyDeviceToHost));
  double e16 = 0, e8 = 0;
  for (int i = 0; i < 128; ++i) { e16 = fmax(e16, fabs(C16[i] - C[i])); e8 = fmax(e8, fabs(C8[i] - C[i])); }
  printf("layout check: m16n8k16 max |diff| %.3g, m16n8k8 (two steps) %.3g\n", e16, e8);
  const int n = prop.multiProcessorCount;
  rate<0, 8>("m8n8k4", n, 1, 256);
  rate<0, 8>("m8n8k4", n, 2, 256);
  rate<0, 16>("m8n8k4", n, 2, 256);
  rate<1, 4>("m16n8k8", n, 1, 1024);
  rate<1, 4>("m16n8k8", n, 2, 1024);
  rate<1, 8>("m16n8k8", n, 2, 1024);
  rate<2, 2>("m16n8k16", n, 1, 2048);
  rate<2, 4>("m16n8k16", n, 2, 2048);
  rate<2, 8>("m16n8k16", n, 2, 2048);
  return 0;
}
