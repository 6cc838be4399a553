// Micro-probe (not part of the library): per-SM throughput of MUFU.LG2 / MUFU.RCP / MUFU.EX2 with 1..16 warps per SM.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/mufu_probe.cu -o tools/_bin/mufu_probe && tools/_bin/mufu_probe
#include <cstdio>
#include <cuda_runtime.h>
template <int OP>
__global__ void k(float* out, int iters, long long* cyc) {
  float x[8];
  for (int i = 0; i < 8; ++i) x[i] = 1.5f + threadIdx.x * 0.001f + i;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (OP == 0) asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
      if (OP == 1) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
      if (OP == 2) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
      if (OP == 3) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(x[i]));
    }
  }
  const long long t1 = clock64();
  float s = 0;
  for (int i = 0; i < 8; ++i) s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
// 8 MUFU + 8*R independent FFMA per iteration: do the two pipes overlap inside one SM sub-partition?
template <int R>
__global__ void kmix(float* out, int iters, long long* cyc) {
  float x[8], y[8 * R];
  for (int i = 0; i < 8; ++i) x[i] = 1.5f + threadIdx.x * 0.001f + i;
  for (int i = 0; i < 8 * R; ++i) y[i] = 0.5f + threadIdx.x * 0.002f + i;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
#pragma unroll
      for (int r = 0; r < R; ++r) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(y[i * R + r]));
    }
  }
  const long long t1 = clock64();
  float s = 0;
  for (int i = 0; i < 8; ++i) s += x[i];
  for (int i = 0; i < 8 * R; ++i) s += y[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int R>
void runmix(float* out, long long* cyc) {
  for (int warps : {4, 8, 16}) {
    const int iters = 2000;
    for (int rep = 0; rep < 2; ++rep) kmix<R><<<148, warps * 32>>>(out, iters, cyc);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("mix 1 lg2 : %d ffma, warps/SM=%2d: %.1f cycles per (lg2 + %d ffma) per warp-slot on an SMSP (mufu alone 8.0, ffma alone %d)\n",
           R, warps, (double)c / (iters * 8.0 * (warps / 4.0)), R, R);
  }
}
int main() {
  { float* o; long long* c; cudaMalloc(&o, 1 << 22); cudaMalloc(&c, 8); runmix<2>(o, c); runmix<4>(o, c); runmix<6>(o, c); }
  float* out; long long* cyc; cudaMalloc(&out, 1 << 22); cudaMalloc(&cyc, 8);
  const char* names[4] = {"lg2", "rcp", "ex2", "ffma"};
  for (int op = 0; op < 4; ++op)
    for (int warps : {1, 4, 8, 16, 32}) {
      const int iters = 2000;
      for (int rep = 0; rep < 2; ++rep) {
        if (op == 0) k<0><<<148, warps * 32>>>(out, iters, cyc);
        if (op == 1) k<1><<<148, warps * 32>>>(out, iters, cyc);
        if (op == 2) k<2><<<148, warps * 32>>>(out, iters, cyc);
        if (op == 3) k<3><<<148, warps * 32>>>(out, iters, cyc);
      }
      long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
      const double ops = (double)iters * 8 * warps * 32;
      printf("%s warps/SM=%2d: %lld cycles, %.2f lane-ops/clk/SM, %.2f cycles per warp-instr per SMSP\n", names[op], warps, c,
             ops / c, (double)c / (iters * 8.0 * (warps > 4 ? warps / 4.0 : 1.0)));
    }
  return 0;
}
