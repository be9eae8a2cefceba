// Throughput probe (development tool): scalar FFMA vs packed fma.rn.f32x2 (FFMA2) on sm_100a, with the
// multiplier taken from a register or from the kernel-parameter constant bank.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probe/ffma2_probe tools/probe/ffma2_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

struct Tab
{
  float m[64];
};

template <int MODE>
__global__ void __launch_bounds__(256) k(const __grid_constant__ Tab t, float* out, int iters)
{
  const int lane = threadIdx.x & 31;
  float x0 = 1.0f + lane * 1e-6f, x1 = 1.0f - lane * 1e-6f;
  float c[32];
#pragma unroll
  for (int i = 0; i < 32; ++i)
    c[i] = i;
  for (int it = 0; it < iters; ++it)
  {
    if constexpr (MODE == 0)
    { // scalar FFMA, constant-bank multiplier: 32 chains
#pragma unroll
      for (int i = 0; i < 32; ++i)
        c[i] = fmaf(t.m[i], (i & 1) ? x1 : x0, c[i]);
    }
    else if constexpr (MODE == 1)
    { // FFMA2: {c[2i], c[2i+1]} += {m, m} * {x0, x1}, multiplier pair built from the constant bank
#pragma unroll
      for (int i = 0; i < 16; ++i)
      {
        unsigned long long cc, mm, xx;
        asm("mov.b64 %0, {%1, %2};" : "=l"(cc) : "f"(c[2 * i]), "f"(c[2 * i + 1]));
        asm("mov.b64 %0, {%1, %1};" : "=l"(mm) : "f"(t.m[i]));
        asm("mov.b64 %0, {%1, %2};" : "=l"(xx) : "f"(x0), "f"(x1));
        asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(cc) : "l"(mm), "l"(xx));
        asm("mov.b64 {%0, %1}, %2;" : "=f"(c[2 * i]), "=f"(c[2 * i + 1]) : "l"(cc));
      }
    }
    else
    { // FFMA2: {c[2i], c[2i+1]} += {m[2i], m[2i+1]} * {x0, x0}: table pairs as they lie in the constant bank
#pragma unroll
      for (int i = 0; i < 16; ++i)
      {
        unsigned long long cc, mm, xx;
        asm("mov.b64 %0, {%1, %2};" : "=l"(cc) : "f"(c[2 * i]), "f"(c[2 * i + 1]));
        asm("mov.b64 %0, {%1, %2};" : "=l"(mm) : "f"(t.m[2 * i]), "f"(t.m[2 * i + 1]));
        asm("mov.b64 %0, {%1, %1};" : "=l"(xx) : "f"(x0));
        asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(cc) : "l"(mm), "l"(xx));
        asm("mov.b64 {%0, %1}, %2;" : "=f"(c[2 * i]), "=f"(c[2 * i + 1]) : "l"(cc));
      }
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 32; ++i)
    s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char* name)
{
  Tab t;
  for (int i = 0; i < 64; ++i)
    t.m[i] = 1.0f + 1e-7f * i;
  float* out;
  const int grid = 148 * 8, iters = 20000;
  cudaMalloc(&out, grid * 256 * sizeof(float));
  k<MODE><<<grid, 256>>>(t, out, 100);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<MODE><<<grid, 256>>>(t, out, iters);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  const double flops = 2.0 * 32 * (double)iters * grid * 256;
  printf("%-44s %8.3f ms  %7.2f TFLOP/s\n", name, ms, flops / ms / 1e9);
  cudaFree(out);
}

int main()
{
  run<0>("FFMA  scalar, constant-bank multiplier");
  run<1>("FFMA2 {m,m} x {x0,x1}");
  run<2>("FFMA2 {m[2i],m[2i+1]} x {x0,x0}");
  return 0;
}
