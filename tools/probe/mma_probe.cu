// Throughput probe (development tool): FP64 FMA vs DMMA m8n8k4 / m16n8k{4,8,16} vs TF32 m16n8k8 vs FFMA on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/probe/mma_probe tools/probe/mma_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256) k(double* out, int iters)
{
  const int lane = threadIdx.x & 31;
  double a = 1.0 + lane * 1e-9, b = 1.0 - lane * 1e-9;
  double c[16];
#pragma unroll
  for (int i = 0; i < 16; ++i)
    c[i] = i;
  float fa = (float)a, fb = (float)b;
  float fc[32];
#pragma unroll
  for (int i = 0; i < 32; ++i)
    fc[i] = i;
  for (int it = 0; it < iters; ++it)
  {
    if constexpr (MODE == 0)
    { // DFMA: 16 independent chains
#pragma unroll
      for (int i = 0; i < 16; ++i)
        c[i] = fma(a, b, c[i]);
    }
    else if constexpr (MODE == 1)
    { // DMMA m8n8k4: 8 independent accumulators (2 regs each)
#pragma unroll
      for (int i = 0; i < 8; ++i)
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                     : "+d"(c[2 * i]), "+d"(c[2 * i + 1]) : "d"(a), "d"(b));
    }
    else if constexpr (MODE == 2)
    { // DMMA m16n8k4: 4 accumulators (4 regs each)
#pragma unroll
      for (int i = 0; i < 4; ++i)
        asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                     : "+d"(c[4 * i]), "+d"(c[4 * i + 1]), "+d"(c[4 * i + 2]), "+d"(c[4 * i + 3]) : "d"(a), "d"(b), "d"(b));
    }
    else if constexpr (MODE == 3)
    { // DMMA m16n8k8
#pragma unroll
      for (int i = 0; i < 4; ++i)
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+d"(c[4 * i]), "+d"(c[4 * i + 1]), "+d"(c[4 * i + 2]), "+d"(c[4 * i + 3])
                     : "d"(a), "d"(b), "d"(a), "d"(b), "d"(a), "d"(b));
    }
    else if constexpr (MODE == 4)
    { // TF32 m16n8k8: 8 accumulators (4 regs each)
      unsigned ua = __float_as_uint(fa), ub = __float_as_uint(fb);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(fc[4 * i]), "+f"(fc[4 * i + 1]), "+f"(fc[4 * i + 2]), "+f"(fc[4 * i + 3])
                     : "r"(ua), "r"(ub), "r"(ua), "r"(ub), "r"(ua), "r"(ub));
    }
    else if constexpr (MODE == 5)
    { // FFMA: 32 chains
#pragma unroll
      for (int i = 0; i < 32; ++i)
        fc[i] = fmaf(fa, fb, fc[i]);
    }
    else if constexpr (MODE == 6)
    { // BF16 m16n8k16 (for scale)
      unsigned ua = 0x3f803f80u;
#pragma unroll
      for (int i = 0; i < 8; ++i)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(fc[4 * i]), "+f"(fc[4 * i + 1]), "+f"(fc[4 * i + 2]), "+f"(fc[4 * i + 3])
                     : "r"(ua), "r"(ua), "r"(ua), "r"(ua), "r"(ua), "r"(ua));
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i)
    s += c[i];
#pragma unroll
  for (int i = 0; i < 32; ++i)
    s += fc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char* name, double macs_per_warp_iter, int wpb)
{
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  double* out;
  const int blocks = sms * (2048 / (wpb * 32));
  cudaMalloc(&out, sizeof(double) * blocks * wpb * 32);
  const int iters = 20000;
  k<MODE><<<blocks, wpb * 32>>>(out, 100);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<MODE><<<blocks, wpb * 32>>>(out, iters);
  cudaEventRecord(e1);
  cudaDeviceSynchronize();
  cudaError_t err = cudaGetLastError();
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  const double warps = (double)blocks * wpb;
  const double macs = warps * iters * macs_per_warp_iter;
  printf("%-22s warps/SM=%2d  %8.3f ms  %8.2f TMAC/s = %7.2f TFLOP/s  (%.1f MAC/clk/SM at 1.9 GHz)  %s\n", name,
         (int)(warps / sms), ms, macs / ms / 1e9, 2 * macs / ms / 1e9, macs / (ms * 1e-3) / sms / 1.9e9,
         err == cudaSuccess ? "" : cudaGetErrorString(err));
  cudaFree(out);
}

int main()
{
  for (int wpb : {8, 2})
  {
    run<0>("DFMA", 16.0 * 32, wpb);
    run<1>("DMMA m8n8k4", 8.0 * 8 * 8 * 4, wpb);
    run<2>("DMMA m16n8k4", 4.0 * 16 * 8 * 4, wpb);
    run<3>("DMMA m16n8k8", 4.0 * 16 * 8 * 8, wpb);
    run<4>("TF32 m16n8k8", 8.0 * 16 * 8 * 8, wpb);
    run<5>("FFMA", 32.0 * 32, wpb);
    run<6>("BF16 m16n8k16", 8.0 * 16 * 8 * 16, wpb);
  }
  return 0;
}
