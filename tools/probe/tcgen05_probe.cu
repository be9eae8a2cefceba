// Development probe: one tcgen05.mma kind::tf32 (M = 128, N = 16, K = 8) on known operands, every layout the tf32 arm
// uses.  Prints the mismatches against a host product.   nvcc -arch=sm_100a -o tcgen05_probe tcgen05_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo, uint32_t sbo, uint64_t layout_type = 0)
{
  return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46) |
         (layout_type << 61);
}
__host__ __device__ constexpr uint32_t instr_desc(int M, int N, int a_mn, int b_mn)
{
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// mode bit 0: A MN-major (else K-major); bit 1: B K-major (else MN-major)
__global__ void __launch_bounds__(128) probe(const float* A, const float* B, float* D, int mode, uint32_t pitch)
{
  __shared__ __align__(1024) unsigned char opA[16384];
  __shared__ __align__(128) float opB[256];
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base;
  const int tid = threadIdx.x;
  const bool a_mn = mode & 1, b_k = mode & 2, sw = mode & 4;
  // swizzled MN-major (SWIZZLE_128B_BASE32B): atoms of 4 k-rows x 128 bytes (32 m); 32-byte chunk index ^= row
  const uint32_t LBO = pitch == 128u ? 1024u : 512u, SBO = pitch == 128u ? 512u : 2048u;
  // A[m][k], m = tid
  for (int k = 0; k < 8; ++k)
  {
    uint32_t off;
    if (sw)
      off = (tid >> 5) * LBO + (k >> 2) * SBO + (k & 3) * 128 + ((((tid & 31) >> 3) ^ (k & 3)) * 32) + (tid & 7) * 4;
    else if (!a_mn)
      off = (tid >> 3) * 256 + (k >> 2) * 128 + (tid & 7) * 16 + (k & 3) * 4;
    else
      off = (tid >> 2) * pitch + k * 16 + (tid & 3) * 4;
    *reinterpret_cast<float*>(opA + off) = A[tid * 8 + k];
  }
  // B[k][n], n < 16
  {
    const int k = tid >> 4, n = tid & 15;
    int at;
    if (!b_k)
      at = (n >> 2) * 32 + k * 4 + (n & 3);           // MN-major: n-chunk stride 128 B, k row 16 B
    else
      at = (n >> 3) * 64 + (k >> 2) * 32 + (n & 7) * 4 + (k & 3); // K-major: 8 n-rows x 16 B core matrices, k-chunk stride 128 B, n-group stride 256 B
    opB[at] = B[k * 16 + n];
  }
  if (tid == 0)
  {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (tid < 32)
  {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&tmem_base)), "r"(32u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base;
  if (tid == 0)
  {
    const uint32_t idesc = instr_desc(128, 16, a_mn, !b_k);
    const uint64_t da = sw ? smem_desc(s32(opA), LBO, SBO, 1) : a_mn ? smem_desc(s32(opA), 128, pitch) : smem_desc(s32(opA), 128, 256);
    const uint64_t db = b_k ? smem_desc(s32(opB), 128, 256) : smem_desc(s32(opB), 128, 128);
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}" ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(0u) : "memory");
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(&bar)) : "memory");
    printf("mode %d tmem base %08x idesc %08x da %016llx db %016llx\n", mode, tmem, idesc, (unsigned long long)da, (unsigned long long)db);
  }
  asm volatile("{\n.reg .pred P1;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n@P1 bra E;\nbra W;\nE:\n}" ::"r"(s32(&bar)), "r"(0u) : "memory");
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t r[16];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                 "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(tmem + ((uint32_t)(tid & ~31) << 16))
               : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  for (int n = 0; n < 16; ++n)
    D[tid * 16 + n] = __uint_as_float(r[n]);
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (tid < 32)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(32u) : "memory");
}

int main()
{
  float hA[128 * 8], hB[8 * 16], hD[128 * 16], *dA, *dB, *dD;
  for (int i = 0; i < 128 * 8; ++i) hA[i] = (float)((i * 7) % 13) - 3;   // small integers: exact in tf32
  for (int i = 0; i < 8 * 16; ++i) hB[i] = (float)((i * 5) % 11) - 2;
  cudaMalloc(&dA, sizeof hA), cudaMalloc(&dB, sizeof hB), cudaMalloc(&dD, sizeof hD);
  cudaMemcpy(dA, hA, sizeof hA, cudaMemcpyHostToDevice), cudaMemcpy(dB, hB, sizeof hB, cudaMemcpyHostToDevice);
  for (int mode : {2, 7})
    for (uint32_t pitch : {128u, 144u})
    {
      if (!(mode & 1) && pitch != 128u) continue;
      cudaMemset(dD, 0xff, sizeof hD);
      probe<<<1, 128>>>(dA, dB, dD, mode, pitch);
      cudaError_t e = cudaDeviceSynchronize();
      cudaMemcpy(hD, dD, sizeof hD, cudaMemcpyDeviceToHost);
      int bad = 0, zeros = 0;
      for (int m = 0; m < 128; ++m)
        for (int n = 0; n < 16; ++n)
        {
          float ref = 0;
          for (int k = 0; k < 8; ++k) ref += hA[m * 8 + k] * hB[k * 16 + n];
          if (hD[m * 16 + n] != ref) ++bad;
          if (hD[m * 16 + n] == 0) ++zeros;
        }
      printf("mode %d (A %s, B %s) pitch %u: %s, %d of 2048 wrong, %d zeros; D[0][0..3] = %g %g %g %g, D[37][5] = %g\n", mode, mode & 1 ? "MN" : "K",
             mode & 2 ? "K" : "MN", pitch, cudaGetErrorString(e), bad, zeros, hD[0], hD[1], hD[2], hD[3], hD[37 * 16 + 5]);
    }
  return 0;
}
