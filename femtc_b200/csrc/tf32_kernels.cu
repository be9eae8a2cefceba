// fp32 tensor-core arm of BASELINE config 4 ("cells grouped into batched tensor-core contractions"): the mass action
// at n1 = 8 (degree 7) with every axis contraction issued as tcgen05.mma kind::tf32, accumulator in tensor memory.
//
// The contraction replaced is linalg::gemm_blocked<T, k = n1, m = n1, n = n1^2> (reference src/mass.hpp:36-47,
// src/linalg.hpp:101-182): Y[i, col] = sum_p M(i,p) X[p, col] over the n1^2 columns of a cell.  Here the CELL DATA is
// the M side of the MMA and the 1-D table the N side:
//     D[128 tensor columns (2 cells) x 16] = A[128 x 8] . B[8 x 16]        (one UTCMMA, M = 128, N = 16, K = 8)
// One TF32 product carries 2^-11 relative error and the fp32 contract is 1e-5, so each operand is split
// x = x_hi + x_lo (x_hi = cvt.rna.tf32, x_lo = x - x_hi exactly) and three products are kept:
//     D[:, 0:8]  = X_hi . M_hi + X_lo . M_hi          D[:, 8:16] = X_hi . M_lo
// i.e. two MMAs (X_hi against [M_hi | M_lo], X_lo against [M_hi | 0], accumulating); the reader adds the two halves.
//
// (MN-major tf32 operands exist only in the SWIZZLE_128B_BASE32B layout; B is kept K-major.)
// Data path of one contraction: the 128 threads of a block each own one tensor column (= one TMEM lane): they read their
// 16 accumulator columns with tcgen05.ld, add the halves, split into hi / lo and store both into the shared-memory
// operand buffers in the canonical no-swizzle UMMA layout of the NEXT contraction (core matrices of 8 x 16 bytes):
//   * the axis to contract lies along the thread's registers (first forward and first backward contraction): K-major A,
//     row = lane, two 16-byte k-chunks per row;
//   * otherwise the thread's 8 registers run along the new output index, which becomes part of the next M index:
//     MN-major A, the thread writes 8 consecutive m at its own k -- the exchange between lanes that the FMA kernels do
//     through their padded scratch IS the operand store here.
// One thread issues the MMAs and commits them to an mbarrier; U / detJ tiles arrive by 1-D TMA bulk copies into a
// two-stage ring, W leaves by a bulk store.  SASS: UTCMMA (tcgen05.mma), LDTM (tcgen05.ld), UBLKCP, SYNCS.
//
// Why it is NOT the default (measured, profiles/r02_tf32_arm.txt): every contraction makes the tensor core read the
// hi and the lo operand from shared memory (8 KB per 128 columns) on top of the 8 KB the threads store, twice the
// shared-memory traffic of the FFMA2 column kernel, whose limiter is already shared-memory bandwidth.
#include "common.cuh"
#include "tma.cuh"

#include <algorithm>
#include <map>
#include <mutex>

namespace femtc
{
long long opt_tc(); // api.cu

namespace
{

constexpr int TN1 = 8, TND = 512;
constexpr int TILE_CELLS = 2;                 // 2 cells x 64 tensor columns = the 128 rows of one MMA
constexpr uint32_t TILE_FLOATS = TILE_CELLS * TND;
constexpr uint32_t TILE_BYTES = TILE_FLOATS * 4;
constexpr uint32_t OP_BYTES = 4096;           // 128 rows x 8 k x 4 bytes, either majorness
// MN-major tf32 operands exist in one shared-memory layout only, SWIZZLE_128B_BASE32B (cute: Layout_MN_SW128_32B_Atom;
// the no-swizzle form silently produces zeros, tools/probe/tcgen05_probe.cu): atoms of 4 k-rows x 128 bytes (32
// consecutive m), the 32-byte chunk index of a row XORed with the row index; atoms along m MN_LBO apart, along k MN_SBO.
constexpr uint32_t MN_LBO = 1024, MN_SBO = 512;
constexpr uint32_t TMEM_COLS = 32;            // smallest allocation; 16 used

struct Tf32Tab
{
  float a[TN1 * TN1];
};

struct Tf32Smem
{
  alignas(128) float inU[2][TILE_FLOATS];
  alignas(128) float inJ[2][TILE_FLOATS];
  alignas(1024) unsigned char opHi[OP_BYTES];
  alignas(1024) unsigned char opLo[OP_BYTES];
  alignas(128) float outW[TILE_FLOATS];
  alignas(128) float tabHL[128]; // B = [M_hi | M_lo], K-major core matrices: 2 n-groups x 2 k-chunks x (8 n x 4 k)
  alignas(128) float tabH0[128]; // B = [M_hi | 0]
  alignas(8) uint64_t full[2];
  alignas(8) uint64_t mma_done;
  uint32_t tmem_base;
};

// round to the nearest tf32 (10 explicit mantissa bits; ties away from zero, like cvt.rna.tf32.f32, which ptxas expands
// into six instructions -- this form is two)
__device__ __forceinline__ float tf32_rna(float x)
{
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xFFFFE000u);
}

// ---- tcgen05 wrappers -------------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t cols)
{
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t cols)
{
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(addr), "r"(cols) : "memory");
}
// shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start address, leading / stride byte offsets in 16-byte
// units, version 1 (sm_100), layout type 0 = no swizzle, 1 = SWIZZLE_128B_BASE32B
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint64_t layout_type = 0)
{
  return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | (1ull << 46) | (layout_type << 61);
}
// instruction descriptor (cute::UMMA::InstrDescriptor): D = f32, A = B = tf32, dense, M x N, majorness of A and B
constexpr uint32_t instr_desc(int M, int N, int a_mn_major, int b_mn_major)
{
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate)
{
  asm volatile("{\n"
               ".reg .pred p;\n"
               "setp.ne.b32 p, %4, 0;\n"
               "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
               "}" ::"r"(d_tmem),
               "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
               : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar)
{
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16])
{
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 "
               "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                 "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(taddr)
               : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void sts128(uint32_t addr, float a, float b, float c, float d)
{
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

enum StageKind
{
  K_MAJOR = 0,  // rows = this lane, k = register index
  MN_ROWS_A = 1, // lane (c, a, b): k = b, the 8 registers become m = (c, a, 0..7)
  MN_ROWS_B = 2  // lane (c, a, b): k = a, the 8 registers become m = (c, b, 0..7)
};

// One axis contraction for the block's 128 tensor columns: x (this lane's 8 values) -> split -> operand buffers ->
// two MMAs -> this lane's 8 results, for the lane the MMA's row mapping gives it (see the stage table in the kernel).
template <int KIND>
__device__ __forceinline__ void contract(Tf32Smem& s, float (&x)[8], uint32_t tmem_lane_base, uint32_t& mma_parity,
                                         bool swapped = false)
{
  const int tid = threadIdx.x;
  float hi[8], lo[8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
  {
    hi[i] = tf32_rna(x[i]);
    lo[i] = x[i] - hi[i]; // exact; the tensor core drops its low 13 bits (<= 2^-21 |x|)
  }
  const uint32_t ph = smem_u32(s.opHi), pl = smem_u32(s.opLo);
  if (KIND == K_MAJOR)
  {
    // row = lane: two 16-byte k-chunks 128 bytes apart; 8 consecutive lanes fill one 128-byte core matrix.
    // `swapped` (first contraction only): x[0..3] hold k = 4..7 and x[4..7] hold k = 0..3 (see the U load).
    const uint32_t o = (uint32_t)(tid >> 3) * 256u + (uint32_t)(tid & 7) * 16u;
    const uint32_t o0 = swapped ? o + 128u : o, o1 = swapped ? o : o + 128u;
    sts128(ph + o0, hi[0], hi[1], hi[2], hi[3]);
    sts128(ph + o1, hi[4], hi[5], hi[6], hi[7]);
    sts128(pl + o0, lo[0], lo[1], lo[2], lo[3]);
    sts128(pl + o1, lo[4], lo[5], lo[6], lo[7]);
  }
  else
  {
    const uint32_t c = tid >> 6, a = (tid >> 3) & 7, b = tid & 7;
    const uint32_t q = KIND == MN_ROWS_A ? a : b, k = KIND == MN_ROWS_A ? b : a;
    // m = c 64 + q 8 + (0..7): atom (m / 32, k / 4), row k % 4, 32-byte chunk ((m % 32) / 8) ^ (k % 4).
    // Lanes b and b + 4 of a quarter-warp differ by a whole atom (same banks): the upper four store their second
    // 16 bytes first, so that the eight lanes of one store instruction cover eight different 16-byte bank groups.
    const uint32_t o = (c * 2u + (q >> 2)) * MN_LBO + (k >> 2) * MN_SBO + (k & 3u) * 128u + (((q & 3u) ^ (k & 3u)) * 32u);
    const bool up = (b & 4u) != 0;
    const uint32_t o0 = up ? o + 16u : o, o1 = up ? o : o + 16u;
    sts128(ph + o0, up ? hi[4] : hi[0], up ? hi[5] : hi[1], up ? hi[6] : hi[2], up ? hi[7] : hi[3]);
    sts128(ph + o1, up ? hi[0] : hi[4], up ? hi[1] : hi[5], up ? hi[2] : hi[6], up ? hi[3] : hi[7]);
    sts128(pl + o0, up ? lo[4] : lo[0], up ? lo[5] : lo[1], up ? lo[6] : lo[2], up ? lo[7] : lo[3]);
    sts128(pl + o1, up ? lo[0] : lo[4], up ? lo[1] : lo[5], up ? lo[2] : lo[6], up ? lo[3] : lo[7]);
  }
  fence_proxy_async(); // generic-proxy stores -> tensor-core (async proxy) reads
  tc_fence_before();   // this lane's tcgen05.ld of the previous stage is ordered before the MMAs that overwrite D
  __syncthreads();
  if (tid == 0)
  {
    tc_fence_after();
    constexpr uint32_t idesc = instr_desc(128, 16, KIND != K_MAJOR, 0);
    const uint64_t bHL = smem_desc(smem_u32(s.tabHL), 128, 256), bH0 = smem_desc(smem_u32(s.tabH0), 128, 256);
    const uint64_t aHi = KIND == K_MAJOR ? smem_desc(ph, 128, 256) : smem_desc(ph, MN_LBO, MN_SBO, 1);
    const uint64_t aLo = KIND == K_MAJOR ? smem_desc(pl, 128, 256) : smem_desc(pl, MN_LBO, MN_SBO, 1);
    umma_tf32(tmem_lane_base, aHi, bHL, idesc, 0);
    umma_tf32(tmem_lane_base, aLo, bH0, idesc, 1);
    umma_commit(smem_u32(&s.mma_done));
  }
  mbar_wait(smem_u32(&s.mma_done), mma_parity);
  mma_parity ^= 1;
  tc_fence_after();
  uint32_t r[16];
  tmem_ld16(tmem_lane_base + ((uint32_t)(tid & ~31) << 16), r);
#pragma unroll
  for (int i = 0; i < 8; ++i)
    x[i] = __uint_as_float(r[i]) + __uint_as_float(r[8 + i]);
}

// Stage table for lane l = (c, a, b) = (l / 64, (l / 8) % 8, l % 8); tensor U[a0][a1][a2], quadrature index i, output j:
//   1  K-major      lane (c, a0, a1), registers a2   -> registers i2
//   2  MN rows A    k = a1 (= b), m = (c, a0, i2)    -> lane (c, a0, i2), registers i1
//   3  MN rows B    k = a0 (= a), m = (c, i2, i1)    -> lane (c, i2, i1), registers i0      then x detJ[c][i0][i1][i2]
//   4  K-major      k = i0 (registers)               -> lane (c, i2, i1), registers j0
//   5  MN rows A    k = i1 (= b), m = (c, i2, j0)    -> lane (c, i2, j0), registers j1
//   6  MN rows B    k = i2 (= a), m = (c, j0, j1)    -> lane (c, j0, j1), registers j2 = the natural row of W
__global__ void __launch_bounds__(128) mass_tf32_kernel(const __grid_constant__ Tf32Tab tab, const float* __restrict__ U,
                                                        float* __restrict__ W, const float* __restrict__ detJ,
                                                        long long ntiles)
{
  // CUDA promises 16-byte alignment of the dynamic shared-memory base, nothing more; the SWIZZLE_128B operand
  // descriptors below are built with base offset 0 and need 1024: round up here (the launch asks for 1 KB of slack)
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  Tf32Smem& s = *reinterpret_cast<Tf32Smem*>(smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u));
  const int tid = threadIdx.x;

  if (tid < 64) // B[k = p][n = i] = A(i, p) = phi[i + p n1] (mass.hpp:20-21), as K-major core matrices (8 n x 16 bytes)
  {
    const int i = tid & 7, p = tid >> 3;
    const float m = tab.a[i + p * TN1];
    const float hi = tf32_rna(m), lo = tf32_rna(m - hi);
    const int at_hi = (p >> 2) * 32 + i * 4 + (p & 3), at_lo = at_hi + 64; // n-group 1 (n = 8..15) is 256 bytes on
    s.tabHL[at_hi] = hi;
    s.tabHL[at_lo] = lo;
    s.tabH0[at_hi] = hi;
    s.tabH0[at_lo] = 0.f;
  }
  if (tid == 0)
  {
    mbar_init(smem_u32(&s.full[0]), 1);
    mbar_init(smem_u32(&s.full[1]), 1);
    mbar_init(smem_u32(&s.mma_done), 1);
    fence_mbar_init();
  }
  if (tid < 32)
    tmem_alloc(smem_u32(&s.tmem_base), TMEM_COLS);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s.tmem_base;

  auto request_U = [&](long long tile, int st)
  {
    // one arrival per ring stage and phase: it announces the bytes of both copies; detJ's copy is issued later
    mbar_expect_tx(smem_u32(&s.full[st]), 2 * TILE_BYTES);
    bulk_g2s(smem_u32(s.inU[st]), U + tile * TILE_FLOATS, TILE_BYTES, smem_u32(&s.full[st]));
  };
  auto request_J = [&](long long tile, int st)
  { bulk_g2s(smem_u32(s.inJ[st]), detJ + tile * TILE_FLOATS, TILE_BYTES, smem_u32(&s.full[st])); };

  const long long stride = gridDim.x;
  if (tid == 0)
    for (int k = 0; k < 2; ++k)
    {
      const long long t = blockIdx.x + k * stride;
      if (t < ntiles)
      {
        request_U(t, k);
        request_J(t, k);
      }
    }

  uint32_t mma_parity = 0;
  long long k = 0;
  for (long long tile = blockIdx.x; tile < ntiles; tile += stride, ++k)
  {
    const int st = (int)(k & 1);
    mbar_wait(smem_u32(&s.full[st]), (uint32_t)((k >> 1) & 1));
    float x[8];
    // this lane's row of U: 32 bytes at a 32-byte lane pitch; the upper four lanes of each eight read their halves in
    // the other order (no bank conflict) and the K-major store puts them back in place
    const bool swapped = (tid & 4) != 0;
    {
      const float4* src = reinterpret_cast<const float4*>(s.inU[st] + tid * 8);
      const float4 v0 = src[swapped ? 1 : 0], v1 = src[swapped ? 0 : 1];
      x[0] = v0.x, x[1] = v0.y, x[2] = v0.z, x[3] = v0.w, x[4] = v1.x, x[5] = v1.y, x[6] = v1.z, x[7] = v1.w;
    }
    if (tid == 0)
      bulk_wait_read0(); // the previous tile's W store has read outW (all threads pass a barrier before writing it)
    contract<K_MAJOR>(s, x, tmem, mma_parity, swapped);
    // inU[st] was read into registers that fed a shared-memory store before the barrier inside contract(): free
    const long long next = tile + 2 * stride;
    if (tid == 0 && next < ntiles)
      request_U(next, st);
    contract<MN_ROWS_A>(s, x, tmem, mma_parity);
    contract<MN_ROWS_B>(s, x, tmem, mma_parity);
    {
      const int c = tid >> 6, i2 = (tid >> 3) & 7, i1 = tid & 7;
      const float* j = s.inJ[st] + c * TND + i1 * 8 + i2;
#pragma unroll
      for (int i0 = 0; i0 < 8; ++i0)
        x[i0] *= j[i0 * 64];
    }
    contract<K_MAJOR>(s, x, tmem, mma_parity);
    if (tid == 0 && next < ntiles)
      request_J(next, st);
    contract<MN_ROWS_A>(s, x, tmem, mma_parity);
    contract<MN_ROWS_B>(s, x, tmem, mma_parity);
    {
      const uint32_t o = smem_u32(s.outW) + (uint32_t)tid * 32u;
      sts128(o, x[0], x[1], x[2], x[3]);
      sts128(o + 16, x[4], x[5], x[6], x[7]);
    }
    fence_proxy_async();
    __syncthreads();
    if (tid == 0)
    {
      bulk_s2g(W + tile * TILE_FLOATS, smem_u32(s.outW), TILE_BYTES);
      bulk_commit();
    }
  }
  if (tid == 0)
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  if (tid < 32)
    tmem_dealloc(tmem, TMEM_COLS);
}

inline bool al16f(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

} // namespace

// Option "tc": 1 = use this arm where it applies (fp32 mass, n1 = 8); -1 (default) = only where it is measured faster
// than the FFMA2 column kernels (nowhere: profiles/r02_tf32_arm.txt); 0 = never.
long long launch_mass_tf32(const MassArgs<float>& a)
{
  if (a.degree + 1 != TN1 || opt_tc() != 1)
    return 0;
  const long long ntiles = (long long)(a.num_cells / TILE_CELLS);
  if (ntiles == 0 || !al16f(a.U) || !al16f(a.W) || !al16f(a.detJ))
    return 0;
  Tf32Tab B;
  for (int i = 0; i < TN1 * TN1; ++i)
    B.a[i] = a.phi_host[i];
  auto kernel = mass_tf32_kernel;
  constexpr size_t smem = sizeof(Tf32Smem) + 1024; // + slack for the in-kernel round-up of the base to 1024 bytes
  static std::mutex mu;
  static std::map<int, int> plans; // device -> resident blocks per SM
  int dev = 0, bps = 0;
  if (cudaGetDevice(&dev) != cudaSuccess)
    return -FEMTC_ERR_CUDA;
  {
    std::lock_guard<std::mutex> lock(mu);
    const auto it = plans.find(dev);
    if (it != plans.end())
      bps = it->second;
    else
    {
      // The occupancy query answers 1 for a kernel that allocates tensor memory (it cannot know how many columns), so
      // the resident blocks are counted here: shared memory (dynamic + 1 KB system share per block) and 512 TMEM columns.
      int smem_sm = 0;
      if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess ||
          cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100) != cudaSuccess ||
          cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev) != cudaSuccess)
      {
        (void)cudaGetLastError();
        set_error("tf32 tensor-core kernel: launch planning failed");
        return -FEMTC_ERR_CUDA;
      }
      bps = (int)std::min<size_t>(smem_sm / (smem + 1024), 512 / TMEM_COLS);
      if (bps < 1)
      {
        set_error("tf32 tensor-core kernel does not fit on an SM");
        return -FEMTC_ERR_CUDA;
      }
      plans[dev] = bps;
    }
  }
  const long long grid = std::min<long long>(ntiles, (long long)sm_count() * bps);
  FEMTC_LAUNCH(FAMILY_TC, 32, TILE_CELLS, bps, kernel, (int)grid, 128, smem, a.stream, B, a.U, a.W, a.detJ, ntiles);
  return ntiles * TILE_CELLS;
}

} // namespace femtc
