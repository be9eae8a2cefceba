// Batched rectangular tensor contraction: the first-class kernel behind tmm.hpp's apply_contractions
// (reference src/tmm.hpp:91-126; SURVEY 8f rank 2: quadrature points != DoFs).
//
//     A[e][a0,a1,a2] = sum_{b0,b1,b2} I0(a0,b0) I1(a1,b1) I2(a2,b2) B[e][b0,b1,b2],     e = 0 .. batch-1
//
// with B tensors of Nb^3 and A tensors of Na^3 values, contiguous per tensor.  The reference contracts one
// axis at a time with explicit transposes in between (contract + transpose, tmm.hpp:65-85, :37-44); here the
// three axis sweeps run inside one kernel on a tile of C tensors held in shared memory, so a tensor crosses
// HBM once in and once out:
//     phase 0   thread (e, b1, b2)   column along b0 of the staged input  ->  S0[e][a0][b1][b2]
//     phase 1   thread (e, a0, b2)   column along b1 of S0                 ->  S1[e][a0][a1][b2]   (rows padded)
//     phase 2   thread (e, a0, a1)   row    along b2 of S1                 ->  out[e][a0][a1][a2]   (natural order)
// Input tiles arrive by one TMA bulk copy (requested for the block's next tile as soon as phase 0 has read the
// current one), output tiles leave by one bulk store.  The three Na x Nb matrices are kernel parameters
// (constant bank), already in I(a,b) = m[b*Na + a] form whatever the caller's `Tr`.  Sums run over b ascending
// from zero, as the reference's accumulate-into-zeroed-buffer does (tmm.hpp:24-32, :65-85).
//
// Algorithmic HBM bytes per tensor: (Nb^3 + Na^3) * sizeof(T).
#include "common.cuh"
#include "tma.cuh"

#include <algorithm>

namespace femtc
{
namespace
{
constexpr int tmax(int a, int b) { return a > b ? a : b; }
constexpr int tup128(int bytes) { return (bytes + 127) & ~127; }

#ifndef FEMTC_FFMA2
#define FEMTC_FFMA2 1
#endif
// Tables in [b][a] order, I(a,b) = m[b*NA + a]: the outputs a, a+1 of one input b are neighbours, which is what
// the packed fp32 FMA (FFMA2) wants as its table operand.
template <typename T, int NA, int NB>
struct alignas(16) TmmTab
{
  T i0[NA * NB];
  T i1[NA * NB];
  T i2[NA * NB];
};

// y[a] = sum_b I(a,b) x[b], b ascending from zero (tmm.hpp:24-32); fp32 computes a, a+1 together
template <typename T, int NA, int NB>
__device__ __forceinline__ void tmm_column(const T (&m)[NA * NB], const T (&x)[NB], T (&y)[NA])
{
  if constexpr (sizeof(T) == 4 && FEMTC_FFMA2)
  {
#pragma unroll
    for (int a = 0; a + 1 < NA; a += 2)
    {
      float2 acc = make_float2(0.f, 0.f);
#pragma unroll
      for (int b = 0; b < NB; ++b)
        acc = __ffma2_rn(make_float2(m[b * NA + a], m[b * NA + a + 1]), make_float2(x[b], x[b]), acc);
      y[a] = acc.x, y[a + 1] = acc.y;
    }
    if constexpr (NA & 1)
    {
      T acc = T(0);
#pragma unroll
      for (int b = 0; b < NB; ++b)
        acc = fma_t(m[b * NA + NA - 1], x[b], acc);
      y[NA - 1] = acc;
    }
  }
  else
  {
#pragma unroll
    for (int a = 0; a < NA; ++a)
    {
      T acc = T(0);
#pragma unroll
      for (int b = 0; b < NB; ++b)
        acc = fma_t(m[b * NA + a], x[b], acc);
      y[a] = acc;
    }
  }
}

template <typename T, int NA, int NB>
struct TmmPlan
{
  static constexpr int NB2 = NB * NB, NB3 = NB2 * NB, NA2 = NA * NA, NA3 = NA2 * NA;
  static constexpr int widest = tmax(tmax(NB2, NA * NB), NA2);   // threads per tensor in the widest phase
  // tensors per tile: enough for about 256 threads in the widest phase, and at least ~16 KB of output so that
  // the four block barriers and the two bulk copies of a tile are amortised for the small shapes too (threads
  // then loop over the tile); always whole 16-byte units for both bulk copies
#ifndef FEMTC_TMM_MIN_OUT_BYTES
#define FEMTC_TMM_MIN_OUT_BYTES 16384
#endif
  static constexpr int c_threads = (256 / widest) & ~3;
  static constexpr int c_bytes = (FEMTC_TMM_MIN_OUT_BYTES / (NA3 * (int)sizeof(T))) & ~3;
  static constexpr int C = tmax(4, tmax(c_threads, c_bytes));
  static constexpr int NT = C * widest >= 512 ? 512 : ((C * widest + 31) & ~31);
  static constexpr int LD1 = NB | 1;                              // padded row of S1: odd stride for phase 2
  // S0 plane stride: phase 1 reads, for fixed b1, runs of NB values from consecutive (e, a0) planes; with the
  // stride = NB (mod 32) those runs tile the banks instead of piling up (NB = 8: 64 -> 72)
  static constexpr int LD0 = NB < 4 ? NB2 : NB2 + ((NB - NB2) % 32 + 32) % 32;
  static constexpr uint32_t in_bytes = C * NB3 * sizeof(T);
  static constexpr uint32_t out_bytes = C * NA3 * sizeof(T);
  static_assert(in_bytes % 16 == 0 && out_bytes % 16 == 0, "tiles must be whole 16-byte units");
  static constexpr int off_in = 0;
  static constexpr int off_s0 = off_in + tup128(in_bytes);
  static constexpr int off_s1 = off_s0 + tup128(C * NA * LD0 * (int)sizeof(T));
  static constexpr int off_out = off_s1 + tup128(C * NA2 * LD1 * (int)sizeof(T));
  static constexpr int off_bar = off_out + tup128(out_bytes);
  static constexpr int smem_bytes = off_bar + 16;
};

template <typename T, int NA, int NB>
__global__ void __launch_bounds__(TmmPlan<T, NA, NB>::NT)
tmm_kernel(const __grid_constant__ TmmTab<T, NA, NB> M, const T* __restrict__ B, T* __restrict__ A, long long ntiles)
{
  using PL = TmmPlan<T, NA, NB>;
  constexpr int C = PL::C, NB2 = PL::NB2, NB3 = PL::NB3, NA2 = PL::NA2, NA3 = PL::NA3, LD1 = PL::LD1, LD0 = PL::LD0;
  extern __shared__ __align__(128) unsigned char smem[];
  T* inS = reinterpret_cast<T*>(smem + PL::off_in);
  T* s0 = reinterpret_cast<T*>(smem + PL::off_s0);
  T* s1 = reinterpret_cast<T*>(smem + PL::off_s1);
  T* outS = reinterpret_cast<T*>(smem + PL::off_out);
  const uint32_t bar = smem_u32(smem + PL::off_bar);
  const int tid = threadIdx.x;

  long long tile = blockIdx.x;
  if (tid == 0)
  {
    mbar_init(bar, 1);
    fence_mbar_init();
    if (tile < ntiles)
    {
      mbar_expect_tx(bar, PL::in_bytes);
      bulk_g2s(smem_u32(inS), B + (size_t)tile * C * NB3, PL::in_bytes, bar);
    }
  }
  __syncthreads();

  uint32_t phase = 0;
  for (; tile < ntiles; tile += gridDim.x, phase ^= 1)
  {
    const long long tn = tile + gridDim.x;
    // ---- phase 0: axis 0 (tmm.hpp:100-104) -----------------------------------------------------------
    mbar_wait(bar, phase);
    for (int idx = tid; idx < C * NB2; idx += PL::NT)
    {
      const int e = idx / NB2, d = idx - e * NB2;
      T x[NB];
#pragma unroll
      for (int b = 0; b < NB; ++b)
        x[b] = inS[e * NB3 + b * NB2 + d];
      T y[NA];
      tmm_column<T, NA, NB>(M.i0, x, y);
#pragma unroll
      for (int a = 0; a < NA; ++a)
        s0[(e * NA + a) * LD0 + d] = y[a];
    }
    __syncthreads(); // (1) input staging consumed, S0 complete
    if (tid == 0 && tn < ntiles)
    {
      mbar_expect_tx(bar, PL::in_bytes);
      bulk_g2s(smem_u32(inS), B + (size_t)tn * C * NB3, PL::in_bytes, bar);
    }
    // ---- phase 1: axis 1 (tmm.hpp:108-112) -----------------------------------------------------------
    for (int idx = tid; idx < C * NA * NB; idx += PL::NT)
    {
      const int e = idx / (NA * NB), r = idx - e * (NA * NB), a0 = r / NB, b2 = r - a0 * NB;
      T x[NB];
#pragma unroll
      for (int b = 0; b < NB; ++b)
        x[b] = s0[(e * NA + a0) * LD0 + b * NB + b2];
      T y[NA];
      tmm_column<T, NA, NB>(M.i1, x, y);
#pragma unroll
      for (int a = 0; a < NA; ++a)
        s1[((e * NA + a0) * NA + a) * LD1 + b2] = y[a];
    }
    __syncthreads(); // (2)
    // ---- phase 2: axis 2 (tmm.hpp:116-120), natural-order output tile -------------------------------------
    if (tid == 0)
      bulk_wait_read0(); // the previous tile's store has finished reading outS
    __syncthreads();     // (3)
    for (int row = tid; row < C * NA2; row += PL::NT)
    {
      T x[NB];
#pragma unroll
      for (int b = 0; b < NB; ++b)
        x[b] = s1[row * LD1 + b];
      T y[NA];
      tmm_column<T, NA, NB>(M.i2, x, y);
      // a thread's NA outputs are contiguous: 16-byte stores where the row allows it cut the bank conflicts of
      // the even row strides (NA = 8 fp32: 8-way on scalar stores, 2-way on 128-bit stores)
      constexpr int V = 16 / (int)sizeof(T);
      if constexpr ((NA * sizeof(T)) % 16 == 0)
      {
#pragma unroll
        for (int a = 0; a < NA; a += V)
        {
          if constexpr (sizeof(T) == 4)
            *reinterpret_cast<float4*>(outS + row * NA + a) = make_float4(y[a], y[a + 1], y[a + 2], y[a + 3]);
          else
            *reinterpret_cast<double2*>(outS + row * NA + a) = make_double2(y[a], y[a + 1]);
        }
      }
      else if constexpr ((NA * sizeof(T)) % 8 == 0 && sizeof(T) == 4)
      {
#pragma unroll
        for (int a = 0; a < NA; a += 2)
          *reinterpret_cast<float2*>(outS + row * NA + a) = make_float2(y[a], y[a + 1]);
      }
      else
      {
#pragma unroll
        for (int a = 0; a < NA; ++a)
          outS[row * NA + a] = y[a];
      }
    }
    fence_proxy_async();
    __syncthreads(); // (4)
    if (tid == 0)
    {
      bulk_s2g(A + (size_t)tile * C * NA3, smem_u32(outS), PL::out_bytes);
      bulk_commit();
    }
  }
  if (tid == 0)
    bulk_wait_read0();
}

inline bool tal16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// I(a,b) in [b][a] form from the caller's layout (tmm.hpp:65-85: Tr ? I[a*Nb+b] : I[b*Na+a])
template <typename T, int NA, int NB>
void fill_tab(T* dst, const T* src, int Tr)
{
  for (int a = 0; a < NA; ++a)
    for (int b = 0; b < NB; ++b)
      dst[b * NA + a] = Tr ? src[a * NB + b] : src[b * NA + a];
}

template <typename T, int NA, int NB>
long long run_tmm(int Tr, const T* I0, const T* I1, const T* I2, const T* dB, T* dA, size_t batch, cudaStream_t s)
{
  using PL = TmmPlan<T, NA, NB>;
  const long long ntiles = (long long)(batch / PL::C);
  if (ntiles == 0 || !tal16(dB) || !tal16(dA))
    return 0;
  TmmTab<T, NA, NB> M;
  fill_tab<T, NA, NB>(M.i0, I0, Tr);
  fill_tab<T, NA, NB>(M.i1, I1, Tr);
  fill_tab<T, NA, NB>(M.i2, I2, Tr);
  auto kernel = tmm_kernel<T, NA, NB>;
  // planned once per shape (thread-safe static): the attribute calls and the occupancy query cost more host time
  // than a small batch spends on the device
  static const int bps = [&]() {
    int b = 0;
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PL::smem_bytes) != cudaSuccess ||
        cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, kernel, PL::NT, PL::smem_bytes) != cudaSuccess || b < 1)
    {
      (void)cudaGetLastError();
      return 0;
    }
    return b;
  }();
  if (bps < 1)
  {
    set_error("tmm kernel: launch planning failed");
    return -FEMTC_ERR_CUDA;
  }
  const int grid = (int)std::min<long long>(ntiles, (long long)sm_count() * bps);
  kernel<<<grid, PL::NT, PL::smem_bytes, s>>>(M, dB, dA, ntiles);
  count_launch();
  if (cudaGetLastError() != cudaSuccess)
    return -FEMTC_ERR_CUDA;
  return ntiles * PL::C;
}

template <typename T, int NB>
long long run_tmm_nb(int Na, int Tr, const T* I0, const T* I1, const T* I2, const T* dB, T* dA, size_t batch, cudaStream_t s)
{
  // Na = Nb - 1 .. Nb + 2: restriction, collocation, over-integration by one and two points per direction
  if constexpr (NB >= 3)
    if (Na == NB - 1) return run_tmm<T, NB - 1, NB>(Tr, I0, I1, I2, dB, dA, batch, s);
  if (Na == NB) return run_tmm<T, NB, NB>(Tr, I0, I1, I2, dB, dA, batch, s);
  if (Na == NB + 1) return run_tmm<T, NB + 1, NB>(Tr, I0, I1, I2, dB, dA, batch, s);
  if (Na == NB + 2) return run_tmm<T, NB + 2, NB>(Tr, I0, I1, I2, dB, dA, batch, s);
  return 0;
}
} // namespace

// Fast path of apply_contractions.  I0, I1, I2 are HOST pointers (they travel as kernel parameters), dB / dA
// device pointers.  Returns the number of leading tensors served (0 = shape or alignment not covered: the
// caller runs the generic kernel), or -(FEMTC_ERR_*).
template <typename T>
long long launch_tmm_tiled(int Na, int Nb, int Tr, const T* I0, const T* I1, const T* I2, const T* dB, T* dA,
                           size_t batch, cudaStream_t s)
{
  switch (Nb)
  {
  case 2: return run_tmm_nb<T, 2>(Na, Tr, I0, I1, I2, dB, dA, batch, s);
  case 3: return run_tmm_nb<T, 3>(Na, Tr, I0, I1, I2, dB, dA, batch, s);
  case 4: return run_tmm_nb<T, 4>(Na, Tr, I0, I1, I2, dB, dA, batch, s);
  case 5: return run_tmm_nb<T, 5>(Na, Tr, I0, I1, I2, dB, dA, batch, s);
  case 6: return run_tmm_nb<T, 6>(Na, Tr, I0, I1, I2, dB, dA, batch, s);
  case 7: return run_tmm_nb<T, 7>(Na, Tr, I0, I1, I2, dB, dA, batch, s);
  case 8: return run_tmm_nb<T, 8>(Na, Tr, I0, I1, I2, dB, dA, batch, s);
  default: return 0;
  }
}

template long long launch_tmm_tiled<double>(int, int, int, const double*, const double*, const double*, const double*,
                                            double*, size_t, cudaStream_t);
template long long launch_tmm_tiled<float>(int, int, int, const float*, const float*, const float*, const float*, float*,
                                           size_t, cudaStream_t);

} // namespace femtc
