// Cell kernels: one THREAD per cell, for the smallest cells (n1 = 2, and n1 = 3 where the slab
// kernels of tiled_kernels.cu cannot stage fp32 rounds with 16-byte bulk copies).
//
// Work decomposition.  A *round* is 32 consecutive cells handled by one warp; lane l owns cell l of
// the round and keeps the whole n1^3 cell tensor in registers.  All three axis contractions are
// then fully unrolled FMA chains on registers whose matrix entries are immediate constant-bank
// operands (the 1-D tables are kernel parameters): there is no exchange through shared memory, no
// index arithmetic and no barrier other than the warp-level ones of the copy pipeline.
//     mass       6 axis contractions + the quadrature scaling            (mass.hpp:36-47)
//     stiffness  REF_EXACT 9 contractions, FULL 16 (shared products once) (stiffness.hpp:68-97)
//
// Memory pipeline (per warp, NS stages).  A round of U, detJ or G is one contiguous, 16-byte
// aligned byte range of HBM.  The warp copies it with coalesced 16-byte cp.async (LDGSTS: lane j
// moves chunk j, j+32, ...), each chunk landing where the owning lane can read its cell without
// bank conflicts:
//     n1 = 2: cells are whole 16-byte chunks -> one pad chunk per cell
//             (U: 48 / 80 B cell stride, G: 208 / 400 B; 128-bit reads, conflict free)
//     n1 = 3: natural layout; the odd cell strides (27 / 54 / 162 / 324 words) already spread the
//             lanes over the banks for 32- / 64- / 128-bit reads
// Rounds k+1 .. k+NS-1 are in flight while round k is computed.  W goes back through a natural-
// layout staging buffer and one TMA bulk store per round, so every byte of U, detJ / G and W
// crosses HBM exactly once, fully coalesced.
//
// Semantics follow the reference as coded (REF_EXACT / FULL: include/femtc_b200.h).
#include "common.cuh"
#include "tma.cuh"

#include <algorithm>

namespace femtc
{
long long opt_variant();

namespace
{
constexpr int up16c(int b) { return (b + 15) & ~15; }
constexpr int up128c(int b) { return (b + 127) & ~127; }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src)
{
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <typename T, int N1>
struct TabA1
{
  T a[N1 * N1];
};
template <typename T, int N1>
struct TabAD1
{
  T a[N1 * N1];
  T d[N1 * N1];
};

// Staged field of one round: CELLV values per cell, 32 cells.
template <typename T, int CELLV>
struct Field
{
  static constexpr int cell_bytes = CELLV * (int)sizeof(T);
  static constexpr bool padded = cell_bytes % 16 == 0;           // n1 = 2
  static constexpr int CH = cell_bytes / 16;                     // chunks per cell (padded only)
  static constexpr int stride_bytes = padded ? cell_bytes + 16 : cell_bytes;
  static constexpr int round_bytes = 32 * cell_bytes;
  static constexpr int chunks = round_bytes / 16;
  static constexpr int smem_bytes = up128c(32 * stride_bytes);
  static_assert(round_bytes % 16 == 0, "a round is a whole number of 16-byte chunks");

  // coalesced copy of a round: lane j moves chunks j, j+32, ...
  __device__ static __forceinline__ void issue(unsigned char* sdst, const T* gsrc_round, int lane)
  {
    const unsigned char* g = reinterpret_cast<const unsigned char*>(gsrc_round);
    const uint32_t s = smem_u32(sdst);
#pragma unroll
    for (int i = 0; i < (chunks + 31) / 32; ++i)
    {
      const int j = lane + 32 * i;
      if ((i + 1) * 32 <= chunks || j < chunks)
      {
        const int dst = padded ? 16 * (j + j / CH) : 16 * j;
        cp_async16(s + dst, g + 16 * (size_t)j);
      }
    }
  }
  __device__ static __forceinline__ const T* cell(const unsigned char* s, int lane)
  {
    return reinterpret_cast<const T*>(s + lane * stride_bytes);
  }
};

// cell tensor <- staging.  Vector width: what the cell's alignment allows (see the header comment).
template <typename T, int N1>
__device__ __forceinline__ void load_cell(const T* p, T (&x)[N1][N1][N1])
{
  constexpr int ND = N1 * N1 * N1;
  T* flat = &x[0][0][0];
  if constexpr ((ND * sizeof(T)) % 16 == 0)
  {
#pragma unroll
    for (int i = 0; i < ND; i += 16 / sizeof(T))
    {
      const float4 v = *reinterpret_cast<const float4*>(p + i);
      const T* e = reinterpret_cast<const T*>(&v);
#pragma unroll
      for (int k = 0; k < (int)(16 / sizeof(T)); ++k)
        flat[i + k] = e[k];
    }
  }
  else if constexpr (sizeof(T) == 8)
  {
#pragma unroll
    for (int i = 0; i < ND; ++i)
      flat[i] = p[i]; // 64-bit reads, stride 54 words: conflict free
  }
  else
  {
#pragma unroll
    for (int i = 0; i < ND; ++i)
      flat[i] = p[i]; // 32-bit reads, stride 27 words: conflict free
  }
}

template <typename T, int N1>
__device__ __forceinline__ void store_cell(T* p, const T (&x)[N1][N1][N1])
{
  constexpr int ND = N1 * N1 * N1;
  const T* flat = &x[0][0][0];
  if constexpr ((ND * sizeof(T)) % 16 == 0)
  {
#pragma unroll
    for (int i = 0; i < ND; i += 16 / sizeof(T))
    {
      float4 v;
      T* e = reinterpret_cast<T*>(&v);
#pragma unroll
      for (int k = 0; k < (int)(16 / sizeof(T)); ++k)
        e[k] = flat[i + k];
      *reinterpret_cast<float4*>(p + i) = v;
    }
  }
  else
  {
#pragma unroll
    for (int i = 0; i < ND; ++i)
      p[i] = flat[i];
  }
}

// out[.. i ..] = sum_p M(i,p) in[.. p ..] along AXIS, M(i,p) = M[i + p*N1]; p ascending starting from
// the plain product (0 + a*b is exact: linalg.hpp:57,128).  in and out must be different arrays.
template <int AXIS, typename T, int N1>
__device__ __forceinline__ void contract3(const T (&M)[N1 * N1], const T (&in)[N1][N1][N1], T (&out)[N1][N1][N1])
{
#pragma unroll
  for (int a = 0; a < N1; ++a)
#pragma unroll
    for (int b = 0; b < N1; ++b)
#pragma unroll
      for (int i = 0; i < N1; ++i)
      {
        auto at = [&](const T(&t)[N1][N1][N1], int r) -> T
        { return AXIS == 0 ? t[r][a][b] : (AXIS == 1 ? t[a][r][b] : t[a][b][r]); };
        T acc = M[i] * at(in, 0);
#pragma unroll
        for (int p = 1; p < N1; ++p)
          acc = fma_t(M[i + p * N1], at(in, p), acc);
        (AXIS == 0 ? out[i][a][b] : (AXIS == 1 ? out[a][i][b] : out[a][b][i])) = acc;
      }
}
// out += M . in along AXIS
template <int AXIS, typename T, int N1>
__device__ __forceinline__ void contract3_add(const T (&M)[N1 * N1], const T (&in)[N1][N1][N1], T (&out)[N1][N1][N1])
{
#pragma unroll
  for (int a = 0; a < N1; ++a)
#pragma unroll
    for (int b = 0; b < N1; ++b)
#pragma unroll
      for (int i = 0; i < N1; ++i)
      {
        auto at = [&](const T(&t)[N1][N1][N1], int r) -> T
        { return AXIS == 0 ? t[r][a][b] : (AXIS == 1 ? t[a][r][b] : t[a][b][r]); };
        T acc = (AXIS == 0 ? out[i][a][b] : (AXIS == 1 ? out[a][i][b] : out[a][b][i]));
#pragma unroll
        for (int p = 0; p < N1; ++p)
          acc = fma_t(M[i + p * N1], at(in, p), acc);
        (AXIS == 0 ? out[i][a][b] : (AXIS == 1 ? out[a][i][b] : out[a][b][i])) = acc;
      }
}

// ---- geometry record of point q of the lane's cell: values [zz,yz,xz,yy,xy,xx] (stiffness.hpp:20-27) ----
template <bool FULL, typename T>
__device__ __forceinline__ void load_g(const T* rec, T (&g)[6])
{
  if constexpr (sizeof(T) == 8)
  {
    const double2 a = *reinterpret_cast<const double2*>(rec);
    const double2 b = *reinterpret_cast<const double2*>(rec + 2);
    g[0] = a.x, g[1] = a.y, g[2] = b.x, g[3] = b.y;
    if constexpr (FULL)
    {
      const double2 c = *reinterpret_cast<const double2*>(rec + 4);
      g[4] = c.x, g[5] = c.y;
    }
  }
  else
  {
    const float2 a = *reinterpret_cast<const float2*>(rec);
    const float2 b = *reinterpret_cast<const float2*>(rec + 2);
    g[0] = a.x, g[1] = a.y, g[2] = b.x, g[3] = b.y;
    if constexpr (FULL)
    {
      const float2 c = *reinterpret_cast<const float2*>(rec + 4);
      g[4] = c.x, g[5] = c.y;
    }
  }
}

template <typename T, int N1, bool STIFF, int NS>
struct CellPlan
{
  static constexpr int ND = N1 * N1 * N1;
  using FU = Field<T, ND>;
  using FG = Field<T, STIFF ? 6 * ND : ND>;   // G, or detJ for the mass operator
  static constexpr int w_bytes = up128c(32 * ND * (int)sizeof(T));
  static constexpr int stage_bytes = FU::smem_bytes + FG::smem_bytes;
  static constexpr int off_w = NS * stage_bytes;
  static constexpr int warp_bytes = off_w + w_bytes;
};

template <typename T, int N1, int NS>
__global__ void __launch_bounds__(128)
mass_cell_kernel(const __grid_constant__ TabA1<T, N1> B, const T* __restrict__ U, T* __restrict__ W,
                 const T* __restrict__ J, long long rounds)
{
  using PL = CellPlan<T, N1, false, NS>;
  constexpr int ND = PL::ND;
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* my = smem + warp * PL::warp_bytes;
  T* wS = reinterpret_cast<T*>(my + PL::off_w);
  const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
  const long long r0 = (long long)blockIdx.x * (blockDim.x >> 5) + warp;

  auto issue = [&](long long k)
  {
    const long long r = r0 + k * nwarps;
    if (r < rounds)
    {
      unsigned char* st = my + (int)(k % NS) * PL::stage_bytes;
      PL::FU::issue(st, U + (size_t)r * 32 * ND, lane);
      PL::FG::issue(st + PL::FU::smem_bytes, J + (size_t)r * 32 * ND, lane);
    }
    cp_async_commit();
  };
#pragma unroll
  for (int k = 0; k < NS - 1; ++k)
    issue(k);

  long long k = 0;
  for (long long r = r0; r < rounds; r += nwarps, ++k)
  {
    issue(k + NS - 1);
    cp_async_wait<NS - 1>();
    __syncwarp();
    const unsigned char* st = my + (int)(k % NS) * PL::stage_bytes;
    T x[N1][N1][N1], y[N1][N1][N1];
    load_cell<T, N1>(PL::FU::cell(st, lane), x);
    contract3<2>(B.a, x, y);                               // mass.hpp:36-38
    contract3<1>(B.a, y, x);
    contract3<0>(B.a, x, y);
    load_cell<T, N1>(PL::FG::cell(st + PL::FU::smem_bytes, lane), x);
    __syncwarp(); // every lane is done with this stage: the next issue() may overwrite it
#pragma unroll
    for (int i = 0; i < ND; ++i)
      (&y[0][0][0])[i] *= (&x[0][0][0])[i];                // mass.hpp:40-41
    contract3<0>(B.a, y, x);                               // mass.hpp:45-47 (same matrix)
    contract3<1>(B.a, x, y);
    contract3<2>(B.a, y, x);

    if (lane == 0)
      bulk_wait_read0(); // the previous round's bulk store has finished reading wS
    __syncwarp();
    store_cell<T, N1>(wS + lane * ND, x);
    fence_proxy_async();
    __syncwarp();
    if (lane == 0)
    {
      bulk_s2g(W + (size_t)r * 32 * ND, smem_u32(wS), 32 * ND * sizeof(T));
      bulk_commit();
    }
  }
  cp_async_wait<0>();
  if (lane == 0)
    bulk_wait_read0();
}

template <typename T, int N1, int MODE, int NS>
__global__ void __launch_bounds__(128)
stiffness_cell_kernel(const __grid_constant__ TabAD1<T, N1> B, const T* __restrict__ U, T* __restrict__ W,
                      const T* __restrict__ G, long long rounds)
{
  constexpr bool FULL = MODE == FEMTC_STIFFNESS_FULL;
  using PL = CellPlan<T, N1, true, NS>;
  constexpr int ND = PL::ND;
  extern __shared__ __align__(128) unsigned char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* my = smem + warp * PL::warp_bytes;
  T* wS = reinterpret_cast<T*>(my + PL::off_w);
  const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
  const long long r0 = (long long)blockIdx.x * (blockDim.x >> 5) + warp;

  auto issue = [&](long long k)
  {
    const long long r = r0 + k * nwarps;
    if (r < rounds)
    {
      unsigned char* st = my + (int)(k % NS) * PL::stage_bytes;
      PL::FU::issue(st, U + (size_t)r * 32 * ND, lane);
      PL::FG::issue(st + PL::FU::smem_bytes, G + (size_t)r * 32 * ND * 6, lane);
    }
    cp_async_commit();
  };
#pragma unroll
  for (int k = 0; k < NS - 1; ++k)
    issue(k);

  long long k = 0;
  for (long long r = r0; r < rounds; r += nwarps, ++k)
  {
    issue(k + NS - 1);
    cp_async_wait<NS - 1>();
    __syncwarp();
    const unsigned char* st = my + (int)(k % NS) * PL::stage_bytes;
    const T* gcell = PL::FG::cell(st + PL::FU::smem_bytes, lane);

    T f2[N1][N1][N1];
    T f0[FULL ? N1 : 1][FULL ? N1 : 1][FULL ? N1 : 1], f1[FULL ? N1 : 1][FULL ? N1 : 1][FULL ? N1 : 1];
    {
      T u[N1][N1][N1], t2[N1][N1][N1], t12[N1][N1][N1];
      load_cell<T, N1>(PL::FU::cell(st, lane), u);
      contract3<2>(B.a, u, t2);                            // A2 u
      contract3<1>(B.a, t2, t12);                          // A1 A2 u
      T wx[N1][N1][N1], wy[N1][N1][N1], wz[N1][N1][N1];
      contract3<0>(B.d, t12, wx);                          // Wx = (D x A x A) u     stiffness.hpp:68-70
      if constexpr (!FULL)
        contract3<0>(B.a, t12, wz);                        // Wz = (A x A x A) u     :76-77 (REF_EXACT)
      contract3<1>(B.d, t2, t12);                          // D1 A2 u
      contract3<0>(B.a, t12, wy);                          // Wy = (A x D x A) u     :72-74
      if constexpr (FULL)
      {
        contract3<2>(B.d, u, t2);                          // D2 u
        contract3<1>(B.a, t2, t12);
        contract3<0>(B.a, t12, wz);                        // Wz = (A x A x D) u     (FULL)
      }
      // geometry transform (stiffness.hpp:14-29)
#pragma unroll
      for (int q = 0; q < ND; ++q)
      {
        T g[6];
        load_g<FULL>(gcell + 6 * q, g);
        const T w0 = (&wx[0][0][0])[q], w1 = (&wy[0][0][0])[q], w2 = (&wz[0][0][0])[q];
        if constexpr (FULL)
        {
          (&f0[0][0][0])[q] = fma_t(g[2], w2, fma_t(g[4], w1, g[5] * w0));
          (&f1[0][0][0])[q] = fma_t(g[1], w2, fma_t(g[3], w1, g[4] * w0));
        }
        (&f2[0][0][0])[q] = fma_t(g[0], w2, fma_t(g[1], w1, g[2] * w0));
      }
    }
    __syncwarp(); // every lane is done with this stage

    T out[N1][N1][N1];
    {
      T h[N1][N1][N1], z[N1][N1][N1];
      if constexpr (!FULL)
      {
        // only W = (A x A x D) f2 is observable (stiffness.hpp:95-97 overwrites :83-91)
        contract3<0>(B.a, f2, h);
        contract3<1>(B.a, h, z);
        contract3<2>(B.d, z, out);
      }
      else
      {
        // W = A2 [ A1 D0 f0 + D1 A0 f1 ] + D2 [ A1 A0 f2 ]
        contract3<0>(B.d, f0, h);
        contract3<1>(B.a, h, z);
        contract3<0>(B.a, f1, h);
        contract3_add<1>(B.d, h, z);
        contract3<2>(B.a, z, out);
        contract3<0>(B.a, f2, h);
        contract3<1>(B.a, h, z);
        contract3_add<2>(B.d, z, out);
      }
    }

    if (lane == 0)
      bulk_wait_read0();
    __syncwarp();
    store_cell<T, N1>(wS + lane * ND, out);
    fence_proxy_async();
    __syncwarp();
    if (lane == 0)
    {
      bulk_s2g(W + (size_t)r * 32 * ND, smem_u32(wS), 32 * ND * sizeof(T));
      bulk_commit();
    }
  }
  cp_async_wait<0>();
  if (lane == 0)
    bulk_wait_read0();
}

// ---- launch -----------------------------------------------------------------------------------------------
struct CellLaunch
{
  int warps_per_block, blocks_per_sm;
  size_t smem;
};

template <typename K>
int plan_cell(K kernel, int warp_bytes, CellLaunch& p)
{
  int dev = 0, max_optin = 0;
  FEMTC_CUDA(cudaGetDevice(&dev));
  FEMTC_CUDA(cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  const int max_wpb = std::min(4, max_optin / warp_bytes);
  if (max_wpb < 1)
  {
    set_error("cell kernel does not fit on an SM");
    return FEMTC_ERR_CUDA;
  }
  FEMTC_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, max_wpb * warp_bytes));
  FEMTC_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
  int best = 0;
  for (int wpb = max_wpb; wpb >= 1; --wpb)
  {
    int occ = 0;
    FEMTC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, wpb * 32, (size_t)wpb * warp_bytes));
    if (occ * wpb > best)
    {
      best = occ * wpb;
      p.warps_per_block = wpb;
      p.blocks_per_sm = occ;
      p.smem = (size_t)wpb * warp_bytes;
    }
  }
  if (best < 1)
  {
    set_error("cell kernel does not fit on an SM");
    return FEMTC_ERR_CUDA;
  }
  return FEMTC_OK;
}

inline bool aligned16c(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <typename T, int N1, int NS>
long long run_mass_cell(const MassArgs<T>& a)
{
  using PL = CellPlan<T, N1, false, NS>;
  const long long rounds = (long long)(a.num_cells / 32);
  if (rounds == 0 || !aligned16c(a.U) || !aligned16c(a.W) || !aligned16c(a.detJ))
    return 0;
  TabA1<T, N1> B;
  for (int i = 0; i < N1 * N1; ++i)
    B.a[i] = a.phi_host[i];
  auto kernel = mass_cell_kernel<T, N1, NS>;
  CellLaunch p;
  int rc = plan_cell(kernel, PL::warp_bytes, p);
  if (rc)
    return -rc;
  const long long blocks_needed = (rounds + p.warps_per_block - 1) / p.warps_per_block;
  const int grid = (int)std::min<long long>(blocks_needed, (long long)sm_count() * p.blocks_per_sm);
  FEMTC_LAUNCH(FAMILY_CELL, NS, 32, p.blocks_per_sm, kernel, grid, p.warps_per_block * 32, p.smem, a.stream, B, a.U, a.W, a.detJ, rounds);
  return rounds * 32;
}

template <typename T, int N1, int MODE, int NS>
long long run_stiff_cell(const StiffnessArgs<T>& a)
{
  using PL = CellPlan<T, N1, true, NS>;
  const long long rounds = (long long)(a.num_cells / 32);
  if (rounds == 0 || !aligned16c(a.U) || !aligned16c(a.W) || !aligned16c(a.G))
    return 0;
  TabAD1<T, N1> B;
  for (int i = 0; i < N1 * N1; ++i)
  {
    B.a[i] = a.A_host[i];
    B.d[i] = a.D_host[i];
  }
  auto kernel = stiffness_cell_kernel<T, N1, MODE, NS>;
  CellLaunch p;
  int rc = plan_cell(kernel, PL::warp_bytes, p);
  if (rc)
    return -rc;
  const long long blocks_needed = (rounds + p.warps_per_block - 1) / p.warps_per_block;
  const int grid = (int)std::min<long long>(blocks_needed, (long long)sm_count() * p.blocks_per_sm);
  FEMTC_LAUNCH(FAMILY_CELL, NS, 32, p.blocks_per_sm, kernel, grid, p.warps_per_block * 32, p.smem, a.stream, B, a.U, a.W, a.G, rounds);
  return rounds * 32;
}

template <typename T, int N1, int NS>
long long run_stiff_cell_mode(const StiffnessArgs<T>& a)
{
  return a.mode == FEMTC_STIFFNESS_FULL ? run_stiff_cell<T, N1, FEMTC_STIFFNESS_FULL, NS>(a)
                                        : run_stiff_cell<T, N1, FEMTC_STIFFNESS_REF_EXACT, NS>(a);
}

} // namespace

// Which (type, n1) the cell family serves.  variant (development knob): 0 = default stage count,
// 1 / 2 = alternative stage counts, 9 = family disabled (fall through to the slab / column kernels).
template <typename T>
long long launch_mass_cell(const MassArgs<T>& a)
{
  const int n1 = a.degree + 1;
  const long long v = opt_variant();
  if (v == 9)
    return 0;
  if (n1 == 2)
  {
    // fp64: the slab kernel is faster (96 % vs 90 % of measured HBM at 8 M cells) unless asked for
    if (sizeof(T) == 8 && v < 1)
      return 0;
    return v == 1 ? run_mass_cell<T, 2, 2>(a) : (v == 2 ? run_mass_cell<T, 2, 3>(a) : run_mass_cell<T, 2, 4>(a));
  }
  if (n1 == 3)
    return v == 1 ? run_mass_cell<T, 3, 2>(a) : (v == 2 ? run_mass_cell<T, 3, 4>(a) : run_mass_cell<T, 3, 3>(a));
  return 0;
}

template <typename T>
long long launch_stiffness_cell(const StiffnessArgs<T>& a)
{
  const int n1 = a.degree + 1;
  const long long v = opt_variant();
  if (v == 9)
    return 0;
  if (n1 == 2)
  {
    if (v == 1)
      return run_stiff_cell_mode<T, 2, 4>(a);
    // measured best: two stages in fp64, three in fp32
    return (sizeof(T) == 8) == (v == 0) ? run_stiff_cell_mode<T, 2, 2>(a) : run_stiff_cell_mode<T, 2, 3>(a);
  }
  if constexpr (sizeof(T) == 4) // fp64 n1 = 3 (six live cell tensors = 324 registers) stays on the slab kernels
  {
    if (n1 == 3)
      return v == 1 ? run_stiff_cell_mode<T, 3, 3>(a) : run_stiff_cell_mode<T, 3, 2>(a);
  }
  return 0;
}

template long long launch_mass_cell<double>(const MassArgs<double>&);
template long long launch_mass_cell<float>(const MassArgs<float>&);
template long long launch_stiffness_cell<double>(const StiffnessArgs<double>&);
template long long launch_stiffness_cell<float>(const StiffnessArgs<float>&);

} // namespace femtc
