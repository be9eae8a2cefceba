// Tensor-core kernels for the high degrees (n1 = 6..9, fp64): BASELINE config 4.
//
// Why.  The column kernels (column_kernels.cu) are instruction-issue bound at high degree: an fp64
// FMA cannot take a constant-bank operand, so every DFMA of a rolled contraction is paired with a
// uniform-register fetch of its matrix entry (ncu: 10.2 warp instructions per DoF at n1 = 9, a quarter
// of them DFMA, issue 56 %, DRAM 49 %).  DMMA.8x8x4 has the same peak as the DFMA pipe on B200
// (measured: 37.1 vs 34.0 TFLOP/s, tools/probe/mma_probe.cu) but performs 256 FMAs per warp
// instruction with the 8x4 matrix fragment held in ONE register pair: the contraction drops from
// ~25 to ~5 issued instructions per 8 tensor columns.
//
// Decomposition.  One BLOCK of NW warps owns one cell at a time; the tiles of every phase are dealt
// round-robin to its warps and a block barrier separates the phases.  A contraction
// along one axis is a GEMM  Y[i, col] = sum_p M(i,p) X[p, col]  over the n1^2 columns of the cell:
//     A operand = the 1-D table M, rows i = 0..7, two k-steps p = 0..3, 4..7 (zero padded below n1 = 8);
//     B operand = 8 consecutive tensor columns; C = Y[i = lane/4][col 2(lane%4), +1].
// n1 = 9 is 8 + 1: the p = 8 term enters as the accumulator's initial value (rank-1, two DMULs), the
// row i = 8 is a short DFMA pass with one lane per column (or a second m-tile where the result must
// stay in fragment layout for the geometry).  The last tile of a cell overlaps its predecessor, so
// all tiles are full (n1^2 = 81 -> 11 tiles).
// Between contractions the tensor lives in the block's shared memory in the layout the next
// contraction reads conflict-free:  [p][col] with plane stride S = 4 or 12 mod 16 doubles, columns
// contiguous -- every fragment load is `lane base + immediate`.
//     stiffness REF_EXACT   A2 | A1,D1 | D0,A0,A0 + geometry on the accumulators | A0 | A1 | D2   (9 GEMMs)
//     mass                  A2 | A1 | A0 * detJ | A0 | A1 | A2                                        (6 GEMMs)
//
// Memory pipeline.  U (and detJ) of the block's next cell are requested by TMA bulk copy as soon as
// the current one has been consumed.  G is consumed tile by tile in the geometry phase, by the warp
// that owns the tile, through that warp's private ring: tile j needs
// the records of its 8 columns in all n1 planes = n1 contiguous 384-byte pieces; the warp copies them
// with coalesced 16-byte cp.async (LDGSTS; per-lane source / destination offsets are loop invariants, so a
// tile costs 7 copies + 7 address adds, a tenth of what n1 bulk-copy issues cost) into a ring of NGS
// tile slots (plane stride padded to 400 bytes: conflict-free 128-bit reads) that is refilled the moment
// a tile has been used, across cell boundaries.  W leaves by bulk store from the exchange
// buffer that is free at that point and is rewritten latest in the next cell (so the store's read of
// shared memory is never actually waited for).  Odd n1: cells are 8 bytes short of 16-byte units -> skewed copies as in
// tiled_kernels.cu (RoundStage, SHIFT).  Every byte crosses HBM once.
//
// Semantics follow the reference as coded (REF_EXACT: include/femtc_b200.h).  FULL (the x + y + z operator, round 2):
//     A2,D2 | A1,D1 ; A1 | D0,A0,A0 + full 3x3 geometry | D0,A0,A0 | A1+D1 ; A1 | A2+D2        (16 GEMMs, 6 exchange buffers)
// the partial products of the three backward chains are merged as soon as they share the rest of their path.
//
// Round 2: the geometry of the stiffness kernels arrives in whole-cell stages filled by n1 bulk copies (TcPlan::GB);
// the cp.async rings described above remain as variant 30 for A/B.
#include "common.cuh"

#include <map>
#include <mutex>
#include "tma.cuh"

#include <algorithm>

namespace femtc
{
long long opt_variant();
long long opt_tc();

namespace
{
constexpr int TC_MASS = 0, TC_STIFF = 1, TC_STIFF_FULL = 2; // FULL: the x + y + z operator (include/femtc_b200.h)

__device__ __forceinline__ void tc_cp_async16(uint32_t dst, const void* src)
{
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void tc_cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tc_cp_wait()
{
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

__device__ __forceinline__ void dmma(double (&c)[2], double a, double b)
{
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c[0]), "+d"(c[1])
               : "d"(a), "d"(b));
}

template <int N1>
struct TcTab
{
  double a[N1 * N1];
  double d[N1 * N1];
};

// Which two contraction indices p the lane (g, t) feeds to the two DMMAs of a tile.  Any split of 0..7 into two sets
// of four works as long as table and data fragments agree.  Round 1 used {t, 4 + t}; with {2t, 2t + 1} (TC_KPERM) the
// data a lane loads and the results it stores are both two NEIGHBOURING planes per t, so one plane stride serves both:
// 64-bit shared-memory accesses are served per half-warp (g = 0..3 / 4..7 x t = 0..3), conflict-free when the word
// stride between consecutive t is 8 or 24 mod 32 -- for loads AND stores that is 2 planes, i.e. S = 2 or 6 mod 8.
// (Round 1: loads 1 plane apart, stores 2 planes apart, S = 4 mod 8: stores two-way conflicted; ncu, n1 = 8 mass kernel:
// l1tex 94 %, 42 % of the shared-memory wavefronts excessive, profiles/r02_ncu_tc_mass_n1_8_before.txt.)
#ifndef TC_KPERM
#define TC_KPERM 1
#endif
__host__ __device__ constexpr int tc_k0(int t) { return TC_KPERM ? 2 * t : t; }
__host__ __device__ constexpr int tc_k1(int t) { return TC_KPERM ? 2 * t + 1 : 4 + t; }

constexpr int tc_plane_stride(int n2)
{
  int s = n2;
  if (TC_KPERM)
    while (s % 8 != 2 && s % 8 != 6)
      ++s;
  else
    while (s % 16 != 4 && s % 16 != 12)
      ++s;
  return s;
}

template <int N1, int OP, int NGS, int NW>
struct TcPlan
{
  static_assert(N1 >= 6 && N1 <= 9, "tensor-core family: n1 = 6..9");
  static constexpr int N2 = N1 * N1, ND = N2 * N1;
  static constexpr bool R8 = N1 == 9;                 // 8 + 1 split of the contraction
  static constexpr int NT = (N2 + 7) / 8;             // tiles of 8 columns, the last one overlapping
  static constexpr int S = tc_plane_stride(N2);
  static constexpr int NP = R8 ? 9 : 8;               // planes held by an exchange buffer
  static constexpr bool SKEW = (ND % 2) == 1;         // cells start on 8-byte boundaries
  static constexpr int cell_copy_bytes = ND * 8 + (SKEW ? 8 : 0);
  static constexpr int stage_bytes = (cell_copy_bytes + 8 * 8 + 127) & ~127; // + slack for padded k reads
  static constexpr int e_bytes = (NP * S * 8 + 127) & ~127;
  static_assert(NP * S >= ND + 2, "an exchange buffer also stages W");
  static constexpr int g_piece_bytes = 8 * 6 * 8;     // 8 records of one plane
  static constexpr int g_plane_pitch = 400;           // bytes; 25 16-byte units: conflict-free 128-bit reads
  static constexpr int g_slot_bytes = N1 * g_plane_pitch;
  static constexpr int off_u = 0;
  static constexpr int off_j = off_u + stage_bytes;
  static constexpr bool STIFF = OP != TC_MASS, FULL = OP == TC_STIFF_FULL;
  static constexpr int NE = FULL ? 6 : 3;             // exchange buffers
  static constexpr int off_e = off_j + (OP == TC_MASS ? stage_bytes : 0);
  static constexpr int off_g = off_e + NE * e_bytes;
  // NGS >= 100: geometry in GST = NGS - 100 block-wide stages that hold a WHOLE cell each, filled by n1 bulk copies
  // (one i0-plane = n2 records = 3 n2 16-byte units; planes are pitched to an odd number of units so that the 128-bit
  // reads of a quarter-warp -- two planes x four column pairs 6 units apart -- fall into eight different bank groups).
  // The cp.async rings cost 11.5 shared-memory wavefronts per 512-byte instruction (sectors land one by one) and were
  // 40 % of all wavefronts of the n1 = 7 kernel (profiles/r02_ncu_tc_stiffness_before.txt).
  static constexpr bool GB = NGS >= 100;
  static constexpr int GST = GB ? NGS - 100 : 0;
  static constexpr int gb_plane_bytes = N2 * 48;
  static constexpr int gb_pitch = 16 * ((3 * N2) % 2 ? 3 * N2 : 3 * N2 + 1);
  static constexpr int gb_stage_bytes = (N1 * gb_pitch + 127) & ~127;
  static_assert(!FULL || GB, "FULL reads all six geometry values from whole-cell stages");
  static constexpr int off_bar = off_g + (STIFF ? (GB ? GST * gb_stage_bytes : NW * NGS * g_slot_bytes) : 0);
  static constexpr int smem_bytes = (off_bar + 16 + 8 * GST + 127) & ~127;
  static constexpr int g_chunks = N1 * (g_piece_bytes / 16);  // 16-byte chunks of one tile's geometry
  static constexpr int g_nj = (g_chunks + 31) / 32;           // cp.async per lane and tile
};

// fragment of one 1-D table for lane (g, t): rows g, columns 4k + t (zero beyond n1), and the p = 8 column
struct MFrag
{
  double a0, a1, m8;
};
template <int N1>
__device__ __forceinline__ MFrag make_frag(const double* tab, int row, int t, bool valid)
{
  MFrag f;
  f.a0 = (valid && row < N1 && tc_k0(t) < N1) ? tab[row + tc_k0(t) * N1] : 0.0;
  f.a1 = (valid && row < N1 && tc_k1(t) < N1) ? tab[row + tc_k1(t) * N1] : 0.0;
  f.m8 = (valid && N1 == 9 && row < N1) ? tab[row + 8 * N1] : 0.0;
  return f;
}

// ---- one contraction phase over the warp's cell: NM tables applied to ONE input tensor -----------------
//   ain(p, col)          element offset of X[p][col] inside `in`
//   sout(m, i, col, v)   stores Y_m[i][col] = v into the layout the next phase reads
template <int N1, int NW, int NM, class AIN, class SOUT>
__device__ __forceinline__ void tc_phase(const double* in, const MFrag (&M)[NM], const double* const (&tab)[NM], AIN ain,
                                         SOUT sout, int g, int t, int warp, int tid)
{
  constexpr int N2 = N1 * N1, NT = (N2 + 7) / 8;
  constexpr bool R8 = N1 == 9;
#pragma unroll
  for (int k = 0; k < (NT + NW - 1) / NW; ++k)
  {
    const int tile = warp + k * NW;
    if ((k + 1) * NW > NT && tile >= NT)
      break;
    const int c0 = (8 * tile + 8 <= N2) ? 8 * tile : N2 - 8;
    const double b0 = in[ain(tc_k0(t), c0 + g)], b1 = in[ain(tc_k1(t), c0 + g)];
    double acc[NM][2];
    if constexpr (R8)
    {
      // p = 8 enters as the initial value of the accumulators (rank-1 term)
      const double x0 = in[ain(8, c0 + 2 * t)], x1 = in[ain(8, c0 + 2 * t + 1)];
#pragma unroll
      for (int m = 0; m < NM; ++m)
        acc[m][0] = M[m].m8 * x0, acc[m][1] = M[m].m8 * x1;
    }
    else
    {
#pragma unroll
      for (int m = 0; m < NM; ++m)
        acc[m][0] = acc[m][1] = 0.0;
    }
#pragma unroll
    for (int m = 0; m < NM; ++m)
    {
      dmma(acc[m], M[m].a0, b0);
      dmma(acc[m], M[m].a1, b1);
    }
    if (N1 >= 8 || g < N1)
    {
#pragma unroll
      for (int m = 0; m < NM; ++m)
      {
        sout(m, g, c0 + 2 * t, acc[m][0]);
        sout(m, g, c0 + 2 * t + 1, acc[m][1]);
      }
    }
  }
  if constexpr (R8)
  {
    // row i = 8: one thread per column, plain DFMA chain over p (table row fetched on the uniform datapath)
#pragma unroll
    for (int j = 0; j < (N2 + 32 * NW - 1) / (32 * NW); ++j)
    {
      const int col = tid + 32 * NW * j;
      if (col < N2)
      {
        double x[9];
#pragma unroll
        for (int p = 0; p < 9; ++p)
          x[p] = in[ain(p, col)];
#pragma unroll
        for (int m = 0; m < NM; ++m)
        {
          double s = tab[m][8] * x[0];
#pragma unroll
          for (int p = 1; p < 9; ++p)
            s = fma(tab[m][8 + 9 * p], x[p], s);
          sout(m, 8, col, s);
        }
      }
    }
  }
}

// Last contraction of a cell, even n1: the two DMMA operands change places, so lane (g, t) ends up with
// Y[i = 2t, 2t + 1][col = c0 + g] -- two neighbouring values of the natural W layout, stored by ONE 128-bit store whose
// quarter-warps (g = 2q, 2q + 1; t = 0..3) cover all 32 banks.  The round-1 form (lane holds two columns of one i)
// stored 64-bit values 16 doubles apart: four-way conflicts, 96 of the 160 excessive wavefronts per cell of the n1 = 8
// mass kernel (profiles/r02_ncu_tc_mass_n1_8.txt).  TWO: Y = M0 . X0 + M1 . X1 (FULL).
template <int N1, int NW, bool TWO, class AIN>
__device__ __forceinline__ void tc_phase_w(const double* in0, const double* in1, const MFrag& M0, const MFrag& M1, double* w,
                                           AIN ain, int g, int t, int warp)
{
  static_assert(N1 % 2 == 0, "natural rows of an odd n1 are not 16-byte aligned");
  constexpr int N2 = N1 * N1, NT = (N2 + 7) / 8;
#pragma unroll
  for (int k = 0; k < (NT + NW - 1) / NW; ++k)
  {
    const int tile = warp + k * NW;
    if ((k + 1) * NW > NT && tile >= NT)
      break;
    const int c0 = (8 * tile + 8 <= N2) ? 8 * tile : N2 - 8;
    double acc[2] = {0.0, 0.0};
    dmma(acc, in0[ain(tc_k0(t), c0 + g)], M0.a0);
    dmma(acc, in0[ain(tc_k1(t), c0 + g)], M0.a1);
    if constexpr (TWO)
    {
      dmma(acc, in1[ain(tc_k0(t), c0 + g)], M1.a0);
      dmma(acc, in1[ain(tc_k1(t), c0 + g)], M1.a1);
    }
    if (N1 >= 8 || 2 * t < N1)
      *reinterpret_cast<double2*>(w + (c0 + g) * N1 + 2 * t) = make_double2(acc[0], acc[1]);
  }
}

// Y = M0 . X0 + M1 . X1: two inputs in the same layout accumulated into one output (the backward sweeps of the
// FULL operator, where two partial products meet before the next axis).
template <int N1, int NW, class AIN, class SOUT>
__device__ __forceinline__ void tc_phase_sum(const double* in0, const double* in1, const MFrag& M0, const MFrag& M1,
                                             const double* tab0, const double* tab1, AIN ain, SOUT sout, int g, int t,
                                             int warp, int tid)
{
  constexpr int N2 = N1 * N1, NT = (N2 + 7) / 8;
  constexpr bool R8 = N1 == 9;
#pragma unroll
  for (int k = 0; k < (NT + NW - 1) / NW; ++k)
  {
    const int tile = warp + k * NW;
    if ((k + 1) * NW > NT && tile >= NT)
      break;
    const int c0 = (8 * tile + 8 <= N2) ? 8 * tile : N2 - 8;
    const double x0 = in0[ain(tc_k0(t), c0 + g)], x1 = in0[ain(tc_k1(t), c0 + g)];
    const double y0 = in1[ain(tc_k0(t), c0 + g)], y1 = in1[ain(tc_k1(t), c0 + g)];
    double acc[2] = {0.0, 0.0};
    if constexpr (R8)
    {
      acc[0] = fma(M1.m8, in1[ain(8, c0 + 2 * t)], M0.m8 * in0[ain(8, c0 + 2 * t)]);
      acc[1] = fma(M1.m8, in1[ain(8, c0 + 2 * t + 1)], M0.m8 * in0[ain(8, c0 + 2 * t + 1)]);
    }
    dmma(acc, M0.a0, x0);
    dmma(acc, M0.a1, x1);
    dmma(acc, M1.a0, y0);
    dmma(acc, M1.a1, y1);
    if (N1 >= 8 || g < N1)
    {
      sout(g, c0 + 2 * t, acc[0]);
      sout(g, c0 + 2 * t + 1, acc[1]);
    }
  }
  if constexpr (R8)
  {
#pragma unroll
    for (int j = 0; j < (N2 + 32 * NW - 1) / (32 * NW); ++j)
    {
      const int col = tid + 32 * NW * j;
      if (col < N2)
      {
        double s = tab0[8] * in0[ain(0, col)];
#pragma unroll
        for (int p = 1; p < 9; ++p)
          s = fma(tab0[8 + 9 * p], in0[ain(p, col)], s);
#pragma unroll
        for (int p = 0; p < 9; ++p)
          s = fma(tab1[8 + 9 * p], in1[ain(p, col)], s);
        sout(8, col, s);
      }
    }
  }
}

template <int N1, int OP, int NGS, int NW>
__global__ void __launch_bounds__(32 * NW)
tc_kernel(const __grid_constant__ TcTab<N1> B, const double* __restrict__ U, double* __restrict__ W,
          const double* __restrict__ Gm, long long ncells)
{
  using PL = TcPlan<N1, OP, NGS, NW>;
  constexpr int N2 = PL::N2, ND = PL::ND, S = PL::S, NT = PL::NT;
  constexpr bool R8 = PL::R8, STIFF = PL::STIFF, FULL = PL::FULL;
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  auto block_sync = [&]()
  {
    if constexpr (NW == 1)
      __syncwarp();
    else
      __syncthreads();
  };
  double* uS = reinterpret_cast<double*>(smem + PL::off_u);
  double* jS = reinterpret_cast<double*>(smem + PL::off_j); // mass only
  double* E0 = reinterpret_cast<double*>(smem + PL::off_e);
  double* E1 = reinterpret_cast<double*>(smem + PL::off_e + PL::e_bytes);
  double* E2 = reinterpret_cast<double*>(smem + PL::off_e + 2 * PL::e_bytes);
  double* E3 = reinterpret_cast<double*>(smem + PL::off_e + (FULL ? 3 : 0) * PL::e_bytes); // FULL only
  double* E4 = reinterpret_cast<double*>(smem + PL::off_e + (FULL ? 4 : 0) * PL::e_bytes);
  double* E5 = reinterpret_cast<double*>(smem + PL::off_e + (FULL ? 5 : 0) * PL::e_bytes);
  unsigned char* gRing = smem + PL::off_g + warp * NGS * PL::g_slot_bytes; // this warp's ring
  const uint32_t barU = smem_u32(smem + PL::off_bar), barJ = barU + 8;

  // zero everything once: padded planes / k-steps are multiplied by zero table entries and must be finite
  for (int i = tid; i < PL::off_bar / 16; i += 32 * NW)
    reinterpret_cast<float4*>(smem)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  const uint32_t barG = barJ + 8; // whole-cell geometry stages (PL::GB)
  if (tid == 0)
  {
    mbar_init(barU, 1);
    mbar_init(barJ, 1);
    for (int i = 0; i < PL::GST; ++i)
      mbar_init(barG + 8 * i, 1);
    fence_mbar_init();
  }
  fence_proxy_async(); // the zero fill (generic proxy) is ordered before the first bulk copies (async proxy)
  block_sync();

  // table fragments of lane (g, t)
  MFrag FA[1] = {make_frag<N1>(B.a, g, t, true)};
  MFrag FD[1] = {make_frag<N1>(B.d, g, t, true)};
  MFrag FAD[2] = {FA[0], FD[0]};
  // second m-tile (n1 = 9, geometry phase): only row 8 is real and it belongs to the lanes with g == 0
  const MFrag FA2 = make_frag<N1>(B.a, 8, t, R8 && g == 0), FD2 = make_frag<N1>(B.d, 8, t, R8 && g == 0);
  const double* const tabA[1] = {B.a};
  const double* const tabD[1] = {B.d};
  const double* const tabAD[2] = {B.a, B.d};

  const long long stride = gridDim.x;
  long long cell = blockIdx.x;

  auto load_cell = [&](const double* field, double* dst, uint32_t bar, long long c)
  {
    if (tid == 0)
    {
      const int sk = PL::SKEW ? (int)(c & 1) : 0;
      mbar_expect_tx(bar, PL::cell_copy_bytes);
      bulk_g2s(smem_u32(dst), field + (size_t)c * ND - sk, PL::cell_copy_bytes, bar);
    }
  };
  // geometry of one tile: chunk k = lane + 32 j of its n1 x 384 bytes; plane i0 = k / 24, 16-byte unit k % 24
  int g_src[PL::g_nj], g_dst[PL::g_nj];
#pragma unroll
  for (int j = 0; j < PL::g_nj; ++j)
  {
    const int k = min(lane + 32 * j, PL::g_chunks - 1); // surplus lanes repeat the last chunk
    const int i0 = k / 24, w = k - 24 * i0;
    g_src[j] = i0 * N2 * 48 + w * 16;
    g_dst[j] = i0 * PL::g_plane_pitch + w * 16;
  }
  // tiles are requested in the order this warp consumes them (tiles warp, warp + NW, ... of the block's
  // cells): (gi_cell, gi_tile) is the next one
  long long gi_cell = cell;
  int gi_tile = warp, gi_slot = 0;
  auto load_tile_g = [&]()
  {
    if (gi_cell < ncells)
    {
      const int c0 = (8 * gi_tile + 8 <= N2) ? 8 * gi_tile : N2 - 8;
      const unsigned char* src = reinterpret_cast<const unsigned char*>(Gm + ((size_t)gi_cell * ND + c0) * 6);
      const uint32_t dst = smem_u32(gRing + gi_slot * PL::g_slot_bytes);
#pragma unroll
      for (int j = 0; j < PL::g_nj; ++j)
        tc_cp_async16(dst + g_dst[j], src + g_src[j]);
      if ((gi_tile += NW) >= NT)
        gi_tile = warp, gi_cell += stride;
      gi_slot = (gi_slot + 1 == NGS) ? 0 : gi_slot + 1;
    }
    tc_cp_commit(); // one group per tile, also when nothing is left to request: keeps the group count uniform
  };

  // whole-cell geometry stage: n1 bulk copies, one per i0-plane, completing on the stage's barrier
  auto load_cell_g = [&](long long c, int stage)
  {
    if (tid == 0)
    {
      mbar_expect_tx(barG + 8 * stage, N1 * PL::gb_plane_bytes);
      const unsigned char* src = reinterpret_cast<const unsigned char*>(Gm + (size_t)c * ND * 6);
      const uint32_t dst = smem_u32(smem + PL::off_g + stage * PL::gb_stage_bytes);
#pragma unroll
      for (int i0 = 0; i0 < N1; ++i0)
        bulk_g2s(dst + i0 * PL::gb_pitch, src + (size_t)i0 * PL::gb_plane_bytes, PL::gb_plane_bytes, barG + 8 * stage);
    }
  };
  int gstage = 0;
  uint32_t gphase = 0;

  if (cell < ncells)
  {
    load_cell(U, uS, barU, cell);
    if constexpr (OP == TC_MASS)
      load_cell(Gm, jS, barJ, cell);
  }
  if constexpr (STIFF && PL::GB)
  {
    for (int i = 0; i < PL::GST; ++i)
      if (cell + i * stride < ncells)
        load_cell_g(cell + i * stride, i);
  }
  int slot = 0; // ring slot of the next tile to consume
  if constexpr (STIFF && !PL::GB)
  {
#pragma unroll
    for (int i = 0; i < NGS - 1; ++i)
      load_tile_g();
  }

  uint32_t phase = 0;
  for (; cell < ncells; cell += stride, phase ^= 1)
  {
    const long long cn = cell + stride;
    const int sk = PL::SKEW ? (int)(cell & 1) : 0;

    // ---- phase A: axis 2.  t2 = A2 u -> E0 as [a1][a0, i2]  (stiffness.hpp:68, mass.hpp:38 order is immaterial) --
    mbar_wait(barU, phase);
    if constexpr (FULL)
      tc_phase<N1, NW, 2>(                                   // A2 u -> E0, D2 u -> E3
          uS + sk, FAD, tabAD, [](int p, int col) { return col * N1 + p; },
          [&](int m, int i, int col, double v)
          {
            const int a0 = col / N1, a1 = col - a0 * N1;
            (m == 0 ? E0 : E3)[a1 * S + a0 * N1 + i] = v;
          },
          g, t, warp, tid);
    else
      tc_phase<N1, NW, 1>(
          uS + sk, FA, tabA, [](int p, int col) { return col * N1 + p; },
          [&](int, int i, int col, double v)
          {
            const int a0 = col / N1, a1 = col - a0 * N1;
            E0[a1 * S + a0 * N1 + i] = v;
          },
          g, t, warp, tid);
    if (tid == 0)
      bulk_wait_read0(); // the previous cell's W store has finished reading its staging (E1 / E2), rewritten below
    block_sync();
    if (cn < ncells)
      load_cell(U, uS, barU, cn);

    // ---- phase B: axis 1.  stiffness: ta = A1 t2 -> E1, td = D1 t2 -> E2; mass: A1 t2 -> E1; layout [a0][i1, a2] ----
    if constexpr (STIFF)
    {
      tc_phase<N1, NW, 2>(
          E0, FAD, tabAD, [](int p, int col) { return p * S + col; },
          [&](int m, int i, int col, double v)
          {
            const int a0 = col / N1, a2 = col - a0 * N1;
            (m == 0 ? E1 : E2)[a0 * S + i * N1 + a2] = v;
          },
          g, t, warp, tid);
      if constexpr (FULL)
        tc_phase<N1, NW, 1>(                                 // A1 (D2 u) -> E4
            E3, FA, tabA, [](int p, int col) { return p * S + col; },
            [&](int, int i, int col, double v)
            {
              const int a0 = col / N1, a2 = col - a0 * N1;
              E4[a0 * S + i * N1 + a2] = v;
            },
            g, t, warp, tid);
    }
    else
      tc_phase<N1, NW, 1>(
          E0, FA, tabA, [](int p, int col) { return p * S + col; },
          [&](int, int i, int col, double v)
          {
            const int a0 = col / N1, a2 = col - a0 * N1;
            E1[a0 * S + i * N1 + a2] = v;
          },
          g, t, warp, tid);
    block_sync();

    // ---- phase C: axis 0 + the quadrature-point stage on the accumulators -------------------------------------------
    // fragment (g, t) of a tile holds the points (i0 = g, col = c0 + 2t, c0 + 2t + 1), col = (a1, a2) flattened
    if constexpr (OP == TC_MASS)
      mbar_wait(barJ, phase);
    if constexpr (STIFF && PL::GB)
      mbar_wait(barG + 8 * gstage, gphase);
    {
      const double* jc = jS + sk; // mass: detJ in natural order
#pragma unroll
      for (int k = 0; k < (NT + NW - 1) / NW; ++k)
      {
        const int tile = warp + k * NW;
        if ((k + 1) * NW > NT && tile >= NT)
          break;
        const int c0 = (8 * tile + 8 <= N2) ? 8 * tile : N2 - 8;
        const int cp = c0 + 2 * t;
        if constexpr (OP == TC_MASS)
        {
          const double b0 = E1[tc_k0(t) * S + c0 + g], b1 = E1[tc_k1(t) * S + c0 + g];
          double v[2] = {0.0, 0.0}, v2[2] = {0.0, 0.0};
          if constexpr (R8)
          {
            const double x0 = E1[8 * S + cp], x1 = E1[8 * S + cp + 1];
            v[0] = FA[0].m8 * x0, v[1] = FA[0].m8 * x1;
            v2[0] = FA2.m8 * x0, v2[1] = FA2.m8 * x1;
          }
          dmma(v, FA[0].a0, b0);
          dmma(v, FA[0].a1, b1);
          if (N1 >= 8 || g < N1)
          {
            E2[g * S + cp] = v[0] * jc[g * N2 + cp];           // mass.hpp:40-41
            E2[g * S + cp + 1] = v[1] * jc[g * N2 + cp + 1];
          }
          if constexpr (R8)
          {
            dmma(v2, FA2.a0, b0);
            dmma(v2, FA2.a1, b1);
            if (g == 0)
            {
              E2[8 * S + cp] = v2[0] * jc[8 * N2 + cp];
              E2[8 * S + cp + 1] = v2[1] * jc[8 * N2 + cp + 1];
            }
          }
        }
        else
        {
          const double ba0 = E1[tc_k0(t) * S + c0 + g], ba1 = E1[tc_k1(t) * S + c0 + g];
          const double bd0 = E2[tc_k0(t) * S + c0 + g], bd1 = E2[tc_k1(t) * S + c0 + g];
          // the z term: REF_EXACT uses A0 A1 A2 u (stiffness.hpp:76-77 as coded), FULL the real A0 A1 D2 u
          const double bz0 = FULL ? E4[tc_k0(t) * S + c0 + g] : ba0, bz1 = FULL ? E4[tc_k1(t) * S + c0 + g] : ba1;
          double wx[2] = {0.0, 0.0}, wy[2] = {0.0, 0.0}, wz[2] = {0.0, 0.0};
          double wx2[2] = {0.0, 0.0}, wy2[2] = {0.0, 0.0}, wz2[2] = {0.0, 0.0};
          if constexpr (R8)
          {
            const double xa0 = E1[8 * S + cp], xa1 = E1[8 * S + cp + 1];
            const double xd0 = E2[8 * S + cp], xd1 = E2[8 * S + cp + 1];
            const double xz0 = FULL ? E4[8 * S + cp] : xa0, xz1 = FULL ? E4[8 * S + cp + 1] : xa1;
            wx[0] = FD[0].m8 * xa0, wx[1] = FD[0].m8 * xa1;
            wz[0] = FA[0].m8 * xz0, wz[1] = FA[0].m8 * xz1;
            wy[0] = FA[0].m8 * xd0, wy[1] = FA[0].m8 * xd1;
            wx2[0] = FD2.m8 * xa0, wx2[1] = FD2.m8 * xa1;
            wz2[0] = FA2.m8 * xz0, wz2[1] = FA2.m8 * xz1;
            wy2[0] = FA2.m8 * xd0, wy2[1] = FA2.m8 * xd1;
          }
          dmma(wx, FD[0].a0, ba0);  dmma(wx, FD[0].a1, ba1);   // Wx = (D x A x A) u   stiffness.hpp:68-70
          dmma(wz, FA[0].a0, bz0);  dmma(wz, FA[0].a1, bz1);   // Wz = (A x A x A) u   :76-77 (REF_EXACT) / (A x A x D) u (FULL)
          dmma(wy, FA[0].a0, bd0);  dmma(wy, FA[0].a1, bd1);   // Wy = (A x D x A) u   :72-74
          if constexpr (R8)
          {
            dmma(wx2, FD2.a0, ba0);  dmma(wx2, FD2.a1, ba1);
            dmma(wz2, FA2.a0, bz0);  dmma(wz2, FA2.a1, bz1);
            dmma(wy2, FA2.a0, bd0);  dmma(wy2, FA2.a1, bd1);
          }
          // geometry (stiffness.hpp:14-29); only f2 is observable in REF_EXACT
          const unsigned char* sl;
          int pitch;
          if constexpr (PL::GB)
          {
            sl = smem + PL::off_g + gstage * PL::gb_stage_bytes + c0 * 48; // the tile's first record of plane 0
            pitch = PL::gb_pitch;
          }
          else
          {
            load_tile_g();              // the tile NGS - 1 ahead, into the slot freed by the previous tile
            tc_cp_wait<(PL::GB ? 0 : NGS - 1)>(); // this lane's chunks of the current tile have landed ...
            __syncwarp();               // ... and so have everybody else's
            sl = gRing + slot * PL::g_slot_bytes;
            pitch = PL::g_plane_pitch;
          }
          // records of the points (plane, cp) and (plane, cp + 1): [zz, yz, xz, yy, xy, xx] each
          auto transform = [&](int plane, const double (&vx)[2], const double (&vy)[2], const double (&vz)[2])
          {
            const double* r = reinterpret_cast<const double*>(sl + plane * pitch) + 12 * t;
            const double2 p0 = *reinterpret_cast<const double2*>(r);      // zz yz
            const double2 p1 = *reinterpret_cast<const double2*>(r + 2);  // xz yy
            const double2 p3 = *reinterpret_cast<const double2*>(r + 6);
            const double2 p4 = *reinterpret_cast<const double2*>(r + 8);
            const double f20 = fma(p0.x, vz[0], fma(p0.y, vy[0], p1.x * vx[0]));
            const double f21 = fma(p3.x, vz[1], fma(p3.y, vy[1], p4.x * vx[1]));
            auto put = [&](double* E, double a, double b)
            {
              if ((c0 & 1) == 0)
                *reinterpret_cast<double2*>(E + plane * S + cp) = make_double2(a, b);
              else
                E[plane * S + cp] = a, E[plane * S + cp + 1] = b; // the overlapping last tile starts on an odd column
            };
            put(E0, f20, f21);
            if constexpr (FULL)
            {
              const double2 p2 = *reinterpret_cast<const double2*>(r + 4);  // xy xx
              const double2 p5 = *reinterpret_cast<const double2*>(r + 10);
              // f0 = xx wx + xy wy + xz wz,  f1 = xy wx + yy wy + yz wz   (tiled_kernels.cu, same expression order)
              put(E3, fma(p1.x, vz[0], fma(p2.x, vy[0], p2.y * vx[0])), fma(p4.x, vz[1], fma(p5.x, vy[1], p5.y * vx[1])));
              put(E5, fma(p0.y, vz[0], fma(p1.y, vy[0], p2.x * vx[0])), fma(p3.y, vz[1], fma(p4.y, vy[1], p5.x * vx[1])));
            }
          };
          if (N1 >= 8 || g < N1)
            transform(g, wx, wy, wz);
          if constexpr (R8)
            if (g == 0)
              transform(8, wx2, wy2, wz2);
          if constexpr (!PL::GB)
          {
            __syncwarp(); // every lane is done with this slot before the next tile's request overwrites it
            slot = (slot + 1 == NGS) ? 0 : slot + 1;
          }
        }
      }
    }
    block_sync();
    if constexpr (OP == TC_MASS)
      if (cn < ncells)
        load_cell(Gm, jS, barJ, cn);
    if constexpr (STIFF && PL::GB)
    {
      // every geometry value of this stage went into an E0 store before the barrier above (tma.cuh, rule 1): refill
      if (cell + PL::GST * stride < ncells)
        load_cell_g(cell + PL::GST * stride, gstage);
      if (++gstage == PL::GST)
        gstage = 0, gphase ^= 1;
    }

    // ---- backward: axis 0, axis 1, axis 2 ----------------------------------------------------------------------------------
    double* wS; // W staging, natural order (+ skew)
    if constexpr (FULL)
    {
      // W = (D x A x A) f0 + (A x D x A) f1 + (A x A x D) f2, partial products merged as soon as they share the rest of
      // their path:  axis 0: X = D0 f0, Y = A0 f1, Z = A0 f2 | axis 1: X' = A1 X + D1 Y, Z' = A1 Z | axis 2: A2 X' + D2 Z'
      auto to_b = [&](double* E)
      {
        return [E](int, int i, int col, double v)
        {
          const int a1 = col / N1, a2 = col - a1 * N1;
          E[a1 * S + i * N1 + a2] = v; // [a1][j0, a2]
        };
      };
      tc_phase<N1, NW, 1>(E3, FD, tabD, [](int p, int col) { return p * S + col; }, to_b(E1), g, t, warp, tid);
      tc_phase<N1, NW, 1>(E5, FA, tabA, [](int p, int col) { return p * S + col; }, to_b(E2), g, t, warp, tid);
      tc_phase<N1, NW, 1>(E0, FA, tabA, [](int p, int col) { return p * S + col; }, to_b(E4), g, t, warp, tid);
      block_sync();
      tc_phase_sum<N1, NW>(
          E1, E2, FA[0], FD[0], B.a, B.d, [](int p, int col) { return p * S + col; },
          [&](int i, int col, double v)
          {
            const int j0 = col / N1, a2 = col - j0 * N1;
            E0[a2 * S + j0 * N1 + i] = v; // [a2][j0, j1]
          },
          g, t, warp, tid);
      tc_phase<N1, NW, 1>(
          E4, FA, tabA, [](int p, int col) { return p * S + col; },
          [&](int, int i, int col, double v)
          {
            const int j0 = col / N1, a2 = col - j0 * N1;
            E3[a2 * S + j0 * N1 + i] = v;
          },
          g, t, warp, tid);
      block_sync();
      wS = E1;
      if constexpr (N1 % 2 == 0)
        tc_phase_w<N1, NW, true>(E0, E3, FA[0], FD[0], wS, [](int p, int col) { return p * S + col; }, g, t, warp);
      else
        tc_phase_sum<N1, NW>(
            E0, E3, FA[0], FD[0], B.a, B.d, [](int p, int col) { return p * S + col; },
            [&](int i, int col, double v) { wS[sk + col * N1 + i] = v; }, g, t, warp, tid);
    }
    else if constexpr (STIFF)
    {
      // only W = (A x A x D) f2 is observable (stiffness.hpp:95-97 overwrites :83-91)
      tc_phase<N1, NW, 1>(
          E0, FA, tabA, [](int p, int col) { return p * S + col; },
          [&](int, int i, int col, double v)
          {
            const int a1 = col / N1, a2 = col - a1 * N1;
            E1[a1 * S + i * N1 + a2] = v; // [a1][j0, a2]
          },
          g, t, warp, tid);
      block_sync();
      tc_phase<N1, NW, 1>(
          E1, FA, tabA, [](int p, int col) { return p * S + col; },
          [&](int, int i, int col, double v)
          {
            const int j0 = col / N1, a2 = col - j0 * N1;
            E2[a2 * S + j0 * N1 + i] = v; // [a2][j0, j1]
          },
          g, t, warp, tid);
      block_sync();
      wS = E1;
      if constexpr (N1 % 2 == 0)
        tc_phase_w<N1, NW, false>(E2, E2, FD[0], FD[0], wS, [](int p, int col) { return p * S + col; }, g, t, warp);
      else
        tc_phase<N1, NW, 1>(
            E2, FD, tabD, [](int p, int col) { return p * S + col; },
            [&](int, int i, int col, double v) { wS[sk + col * N1 + i] = v; }, g, t, warp, tid);
    }
    else
    {
      tc_phase<N1, NW, 1>(
          E2, FA, tabA, [](int p, int col) { return p * S + col; },
          [&](int, int i, int col, double v)
          {
            const int a1 = col / N1, a2 = col - a1 * N1;
            E0[a1 * S + i * N1 + a2] = v;
          },
          g, t, warp, tid);
      block_sync();
      tc_phase<N1, NW, 1>(
          E0, FA, tabA, [](int p, int col) { return p * S + col; },
          [&](int, int i, int col, double v)
          {
            const int j0 = col / N1, a2 = col - j0 * N1;
            E1[a2 * S + j0 * N1 + i] = v;
          },
          g, t, warp, tid);
      block_sync();
      wS = E2;
      if constexpr (N1 % 2 == 0)
        tc_phase_w<N1, NW, false>(E1, E1, FA[0], FA[0], wS, [](int p, int col) { return p * S + col; }, g, t, warp);
      else
        tc_phase<N1, NW, 1>(
            E1, FA, tabA, [](int p, int col) { return p * S + col; },
            [&](int, int i, int col, double v) { wS[sk + col * N1 + i] = v; }, g, t, warp, tid);
    }
    fence_proxy_async();
    block_sync();
    if (tid == 0)
    {
      double* dst = W + (size_t)cell * ND;
      if constexpr (PL::SKEW)
      {
        // 16-byte aligned interior by bulk copy, the remaining value (first or last) by a plain store
        bulk_s2g(dst + sk, smem_u32(wS + 2 * sk), (ND - 1) * 8);
        bulk_commit();
        const int e = sk ? 0 : ND - 1;
        dst[e] = wS[sk + e];
      }
      else
      {
        bulk_s2g(dst, smem_u32(wS), ND * 8);
        bulk_commit();
      }
    }
  }
  tc_cp_wait<0>();
  if (tid == 0)
    bulk_wait_read0();
}

// ---- launch -----------------------------------------------------------------------------------------------
inline bool al16t(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <int N1, int OP, int NGS, int NW>
long long run_tc(const double* A, const double* D, const double* U, double* W, const double* geom, size_t num_cells,
                 cudaStream_t stream)
{
  using PL = TcPlan<N1, OP, NGS, NW>;
  long long ncells = (long long)num_cells;
  if (PL::SKEW)
    ncells &= ~1LL; // the last cell must be an odd one: skewed copies never read past a field
  if (ncells == 0 || !al16t(U) || !al16t(W) || !al16t(geom))
    return 0;
  TcTab<N1> B;
  for (int i = 0; i < N1 * N1; ++i)
  {
    B.a[i] = A[i];
    B.d[i] = D[i];
  }
  auto kernel = tc_kernel<N1, OP, NGS, NW>;
  // the plan of a kernel on a device never changes: attributes and the occupancy query are paid on the first launch only
  static std::mutex mu;
  static std::map<int, int> plans; // device -> resident blocks per SM
  int dev = 0, bps = 0;
  FEMTC_CUDA(cudaGetDevice(&dev));
  {
    std::lock_guard<std::mutex> lock(mu);
    const auto it = plans.find(dev);
    if (it != plans.end())
      bps = it->second;
    else
    {
      int max_optin = 0;
      FEMTC_CUDA(cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
      if (max_optin < PL::smem_bytes)
      {
        set_error("tensor-core kernel does not fit on an SM");
        return -FEMTC_ERR_CUDA;
      }
      if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PL::smem_bytes) != cudaSuccess ||
          cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100) != cudaSuccess ||
          cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kernel, 32 * NW, PL::smem_bytes) != cudaSuccess || bps < 1)
      {
        (void)cudaGetLastError();
        set_error("tensor-core kernel: launch planning failed");
        return -FEMTC_ERR_CUDA;
      }
      plans[dev] = bps;
    }
  }
  const int grid = (int)std::min<long long>(ncells, (long long)sm_count() * bps);
  FEMTC_LAUNCH(FAMILY_TC, NW + 10 * NGS, 1, bps, kernel, grid, 32 * NW, PL::smem_bytes, stream, B, U, W, geom, ncells);
  return ncells;
}

// NGS = geometry ring slots per warp, NW = warps per cell.  Defaults are the measured best
// (profiles/r01_tc_sweep.txt); `variant` (development knob) picks alternatives.
template <int OP>
long long run_tc_n1(int n1, const double* A, const double* D, const double* U, double* W, const double* geom,
                    size_t num_cells, cudaStream_t stream)
{
  const long long v = opt_variant();
  if constexpr (OP == TC_STIFF_FULL)
  {
    // whole-cell geometry stages only; variant 31 / 33: the other stage count / more warps per cell
    switch (n1)
    {
    case 6: return v == 31 ? run_tc<6, OP, 101, 5>(A, D, U, W, geom, num_cells, stream)
                 : v == 33 ? run_tc<6, OP, 102, 2>(A, D, U, W, geom, num_cells, stream)
                           : run_tc<6, OP, 102, 5>(A, D, U, W, geom, num_cells, stream);
    case 7: return v == 31 ? run_tc<7, OP, 101, 4>(A, D, U, W, geom, num_cells, stream)
                 : v == 33 ? run_tc<7, OP, 102, 7>(A, D, U, W, geom, num_cells, stream)
                           : run_tc<7, OP, 102, 4>(A, D, U, W, geom, num_cells, stream);
    case 8: return v == 31 ? run_tc<8, OP, 102, 4>(A, D, U, W, geom, num_cells, stream)
                 : v == 33 ? run_tc<8, OP, 101, 8>(A, D, U, W, geom, num_cells, stream)
                           : run_tc<8, OP, 101, 4>(A, D, U, W, geom, num_cells, stream);
    case 9: return v == 31 ? run_tc<9, OP, 102, 4>(A, D, U, W, geom, num_cells, stream)
                 : v == 33 ? run_tc<9, OP, 101, 4>(A, D, U, W, geom, num_cells, stream)
                           : run_tc<9, OP, 101, 6>(A, D, U, W, geom, num_cells, stream);
    default: return 0;
    }
  }
  else if constexpr (OP == TC_MASS)
  {
    switch (n1)
    {
    case 6: return run_tc<6, OP, 2, 2>(A, D, U, W, geom, num_cells, stream);
    case 7: return run_tc<7, OP, 2, 2>(A, D, U, W, geom, num_cells, stream);
    case 8: return v == 1 ? run_tc<8, OP, 2, 4>(A, D, U, W, geom, num_cells, stream)
                          : run_tc<8, OP, 2, 1>(A, D, U, W, geom, num_cells, stream);
    case 9: return run_tc<9, OP, 2, 4>(A, D, U, W, geom, num_cells, stream);
    default: return 0;
    }
  }
  else
  {
    switch (n1)
    {
    case 6: return v == 31 ? run_tc<6, OP, 102, 2>(A, D, U, W, geom, num_cells, stream)
                 : v == 30 ? run_tc<6, OP, 3, 2>(A, D, U, W, geom, num_cells, stream)
                           : run_tc<6, OP, 102, 5>(A, D, U, W, geom, num_cells, stream);
    // Geometry transport (profiles/r02_ab_tc_geometry_stages.txt; % of measured HBM, burst): whole-cell stages filled by
    // bulk copies replace the per-warp cp.async rings of round 1 (variant 30):  n1 = 7  74.5 -> 87.2 (2 stages; 3: 81.0),
    // n1 = 8  92.4 -> 102.2 (2),  n1 = 9  67.2 -> 74.2 (1 stage; 2 stages cost a resident block: 66.4).
    case 7: return v == 1   ? run_tc<7, OP, 2, 2>(A, D, U, W, geom, num_cells, stream)
                 : v == 30 ? run_tc<7, OP, 3, 4>(A, D, U, W, geom, num_cells, stream)
                 : v == 31 ? run_tc<7, OP, 103, 4>(A, D, U, W, geom, num_cells, stream)
                 : v == 33 ? run_tc<7, OP, 102, 7>(A, D, U, W, geom, num_cells, stream)
                           : run_tc<7, OP, 102, 4>(A, D, U, W, geom, num_cells, stream);
    case 8: return v == 1   ? run_tc<8, OP, 2, 2>(A, D, U, W, geom, num_cells, stream)
                 : v == 30 ? run_tc<8, OP, 3, 4>(A, D, U, W, geom, num_cells, stream)
                 : v == 33 ? run_tc<8, OP, 102, 8>(A, D, U, W, geom, num_cells, stream)
                           : run_tc<8, OP, 102, 4>(A, D, U, W, geom, num_cells, stream);
    case 9: return v == 1   ? run_tc<9, OP, 3, 4>(A, D, U, W, geom, num_cells, stream)
                 : v == 30 ? run_tc<9, OP, 2, 4>(A, D, U, W, geom, num_cells, stream)
                 : v == 31 ? run_tc<9, OP, 102, 4>(A, D, U, W, geom, num_cells, stream)
                 : v == 33 ? run_tc<9, OP, 101, 6>(A, D, U, W, geom, num_cells, stream)
                 : v == 34 ? run_tc<9, OP, 102, 6>(A, D, U, W, geom, num_cells, stream)
                           : run_tc<9, OP, 101, 4>(A, D, U, W, geom, num_cells, stream);
    default: return 0;
    }
  }
}

} // namespace

// Which shapes the tensor-core family serves: fp64, n1 = 6..9, mass and stiffness REF_EXACT.
// Option "tc": 0 = never, 1 = wherever applicable, -1 (default) = where measured faster than the column kernels.
template <typename T>
long long launch_mass_tc(const MassArgs<T>& a)
{
  if constexpr (sizeof(T) == 8)
  {
    const int n1 = a.degree + 1;
    const long long tc = opt_tc();
    if (tc == 0 || n1 < 6 || n1 > 9)
      return 0;
    if (tc < 0 && n1 != 8) // measured: only n1 = 8 beats the column kernel (71 % vs 59 % of the HBM peak)
      return 0;
    return run_tc_n1<TC_MASS>(n1, a.phi_host, a.phi_host, a.U, a.W, a.detJ, a.num_cells, a.stream);
  }
  else
    return launch_mass_tf32(a); // fp32: tcgen05 kind::tf32 arm (tf32_kernels.cu), option tc = 1
}

template <typename T>
long long launch_stiffness_tc(const StiffnessArgs<T>& a)
{
  if constexpr (sizeof(T) == 8)
  {
    const int n1 = a.degree + 1;
    const long long tc = opt_tc();
    if (tc == 0 || n1 < 6 || n1 > 9)
      return 0;
    if (a.mode == FEMTC_STIFFNESS_FULL)
    {
      // measured (profiles/r02_ab_tc_full.txt; % of measured HBM, column kernel -> DMMA): n1 = 6  57 -> 52 (stays on the
      // column kernel), n1 = 7  45 -> 59, n1 = 8  51 -> 73, n1 = 9  37 -> 41
      if (tc < 0 && n1 < 7)
        return 0;
      return run_tc_n1<TC_STIFF_FULL>(n1, a.A_host, a.D_host, a.U, a.W, a.G, a.num_cells, a.stream);
    }
    // measured against the column kernels (% of measured HBM, burst): n1 = 6  88 -> 92 (whole-cell stages, 5 warps per
    // cell), n1 = 7  69 -> 94, n1 = 8  73 -> 107, n1 = 9  68 -> 79: the family serves every REF_EXACT shape it covers
    return run_tc_n1<TC_STIFF>(n1, a.A_host, a.D_host, a.U, a.W, a.G, a.num_cells, a.stream);
  }
  return 0;
}

template long long launch_mass_tc<double>(const MassArgs<double>&);
template long long launch_mass_tc<float>(const MassArgs<float>&);
template long long launch_stiffness_tc<double>(const StiffnessArgs<double>&);
template long long launch_stiffness_tc<float>(const StiffnessArgs<float>&);

} // namespace femtc
