// Column kernels: degrees 5..8 (n1 = 6..9) and the low-degree shapes the slab kernels of
// tiled_kernels.cu do not cover (fp32 with odd n1, stiffness at n1 = 2).
//
// Work decomposition.  A block owns a *tile* of C consecutive cells and runs n1^2 threads per cell.
// In every phase a thread holds one *column* of its cell tensor X[i0][i1][i2] in registers: all
// values along the phase's register axis for fixed indices on the two thread axes.
//     phase C: register axis 0, threads (i1, i2)  -- natural order, so staged U / detJ / G / W are
//                                                     touched with consecutive lanes (coalesced)
//     phase B: register axis 1, threads (i0, i2)
//     phase A: register axis 2, threads (i0, i1)
// A contraction along the register axis is a fully unrolled FMA chain whose matrix entries are
// immediate constant-bank operands (the 1-D tables are kernel parameters).  Between phases the
// tensor is exchanged through padded shared-memory scratch whose strides make both the writing and
// the reading phase bank-conflict free (tools/gen_column_layouts.py -> column_layouts.cuh).
//     mass       C -> B -> A (x detJ) -> B -> C           6 contractions, 4 exchanges   (mass.hpp:36-47)
//     stiffness  A -> B -> C (geometry) -> C -> B -> A    REF_EXACT 9 contractions, FULL 16
//                                                          (stiffness.hpp:68-97; shared products formed once)
//
// Memory pipeline.  U (and detJ) of a tile arrive by one TMA bulk copy per tile into block staging,
// re-requested for the block's next tile right after the barrier that follows their last read.
// G arrives through *warp-private* rings: for plane i0 the records of a warp's 32 points are
// contiguous in HBM (one or two pieces when the warp straddles a cell boundary), so each warp runs
// its own cp.async.bulk + mbarrier pipeline NG planes deep, across tile boundaries, with no block
// barrier in the geometry loop.  W leaves by TMA bulk store from staging (or by coalesced stores
// after a C-order exchange where the phase-A staging access would conflict, n1 = 4, 8).
// Every byte of U, detJ / G and W crosses HBM exactly once.
//
// Semantics follow the reference as coded (REF_EXACT / FULL: include/femtc_b200.h).
#include "column_layouts.cuh"
#include "common.cuh"
#include "tma.cuh"

#include <algorithm>

namespace femtc
{
long long opt_variant();

namespace
{
constexpr int cmaxi(int a, int b) { return a > b ? a : b; }
constexpr int up16(int bytes) { return (bytes + 15) & ~15; }
constexpr int up128(int bytes) { return (bytes + 127) & ~127; }

// The 1-D tables as kernel parameters.  Column-major like the reference (A(i,p) = a[i + p*LD]) but
// with the leading dimension padded to an even count, plus the transposes (at[p + i*LD] = A(i,p)),
// so that both a column and a row of a matrix are contiguous 16-byte-aligned runs: fp64 entries
// then reach the uniform registers two per LDCU.128.
template <int N1>
constexpr int tab_ld()
{
  return (N1 + 1) & ~1;
}
template <typename T, int N1>
struct alignas(16) TabA
{
  static constexpr int LD = tab_ld<N1>();
  T a[N1 * LD];
  T at[N1 * LD];
};
template <typename T, int N1>
struct alignas(16) TabAD
{
  static constexpr int LD = tab_ld<N1>();
  T a[N1 * LD];
  T d[N1 * LD];
  T at[N1 * LD];
  T dt[N1 * LD];
};
template <typename T, int N1>
void fill_table(T* dst, const T* src_colmajor, bool transpose)
{
  constexpr int LD = tab_ld<N1>();
  for (int i = 0; i < N1 * LD; ++i)
    dst[i] = T(0);
  for (int i = 0; i < N1; ++i)
    for (int p = 0; p < N1; ++p)
      (transpose ? dst[p + i * LD] : dst[i + p * LD]) = src_colmajor[i + p * N1];
}

// ---- contractions -------------------------------------------------------------------------------------
// One side of every contraction is shared memory (the column being read, or the column being
// written), the other a register array.  The loop over the shared-memory side is what carries the
// matrix index, so it may stay rolled: the matrix entries are then fetched from the kernel-parameter
// constant bank at a *dynamic uniform* offset.  fp32 FMAs take constant-bank operands directly, so
// the loops are fully unrolled there; fp64 FMAs need the entries in uniform registers, and a fully
// unrolled kernel makes ptxas hoist whole matrices into (spilled) registers -- the rolled form keeps
// the working set to one matrix row / column.  Sums run over p ascending, starting from zero
// (0 + a*b is exact, so this equals the reference's "Ctemp = 0; Ctemp += a*b": linalg.hpp:57,128).
template <typename T, int N1>
struct Roll
{
  // matrix columns / rows handled per trip of a rolled loop: fp64 keeps at most ~27 entries
  // (54 uniform registers) in flight; fp32 (and tiny n1) is fully unrolled
  static constexpr int in = (sizeof(T) == 4 || N1 <= 4) ? N1 : (N1 % 3 == 0 ? 3 : 2);
  static constexpr int plane = (sizeof(T) == 4 || N1 <= 4) ? N1 : 1;
};

// N1 consecutive table entries starting at the even offset `off` (a 16-byte boundary): fp64 pairs
// are fetched as one 128-bit constant load
template <typename T, int N1>
__device__ __forceinline__ void tab_run(const T* M, int off, T (&m)[N1])
{
  if constexpr (sizeof(T) == 8)
  {
#pragma unroll
    for (int i = 0; i < N1; i += 2)
    {
      const double2 v = *reinterpret_cast<const double2*>(M + off + i);
      m[i] = v.x;
      if (i + 1 < N1)
        m[i + 1] = v.y;
    }
  }
  else
  {
#pragma unroll
    for (int i = 0; i < N1; ++i)
      m[i] = M[off + i];
  }
}

// y[i] += m[i] * x for a whole column.  fp32: two outputs per instruction (FFMA2, fma.rn.f32x2 -- the table pair
// {m[i], m[i+1]} is a 64-bit uniform-register operand, x is broadcast); each lane result is the same IEEE fma as
// the scalar form, so nothing changes bitwise, but the FMA instruction count of the issue-bound fp32 kernels halves
// (probe: tools/probe/ffma2_probe.cu, same 74 TFLOP/s from half the issue slots).
#ifndef FEMTC_FFMA2
#define FEMTC_FFMA2 1
#endif
// PACK = false keeps the scalar form: the one-cell kernels at n1 = 9 (FULL stiffness fp32: 161 registers at one
// resident block) spill with the packed form and lose 5 points
template <typename T, int N1, bool PACK = true>
__device__ __forceinline__ void axpy_n(const T (&m)[N1], T x, T (&y)[N1])
{
  if constexpr (sizeof(T) == 4 && FEMTC_FFMA2 && PACK)
  {
    const float2 xx = make_float2(x, x);
#pragma unroll
    for (int i = 0; i + 1 < N1; i += 2)
    {
      const float2 r = __ffma2_rn(make_float2(m[i], m[i + 1]), xx, make_float2(y[i], y[i + 1]));
      y[i] = r.x, y[i + 1] = r.y;
    }
    if constexpr (N1 & 1)
      y[N1 - 1] = fma_t(m[N1 - 1], x, y[N1 - 1]);
  }
  else
  {
#pragma unroll
    for (int i = 0; i < N1; ++i)
      y[i] = fma_t(m[i], x, y[i]);
  }
}
// {r0, r1} = sum_p m[p] * {z0[p], z1[p]}: one table row against the same column of two cells
template <typename T, int N1, bool PACK = true>
__device__ __forceinline__ void dot2_n(const T (&m)[N1], const T (&z0)[N1], const T (&z1)[N1], T& r0, T& r1)
{
  if constexpr (sizeof(T) == 4 && FEMTC_FFMA2 && PACK)
  {
    float2 acc = __fmul2_rn(make_float2(m[0], m[0]), make_float2(z0[0], z1[0]));
#pragma unroll
    for (int p = 1; p < N1; ++p)
      acc = __ffma2_rn(make_float2(m[p], m[p]), make_float2(z0[p], z1[p]), acc);
    r0 = acc.x, r1 = acc.y;
  }
  else
  {
    T a0 = m[0] * z0[0], a1 = m[0] * z1[0];
#pragma unroll
    for (int p = 1; p < N1; ++p)
    {
      a0 = fma_t(m[p], z0[p], a0);
      a1 = fma_t(m[p], z1[p], a1);
    }
    r0 = a0, r1 = a1;
  }
}

// One trip: columns p0 .. p0+CNT-1 of the tables (table offset `off` = p0*LD) applied to the values
// at ptr[0], ptr[stride], ...
template <int CNT, bool DUAL, typename T, int N1>
__device__ __forceinline__ void contract_in_trip(const T* Ma, const T* Mb, int off, const T* ptr, int stride,
                                                 T (&ya)[N1], T (&yb)[N1])
{
  constexpr int LD = tab_ld<N1>();
  T xp[CNT];
#pragma unroll
  for (int u = 0; u < CNT; ++u)
    xp[u] = ptr[u * stride];
#pragma unroll
  for (int u = 0; u < CNT; ++u)
  {
    T m[N1];
    tab_run<T, N1>(Ma, off + u * LD, m);
    axpy_n<T, N1, N1 != 9>(m, xp[u], ya);
    if constexpr (DUAL)
    {
      tab_run<T, N1>(Mb, off + u * LD, m);
      axpy_n<T, N1, N1 != 9>(m, xp[u], yb);
    }
  }
}
template <bool DUAL, typename T, int N1>
__device__ __forceinline__ void contract_in_loop(const T* Ma, const T* Mb, const T* ptr, int stride, T (&ya)[N1],
                                                 T (&yb)[N1])
{
  constexpr int LD = tab_ld<N1>();
  constexpr int U = Roll<T, N1>::in;
  constexpr int TRIPS = N1 / U, REM = N1 - TRIPS * U;
  if constexpr (TRIPS == 1)
    contract_in_trip<U, DUAL>(Ma, Mb, 0, ptr, stride, ya, yb);
  else
  {
    // the table offset is a separate, uniform induction variable; the shared-memory address is a
    // pointer increment -- so the matrix fetches stay on the uniform datapath (LDCU)
    int off = 0;
#pragma unroll 1
    for (int t = 0; t < TRIPS; ++t)
    {
      contract_in_trip<U, DUAL>(Ma, Mb, off, ptr, stride, ya, yb);
      ptr += U * stride;
      off += U * LD;
    }
  }
  if constexpr (REM > 0 && TRIPS != 1)
    contract_in_trip<REM, DUAL>(Ma, Mb, TRIPS * U * LD, ptr, stride, ya, yb);
}

// y[i] = sum_p A(i,p) * buf[base + p*stride]                  (shared memory -> registers)
template <typename T, int N1>
__device__ __forceinline__ void contract_in(const T (&M)[N1 * tab_ld<N1>()], const T* buf, int base, int stride, T (&y)[N1])
{
#pragma unroll
  for (int i = 0; i < N1; ++i)
    y[i] = T(0);
  contract_in_loop<false>(M, M, buf + base, stride, y, y);
}
// y[i] += sum_p A(i,p) * buf[base + p*stride]
template <typename T, int N1>
__device__ __forceinline__ void contract_in_add(const T (&M)[N1 * tab_ld<N1>()], const T* buf, int base, int stride, T (&y)[N1])
{
  contract_in_loop<false>(M, M, buf + base, stride, y, y);
}
// two matrices applied to the same column
template <typename T, int N1>
__device__ __forceinline__ void contract_in2(const T (&Ma)[N1 * tab_ld<N1>()], const T (&Mb)[N1 * tab_ld<N1>()], const T* buf, int base,
                                             int stride, T (&ya)[N1], T (&yb)[N1])
{
#pragma unroll
  for (int i = 0; i < N1; ++i)
    ya[i] = yb[i] = T(0);
  contract_in_loop<true>(Ma, Mb, buf + base, stride, ya, yb);
}
// row `i` of A times a register column: sum_p A(i,p) * z[p]      (MT = transposed table; off = i*LD)
template <typename T, int N1>
__device__ __forceinline__ T row_dot_off(const T* MT, int off, const T (&z)[N1])
{
  T m[N1];
  tab_run<T, N1>(MT, off, m);
  T acc = m[0] * z[0];
#pragma unroll
  for (int p = 1; p < N1; ++p)
    acc = fma_t(m[p], z[p], acc);
  return acc;
}
template <typename T, int N1>
__device__ __forceinline__ T row_dot(const T (&MT)[N1 * tab_ld<N1>()], int i, const T (&z)[N1])
{
  return row_dot_off<T, N1>(MT, i * tab_ld<N1>(), z);
}
// buf[base + i*stride] = sum_p A(i,p) * z[p]                   (registers -> shared memory)
// MT is the transposed table, MT[p + i*LD] = A(i,p): row i is a contiguous run.
template <typename T, int N1>
__device__ __forceinline__ void contract_out(const T (&MT)[N1 * tab_ld<N1>()], const T (&z)[N1], T* buf, int base, int stride)
{
  constexpr int LD = tab_ld<N1>();
  if constexpr (sizeof(T) == 4 && FEMTC_FFMA2 && N1 != 9)
  {
    // fp32: rows i, i+1 together.  MT[p + i*LD] = A(i,p); the pair {A(i,p), A(i+1,p)} is gathered from two rows of
    // the transposed table (constant-bank operands), z[p] is broadcast; lane results equal the scalar form's.
    T* ptr = buf + base;
#pragma unroll
    for (int i = 0; i + 1 < N1; i += 2)
    {
      float2 acc = __fmul2_rn(make_float2(MT[i * LD], MT[(i + 1) * LD]), make_float2(z[0], z[0]));
#pragma unroll
      for (int p = 1; p < N1; ++p)
        acc = __ffma2_rn(make_float2(MT[p + i * LD], MT[p + (i + 1) * LD]), make_float2(z[p], z[p]), acc);
      ptr[i * stride] = acc.x;
      ptr[(i + 1) * stride] = acc.y;
    }
    if constexpr (N1 & 1)
      ptr[(N1 - 1) * stride] = row_dot_off<T, N1>(MT, (N1 - 1) * LD, z);
    return;
  }
  constexpr int U = Roll<T, N1>::in;
  constexpr int TRIPS = N1 / U, REM = N1 - TRIPS * U;
  T* ptr = buf + base;
  if constexpr (TRIPS == 1)
  {
#pragma unroll
    for (int i = 0; i < N1; ++i)
      ptr[i * stride] = row_dot_off<T, N1>(MT, i * LD, z);
  }
  else
  {
    int off = 0;
#pragma unroll 1
    for (int t = 0; t < TRIPS; ++t)
    {
#pragma unroll
      for (int u = 0; u < U; ++u)
        ptr[u * stride] = row_dot_off<T, N1>(MT, off + u * LD, z);
      ptr += U * stride;
      off += U * LD;
    }
#pragma unroll
    for (int u = 0; u < REM; ++u)
      ptr[u * stride] = row_dot_off<T, N1>(MT, (TRIPS * U + u) * LD, z);
  }
}
// y[i] += A(i,p) * f for all i                                   (off = p*LD, may be dynamic)
template <typename T, int N1>
__device__ __forceinline__ void col_axpy(const T (&M)[N1 * tab_ld<N1>()], int p, T f, T (&y)[N1])
{
  T m[N1];
  tab_run<T, N1>(M, p * tab_ld<N1>(), m);
  axpy_n<T, N1, N1 != 9>(m, f, y);
}

// ---- two cells per thread -------------------------------------------------------------------------------
// An fp64 FMA cannot take a constant-bank operand: every DFMA of the contractions above is paired with the
// LDCU that brings its matrix entry into a uniform register (ncu, fp64 mass n1 = 9: LDCU 38 % = DFMA 37 % of
// the executed instructions).  A thread that owns the same column of TWO cells uses each fetched entry
// twice.  `delta` = distance between the two cells' columns in the buffer.
template <int CNT, typename T, int N1>
__device__ __forceinline__ void contract_in_trip_x2(const T* M, int off, const T* ptr, int delta, int stride, T (&y0)[N1],
                                                    T (&y1)[N1])
{
  constexpr int LD = tab_ld<N1>();
  T xa[CNT], xb[CNT];
#pragma unroll
  for (int u = 0; u < CNT; ++u)
  {
    xa[u] = ptr[u * stride];
    xb[u] = ptr[u * stride + delta];
  }
#pragma unroll
  for (int u = 0; u < CNT; ++u)
  {
    T m[N1];
    tab_run<T, N1>(M, off + u * LD, m);
    axpy_n<T, N1>(m, xa[u], y0);
    axpy_n<T, N1>(m, xb[u], y1);
  }
}
// y0[i] = sum_p A(i,p) buf[base + p*stride],  y1[i] = sum_p A(i,p) buf[base + delta + p*stride]
template <typename T, int N1>
__device__ __forceinline__ void contract_in_x2(const T (&M)[N1 * tab_ld<N1>()], const T* buf, int base, int delta, int stride,
                                               T (&y0)[N1], T (&y1)[N1])
{
  constexpr int LD = tab_ld<N1>();
  constexpr int U = Roll<T, N1>::in;
  constexpr int TRIPS = N1 / U, REM = N1 - TRIPS * U;
#pragma unroll
  for (int i = 0; i < N1; ++i)
    y0[i] = y1[i] = T(0);
  const T* ptr = buf + base;
  if constexpr (TRIPS == 1)
    contract_in_trip_x2<U>(M, 0, ptr, delta, stride, y0, y1);
  else
  {
    int off = 0;
#pragma unroll 1
    for (int t = 0; t < TRIPS; ++t)
    {
      contract_in_trip_x2<U>(M, off, ptr, delta, stride, y0, y1);
      ptr += U * stride;
      off += U * LD;
    }
  }
  if constexpr (REM > 0 && TRIPS != 1)
    contract_in_trip_x2<REM>(M, TRIPS * U * LD, ptr, delta, stride, y0, y1);
}
// two matrices applied to the same column of two cells (four accumulator columns per thread)
template <int CNT, typename T, int N1>
__device__ __forceinline__ void contract_in2_trip_x2(const T* Ma, const T* Mb, int off, const T* ptr, int delta, int stride,
                                                     T (&ya0)[N1], T (&yb0)[N1], T (&ya1)[N1], T (&yb1)[N1])
{
  constexpr int LD = tab_ld<N1>();
  T xa[CNT], xb[CNT];
#pragma unroll
  for (int u = 0; u < CNT; ++u)
  {
    xa[u] = ptr[u * stride];
    xb[u] = ptr[u * stride + delta];
  }
#pragma unroll
  for (int u = 0; u < CNT; ++u)
  {
    T m[N1];
    tab_run<T, N1>(Ma, off + u * LD, m);
    axpy_n<T, N1>(m, xa[u], ya0);
    axpy_n<T, N1>(m, xb[u], ya1);
    tab_run<T, N1>(Mb, off + u * LD, m);
    axpy_n<T, N1>(m, xa[u], yb0);
    axpy_n<T, N1>(m, xb[u], yb1);
  }
}
template <typename T, int N1>
__device__ __forceinline__ void contract_in2_x2(const T (&Ma)[N1 * tab_ld<N1>()], const T (&Mb)[N1 * tab_ld<N1>()], const T* buf,
                                                int base, int delta, int stride, T (&ya0)[N1], T (&yb0)[N1], T (&ya1)[N1],
                                                T (&yb1)[N1])
{
  constexpr int LD = tab_ld<N1>();
  constexpr int U = Roll<T, N1>::in;
  constexpr int TRIPS = N1 / U, REM = N1 - TRIPS * U;
#pragma unroll
  for (int i = 0; i < N1; ++i)
    ya0[i] = yb0[i] = ya1[i] = yb1[i] = T(0);
  const T* ptr = buf + base;
  if constexpr (TRIPS == 1)
    contract_in2_trip_x2<U>(Ma, Mb, 0, ptr, delta, stride, ya0, yb0, ya1, yb1);
  else
  {
    int off = 0;
#pragma unroll 1
    for (int t = 0; t < TRIPS; ++t)
    {
      contract_in2_trip_x2<U>(Ma, Mb, off, ptr, delta, stride, ya0, yb0, ya1, yb1);
      ptr += U * stride;
      off += U * LD;
    }
  }
  if constexpr (REM > 0 && TRIPS != 1)
    contract_in2_trip_x2<REM>(Ma, Mb, TRIPS * U * LD, ptr, delta, stride, ya0, yb0, ya1, yb1);
}
// buf[base + i*stride] = sum_p A(i,p) z0[p],  buf[base + delta + i*stride] = sum_p A(i,p) z1[p]   (MT transposed table)
template <typename T, int N1>
__device__ __forceinline__ void row_dot_x2(const T* MT, int off, const T (&z0)[N1], const T (&z1)[N1], T& r0, T& r1)
{
  T m[N1];
  tab_run<T, N1>(MT, off, m);
  dot2_n<T, N1>(m, z0, z1, r0, r1);
}
template <typename T, int N1>
__device__ __forceinline__ void contract_out_x2(const T (&MT)[N1 * tab_ld<N1>()], const T (&z0)[N1], const T (&z1)[N1], T* buf,
                                                int base, int delta, int stride)
{
  constexpr int LD = tab_ld<N1>();
  constexpr int U = Roll<T, N1>::in;
  constexpr int TRIPS = N1 / U, REM = N1 - TRIPS * U;
  T* ptr = buf + base;
  if constexpr (TRIPS == 1)
  {
#pragma unroll
    for (int i = 0; i < N1; ++i)
      row_dot_x2<T, N1>(MT, i * LD, z0, z1, ptr[i * stride], ptr[i * stride + delta]);
  }
  else
  {
    int off = 0;
#pragma unroll 1
    for (int t = 0; t < TRIPS; ++t)
    {
#pragma unroll
      for (int u = 0; u < U; ++u)
        row_dot_x2<T, N1>(MT, off + u * LD, z0, z1, ptr[u * stride], ptr[u * stride + delta]);
      ptr += U * stride;
      off += U * LD;
    }
#pragma unroll
    for (int u = 0; u < REM; ++u)
      row_dot_x2<T, N1>(MT, (TRIPS * U + u) * LD, z0, z1, ptr[u * stride], ptr[u * stride + delta]);
  }
}

template <class X, int AXIS>
constexpr int xstride()
{
  return AXIS == 0 ? X::s0 : (AXIS == 1 ? X::s1 : X::s2);
}
// A phase = (register axis R; tensor axes H, L behind the thread indices hi, lo).
template <int R_, int H_, int L_>
struct Phase
{
  static constexpr int R = R_, H = H_, L = L_;
};

// offset of a thread's column inside an exchange buffer with layout X, and its register stride
template <class X, class PH>
__device__ __forceinline__ int col_base(int c, int hi, int lo)
{
  return c * X::Sc + hi * xstride<X, PH::H>() + lo * xstride<X, PH::L>();
}
template <class X, class PH, typename T, int N1>
__device__ __forceinline__ void col_load(const T* buf, int base, T (&x)[N1])
{
#pragma unroll
  for (int r = 0; r < N1; ++r)
    x[r] = buf[base + r * xstride<X, PH::R>()];
}
template <class X, class PH, typename T, int N1>
__device__ __forceinline__ void col_store(T* buf, int base, const T (&x)[N1])
{
#pragma unroll
  for (int r = 0; r < N1; ++r)
    buf[base + r * xstride<X, PH::R>()] = x[r];
}
template <int N1>
struct Nat // natural layout as an exchange layout
{
  static constexpr int Sc = N1 * N1 * N1, s0 = N1 * N1, s1 = N1, s2 = 1;
};

// cells per tile.  Constraints: a tile of U must be a whole number of 16-byte units (fp64: C*n1^3
// even; fp32: C*n1^3 % 4 == 0); C*n1^2 close to a multiple of 32 keeps lanes busy.
template <typename T, int N1>
constexpr int default_cells()
{
  if (N1 == 9) return sizeof(T) == 8 ? 2 : 4;
  if (N1 == 8) return 2;
  if (N1 == 7) return 4;
  if (N1 == 6) return 7;
  if (N1 == 5) return sizeof(T) == 8 ? 10 : 8;  // slab kernel serves fp64
  if (N1 == 4) return 8;
  if (N1 == 3) return 28;
  return 64;                                     // n1 = 2: 4 threads per cell
}

template <typename T, int N1, int C>
struct ColGeom
{
  static constexpr int ND = N1 * N1 * N1;
  static constexpr int N2 = N1 * N1;
  static constexpr int ACTIVE = C * N2;
  static constexpr int NT = (ACTIVE + 31) & ~31;
  static constexpr int NWARPS = NT / 32;
  static constexpr uint32_t tile_bytes = C * ND * sizeof(T);
  static_assert(tile_bytes % 16 == 0, "a tile must be a whole number of 16-byte units");
  using L = ColumnLayout<T, N1>;
  using PC = Phase<0, 1, 2>;
  using PB = Phase<1, L::hB, L::lB>;
  using PA = Phase<2, L::hA, L::lA>;
  // phase A reads / writes natural-layout staging directly unless that conflicts badly
  static constexpr bool viaC = (L::stageA_wavefronts > 3 * L::stageA_ideal);
  static constexpr int e_elems = C * cmaxi(cmaxi(L::XBC::Sc, L::XAB::Sc), cmaxi(viaC ? L::XCA::Sc : 0, ND));
  static constexpr int e_bytes = up128(e_elems * (int)sizeof(T));
  static constexpr int stage_bytes = up128(C * ND * (int)sizeof(T));
};

// Resident blocks per SM the register allocator is asked to allow for (__launch_bounds__): what
// shared memory permits, but never so many that a thread gets fewer than `target_regs` registers.
constexpr int min_blocks_for(int threads, int smem_bytes, int target_regs)
{
  int by_smem = (227 * 1024) / (smem_bytes + 1024);
  int by_regs = 65536 / (threads * target_regs);
  int by_threads = 2048 / threads;
  int b = by_smem < by_regs ? by_smem : by_regs;
  b = b < by_threads ? b : by_threads;
  return b < 1 ? 1 : b;
}

// =========================================== mass =====================================================
template <typename T, int N1, int C>
struct MassCol
{
  using G = ColGeom<T, N1, C>;
  static constexpr int NE = G::viaC ? 3 : 2;
  static constexpr int off_u = 0;
  static constexpr int off_j = off_u + G::stage_bytes;
  static constexpr int off_e = off_j + G::stage_bytes;
  static constexpr int off_bar = off_e + NE * G::e_bytes;
  static constexpr int smem_bytes = off_bar + 16;
  static constexpr int min_blocks = min_blocks_for(G::NT, smem_bytes, sizeof(T) == 8 ? 64 : 48);
};

template <typename T, int N1, int C>
__global__ void __launch_bounds__(ColGeom<T, N1, C>::NT, MassCol<T, N1, C>::min_blocks)
mass_column_kernel(const __grid_constant__ TabA<T, N1> B, const T* __restrict__ U, T* __restrict__ W,
                   const T* __restrict__ J, long long ntiles)
{
  using MC = MassCol<T, N1, C>;
  using G = ColGeom<T, N1, C>;
  using L = typename G::L;
  using PA = typename G::PA;
  using PB = typename G::PB;
  using PC = typename G::PC;
  using XBC = typename L::XBC;
  using XAB = typename L::XAB;
  using XCA = typename L::XCA;
  constexpr int ND = G::ND, N2 = G::N2;

  extern __shared__ __align__(128) unsigned char smem[];
  T* uS = reinterpret_cast<T*>(smem + MC::off_u);
  T* jS = reinterpret_cast<T*>(smem + MC::off_j);
  T* E0 = reinterpret_cast<T*>(smem + MC::off_e);
  T* E1 = reinterpret_cast<T*>(smem + MC::off_e + G::e_bytes);
  T* E2 = reinterpret_cast<T*>(smem + MC::off_e + 2 * G::e_bytes); // only with viaC
  const uint32_t barU = smem_u32(smem + MC::off_bar), barJ = barU + 8;

  const int tid = threadIdx.x;
  const int te = min(tid, G::ACTIVE - 1); // idle lanes clone the last active thread
  const int c = te / N2, p = te - c * N2;
  const int hi = p / N1, lo = p - hi * N1;

  long long tile = blockIdx.x;
  if (tid == 0)
  {
    mbar_init(barU, 1);
    mbar_init(barJ, 1);
    fence_mbar_init();
    if (tile < ntiles)
    {
      mbar_expect_tx(barU, G::tile_bytes);
      bulk_g2s(smem_u32(uS), U + (size_t)tile * C * ND, G::tile_bytes, barU);
      mbar_expect_tx(barJ, G::tile_bytes);
      bulk_g2s(smem_u32(jS), J + (size_t)tile * C * ND, G::tile_bytes, barJ);
    }
  }
  __syncthreads();

  uint32_t phase = 0;
  for (; tile < ntiles; tile += gridDim.x, phase ^= 1)
  {
    const long long tn = tile + gridDim.x;
    T x[N1];
    // ---- phase C: forward along axis 0 (mass.hpp:36) -------------------------------------------------
    mbar_wait(barU, phase);
    if constexpr (G::viaC)
    {
      T d[N1];
      mbar_wait(barJ, phase);
      col_load<Nat<N1>, PC>(jS, c * ND + p, d);
      col_store<XCA, PC>(E2, col_base<XCA, PC>(c, hi, lo), d);
    }
    contract_in(B.a, uS, c * ND + p, N2, x);
    col_store<XBC, PC>(E0, col_base<XBC, PC>(c, hi, lo), x);
    __syncthreads(); // (1) staging consumed, E0 complete
    if (tid == 0 && tn < ntiles)
    {
      mbar_expect_tx(barU, G::tile_bytes);
      bulk_g2s(smem_u32(uS), U + (size_t)tn * C * ND, G::tile_bytes, barU);
      if constexpr (G::viaC)
      {
        mbar_expect_tx(barJ, G::tile_bytes);
        bulk_g2s(smem_u32(jS), J + (size_t)tn * C * ND, G::tile_bytes, barJ);
      }
    }
    // ---- phase B: axis 1 (mass.hpp:37) ------------------------------------------------------------------
    contract_in(B.a, E0, col_base<XBC, PB>(c, hi, lo), xstride<XBC, 1>(), x);
    col_store<XAB, PB>(E1, col_base<XAB, PB>(c, hi, lo), x);
    __syncthreads(); // (2)
    // ---- phase A: axis 2, quadrature scaling (mass.hpp:38-41), back along axis 2 (mass.hpp:45, same matrix) --
    contract_in(B.a, E1, col_base<XAB, PA>(c, hi, lo), xstride<XAB, 2>(), x);
    {
      T d[N1];
      if constexpr (G::viaC)
        col_load<XCA, PA>(E2, col_base<XCA, PA>(c, hi, lo), d);
      else
      {
        mbar_wait(barJ, phase);
        col_load<Nat<N1>, PA>(jS, col_base<Nat<N1>, PA>(c, hi, lo), d);
      }
#pragma unroll
      for (int r = 0; r < N1; ++r)
        x[r] = x[r] * d[r];
    }
    contract_out(B.at, x, E0, col_base<XAB, PA>(c, hi, lo), xstride<XAB, 2>());
    __syncthreads(); // (3)
    if constexpr (!G::viaC)
    {
      if (tid == 0 && tn < ntiles)
      {
        mbar_expect_tx(barJ, G::tile_bytes);
        bulk_g2s(smem_u32(jS), J + (size_t)tn * C * ND, G::tile_bytes, barJ);
      }
    }
    // ---- phase B': back along axis 1 (mass.hpp:46) ---------------------------------------------------
    contract_in(B.a, E0, col_base<XAB, PB>(c, hi, lo), xstride<XAB, 1>(), x);
    col_store<XBC, PB>(E1, col_base<XBC, PB>(c, hi, lo), x);
    __syncthreads(); // (4)
    // ---- phase C': back along axis 0 (mass.hpp:47), coalesced store of W ---------------------------------
    contract_in(B.a, E1, col_base<XBC, PC>(c, hi, lo), xstride<XBC, 0>(), x);
    if (tid < G::ACTIVE)
    {
      T* w = W + (size_t)tile * C * ND + c * ND + p;
#pragma unroll
      for (int r = 0; r < N1; ++r)
        w[r * N2] = x[r];
    }
  }
}

// ---- mass, two cells per thread (fp64 high degree) -----------------------------------------------------------
// Same phases, staging and exchange layouts as mass_column_kernel on a tile of C cells, but only C/2 cells'
// worth of threads: thread (c, hi, lo) owns that column in cells c and c + C/2 and shares every matrix fetch
// between the two (contract_in_x2 / contract_out_x2).
template <typename T, int N1, int C>
struct MassCol2
{
  static_assert(C % 2 == 0, "two cells per thread");
  using G = ColGeom<T, N1, C>;      // memory geometry: the whole tile
  using M1 = MassCol<T, N1, C>;
  static constexpr int ACTIVE = (C / 2) * N1 * N1;  // thread geometry: half of the tile's cells
  static constexpr int NT = (ACTIVE + 31) & ~31;
  static constexpr int smem_bytes = M1::smem_bytes;
  static constexpr int min_blocks = min_blocks_for(NT, smem_bytes, 96);
};

template <typename T, int N1, int C>
__global__ void __launch_bounds__(MassCol2<T, N1, C>::NT, MassCol2<T, N1, C>::min_blocks)
mass_column2_kernel(const __grid_constant__ TabA<T, N1> B, const T* __restrict__ U, T* __restrict__ W,
                    const T* __restrict__ J, long long ntiles)
{
  using MC = MassCol<T, N1, C>;
  using G = ColGeom<T, N1, C>;
  using H = MassCol2<T, N1, C>;
  using L = typename G::L;
  using PA = typename G::PA;
  using PB = typename G::PB;
  using PC = typename G::PC;
  using XBC = typename L::XBC;
  using XAB = typename L::XAB;
  using XCA = typename L::XCA;
  constexpr int ND = G::ND, N2 = G::N2, CH = C / 2;

  extern __shared__ __align__(128) unsigned char smem[];
  T* uS = reinterpret_cast<T*>(smem + MC::off_u);
  T* jS = reinterpret_cast<T*>(smem + MC::off_j);
  T* E0 = reinterpret_cast<T*>(smem + MC::off_e);
  T* E1 = reinterpret_cast<T*>(smem + MC::off_e + G::e_bytes);
  T* E2 = reinterpret_cast<T*>(smem + MC::off_e + 2 * G::e_bytes); // only with viaC
  const uint32_t barU = smem_u32(smem + MC::off_bar), barJ = barU + 8;

  const int tid = threadIdx.x;
  const int te = min(tid, H::ACTIVE - 1); // idle lanes clone the last active thread
  const int c = te / N2, p = te - c * N2;
  const int hi = p / N1, lo = p - hi * N1;

  long long tile = blockIdx.x;
  if (tid == 0)
  {
    mbar_init(barU, 1);
    mbar_init(barJ, 1);
    fence_mbar_init();
    if (tile < ntiles)
    {
      mbar_expect_tx(barU, G::tile_bytes);
      bulk_g2s(smem_u32(uS), U + (size_t)tile * C * ND, G::tile_bytes, barU);
      mbar_expect_tx(barJ, G::tile_bytes);
      bulk_g2s(smem_u32(jS), J + (size_t)tile * C * ND, G::tile_bytes, barJ);
    }
  }
  __syncthreads();

  uint32_t phase = 0;
  for (; tile < ntiles; tile += gridDim.x, phase ^= 1)
  {
    const long long tn = tile + gridDim.x;
    T x0[N1], x1[N1];
    // ---- phase C: forward along axis 0 (mass.hpp:36) -------------------------------------------------
    mbar_wait(barU, phase);
    if constexpr (G::viaC)
    {
      T d[N1];
      mbar_wait(barJ, phase);
#pragma unroll
      for (int r = 0; r < 2; ++r)
      {
        col_load<Nat<N1>, PC>(jS, (c + r * CH) * ND + p, d);
        col_store<XCA, PC>(E2, col_base<XCA, PC>(c + r * CH, hi, lo), d);
      }
    }
    contract_in_x2(B.a, uS, c * ND + p, CH * ND, N2, x0, x1);
    col_store<XBC, PC>(E0, col_base<XBC, PC>(c, hi, lo), x0);
    col_store<XBC, PC>(E0, col_base<XBC, PC>(c + CH, hi, lo), x1);
    __syncthreads(); // (1) staging consumed, E0 complete
    if (tid == 0 && tn < ntiles)
    {
      mbar_expect_tx(barU, G::tile_bytes);
      bulk_g2s(smem_u32(uS), U + (size_t)tn * C * ND, G::tile_bytes, barU);
      if constexpr (G::viaC)
      {
        mbar_expect_tx(barJ, G::tile_bytes);
        bulk_g2s(smem_u32(jS), J + (size_t)tn * C * ND, G::tile_bytes, barJ);
      }
    }
    // ---- phase B: axis 1 (mass.hpp:37) ------------------------------------------------------------------
    contract_in_x2(B.a, E0, col_base<XBC, PB>(c, hi, lo), CH * XBC::Sc, xstride<XBC, 1>(), x0, x1);
    col_store<XAB, PB>(E1, col_base<XAB, PB>(c, hi, lo), x0);
    col_store<XAB, PB>(E1, col_base<XAB, PB>(c + CH, hi, lo), x1);
    __syncthreads(); // (2)
    // ---- phase A: axis 2, quadrature scaling (mass.hpp:38-41), back along axis 2 (mass.hpp:45) ----------
    contract_in_x2(B.a, E1, col_base<XAB, PA>(c, hi, lo), CH * XAB::Sc, xstride<XAB, 2>(), x0, x1);
    {
      T d0[N1], d1[N1];
      if constexpr (G::viaC)
      {
        col_load<XCA, PA>(E2, col_base<XCA, PA>(c, hi, lo), d0);
        col_load<XCA, PA>(E2, col_base<XCA, PA>(c + CH, hi, lo), d1);
      }
      else
      {
        mbar_wait(barJ, phase);
        col_load<Nat<N1>, PA>(jS, col_base<Nat<N1>, PA>(c, hi, lo), d0);
        col_load<Nat<N1>, PA>(jS, col_base<Nat<N1>, PA>(c + CH, hi, lo), d1);
      }
#pragma unroll
      for (int r = 0; r < N1; ++r)
      {
        x0[r] = x0[r] * d0[r];
        x1[r] = x1[r] * d1[r];
      }
    }
    contract_out_x2(B.at, x0, x1, E0, col_base<XAB, PA>(c, hi, lo), CH * XAB::Sc, xstride<XAB, 2>());
    __syncthreads(); // (3)
    if constexpr (!G::viaC)
    {
      if (tid == 0 && tn < ntiles)
      {
        mbar_expect_tx(barJ, G::tile_bytes);
        bulk_g2s(smem_u32(jS), J + (size_t)tn * C * ND, G::tile_bytes, barJ);
      }
    }
    // ---- phase B': back along axis 1 (mass.hpp:46) ---------------------------------------------------
    contract_in_x2(B.a, E0, col_base<XAB, PB>(c, hi, lo), CH * XAB::Sc, xstride<XAB, 1>(), x0, x1);
    col_store<XBC, PB>(E1, col_base<XBC, PB>(c, hi, lo), x0);
    col_store<XBC, PB>(E1, col_base<XBC, PB>(c + CH, hi, lo), x1);
    __syncthreads(); // (4)
    // ---- phase C': back along axis 0 (mass.hpp:47), coalesced store of W ---------------------------------
    contract_in_x2(B.a, E1, col_base<XBC, PC>(c, hi, lo), CH * XBC::Sc, xstride<XBC, 0>(), x0, x1);
    if (tid < H::ACTIVE)
    {
      T* w = W + (size_t)tile * C * ND + c * ND + p;
#pragma unroll
      for (int r = 0; r < N1; ++r)
      {
        w[r * N2] = x0[r];
        w[r * N2 + CH * ND] = x1[r];
      }
    }
  }
}

// ========================================= stiffness ==================================================
// Warp-private geometry ring.  For plane i0 the 6-value records of a warp's lanes (thread order =
// point order) form `pieces` contiguous runs in HBM, one per cell the warp touches.  Piece k is
// issued by lane k.  fp32 records are 24 bytes, so a piece may start / end on an 8-byte boundary:
// the copy then covers the enclosing 16-byte units and readers skip the leading 8 bytes (`shift`);
// tiles are whole 16-byte units, so the over-read never leaves the tile's own records.
//
// PW = true (producer-warp variant): one extra warp per block requests the geometry for the whole
// block -- per plane one bulk copy per cell of the tile (n1^2 records, contiguous) into a block-wide
// ring slot -- and the compute warps only wait on the slot's `full` barrier and arrive on its `empty`
// barrier.  The per-plane request code (address arithmetic, election loop around the bulk copy) then
// runs C times per block instead of once per warp, which is what the issue-bound high-degree fp32
// kernels need (ncu: profiles/r01_ncu_weak_configs_before.txt).
template <typename T, int N1, int C, int MODE, int NG, bool PW = false>
struct StiffCol
{
  using G = ColGeom<T, N1, C>;
  static constexpr bool FULL = MODE == FEMTC_STIFFNESS_FULL;
  static constexpr int RB = 6 * (int)sizeof(T);                          // record bytes
  static constexpr int MAXP = (32 + G::N2 - 2) / G::N2 + 1;              // pieces a warp can have
  static constexpr bool F32 = sizeof(T) == 4;
  static constexpr int piece_stride = up16(G::N2 * RB + (F32 ? 8 : 0));  // PW: one cell's plane (+ fp32 skew)
  static constexpr int slot_bytes = PW ? C * piece_stride : up16(32 * RB) + (F32 ? 32 * MAXP + 16 : 0);
  static constexpr int ring_bytes = (PW ? 1 : G::NWARPS) * NG * slot_bytes;
  static constexpr int n_bars = PW ? 1 + 2 * NG : 1 + 2 * G::NWARPS * NG; // U, then `full` and `empty` per ring slot
  static constexpr int threads = G::NT + (PW ? 32 : 0);
  static constexpr int NE = G::viaC ? 4 : 3;
  static constexpr int off_u = 0;
  static constexpr int off_e = off_u + G::stage_bytes;
  static constexpr int off_g = off_e + NE * G::e_bytes;
  static constexpr int off_bar = off_g + ring_bytes;
  static constexpr int smem_bytes = off_bar + 8 * n_bars + 8;
  static constexpr int min_blocks = min_blocks_for(threads, smem_bytes, sizeof(T) == 8 ? (FULL ? 128 : 104) : (FULL ? ((N1 == 7 || N1 == 8) ? 88 : 104) : 72));
  // (FULL fp32 n1 = 7, 8: one more resident block is worth 28 bytes of spills: 42 -> 45 %, 56 -> 60 % of measured HBM)
};

template <typename T, int N1, int C, int MODE, int NG, bool PW>
__global__ void __launch_bounds__(StiffCol<T, N1, C, MODE, NG, PW>::threads, StiffCol<T, N1, C, MODE, NG, PW>::min_blocks)
stiffness_column_kernel(const __grid_constant__ TabAD<T, N1> B, const T* __restrict__ U, T* __restrict__ W,
                        const T* __restrict__ Gm, long long ntiles)
{
  using SC = StiffCol<T, N1, C, MODE, NG, PW>;
  using G = ColGeom<T, N1, C>;
  using L = typename G::L;
  using PA = typename G::PA;
  using PB = typename G::PB;
  using PC = typename G::PC;
  using XBC = typename L::XBC;
  using XAB = typename L::XAB;
  using XCA = typename L::XCA;
  constexpr bool FULL = SC::FULL;
  constexpr int ND = G::ND, N2 = G::N2, RB = SC::RB;

  extern __shared__ __align__(128) unsigned char smem[];
  T* uS = reinterpret_cast<T*>(smem + SC::off_u);
  T* E0 = reinterpret_cast<T*>(smem + SC::off_e);
  T* E1 = reinterpret_cast<T*>(smem + SC::off_e + G::e_bytes);
  T* E2 = reinterpret_cast<T*>(smem + SC::off_e + 2 * G::e_bytes);
  T* E3 = reinterpret_cast<T*>(smem + SC::off_e + 3 * G::e_bytes); // only with viaC
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  unsigned char* gRing = smem + SC::off_g + (PW ? 0 : warp * NG * SC::slot_bytes);
  const uint32_t barU = smem_u32(smem + SC::off_bar);
  const uint32_t barG = barU + 8 + (PW ? 0 : warp * 2 * NG * 8); // `full` barriers of the ring slots
  const uint32_t barE = barG + 8 * NG;                            // their `empty` barriers (PW: one arrival per compute warp)
  // the compute threads' block barrier (the producer warp never takes part in it)
  auto block_sync = [&]()
  {
    if constexpr (PW)
      asm volatile("bar.sync 1, %0;" ::"n"(G::NT) : "memory");
    else
      __syncthreads();
  };

  const int te = min(tid, G::ACTIVE - 1);
  const int c = te / N2, p = te - c * N2;
  const int hi = p / N1, lo = p - hi * N1;

  // ---- this warp's geometry pieces (tile independent) -------------------------------------------------
  const int q0 = warp * 32;                                  // first thread of the warp
  const int nAct = min(32, G::ACTIVE - q0);                  // active lanes of the warp
  const int cFirst = q0 / N2;
  const int npieces = (q0 + nAct - 1) / N2 - cFirst + 1;
  // as issuer of piece `lane`
  int is_cell = cFirst + lane;
  int is_qs = max(q0, is_cell * N2), is_qe = min(q0 + nAct, (is_cell + 1) * N2);
  const bool issuer = lane < npieces;
  const int is_rec = is_cell * ND + (is_qs - is_cell * N2);  // record index inside the tile, plane 0
  const int is_cnt = is_qe - is_qs;
  const int is_dst = up16(RB * (is_qs - q0)) + (SC::F32 ? 32 * lane : 0);
  // as reader: my piece is the one of my (cloned) thread
  const int my_k = c - cFirst;
  const int my_qs = max(q0, c * N2);
  // fp32: parity of the piece's first record; PW: a piece is a whole cell plane
  const int my_par = PW ? ((c * ND) & 1) : ((c * ND + (my_qs - c * N2)) & 1);
  const int my_off = PW ? c * SC::piece_stride + RB * p
                        : up16(RB * (my_qs - q0)) + (SC::F32 ? 32 * my_k : 0) + RB * (te - my_qs);

  const long long my_tiles = (long long)blockIdx.x < ntiles ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;

  // Ring bookkeeping without any division: the issue side runs exactly NG planes ahead of the consume
  // side, so while plane i0 of a tile is consumed the plane to request is i0 + NG of the same tile, or
  // i0 + NG - N1 of the block's next tile -- a compile-time distinction wherever the plane loop is
  // unrolled.  Only (slot, parity) and one byte pointer per tile are carried.
  static_assert(PW || NG <= N1, "the warp-private geometry ring may run at most one tile ahead");
  const unsigned char* g_tile = reinterpret_cast<const unsigned char*>(Gm) +
                                ((size_t)blockIdx.x * C * ND + is_rec) * RB;   // my piece, plane 0, current tile
  const size_t g_stride = (size_t)gridDim.x * C * ND * RB;                      // to the block's next tile
  const int is_par = SC::F32 ? (is_rec & 1) : 0;              // tile bases are even record counts
  const uint32_t is_bytes0 = (uint32_t)up16(RB * is_cnt), is_bytes1 = (uint32_t)up16(RB * is_cnt + 8);
  const uint32_t is_dst_u32 = smem_u32(gRing) + is_dst;

  auto request_plane = [&](int plane, int slot, const unsigned char* tile_src)
  {
    if (!PW && issuer)
    {
      const int sh = SC::F32 ? ((is_par + plane * (N2 & 1)) & 1) : 0;
      const uint32_t bytes = sh ? is_bytes1 : is_bytes0;
      mbar_expect_tx(barG + 8 * slot, bytes);
      bulk_g2s(is_dst_u32 + slot * SC::slot_bytes, tile_src + plane * (N2 * RB) - 8 * sh, bytes, barG + 8 * slot);
    }
  };

  long long tile = blockIdx.x;
  if (tid == 0)
  {
    mbar_init(barU, 1);
    fence_mbar_init();
  }
  if constexpr (PW)
  {
    if (tid == 0)
    {
      for (int s = 0; s < NG; ++s)
      {
        mbar_init(barG + 8 * s, C);          // one arrive.expect_tx per cell piece
        mbar_init(barE + 8 * s, G::NWARPS);  // one arrive per compute warp
      }
      fence_mbar_init();
    }
  }
  else if (lane == 0)
  {
    for (int s = 0; s < NG; ++s)
    {
      mbar_init(barG + 8 * s, npieces);
      mbar_init(barE + 8 * s, FEMTC_RELEASE_LANES);
    }
    fence_mbar_init();
  }
  __syncthreads();
  if constexpr (PW)
  {
    if (warp == G::NWARPS)
    {
      // ---- producer warp: lane k streams the planes of cell k of every tile of this block ------------------
      if (lane < C)
      {
        const unsigned char* src = reinterpret_cast<const unsigned char*>(Gm) +
                                   ((size_t)blockIdx.x * C + lane) * ND * RB;   // cell `lane`, plane 0
        const int par0 = SC::F32 ? ((lane * ND) & 1) : 0;                       // tile bases are even record counts
        const uint32_t dst = smem_u32(gRing) + lane * SC::piece_stride;
        constexpr uint32_t bytes0 = (uint32_t)up16(N2 * RB), bytes1 = (uint32_t)up16(N2 * RB + 8);
        int slot = 0;
        uint32_t epar = 1;                                                      // a fresh barrier passes parity 1
        for (long long t = 0; t < my_tiles; ++t)
        {
#pragma unroll
          for (int pl = 0; pl < N1; ++pl)
          {
            const int sh = SC::F32 ? ((par0 + pl * (N2 & 1)) & 1) : 0;
            const uint32_t bytes = sh ? bytes1 : bytes0;
            mbar_wait(barE + 8 * slot, epar);
            mbar_expect_tx(barG + 8 * slot, bytes);
            bulk_g2s(dst + slot * SC::slot_bytes, src + pl * (N2 * RB) - 8 * sh, bytes, barG + 8 * slot);
            if (++slot == NG)
            {
              slot = 0;
              epar ^= 1;
            }
          }
          src += g_stride;
        }
      }
      return;
    }
  }
  if (tid == 0 && tile < ntiles)
  {
    mbar_expect_tx(barU, G::tile_bytes);
    bulk_g2s(smem_u32(uS), U + (size_t)tile * C * ND, G::tile_bytes, barU);
  }
  if (!PW && my_tiles > 0)
    for (int s = 0; s < NG; ++s)
      request_plane(s, s, g_tile);
  int c_slot = 0;
  uint32_t c_par = 0;

  uint32_t phase = 0;
  for (; tile < ntiles; tile += gridDim.x, phase ^= 1)
  {
    const long long tn = tile + gridDim.x;
    auto next_u = [&]()
    {
      if (tid == 0 && tn < ntiles)
      {
        mbar_expect_tx(barU, G::tile_bytes);
        bulk_g2s(smem_u32(uS), U + (size_t)tn * C * ND, G::tile_bytes, barU);
      }
    };
    // ---- phase A: u columns along axis 2; t = A2 u (FULL: s = D2 u) ------------------------------------
    {
      T y[N1], z[N1];
      mbar_wait(barU, phase);
      const T* src = uS;
      int sbase = col_base<Nat<N1>, PA>(c, hi, lo), sstride = 1;
      if constexpr (G::viaC)
      {
        col_load<Nat<N1>, PC>(uS, c * ND + p, y);
        col_store<XCA, PC>(E3, col_base<XCA, PC>(c, hi, lo), y);
        block_sync(); // (0)
        next_u();
        src = E3, sbase = col_base<XCA, PA>(c, hi, lo), sstride = xstride<XCA, 2>();
      }
      if constexpr (!FULL)
        contract_in(B.a, src, sbase, sstride, y);
      else
      {
        contract_in2(B.a, B.d, src, sbase, sstride, y, z);
        col_store<XAB, PA>(E1, col_base<XAB, PA>(c, hi, lo), z);
      }
      col_store<XAB, PA>(E0, col_base<XAB, PA>(c, hi, lo), y);
    }
    if constexpr (!G::viaC)
      if (tid == 0)
        bulk_wait_read0(); // the previous tile's W store has finished reading E2
    block_sync(); // (1)
    if constexpr (!G::viaC)
      next_u();
    // ---- phase B: axis 1 ------------------------------------------------------------------------------------
    {
      T ya[N1], yd[N1];
      contract_in2(B.a, B.d, E0, col_base<XAB, PB>(c, hi, lo), xstride<XAB, 1>(), ya, yd); // A1 A2 u, D1 A2 u
      if constexpr (!FULL)
      {
        col_store<XBC, PB>(E1, col_base<XBC, PB>(c, hi, lo), ya);
        col_store<XBC, PB>(E2, col_base<XBC, PB>(c, hi, lo), yd);
      }
      else
      {
        T ys[N1];
        contract_in(B.a, E1, col_base<XAB, PB>(c, hi, lo), xstride<XAB, 1>(), ys);           // A1 D2 u
        block_sync(); // (1b) E0, E1 are rewritten below
        col_store<XBC, PB>(E0, col_base<XBC, PB>(c, hi, lo), ya);
        col_store<XBC, PB>(E1, col_base<XBC, PB>(c, hi, lo), yd);
        col_store<XBC, PB>(E2, col_base<XBC, PB>(c, hi, lo), ys);
      }
    }
    block_sync(); // (2)
    // ---- phase C: axis 0 forward, geometry, axis 0 backward -- fused plane by plane ---------------------------
    // For plane i0:  w0 = D(i0,:) ta, w1 = A(i0,:) td, w2 = A(i0,:) ta   [REF_EXACT: stiffness.hpp:68-77]
    //                                                  w2 = A(i0,:) sa   [FULL]
    // then the 3x3 symmetric geometry map (stiffness.hpp:14-29) and h(:) += M(:,i0) f.
    {
      T ta[N1], td[N1], h2[N1];
      T sa[FULL ? N1 : 1], h0[FULL ? N1 : 1], h1[FULL ? N1 : 1];
      if constexpr (!FULL)
      {
        col_load<XBC, PC>(E1, col_base<XBC, PC>(c, hi, lo), ta);
        col_load<XBC, PC>(E2, col_base<XBC, PC>(c, hi, lo), td);
      }
      else
      {
        col_load<XBC, PC>(E0, col_base<XBC, PC>(c, hi, lo), ta);
        col_load<XBC, PC>(E1, col_base<XBC, PC>(c, hi, lo), td);
        col_load<XBC, PC>(E2, col_base<XBC, PC>(c, hi, lo), sa);
      }
#pragma unroll
      for (int i = 0; i < N1; ++i)
      {
        h2[i] = T(0);
        if constexpr (FULL)
          h0[i] = h1[i] = T(0);
      }
#pragma unroll(Roll<T, N1>::in)
      for (int i0 = 0; i0 < N1; ++i0)
      {
        const T w0 = row_dot(B.dt, i0, ta);
        T w1, w2; // one table row against two columns (fp32: one FFMA2 per entry)
        {
          T mrow[N1];
          tab_run<T, N1>(B.at, i0 * tab_ld<N1>(), mrow);
          if constexpr (FULL)
            dot2_n<T, N1, N1 != 9>(mrow, td, sa, w1, w2);
          else
            dot2_n<T, N1, N1 != 9>(mrow, td, ta, w1, w2);
        }
        mbar_wait(barG + 8 * c_slot, c_par);
        const int sh = SC::F32 ? ((my_par + i0 * (N2 & 1)) & 1) : 0;
        const T* gp = reinterpret_cast<const T*>(gRing + c_slot * SC::slot_bytes + my_off + 8 * sh);
        T g[6];
        if constexpr (sizeof(T) == 8)
        {
          const double2 a = *reinterpret_cast<const double2*>(gp);
          g[0] = a.x, g[1] = a.y;
          if constexpr (FULL)
          {
            const double2 b = *reinterpret_cast<const double2*>(gp + 2);
            const double2 cc = *reinterpret_cast<const double2*>(gp + 4);
            g[2] = b.x, g[3] = b.y, g[4] = cc.x, g[5] = cc.y;
          }
          else
            g[2] = gp[2];
        }
        else
        {
          const float2 a = *reinterpret_cast<const float2*>(gp);
          g[0] = a.x, g[1] = a.y;
          if constexpr (FULL)
          {
            const float2 b = *reinterpret_cast<const float2*>(gp + 2);
            const float2 cc = *reinterpret_cast<const float2*>(gp + 4);
            g[2] = b.x, g[3] = b.y, g[4] = cc.x, g[5] = cc.y;
          }
          else
            g[2] = gp[2];
        }
        const T f2 = fma_t(g[0], w2, fma_t(g[1], w1, g[2] * w0));
        if constexpr (FULL)
        {
          const T f0 = fma_t(g[2], w2, fma_t(g[4], w1, g[5] * w0));
          const T f1 = fma_t(g[1], w2, fma_t(g[3], w1, g[4] * w0));
          col_axpy(B.d, i0, f0, h0);                            // D0 f0
          col_axpy(B.a, i0, f1, h1);                            // A0 f1
        }
        // REF_EXACT: only W = (A x A x D) f2 is observable (stiffness.hpp:95-97 overwrites :83-91)
        col_axpy(B.a, i0, f2, h2);                              // A0 f2
        // every lane is done with this slot: hand it back through its `empty` barrier (tma.cuh, rule 2) -- to the
        // producer warp (PW), or to this warp's own issuing lanes, which then request the chunk NG ahead into it
        if constexpr (PW)
        {
          __syncwarp();
          if (lane == 0)
            mbar_arrive(barE + 8 * c_slot);
        }
        else
        {
          slot_release_and_acquire(barE + 8 * c_slot, c_par, lane);
          if (i0 + NG < N1)
            request_plane(i0 + NG, c_slot, g_tile);
          else if (tn < ntiles)
            request_plane(i0 + NG - N1, c_slot, g_tile + g_stride);
        }
        if (++c_slot == NG)
        {
          c_slot = 0;
          c_par ^= 1;
        }
      }
      if constexpr (!FULL)
        col_store<XBC, PC>(E0, col_base<XBC, PC>(c, hi, lo), h2);
      else
      {
        block_sync(); // (2b) E0..E2 were read by this phase
        col_store<XBC, PC>(E0, col_base<XBC, PC>(c, hi, lo), h0);
        col_store<XBC, PC>(E1, col_base<XBC, PC>(c, hi, lo), h1);
        col_store<XBC, PC>(E2, col_base<XBC, PC>(c, hi, lo), h2);
      }
    }
    g_tile += g_stride;
    block_sync(); // (3)
    // ---- phase B': back along axis 1 --------------------------------------------------------------------------
    {
      T m[N1];
      contract_in(B.a, E0, col_base<XBC, PB>(c, hi, lo), xstride<XBC, 1>(), m);              // A1 (.)
      if constexpr (!FULL)
        col_store<XAB, PB>(E1, col_base<XAB, PB>(c, hi, lo), m);
      else
      {
        T n[N1];
        contract_in_add(B.d, E1, col_base<XBC, PB>(c, hi, lo), xstride<XBC, 1>(), m);        // + D1 A0 f1
        contract_in(B.a, E2, col_base<XBC, PB>(c, hi, lo), xstride<XBC, 1>(), n);            // A1 A0 f2
        block_sync(); // (3b)
        col_store<XAB, PB>(E0, col_base<XAB, PB>(c, hi, lo), m);
        col_store<XAB, PB>(E1, col_base<XAB, PB>(c, hi, lo), n);
      }
    }
    block_sync(); // (4)
    // ---- phase A': back along axis 2, W ---------------------------------------------------------------------------
    {
      T out[N1];
      if constexpr (!FULL)
        contract_in(B.d, E1, col_base<XAB, PA>(c, hi, lo), xstride<XAB, 2>(), out);          // D2 (.)
      else
      {
        contract_in(B.a, E0, col_base<XAB, PA>(c, hi, lo), xstride<XAB, 2>(), out);
        contract_in_add(B.d, E1, col_base<XAB, PA>(c, hi, lo), xstride<XAB, 2>(), out);
      }
      if constexpr (G::viaC)
      {
        col_store<XCA, PA>(E2, col_base<XCA, PA>(c, hi, lo), out);
        block_sync(); // (5)
        col_load<XCA, PC>(E2, col_base<XCA, PC>(c, hi, lo), out);
        if (tid < G::ACTIVE)
        {
          T* w = W + (size_t)tile * C * ND + c * ND + p;
#pragma unroll
          for (int r = 0; r < N1; ++r)
            w[r * N2] = out[r];
        }
      }
      else
      {
        col_store<Nat<N1>, PA>(E2, col_base<Nat<N1>, PA>(c, hi, lo), out);
        fence_proxy_async();
        block_sync(); // (5)
        if (tid == 0)
        {
          bulk_s2g(W + (size_t)tile * C * ND, smem_u32(E2), G::tile_bytes);
          bulk_commit();
        }
      }
    }
  }
  if constexpr (!G::viaC)
    if (tid == 0)
      bulk_wait_read0();
}

// ---- stiffness REF_EXACT, two cells per thread -------------------------------------------------------------
// Same phases, exchange layouts and results as stiffness_column_kernel<.., REF_EXACT, .., PW = true> on a tile of
// C cells, with C/2 cells' worth of compute threads: thread (c, hi, lo) owns that column in cells c and c + C/2
// and shares every matrix fetch between the two.  Geometry arrives through the block-wide producer-warp ring
// (a piece = one cell's plane, so the second cell is one constant offset away).  n1 without the viaC detour only.
template <typename T, int N1, int C, int NG>
struct StiffCol2
{
  static_assert(C % 2 == 0, "two cells per thread");
  using G = ColGeom<T, N1, C>;
  static_assert(!G::viaC, "shapes that stage phase A through a C-order exchange are not served");
  static constexpr bool F32 = sizeof(T) == 4;
  static constexpr int RB = 6 * (int)sizeof(T);
  static constexpr int CH = C / 2;
  static constexpr int ACTIVE = CH * G::N2;
  static constexpr int NT = (ACTIVE + 31) & ~31;
  static constexpr int NWARPS = NT / 32;
  static constexpr int threads = NT + 32;
  static constexpr int piece_stride = up16(G::N2 * RB + (F32 ? 8 : 0));
  static constexpr int slot_bytes = C * piece_stride;
  static constexpr int off_u = 0;
  static constexpr int off_e = off_u + G::stage_bytes;
  static constexpr int off_g = off_e + 3 * G::e_bytes;
  static constexpr int off_bar = off_g + NG * slot_bytes;
  static constexpr int smem_bytes = off_bar + 8 * (1 + 2 * NG) + 8;
  static constexpr int min_blocks = min_blocks_for(threads, smem_bytes, F32 ? 112 : 168);
};

template <typename T, int N1, int C, int NG>
__global__ void __launch_bounds__(StiffCol2<T, N1, C, NG>::threads, StiffCol2<T, N1, C, NG>::min_blocks)
stiffness_column2_kernel(const __grid_constant__ TabAD<T, N1> B, const T* __restrict__ U, T* __restrict__ W,
                         const T* __restrict__ Gm, long long ntiles)
{
  using SC = StiffCol2<T, N1, C, NG>;
  using G = ColGeom<T, N1, C>;
  using L = typename G::L;
  using PA = typename G::PA;
  using PB = typename G::PB;
  using PC = typename G::PC;
  using XBC = typename L::XBC;
  using XAB = typename L::XAB;
  constexpr int ND = G::ND, N2 = G::N2, RB = SC::RB, CH = SC::CH, LD = tab_ld<N1>();

  extern __shared__ __align__(128) unsigned char smem[];
  T* uS = reinterpret_cast<T*>(smem + SC::off_u);
  T* E0 = reinterpret_cast<T*>(smem + SC::off_e);
  T* E1 = reinterpret_cast<T*>(smem + SC::off_e + G::e_bytes);
  T* E2 = reinterpret_cast<T*>(smem + SC::off_e + 2 * G::e_bytes);
  unsigned char* gRing = smem + SC::off_g;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t barU = smem_u32(smem + SC::off_bar);
  const uint32_t barG = barU + 8;            // `full` barriers of the ring slots
  const uint32_t barE = barG + 8 * NG;       // `empty` barriers
  auto block_sync = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(SC::NT) : "memory"); };

  const int te = min(tid, SC::ACTIVE - 1);
  const int c = te / N2, p = te - c * N2;
  const int hi = p / N1, lo = p - hi * N1;
  const long long my_tiles = (long long)blockIdx.x < ntiles ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const size_t g_stride = (size_t)gridDim.x * C * ND * RB;

  if (tid == 0)
  {
    mbar_init(barU, 1);
    for (int s = 0; s < NG; ++s)
    {
      mbar_init(barG + 8 * s, C);
      mbar_init(barE + 8 * s, SC::NWARPS);
    }
    fence_mbar_init();
  }
  __syncthreads();
  if (warp == SC::NWARPS)
  {
    // ---- producer warp: lane k streams the planes of cell k of every tile of this block ----------------------
    if (lane < C)
    {
      const unsigned char* src = reinterpret_cast<const unsigned char*>(Gm) + ((size_t)blockIdx.x * C + lane) * ND * RB;
      const int par0 = SC::F32 ? ((lane * ND) & 1) : 0;
      const uint32_t dst = smem_u32(gRing) + lane * SC::piece_stride;
      constexpr uint32_t bytes0 = (uint32_t)up16(N2 * RB), bytes1 = (uint32_t)up16(N2 * RB + 8);
      int slot = 0;
      uint32_t epar = 1;
      for (long long t = 0; t < my_tiles; ++t)
      {
#pragma unroll
        for (int pl = 0; pl < N1; ++pl)
        {
          const int sh = SC::F32 ? ((par0 + pl * (N2 & 1)) & 1) : 0;
          const uint32_t bytes = sh ? bytes1 : bytes0;
          mbar_wait(barE + 8 * slot, epar);
          mbar_expect_tx(barG + 8 * slot, bytes);
          bulk_g2s(dst + slot * SC::slot_bytes, src + pl * (N2 * RB) - 8 * sh, bytes, barG + 8 * slot);
          if (++slot == NG)
          {
            slot = 0;
            epar ^= 1;
          }
        }
        src += g_stride;
      }
    }
    return;
  }
  long long tile = blockIdx.x;
  if (tid == 0 && tile < ntiles)
  {
    mbar_expect_tx(barU, G::tile_bytes);
    bulk_g2s(smem_u32(uS), U + (size_t)tile * C * ND, G::tile_bytes, barU);
  }
  // geometry reader offsets of my two cells (a piece = one cell's plane; fp32: 8-byte skew by record parity)
  const int my_off0 = c * SC::piece_stride + RB * p;
  const int my_off1 = my_off0 + CH * SC::piece_stride;
  const int my_par0 = SC::F32 ? ((c * ND) & 1) : 0;
  const int my_par1 = SC::F32 ? (((c + CH) * ND) & 1) : 0;
  int c_slot = 0;
  uint32_t c_par = 0;

  uint32_t phase = 0;
  for (; tile < ntiles; tile += gridDim.x, phase ^= 1)
  {
    const long long tn = tile + gridDim.x;
    // ---- phase A: t = A2 u ------------------------------------------------------------------------------------
    {
      T y0[N1], y1[N1];
      mbar_wait(barU, phase);
      contract_in_x2(B.a, uS, col_base<Nat<N1>, PA>(c, hi, lo), CH * ND, 1, y0, y1);
      col_store<XAB, PA>(E0, col_base<XAB, PA>(c, hi, lo), y0);
      col_store<XAB, PA>(E0, col_base<XAB, PA>(c + CH, hi, lo), y1);
    }
    if (tid == 0)
      bulk_wait_read0(); // the previous tile's W store has finished reading E2
    block_sync(); // (1)
    if (tid == 0 && tn < ntiles)
    {
      mbar_expect_tx(barU, G::tile_bytes);
      bulk_g2s(smem_u32(uS), U + (size_t)tn * C * ND, G::tile_bytes, barU);
    }
    // ---- phase B: A1 A2 u and D1 A2 u ----------------------------------------------------------------------------
    {
      T ya0[N1], yd0[N1], ya1[N1], yd1[N1];
      contract_in2_x2(B.a, B.d, E0, col_base<XAB, PB>(c, hi, lo), CH * XAB::Sc, xstride<XAB, 1>(), ya0, yd0, ya1, yd1);
      col_store<XBC, PB>(E1, col_base<XBC, PB>(c, hi, lo), ya0);
      col_store<XBC, PB>(E2, col_base<XBC, PB>(c, hi, lo), yd0);
      col_store<XBC, PB>(E1, col_base<XBC, PB>(c + CH, hi, lo), ya1);
      col_store<XBC, PB>(E2, col_base<XBC, PB>(c + CH, hi, lo), yd1);
    }
    block_sync(); // (2)
    // ---- phase C: axis 0 forward, geometry, axis 0 backward, plane by plane (stiffness.hpp:68-77, :14-29, :95-97) --
    {
      T ta0[N1], td0[N1], ta1[N1], td1[N1], h0[N1], h1[N1];
      col_load<XBC, PC>(E1, col_base<XBC, PC>(c, hi, lo), ta0);
      col_load<XBC, PC>(E2, col_base<XBC, PC>(c, hi, lo), td0);
      col_load<XBC, PC>(E1, col_base<XBC, PC>(c + CH, hi, lo), ta1);
      col_load<XBC, PC>(E2, col_base<XBC, PC>(c + CH, hi, lo), td1);
#pragma unroll
      for (int i = 0; i < N1; ++i)
        h0[i] = h1[i] = T(0);
#pragma unroll(Roll<T, N1>::in)
      for (int i0 = 0; i0 < N1; ++i0)
      {
        T w0a, w0b, w1a, w1b, w2a, w2b;
        row_dot_x2<T, N1>(B.dt, i0 * LD, ta0, ta1, w0a, w0b);   // D(i0,:) ta
        row_dot_x2<T, N1>(B.at, i0 * LD, td0, td1, w1a, w1b);   // A(i0,:) td
        row_dot_x2<T, N1>(B.at, i0 * LD, ta0, ta1, w2a, w2b);   // A(i0,:) ta
        mbar_wait(barG + 8 * c_slot, c_par);
        const int sha = SC::F32 ? ((my_par0 + i0 * (N2 & 1)) & 1) : 0;
        const int shb = SC::F32 ? ((my_par1 + i0 * (N2 & 1)) & 1) : 0;
        const T* ga = reinterpret_cast<const T*>(gRing + c_slot * SC::slot_bytes + my_off0 + 8 * sha);
        const T* gb = reinterpret_cast<const T*>(gRing + c_slot * SC::slot_bytes + my_off1 + 8 * shb);
        T g0a, g1a, g2a, g0b, g1b, g2b;
        if constexpr (sizeof(T) == 8)
        {
          const double2 a = *reinterpret_cast<const double2*>(ga);
          const double2 b = *reinterpret_cast<const double2*>(gb);
          g0a = a.x, g1a = a.y, g0b = b.x, g1b = b.y;
        }
        else
        {
          const float2 a = *reinterpret_cast<const float2*>(ga);
          const float2 b = *reinterpret_cast<const float2*>(gb);
          g0a = a.x, g1a = a.y, g0b = b.x, g1b = b.y;
        }
        g2a = ga[2], g2b = gb[2];
        __syncwarp(); // every lane holds its records: hand the slot back to the producer
        if (lane == 0)
          mbar_arrive(barE + 8 * c_slot);
        if (++c_slot == NG)
        {
          c_slot = 0;
          c_par ^= 1;
        }
        const T fa = fma_t(g0a, w2a, fma_t(g1a, w1a, g2a * w0a));
        const T fb = fma_t(g0b, w2b, fma_t(g1b, w1b, g2b * w0b));
        T m[N1];
        tab_run<T, N1>(B.a, i0 * LD, m);                        // A(:, i0)
        axpy_n<T, N1>(m, fa, h0);
        axpy_n<T, N1>(m, fb, h1);
      }
      col_store<XBC, PC>(E0, col_base<XBC, PC>(c, hi, lo), h0);
      col_store<XBC, PC>(E0, col_base<XBC, PC>(c + CH, hi, lo), h1);
    }
    block_sync(); // (3)
    // ---- phase B': back along axis 1 ------------------------------------------------------------------------------
    {
      T m0[N1], m1[N1];
      contract_in_x2(B.a, E0, col_base<XBC, PB>(c, hi, lo), CH * XBC::Sc, xstride<XBC, 1>(), m0, m1);
      col_store<XAB, PB>(E1, col_base<XAB, PB>(c, hi, lo), m0);
      col_store<XAB, PB>(E1, col_base<XAB, PB>(c + CH, hi, lo), m1);
    }
    block_sync(); // (4)
    // ---- phase A': back along axis 2 with the derivative matrix, W by bulk store -------------------------------------
    {
      T o0[N1], o1[N1];
      contract_in_x2(B.d, E1, col_base<XAB, PA>(c, hi, lo), CH * XAB::Sc, xstride<XAB, 2>(), o0, o1);
      col_store<Nat<N1>, PA>(E2, col_base<Nat<N1>, PA>(c, hi, lo), o0);
      col_store<Nat<N1>, PA>(E2, col_base<Nat<N1>, PA>(c + CH, hi, lo), o1);
      fence_proxy_async();
      block_sync(); // (5)
      if (tid == 0)
      {
        bulk_s2g(W + (size_t)tile * C * ND, smem_u32(E2), G::tile_bytes);
        bulk_commit();
      }
    }
  }
  if (tid == 0)
    bulk_wait_read0();
}

// ---- launch helpers --------------------------------------------------------------------------------------
inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <typename K>
int plan_blocks(K kernel, int threads, size_t smem, int& blocks_per_sm)
{
  int dev = 0, max_optin = 0;
  FEMTC_CUDA(cudaGetDevice(&dev));
  FEMTC_CUDA(cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  if ((size_t)max_optin < smem)
  {
    set_error("column kernel does not fit on an SM");
    return FEMTC_ERR_CUDA;
  }
  FEMTC_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  FEMTC_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
  FEMTC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, kernel, threads, smem));
  if (blocks_per_sm < 1)
  {
    set_error("column kernel does not fit on an SM");
    return FEMTC_ERR_CUDA;
  }
  return FEMTC_OK;
}

template <typename T, int N1, int C>
long long run_mass_col(const MassArgs<T>& a)
{
  using MC = MassCol<T, N1, C>;
  using G = ColGeom<T, N1, C>;
  const long long ntiles = (long long)(a.num_cells / C);
  if (ntiles == 0 || !al16(a.U) || !al16(a.W) || !al16(a.detJ))
    return 0;
  TabA<T, N1> B;
  fill_table<T, N1>(B.a, a.phi_host, false);
  fill_table<T, N1>(B.at, a.phi_host, true);
  auto kernel = mass_column_kernel<T, N1, C>;
  int bps = 0;
  int rc = plan_blocks(kernel, G::NT, MC::smem_bytes, bps);
  if (rc)
    return -rc;
  const int grid = (int)std::min<long long>(ntiles, (long long)sm_count() * bps);
  FEMTC_LAUNCH(FAMILY_COLUMN, 1, C, bps, kernel, grid, G::NT, MC::smem_bytes, a.stream, B, a.U, a.W, a.detJ, ntiles);
  return ntiles * C;
}

template <typename T, int N1, int C>
long long run_mass_col2(const MassArgs<T>& a)
{
  using MC = MassCol2<T, N1, C>;
  using H = MassCol2<T, N1, C>;
  const long long ntiles = (long long)(a.num_cells / C);
  if (ntiles == 0 || !al16(a.U) || !al16(a.W) || !al16(a.detJ))
    return 0;
  TabA<T, N1> B;
  fill_table<T, N1>(B.a, a.phi_host, false);
  fill_table<T, N1>(B.at, a.phi_host, true);
  auto kernel = mass_column2_kernel<T, N1, C>;
  int bps = 0;
  int rc = plan_blocks(kernel, H::NT, MC::smem_bytes, bps);
  if (rc)
    return -rc;
  const int grid = (int)std::min<long long>(ntiles, (long long)sm_count() * bps);
  FEMTC_LAUNCH(FAMILY_COLUMN, 2, C, bps, kernel, grid, H::NT, MC::smem_bytes, a.stream, B, a.U, a.W, a.detJ, ntiles);
  return ntiles * C;
}

template <typename T, int N1, int C, int MODE, int NG, bool PW>
long long run_stiff_col(const StiffnessArgs<T>& a)
{
  using SC = StiffCol<T, N1, C, MODE, NG, PW>;
  using G = ColGeom<T, N1, C>;
  const long long ntiles = (long long)(a.num_cells / C);
  if (ntiles == 0 || !al16(a.U) || !al16(a.W) || !al16(a.G))
    return 0;
  TabAD<T, N1> B;
  fill_table<T, N1>(B.a, a.A_host, false);
  fill_table<T, N1>(B.d, a.D_host, false);
  fill_table<T, N1>(B.at, a.A_host, true);
  fill_table<T, N1>(B.dt, a.D_host, true);
  auto kernel = stiffness_column_kernel<T, N1, C, MODE, NG, PW>;
  int bps = 0;
  int rc = plan_blocks(kernel, SC::threads, SC::smem_bytes, bps);
  if (rc)
    return -rc;
  const int grid = (int)std::min<long long>(ntiles, (long long)sm_count() * bps);
  FEMTC_LAUNCH(FAMILY_COLUMN, 1 + 10 * NG + (PW ? 1000 : 0), C, bps, kernel, grid, SC::threads, SC::smem_bytes, a.stream, B, a.U, a.W, a.G, ntiles);
  return ntiles * C;
}

template <typename T, int N1, int C, int NG>
long long run_stiff_col2(const StiffnessArgs<T>& a)
{
  using SC = StiffCol2<T, N1, C, NG>;
  const long long ntiles = (long long)(a.num_cells / C);
  if (ntiles == 0 || !al16(a.U) || !al16(a.W) || !al16(a.G))
    return 0;
  TabAD<T, N1> B;
  fill_table<T, N1>(B.a, a.A_host, false);
  fill_table<T, N1>(B.d, a.D_host, false);
  fill_table<T, N1>(B.at, a.A_host, true);
  fill_table<T, N1>(B.dt, a.D_host, true);
  auto kernel = stiffness_column2_kernel<T, N1, C, NG>;
  int bps = 0;
  int rc = plan_blocks(kernel, SC::threads, SC::smem_bytes, bps);
  if (rc)
    return -rc;
  const int grid = (int)std::min<long long>(ntiles, (long long)sm_count() * bps);
  FEMTC_LAUNCH(FAMILY_COLUMN, 2 + 10 * NG + 1000, C, bps, kernel, grid, SC::threads, SC::smem_bytes, a.stream, B, a.U, a.W, a.G, ntiles);
  return ntiles * C;
}
// two cells per thread: tile sizes per shape (fp32 tiles must be whole 16-byte units)
template <typename T, int N1>
constexpr int cells2()
{
  if (N1 == 9) return sizeof(T) == 8 ? 2 : 4;
  if (N1 == 7) return 4;
  return 8; // n1 = 6
}

// Geometry transport per shape (REF_EXACT): depth of the block-wide producer-warp ring, 0 = warp-private rings.
// Measured (profiles/r01_ab_producer_warp.txt, r01_ab_producer_warp_depth.txt; % of measured HBM, 1 M cells):
//   fp32 n1=6  84 (private) -> 85 (3) -> 96 (5) -> 101 (8)      fp32 n1=7  69 -> 70 -> 78 -> 80
//   fp32 n1=8  90 -> 91..92 (3) -> 88 (5) -> 66 (8)              fp32 n1=9  63 (private) vs 58..60: stays private
//   fp64 n1=6  85 -> 90..91 (3) -> 70 (5), private depth 4: 94   fp64 n1=7..9 and FULL mode: private is faster
// variant 3 / 4 force the depth-3 ring / the depth-2 private rings, 5 / 6 the depth-5 / depth-8 ring,
// 7 / 8 private rings of depth 4 / 6 (development).
template <typename T, int N1>
constexpr int default_pw_depth()
{
  if (sizeof(T) == 4)
    return N1 == 6 || N1 == 7 ? 8 : (N1 == 8 ? 3 : 0);
  return 0;
}
// depth of the warp-private rings where those serve (fp64 n1 = 6: 85 % at 2, 94 % at 4 -- better than the
// block-wide ring's 91 %; deeper rings cost a resident block everywhere else: profiles/r01_ab_private_ring_depth.txt)
template <typename T, int N1>
constexpr int default_private_depth()
{
  return sizeof(T) == 8 && N1 == 6 ? 4 : 2;
}
template <typename T, int N1, int C, int NG>
long long run_stiff_col_mode(const StiffnessArgs<T>& a)
{
  if (a.mode == FEMTC_STIFFNESS_FULL)
  {
    if constexpr (N1 >= 6)
      if (opt_variant() == 3)
        return run_stiff_col<T, N1, C, FEMTC_STIFFNESS_FULL, 3, true>(a);
    return run_stiff_col<T, N1, C, FEMTC_STIFFNESS_FULL, NG, false>(a);
  }
  if constexpr (N1 == 6 || N1 == 7 || N1 == 9)
  {
    // Two cells per thread, n1 = 9 (profiles/r01_ab_stiffness_two_cells.txt; % of measured HBM, one -> two cells):
    //   fp32 63 -> 66 (ring depth 3) -> 70 (5);  fp64 62 -> 68 (3) -> 64 (5).  n1 = 6, 7: within +-3 points, not used.
    // variant 13 / 14 / 15 force it with ring depth 3 / 5 / 8; variant 4 forces the one-cell kernel.
    const long long v = opt_variant();
    if (v == 13) return run_stiff_col2<T, N1, cells2<T, N1>(), 3>(a);
    if (v == 14) return run_stiff_col2<T, N1, cells2<T, N1>(), 5>(a);
    if (v == 15) return run_stiff_col2<T, N1, cells2<T, N1>(), 8>(a);
    if (N1 == 9 && (v == 0 || v == 1 || v == 2 || v == 9))
      return sizeof(T) == 4 ? run_stiff_col2<T, N1, cells2<T, N1>(), 5>(a) : run_stiff_col2<T, N1, cells2<T, N1>(), 3>(a);
  }
  if constexpr (N1 >= 6)
  {
    int depth = default_pw_depth<T, N1>();
    switch (opt_variant())
    {
    case 3: depth = 3; break;
    case 4: depth = 0; break;
    case 5: depth = 5; break;
    case 6: depth = 8; break;
    default: break;
    }
    if (opt_variant() == 7) return run_stiff_col<T, N1, C, FEMTC_STIFFNESS_REF_EXACT, 4, false>(a); // deeper private rings
    if (opt_variant() == 8) return run_stiff_col<T, N1, C, FEMTC_STIFFNESS_REF_EXACT, 6, false>(a);
    if (depth == 3) return run_stiff_col<T, N1, C, FEMTC_STIFFNESS_REF_EXACT, 3, true>(a);
    if (depth == 5) return run_stiff_col<T, N1, C, FEMTC_STIFFNESS_REF_EXACT, 5, true>(a);
    if (depth == 8) return run_stiff_col<T, N1, C, FEMTC_STIFFNESS_REF_EXACT, 8, true>(a);
  }
  if constexpr (N1 >= 6 && default_private_depth<T, N1>() == 4)
    if (opt_variant() != 4)
      return run_stiff_col<T, N1, C, FEMTC_STIFFNESS_REF_EXACT, 4, false>(a);
  return run_stiff_col<T, N1, C, FEMTC_STIFFNESS_REF_EXACT, NG, false>(a);
}

} // namespace

// Which (type, n1) the column family serves: everything the slab kernels do not.
template <typename T>
long long launch_mass_column(const MassArgs<T>& a)
{
  const int n1 = a.degree + 1;
  {
    // Two cells per thread.  Measured (profiles/r01_ab_mass_two_cells.txt, % of measured HBM, one -> two cells):
    //   fp64 n1 = 6: 78 -> 84 (tile of 14), n1 = 7: 58..64 -> 68, n1 = 9: 47 -> 57; n1 = 8 stays on DMMA (71).
    // variant 10 / 11 force it with the one-cell kernel's tile / twice that tile, variant 12 forces one cell.
    const long long v = opt_variant();
    constexpr bool F64 = sizeof(T) == 8;
    int two = 0; // 0 = one cell per thread, 1 = tile as the one-cell kernel, 2 = twice that
    if (F64)
      two = n1 == 6 ? 2 : (n1 == 7 || n1 == 9 ? 1 : 0);
    else
      two = n1 == 8 ? 1 : (n1 == 9 ? 2 : 0); // fp32: n1 = 8 63 -> 73, n1 = 9 57 -> 60; n1 = 6, 7 lose or tie
    if (v == 10) two = 1;
    if (v == 11) two = 2;
    if (v == 12) two = 0;
    if (v == 22 && n1 == 7) // development: one cell per thread, tile of 8 cells (392 of 416 lanes busy)
      return run_mass_col<T, 7, 8>(a);
    if (two == 1)
      switch (n1)
      {
      case 6: return run_mass_col2<T, 6, 8>(a);
      case 7: return run_mass_col2<T, 7, 4>(a);
      case 8: return run_mass_col2<T, 8, 2>(a);
      case 9: return run_mass_col2<T, 9, F64 ? 2 : 4>(a);
      default: break;
      }
    if (two == 2)
      switch (n1)
      {
      case 6: return run_mass_col2<T, 6, F64 ? 14 : 16>(a);
      case 7: return run_mass_col2<T, 7, 8>(a);
      case 8: return run_mass_col2<T, 8, 4>(a);
      case 9: return run_mass_col2<T, 9, F64 ? 4 : 8>(a);
      default: break;
      }
  }
  switch (n1)
  {
  case 6: return run_mass_col<T, 6, default_cells<T, 6>()>(a);
  case 7: return run_mass_col<T, 7, default_cells<T, 7>()>(a);
  case 8: return run_mass_col<T, 8, default_cells<T, 8>()>(a);
  case 9: return run_mass_col<T, 9, default_cells<T, 9>()>(a);
  default: break;
  }
  if constexpr (sizeof(T) == 4)
  {
    if (n1 == 3) return run_mass_col<T, 3, default_cells<T, 3>()>(a);
    if (n1 == 5) return run_mass_col<T, 5, default_cells<T, 5>()>(a);
  }
  return 0;
}

template <typename T>
long long launch_stiffness_column(const StiffnessArgs<T>& a)
{
  const int n1 = a.degree + 1;
  switch (n1)
  {
  case 2: return run_stiff_col_mode<T, 2, default_cells<T, 2>(), 2>(a);
  case 6: return run_stiff_col_mode<T, 6, default_cells<T, 6>(), 2>(a);
  case 7: return run_stiff_col_mode<T, 7, default_cells<T, 7>(), 2>(a);
  case 8: return run_stiff_col_mode<T, 8, default_cells<T, 8>(), 2>(a);
  case 9: return run_stiff_col_mode<T, 9, default_cells<T, 9>(), 2>(a);
  default: break;
  }
  if constexpr (sizeof(T) == 4)
  {
    if (n1 == 3) return run_stiff_col_mode<T, 3, default_cells<T, 3>(), 2>(a);
    if (n1 == 5) return run_stiff_col_mode<T, 5, default_cells<T, 5>(), 2>(a);
  }
  return 0;
}

template long long launch_mass_column<double>(const MassArgs<double>&);
template long long launch_mass_column<float>(const MassArgs<float>&);
template long long launch_stiffness_column<double>(const StiffnessArgs<double>&);
template long long launch_stiffness_column<float>(const StiffnessArgs<float>&);

} // namespace femtc
