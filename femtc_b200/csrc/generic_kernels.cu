// Generic kernels: any degree 1..25, any pointer alignment, any cell count.
//
// These serve (a) degrees 9..25, which the reference dispatches (mass.hpp:83-133) but which
// are outside the tuned range, (b) the ragged tail and unaligned inputs of the tiled
// TMA kernels (tiled_kernels.cu), and (c) the standalone small-GEMM / tensor-contraction
// entry points behind linalg::gemm_blocked, linalg::gemm_micro and apply_contractions.
// One thread block owns a group of cells; the per-axis contraction is written exactly as
// gemm_blocked defines it (linalg.hpp:101-182 with k = m = n1, n = n1^2):
//     Y[j1,j2,i] = sum_p M(i,p) X[p,j1,j2],   M(i,p) = M[i + p*n1]
// so three sweeps walk the three axes and come back to the natural index order.
#include "common.cuh"

#include <algorithm>
#include <mutex>
#include <set>
#include <utility>

namespace femtc
{
namespace
{

constexpr int kThreads = 256;

// Y[c][j*n1 + i] = sum_p M(i,p) * X[c][p*n2 + j]   for a group of `cells` cells.
// Optional fused epilogues:  * scale[c][idx]   or   (add0 + add1) + result.
// A thread owns IB = 4 consecutive outputs i of one column j: per p it loads X once and the four table entries as one
// or two 128-bit words (the table is staged with rows padded to a multiple of 4, M(i,p) = Mp[p * n1p + i]), i.e.
// 2-3 shared-memory loads per 4 FMAs instead of the 8 of the one-output-per-thread form (round 1) -- the sweep of
// degrees 9..25 and of every ragged tail.  p ascends exactly as before: results are bit-identical to round 1.
constexpr int kIB = 4;
__host__ __device__ constexpr int pad4(int n) { return (n + kIB - 1) / kIB * kIB; }

template <typename T>
__device__ __forceinline__ void sweep(const T* __restrict__ Mp, const T* X, size_t xstride, T* Y,
                                      size_t ystride, int n1, int cells, const T* scale = nullptr,
                                      size_t sstride = 0, const T* add0 = nullptr,
                                      const T* add1 = nullptr, size_t astride = 0)
{
  const int n2 = n1 * n1, nd = n2 * n1, n1p = pad4(n1), nib = n1p / kIB;
  const int per_cell = n2 * nib;
  for (int t = threadIdx.x; t < cells * per_cell; t += blockDim.x)
  {
    const int c = t / per_cell, r = t - c * per_cell;
    const int j = r / nib, i0 = (r - j * nib) * kIB;
    const T* x = X + (size_t)c * xstride + j;
    const T* m = Mp + i0;
    T acc[kIB] = {T(0), T(0), T(0), T(0)};
    for (int p = 0; p < n1; ++p)
    {
      const T xv = x[(size_t)p * n2];
      T mv[kIB];
      if constexpr (sizeof(T) == 4)
      {
        const float4 q = *reinterpret_cast<const float4*>(m + p * n1p);
        mv[0] = q.x, mv[1] = q.y, mv[2] = q.z, mv[3] = q.w;
      }
      else
      {
        const double2 q0 = *reinterpret_cast<const double2*>(m + p * n1p);
        const double2 q1 = *reinterpret_cast<const double2*>(m + p * n1p + 2);
        mv[0] = q0.x, mv[1] = q0.y, mv[2] = q1.x, mv[3] = q1.y;
      }
#pragma unroll
      for (int k = 0; k < kIB; ++k)
        acc[k] = fma_t(mv[k], xv, acc[k]);
    }
    const int base = j * n1 + i0;
#pragma unroll
    for (int k = 0; k < kIB; ++k)
      if (i0 + k < n1)
      {
        T v = acc[k];
        const int idx = base + k;
        if (scale)
          v = v * scale[(size_t)c * sstride + idx];
        if (add0)
          v = (add0[(size_t)c * astride + idx] + add1[(size_t)c * astride + idx]) + v;
        Y[(size_t)c * ystride + idx] = v;
      }
  }
  __syncthreads();
}

// table staging: M(i,p) = src[i + p*n1] -> dst[p*n1p + i], rows zero-padded to n1p = pad4(n1)
template <typename T>
__device__ __forceinline__ void load_matrix(T* dst, const T* __restrict__ src, int n1)
{
  const int n1p = pad4(n1);
  for (int t = threadIdx.x; t < n1 * n1p; t += blockDim.x)
  {
    const int p = t / n1p, i = t - p * n1p;
    dst[t] = i < n1 ? src[i + p * n1] : T(0);
  }
}

// The 1-D tables of the tuned degrees (n1 <= 9) travel as kernel parameters like in every other family: the ragged tail
// of a batch then costs one launch and nothing else (no stream-ordered allocation, no pageable host copy), and the whole
// operator call can be captured in a CUDA graph.  Higher degrees pass a device pointer instead (`phi` != null).
constexpr int kSmallN1 = 9;
template <typename T>
struct SmallTab
{
  T a[kSmallN1 * kSmallN1], d[kSmallN1 * kSmallN1];
};
template <typename T>
__device__ __forceinline__ void load_matrix_param(T* dst, const T* src, int n1)
{
  const int n1p = pad4(n1);
  for (int t = threadIdx.x; t < n1 * n1p; t += blockDim.x)
  {
    const int p = t / n1p, i = t - p * n1p;
    dst[t] = i < n1 ? src[i + p * n1] : T(0);
  }
}

// mass.hpp:10-49.  Scratch: 2 buffers of cpb*nd values, in shared memory when `ws` is null,
// else in the global workspace slice of this block.
template <typename T>
__global__ void __launch_bounds__(kThreads)
mass_generic_kernel(const T* __restrict__ phi, const __grid_constant__ SmallTab<T> tab, const T* __restrict__ U,
                    T* __restrict__ W, const T* __restrict__ detJ, size_t num_cells, int n1, int cpb, T* ws)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* A = reinterpret_cast<T*>(smem_raw);
  const int n2 = n1 * n1;
  const size_t nd = (size_t)n2 * n1;
  const int nt = n1 * pad4(n1); // staged table: rows padded to a multiple of 4 (16-byte aligned rows)
  T* b0 = ws ? ws + (size_t)blockIdx.x * 2 * cpb * nd : A + nt;
  T* b1 = b0 + (size_t)cpb * nd;
  if (phi)
    load_matrix(A, phi, n1);
  else
    load_matrix_param(A, tab.a, n1);
  __syncthreads();

  const size_t groups = (num_cells + cpb - 1) / cpb;
  for (size_t g = blockIdx.x; g < groups; g += gridDim.x)
  {
    const size_t cell0 = g * cpb;
    const int cells = (int)min((size_t)cpb, num_cells - cell0);
    const T* u = U + cell0 * nd;
    sweep(A, u, nd, b0, nd, n1, cells);                            // mass.hpp:36
    sweep(A, b0, nd, b1, nd, n1, cells);                           // :37
    sweep(A, b1, nd, b0, nd, n1, cells, detJ + cell0 * nd, nd);    // :38 and the scaling :40-41
    sweep(A, b0, nd, b1, nd, n1, cells);                           // :45 (same matrix)
    sweep(A, b1, nd, b0, nd, n1, cells);                           // :46
    sweep(A, b0, nd, W + cell0 * nd, nd, n1, cells);               // :47
  }
}

// stiffness.hpp:34-99.  Scratch: 5 buffers of cpb*nd values.
template <typename T>
__global__ void __launch_bounds__(kThreads)
stiffness_generic_kernel(const T* __restrict__ phiA, const T* __restrict__ phiD, const __grid_constant__ SmallTab<T> tab,
                         const T* __restrict__ U, T* __restrict__ W, const T* __restrict__ G,
                         size_t num_cells, int n1, int cpb, int mode, T* ws)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* A = reinterpret_cast<T*>(smem_raw);
  const int n2 = n1 * n1;
  const int nt = n1 * pad4(n1);
  T* D = A + nt;
  const size_t nd = (size_t)n2 * n1;
  const size_t bs = (size_t)cpb * nd;
  T* t0 = ws ? ws + (size_t)blockIdx.x * 5 * bs : D + nt;
  T* t1 = t0 + bs;
  T* wx = t1 + bs;
  T* wy = wx + bs;
  T* wz = wy + bs;
  if (phiA)
  {
    load_matrix(A, phiA, n1);
    load_matrix(D, phiD, n1);
  }
  else
  {
    load_matrix_param(A, tab.a, n1);
    load_matrix_param(D, tab.d, n1);
  }
  __syncthreads();

  const size_t groups = (num_cells + cpb - 1) / cpb;
  for (size_t g = blockIdx.x; g < groups; g += gridDim.x)
  {
    const size_t cell0 = g * cpb;
    const int cells = (int)min((size_t)cpb, num_cells - cell0);
    const T* u = U + cell0 * nd;
    T* w = W + cell0 * nd;

    sweep(D, u, nd, t0, nd, n1, cells); // stiffness.hpp:68-70   Wx = (D x A x A) U
    sweep(A, t0, nd, t1, nd, n1, cells);
    sweep(A, t1, nd, wx, nd, n1, cells);
    sweep(A, u, nd, t0, nd, n1, cells); // :72-74                Wy = (A x D x A) U
    sweep(D, t0, nd, t1, nd, n1, cells);
    sweep(A, t1, nd, wy, nd, n1, cells);
    sweep(A, t0, nd, t1, nd, n1, cells); // :76-77 (REF_EXACT: no derivative)
    sweep(mode == FEMTC_STIFFNESS_FULL ? D : A, t1, nd, wz, nd, n1, cells);

    // stiffness.hpp:14-29
    const T* g6 = G + cell0 * nd * 6;
    for (int t = threadIdx.x; t < cells * (int)nd; t += blockDim.x)
    {
      const T* gq = g6 + (size_t)t * 6;
      const T w0 = wx[t], w1 = wy[t], w2 = wz[t];
      wx[t] = (gq[5] * w0 + gq[4] * w1 + gq[2] * w2);
      wy[t] = (gq[4] * w0 + gq[3] * w1 + gq[1] * w2);
      wz[t] = (gq[2] * w0 + gq[1] * w1 + gq[0] * w2);
    }
    __syncthreads();

    if (mode == FEMTC_STIFFNESS_REF_EXACT)
    {
      // :95-97 is the only back-contraction whose result survives (each one assigns W)
      sweep(A, wz, nd, t0, nd, n1, cells);
      sweep(A, t0, nd, t1, nd, n1, cells);
      sweep(D, t1, nd, w, nd, n1, cells);
    }
    else
    {
      sweep(D, wx, nd, t0, nd, n1, cells);
      sweep(A, t0, nd, t1, nd, n1, cells);
      sweep(A, t1, nd, wx, nd, n1, cells);
      sweep(A, wy, nd, t0, nd, n1, cells);
      sweep(D, t0, nd, t1, nd, n1, cells);
      sweep(A, t1, nd, wy, nd, n1, cells);
      sweep(A, wz, nd, t0, nd, n1, cells);
      sweep(A, t0, nd, t1, nd, n1, cells);
      sweep(D, t1, nd, w, nd, n1, cells, (const T*)nullptr, 0, wx, wy, nd);
    }
  }
}

// linalg.hpp:101-182: C[i + j*m] = sum_p A[i + p*m] * B[p*n + j]  (C overwritten)
template <typename T>
__global__ void gemm_blocked_kernel(int k, int m, int n, const T* __restrict__ A,
                                    const T* __restrict__ B, T* __restrict__ C)
{
  const size_t total = (size_t)m * n;
  for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total;
       t += (size_t)gridDim.x * blockDim.x)
  {
    const int j = (int)(t / m), i = (int)(t - (size_t)j * m);
    T acc = T(0);
    for (int p = 0; p < k; ++p)
      acc = fma_t(A[i + (size_t)p * m], B[(size_t)p * n + j], acc);
    C[t] = acc;
  }
}

// linalg.hpp:45-94 in strided form: c[i*ldc + j] += sum_p a[i*sai + p*sap] * b[p*sbp + j*sbj]
// (accumulates, p ascending).  gemm_micro is sai=1, sap=lda, sbp=ldb, sbj=1; tmm.hpp's `contract`
// variants (tmm.hpp:24-32, :65-85) are other stride choices.
template <typename T>
__global__ void gemm_strided_kernel(int k, int m, int nc, int sai, int sap, int sbp, int sbj, int ldc,
                                    const T* __restrict__ a, const T* __restrict__ b, T* c)
{
  const size_t total = (size_t)m * nc;
  for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total;
       t += (size_t)gridDim.x * blockDim.x)
  {
    const int i = (int)(t / nc), j = (int)(t - (size_t)i * nc);
    T acc = c[(size_t)i * ldc + j];
    for (int p = 0; p < k; ++p)
      acc = fma_t(a[(size_t)i * sai + (size_t)p * sap], b[(size_t)p * sbp + (size_t)j * sbj], acc);
    c[(size_t)i * ldc + j] = acc;
  }
}

// tmm.hpp:91-126, one tensor per block iteration; scratch s0 [Na,Nb,Nb], s1 [Na,Na,Nb].
template <typename T>
__global__ void __launch_bounds__(kThreads)
apply_contractions_kernel(int Na, int Nb, int Tr, const T* __restrict__ I0,
                          const T* __restrict__ I1, const T* __restrict__ I2,
                          const T* __restrict__ B, T* __restrict__ A, size_t batch, T* ws)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const size_t na = Na, nb = Nb;
  const size_t n0 = na * nb * nb, n1 = na * na * nb, n3 = na * na * na;
  T* s0 = ws ? ws + (size_t)blockIdx.x * (n0 + n1) : reinterpret_cast<T*>(smem_raw);
  T* s1 = s0 + n0;
  auto mat = [&](const T* I, size_t a, size_t b) { return Tr ? I[a * nb + b] : I[b * na + a]; };
  for (size_t e = blockIdx.x; e < batch; e += gridDim.x)
  {
    const T* Be = B + e * nb * nb * nb;
    T* Ae = A + e * n3;
    for (size_t t = threadIdx.x; t < n0; t += blockDim.x) // s0[a0,b1,b2]
    {
      const size_t a0 = t / (nb * nb), d = t - a0 * nb * nb;
      T acc = T(0);
      for (size_t b0 = 0; b0 < nb; ++b0)
        acc = fma_t(mat(I0, a0, b0), Be[b0 * nb * nb + d], acc);
      s0[t] = acc;
    }
    __syncthreads();
    for (size_t t = threadIdx.x; t < n1; t += blockDim.x) // s1[a1,a0,b2]
    {
      const size_t a1 = t / (na * nb), r = t - a1 * na * nb, a0 = r / nb, b2 = r - a0 * nb;
      T acc = T(0);
      for (size_t b1 = 0; b1 < nb; ++b1)
        acc = fma_t(mat(I1, a1, b1), s0[(a0 * nb + b1) * nb + b2], acc);
      s1[t] = acc;
    }
    __syncthreads();
    for (size_t t = threadIdx.x; t < n3; t += blockDim.x) // A[a0,a1,a2]
    {
      const size_t a0 = t / (na * na), r = t - a0 * na * na, a1 = r / na, a2 = r - a1 * na;
      T acc = T(0);
      for (size_t b2 = 0; b2 < nb; ++b2)
        acc = fma_t(mat(I2, a2, b2), s1[(a1 * na + a0) * nb + b2], acc);
      Ae[t] = acc;
    }
    __syncthreads();
  }
}

// ---- checksum: deterministic two-stage reduction ----------------------------------------------
constexpr int kChecksumBlocks = kChecksumStage1Triples; // 148 SMs x 8
// per device: [0, kChecksumFused) triples written by operator kernels with a fused checksum epilogue (one per warp),
// then kChecksumBlocks triples of the stand-alone first stage
__device__ double g_checksum_partials[(kChecksumFused + kChecksumBlocks) * 3];

template <typename T>
__global__ void __launch_bounds__(256) checksum_stage1(const T* __restrict__ x, size_t n, double* partials)
{
  double s = 0, s2 = 0, mx = 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
  {
    const double v = (double)x[i];
    s += v;
    s2 = fma(v, v, s2);
    mx = fmax(mx, fabs(v));
  }
  __shared__ double sh[3][256];
  sh[0][threadIdx.x] = s;
  sh[1][threadIdx.x] = s2;
  sh[2][threadIdx.x] = mx;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1)
  {
    if ((int)threadIdx.x < w)
    {
      sh[0][threadIdx.x] += sh[0][threadIdx.x + w];
      sh[1][threadIdx.x] += sh[1][threadIdx.x + w];
      sh[2][threadIdx.x] = fmax(sh[2][threadIdx.x], sh[2][threadIdx.x + w]);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0)
  {
    partials[blockIdx.x * 3 + 0] = sh[0][0];
    partials[blockIdx.x * 3 + 1] = sh[1][0];
    partials[blockIdx.x * 3 + 2] = sh[2][0];
  }
}

__global__ void __launch_bounds__(256) checksum_stage2(const double* partials, int nblocks, size_t n, double* out4)
{
  __shared__ double sh[3][256];
  double s = 0, s2 = 0, mx = 0;
  for (int i = threadIdx.x; i < nblocks; i += 256)
  {
    s += partials[i * 3 + 0];
    s2 += partials[i * 3 + 1];
    mx = fmax(mx, partials[i * 3 + 2]);
  }
  sh[0][threadIdx.x] = s;
  sh[1][threadIdx.x] = s2;
  sh[2][threadIdx.x] = mx;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1)
  {
    if ((int)threadIdx.x < w)
    {
      sh[0][threadIdx.x] += sh[0][threadIdx.x + w];
      sh[1][threadIdx.x] += sh[1][threadIdx.x + w];
      sh[2][threadIdx.x] = fmax(sh[2][threadIdx.x], sh[2][threadIdx.x + w]);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0)
  {
    out4[0] = sh[0][0];
    out4[1] = sh[1][0];
    out4[2] = sh[2][0];
    out4[3] = (double)n;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
fill_uniform_kernel(T* x, size_t n, uint64_t base, int centred)
{
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
  {
    const uint64_t r = splitmix64(base + i);
    T v;
    if constexpr (sizeof(T) == 8)
      v = (double)(r >> 11) * 0x1.0p-53;
    else
      v = (float)(r >> 40) * 0x1.0p-24f;
    x[i] = centred ? v - T(0.5) : v;
  }
}

// shared-memory budget we are willing to request for the generic kernels
constexpr size_t kGenericSmemMax = 160 * 1024;

// The opt-in ceiling is a property of (kernel, device): it is raised ONCE to the largest size any degree asks for,
// so two host threads launching different degrees never interleave an attribute change with a launch.
template <typename K>
int opt_in_smem(K kernel)
{
  static std::mutex mu;
  static std::set<std::pair<const void*, int>> done;
  int dev = 0;
  FEMTC_CUDA(cudaGetDevice(&dev));
  const auto key = std::make_pair(reinterpret_cast<const void*>(kernel), dev);
  std::lock_guard<std::mutex> lock(mu);
  if (done.count(key))
    return FEMTC_OK;
  FEMTC_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGenericSmemMax));
  done.insert(key);
  return FEMTC_OK;
}

// stream-ordered scratch that is handed back on every exit path
struct AsyncBuf
{
  void* p = nullptr;
  cudaStream_t s = nullptr;
  AsyncBuf() = default;
  AsyncBuf(const AsyncBuf&) = delete;
  AsyncBuf& operator=(const AsyncBuf&) = delete;
  ~AsyncBuf()
  {
    if (p && cudaFreeAsync(p, s) != cudaSuccess)
      (void)cudaGetLastError();
  }
  int alloc(size_t bytes, cudaStream_t stream)
  {
    s = stream;
    FEMTC_CUDA(cudaMallocAsync(&p, bytes, stream));
    return FEMTC_OK;
  }
  template <typename T>
  T* as() const { return static_cast<T*>(p); }
};

} // namespace

// ---- launchers -------------------------------------------------------------------------------------

template <typename T>
int launch_mass_generic(const MassArgs<T>& a)
{
  if (a.num_cells == 0)
    return FEMTC_OK;
  const int n1 = a.degree + 1, n2 = n1 * n1;
  const size_t nd = (size_t)n2 * n1;
  int cpb = (int)max((size_t)1, (size_t)kThreads / nd);
  const size_t nt = (size_t)n1 * pad4(n1);
  size_t smem = (nt + 2 * cpb * nd) * sizeof(T);
  const size_t groups = (a.num_cells + cpb - 1) / cpb;
  int grid = (int)min(groups, (size_t)sm_count() * 8);
  const bool spill = smem > kGenericSmemMax; // the per-cell work arrays live in global scratch instead
  if (spill)
  {
    cpb = 1;
    smem = nt * sizeof(T);
    grid = (int)min(a.num_cells, (size_t)sm_count() * 4);
  }
  note_launch(FAMILY_GENERIC, spill, mass_generic_kernel<T>, cpb, 0, grid, kThreads, smem);
  if (dry_run())
    return FEMTC_OK;
  AsyncBuf dphi, ws;
  SmallTab<T> tab;
  int rc = FEMTC_OK;
  if (n1 <= kSmallN1)
    std::copy(a.phi_host, a.phi_host + n2, tab.a);
  else
  {
    if ((rc = dphi.alloc(n2 * sizeof(T), a.stream)))
      return rc;
    FEMTC_CUDA(cudaMemcpyAsync(dphi.p, a.phi_host, n2 * sizeof(T), cudaMemcpyHostToDevice, a.stream));
  }
  if (spill && (rc = ws.alloc((size_t)grid * 2 * nd * sizeof(T), a.stream)))
    return rc;
  if ((rc = opt_in_smem(mass_generic_kernel<T>)))
    return rc;
  mass_generic_kernel<T><<<grid, kThreads, smem, a.stream>>>(dphi.as<T>(), tab, a.U, a.W, a.detJ, a.num_cells, n1, cpb,
                                                             ws.as<T>());
  count_launch();
  FEMTC_CUDA(cudaGetLastError());
  return FEMTC_OK;
}

template <typename T>
int launch_stiffness_generic(const StiffnessArgs<T>& a)
{
  if (a.num_cells == 0)
    return FEMTC_OK;
  const int n1 = a.degree + 1, n2 = n1 * n1;
  const size_t nd = (size_t)n2 * n1;
  int cpb = (int)max((size_t)1, (size_t)kThreads / nd);
  const size_t nt = (size_t)n1 * pad4(n1);
  size_t smem = (2 * nt + 5 * cpb * nd) * sizeof(T);
  const size_t groups = (a.num_cells + cpb - 1) / cpb;
  int grid = (int)min(groups, (size_t)sm_count() * 8);
  const bool spill = smem > kGenericSmemMax;
  if (spill)
  {
    cpb = 1;
    smem = 2 * nt * sizeof(T);
    grid = (int)min(a.num_cells, (size_t)sm_count() * 4);
  }
  note_launch(FAMILY_GENERIC, spill, stiffness_generic_kernel<T>, cpb, 0, grid, kThreads, smem);
  if (dry_run())
    return FEMTC_OK;
  AsyncBuf dphi, ws;
  SmallTab<T> tab;
  int rc = FEMTC_OK;
  if (n1 <= kSmallN1)
  {
    std::copy(a.A_host, a.A_host + n2, tab.a);
    std::copy(a.D_host, a.D_host + n2, tab.d);
  }
  else
  {
    if ((rc = dphi.alloc(2 * n2 * sizeof(T), a.stream)))
      return rc;
    FEMTC_CUDA(cudaMemcpyAsync(dphi.p, a.A_host, n2 * sizeof(T), cudaMemcpyHostToDevice, a.stream));
    FEMTC_CUDA(cudaMemcpyAsync(dphi.as<T>() + n2, a.D_host, n2 * sizeof(T), cudaMemcpyHostToDevice, a.stream));
  }
  if (spill && (rc = ws.alloc((size_t)grid * 5 * nd * sizeof(T), a.stream)))
    return rc;
  if ((rc = opt_in_smem(stiffness_generic_kernel<T>)))
    return rc;
  stiffness_generic_kernel<T><<<grid, kThreads, smem, a.stream>>>(dphi.as<T>(), dphi.p ? dphi.as<T>() + n2 : nullptr, tab,
                                                                  a.U, a.W, a.G, a.num_cells, n1, cpb, a.mode, ws.as<T>());
  count_launch();
  FEMTC_CUDA(cudaGetLastError());
  return FEMTC_OK;
}

template <typename T>
int launch_gemm_blocked(int k, int m, int n, const T* dA, const T* dB, T* dC, cudaStream_t s)
{
  const size_t total = (size_t)m * n;
  const int grid = (int)min((total + 255) / 256, (size_t)sm_count() * 8);
  gemm_blocked_kernel<T><<<grid, 256, 0, s>>>(k, m, n, dA, dB, dC);
  count_launch();
  FEMTC_CUDA(cudaGetLastError());
  return FEMTC_OK;
}

template <typename T>
int launch_gemm_strided(int k, int m, int nc, int sai, int sap, int sbp, int sbj, int ldc,
                        const T* dA, const T* dB, T* dC, cudaStream_t s)
{
  const size_t total = (size_t)m * nc;
  const int grid = (int)min((total + 255) / 256, (size_t)sm_count() * 8);
  gemm_strided_kernel<T><<<grid, 256, 0, s>>>(k, m, nc, sai, sap, sbp, sbj, ldc, dA, dB, dC);
  count_launch();
  FEMTC_CUDA(cudaGetLastError());
  return FEMTC_OK;
}

template <typename T>
int launch_apply_contractions(int Na, int Nb, int Tr, const T* dI0, const T* dI1, const T* dI2,
                              const T* dB, T* dA, size_t batch, cudaStream_t s)
{
  if (batch == 0)
    return FEMTC_OK;
  const size_t na = Na, nb = Nb;
  size_t smem = (na * nb * nb + na * na * nb) * sizeof(T);
  int grid = (int)min(batch, (size_t)sm_count() * 8);
  AsyncBuf ws;
  int rc = FEMTC_OK;
  if (smem > kGenericSmemMax)
  {
    if ((rc = ws.alloc((size_t)grid * smem, s)))
      return rc;
    smem = 0;
  }
  if ((rc = opt_in_smem(apply_contractions_kernel<T>)))
    return rc;
  apply_contractions_kernel<T><<<grid, kThreads, smem, s>>>(Na, Nb, Tr, dI0, dI1, dI2, dB, dA, batch, ws.as<T>());
  count_launch();
  FEMTC_CUDA(cudaGetLastError());
  return FEMTC_OK;
}

// The partials buffer is per device and reused: calls on one device must be stream-ordered with respect to each other
// (they are in every use this library makes of it).
int checksum_scratch(double** partials)
{
  FEMTC_CUDA(cudaGetSymbolAddress((void**)partials, g_checksum_partials));
  return FEMTC_OK;
}

template <typename T>
int launch_checksum_stage1(const T* dX, size_t n, double* partials, cudaStream_t s)
{
  checksum_stage1<T><<<kChecksumBlocks, 256, 0, s>>>(dX, n, partials);
  count_launch();
  FEMTC_CUDA(cudaGetLastError());
  return FEMTC_OK;
}

int launch_checksum_stage2(const double* partials, int ntriples, size_t n, double* d_out4, cudaStream_t s)
{
  checksum_stage2<<<1, 256, 0, s>>>(partials, ntriples, n, d_out4);
  count_launch();
  FEMTC_CUDA(cudaGetLastError());
  return FEMTC_OK;
}

template <typename T>
int launch_checksum(const T* dX, size_t n, double* d_out4, cudaStream_t s)
{
  double* partials = nullptr;
  int rc = checksum_scratch(&partials);
  if (rc)
    return rc;
  partials += 3 * kChecksumFused;
  if ((rc = launch_checksum_stage1<T>(dX, n, partials, s)))
    return rc;
  return launch_checksum_stage2(partials, kChecksumBlocks, n, d_out4, s);
}

template <typename T>
int launch_fill_uniform(T* dX, size_t n, uint64_t seed, uint64_t stream_id, uint64_t offset,
                        int centred, cudaStream_t s)
{
  if (n == 0)
    return FEMTC_OK;
  const uint64_t base = seed + stream_id * 0x9E3779B97F4A7C15ull + offset;
  const int grid = (int)min((n + 255) / 256, (size_t)sm_count() * 16);
  fill_uniform_kernel<T><<<grid, 256, 0, s>>>(dX, n, base, centred);
  count_launch();
  FEMTC_CUDA(cudaGetLastError());
  return FEMTC_OK;
}

// ---- streaming probe: the operators' traffic mix without their arithmetic --------------------------------
// out[i] = in_0[i] + ... + in_{R-1}[i] over 16-byte words, streams back to back in `in`.  R reads : 1 write per word is
// the mix of the stiffness (7:1) and mass (2:1) action; a device copy is 1:1.  bench.py times it as the roofline
// denominator for a read-dominated stream next to the copy bandwidth of MEASURED_PEAKS.json.
template <int R>
__global__ void __launch_bounds__(256) stream_probe_kernel(const float4* __restrict__ in, float4* __restrict__ out,
                                                           size_t n)
{
  const size_t stride = (size_t)gridDim.x * 256;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += stride)
  {
    float4 v[R];
#pragma unroll
    for (int r = 0; r < R; ++r)
      v[r] = __ldcs(in + (size_t)r * n + i);
    float4 acc = v[0];
#pragma unroll
    for (int r = 1; r < R; ++r)
      acc.x += v[r].x, acc.y += v[r].y, acc.z += v[r].z, acc.w += v[r].w;
    __stcs(out + i, acc);
  }
}

int launch_stream_probe(const void* d_in, void* d_out, size_t bytes_per_stream, int reads, cudaStream_t s)
{
  const size_t n = bytes_per_stream / 16;
  if (n == 0)
    return FEMTC_OK;
  const int grid = (int)min((n + 255) / 256, (size_t)sm_count() * 8);
  const float4* in = static_cast<const float4*>(d_in);
  float4* out = static_cast<float4*>(d_out);
  switch (reads)
  {
  case 1: stream_probe_kernel<1><<<grid, 256, 0, s>>>(in, out, n); break;
  case 2: stream_probe_kernel<2><<<grid, 256, 0, s>>>(in, out, n); break;
  case 7: stream_probe_kernel<7><<<grid, 256, 0, s>>>(in, out, n); break;
  default: set_error("stream probe: reads must be 1, 2 or 7"); return FEMTC_ERR_ARG;
  }
  count_launch();
  FEMTC_CUDA(cudaGetLastError());
  return FEMTC_OK;
}

// ---- apply_contractions for callers that cannot allocate device memory -------------------------------------------
// launch_apply_contractions above takes its three matrices and (for large shapes) its per-block work arrays from
// cudaMallocAsync.  Some users of this library cannot afford that: a solver that handed the whole of HBM to its own arena
// at start-up (an allocation made behind its back fails, or eats memory the arena has promised to someone else), and a
// time-stepping loop that must not meet the device-wide synchronisation cudaMalloc / cudaFree may imply.  For them the
// same kernel runs with every scratch byte in the array below: it belongs to the library image, so each GPU that loads
// libfemtc_b200.so carries these 4 MiB from load to exit, and nothing is requested from the allocator per call.
//
// Layout of a call:  [ I0 | I1 | I2 ]  3*Na*Nb values, rounded up to 256 bytes, then one (Na*Nb*Nb + Na*Na*Nb)-value
// work slice per block.  Every byte a call reads it has written itself (the matrices by the copies below, a work slice by
// the sweep before the one that reads it), so whatever the previous call left behind is harmless and nothing has to be
// cleared.  There is ONE array per device: two calls that run on the GPU at the same time (different streams, no
// dependency) overwrite each other's matrices and work slices.
constexpr size_t kStaticScratchBytes = size_t(4) << 20;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

// `g_static_scratch` in host code is only a handle for the runtime's symbol table, not an address the GPU understands;
// the device address differs from device to device and is asked for once per device, with that device current.
constexpr int kStaticMaxDevices = 64;
static std::mutex g_static_mu;
static unsigned char* g_static_addr[kStaticMaxDevices];

static int static_scratch_address(unsigned char** out)
{
  int dev = 0;
  FEMTC_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kStaticMaxDevices)
  {
    set_error("static scratch: device ordinal beyond 63");
    return FEMTC_ERR_ARG;
  }
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_addr[dev])
  {
    void* p = nullptr;
    FEMTC_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
    g_static_addr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_static_addr[dev];
  return FEMTC_OK;
}

// I0..I2 host pointers, dB / dA device pointers.  No cudaMalloc*, cudaFree* or other allocating call on this path; the
// first call on a device looks the array's address up (static_scratch_address).
int launch_apply_contractions_static(int Na, int Nb, int Tr, const double* I0, const double* I1, const double* I2,
                                     const double* dB, double* dA, size_t batch, cudaStream_t s)
{
  using T = double;
  const size_t na = Na, nb = Nb, nm = na * nb;
  const size_t tables = (3 * nm * sizeof(T) + 255) & ~size_t(255);
  const size_t slice = (na * nb * nb + na * na * nb) * sizeof(T);
  if (tables + slice > kStaticScratchBytes)
  {
    char msg[224];
    std::snprintf(msg, sizeof msg,
                  "apply_contractions (no-allocation entry): Na = %d, Nb = %d needs %zu bytes of scratch, the built-in "
                  "array holds %zu; use femtc_apply_contractions_device_f64 for this shape",
                  Na, Nb, tables + slice, kStaticScratchBytes);
    set_error(msg);
    return FEMTC_ERR_ARG;
  }
  if (batch == 0)
    return FEMTC_OK;
  unsigned char* base = nullptr;
  int rc = static_scratch_address(&base);
  if (rc)
    return rc;
  T* dI = reinterpret_cast<T*>(base);
  T* ws = reinterpret_cast<T*>(base + tables);
  const size_t fit = (kStaticScratchBytes - tables) / slice; // work slices the array has room for
  const int grid = (int)min(min(batch, fit), (size_t)sm_count() * 8);
  FEMTC_CUDA(cudaMemcpyAsync(dI, I0, nm * sizeof(T), cudaMemcpyHostToDevice, s));
  FEMTC_CUDA(cudaMemcpyAsync(dI + nm, I1, nm * sizeof(T), cudaMemcpyHostToDevice, s));
  FEMTC_CUDA(cudaMemcpyAsync(dI + 2 * nm, I2, nm * sizeof(T), cudaMemcpyHostToDevice, s));
  apply_contractions_kernel<T><<<grid, kThreads, 0, s>>>(Na, Nb, Tr, dI, dI + nm, dI + 2 * nm, dB, dA, batch, ws);
  count_launch();
  FEMTC_CUDA(cudaGetLastError());
  return FEMTC_OK;
}

#define INSTANTIATE(T)                                                                             \
  template int launch_mass_generic<T>(const MassArgs<T>&);                                     \
  template int launch_stiffness_generic<T>(const StiffnessArgs<T>&);                               \
  template int launch_gemm_blocked<T>(int, int, int, const T*, const T*, T*, cudaStream_t);        \
  template int launch_gemm_strided<T>(int, int, int, int, int, int, int, int, const T*, const T*,  \
                                      T*, cudaStream_t);                                           \
  template int launch_apply_contractions<T>(int, int, int, const T*, const T*, const T*, const T*, \
                                            T*, size_t, cudaStream_t);                             \
  template int launch_checksum<T>(const T*, size_t, double*, cudaStream_t);                        \
  template int launch_checksum_stage1<T>(const T*, size_t, double*, cudaStream_t);                 \
  template int launch_fill_uniform<T>(T*, size_t, uint64_t, uint64_t, uint64_t, int, cudaStream_t);
INSTANTIATE(double)
INSTANTIATE(float)

} // namespace femtc
