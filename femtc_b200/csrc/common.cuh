// Shared device/host helpers for libfemtc_b200.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <string>

#include "../../include/femtc_b200.h"

namespace femtc
{

// ---- error plumbing -------------------------------------------------------------------
void set_error(const std::string& msg);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define FEMTC_CUDA(call)                                                                          \
  do                                                                                              \
  {                                                                                               \
    cudaError_t e__ = (call);                                                                     \
    if (e__ != cudaSuccess)                                                                       \
      return ::femtc::cuda_fail(e__, #call, __FILE__, __LINE__);                                  \
  } while (0)

void count_launch(int n = 1);

// ---- launch record (femtc_kernel_info) -------------------------------------------------------------
// Every operator launcher goes through FEMTC_LAUNCH: it notes what is about to run in a thread-local record and,
// unless the calling thread is in a dry run (femtc_kernel_info walks the dispatcher that way), launches it.
enum KernelFamily
{
  FAMILY_GENERIC = 0,
  FAMILY_SLAB = 1,   // tiled_kernels.cu
  FAMILY_CELL = 2,   // cell_kernels.cu
  FAMILY_COLUMN = 3, // column_kernels.cu
  FAMILY_TC = 4      // tc_kernels.cu (DMMA) / tf32_kernels.cu
};
struct LaunchRecord
{
  int family = 0, cells_per_tile = 0, threads = 0, blocks_per_sm = 0, regs = 0, detail = 0;
  long long smem = 0, grid = 0;
};
bool dry_run();
LaunchRecord& launch_record();
template <typename K>
inline void note_launch(int family, int detail, K kernel, int cells_per_tile, int blocks_per_sm, long long grid,
                        int threads, size_t smem)
{
  LaunchRecord& r = launch_record();
  r.family = family, r.detail = detail, r.cells_per_tile = cells_per_tile, r.blocks_per_sm = blocks_per_sm;
  r.grid = grid, r.threads = threads, r.smem = (long long)smem, r.regs = 0;
  if (dry_run())
  {
    cudaFuncAttributes fa{};
    if (cudaFuncGetAttributes(&fa, kernel) == cudaSuccess)
      r.regs = fa.numRegs;
    else
      (void)cudaGetLastError();
  }
}
// `detail`: family-specific variant code for reports (cells per thread, ring depth, ...)
#define FEMTC_LAUNCH(family, detail, cells_per_tile, bps, kernel, grid, threads, smem, stream, ...)             \
  do                                                                                                            \
  {                                                                                                             \
    ::femtc::note_launch(family, detail, kernel, cells_per_tile, bps, grid, threads, smem);                     \
    if (!::femtc::dry_run())                                                                                    \
    {                                                                                                           \
      kernel<<<grid, threads, smem, stream>>>(__VA_ARGS__);                                                     \
      ::femtc::count_launch();                                                                                  \
      const cudaError_t le__ = cudaGetLastError();                                                              \
      if (le__ != cudaSuccess)                                                                                  \
        return -::femtc::cuda_fail(le__, "launch of " #kernel, __FILE__, __LINE__);                             \
    }                                                                                                           \
  } while (0)

// ---- problem description handed to the launchers ----------------------------------------
template <typename T>
struct MassArgs
{
  const T* phi_host; // n1*n1 values, host
  const T* U;
  T* W;
  const T* detJ;
  size_t num_cells;
  int degree;
  cudaStream_t stream;
  // checksum fused into the epilogue (femtc_*_apply_device_ex): a launcher whose kernel family supports it leaves one
  // {sum, sum of squares, max |w|} triple per warp in cs_partials (room for cs_capacity triples) and their number in
  // *cs_count; a family that does not leaves *cs_count at 0 and the caller reduces W with the stand-alone kernel
  double* cs_partials = nullptr;
  int cs_capacity = 0;
  int* cs_count = nullptr;
};

template <typename T>
struct StiffnessArgs
{
  const T* A_host; // n1*n1 values, host
  const T* D_host; // n1*n1 values, host (== A_host under the short-phi policy)
  const T* U;
  T* W;
  const T* G;
  size_t num_cells;
  int degree;
  int mode;
  cudaStream_t stream;
  double* cs_partials = nullptr; // see MassArgs
  int cs_capacity = 0;
  int* cs_count = nullptr;
};

// generic_kernels.cu -- any degree 1..25, any alignment
template <typename T>
int launch_mass_generic(const MassArgs<T>& a);
template <typename T>
int launch_stiffness_generic(const StiffnessArgs<T>& a);
template <typename T>
int launch_gemm_blocked(int k, int m, int n, const T* dA, const T* dB, T* dC, cudaStream_t s);
template <typename T>
int launch_gemm_strided(int k, int m, int nc, int sai, int sap, int sbp, int sbj, int ldc,
                        const T* dA, const T* dB, T* dC, cudaStream_t s);
template <typename T>
int launch_apply_contractions(int Na, int Nb, int Tr, const T* dI0, const T* dI1, const T* dI2,
                              const T* dB, T* dA, size_t batch, cudaStream_t s);
template <typename T>
int launch_checksum(const T* dX, size_t n, double* d_out4, cudaStream_t s);
// pieces of it, for the operators' fused checksum (api.cu): the per-device scratch holds kChecksumFused triples for
// operator kernels, then the kChecksumBlocks triples launch_checksum_stage1 writes; stage 2 sums any number of triples
constexpr int kChecksumFused = 16384;
constexpr int kChecksumStage1Triples = 1184;
int checksum_scratch(double** partials);
template <typename T>
int launch_checksum_stage1(const T* dX, size_t n, double* partials, cudaStream_t s);
int launch_checksum_stage2(const double* partials, int ntriples, size_t n, double* d_out4, cudaStream_t s);
template <typename T>
int launch_fill_uniform(T* dX, size_t n, uint64_t seed, uint64_t stream_id, uint64_t offset,
                        int centred, cudaStream_t s);

// tf32_kernels.cu -- tcgen05.mma kind::tf32 arm (fp32 mass n1 = 8), selected with option tc = 1; same convention
long long launch_mass_tf32(const MassArgs<float>& a);
int launch_stream_probe(const void* d_in, void* d_out, size_t bytes_per_stream, int reads, cudaStream_t s);
// generic_kernels.cu -- the general contraction kernel with all of its scratch in a file-scope __device__ array (no
// allocation on the call path); I0..I2 host pointers
int launch_apply_contractions_static(int Na, int Nb, int Tr, const double* I0, const double* I1, const double* I2,
                                     const double* dB, double* dA, size_t batch, cudaStream_t s);

// tiled_kernels.cu -- degrees 1..8, TMA-staged warp tiles.  Return value:
//   >= 0 : number of leading cells processed on the fast path (the caller runs the generic
//          kernel on the remainder); < 0 : -(FEMTC_ERR_*).
template <typename T>
long long launch_mass_tiled(const MassArgs<T>& a);
template <typename T>
long long launch_stiffness_tiled(const StiffnessArgs<T>& a);
// column_kernels.cu -- n1^2 threads per cell, degrees 5..8 and the shapes the slab kernels skip;
// same return convention as the tiled launchers
template <typename T>
long long launch_mass_column(const MassArgs<T>& a);
template <typename T>
long long launch_stiffness_column(const StiffnessArgs<T>& a);
// cell_kernels.cu -- one thread per cell, n1 = 2 and (fp32) n1 = 3; same return convention
template <typename T>
long long launch_mass_cell(const MassArgs<T>& a);
template <typename T>
long long launch_stiffness_cell(const StiffnessArgs<T>& a);
// tc_kernels.cu -- DMMA tensor-core family, fp64 n1 = 6..9 (mass, stiffness REF_EXACT); same convention
template <typename T>
long long launch_mass_tc(const MassArgs<T>& a);
template <typename T>
long long launch_stiffness_tc(const StiffnessArgs<T>& a);

int sm_count();

// ---- device helpers ------------------------------------------------------------------------
__host__ __device__ inline uint64_t splitmix64(uint64_t x)
{
  uint64_t z = x + 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

template <typename T>
__device__ __forceinline__ T fma_t(T a, T b, T c);
template <>
__device__ __forceinline__ double fma_t<double>(double a, double b, double c)
{
  return fma(a, b, c);
}
template <>
__device__ __forceinline__ float fma_t<float>(float a, float b, float c)
{
  return fmaf(a, b, c);
}

} // namespace femtc
