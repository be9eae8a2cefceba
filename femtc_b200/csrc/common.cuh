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
template <typename T>
int launch_fill_uniform(T* dX, size_t n, uint64_t seed, uint64_t stream_id, uint64_t offset,
                        int centred, cudaStream_t s);

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
int tiled_kernel_info(int op, int dtype_bytes, int degree, int mode, long long info[8]);

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
