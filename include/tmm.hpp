// tmm.hpp -- drop-in replacement for IgorBaratta/femtc src/tmm.hpp, B200 backend.
//
// The reference's alternative formulation of the per-axis contraction (no file of the
// reference includes it): transpose / contract building blocks, the Buffer scratch, and
// apply_contractions = three contract+transpose rounds (src/tmm.hpp:91-126).  Here
// apply_contractions is ONE fused device kernel, contract forwards to the strided
// accumulate-GEMM kernel, and transpose -- pure data movement on caller-owned host arrays, no
// arithmetic -- is an index permutation.  Buffer is kept so existing call sites compile; the
// device kernels keep their scratch in shared memory and never touch it.
#pragma once

#include <array>

#include "linalg.hpp"

/// B[a*offa + b*offb] = A[a,b]      reference src/tmm.hpp:9-15
template <typename T, int Na, int Nb, int offa, int offb>
void transpose(const T* __restrict__ A, T* __restrict__ B)
{
  for (int x = 0; x < Na * Nb; ++x)
    B[(x / Nb) * offa + (x % Nb) * offb] = A[x];
}

/// C[a,b] += sum_k A[a,k] B[b,k]    reference src/tmm.hpp:24-32
template <typename T, int Na, int Nb, int Nk>
void contract(const T* __restrict__ A, const T* __restrict__ B, T* __restrict__ C)
{
  const int rc = linalg::detail::abi<T>::gemm_strided(Nk, Na, Nb, /*a(i,p)*/ Nk, 1, /*b(p,j)*/ 1, Nk,
                                                      /*ldc*/ Nb, A, B, C);
  if (rc != FEMTC_OK)
    linalg::detail::fail(rc);
}

/// B[offa*a + offb*b + offc*c] = A[a,b,c]      reference src/tmm.hpp:37-44
template <typename T, int Na, int Nb, int Nc, int offa, int offb, int offc>
void transpose(T* __restrict__ A, T* __restrict__ B)
{
  for (int x = 0; x < Na * Nb * Nc; ++x)
    B[offa * (x / (Nb * Nc)) + offb * ((x / Nc) % Nb) + offc * (x % Nc)] = A[x];
}

/// Scratch of the reference's CPU formulation (src/tmm.hpp:46-60); unused by the device kernels.
template <typename T, int Na, int Nb>
struct Buffer
{
  std::array<T, Na * Nb * Nb> T0{0};
  std::array<T, Na * Nb * Nb> T0_t{0};
  std::array<T, Na * Na * Nb> T1{0};
  std::array<T, Na * Na * Nb> T1_t{0};
  std::array<T, Na * Na * Na> T2{0};
  void zero()
  {
    T0.fill(T(0));
    T1.fill(T(0));
    T2.fill(T(0));
  }
};

/// C[a,(b,c)] += sum_k M(a,k) B[k,(b,c)],  M(a,k) = transpose ? A[a*Nk+k] : A[k*Na+a]
/// reference src/tmm.hpp:65-85
template <typename T, int Nk, int Na, int Nb, int Nc, bool transpose>
void contract(const T* __restrict__ A, const T* __restrict__ B, T* __restrict__ C)
{
  constexpr int Nd = Nb * Nc;
  const int rc = linalg::detail::abi<T>::gemm_strided(Nk, Na, Nd, transpose ? Nk : 1, transpose ? 1 : Na,
                                                      Nd, 1, Nd, A, B, C);
  if (rc != FEMTC_OK)
    linalg::detail::fail(rc);
}

/// A[a0,a1,a2] = sum I0(a0,b0) I1(a1,b1) I2(a2,b2) B[b0,b1,b2]   reference src/tmm.hpp:91-126
template <typename T, int Na, int Nb, bool Tr>
void apply_contractions(const T* I0, const T* I1, const T* I2, T* __restrict__ B, T* __restrict__ A,
                        Buffer<T, Na, Nb>& /*buffer*/)
{
  const int rc = linalg::detail::abi<T>::contractions(Na, Nb, Tr ? 1 : 0, I0, I1, I2, B, A, 1);
  if (rc != FEMTC_OK)
    linalg::detail::fail(rc);
}

/// Extension (no counterpart in the reference): the same contraction applied to `batch` tensors stored back to
/// back -- B holds batch*Nb^3 values, A receives batch*Na^3 -- in one call, i.e. one pipelined pass of the tiled
/// device kernel (femtc_b200/csrc/tmm_kernels.cu) instead of `batch` single-tensor round trips.  Host pointers.
template <typename T, int Na, int Nb, bool Tr>
void apply_contractions_batched(const T* I0, const T* I1, const T* I2, const T* B, T* A, std::size_t batch)
{
  const int rc = linalg::detail::abi<T>::contractions(Na, Nb, Tr ? 1 : 0, I0, I1, I2, B, A, batch);
  if (rc != FEMTC_OK)
    linalg::detail::fail(rc);
}
