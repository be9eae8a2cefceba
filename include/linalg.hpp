// linalg.hpp -- drop-in replacement for IgorBaratta/femtc src/linalg.hpp, B200 backend.
//
// Same namespace, names, template signatures and enumerator values as the reference
// (src/linalg.hpp:11-19 Order, :21-37 string2order, :45-94 gemm_micro, :101-182
// gemm_blocked); the bodies forward to libfemtc_b200.so (include/femtc_b200.h) where
// hand-written sm_100a kernels do the arithmetic.  Compile-time shapes become run-time
// integers of the C ABI.  Unlike the reference header this one has an include guard and only
// inline definitions, so it can be included from several translation units.
//
// Link with -lfemtc_b200.  There is no CPU implementation behind these calls: a CUDA failure
// (including "no B200 present") is reported as std::runtime_error.
#pragma once

#include <stdexcept>
#include <string>

#include "femtc_b200.h"

#ifndef restrict
#define restrict __restrict__ // the reference header defines this macro (src/linalg.hpp:6)
#endif

namespace linalg
{
/// Loop order of the CPU micro-kernel.  Kept because it is part of every signature; the
/// values 0..5 are the C ABI's `order` codes.  On the GPU it has no effect on the result.
enum class Order
{
  ijk,
  ikj,
  jik,
  jki,
  kij,
  kji
};

/// @throws std::runtime_error("Invalid loop order"), like the reference.
inline Order string2order(std::string order_str)
{
  const int code = femtc_string2order(order_str.c_str());
  if (code < 0)
    throw std::runtime_error("Invalid loop order");
  return static_cast<Order>(code);
}

namespace detail
{
[[noreturn]] inline void fail(int code)
{
  throw std::runtime_error("femtc_b200 error " + std::to_string(code) + ": " + femtc_last_error());
}

template <typename T>
struct abi; // only float and double exist (bench_mass.cpp:192-195)

template <>
struct abi<double>
{
  static constexpr auto gemm_blocked = femtc_gemm_blocked_f64;
  static constexpr auto gemm_micro = femtc_gemm_micro_f64;
  static constexpr auto gemm_strided = femtc_gemm_strided_acc_f64;
  static constexpr auto mass = femtc_mass_apply_f64;
  static constexpr auto stiffness = femtc_stiffness_apply_f64;
  static constexpr auto contractions = femtc_apply_contractions_f64;
};
template <>
struct abi<float>
{
  static constexpr auto gemm_blocked = femtc_gemm_blocked_f32;
  static constexpr auto gemm_micro = femtc_gemm_micro_f32;
  static constexpr auto gemm_strided = femtc_gemm_strided_acc_f32;
  static constexpr auto mass = femtc_mass_apply_f32;
  static constexpr auto stiffness = femtc_stiffness_apply_f32;
  static constexpr auto contractions = femtc_apply_contractions_f32;
};
} // namespace detail

/// c += a.b ; a (m x k) column-major lda, b (k x nc) row-major ldb, c (m x nc) row-major ldc.
/// Accumulates into c exactly like the reference (src/linalg.hpp:57).
template <typename T, int k, int m, int nc, int lda, int ldb, int ldc, Order order = Order::ijk>
void gemm_micro(const T* restrict a, const T* restrict b, T* restrict c)
{
  const int rc = detail::abi<T>::gemm_micro(k, m, nc, lda, ldb, ldc, static_cast<int>(order), a, b, c);
  if (rc != FEMTC_OK)
    detail::fail(rc);
}

/// C = A.B ; A (m x k) column-major, B (k x n) row-major, C (m x n) column-major.
/// C is overwritten, as the reference does (src/linalg.hpp:128,134-136).  MB x NB is the CPU
/// tile shape; it is checked for sanity and has no influence on the GPU result.
template <typename T, int k, int m, int n, Order order, int MB, int NB>
void gemm_blocked(const T* restrict a, const T* restrict b, T* restrict c)
{
  const int rc = detail::abi<T>::gemm_blocked(k, m, n, static_cast<int>(order), MB, NB, a, b, c);
  if (rc != FEMTC_OK)
    detail::fail(rc);
}
} // namespace linalg
