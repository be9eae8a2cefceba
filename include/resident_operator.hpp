// resident_operator.hpp -- extension header (no counterpart in the reference): the mass / stiffness
// operator with its geometry resident in HBM.
//
// The reference's drivers apply the operator nreps times to the same detJ / G arrays
// (bench_mass.cpp:95-105, bench_stiffness.cpp:95-105); a solver would do the same every iteration.
// The drop-in entry points operators::mass_operator / stiffness_operator (include/mass.hpp,
// include/stiffness.hpp) take detJ / G on every call, as the reference does, and therefore move them
// across PCIe every time.  An operators::ResidentOperator uploads them once:
//
//     operators::ResidentOperator<double> K(operators::Kind::stiffness, phi, G, num_cells, P);
//     for (int it = 0; it < nreps; ++it) K.apply(U, W);     // moves U in and W out only
//
// Results are bit-identical to the drop-in entry points (same kernels).
#pragma once

#include "stiffness.hpp"
#include <span>
#include <stdexcept>
#include <utility>

namespace operators
{
enum class Kind : int
{
  mass = 0,
  stiffness = 1
};

template <typename T>
class ResidentOperator
{
  static_assert(sizeof(T) == 4 || sizeof(T) == 8, "float or double");

public:
  ResidentOperator(Kind kind, std::span<const T> phi, std::span<const T> geom, std::size_t num_cells, int degree,
                   StiffnessMode mode = StiffnessMode::ref_exact)
  {
    int rc;
    if constexpr (sizeof(T) == 8)
      rc = femtc_operator_create_f64(&h_, static_cast<int>(kind), phi.data(), phi.size(), geom.data(), num_cells,
                                     degree, static_cast<int>(mode));
    else
      rc = femtc_operator_create_f32(&h_, static_cast<int>(kind), phi.data(), phi.size(), geom.data(), num_cells,
                                     degree, static_cast<int>(mode));
    if (rc != FEMTC_OK)
      throw std::runtime_error(femtc_last_error());
  }
  ResidentOperator(const ResidentOperator&) = delete;
  ResidentOperator& operator=(const ResidentOperator&) = delete;
  ResidentOperator(ResidentOperator&& o) noexcept : h_(std::exchange(o.h_, nullptr)) {}
  ~ResidentOperator() { femtc_operator_destroy(h_); }

  /// W = op(U); host spans, synchronous
  void apply(std::span<const T> U, std::span<T> W) const
  {
    if (femtc_operator_apply(h_, U.data(), W.data()) != FEMTC_OK)
      throw std::runtime_error(femtc_last_error());
  }
  /// device pointers, asynchronous on `stream` (cudaStream_t as void*)
  void apply_device(const T* dU, T* dW, void* stream = nullptr) const
  {
    if (femtc_operator_apply_device(h_, dU, dW, stream) != FEMTC_OK)
      throw std::runtime_error(femtc_last_error());
  }
  std::size_t device_bytes() const { return femtc_operator_device_bytes(h_); }

private:
  femtc_operator* h_ = nullptr;
};
} // namespace operators
