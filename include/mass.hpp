// mass.hpp -- drop-in replacement for IgorBaratta/femtc src/mass.hpp, B200 backend.
//
// operators::mass_operator keeps the reference's overload set: the fully templated kernel
// (src/mass.hpp:10-49) and the four run-time dispatch levels above it (degree :52-138,
// order :140-169, Mb :171-194, Nb :196-219).  The 2400 template instantiations of the
// reference collapse to one call into libfemtc_b200.so; knobs are validated in the same
// order and with the same std::cout messages, after which W is left untouched and the call
// returns without throwing -- exactly the reference's behaviour.
#pragma once

#include "linalg.hpp"
#include <cstddef>
#include <iostream>
#include <span>

using namespace linalg; // the reference header does this at global scope (src/mass.hpp:6)

namespace operators
{
namespace detail
{
/// Prints what the reference prints for a rejected knob; throws for real failures.
inline void report(int rc)
{
  switch (rc)
  {
  case FEMTC_OK:
    return;
  case FEMTC_ERR_DEGREE:
    std::cout << "degree not supported" << std::endl;
    return;
  case FEMTC_ERR_ORDER:
    std::cout << "order not supported" << std::endl;
    return;
  case FEMTC_ERR_BLOCK:
    std::cout << "Block size not supported" << std::endl;
    return;
  default:
    linalg::detail::fail(rc);
  }
}

template <typename T>
void mass(std::span<const T> phi, std::span<const T> U, std::span<T> W, std::span<const T> detJ,
          std::size_t num_cells, int degree, Order order, int Mb, int Nb)
{
  report(linalg::detail::abi<T>::mass(phi.data(), phi.size(), U.data(), W.data(), detJ.data(),
                                      num_cells, degree, static_cast<int>(order), Mb, Nb));
}
} // namespace detail

/// W = (A x A x A) diag(detJ) (A x A x A) U for every cell; reference src/mass.hpp:10-49.
template <typename T, int P, Order layout = Order::ijk, int Mb, int Nb>
void mass_operator(std::span<const T> phi, std::span<T> U, std::span<T> W, std::span<const T> detJ,
                   std::size_t num_cells)
{
  detail::mass<T>(phi, U, W, detJ, num_cells, P, layout, Mb, Nb);
}

/// reference src/mass.hpp:52-138
template <typename T, Order layout, int Mb, int Nb>
void mass_operator(std::span<T> a, std::span<T> b, std::span<T> c, std::span<T> detJ,
                   std::size_t num_cells, int degree)
{
  detail::mass<T>(a, b, c, detJ, num_cells, degree, layout, Mb, Nb);
}

/// reference src/mass.hpp:140-169
template <typename T, int Mb, int Nb>
void mass_operator(std::span<T> a, std::span<T> b, std::span<T> c, std::span<T> detJ,
                   std::size_t num_cells, int degree, Order order)
{
  detail::mass<T>(a, b, c, detJ, num_cells, degree, order, Mb, Nb);
}

/// reference src/mass.hpp:171-194
template <typename T, int Nb>
void mass_operator(std::span<T> a, std::span<T> b, std::span<T> c, std::span<T> detJ,
                   std::size_t num_cells, int degree, Order order, int Mb)
{
  detail::mass<T>(a, b, c, detJ, num_cells, degree, order, Mb, Nb);
}

/// The public entry point: reference src/mass.hpp:196-219, call site bench_mass.cpp:98.
template <typename T>
void mass_operator(std::span<T> a, std::span<T> b, std::span<T> c, std::span<T> detJ,
                   std::size_t num_cells, int degree, Order order, int Mb, int Nb)
{
  detail::mass<T>(a, b, c, detJ, num_cells, degree, order, Mb, Nb);
}
} // namespace operators
