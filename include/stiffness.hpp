// stiffness.hpp -- drop-in replacement for IgorBaratta/femtc src/stiffness.hpp, B200 backend.
//
// operators::stiffness_operator keeps the reference's overload set (kernel src/stiffness.hpp:34-99,
// dispatch levels :102-188, :190-218, :220-243, :245-268).  By default the result is the
// reference's AS CODED (mode REF_EXACT: only the last back-contraction survives, Wz without the
// derivative -- see include/femtc_b200.h); pass operators::StiffnessMode::full to the extra
// overload for the intended x+y+z operator.
//
// phi must hold 2*(P+1)^2 values [basis | derivative] (src/stiffness.hpp:44-45).  The shipped
// bench_stiffness.cpp:74 passes (P+1)^2; the reference then reads past the end of the vector,
// this backend does not (it warns once and uses derivative := basis).
#pragma once

#include "linalg.hpp"
#include <cstddef>
#include <iostream>
#include <span>

using namespace linalg; // as the reference does (src/stiffness.hpp:8)

namespace operators
{
enum class StiffnessMode : int
{
  ref_exact = FEMTC_STIFFNESS_REF_EXACT,
  full = FEMTC_STIFFNESS_FULL
};

namespace detail
{
inline void report_stiffness(int rc)
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
void stiffness(std::span<const T> phi, std::span<const T> U, std::span<T> W, std::span<const T> G,
               std::size_t num_cells, int degree, Order order, int Mb, int Nb,
               StiffnessMode mode = StiffnessMode::ref_exact)
{
  report_stiffness(linalg::detail::abi<T>::stiffness(phi.data(), phi.size(), U.data(), W.data(),
                                                     G.data(), num_cells, degree,
                                                     static_cast<int>(order), Mb, Nb,
                                                     static_cast<int>(mode)));
}
} // namespace detail

/// reference src/stiffness.hpp:34-99
template <typename T, int P, Order order, int Mb, int Nb>
void stiffness_operator(std::span<const T> phi, std::span<T> U, std::span<T> W, std::span<const T> G,
                        std::size_t num_cells)
{
  detail::stiffness<T>(phi, U, W, G, num_cells, P, order, Mb, Nb);
}

/// reference src/stiffness.hpp:102-188
template <typename T, Order order, int Mb, int Nb>
void stiffness_operator(std::span<T> a, std::span<T> b, std::span<T> c, std::span<T> G, int num_cells,
                        int degree)
{
  detail::stiffness<T>(a, b, c, G, static_cast<std::size_t>(num_cells), degree, order, Mb, Nb);
}

/// reference src/stiffness.hpp:190-218
template <typename T, int Mb, int Nb>
void stiffness_operator(std::span<T> a, std::span<T> b, std::span<T> c, std::span<T> G, int num_cells,
                        int degree, Order order)
{
  detail::stiffness<T>(a, b, c, G, static_cast<std::size_t>(num_cells), degree, order, Mb, Nb);
}

/// reference src/stiffness.hpp:220-243
template <typename T, int Nb>
void stiffness_operator(std::span<T> a, std::span<T> b, std::span<T> c, std::span<T> G, int num_cells,
                        int degree, Order order, int Mb)
{
  detail::stiffness<T>(a, b, c, G, static_cast<std::size_t>(num_cells), degree, order, Mb, Nb);
}

/// The public entry point: reference src/stiffness.hpp:245-268, call site bench_stiffness.cpp:98.
template <typename T>
void stiffness_operator(std::span<T> a, std::span<T> b, std::span<T> c, std::span<T> G, int num_cells,
                        int degree, Order order, int Mb, int Nb)
{
  detail::stiffness<T>(a, b, c, G, static_cast<std::size_t>(num_cells), degree, order, Mb, Nb);
}

/// Extension: same call with an explicit mode and a 64-bit cell count.
template <typename T>
void stiffness_operator(std::span<T> a, std::span<T> b, std::span<T> c, std::span<T> G,
                        std::size_t num_cells, int degree, Order order, int Mb, int Nb,
                        StiffnessMode mode)
{
  detail::stiffness<T>(a, b, c, G, num_cells, degree, order, Mb, Nb, mode);
}
} // namespace operators
