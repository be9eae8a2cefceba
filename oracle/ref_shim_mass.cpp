// TEST INFRASTRUCTURE ONLY (oracle).
// Compiles the reference's mass operator (reference src/mass.hpp:10-49, the fully
// templated per-cell kernel) for degrees 1..10 and four (order, Mb, Nb) triples.
// mass.hpp has no include guard and needs <iostream> from its includer
// (SURVEY.md F7), hence its own translation unit.
#include <iostream>
#include <span>
#include "mass.hpp"
#include "ref_shim_common.hpp"

namespace
{
template <typename T, linalg::Order order, int Mb, int Nb>
int run_degree(const T* phi, const T* U, T* W, const T* detJ, std::size_t nc, int degree)
{
  auto go = [&]<int P>()
  {
    constexpr std::size_t nd = (P + 1) * (P + 1) * (P + 1);
    operators::mass_operator<T, P, order, Mb, Nb>(
        std::span<const T>(phi, (P + 1) * (P + 1)), std::span<T>(const_cast<T*>(U), nc * nd),
        std::span<T>(W, nc * nd), std::span<const T>(detJ, nc * nd), nc);
  };
  switch (degree)
  {
#define X(P) case P: go.template operator()<P>(); return 0;
    FEMTC_REF_FOR_EACH_DEGREE(X)
#undef X
  default:
    return 1;
  }
}

template <typename T>
int run(const T* phi, const T* U, T* W, const T* detJ, std::size_t nc, int degree, int cfg)
{
  using linalg::Order;
  switch (cfg)
  {
  case 0: return run_degree<T, Order::ijk, 0, 0>(phi, U, W, detJ, nc, degree);
  case 1: return run_degree<T, Order::ikj, 0, 0>(phi, U, W, detJ, nc, degree);
  case 2: return run_degree<T, Order::kij, 8, 16>(phi, U, W, detJ, nc, degree);
  case 3: return run_degree<T, Order::jki, 4, 8>(phi, U, W, detJ, nc, degree);
  default: return 2;
  }
}
} // namespace

extern "C"
{
int femtc_ref_mass_f64(const double* phi, const double* U, double* W, const double* detJ,
                       std::size_t num_cells, int degree, int cfg)
{
  return run<double>(phi, U, W, detJ, num_cells, degree, cfg);
}
int femtc_ref_mass_f32(const float* phi, const float* U, float* W, const float* detJ,
                       std::size_t num_cells, int degree, int cfg)
{
  return run<float>(phi, U, W, detJ, num_cells, degree, cfg);
}
}
