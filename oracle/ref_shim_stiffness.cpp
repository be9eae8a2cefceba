// TEST INFRASTRUCTURE ONLY (oracle).
// Compiles the reference's stiffness operator (reference src/stiffness.hpp:34-99)
// for degrees 1..10 and four (order, Mb, Nb) triples.  Calls are chunked so that
// cells*ndofs*6 stays below 2^31: the reference indexes with `int stride`
// (stiffness.hpp:55-57); chunking is exact because the action is
// cell-local.  phi must hold 2*(P+1)^2 values [basis | derivative].
#include <iostream>
#include <span>
#include <algorithm>
#include "stiffness.hpp"
#include "ref_shim_common.hpp"

namespace
{
template <typename T, linalg::Order order, int Mb, int Nb>
int run_degree(const T* phi, const T* U, T* W, const T* G, std::size_t nc, int degree)
{
  auto go = [&]<int P>()
  {
    constexpr std::size_t nd = (P + 1) * (P + 1) * (P + 1);
    constexpr std::size_t max_cells = ((std::size_t(1) << 31) - 1) / (nd * 6) - 1;
    for (std::size_t c0 = 0; c0 < nc; c0 += max_cells)
    {
      const std::size_t n = std::min(max_cells, nc - c0);
      operators::stiffness_operator<T, P, order, Mb, Nb>(
          std::span<const T>(phi, 2 * (P + 1) * (P + 1)),
          std::span<T>(const_cast<T*>(U) + c0 * nd, n * nd), std::span<T>(W + c0 * nd, n * nd),
          std::span<const T>(G + c0 * nd * 6, n * nd * 6), n);
    }
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
int run(const T* phi, const T* U, T* W, const T* G, std::size_t nc, int degree, int cfg)
{
  using linalg::Order;
  switch (cfg)
  {
  case 0: return run_degree<T, Order::ijk, 0, 0>(phi, U, W, G, nc, degree);
  case 1: return run_degree<T, Order::ikj, 0, 0>(phi, U, W, G, nc, degree);
  case 2: return run_degree<T, Order::kij, 8, 16>(phi, U, W, G, nc, degree);
  case 3: return run_degree<T, Order::jki, 4, 8>(phi, U, W, G, nc, degree);
  default: return 2;
  }
}
} // namespace

extern "C"
{
int femtc_ref_stiffness_f64(const double* phi, const double* U, double* W, const double* G,
                            std::size_t num_cells, int degree, int cfg)
{
  return run<double>(phi, U, W, G, num_cells, degree, cfg);
}
int femtc_ref_stiffness_f32(const float* phi, const float* U, float* W, const float* G,
                            std::size_t num_cells, int degree, int cfg)
{
  return run<float>(phi, U, W, G, num_cells, degree, cfg);
}
}
