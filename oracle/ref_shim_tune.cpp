// TEST / MEASUREMENT INFRASTRUCTURE ONLY (oracle).
// The reference is tuned per problem by sweeping its loop order and tile knobs (reference run.sh:13-23 sweeps the six
// orders; bench_mass.cpp:60-62 takes order, Mb, Nb from the command line).  This shim instantiates, for the degrees the
// benchmark headlines use (3: BASELINE configs 1-2, 4: config 5), the six orders x three tile pairs of the UNMODIFIED
// reference operators, so that bench.py's cpu_baseline leg can pick the reference's best triple on the host it runs on
//  instead of assuming one.  Built -Ofast only (oracle/Makefile).
//   order: 0..5 = ijk ikj jik jki kij kji (linalg.hpp:11-19); tile: 0 = (0,0), 1 = (4,8), 2 = (8,16)
#include <iostream>
#include <span>
#include <algorithm>
#include <cstddef>
#ifdef FEMTC_TUNE_MASS
#include "mass.hpp"
#else
#include "stiffness.hpp"
#endif

namespace
{
#ifdef FEMTC_TUNE_MASS
template <typename T, int P, linalg::Order order, int Mb, int Nb>
void apply(const T* phi, const T* U, T* W, const T* geom, std::size_t nc)
{
  constexpr std::size_t nd = (P + 1) * (P + 1) * (P + 1);
  operators::mass_operator<T, P, order, Mb, Nb>(std::span<const T>(phi, (P + 1) * (P + 1)),
                                                 std::span<T>(const_cast<T*>(U), nc * nd), std::span<T>(W, nc * nd),
                                                 std::span<const T>(geom, nc * nd), nc);
}
#else
template <typename T, int P, linalg::Order order, int Mb, int Nb>
void apply(const T* phi, const T* U, T* W, const T* geom, std::size_t nc)
{
  constexpr std::size_t nd = (P + 1) * (P + 1) * (P + 1);
  constexpr std::size_t max_cells = ((std::size_t(1) << 31) - 1) / (nd * 6) - 1; // int offsets, stiffness.hpp:55-57
  for (std::size_t c0 = 0; c0 < nc; c0 += max_cells)
  {
    const std::size_t n = std::min(max_cells, nc - c0);
    operators::stiffness_operator<T, P, order, Mb, Nb>(
        std::span<const T>(phi, 2 * (P + 1) * (P + 1)), std::span<T>(const_cast<T*>(U) + c0 * nd, n * nd),
        std::span<T>(W + c0 * nd, n * nd), std::span<const T>(geom + c0 * nd * 6, n * nd * 6), n);
  }
}
#endif

template <typename T, int P, linalg::Order order>
int by_tile(const T* phi, const T* U, T* W, const T* geom, std::size_t nc, int tile)
{
  switch (tile)
  {
  case 0: apply<T, P, order, 0, 0>(phi, U, W, geom, nc); return 0;
  case 1: apply<T, P, order, 4, 8>(phi, U, W, geom, nc); return 0;
  case 2: apply<T, P, order, 8, 16>(phi, U, W, geom, nc); return 0;
  default: return 2;
  }
}

template <typename T, int P>
int by_order(const T* phi, const T* U, T* W, const T* geom, std::size_t nc, int order, int tile)
{
  using linalg::Order;
  switch (order)
  {
  case 0: return by_tile<T, P, Order::ijk>(phi, U, W, geom, nc, tile);
  case 1: return by_tile<T, P, Order::ikj>(phi, U, W, geom, nc, tile);
  case 2: return by_tile<T, P, Order::jik>(phi, U, W, geom, nc, tile);
  case 3: return by_tile<T, P, Order::jki>(phi, U, W, geom, nc, tile);
  case 4: return by_tile<T, P, Order::kij>(phi, U, W, geom, nc, tile);
  case 5: return by_tile<T, P, Order::kji>(phi, U, W, geom, nc, tile);
  default: return 3;
  }
}

template <typename T>
int run(const T* phi, const T* U, T* W, const T* geom, std::size_t nc, int degree, int order, int tile)
{
  switch (degree)
  {
  case 3: return by_order<T, 3>(phi, U, W, geom, nc, order, tile);
  case 4: return by_order<T, 4>(phi, U, W, geom, nc, order, tile);
  default: return 1; // not instantiated: the caller keeps the default triple
  }
}
} // namespace

extern "C"
{
#ifdef FEMTC_TUNE_MASS
int femtc_ref_tune_mass_f64(const double* phi, const double* U, double* W, const double* g, std::size_t nc, int degree,
                            int order, int tile)
{
  return run<double>(phi, U, W, g, nc, degree, order, tile);
}
int femtc_ref_tune_mass_f32(const float* phi, const float* U, float* W, const float* g, std::size_t nc, int degree,
                            int order, int tile)
{
  return run<float>(phi, U, W, g, nc, degree, order, tile);
}
#else
int femtc_ref_tune_stiffness_f64(const double* phi, const double* U, double* W, const double* g, std::size_t nc,
                                 int degree, int order, int tile)
{
  return run<double>(phi, U, W, g, nc, degree, order, tile);
}
int femtc_ref_tune_stiffness_f32(const float* phi, const float* U, float* W, const float* g, std::size_t nc, int degree,
                                 int order, int tile)
{
  return run<float>(phi, U, W, g, nc, degree, order, tile);
}
#endif
}
