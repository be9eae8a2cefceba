// TEST INFRASTRUCTURE ONLY (oracle).
// Compiles the reference's small-GEMM layer (reference src/linalg.hpp:45-182) and
// the orphaned tensor-contraction layer (reference src/tmm.hpp:91-126) behind a C ABI.
//   gemm_blocked: the operator shape k = m = P+1, n = (P+1)^2 for P = 1..25
//                 (the shapes tests/test_gemm.cpp:65-89 runs), four knob triples.
//   gemm_micro:   the one shape tests/test_micro.cpp:44 runs, all six loop orders.
//   apply_contractions: a few square and rectangular (Na, Nb) pairs, Tr on/off.
#include <array>
#include <cstddef>
#include "linalg.hpp"
#include "tmm.hpp"
#include "ref_shim_common.hpp"

namespace
{
using linalg::Order;

template <typename T, Order order, int MB, int NB>
int blocked_degree(int P, const T* a, const T* b, T* c)
{
  auto go = [&]<int Q>()
  { linalg::gemm_blocked<T, Q + 1, Q + 1, (Q + 1) * (Q + 1), order, MB, NB>(a, b, c); };
  switch (P)
  {
#define X(Q) case Q: go.template operator()<Q>(); return 0;
    X(1) X(2) X(3) X(4) X(5) X(6) X(7) X(8) X(9) X(10) X(11) X(12) X(13)
    X(14) X(15) X(16) X(17) X(18) X(19) X(20) X(21) X(22) X(23) X(24) X(25)
#undef X
  default:
    return 1;
  }
}

template <typename T>
int blocked(int P, int cfg, const T* a, const T* b, T* c)
{
  switch (cfg)
  {
  case 0: return blocked_degree<T, Order::ijk, 0, 0>(P, a, b, c);
  case 1: return blocked_degree<T, Order::ikj, 0, 0>(P, a, b, c);
  case 2: return blocked_degree<T, Order::kij, 8, 16>(P, a, b, c);
  case 3: return blocked_degree<T, Order::jki, 4, 8>(P, a, b, c);
  default: return 2;
  }
}

template <typename T>
int micro(int order, const T* a, const T* b, T* c)
{
  switch (order)
  {
  case 0: linalg::gemm_micro<T, 26, 4, 16, 4, 16, 16, Order::ijk>(a, b, c); return 0;
  case 1: linalg::gemm_micro<T, 26, 4, 16, 4, 16, 16, Order::ikj>(a, b, c); return 0;
  case 2: linalg::gemm_micro<T, 26, 4, 16, 4, 16, 16, Order::jik>(a, b, c); return 0;
  case 3: linalg::gemm_micro<T, 26, 4, 16, 4, 16, 16, Order::jki>(a, b, c); return 0;
  case 4: linalg::gemm_micro<T, 26, 4, 16, 4, 16, 16, Order::kij>(a, b, c); return 0;
  case 5: linalg::gemm_micro<T, 26, 4, 16, 4, 16, 16, Order::kji>(a, b, c); return 0;
  default: return 2;
  }
}

template <typename T, int Na, int Nb>
int contractions_tr(int Tr, const T* I0, const T* I1, const T* I2, T* B, T* A)
{
  Buffer<T, Na, Nb> buffer;
  if (Tr)
    apply_contractions<T, Na, Nb, true>(I0, I1, I2, B, A, buffer);
  else
    apply_contractions<T, Na, Nb, false>(I0, I1, I2, B, A, buffer);
  return 0;
}

template <typename T>
int contractions(int Na, int Nb, int Tr, const T* I0, const T* I1, const T* I2, T* B, T* A)
{
#define X(NA, NB) if (Na == NA && Nb == NB) return contractions_tr<T, NA, NB>(Tr, I0, I1, I2, B, A);
  X(2, 2) X(3, 3) X(4, 4) X(5, 5) X(9, 9) X(6, 4) X(3, 5) X(8, 6) X(5, 4) X(4, 7)
#undef X
  return 1;
}
} // namespace

extern "C"
{
int femtc_ref_gemm_blocked_f64(int P, int cfg, const double* a, const double* b, double* c)
{ return blocked<double>(P, cfg, a, b, c); }
int femtc_ref_gemm_blocked_f32(int P, int cfg, const float* a, const float* b, float* c)
{ return blocked<float>(P, cfg, a, b, c); }
int femtc_ref_gemm_micro_f64(int order, const double* a, const double* b, double* c)
{ return micro<double>(order, a, b, c); }
int femtc_ref_gemm_micro_f32(int order, const float* a, const float* b, float* c)
{ return micro<float>(order, a, b, c); }
int femtc_ref_apply_contractions_f64(int Na, int Nb, int Tr, const double* I0, const double* I1,
                                     const double* I2, double* B, double* A)
{ return contractions<double>(Na, Nb, Tr, I0, I1, I2, B, A); }
int femtc_ref_apply_contractions_f32(int Na, int Nb, int Tr, const float* I0, const float* I1,
                                     const float* I2, float* B, float* A)
{ return contractions<float>(Na, Nb, Tr, I0, I1, I2, B, A); }
int femtc_ref_string2order(const char* s)
{
  try { return static_cast<int>(linalg::string2order(s)); }
  catch (const std::runtime_error&) { return -1; }
}
}
