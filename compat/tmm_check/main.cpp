// Run-time parity of the tmm.hpp building blocks: include/tmm.hpp (this TU, B200
// backend through libfemtc_b200.so) against the reference's src/tmm.hpp (ref_side.cpp, CPU) on the same random data:
// contract 2-D, contract 3-D with both `transpose` values, transpose 2-D / 3-D, apply_contractions square and
// rectangular with both Tr values; double at 1e-12, float at 1e-5 (relative to the largest reference value),
// transposes bit-exact.  Accumulating entry points start from a non-zero C.  Prints one line per block and exits
// non-zero on any mismatch.  Built by compat/build_reference_drivers.sh, run by tests/test_gpu_parity.py.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <type_traits>
#include <vector>

#include "tmm.hpp" // include/tmm.hpp
#include "shapes.hpp"

namespace
{
uint64_t g_state = 0x9E3779B97F4A7C15ull;
double next_uniform() // splitmix64, centred: exercises signs
{
  uint64_t z = (g_state += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (double)(z >> 11) * (1.0 / 9007199254740992.0) - 0.5;
}
template <typename T>
std::vector<T> rnd(size_t n)
{
  std::vector<T> v(n);
  for (auto& x : v)
    x = (T)next_uniform();
  return v;
}
template <typename T>
constexpr double tol() { return std::is_same<T, double>::value ? 1e-12 : 1e-5; }

int g_fail = 0;
double g_worst[2] = {0, 0}; // worst error / tolerance per dtype

template <typename T>
void compare(const char* what, int id, int flag, const std::vector<T>& got, const std::vector<T>& ref, bool exact)
{
  double scale = 0, err = 0;
  for (size_t i = 0; i < ref.size(); ++i)
    scale = std::fmax(scale, std::fabs((double)ref[i]));
  bool bits = std::memcmp(got.data(), ref.data(), ref.size() * sizeof(T)) == 0;
  for (size_t i = 0; i < ref.size(); ++i)
    err = std::fmax(err, std::fabs((double)got[i] - (double)ref[i]));
  const double rel = scale > 0 ? err / scale : err;
  const bool ok = exact ? bits : (rel <= tol<T>() && std::isfinite(rel));
  if (!exact)
    g_worst[sizeof(T) == 8 ? 0 : 1] = std::fmax(g_worst[sizeof(T) == 8 ? 0 : 1], rel / tol<T>());
  if (!ok)
  {
    ++g_fail;
    std::printf("MISMATCH %s<%s> id %d flag %d: rel err %.3e\n", what, sizeof(T) == 8 ? "double" : "float", id, flag, rel);
  }
}

template <typename T>
struct Ref;
template <>
struct Ref<double>
{
  static constexpr auto contract2d = tmm_ref_contract2d_f64;
  static constexpr auto contract3d = tmm_ref_contract3d_f64;
  static constexpr auto transpose2d = tmm_ref_transpose2d_f64;
  static constexpr auto transpose3d = tmm_ref_transpose3d_f64;
  static constexpr auto apply = tmm_ref_apply_f64;
};
template <>
struct Ref<float>
{
  static constexpr auto contract2d = tmm_ref_contract2d_f32;
  static constexpr auto contract3d = tmm_ref_contract3d_f32;
  static constexpr auto transpose2d = tmm_ref_transpose2d_f32;
  static constexpr auto transpose3d = tmm_ref_transpose3d_f32;
  static constexpr auto apply = tmm_ref_apply_f32;
};

template <typename T>
int run_all()
{
  int n = 0;
  // ---- contract, 2-D (accumulates into C) ----
#define X(ID, NA, NB, NK)                                                                                               \
  {                                                                                                                     \
    auto A = rnd<T>(NA * NK), B = rnd<T>(NB * NK), C0 = rnd<T>(NA * NB);                                                \
    auto Cg = C0, Cr = C0;                                                                                              \
    contract<T, NA, NB, NK>(A.data(), B.data(), Cg.data());                                                             \
    if (Ref<T>::contract2d(ID, A.data(), B.data(), Cr.data())) ++g_fail;                                                \
    compare<T>("contract2d", ID, 0, Cg, Cr, false);                                                                     \
    ++n;                                                                                                                \
  }
  TMM_CONTRACT2D(X)
#undef X
  // ---- contract, 3-D, both layouts of the matrix (accumulates into C) ----
#define X(ID, NK, NA, NB, NC)                                                                                           \
  for (int tr = 0; tr < 2; ++tr)                                                                                        \
  {                                                                                                                     \
    auto A = rnd<T>(NA * NK), B = rnd<T>(NK * NB * NC), C0 = rnd<T>(NA * NB * NC);                                      \
    auto Cg = C0, Cr = C0;                                                                                              \
    if (tr) contract<T, NK, NA, NB, NC, true>(A.data(), B.data(), Cg.data());                                           \
    else contract<T, NK, NA, NB, NC, false>(A.data(), B.data(), Cg.data());                                             \
    if (Ref<T>::contract3d(ID, tr, A.data(), B.data(), Cr.data())) ++g_fail;                                            \
    compare<T>("contract3d", ID, tr, Cg, Cr, false);                                                                    \
    ++n;                                                                                                                \
  }
  TMM_CONTRACT3D(X)
#undef X
  // ---- transposes: pure data movement, bit-exact ----
#define X(ID, NA, NB, OA, OB)                                                                                           \
  {                                                                                                                     \
    auto A = rnd<T>(NA * NB);                                                                                           \
    std::vector<T> Bg(NA * NB, T(-7)), Br(NA * NB, T(-7));                                                              \
    transpose<T, NA, NB, OA, OB>(A.data(), Bg.data());                                                                  \
    if (Ref<T>::transpose2d(ID, A.data(), Br.data())) ++g_fail;                                                         \
    compare<T>("transpose2d", ID, 0, Bg, Br, true);                                                                     \
    ++n;                                                                                                                \
  }
  TMM_TRANSPOSE2D(X)
#undef X
#define X(ID, NA, NB, NC, OA, OB, OC)                                                                                   \
  {                                                                                                                     \
    auto A = rnd<T>(NA * NB * NC);                                                                                      \
    std::vector<T> Bg(NA * NB * NC, T(-7)), Br(NA * NB * NC, T(-7));                                                    \
    transpose<T, NA, NB, NC, OA, OB, OC>(A.data(), Bg.data());                                                          \
    if (Ref<T>::transpose3d(ID, A.data(), Br.data())) ++g_fail;                                                         \
    compare<T>("transpose3d", ID, 0, Bg, Br, true);                                                                     \
    ++n;                                                                                                                \
  }
  TMM_TRANSPOSE3D(X)
#undef X
  // ---- apply_contractions (overwrites A), square and rectangular, both Tr ----
#define X(ID, NA, NB)                                                                                                   \
  for (int tr = 0; tr < 2; ++tr)                                                                                        \
  {                                                                                                                     \
    auto I0 = rnd<T>(NA * NB), I1 = rnd<T>(NA * NB), I2 = rnd<T>(NA * NB), B = rnd<T>(NB * NB * NB);                    \
    std::vector<T> Ag(NA * NA * NA, T(-7)), Ar(NA * NA * NA, T(-7));                                                    \
    Buffer<T, NA, NB> buf;                                                                                              \
    auto Bc = B;                                                                                                        \
    if (tr) apply_contractions<T, NA, NB, true>(I0.data(), I1.data(), I2.data(), Bc.data(), Ag.data(), buf);            \
    else apply_contractions<T, NA, NB, false>(I0.data(), I1.data(), I2.data(), Bc.data(), Ag.data(), buf);              \
    if (Ref<T>::apply(ID, tr, I0.data(), I1.data(), I2.data(), B.data(), Ar.data())) ++g_fail;                          \
    compare<T>("apply_contractions", ID, tr, Ag, Ar, false);                                                            \
    ++n;                                                                                                                \
  }
  TMM_APPLY(X)
#undef X
  return n;
}
} // namespace

int main()
{
  int n = 0;
  try
  {
    n += run_all<double>();
    n += run_all<float>();
  }
  catch (const std::exception& e)
  {
    std::printf("EXCEPTION %s\n", e.what());
    return 2;
  }
  std::printf("tmm building blocks: %d comparisons, %d failures, worst error/tolerance double %.3g float %.3g\n", n, g_fail,
              g_worst[0], g_worst[1]);
  return g_fail ? 1 : 0;
}
