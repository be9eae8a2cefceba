// TEST INFRASTRUCTURE ONLY.  The reference's src/tmm.hpp, compiled UNMODIFIED from where it lies (-I $REF/src) inside
// its own namespace so that its templates (contract, transpose, Buffer, apply_contractions -- the very names
// include/tmm.hpp defines) get different symbols than the replacement's in the other TU of the same binary.
// The header includes nothing itself (it uses std::array without <array>), so wrapping is safe.
#include <array>
#include <cstddef>

namespace reference_tmm
{
#include "tmm.hpp" // -I /root/reference/src
}

#include "shapes.hpp"

namespace
{
namespace R = reference_tmm;

template <typename T>
int contract2d(int id, const T* A, const T* B, T* C)
{
  switch (id)
  {
#define X(ID, NA, NB, NK) case ID: R::contract<T, NA, NB, NK>(A, B, C); return 0;
    TMM_CONTRACT2D(X)
#undef X
  }
  return 1;
}
template <typename T>
int contract3d(int id, int tr, const T* A, const T* B, T* C)
{
  switch (id)
  {
#define X(ID, NK, NA, NB, NC)                                                                                           \
  case ID:                                                                                                              \
    if (tr) R::contract<T, NK, NA, NB, NC, true>(A, B, C);                                                              \
    else R::contract<T, NK, NA, NB, NC, false>(A, B, C);                                                                \
    return 0;
    TMM_CONTRACT3D(X)
#undef X
  }
  return 1;
}
template <typename T>
int transpose2d(int id, const T* A, T* B)
{
  switch (id)
  {
#define X(ID, NA, NB, OA, OB) case ID: R::transpose<T, NA, NB, OA, OB>(A, B); return 0;
    TMM_TRANSPOSE2D(X)
#undef X
  }
  return 1;
}
template <typename T>
int transpose3d(int id, T* A, T* B)
{
  switch (id)
  {
#define X(ID, NA, NB, NC, OA, OB, OC) case ID: R::transpose<T, NA, NB, NC, OA, OB, OC>(A, B); return 0;
    TMM_TRANSPOSE3D(X)
#undef X
  }
  return 1;
}
template <typename T>
int apply(int id, int Tr, const T* I0, const T* I1, const T* I2, T* B, T* A)
{
  switch (id)
  {
#define X(ID, NA, NB)                                                                                                   \
  case ID:                                                                                                              \
  {                                                                                                                     \
    R::Buffer<T, NA, NB> buf;                                                                                           \
    if (Tr) R::apply_contractions<T, NA, NB, true>(I0, I1, I2, B, A, buf);                                              \
    else R::apply_contractions<T, NA, NB, false>(I0, I1, I2, B, A, buf);                                                \
    return 0;                                                                                                           \
  }
    TMM_APPLY(X)
#undef X
  }
  return 1;
}
} // namespace

extern "C"
{
int tmm_ref_contract2d_f64(int id, const double* A, const double* B, double* C) { return contract2d(id, A, B, C); }
int tmm_ref_contract2d_f32(int id, const float* A, const float* B, float* C) { return contract2d(id, A, B, C); }
int tmm_ref_contract3d_f64(int id, int t, const double* A, const double* B, double* C) { return contract3d(id, t, A, B, C); }
int tmm_ref_contract3d_f32(int id, int t, const float* A, const float* B, float* C) { return contract3d(id, t, A, B, C); }
int tmm_ref_transpose2d_f64(int id, const double* A, double* B) { return transpose2d(id, A, B); }
int tmm_ref_transpose2d_f32(int id, const float* A, float* B) { return transpose2d(id, A, B); }
int tmm_ref_transpose3d_f64(int id, double* A, double* B) { return transpose3d(id, A, B); }
int tmm_ref_transpose3d_f32(int id, float* A, float* B) { return transpose3d(id, A, B); }
int tmm_ref_apply_f64(int id, int Tr, const double* I0, const double* I1, const double* I2, double* B, double* A)
{ return apply(id, Tr, I0, I1, I2, B, A); }
int tmm_ref_apply_f32(int id, int Tr, const float* I0, const float* I1, const float* I2, float* B, float* A)
{ return apply(id, Tr, I0, I1, I2, B, A); }
}
