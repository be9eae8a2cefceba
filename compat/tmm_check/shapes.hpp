// Shapes both TUs of the tmm.hpp building-block check instantiate (the templates take compile-time sizes).
// One X-macro per building block; the index of an entry is its id in the C interface between the two TUs.
#pragma once
// 2-D contract C[a,b] += sum_k A[a,k] B[b,k]: (Na, Nb, Nk)            reference src/tmm.hpp:24-32
#define TMM_CONTRACT2D(X) X(0, 4, 4, 4) X(1, 6, 4, 5) X(2, 3, 5, 2) X(3, 9, 9, 9) X(4, 1, 7, 3)
// 3-D contract C[a,(b,c)] += sum_k M(a,k) B[k,(b,c)]: (Nk, Na, Nb, Nc), both `transpose` values   :65-85
#define TMM_CONTRACT3D(X) X(0, 4, 4, 4, 4) X(1, 4, 6, 4, 4) X(2, 5, 3, 5, 5) X(3, 9, 9, 9, 9) X(4, 3, 5, 2, 7) X(5, 4, 6, 6, 4)
// 2-D transpose B[a*offa + b*offb] = A[a,b]: (Na, Nb, offa, offb)      :9-15
#define TMM_TRANSPOSE2D(X) X(0, 4, 6, 1, 4) X(1, 5, 3, 3, 1) X(2, 9, 9, 1, 9) X(3, 2, 7, 1, 2)
// 3-D transpose B[offa*a + offb*b + offc*c] = A[a,b,c]: (Na, Nb, Nc, offa, offb, offc)   :37-44
// (the three permutations apply_contractions uses, :105,:115,:125, at rectangular sizes, plus the identity)
#define TMM_TRANSPOSE3D(X)                                                                                              \
  X(0, 6, 4, 4, 4, 24, 1) X(1, 6, 6, 4, 1, 6, 36) X(2, 6, 6, 6, 1, 36, 6) X(3, 3, 5, 5, 5, 15, 1) X(4, 3, 3, 5, 1, 3, 9)  \
  X(5, 2, 3, 4, 12, 4, 1) X(6, 9, 9, 9, 1, 81, 9)
// apply_contractions: (Na, Nb), both Tr values                         :91-126
#define TMM_APPLY(X) X(0, 4, 4) X(1, 6, 4) X(2, 3, 5) X(3, 9, 9) X(4, 2, 2) X(5, 5, 4)

extern "C"
{
// implemented by ref_side.cpp from the UNMODIFIED reference header; return 0, or 1 for an unknown id
int tmm_ref_contract2d_f64(int id, const double* A, const double* B, double* C);
int tmm_ref_contract2d_f32(int id, const float* A, const float* B, float* C);
int tmm_ref_contract3d_f64(int id, int transpose, const double* A, const double* B, double* C);
int tmm_ref_contract3d_f32(int id, int transpose, const float* A, const float* B, float* C);
int tmm_ref_transpose2d_f64(int id, const double* A, double* B);
int tmm_ref_transpose2d_f32(int id, const float* A, float* B);
int tmm_ref_transpose3d_f64(int id, double* A, double* B);
int tmm_ref_transpose3d_f32(int id, float* A, float* B);
int tmm_ref_apply_f64(int id, int Tr, const double* I0, const double* I1, const double* I2, double* B, double* A);
int tmm_ref_apply_f32(int id, int Tr, const float* I0, const float* I1, const float* I2, float* B, float* A);
}
