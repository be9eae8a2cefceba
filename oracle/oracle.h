/* TEST INFRASTRUCTURE ONLY.
 *
 * CPU oracle: a plain-C restatement of the arithmetic of IgorBaratta/femtc's
 * mass / stiffness operator action and the small-GEMM layer beneath it.  It exists
 * to CHECK the CUDA path; nothing under femtc_b200/ or include/ may call it.
 * Allowed users: tests/, __graft_entry__.smoke(), and bench.py's cpu_baseline /
 * --impl reference legs.
 *
 * Parity status: PINNED.  Every entry point below is compared bit-for-bit (fp64 and
 * fp32, -O2 -ffp-contract=off on both sides) against the unmodified reference
 * headers compiled into oracle/_ref/libfemtc_ref.so (tests/test_oracle.py), against
 * the committed fixtures in tests/golden/ that were produced by that library, and
 * against the known-answer table of SURVEY.md section 8c.
 *
 * Conventions (n1 = P+1, nd = n1^3):
 *   basis   A(i,p) = phi[i + p*n1]            (column-major; linalg.hpp:49 with lda=m, :113)
 *   deriv   D(i,p) = phi[n1*n1 + i + p*n1]    (stiffness.hpp:45)
 *   cell tensors are row-major X[b0,b1,b2] = X[(b0*n1+b1)*n1+b2], cells contiguous
 *   loop order codes 0..5 = ijk, ikj, jik, jki, kij, kji  (linalg.hpp:11-19)
 */
#ifndef FEMTC_ORACLE_H
#define FEMTC_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum { ORACLE_STIFFNESS_REF_EXACT = 0, ORACLE_STIFFNESS_FULL = 1 };

/* c(m x nc, row-major, ldc) += a(m x k, col-major, lda) * b(k x nc, row-major, ldb).  linalg.hpp:45-94 */
int oracle_gemm_micro_f64(int k, int m, int nc, int lda, int ldb, int ldc, int order,
                          const double* a, const double* b, double* c);
int oracle_gemm_micro_f32(int k, int m, int nc, int lda, int ldb, int ldc, int order,
                          const float* a, const float* b, float* c);

/* C(m x n, col-major ld=m) = A(m x k, col-major ld=m) * B(k x n, row-major ld=n); C is
 * OVERWRITTEN (linalg.hpp:128,134-136), tiled MB x NB (0 = whole extent).  linalg.hpp:101-182 */
int oracle_gemm_blocked_f64(int k, int m, int n, int order, int MB, int NB,
                            const double* a, const double* b, double* c);
int oracle_gemm_blocked_f32(int k, int m, int n, int order, int MB, int NB,
                            const float* a, const float* b, float* c);

/* W = (A x A x A) diag(detJ) (A x A x A) U per cell, same A both ways.  mass.hpp:10-49 */
int oracle_mass_f64(const double* phi, const double* U, double* W, const double* detJ,
                    size_t num_cells, int degree, int order, int Mb, int Nb);
int oracle_mass_f32(const float* phi, const float* U, float* W, const float* detJ,
                    size_t num_cells, int degree, int order, int Mb, int Nb);

/* mode REF_EXACT: the reference as coded (stiffness.hpp:34-99): only the last of the three
 * back-contractions survives and Wz carries no derivative:
 *     W = (A x A x D) [ g2 (D x A x A)U + g1 (A x D x A)U + g0 (A x A x A)U ]
 * mode FULL: the intended operator, x+y+z accumulated, Wz = (A x A x D)U. */
int oracle_stiffness_f64(const double* phi, const double* U, double* W, const double* G,
                         size_t num_cells, int degree, int order, int Mb, int Nb, int mode);
int oracle_stiffness_f32(const float* phi, const float* U, float* W, const float* G,
                         size_t num_cells, int degree, int order, int Mb, int Nb, int mode);

/* A[a0,a1,a2] = sum I0(a0,b0) I1(a1,b1) I2(a2,b2) B[b0,b1,b2]; I(a,b) = Tr ? I[a*Nb+b] : I[b*Na+a].
 * tmm.hpp:91-126 */
int oracle_apply_contractions_f64(int Na, int Nb, int Tr, const double* I0, const double* I1,
                                  const double* I2, const double* B, double* A);
int oracle_apply_contractions_f32(int Na, int Nb, int Tr, const float* I0, const float* I1,
                                  const float* I2, const float* B, float* A);

/* string -> loop-order code, -1 where linalg::string2order throws (linalg.hpp:21-37). */
int oracle_string2order(const char* s);

/* Synthetic inputs.
 * rand: x = rand()/(T)RAND_MAX, the benches' generator (bench_mass.cpp:74-88); glibc rand().
 * splitmix: counter-based, x_i = top bits of splitmix64(seed + stream*GAMMA + offset + i),
 *           uniform in [0,1) (or [-0.5,0.5) when centred != 0); the CUDA library has the
 *           same generator so any cell range can be regenerated on either side. */
void oracle_srand(unsigned seed);
void oracle_fill_rand_f64(double* x, size_t n);
void oracle_fill_rand_f32(float* x, size_t n);
void oracle_fill_splitmix_f64(double* x, size_t n, uint64_t seed, uint64_t stream, uint64_t offset, int centred);
void oracle_fill_splitmix_f32(float* x, size_t n, uint64_t seed, uint64_t stream, uint64_t offset, int centred);

/* Kahan-free plain checksums in double: out = {sum, sum of squares, max |x|, n}. */
void oracle_checksum_f64(const double* x, size_t n, double out[4]);
void oracle_checksum_f32(const float* x, size_t n, double out[4]);

#ifdef __cplusplus
}
#endif
#endif
