/* femtc_b200.h -- C ABI of libfemtc_b200.so
 *
 * B200 (sm_100a) implementation of the cell-batched, sum-factorised mass and stiffness
 * operator action of IgorBaratta/femtc.  The reference is header-only C++ with no FFI
 * layer; its callers include src/mass.hpp, src/stiffness.hpp, src/linalg.hpp (and, for the
 * orphaned alternative formulation, src/tmm.hpp) and call templates.  The replacement
 * headers shipped next to this file (include/{mass,stiffness,linalg,tmm}.hpp) keep those
 * names and signatures and forward to the entry points declared here; every declaration
 * cites the reference interface it stands in for.
 *
 * Conventions shared by all entry points (n1 = degree+1, nd = n1^3):
 *   basis    A(i,p) = phi[i + p*n1]          column-major      (linalg.hpp:49 with lda=m, :113)
 *   deriv    D(i,p) = phi[n1*n1 + i + p*n1]                     (stiffness.hpp:44-45)
 *   U, W, detJ : num_cells*nd values, cell-major contiguous, each cell row-major
 *                X[b0,b1,b2] = X[(b0*n1+b1)*n1+b2]              (mass.hpp:25-29)
 *   G        : num_cells*nd*6 values, six per point, packed [zz,yz,xz,yy,xy,xx]
 *                                                               (stiffness.hpp:20-27)
 *   order    : 0..5 = ijk, ikj, jik, jki, kij, kji              (linalg.hpp:11-19)
 *   Mb, Nb   : 0, 4, 8 or 16                                    (mass.hpp:171-219)
 * order/Mb/Nb are CPU loop-nest knobs that change nothing but (on a fast-math CPU build)
 * rounding.  They are validated exactly like the reference validates them and otherwise
 * ignored by the GPU kernels.
 *
 * Ownership: the caller owns every buffer; nothing is retained after return.  W is pure
 * output and is overwritten for every cell; inputs are read-only; W must not alias inputs
 * (the reference declares them __restrict__, linalg.hpp:6,47,102).
 *
 * Errors: 0 on success, a FEMTC_ERR_* code otherwise; W is untouched when an argument is
 * rejected.  Nothing throws across this boundary.  femtc_last_error() describes the last
 * failure on the calling thread.  There is no CPU fallback: without a usable sm_100 device
 * every compute entry point returns FEMTC_ERR_CUDA.
 *
 * Threading: entry points are thread-safe and re-entrant.  Host-pointer entry points are
 * synchronous (results are in the caller's buffers on return).  *_device_* entry points
 * only enqueue work on the given stream (cudaStream_t passed as void*; NULL = the legacy
 * default stream) of the current device and return immediately.
 */
#ifndef FEMTC_B200_H
#define FEMTC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FEMTC_OK 0
#define FEMTC_ERR_DEGREE 1 /* degree outside 1..25     -> reference prints "degree not supported"  (mass.hpp:135) */
#define FEMTC_ERR_ORDER 2  /* order outside 0..5       -> "order not supported"                    (mass.hpp:166) */
#define FEMTC_ERR_BLOCK 3  /* Mb/Nb not in {0,4,8,16}  -> "Block size not supported"               (mass.hpp:191,216) */
#define FEMTC_ERR_MODE 4   /* unknown stiffness mode */
#define FEMTC_ERR_LENGTH 5 /* phi shorter than the operator reads */
#define FEMTC_ERR_CUDA 6   /* CUDA runtime failure, or no sm_100 device */
#define FEMTC_ERR_ARG 7    /* null pointer / bad shape */

/* Stiffness semantics.  REF_EXACT reproduces the reference as coded (stiffness.hpp:68-97):
 * the three back-contractions each overwrite W, so only the last survives, and Wz is built
 * without the derivative matrix:
 *     W = (A x A x D) [ g2 (D x A x A)U + g1 (A x D x A)U + g0 (A x A x A)U ].
 * FULL is the operator the reference's flop model (bench_stiffness.cpp:122) describes:
 *     W = (D x A x A)f0 + (A x D x A)f1 + (A x A x D)f2,  f = G (DxAxA U, AxDxA U, AxAxD U). */
#define FEMTC_STIFFNESS_REF_EXACT 0
#define FEMTC_STIFFNESS_FULL 1

/* ---- operators, host pointers, synchronous ------------------------------------------
 * Replaces operators::mass_operator<T>(phi,U,W,detJ,num_cells,degree,order,Mb,Nb)
 * (reference src/mass.hpp:196-199; call site bench_mass.cpp:98).  phi_len >= n1*n1. */
int femtc_mass_apply_f64(const double* phi, size_t phi_len, const double* U, double* W,
                         const double* detJ, size_t num_cells, int degree, int order, int Mb, int Nb);
int femtc_mass_apply_f32(const float* phi, size_t phi_len, const float* U, float* W,
                         const float* detJ, size_t num_cells, int degree, int order, int Mb, int Nb);

/* Replaces operators::stiffness_operator<T>(phi,U,W,G,num_cells,degree,order,Mb,Nb)
 * (reference src/stiffness.hpp:245-248; call site bench_stiffness.cpp:98).
 * The operator reads phi[0 .. 2*n1*n1) (stiffness.hpp:44-45).  bench_stiffness.cpp:74 passes
 * only n1*n1 values and the reference then reads out of bounds; here a short phi is never
 * over-read: with n1*n1 <= phi_len < 2*n1*n1 the derivative matrix is taken equal to the basis
 * matrix and a warning is printed once on stderr; below n1*n1 the call fails with
 * FEMTC_ERR_LENGTH.  num_cells is size_t and indexing is 64-bit (the reference overflows `int`
 * offsets beyond 2^31/(6 nd) cells, stiffness.hpp:55-57). */
int femtc_stiffness_apply_f64(const double* phi, size_t phi_len, const double* U, double* W,
                              const double* G, size_t num_cells, int degree, int order, int Mb,
                              int Nb, int mode);
int femtc_stiffness_apply_f32(const float* phi, size_t phi_len, const float* U, float* W,
                              const float* G, size_t num_cells, int degree, int order, int Mb,
                              int Nb, int mode);

/* ---- operators, device pointers, asynchronous on `stream` ---------------------------
 * Same operators (mass.hpp:10-49, stiffness.hpp:34-99) for callers whose fields already
 * live in HBM.  U, W, detJ/G are device pointers on the current device; phi is a HOST
 * pointer (the tables are tiny and travel as kernel parameters).  16-byte aligned field
 * pointers take the TMA fast path; others are served by a slower kernel. */
int femtc_mass_apply_device_f64(const double* phi_host, size_t phi_len, const double* dU, double* dW,
                                const double* ddetJ, size_t num_cells, int degree, void* stream);
int femtc_mass_apply_device_f32(const float* phi_host, size_t phi_len, const float* dU, float* dW,
                                const float* ddetJ, size_t num_cells, int degree, void* stream);
int femtc_stiffness_apply_device_f64(const double* phi_host, size_t phi_len, const double* dU,
                                     double* dW, const double* dG, size_t num_cells, int degree,
                                     int mode, void* stream);
int femtc_stiffness_apply_device_f32(const float* phi_host, size_t phi_len, const float* dU,
                                     float* dW, const float* dG, size_t num_cells, int degree,
                                     int mode, void* stream);

/* Same, with the checksum of W produced by the operator itself: d_checksum4 (DEVICE pointer to 4 doubles, may be
 * NULL) receives {sum, sum of squares, max |w|, n} of the whole W on the same stream -- what the reference's benches
 * reduce after their timed loop (bench_mass.cpp:135-137: three MPI_Reduce of per-rank figures; here the per-GPU four
 * are combined by one all-gather, femtc_b200/distributed.py).  The slab kernels (degrees 1..4) accumulate it in their
 * epilogue from the W values they hold in registers, one partial per warp, without reading W again; other kernel
 * families and the ragged tail fall back to the stand-alone reduction (femtc_checksum_device_*).  A second-stage kernel
 * sums the partials in a fixed order: deterministic for a given batch size, equal to femtc_checksum_device_* to
 * rounding.  Calls on one device must be stream-ordered with respect to each other (one scratch buffer per device). */
int femtc_mass_apply_device_ex_f64(const double* phi_host, size_t phi_len, const double* dU, double* dW,
                                   const double* ddetJ, size_t num_cells, int degree, double* d_checksum4, void* stream);
int femtc_mass_apply_device_ex_f32(const float* phi_host, size_t phi_len, const float* dU, float* dW,
                                   const float* ddetJ, size_t num_cells, int degree, double* d_checksum4, void* stream);
int femtc_stiffness_apply_device_ex_f64(const double* phi_host, size_t phi_len, const double* dU, double* dW,
                                        const double* dG, size_t num_cells, int degree, int mode, double* d_checksum4,
                                        void* stream);
int femtc_stiffness_apply_device_ex_f32(const float* phi_host, size_t phi_len, const float* dU, float* dW,
                                        const float* dG, size_t num_cells, int degree, int mode, double* d_checksum4,
                                        void* stream);

/* ---- operator handle: geometry resident in HBM (set up once, apply many) ---------------
 * The reference's drivers apply the operator nreps times to the SAME detJ / G
 * (bench_mass.cpp:95-105, bench_stiffness.cpp:95-105); in a matrix-free solver these arrays are
 * set-up data.  A handle uploads them once (num_cells*nd*{1|6} values, HOST pointer, copied --
 * the caller may free `geom` after create) together with phi; an application then moves only U in
 * and W out.  kind: 0 = mass (geom = detJ), 1 = stiffness (geom = G); mode: FEMTC_STIFFNESS_* (ignored
 * for mass).  Same kernels, same results bit for bit as femtc_{mass,stiffness}_apply_*.
 *   femtc_operator_apply         U, W HOST pointers (dtype of the handle), synchronous, pipelined
 *                                H2D(U) -> kernel -> D2H(W) in chunks of "host_chunk_mb"
 *   femtc_operator_apply_device  U, W DEVICE pointers, asynchronous on `stream`
 * A handle is bound to the device that was current at creation; it may be applied from any
 * thread whose current device is that one, but not concurrently with its own destruction. */
typedef struct femtc_operator femtc_operator;
int femtc_operator_create_f64(femtc_operator** op, int kind, const double* phi, size_t phi_len,
                              const double* geom, size_t num_cells, int degree, int mode);
int femtc_operator_create_f32(femtc_operator** op, int kind, const float* phi, size_t phi_len,
                              const float* geom, size_t num_cells, int degree, int mode);
int femtc_operator_apply(femtc_operator* op, const void* U, void* W);
int femtc_operator_apply_device(femtc_operator* op, const void* dU, void* dW, void* stream);
size_t femtc_operator_device_bytes(const femtc_operator* op); /* HBM held by the handle */
void femtc_operator_destroy(femtc_operator* op);

/* ---- small-GEMM layer, host pointers, synchronous ------------------------------------
 * Replaces linalg::gemm_blocked<T,k,m,n,order,MB,NB>(a,b,c) (reference src/linalg.hpp:101-182;
 * call sites tests/test_gemm.cpp:41,52): C(m x n, column-major ld=m) = A(m x k, column-major
 * ld=m) . B(k x n, row-major ld=n).  C is OVERWRITTEN, as in the reference (linalg.hpp:128,
 * 134-136), whatever its comment says. */
int femtc_gemm_blocked_f64(int k, int m, int n, int order, int MB, int NB, const double* a,
                           const double* b, double* c);
int femtc_gemm_blocked_f32(int k, int m, int n, int order, int MB, int NB, const float* a,
                           const float* b, float* c);

/* Replaces linalg::gemm_micro<T,k,m,nc,lda,ldb,ldc,order>(a,b,c) (reference
 * src/linalg.hpp:45-94; call site tests/test_micro.cpp:44): c(m x nc, row-major ldc) +=
 * a(m x k, column-major lda) . b(k x nc, row-major ldb).  ACCUMULATES into c (linalg.hpp:57). */
int femtc_gemm_micro_f64(int k, int m, int nc, int lda, int ldb, int ldc, int order,
                         const double* a, const double* b, double* c);
int femtc_gemm_micro_f32(int k, int m, int nc, int lda, int ldb, int ldc, int order,
                         const float* a, const float* b, float* c);

/* c[i*ldc + j] += sum_p a[i*sai + p*sap] * b[p*sbp + j*sbj], i < m, j < nc, p ascending.
 * The strided form behind gemm_micro (sai=1, sap=lda, sbp=ldb, sbj=1); replaces the two
 * `contract` building blocks of reference src/tmm.hpp:24-32 and :65-85. */
int femtc_gemm_strided_acc_f64(int k, int m, int nc, int sai, int sap, int sbp, int sbj, int ldc,
                               const double* a, const double* b, double* c);
int femtc_gemm_strided_acc_f32(int k, int m, int nc, int sai, int sap, int sbp, int sbj, int ldc,
                               const float* a, const float* b, float* c);

/* Replaces apply_contractions<T,Na,Nb,Tr>(I0,I1,I2,B,A,buffer) (reference src/tmm.hpp:91-126):
 * A[a0,a1,a2] = sum I0(a0,b0) I1(a1,b1) I2(a2,b2) B[b0,b1,b2], I(a,b) = Tr ? I[a*Nb+b] : I[b*Na+a];
 * A (Na^3) is overwritten, B (Nb^3) is read.  `batch` independent tensors are processed,
 * B and A strided by Nb^3 and Na^3 (the reference is the batch == 1 case). */
int femtc_apply_contractions_f64(int Na, int Nb, int Tr, const double* I0, const double* I1,
                                 const double* I2, const double* B, double* A, size_t batch);
int femtc_apply_contractions_f32(int Na, int Nb, int Tr, const float* I0, const float* I1,
                                 const float* I2, const float* B, float* A, size_t batch);

/* The same contraction for tensors already in HBM: dB (batch*Nb^3) and dA (batch*Na^3) are DEVICE pointers,
 * I0, I1, I2 stay HOST pointers (they travel as kernel parameters); asynchronous on `stream`.  Shapes
 * Nb = 2..8 with Na = Nb-1 .. Nb+2 (restriction, collocation, over-integration by one or two points per
 * direction -- quadrature points != DoFs) and 16-byte aligned pointers run on a tiled TMA kernel in which every
 * tensor crosses HBM once in and once out ((Nb^3 + Na^3)*sizeof(T) bytes per tensor); every other shape is
 * served by a slower general kernel.  The host-pointer entry points above use the same kernels. */
int femtc_apply_contractions_device_f64(int Na, int Nb, int Tr, const double* I0, const double* I1,
                                        const double* I2, const double* dB, double* dA, size_t batch,
                                        void* stream);
int femtc_apply_contractions_device_f32(int Na, int Nb, int Tr, const float* I0, const float* I1,
                                        const float* I2, const float* dB, float* dA, size_t batch,
                                        void* stream);

/* No counterpart in the reference (its apply_contractions takes a caller-made Buffer on the CPU, tmm.hpp:46-60; the
 * question of who allocates GPU scratch does not arise there).
 *
 * femtc_apply_contractions_device_f64 for callers that cannot let the library allocate device memory.  The general
 * kernel behind that entry point asks the stream-ordered allocator for room for the three matrices and, for large
 * shapes, for its work arrays.  That is unwelcome in two kinds of program that use this library: a solver that reserved
 * all of HBM for its own memory arena when it started, where a hidden allocation either fails or takes memory the arena
 * has already promised; and a time-stepping loop with a latency budget, which cannot risk the device-wide
 * synchronisation an allocation or a free may bring.  This entry point calls no cudaMalloc / cudaFree of any flavour:
 * its scratch is a 4 MiB (4 194 304 byte) array that is part of the library image, so every GPU that loads
 * libfemtc_b200.so holds those 4 MiB until the process exits, used or not.
 *
 * Arguments, layouts and result as femtc_apply_contractions_device_f64 (I0, I1, I2 HOST pointers, dB / dA DEVICE
 * pointers, asynchronous on `stream`), always on the general kernel (the tiled TMA kernel is not used here), so results
 * agree with the other entry point to rounding.
 * Limit: 8 * (3*Na*Nb rounded up to 32 values + Na*Nb*(Na+Nb)) bytes must fit in the array -- e.g. Na = Nb <= 63; a
 * larger shape returns FEMTC_ERR_ARG, says so in femtc_last_error() and leaves dA untouched.  Large shapes run on
 * fewer blocks (as many work slices as fit), i.e. slower than the allocating entry point.
 * First call on each device: looks up the array's device address (once per device, remembered).  Make that first call
 * ahead of any stream capture.
 * Sharing: all calls on a device use the same array and it keeps what the last call wrote.  Nothing needs to start at
 * zero (a call reads only what it wrote itself), but two calls that can run on the GPU at the same time -- different
 * streams with no dependency between them -- corrupt each other: keep calls on one device stream-ordered. */
int femtc_apply_contractions_device_static_f64(int Na, int Nb, int Tr, const double* I0, const double* I1,
                                               const double* I2, const double* dB, double* dA, size_t batch,
                                               void* stream);

/* linalg::string2order (reference src/linalg.hpp:21-37): returns 0..5, or -1 where the
 * reference throws std::runtime_error("Invalid loop order"). */
int femtc_string2order(const char* name);

/* ---- measurement / multi-GPU helpers (device pointers, asynchronous) ------------------
 * Checksum of a device array in double precision: out4 (DEVICE pointer to 4 doubles) =
 * {sum, sum of squares, max |x|, n}.  The per-GPU partials of a cell-partitioned run are
 * combined by one all-reduce of these four numbers (the analogue of the three MPI_Reduce
 * calls in bench_mass.cpp:135-137).  Deterministic for a given n. */
int femtc_checksum_device_f64(const double* dX, size_t n, double* d_out4, void* stream);
int femtc_checksum_device_f32(const float* dX, size_t n, double* d_out4, void* stream);

/* Counter-based uniform inputs generated in place on the device:
 *   x[i] = top bits of splitmix64(seed + stream_id*0x9E3779B97F4A7C15 + offset + i), in [0,1)
 *   (minus 0.5 when centred != 0).  Same generator as oracle_fill_splitmix_* so any cell
 * range can be regenerated on the host.  Stands in for the std::rand inputs of
 * bench_mass.cpp:74-88 at sizes where a host generator is impractical. */
int femtc_fill_uniform_device_f64(double* dX, size_t n, uint64_t seed, uint64_t stream_id,
                                  uint64_t offset, int centred, void* stream);
int femtc_fill_uniform_device_f32(float* dX, size_t n, uint64_t seed, uint64_t stream_id,
                                  uint64_t offset, int centred, void* stream);

/* Streaming probe for the roofline denominator: d_out[i] = d_in[i] + d_in[n+i] + ... over `reads` back-to-back
 * streams of bytes_per_stream bytes each (16-byte words, fp32 adds).  reads = 7 / 2 is the read:write mix of the
 * stiffness / mass action (bench_stiffness.cpp:123-124 counts the same streams), 1 is a device copy.  No
 * reference counterpart; measurement only. */
int femtc_stream_probe_device(const void* d_in, void* d_out, size_t bytes_per_stream, int reads, void* stream);

/* ---- runtime --------------------------------------------------------------------------- */
int femtc_device_count(void);             /* number of usable sm_100 devices, 0 if none */
int femtc_set_device(int device);         /* device used by the calling thread */
int femtc_get_device(void);
const char* femtc_last_error(void);       /* thread-local, never NULL */
const char* femtc_version(void);
void* femtc_host_alloc(size_t bytes);     /* page-locked host memory for the host-pointer path */
void femtc_host_free(void* p);
void femtc_release_cache(void);           /* drop cached device staging buffers and host registrations */

/* Tuning / introspection knobs, by name.  Unknown names return FEMTC_ERR_ARG / -1.
 *   "kernel"          0 = automatic, 1 = force the generic kernel, 2 = force the tiled kernel
 *   "host_chunk_mb"   input megabytes per pipelined chunk of the host-pointer path
 *   "host_register"   1 = page-lock pageable caller buffers (>= 1 MB) in place the first time the
 *                     host-pointer path sees them and remember the registration by address, so repeated
 *                     calls on the same buffers (bench_mass.cpp:95-105) DMA at the PCIe rate.  The
 *                     caller must call femtc_release_cache() before freeing such a buffer.  Default 0,
 *                     or 1 when FEMTC_HOST_REGISTER=1 is in the environment at load time.
 *   "host_devices"    n > 1: ONE host-pointer operator call is cut into n contiguous cell ranges served by devices
 *                     0..n-1 concurrently (one worker thread and one three-slot copy/compute pipeline per device), so
 *                     the unchanged drivers (bench_stiffness.cpp:95-105) move their fields over n PCIe links.  The
 *                     action is cell-local: the result is bit-identical to the one-device call.  Default 1, or
 *                     FEMTC_HOST_DEVICES from the environment at load time; clamped to the visible devices.
 *                     Every process that inherits the variable opens contexts on devices 0..n-1: leave it unset under
 *                     a one-rank-per-GPU launcher (mpirun, torchrun), where each rank already owns a device.
 *   "host_devices_oversubscribe"  1: the share count is not clamped and share r runs on device r % visible -- lets a
 *                     one-GPU box exercise the fan-out (cut points, worker threads, error hand-over).  Default 0, or
 *                     FEMTC_HOST_DEVICES_OVERSUBSCRIBE from the environment.
 *   "geom_cache"      1: the host-pointer operator entry points keep detJ / G resident in HBM between calls, keyed by
 *                     (device, host address, bytes) and checked against a 64-bit hash of 4 KB sampled across the array;
 *                     a repeat call (bench_stiffness.cpp:95-105 applies the operator to the same G every repetition)
 *                     then moves only U in and W out -- same kernels, bit-identical.  A hash mismatch re-uploads;
 *                     femtc_release_cache() frees the copies.  A caller that rewrites geometry in place between calls
 *                     must call femtc_release_cache() (the samples can miss a local change) or leave this off.  Not
 *                     combined with "host_devices" > 1.  Default 0, or FEMTC_GEOM_CACHE from the environment.
 *   "geom_cache_uploads"  (read-only) geometry uploads the cache has made so far
 *   "tc"              DMMA tensor-core family (fp64, degrees 5..8, mass / REF_EXACT): -1 = where it is
 *                     measured faster than the FMA kernels (default), 0 = never, 1 = wherever applicable
 *   "variant"         development knob for A/B measurements (0 = tuned defaults).  1, 2: alternative pipeline
 *                     depths; 9: one-thread-per-cell family off; 3..8: geometry transport of the column
 *                     stiffness kernels (producer warp / warp-private rings, ring depths); 10..12: mass column
 *                     kernels with two cells / one cell per thread; 13..15: two-cell stiffness kernel;
 *                     20, 21: slab kernel on / off at n1 = 6..8.  Results never depend on it beyond rounding
 *                     (kernels of one family are bit-identical; tests/test_gpu_parity.py).
 *   "launches"        (read-only) kernels launched by this library so far in this process */
int femtc_set_option(const char* name, long long value);
long long femtc_get_option(const char* name);

/* Which kernel serves a large aligned batch of (op, dtype, degree, mode) on the calling thread's device, under the
 * current options -- for reports.  The real dispatcher is walked in a dry run (nothing is launched).
 * op 0 = mass, 1 = stiffness; dtype_bytes 4 or 8.  Fills info[0..7] =
 * {kernel family (0 generic, 1 slab, 2 one-thread-per-cell, 3 column, 4 tensor core), cells per tile (slab: per warp round),
 *  threads per block, dynamic smem bytes, resident blocks per SM (0: not planned), registers per thread, grid blocks,
 *  family-specific detail (column: cells per thread + 10 x ring depth + 1000 if block-wide ring; tensor core: warps
 *  per cell + 10 x ring slots; slab: geometry ring depth; cell: pipeline stages)}. */
int femtc_kernel_info(int op, int dtype_bytes, int degree, int mode, long long info[8]);

#ifdef __cplusplus
}
#endif
#endif /* FEMTC_B200_H */
