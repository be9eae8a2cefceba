// TEST INFRASTRUCTURE ONLY (oracle). Not part of the shipped product path.
//
// Shared helpers for the shims that compile the UNMODIFIED reference headers
// (/root/reference/src/*.hpp, found through -I) into oracle/_ref/*.so.
// No reference source is copied: the shims only #include the headers from where
// they lie and instantiate a handful of the reference's templates behind a C ABI.
#pragma once
#include <cstddef>

// Tuning-knob triples (loop order, Mb, Nb) of the reference that the shim
// instantiates.  They only change rounding, so a few suffice.
//   cfg 0: ijk, 0, 0   (the triple used for the known-answer table)
//   cfg 1: ikj, 0, 0   (fastest loop order on the survey host)
//   cfg 2: kij, 8, 16  (the tiling tests/test_gemm.cpp:41 exercises)
//   cfg 3: jki, 4, 8   (exercises the remainder paths of gemm_blocked)
#define FEMTC_REF_NUM_CFG 4

#define FEMTC_REF_FOR_EACH_DEGREE(X) X(1) X(2) X(3) X(4) X(5) X(6) X(7) X(8) X(9) X(10)
