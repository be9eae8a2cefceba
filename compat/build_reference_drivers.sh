#!/bin/bash
# Builds the reference's drivers and tests UNCHANGED (sources are compiled from where they lie
# under $REF, nothing is copied) against include/*.hpp + libfemtc_b200.so.  Outputs go to
# compat/_bin (git-ignored, travels to the GPU box).  Warning flags are the reference's own
# (tests/CMakeLists.txt:11, CMakeLists.txt:11).
set -euo pipefail
REF=${REF:-/root/reference}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=$ROOT/compat/_bin
[ -d "$REF" ] || { echo "reference tree $REF absent: keeping prebuilt $OUT"; exit 0; }
mkdir -p "$OUT"
CXX=${CXX:-g++}
COMMON="-std=c++20 -O2 -DNDEBUG -Wall -Wextra -pedantic -I$ROOT/include -L$ROOT/femtc_b200 -Wl,-rpath,\$ORIGIN/../../femtc_b200"
$CXX $COMMON -Wfatal-errors -Werror "$REF/tests/test_gemm.cpp"  -o "$OUT/test_gemm"  -lfemtc_b200 &
$CXX $COMMON -Wfatal-errors -Werror "$REF/tests/test_micro.cpp" -o "$OUT/test_micro" -lfemtc_b200 &
$CXX $COMMON -I"$ROOT/compat/mpi_stub" "$REF/bench_mass.cpp"      -o "$OUT/bench_mass"      -lfemtc_b200 &
$CXX $COMMON -I"$ROOT/compat/mpi_stub" "$REF/bench_stiffness.cpp" -o "$OUT/bench_stiffness" -lfemtc_b200 &
# run-time parity of the tmm.hpp building blocks: include/tmm.hpp (main.cpp) against the reference's src/tmm.hpp
# (ref_side.cpp, the header compiled unmodified inside its own namespace), two TUs of one binary
$CXX -std=c++20 -O2 -ffp-contract=off -DNDEBUG -I"$REF/src" -I"$ROOT/compat/tmm_check" -c "$ROOT/compat/tmm_check/ref_side.cpp" -o "$OUT/tmm_ref_side.o" &
wait
$CXX $COMMON -Wfatal-errors -Werror -I"$ROOT/compat/tmm_check" "$ROOT/compat/tmm_check/main.cpp" "$OUT/tmm_ref_side.o" -o "$OUT/test_tmm_blocks" -lfemtc_b200
rm -f "$OUT/tmm_ref_side.o"
ls -la "$OUT"
