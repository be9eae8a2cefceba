#!/bin/bash
# The reference's UNCHANGED bench_stiffness / bench_mass (compat/_bin, built from /root/reference by
# compat/build_reference_drivers.sh) on 64 M DoFs, P = 3, fp64, with the host-pointer path fanned out over
# 1, 2, 4, ... visible GPUs (FEMTC_HOST_DEVICES) and the caller's std::vector fields page-locked in place
# (FEMTC_HOST_REGISTER=1).  Output columns: bench_mass.cpp:140-146 (last column = GDoF/s).
cd "$(dirname "$0")/../.."
N=$(nvidia-smi -L | wc -l)
echo "visible GPUs: $N; host CPUs: $(nproc)"
for bench in bench_stiffness bench_mass; do
  echo "== $bench 64000000 3 ikj 0 0 double, pageable std::vector (no registration), 1 GPU"
  ./compat/_bin/$bench 64000000 3 ikj 0 0 double 2>&1 | tail -1
  for n in 1 2 4 8; do
    [ $n -le $N ] || continue
    echo "== $bench 64000000 3 ikj 0 0 double  FEMTC_HOST_REGISTER=1 FEMTC_HOST_DEVICES=$n"
    FEMTC_HOST_REGISTER=1 FEMTC_HOST_DEVICES=$n ./compat/_bin/$bench 64000000 3 ikj 0 0 double 2>&1 | tail -1
  done
done
