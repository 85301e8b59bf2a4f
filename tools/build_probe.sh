#!/bin/bash
# Measurement build: the product sources with -DCILQR_PROBE_ENV (environment knobs for probes) -> build/libcilqr_probe.so,
# loaded with CILQR_B200_LIB=build/libcilqr_probe.so.  The product library is rebuilt afterwards.
set -e
cd "$(dirname "$0")/.."
P=uncertainty-aware-cilqr-for-trajectory-optimization_b200
OUT=${CILQR_PROBE_OUT:-build/libcilqr_probe.so}
CILQR_NVCC_EXTRA="-DCILQR_PROBE_ENV $*" python $P/build.py --force > /dev/null
mkdir -p build
cp $P/libcilqr_b200.so $OUT
python $P/build.py --force > /dev/null
ls -la $OUT $P/libcilqr_b200.so
