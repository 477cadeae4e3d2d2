#!/bin/bash
# r02: time (and parity-check the first of) the early-count variants of the sort pass built into
# cuda_samples_b200/variants/ by B200_EXTRA_FLAGS=... B200_VARIANT=<name> python cuda_samples_b200/build.py
cd "$(dirname "$0")/.."
first=1
for n in "$@"; do
  if [ $first = 1 ]; then
    B200PRIM_LIB=cuda_samples_b200/variants/lib_$n.so python profiles/sort_check.py $n 2>&1 | tail -3
    first=0
  else
    SORT_CHECK_SKIP_PARITY=1 B200PRIM_LIB=cuda_samples_b200/variants/lib_$n.so python profiles/sort_check.py $n 2>&1 | tail -2
  fi
done
