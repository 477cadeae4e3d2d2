#!/bin/bash
# Build compile-time variants of the sort pass (cuda_samples_b200/variants/lib_<name>.so) and, on a GPU box,
# check + time each:   profiles/sort_variants.sh build | run
set -e
cd "$(dirname "$0")/.."
declare -A V=(
  [t256c4]=""
  [t256c4_ldsts]="-DB200_SORT_RANK_MODE=1"
  [t256c5]="-DB200_SORT_CTAS=5"
  [t512c2]="-DB200_SORT_THREADS=512 -DB200_SORT_CTAS=2"
  [t512c2_i8]="-DB200_SORT_THREADS=512 -DB200_SORT_CTAS=2 -DB200_SORT_ITEMS=8"
  [t256c4_i12]="-DB200_SORT_ITEMS=12"
)
if [ "$1" = build ]; then
  for n in "${!V[@]}"; do
    B200_EXTRA_FLAGS="${V[$n]} -DB200_VARIANT_BUILD" B200_VARIANT=$n python cuda_samples_b200/build.py >/dev/null
    grep -A2 "sort_pass_kernelILb0ELi4" cuda_samples_b200/build_$n/radix_sort.o.ptxas.txt | grep -E "Used|spill" | tr '\n' ' '; echo " <- $n"
  done
else
  for n in "${!V[@]}"; do
    B200PRIM_LIB=cuda_samples_b200/variants/lib_$n.so python profiles/sort_check.py $n 2>&1 | tail -4
  done
fi
