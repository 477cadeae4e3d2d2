for v in default lw2 lw3; do
  echo "== look-back warps: $v"
  if [ $v = default ]; then unset B200PRIM_LIB; else export B200PRIM_LIB=cuda_samples_b200/variants/lib_$v.so; fi
  python profiles/scan_park.py
done
unset B200PRIM_LIB
for m in 0 2; do SCAN_MODE=$m B200PRIM_LIB=cuda_samples_b200/libb200prim_trace.so python profiles/scan_trace.py; done
