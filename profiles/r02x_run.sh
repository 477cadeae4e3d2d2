for v in default compact_old; do
  echo "== compaction: $v"
  if [ $v = default ]; then unset B200PRIM_LIB; else export B200PRIM_LIB=cuda_samples_b200/variants/lib_$v.so; fi
  python profiles/compact_time.py
done
