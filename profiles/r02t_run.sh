SORT_PAIRS=1 python profiles/sort_check.py fold 2>&1 | tee gpurun_out/r02t_sort_fold.log
python -m pytest tests -x -q -m gpu 2>&1 | tail -4
export PROFILE_REPS=2
python profiles/profile_kernels.py all > gpurun_out/r02t_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -o gpurun_out/r02t_full python profiles/profile_kernels.py all > gpurun_out/r02t_ncu.log 2>&1
echo ncu-full rc=$?
python bench.py --steps 2 --warmup 1 > gpurun_out/r02t_bench_plain.json 2> gpurun_out/r02t_bench_plain.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r02t_bench_launches.csv python bench.py --steps 2 --warmup 1 > gpurun_out/r02t_bench_under_ncu.log 2>&1
echo ncu-launches rc=$?
ls -la gpurun_out/
