python bench.py --steps 20 --warmup 5 > gpurun_out/r02z_bench.json 2> gpurun_out/r02z_bench.err; echo bench rc=$?
export PROFILE_REPS=2
python profiles/profile_kernels.py all > gpurun_out/r02z_plain.log 2>&1 && ncu --set full --clock-control none -o /tmp/r02z_full python profiles/profile_kernels.py all > gpurun_out/r02z_ncu.log 2>&1
echo ncu-full rc=$?
python profiles/summarize.py /tmp/r02z_full.ncu-rep gpurun_out/r02z_ncu_full_summary.csv
python profiles/profile_kernels.py sort > gpurun_out/r02z_plain_sort.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sort_pass_kernel -s 5 -c 2 -o gpurun_out/r02z_sort_pass python profiles/profile_kernels.py sort >> gpurun_out/r02z_ncu.log 2>&1
echo ncu-sort rc=$?
python bench.py --steps 2 --warmup 1 > gpurun_out/r02z_bench_plain.json 2> gpurun_out/r02z_bench_plain.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r02z_bench_launches.csv python bench.py --steps 2 --warmup 1 > gpurun_out/r02z_bench_under_ncu.log 2>&1
echo ncu-launches rc=$?
ls -la gpurun_out/ /tmp/r02z_full.ncu-rep
