# round 3, one B200: kept bench (both arms), ncu --set full of every hot-path kernel, launch list of the bench
python bench.py --impl reference > gpurun_out/r03_bench_reference.json 2> gpurun_out/r03_bench_reference.err; echo reference rc=$?
python bench.py > gpurun_out/r03_bench.json 2> gpurun_out/r03_bench.err; echo bench rc=$?
python profiles/hist_isolated.py > gpurun_out/r03_hist_isolated.log 2>&1; echo hist-isolated rc=$?
export PROFILE_REPS=2
python profiles/profile_kernels.py all > gpurun_out/r03_plain.log 2>&1 && ncu --set full --clock-control none -o /tmp/r03_full python profiles/profile_kernels.py all > gpurun_out/r03_ncu.log 2>&1
echo ncu-full rc=$?
python profiles/summarize.py /tmp/r03_full.ncu-rep gpurun_out/r03_ncu_full_summary.csv
python bench.py --steps 2 --warmup 1 --skip-table > gpurun_out/r03_bench_plain.json 2> gpurun_out/r03_bench_plain.err && ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r03_bench_launches.csv python bench.py --steps 2 --warmup 1 --skip-table > gpurun_out/r03_bench_under_ncu.log 2>&1
echo ncu-launches rc=$?
python baseline/run.py gpurun_out/r03_baseline > gpurun_out/r03_baseline.log 2>&1; echo baseline rc=$?
ls -la gpurun_out/ | tail -20
