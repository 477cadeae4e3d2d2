timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r02z_bench_8gpu.json 2> gpurun_out/r02z_bench_8gpu.err; echo bench8 rc=$?
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 tests/multi_gpu_check.py > gpurun_out/r02z_multi_gpu_check_8gpu.log 2>&1; echo check8 rc=$?; tail -3 gpurun_out/r02z_multi_gpu_check_8gpu.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 profiles/sort_sharded_phases.py > gpurun_out/r02z_sort_phases_8gpu.log 2>&1; echo phases rc=$?; tail -7 gpurun_out/r02z_sort_phases_8gpu.log
python - <<'P'
import json
d=json.load(open("gpurun_out/r02z_bench_8gpu.json"))
print(d["value"], d["e2e"]["value"], d["parity"]["all_ok"], d["parity"].get("failures"))
for k,v in d["primitives"].items(): print(k, {a:b for a,b in v.items() if a in ("GB/s","Gkeys/s","us_per_call","ms_per_call","exchange")})
P
