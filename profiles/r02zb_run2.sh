timeout 300 python -m pytest tests/test_gpu_sharded_capi.py tests/test_gpu_multi.py -x -q -m gpu 2>&1 | tail -3
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29522 tests/multi_gpu_check.py > gpurun_out/r02zb_multi_gpu_check_2gpu.log 2>&1; echo check2 rc=$?; tail -2 gpurun_out/r02zb_multi_gpu_check_2gpu.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02zb_bench_2gpu.json 2> gpurun_out/r02zb_bench_2gpu.err; echo bench2 rc=$?
python - <<'P'
import json
d=json.load(open("gpurun_out/r02zb_bench_2gpu.json"))
print(d["value"], d["e2e"]["value"], d["parity"]["all_ok"], d["parity"].get("failures"))
for k,v in d["primitives"].items(): print(k, {a:b for a,b in v.items() if a in ("GB/s","Gkeys/s","us_per_call","ms_per_call")})
P
