nvidia-smi topo -m > gpurun_out/r02s_topo_8gpu.log 2>&1; lscpu | grep -i "numa\|socket\|model name\|^CPU(s)" >> gpurun_out/r02s_topo_8gpu.log
timeout 300 python -m pytest tests/test_gpu_sharded_capi.py -x -q -m gpu 2>&1 | tail -3
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r02s_bench_8gpu.json 2> gpurun_out/r02s_bench_8gpu.err; echo bench8 rc=$?
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 tests/multi_gpu_check.py > gpurun_out/r02s_multi_gpu_check_8gpu.log 2>&1; echo check8 rc=$?; tail -4 gpurun_out/r02s_multi_gpu_check_8gpu.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 profiles/sort_sharded_phases.py > gpurun_out/r02s_sort_phases_8gpu.log 2>&1; echo phases rc=$?; tail -12 gpurun_out/r02s_sort_phases_8gpu.log
CUDA_VISIBLE_DEVICES=0,1,2,3 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29524 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/r02s_bench_4gpu.json 2> gpurun_out/r02s_bench_4gpu.err; echo bench4 rc=$?
tail -c 300 gpurun_out/r02s_bench_8gpu.err
