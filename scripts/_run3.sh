timeout 600 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_multi.py tests/test_gpu_parity.py -x -q 2>&1 | tail -5 > gpurun_out/r2_pytest12.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --scale 0.02 --steps 3 --warmup 3 --only c4 --no-e2e > gpurun_out/r2_bench_n2_small.json 2> gpurun_out/r2_bench_n2_small.err
echo rc=$? >> gpurun_out/r2_bench_n2_small.err
