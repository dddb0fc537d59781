# gpurun job: tests that need several GPUs + the contract bench on 8 GPUs (usage: gpurun --gpus 8 -- bash scripts/gpu_job_n8.sh)
python -m pytest tests/test_gpu_multi.py tests/test_gpu_sharded.py -x -q 2>&1 | tail -6 > gpurun_out/r2_pytest_n8.log
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 > gpurun_out/r2_bench_n8_a.json 2> gpurun_out/r2_bench_n8_a.err ) 2> gpurun_out/r2_bench_n8_a.time
echo rc=$? >> gpurun_out/r2_bench_n8_a.err
free -g | head -2 > gpurun_out/r2_box_n8.txt; nproc >> gpurun_out/r2_box_n8.txt; df -h /dev/shm >> gpurun_out/r2_box_n8.txt
