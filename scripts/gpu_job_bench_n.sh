# gpurun job: contract bench and reference arm under torchrun on N GPUs (usage: gpurun --gpus N -- bash scripts/gpu_job_bench_n.sh N)
N=$1
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N > gpurun_out/r2_bench_n${N}.json 2> gpurun_out/r2_bench_n${N}.err ) 2> gpurun_out/r2_bench_n${N}.time
echo rc=$? >> gpurun_out/r2_bench_n${N}.err
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29545 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/r2_bench_ref_n${N}.json 2> gpurun_out/r2_bench_ref_n${N}.err ) 2> gpurun_out/r2_bench_ref_n${N}.time
