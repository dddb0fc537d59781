set -x
python -m pytest tests/test_gpu_sharded.py -x -q 2>&1 | tail -8 > gpurun_out/r2_pytest2.log
python bench.py --scale 0.05 --steps 3 --warmup 3 > gpurun_out/r2_bench_small.json 2> gpurun_out/r2_bench_small.err
echo rc=$? >> gpurun_out/r2_bench_small.err
