( time python bench.py > gpurun_out/r2_bench_n1_a.json 2> gpurun_out/r2_bench_n1_a.err ) 2> gpurun_out/r2_bench_n1_a.time
echo rc=$? >> gpurun_out/r2_bench_n1_a.err
( time python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2_bench_ref_a.json 2> gpurun_out/r2_bench_ref_a.err ) 2> gpurun_out/r2_bench_ref_a.time
