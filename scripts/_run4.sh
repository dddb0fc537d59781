( time python bench.py > gpurun_out/r2_bench_n1_b.json 2> gpurun_out/r2_bench_n1_b.err ) 2> gpurun_out/r2_bench_n1_b.time
echo rc=$? >> gpurun_out/r2_bench_n1_b.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1
