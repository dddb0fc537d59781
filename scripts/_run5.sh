timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2_pytest10.log
timeout 300 python scripts/variant_bench.py lanes8 > gpurun_out/r2_variants3.jsonl 2>&1
XETGPU_TWO_LANES=1 timeout 300 python scripts/variant_bench.py lanes8 >> gpurun_out/r2_variants3.jsonl 2>&1
timeout 300 python scripts/bench_configs.py c5 >> gpurun_out/r2_variants3.jsonl 2>&1
