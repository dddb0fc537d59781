timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2_pytest13.log
XETGPU_STAGED=0 timeout 300 python scripts/variant_bench.py unstaged > gpurun_out/r2_variants4.jsonl 2>&1
timeout 300 python scripts/variant_bench.py staged >> gpurun_out/r2_variants4.jsonl 2>&1
timeout 300 python scripts/bench_configs.py c1 c2 c3slice >> gpurun_out/r2_variants4.jsonl 2>&1
