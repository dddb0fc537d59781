timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2_pytest8.log
timeout 300 python scripts/bench_configs.py c1 c2 > gpurun_out/r2_variants2.jsonl 2>&1
timeout 300 python scripts/profile_workload.py 2 3 > gpurun_out/r2_plain_prof.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_launches_a.csv python scripts/profile_workload.py 2 3 > gpurun_out/r2_ncu_a.log 2>&1
