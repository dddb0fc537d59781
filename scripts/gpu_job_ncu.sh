# gpurun job: plain run, ncu launch list, plain run, ncu --set full of the hot kernels (B200_PROFILING.md recipe), single-GPU configs
timeout 300 python scripts/profile_workload.py 2 3 > gpurun_out/r2_plain_prof.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_launches.csv python scripts/profile_workload.py 2 3 > gpurun_out/r2_ncu_l.log 2>&1
timeout 300 python scripts/profile_workload.py 2 3 > gpurun_out/r2_plain_prof2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"gear_scan_kernel|leaf_group_kernel|merkle_coop_kernel|chunk_tree_kernel" -s 4 -c 4 -o gpurun_out/prof_r2 -f python scripts/profile_workload.py 2 3 > gpurun_out/r2_ncu_f.log 2>&1
timeout 300 python scripts/bench_configs.py c1 c2 c5 > gpurun_out/r2_configs_b.jsonl 2>&1
