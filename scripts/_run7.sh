for s in 16 32 64; do for t in 16 32; do echo "slots=$s threads=$t"; XETGPU_READ_SLOTS=$s XETGPU_READ_THREADS=$t timeout 300 python scripts/bench_files.py 5 5 5 1 2>&1 | grep xetgpu | tail -2; done; done > gpurun_out/r2_files_sweep.log 2>&1
timeout 300 python scripts/bench_files.py 5 5 5 1 >> gpurun_out/r2_files_sweep.log 2>&1
