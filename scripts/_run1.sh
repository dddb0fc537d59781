set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r2_pytest1.log
V=variants_tmp
run() { XETGPU_LIBRARY=$1 XETGPU_TWO_LANES=$2 python scripts/variant_bench.py $3 2>&1 | tail -1 >> gpurun_out/r2_variants1.jsonl; }
rm -f gpurun_out/r2_variants1.jsonl
python scripts/variant_bench.py default >> gpurun_out/r2_variants1.jsonl 2>&1
run ome_b200/libxetgpu.so 1 default
run $V/roll0/libxetgpu.so 0 roll0
run $V/roll2/libxetgpu.so 0 roll2
run $V/co6/libxetgpu.so 1 co6
run $V/co6/libxetgpu.so 0 co6
run $V/co8/libxetgpu.so 1 co8
run $V/co4/libxetgpu.so 1 co4
run $V/co6r0/libxetgpu.so 1 co6r0
cat gpurun_out/r2_variants1.jsonl
