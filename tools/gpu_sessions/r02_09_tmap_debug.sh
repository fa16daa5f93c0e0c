#!/bin/bash
# GPU session r02-09 (1 GPU): tensor-map pass A, second look: parity of the uniform variant (div(c grad) gated off),
# ncu --set full of one tensor-map launch and one flat launch of the same iteration on the C5 slab, and -- last,
# because a fault kills its process -- the stand-alone probe of the box shapes (tools/microbench/tmap_probe.cu).
set -x
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "tensor_map" -rA ) > gpurun_out/r02i_pytest_tmap.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02i_pytest_tmap.log
grep -E "passed|failed|rc=|Error" gpurun_out/r02i_pytest_tmap.log | tail -8
export NLS_PASS_A_TMAP=1
timeout 300 python bench.py --workload c5 --no-others --no-cpu --steps 1 --warmup 3 > gpurun_out/r02i_plain_c5_tmap.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_lanczos_a_tm -s 30 -c 2 -o gpurun_out/r02i_c5_lanczos_a_tm python bench.py --workload c5 --no-others --no-cpu --steps 1 --warmup 3 > gpurun_out/r02i_ncu_tm.log 2>&1
unset NLS_PASS_A_TMAP
timeout 300 python bench.py --workload c5 --no-others --no-cpu --steps 1 --warmup 3 > gpurun_out/r02i_plain_c5_flat.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_lanczos_a<" -s 30 -c 2 -o gpurun_out/r02i_c5_lanczos_a_flat python bench.py --workload c5 --no-others --no-cpu --steps 1 --warmup 3 > gpurun_out/r02i_ncu_flat.log 2>&1
ls -la gpurun_out/*.ncu-rep
for c in 0 1 2 3; do timeout 60 ./tools/microbench/tmap_probe_bench $c >> gpurun_out/r02i_tmap_probe.log 2>&1; echo "case $c rc=$?" >> gpurun_out/r02i_tmap_probe.log; done
cat gpurun_out/r02i_tmap_probe.log
