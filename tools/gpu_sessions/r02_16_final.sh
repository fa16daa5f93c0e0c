#!/bin/bash
# GPU session r02-16 (1 GPU): what the driver runs at round end, on the final library -- the whole GPU suite, smoke(),
# the default bench line and the reference arm -- plus the launch list (time + DRAM bytes per launch) of the C5 headline.
set -x
mkdir -p gpurun_out
( timeout 1800 python -m pytest tests -m gpu -q -x ) > gpurun_out/r02p_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02p_pytest_gpu.log
tail -4 gpurun_out/r02p_pytest_gpu.log
( timeout 600 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/r02p_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r02p_smoke.log
tail -6 gpurun_out/r02p_smoke.log
timeout 900 python bench.py > gpurun_out/r02p_bench_default.log 2>&1; echo "rc=$?" >> gpurun_out/r02p_bench_default.log
grep '^{' gpurun_out/r02p_bench_default.log | cut -c1-600
timeout 900 python bench.py --impl reference > gpurun_out/r02p_bench_reference.log 2>&1; echo "rc=$?" >> gpurun_out/r02p_bench_reference.log
grep '^{' gpurun_out/r02p_bench_reference.log | cut -c1-600
timeout 300 python bench.py --workload c5 --no-others --no-cpu --steps 2 --warmup 3 > gpurun_out/r02p_plain_c5.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 260 --csv --log-file gpurun_out/r02p_launches_c5.csv python bench.py --workload c5 --no-others --no-cpu --steps 2 --warmup 3 > gpurun_out/r02p_ncu_c5.log 2>&1
tail -2 gpurun_out/r02p_ncu_c5.log
