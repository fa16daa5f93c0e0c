#!/bin/bash
# GPU session r02-07 (1 GPU): tensor-map TMA pass A (nls_tmap.cuh): parity, A/B timing on C3 and C5, ncu launch lists
# (duration + DRAM bytes per launch) of one C3 step with and without it.
set -x
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_fieldgen.py tests/test_downsample.py tests/test_launcher_shim.py tests/test_drivers.py -m gpu -q -x -k "tensor_map or mgs or fieldgen or device_ or fft or batched_launch or downsampled" -rA ) > gpurun_out/r02g_pytest_new.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02g_pytest_new.log
grep -E "passed|failed|rc=|Error|MGS|max abs err" gpurun_out/r02g_pytest_new.log | tail -30
for wl in c3 c5; do
  for v in flat tmap; do
    if [ $v = tmap ]; then export NLS_PASS_A_TMAP=1; else unset NLS_PASS_A_TMAP; fi
    timeout 600 python bench.py --workload $wl --no-others --no-cpu --steps 10 --warmup 3 > gpurun_out/r02g_bench_${wl}_$v.log 2>&1
    python - <<PY
import json
for ln in open("gpurun_out/r02g_bench_${wl}_$v.log"):
    if ln.startswith("{"):
        d = json.loads(ln); print("$wl $v ms/step", d["ms_per_step"], "value", d["value"], "frac", d["roofline"]["frac"])
PY
  done
done
export NLS_PASS_A_TMAP=1
timeout 300 python bench.py --workload c3 --no-others --no-cpu --steps 2 --warmup 3 > gpurun_out/r02g_plain_c3_tmap.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"k_lanczos|k_nlse|k_tridiag" -c 400 --csv --log-file gpurun_out/r02g_launches_c3_tmap.csv python bench.py --workload c3 --no-others --no-cpu --steps 2 --warmup 3 > gpurun_out/r02g_ncu_c3_tmap.log 2>&1
unset NLS_PASS_A_TMAP
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"k_lanczos|k_nlse|k_tridiag" -c 400 --csv --log-file gpurun_out/r02g_launches_c3_flat.csv python bench.py --workload c3 --no-others --no-cpu --steps 2 --warmup 3 > gpurun_out/r02g_ncu_c3_flat.log 2>&1
ls -la gpurun_out/r02g_launches_c3_*.csv
