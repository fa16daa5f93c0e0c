#!/bin/bash
# GPU session r02-20 (1 GPU): the round-end sequence on the final library (whole GPU suite, smoke(), default bench line,
# reference arm), then a C4 sweep of the first-chunk width and the C4 launch list (time + DRAM bytes per launch).
set -x
mkdir -p gpurun_out
( timeout 1800 python -m pytest tests -m gpu -q -x ) > gpurun_out/r02t_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02t_pytest_gpu.log
tail -4 gpurun_out/r02t_pytest_gpu.log
( timeout 600 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/r02t_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r02t_smoke.log
tail -4 gpurun_out/r02t_smoke.log
timeout 900 python bench.py > gpurun_out/r02t_bench_default.log 2>&1; echo "rc=$?" >> gpurun_out/r02t_bench_default.log
grep '^{' gpurun_out/r02t_bench_default.log | cut -c1-400
timeout 900 python bench.py --impl reference > gpurun_out/r02t_bench_reference.log 2>&1; echo "rc=$?" >> gpurun_out/r02t_bench_reference.log
grep '^{' gpurun_out/r02t_bench_reference.log | cut -c1-300
line() { python - "$1" "$2" <<'PY'
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        d = json.loads(ln); print(sys.argv[2], "ms/step %.4f" % d["ms_per_step"], "value %.4g" % d["value"], "frac %.3f" % d["roofline"]["frac"])
PY
}
for cap in 0 6 8 12; do
  NLS_FIRST_CAP=$cap timeout 300 python bench.py --workload c4 --no-others --no-cpu --steps 10 --warmup 3 > gpurun_out/r02t_c4_firstcap$cap.log 2>&1
  line gpurun_out/r02t_c4_firstcap$cap.log "c4 first_cap=$cap"
done
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 220 --csv --log-file gpurun_out/r02t_launches_c4.csv python bench.py --workload c4 --no-others --no-cpu --steps 2 --warmup 3 > gpurun_out/r02t_ncu_c4.log 2>&1
tail -2 gpurun_out/r02t_ncu_c4.log
