#!/bin/bash
# GPU session r02-11 (1 GPU): tensor-map pass A version 2 (64x8 tile, two sites per thread, column loads issued before
# the mbarrier wait): (1) parity, (2) A/B against the flat pass on C3 / C5, (3) launch list + ncu --set full of 3 launches,
# (4) the whole GPU suite on this library.
set -x
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "tensor_map" -rA ) > gpurun_out/r02k_pytest_tmap.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02k_pytest_tmap.log
grep -E "passed|failed|rc=|Error" gpurun_out/r02k_pytest_tmap.log | tail -6
line() { python - "$1" "$2" <<'PY'
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        d = json.loads(ln); print(sys.argv[2], "ms/step %.4f" % d["ms_per_step"], "value %.4g" % d["value"], "frac %.3f" % d["roofline"]["frac"])
PY
}
for wl in c3 c5; do
  timeout 600 python bench.py --workload $wl --no-others --no-cpu --steps 10 --warmup 3 > gpurun_out/r02k_flat_$wl.log 2>&1
  line gpurun_out/r02k_flat_$wl.log "flat $wl"
  NLS_PASS_A_TMAP=1 timeout 600 python bench.py --workload $wl --no-others --no-cpu --steps 10 --warmup 3 > gpurun_out/r02k_tmap_$wl.log 2>&1
  line gpurun_out/r02k_tmap_$wl.log "tmap-v2 $wl"
done
NLS_PASS_A_TMAP=1 timeout 300 python bench.py --workload c3 --no-others --no-cpu --steps 1 --warmup 3 > gpurun_out/r02k_plain_c3_tmap.log 2>&1 && \
NLS_PASS_A_TMAP=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_lanczos_a_tm -s 24 -c 3 -o gpurun_out/r02k_c3_lanczos_a_tm_v2 python bench.py --workload c3 --no-others --no-cpu --steps 1 --warmup 3 > gpurun_out/r02k_ncu_tm.log 2>&1
ls -la gpurun_out/r02k*.ncu-rep
( timeout 1500 python -m pytest tests -m gpu -q -x ) > gpurun_out/r02k_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02k_pytest_gpu.log
tail -5 gpurun_out/r02k_pytest_gpu.log
