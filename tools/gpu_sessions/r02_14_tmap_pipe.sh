#!/bin/bash
# GPU session r02-14 (1 GPU): tensor-map pass A version 4 knobs -- per-warp ring hand-back through EMPTY mbarriers
# (NLS_TM_PIPE=1, 6 plane slots, no CTA barrier per plane) and L2 prefetch of the column values (NLS_TM_PREFETCH=planes):
# parity under each, then A/B on C3 / C5.
set -x
mkdir -p gpurun_out
for pipe in 0 1; do
  ( NLS_TM_PIPE=$pipe NLS_TM_PREFETCH=$((2*pipe)) timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "tensor_map" -rA ) > gpurun_out/r02n_pytest_tmap_pipe$pipe.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02n_pytest_tmap_pipe$pipe.log
  grep -E "passed|failed|rc=|Error" gpurun_out/r02n_pytest_tmap_pipe$pipe.log | tail -4
done
line() { python - "$1" "$2" <<'PY'
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        d = json.loads(ln); print(sys.argv[2], "ms/step %.4f" % d["ms_per_step"], "value %.4g" % d["value"], "frac %.3f" % d["roofline"]["frac"])
PY
}
run() { # run <tag> <workload> [ENV=VAL ...]
  local tag=$1 wl=$2; shift 2
  env "$@" timeout 600 python bench.py --workload $wl --no-others --no-cpu --steps 10 --warmup 3 > gpurun_out/r02n_${tag}_$wl.log 2>&1
  line gpurun_out/r02n_${tag}_$wl.log "$tag $wl"
}
for wl in c3 c5; do
  run default $wl NLS_X=0
  run pipe $wl NLS_TM_PIPE=1
  run pipe_pf1 $wl NLS_TM_PIPE=1 NLS_TM_PREFETCH=1
  run pipe_pf2 $wl NLS_TM_PIPE=1 NLS_TM_PREFETCH=2
  run pf2 $wl NLS_TM_PREFETCH=2
  run pipe_body6 $wl NLS_TM_PIPE=1 NLS_TM_BODY=6
done
NLS_TM_PIPE=1 timeout 300 python bench.py --workload c3 --no-others --no-cpu --steps 1 --warmup 3 > gpurun_out/r02n_plain_c3.log 2>&1 && \
NLS_TM_PIPE=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_lanczos_a_tm -s 24 -c 4 -o gpurun_out/r02n_c3_lanczos_a_tm_v4pipe python bench.py --workload c3 --no-others --no-cpu --steps 1 --warmup 3 > gpurun_out/r02n_ncu_tm.log 2>&1
ls -la gpurun_out/r02n*.ncu-rep
