#!/bin/bash
# GPU session r02-13 (1 GPU): (1) the whole GPU suite on the library with prefetch lead 1 + tensor-map default (the
# allocation-free entry point's file runs last), then that file alone in a fresh process; (2) knob A/B: prefetch of the
# stencil-carrying flat pass vs the other passes, split closing pass B; (3) the default bench line; (4) C3 launch list.
set -x
mkdir -p gpurun_out
( timeout 1800 python -m pytest tests -m gpu -q -x ) > gpurun_out/r02m_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02m_pytest_gpu.log
tail -5 gpurun_out/r02m_pytest_gpu.log
( timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -rA -s ) > gpurun_out/r02m_pytest_noalloc_alone.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02m_pytest_noalloc_alone.log
grep -E "passed|failed|rc=|Error|worst" gpurun_out/r02m_pytest_noalloc_alone.log | tail -6
line() { python - "$1" "$2" <<'PY'
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        d = json.loads(ln); print(sys.argv[2], "ms/step %.4f" % d["ms_per_step"], "value %.4g" % d["value"], "frac %.3f" % d["roofline"]["frac"])
PY
}
run() { # run <tag> <workload> [ENV=VAL ...]
  local tag=$1 wl=$2; shift 2
  env "$@" timeout 600 python bench.py --workload $wl --no-others --no-cpu --steps 10 --warmup 3 > gpurun_out/r02m_${tag}_$wl.log 2>&1
  line gpurun_out/r02m_${tag}_$wl.log "$tag $wl"
}
for wl in c3 c5 c4; do
  run default $wl NLS_X=0
  run pf0 $wl NLS_PREFETCH=0
  run pffirst0 $wl NLS_PREFETCH_FIRST=0
  run bsplit $wl NLS_B_SPLIT=1
done
run flat_default c5 NLS_PASS_A_TMAP=0
run flat_pffirst0 c5 NLS_PASS_A_TMAP=0 NLS_PREFETCH_FIRST=0
run flat_pf0 c5 NLS_PASS_A_TMAP=0 NLS_PREFETCH=0
timeout 900 python bench.py > gpurun_out/r02m_bench_default.log 2>&1; echo "rc=$?" >> gpurun_out/r02m_bench_default.log
grep '^{' gpurun_out/r02m_bench_default.log | cut -c1-400
timeout 300 python bench.py --workload c3 --no-others --no-cpu --steps 2 --warmup 3 > gpurun_out/r02m_plain_c3.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02m_launches_c3.csv python bench.py --workload c3 --no-others --no-cpu --steps 2 --warmup 3 > gpurun_out/r02m_ncu_c3.log 2>&1
tail -2 gpurun_out/r02m_ncu_c3.log
