#!/bin/bash
# GPU session r02-12 (1 GPU): (1) the allocation-free mass reduction alone in a fresh process; (2) tensor-map pass A
# version 3 (interior / boundary split, compile-time column roles; now the default for complex 3D grids): parity;
# (3) A/B benches: flat vs tensor-map, body-chunk width, wide pass B, L2 prefetch of the flat passes;
# (4) the whole GPU suite (the no-alloc file runs last); (5) launch list + ncu --set full of the tensor-map pass on C3.
set -x
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -rA -s ) > gpurun_out/r02l_pytest_noalloc_alone.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02l_pytest_noalloc_alone.log
grep -E "passed|failed|rc=|Error|worst" gpurun_out/r02l_pytest_noalloc_alone.log | tail -6
( timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "tensor_map" -rA ) > gpurun_out/r02l_pytest_tmap.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02l_pytest_tmap.log
grep -E "passed|failed|rc=|Error" gpurun_out/r02l_pytest_tmap.log | tail -6
line() { python - "$1" "$2" <<'PY'
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        d = json.loads(ln); print(sys.argv[2], "ms/step %.4f" % d["ms_per_step"], "value %.4g" % d["value"], "frac %.3f" % d["roofline"]["frac"])
PY
}
run() { # run <tag> <workload> [ENV=VAL ...]
  local tag=$1 wl=$2; shift 2
  env "$@" timeout 600 python bench.py --workload $wl --no-others --no-cpu --steps 10 --warmup 3 > gpurun_out/r02l_${tag}_$wl.log 2>&1
  line gpurun_out/r02l_${tag}_$wl.log "$tag $wl"
}
for wl in c3 c5; do
  run flat $wl NLS_PASS_A_TMAP=0
  run tmap3 $wl NLS_X=0
  run tmap3_body6 $wl NLS_TM_BODY=6
  run tmap3_body2 $wl NLS_TM_BODY=2
  run tmap3_bwide $wl NLS_B_WIDE=1
  run tmap3_pf1 $wl NLS_PREFETCH=1
  run flat_pf1 $wl NLS_PASS_A_TMAP=0 NLS_PREFETCH=1
  run flat_pf2 $wl NLS_PASS_A_TMAP=0 NLS_PREFETCH=2
done
run base c4 NLS_X=0
run pf1 c4 NLS_PREFETCH=1
run pf2 c4 NLS_PREFETCH=2
( timeout 1500 python -m pytest tests -m gpu -q -x ) > gpurun_out/r02l_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02l_pytest_gpu.log
tail -5 gpurun_out/r02l_pytest_gpu.log
timeout 300 python bench.py --workload c3 --no-others --no-cpu --steps 1 --warmup 3 > gpurun_out/r02l_plain_c3.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_lanczos_a_tm -s 24 -c 4 -o gpurun_out/r02l_c3_lanczos_a_tm_v3 python bench.py --workload c3 --no-others --no-cpu --steps 1 --warmup 3 > gpurun_out/r02l_ncu_tm.log 2>&1
ls -la gpurun_out/r02l*.ncu-rep
