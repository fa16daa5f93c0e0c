#!/bin/bash
# GPU session r02-21 (1 GPU): stencil-carrying flat pass A with the grid dimension compiled in (NLS_A_SPEC=1) and with the
# NOTE: the two variants measured here lost (profiles/r02u_flat_pass_a_occupancy_ab.txt) and their knobs (NLS_A_SPEC, NLS_A_OCC) were removed again.
# 3-CTAs/SM register allocation for its narrow chunks (NLS_A_OCC=3): parity subset under the knob, then C4 A/B.
set -x
mkdir -p gpurun_out
( NLS_A_OCC=3 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "stencil or lanczos_basis or nlse_2d_steps or batched or three_term or nlse_3d_cubic or real_2d or kge_3d or tensor_map_pass_matches or persistent_kernel" ) > gpurun_out/r02u_pytest_a_occ3.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02u_pytest_a_occ3.log
tail -3 gpurun_out/r02u_pytest_a_occ3.log
line() { python - "$1" "$2" <<'PY'
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        d = json.loads(ln); print(sys.argv[2], "ms/step %.4f" % d["ms_per_step"], "value %.4g" % d["value"], "frac %.3f" % d["roofline"]["frac"])
PY
}
for cfg in "NLS_X=0" "NLS_A_SPEC=1" "NLS_A_OCC=3"; do
  env $cfg timeout 300 python bench.py --workload c4 --no-others --no-cpu --steps 10 --warmup 3 > gpurun_out/r02u_c4_$cfg.log 2>&1
  line gpurun_out/r02u_c4_$cfg.log "c4 $cfg"
done
