#!/bin/bash
# GPU session r02-15 (1 GPU): persistent kernel with fewer, larger CTAs on small grids (NLS_PK_MIN_CSZ = minimum sites
# per CTA): parity of the persistent-kernel tests under two settings, then C1 / C2 (Gautschi) timings over the knob.
set -x
mkdir -p gpurun_out
for csz in 1024 4096; do
  ( NLS_PK_MIN_CSZ=$csz timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "persistent_kernel or config1_full or config2_sge or nlse_2d_steps or real_2d_steps or odd_grid" ) > gpurun_out/r02o_pytest_pk_csz$csz.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02o_pytest_pk_csz$csz.log
  tail -3 gpurun_out/r02o_pytest_pk_csz$csz.log
done
line() { python - "$1" "$2" <<'PY'
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        d = json.loads(ln); print(sys.argv[2], "ms/step %.5f" % d["ms_per_step"], "value %.4g" % d["value"], "frac %.3f" % d["roofline"]["frac"])
PY
}
for wl in c1 c2_gautschi; do
  for csz in 64 886 1024 1772 2048 3072 4096 8192; do
    NLS_PK_MIN_CSZ=$csz timeout 300 python bench.py --workload $wl --no-others --no-cpu --steps 200 --warmup 20 > gpurun_out/r02o_${wl}_csz$csz.log 2>&1
    line gpurun_out/r02o_${wl}_csz$csz.log "$wl min_csz=$csz"
  done
done
