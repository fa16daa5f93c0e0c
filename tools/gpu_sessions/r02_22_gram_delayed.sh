#!/bin/bash
# GPU session r02-22 (1 GPU): NLS_REORTH_GRAM (Gram-delayed re-orthogonalisation, j + 6 passes per iteration): parity
# tests, then A/B against the default two-pass form on C5 / C3 / C4 (NLS_REORTH=3 switches the bench's handles).
set -x
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "gram_delayed" -rA -s ) > gpurun_out/r02v_pytest_gram.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02v_pytest_gram.log
grep -E "gram-delayed|passed|failed|rc=|Error" gpurun_out/r02v_pytest_gram.log | tail -24
line() { python - "$1" "$2" <<'PY'
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        d = json.loads(ln); print(sys.argv[2], "ms/step %.4f" % d["ms_per_step"], "value %.4g" % d["value"], "frac %.3f" % d["roofline"]["frac"], "B/site-step %.0f" % d["roofline"]["bytes_per_site_step"])
PY
}
for wl in c5 c3 c4; do
  for ro in 0 3; do
    NLS_REORTH=$ro timeout 200 python bench.py --workload $wl --no-others --no-cpu --steps 10 --warmup 3 > gpurun_out/r02v_${wl}_reorth$ro.log 2>&1
    line gpurun_out/r02v_${wl}_reorth$ro.log "$wl NLS_REORTH=$ro"
  done
done
