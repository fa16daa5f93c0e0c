#!/bin/bash
# GPU session r02-10 (1 GPU): (1) tensor-map pass A with 32-byte-granular box rows: parity; (2) column-stride padding
# A/B (NLS_COL_PAD) on C3 / C5 / C4; (3) tensor-map vs flat pass A on C3 / C5; (4) ncu --set full of one launch of each;
# (5) the box-shape probe with 66- and 68-double rows (last: the 66 case faults its own process).
set -x
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "tensor_map" -rA ) > gpurun_out/r02j_pytest_tmap.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02j_pytest_tmap.log
grep -E "passed|failed|rc=|Error" gpurun_out/r02j_pytest_tmap.log | tail -6
line() { python - "$1" "$2" <<'PY'
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        d = json.loads(ln); print(sys.argv[2], "ms/step %.4f" % d["ms_per_step"], "value %.4g" % d["value"], "frac %.3f" % d["roofline"]["frac"])
PY
}
for wl in c3 c5 c4; do
  for pad in 0 8448 74240 1122816; do
    NLS_COL_PAD=$pad timeout 600 python bench.py --workload $wl --no-others --no-cpu --steps 10 --warmup 3 > gpurun_out/r02j_pad_${wl}_$pad.log 2>&1
    line gpurun_out/r02j_pad_${wl}_$pad.log "colpad $wl pad=$pad"
  done
done
for wl in c3 c5; do
  NLS_PASS_A_TMAP=1 timeout 600 python bench.py --workload $wl --no-others --no-cpu --steps 10 --warmup 3 > gpurun_out/r02j_tmap_$wl.log 2>&1
  line gpurun_out/r02j_tmap_$wl.log "tmap $wl"
done
NLS_PASS_A_TMAP=1 timeout 300 python bench.py --workload c3 --no-others --no-cpu --steps 1 --warmup 3 > gpurun_out/r02j_plain_c3_tmap.log 2>&1 && \
NLS_PASS_A_TMAP=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_lanczos_a_tm -s 24 -c 3 -o gpurun_out/r02j_c3_lanczos_a_tm python bench.py --workload c3 --no-others --no-cpu --steps 1 --warmup 3 > gpurun_out/r02j_ncu_tm.log 2>&1
timeout 300 python bench.py --workload c3 --no-others --no-cpu --steps 1 --warmup 3 > gpurun_out/r02j_plain_c3_flat.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'^k_lanczos_a$' -s 36 -c 3 -o gpurun_out/r02j_c3_lanczos_a_flat python bench.py --workload c3 --no-others --no-cpu --steps 1 --warmup 3 > gpurun_out/r02j_ncu_flat.log 2>&1
ls -la gpurun_out/r02j*.ncu-rep
for r in 68 66; do timeout 60 ./tools/microbench/tmap_probe_bench 1 $r >> gpurun_out/r02j_tmap_probe.log 2>&1; echo "real row $r rc=$?" >> gpurun_out/r02j_tmap_probe.log; done
timeout 60 ./tools/microbench/tmap_probe_bench 3 68 >> gpurun_out/r02j_tmap_probe.log 2>&1; echo "both loads, row 68 rc=$?" >> gpurun_out/r02j_tmap_probe.log
cat gpurun_out/r02j_tmap_probe.log
