#!/bin/bash
# GPU session r02-17 (1 GPU): prefetch in the recombination kernel on / off (NLS_PREFETCH_FINISH), C3 / C4 / C5.
set -x
mkdir -p gpurun_out
line() { python - "$1" "$2" <<'PY'
import json, sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        d = json.loads(ln); print(sys.argv[2], "ms/step %.4f" % d["ms_per_step"], "value %.4g" % d["value"], "frac %.3f" % d["roofline"]["frac"])
PY
}
for rep in 1 2; do
for wl in c3 c4 c5; do
  for pf in 1 0; do
    NLS_PREFETCH_FINISH=$pf timeout 600 python bench.py --workload $wl --no-others --no-cpu --steps 10 --warmup 3 > gpurun_out/r02q_finpf${pf}_${wl}_$rep.log 2>&1
    line gpurun_out/r02q_finpf${pf}_${wl}_$rep.log "finish_pf=$pf $wl rep$rep"
  done
done
done
# DRAM traffic of the recombination kernel without its prefetch (with: profiles/r02m_launches_c3.csv, 4.60 GB read)
NLS_PREFETCH_FINISH=0 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:k_nlse_finish -s 4 -c 3 --csv --log-file gpurun_out/r02q_finish_nopf_c3.csv python bench.py --workload c3 --no-others --no-cpu --steps 2 --warmup 3 > gpurun_out/r02q_ncu_finish.log 2>&1
grep -v "^==" gpurun_out/r02q_finish_nopf_c3.csv | cut -d, -f5,13- | tail -9
