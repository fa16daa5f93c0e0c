#!/bin/bash
# GPU session 39 (1 GPU): ncu --set full captures of the streaming (HBM-bound) Lanczos passes on C3 (256^3 complex, div(c grad), m=12):
# one k_lanczos_a launch (stencil + dots, 10-column chunk) and one k_lanczos_b launch (orthogonalisation + norm)
set -x
mkdir -p gpurun_out
timeout 300 python bench.py --workload c3 --steps 2 --warmup 1 --no-cpu > gpurun_out/r39_plain_c3.log 2>&1 && {
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_lanczos_a -s 40 -c 1 -f -o gpurun_out/r39_c3_lanczos_a_full python bench.py --workload c3 --steps 2 --warmup 1 --no-cpu > gpurun_out/r39_ncu_a.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_lanczos_b -s 40 -c 1 -f -o gpurun_out/r39_c3_lanczos_b_full python bench.py --workload c3 --steps 2 --warmup 1 --no-cpu > gpurun_out/r39_ncu_b.log 2>&1
}
ls -la gpurun_out | grep r39
