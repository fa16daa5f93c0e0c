#!/bin/bash
# GPU session 16: finer phase trace (site loop vs publish)
set -x
mkdir -p gpurun_out
NLS_B200_LIB=$PWD/nonlinear-solvers_b200/variants/libnls_trace.so timeout 300 python tools/pk_trace.py > gpurun_out/r16_pk_trace.txt 2>&1
wc -l gpurun_out/r16_pk_trace.txt
