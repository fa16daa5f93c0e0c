#!/bin/bash
# GPU session r02-19 (8 GPUs): the C5 line at N = 8 and N = 4 on the final library (finish prefetch off, 10-step e2e
# segments) -- the driver's launch form.  Short: box time is charged 8x.
set -x
mkdir -p gpurun_out
for n in 8 4; do
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n)) bench.py --gpus $n --no-cpu --no-others > gpurun_out/r02s_bench_c5_n$n.log 2>&1; echo "rc=$?" >> gpurun_out/r02s_bench_c5_n$n.log
  grep '^{' gpurun_out/r02s_bench_c5_n$n.log | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('N', d['n_gpus'], 'ms/step %.3f'%d['ms_per_step'], 'value %.4g'%d['value'], 'e2e %.4g'%d['e2e']['value'], 'ratio %.3f'%(d['e2e']['value']/d['value']), 'seg', d['e2e']['segment_steps'], d['clocks']['reasons'])"
done
