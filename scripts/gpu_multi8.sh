#!/bin/bash
# N-GPU box: torchrun bench at N = 1, 2, 4, 8 (whatever the box has), NCCL gather over NVLink.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l); echo "gpus: $NG"
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "two_gpus" 2>&1 | tail -2
for n in ${NLIST:-1 2 4 8}; do
  [ $n -gt $NG ] && break
  if [ $n -eq 1 ]; then timeout 600 python bench.py --steps 16 --warmup 4 --no-cpu > gpurun_out/scale8_n$n.json 2> gpurun_out/scale8_n$n.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520+n)) bench.py --gpus $n --steps 16 --warmup 4 --gather ${GATHER:-peer} > gpurun_out/scale8_n$n.json 2> gpurun_out/scale8_n$n.err; fi
  echo "n=$n rc=$?"; python -c "
import json,sys
d=json.loads(open('gpurun_out/scale8_n$n.json').read().strip().splitlines()[-1])
print('n',d['n_gpus'],'ms/frame',round(d['ms_per_step'],3),'Mrays/s',round(d['value']/1e6,1),'e2e',round(d['e2e']['value']/1e6,1))" 2>&1 | tail -1
  tail -2 gpurun_out/scale8_n$n.err | cut -c1-200
done
