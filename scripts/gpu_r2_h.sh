#!/bin/bash
# round 2, call H (8 GPUs): C3 strong scaling at N = 1, 2, 4, 8 with the one-sided peer fence (the driver's launch), the NCCL
# gather and the in-process mode at N = 4, 8, C4 tile- vs pose-parallel at N = 8, the reference arm under torchrun, 2-GPU tests.
mkdir -p gpurun_out
nvidia-smi -L | wc -l > gpurun_out/r2h_ngpus.txt
TR() { n=$1; shift; timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n "$@"; }
(timeout 300 python bench.py --no-cpu 2>&1 | tail -1) > gpurun_out/r2h_scale_n1.json
for n in 2 4 8; do (TR $n bench.py --gpus $n 2>&1 | tail -1) > gpurun_out/r2h_scale_n$n.json; done
(TR 8 bench.py --gpus 8 --gather nccl 2>&1 | tail -1) > gpurun_out/r2h_scale_n8_nccl.json
(timeout 300 python bench.py --gpus 8 --mode inproc 2>&1 | tail -1) > gpurun_out/r2h_scale_n8_inproc.json
(timeout 300 python bench.py --gpus 4 --mode inproc 2>&1 | tail -1) > gpurun_out/r2h_scale_n4_inproc.json
(TR 8 bench.py --gpus 8 --workload C4 --steps 32 2>&1 | tail -1) > gpurun_out/r2h_c4_n8_tiles.json
(TR 8 bench.py --gpus 8 --workload C4 --shard poses --steps 8 2>&1 | tail -1) > gpurun_out/r2h_c4_n8_poses.json
(TR 8 bench.py --gpus 8 --impl reference --steps 8 --warmup 2 2>&1 | tail -1) > gpurun_out/r2h_ref_n8.json
(timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "two_gpus or peer or sharded" 2>&1 | tail -5) > gpurun_out/r2h_pytest_multi.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2h_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r2h_')[1], d.get('n_gpus'), round(d['ms_per_step'],3), round(d['value']/1e6,1), 'e2e', round(d['e2e']['value']/1e6,1), d['e2e'].get('batch_call_value'))
    except Exception as e: print(f, 'ERR', e, open(f).read()[-300:])
PY
cat gpurun_out/r2h_pytest_multi.log
