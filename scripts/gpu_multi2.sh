#!/bin/bash
# 2-GPU box: single-process peer-frame test, torchrun bench at N=1 and N=2 (NCCL gather over NVLink).
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader
nvidia-smi topo -m | head -6
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "two_gpus or sharded" 2>&1 | tail -5
timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err; echo rc=$?; cut -c1-400 gpurun_out/scale_n1.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 8 --warmup 3 > gpurun_out/scale_n2.json 2> gpurun_out/scale_n2.err; echo rc=$?; cut -c1-1200 gpurun_out/scale_n2.json; tail -5 gpurun_out/scale_n2.err
