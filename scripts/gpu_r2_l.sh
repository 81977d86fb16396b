#!/bin/bash
# round 2, call L (2 GPUs): the multi-GPU tests (incl. the two-device batch pipeline), N=2 lines with the final kernels.
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q -k "two_gpus or peer or sharded or snapshot_written" 2>&1 | tail -6) > gpurun_out/r2l_pytest_multi.log
TR="timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
($TR bench.py --gpus 2 2>&1 | tail -1) > gpurun_out/r2l_scale_n2.json
(timeout 400 python bench.py --gpus 2 --mode inproc 2>&1 | tail -1) > gpurun_out/r2l_scale_n2_inproc.json
cat gpurun_out/r2l_pytest_multi.log; cut -c1-200 gpurun_out/r2l_scale_n2.json; cut -c1-200 gpurun_out/r2l_scale_n2_inproc.json
