#!/bin/bash
# round 2, call D (2 GPUs): whole GPU suite incl. the 2-GPU tests, then C3 at N=1 and N=2 (ranks/peer, ranks/nccl, inproc).
nvidia-smi -L > gpurun_out/r2d_gpus.txt
(timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -60) > gpurun_out/r2d_pytest.log
(timeout 400 python bench.py --no-cpu 2>&1 | tail -3) > gpurun_out/r2d_bench_n1.json
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
(timeout 400 $TR bench.py --gpus 2 2>&1 | tail -3) > gpurun_out/r2d_bench_n2_peer.json
(timeout 400 $TR bench.py --gpus 2 --gather nccl 2>&1 | tail -3) > gpurun_out/r2d_bench_n2_nccl.json
(timeout 400 python bench.py --gpus 2 --mode inproc 2>&1 | tail -3) > gpurun_out/r2d_bench_n2_inproc.json
(timeout 400 $TR bench.py --gpus 2 --workload C4 --shard poses 2>&1 | tail -3) > gpurun_out/r2d_bench_c4_n2_poses.json
tail -15 gpurun_out/r2d_pytest.log; for f in gpurun_out/r2d_bench_*.json; do echo $f; tail -1 $f | cut -c1-300; done
