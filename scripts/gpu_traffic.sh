#!/bin/bash
# Full-size capture of ONE k_network<2> launch of the default bench (C3 1080p): dram traffic + L1 wavefronts per launch.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python scripts/pose_samples.py C3 > gpurun_out/pose_samples_c3.json 2>gpurun_out/pose_samples.err; cat gpurun_out/pose_samples_c3.json
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/traffic_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_network -s 11 -c 1 -o gpurun_out/prof_c3_full -f \
    python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/traffic_ncu.log 2>&1
echo rc=$?; tail -2 gpurun_out/traffic_ncu.log
