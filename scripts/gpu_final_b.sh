#!/bin/bash
# Round-end evidence, part B: one ncu --set full capture of every frame kernel of one C3 1080p frame (bench command).
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain_b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_network|k_rays_count|k_emit|k_composite' -s 44 -c 4 -o gpurun_out/prof_final -f python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_f.log 2>&1; echo rc=$?; tail -2 gpurun_out/ncu_f.log
