#!/bin/bash
# ncu --set full of the network kernel (all three modes) and the march kernels on a reduced C2 / C3 frame.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CFG=${1:-C2}; W=${2:-400}
python scripts/profile_stage.py $CFG $W > gpurun_out/prof_plain_$CFG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_network|k_rays_count|k_emit|k_composite' -s 5 -c 8 -o gpurun_out/prof_${CFG} -f python scripts/profile_stage.py $CFG $W > gpurun_out/prof_ncu_$CFG.log 2>&1
echo rc=$?; cat gpurun_out/prof_plain_$CFG.log; tail -3 gpurun_out/prof_ncu_$CFG.log
