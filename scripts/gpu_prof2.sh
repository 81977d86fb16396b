#!/bin/bash
# ncu --set full of every frame kernel + the three k_network modes, on a C2 (400x400) and a C3 (960x540) frame.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for spec in "C2 400 400" "C3 960 540"; do
  set -- $spec
  python scripts/profile_stage.py $1 $2 $3 > gpurun_out/prof_plain_$1.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:'k_network|k_rays_count|k_emit|k_composite' -s 4 -c 10 \
      -o gpurun_out/prof_$1 -f python scripts/profile_stage.py $1 $2 $3 > gpurun_out/prof_ncu_$1.log 2>&1
  echo "$1 rc=$?"; cat gpurun_out/prof_plain_$1.log; tail -2 gpurun_out/prof_ncu_$1.log
done
