#!/bin/bash
# A/B of prebuilt library variants x one environment variable: VARIANTS="base pair" VAR=NRF_NET_GROUPS VALUES="7 6"
cd $GRAFT_REPO_ROOT
cp nerf_cuda_b200/libnerf_b200.so /tmp/lib_orig.so
for v in ${VARIANTS}; do
  cp nerf_cuda_b200/lib_v$v.so nerf_cuda_b200/libnerf_b200.so
  for x in ${VALUES}; do
    export $VAR=$x
    echo "== variant $v $VAR=$x"; python scripts/pose_samples.py C3 | python -c "
import json,sys
P=json.load(sys.stdin)
m=lambda k: round(sum(p[k] for p in P)/len(P),3)
print('mean total',m('ms_total'),'setup',m('ms_setup'),'march',m('ms_march'),'network',m('ms_network'), 'typical net', round(P[1]['ms_network'],3), 'heavy net', round(P[2]['ms_network'],3))"
  done
done
cp /tmp/lib_orig.so nerf_cuda_b200/libnerf_b200.so
