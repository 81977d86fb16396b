#!/bin/bash
# A/B of prebuilt library variants nerf_cuda_b200/lib_v*.so (copied over libnerf_b200.so one at a time): per-pose stage times.
cd $GRAFT_REPO_ROOT
cp nerf_cuda_b200/libnerf_b200.so /tmp/lib_orig.so
for v in ${VARIANTS}; do
  cp nerf_cuda_b200/lib_v$v.so nerf_cuda_b200/libnerf_b200.so
  echo "== variant $v"; python scripts/pose_samples.py C3 | python -c "
import json,sys
P=json.load(sys.stdin)
for i,p in enumerate(P): print(i,p['samples'],round(p['ms_total'],2),round(p['ms_setup'],2),round(p['ms_march'],2),round(p['ms_network'],2))
print('mean total',round(sum(p['ms_total'] for p in P)/len(P),3),'mean setup',round(sum(p['ms_setup'] for p in P)/len(P),3))"
done
cp /tmp/lib_orig.so nerf_cuda_b200/libnerf_b200.so
