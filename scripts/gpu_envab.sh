#!/bin/bash
# A/B over values of one environment variable: per-pose stage means.  usage: VAR=NAME VALUES="a b c" bash scripts/gpu_envab.sh
cd $GRAFT_REPO_ROOT
for v in ${VALUES}; do
  export $VAR=$v
  echo "== $VAR=$v"; python scripts/pose_samples.py C3 | python -c "
import json,sys
P=json.load(sys.stdin)
m=lambda k: round(sum(p[k] for p in P)/len(P),3)
print('mean total',m('ms_total'),'setup',m('ms_setup'),'march',m('ms_march'),'network',m('ms_network'))"
done
