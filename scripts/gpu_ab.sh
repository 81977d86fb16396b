#!/bin/bash
# parity tests + stage timings of the network kernel on two scenes
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
for cfg in "C2 800" "C3 1000"; do echo "== $cfg"; timeout 300 python scripts/profile_stage.py $cfg 2>&1 | tail -5; done
