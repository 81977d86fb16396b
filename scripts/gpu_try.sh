cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for n in 3 4 6; do echo "== CTAS $n"; NRF_NET_CTAS_PER_SM=$n python scripts/profile_stage.py C2 400 2>&1 | tail -4; done
