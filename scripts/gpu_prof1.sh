cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python scripts/profile_stage.py C2 400 > gpurun_out/prof_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_network -s 2 -c 9 -o gpurun_out/prof_net_r1b -f python scripts/profile_stage.py C2 400 > gpurun_out/prof_ncu.log 2>&1
echo rc=$?; cat gpurun_out/prof_plain.log; tail -3 gpurun_out/prof_ncu.log
