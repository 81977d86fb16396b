#!/bin/bash
# Round-end evidence: reference arm, default bench, ncu launch list of the same command, one full-set capture per kernel.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== reference arm"; ( time timeout 900 python bench.py --impl reference --steps 4 --warmup 1 ) > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo rc=$?; cut -c1-500 gpurun_out/bench_ref.json; tail -4 gpurun_out/bench_ref.err
echo "== bench default"; ( time timeout 900 python bench.py ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo rc=$?; cut -c1-200 gpurun_out/bench_default.json; tail -4 gpurun_out/bench_default.err
echo "== ncu launches"; python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 40 --csv --log-file gpurun_out/launches_final.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_l.log 2>&1; echo rc=$?
echo "== ncu full"; ncu --set full --clock-control none --import-source on -k regex:'k_network|k_rays_count|k_emit|k_composite' -s 44 -c 4 -o gpurun_out/prof_final -f python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_f.log 2>&1; echo rc=$?; tail -2 gpurun_out/ncu_f.log
