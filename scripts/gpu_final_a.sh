#!/bin/bash
# Round-end evidence, part A: reference arm, default bench, other configs, ncu launch list of the bench command.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
echo "== reference arm"; ( time timeout 900 python bench.py --impl reference --steps 4 --warmup 1 ) > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo rc=$?; cut -c1-300 gpurun_out/bench_ref.json; tail -4 gpurun_out/bench_ref.err
echo "== bench default"; ( time timeout 900 python bench.py ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo rc=$?; cut -c1-200 gpurun_out/bench_default.json; tail -4 gpurun_out/bench_default.err
for w in C2 C4 C5; do
  echo "== bench $w"; timeout 900 python bench.py --workload $w --steps 8 --warmup 3 --no-cpu > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo rc=$?; cut -c1-200 gpurun_out/bench_$w.json; tail -2 gpurun_out/bench_$w.err
done
echo "== ncu launches"; python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 40 --csv --log-file gpurun_out/launches_final.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_l.log 2>&1; echo rc=$?
