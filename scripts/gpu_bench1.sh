set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python bench.py --steps 8 --warmup 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo rc=$?; cat gpurun_out/bench_c3.json; tail -5 gpurun_out/bench_c3.err
python bench.py --steps 8 --warmup 3 --workload C2 --no-cpu > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo rc=$?; cat gpurun_out/bench_c2.json; tail -5 gpurun_out/bench_c2.err
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_r1.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu.log 2>&1
echo ncu rc=$?
tail -3 gpurun_out/ncu.log
