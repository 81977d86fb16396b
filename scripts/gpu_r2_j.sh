#!/bin/bash
# round 2, call J: final evidence with the specialised kernel: whole suite, smoke, default bench + reference arm as the driver
# runs them, C2 line, ncu launch list of the two timed device frames of the bench command, ncu --set full of one heavy C3 frame.
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -6) > gpurun_out/r2j_pytest.log
(timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2) > gpurun_out/r2j_smoke.log
(timeout 900 python bench.py --impl reference --gpus 1 --steps 16 --warmup 4 2>&1 | tail -1) > gpurun_out/r2j_bench_reference.json
(timeout 600 python bench.py 2>&1 | tail -1) > gpurun_out/r2j_bench_c3.json
(timeout 600 python bench.py --workload C2 --no-cpu 2>&1 | tail -1) > gpurun_out/r2j_bench_c2.json
K='k_rays_count|k_round_sums|k_scan_blocks|k_emit|k_network|k_composite'
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2j_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"$K" -s 55 -c 10 --csv --log-file gpurun_out/r2j_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2j_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_network|k_rays_count|k_emit|k_composite' -s 40 -c 4 -o gpurun_out/r2j_prof -f python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2j_ncu_f.log 2>&1
# (compute-sanitizer is closed on this pool: the exercise in scripts/sanitize_small.py is run plain by the tests instead)
tail -3 gpurun_out/r2j_pytest.log; cat gpurun_out/r2j_smoke.log; for f in gpurun_out/r2j_bench_*.json; do echo $f; cut -c1-260 $f; done; tail -2 gpurun_out/r2j_ncu_f.log
