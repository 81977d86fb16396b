#!/bin/bash
# round 2, call G: whole GPU suite, smoke, bench lines of every config, launch list + ncu --set full of one heavy C3 frame.
mkdir -p gpurun_out
(timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -15) > gpurun_out/r2g_pytest.log
(timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3) > gpurun_out/r2g_smoke.log
(timeout 600 python bench.py 2>&1 | tail -1) > gpurun_out/r2g_bench_c3.json
(timeout 600 python bench.py --impl reference --steps 16 --warmup 4 2>&1 | tail -1) > gpurun_out/r2g_bench_reference.json
for w in C2 C4 C5; do (timeout 600 python bench.py --workload $w --steps 8 --warmup 3 --no-cpu 2>&1 | tail -1) > gpurun_out/r2g_bench_$w.json; done
(timeout 600 python bench.py --scene opaque --no-cpu 2>&1 | tail -1) > gpurun_out/r2g_bench_c3_opaque.json
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2g_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 40 -c 80 --csv --log-file gpurun_out/r2g_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2g_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_network|k_rays_count|k_emit|k_composite' -s 40 -c 4 -o gpurun_out/r2g_prof -f python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2g_ncu_f.log 2>&1
tail -4 gpurun_out/r2g_pytest.log; cat gpurun_out/r2g_smoke.log; for f in gpurun_out/r2g_bench_*.json; do echo $f; cut -c1-250 $f; done; tail -2 gpurun_out/r2g_ncu_f.log; ls -la gpurun_out/r2g_prof.ncu-rep
