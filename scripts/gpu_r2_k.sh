#!/bin/bash
# round 2, call K: C3 with its own specialised k_network instantiation: parity at full size + the small exercise, A/B against
# the generic kernel, the default bench line, launch list of the two timed device frames, ncu --set full of one heavy frame.
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -q -k "full_size or exercise or termination" 2>&1 | tail -4) > gpurun_out/r2k_pytest.log
for wl in C3 C4; do
  (timeout 400 python bench.py --no-cpu --workload $wl --steps 16 2>&1 | tail -1) > gpurun_out/r2k_spec_$wl.json
  (NRF_NO_SPECIALISE=1 timeout 400 python bench.py --no-cpu --workload $wl --steps 16 2>&1 | tail -1) > gpurun_out/r2k_generic_$wl.json
done
(timeout 600 python bench.py 2>&1 | tail -1) > gpurun_out/r2k_bench_c3.json
(timeout 600 python bench.py --scene opaque --no-cpu 2>&1 | tail -1) > gpurun_out/r2k_bench_c3_opaque.json
K='k_rays_count|k_round_sums|k_scan_blocks|k_emit|k_network|k_composite'
python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2k_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"$K" -s 55 -c 10 --csv --log-file gpurun_out/r2k_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2k_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_network|k_rays_count|k_emit|k_composite' -s 40 -c 4 -o gpurun_out/r2k_prof -f python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r2k_ncu_f.log 2>&1
cat gpurun_out/r2k_pytest.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2k_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r2k_')[1], round(d['ms_per_step'],3), 'net', round(d['roofline']['ms_per_launch'],3), 'e2e', round(d['e2e']['ms_per_step'],3))
    except Exception as e: print(f, 'ERR', e, open(f).read()[-300:])
PY
grep -c k_network gpurun_out/r2k_launches.csv; grep k_network gpurun_out/r2k_launches.csv | head -2 | cut -c1-120
