#!/bin/bash
# round 2, call Z: k_network group-synchronisation experiments (no tail barrier; mbarrier arrive instead of bar.sync) vs base.
mkdir -p gpurun_out
L=$PWD/nerf_cuda_b200
run() { name=$1; shift; for wl in C3 C2; do (env "$@" timeout 120 python bench.py --no-cpu --workload $wl 2>&1 | tail -1) > gpurun_out/r2z_${name}_$wl.json; done; }
par() { name=$1; shift; (env "$@" timeout 200 python -m pytest tests -m gpu -q -x -k "network or samples or golden or full_size_c3 or frame_parity or no_alloc" 2>&1 | tail -2) > gpurun_out/r2z_pytest_$name.log; cat gpurun_out/r2z_pytest_$name.log; }
run base A=1
par no_tail_sync NRF_LIB=$L/libnerf_b200_no_tail_sync.so
run no_tail_sync NRF_LIB=$L/libnerf_b200_no_tail_sync.so
par arrive_sync NRF_LIB=$L/libnerf_b200_arrive_sync.so
run arrive_sync NRF_LIB=$L/libnerf_b200_arrive_sync.so
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2z_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r2z_')[1], round(d['ms_per_step'],3), 'net', round(d['roofline']['ms_per_launch'],3), 'mlp', round(d['stages_alone']['mlp_ms'],3))
    except Exception as e: print(f, 'ERR', e, open(f).read()[-200:])
PY
