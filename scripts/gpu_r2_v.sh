#!/bin/bash
# round 2, call V: t-cache size A/B (32 / 64 / 128 cached sample t's per ray) on C3.
mkdir -p gpurun_out
L=$PWD/nerf_cuda_b200
(timeout 300 python bench.py --no-cpu 2>&1 | tail -1) > gpurun_out/r2v_tc64.json
(NRF_LIB=$L/libnerf_b200_tcache128.so timeout 300 python bench.py --no-cpu 2>&1 | tail -1) > gpurun_out/r2v_tc128.json
(NRF_LIB=$L/libnerf_b200_tcache32.so timeout 300 python bench.py --no-cpu 2>&1 | tail -1) > gpurun_out/r2v_tc32.json
(NRF_LIB=$L/libnerf_b200_tcache128.so timeout 600 python -m pytest tests -m gpu -q -k "t_cache or full_size_c3 or termination or frame_parity" 2>&1 | tail -2) > gpurun_out/r2v_pytest_tc128.log
cat gpurun_out/r2v_pytest_tc128.log
python - <<'PY'
import json
for f in ('tc32','tc64','tc128'):
    d=json.loads(open(f'gpurun_out/r2v_{f}.json').read().strip().splitlines()[-1])
    print(f, round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['per_rank'][0]['stage_ms'].items()})
PY
