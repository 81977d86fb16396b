#!/bin/bash
# round 2, call N: A/B of the march-step experiment (F2I.TRUNC -> FADD.RZ magic) on C3 and C2; parity of the variant.
mkdir -p gpurun_out
L=$PWD/nerf_cuda_b200
for wl in C3 C2; do
  (timeout 300 python bench.py --no-cpu --workload $wl 2>&1 | tail -1) > gpurun_out/r2n_base_$wl.json
  (NRF_LIB=$L/libnerf_b200_march_magic.so timeout 300 python bench.py --no-cpu --workload $wl 2>&1 | tail -1) > gpurun_out/r2n_magic_$wl.json
  (timeout 300 python bench.py --no-cpu --workload $wl 2>&1 | tail -1) > gpurun_out/r2n_base2_$wl.json
done
(NRF_LIB=$L/libnerf_b200_march_magic.so timeout 600 python -m pytest tests/test_ref_cuda_golden.py tests/test_gpu_parity.py -m gpu -q -k "bit_exact or march or full_size or stages" 2>&1 | tail -3) > gpurun_out/r2n_pytest_magic.log
cat gpurun_out/r2n_pytest_magic.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2n_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r2n_')[1], round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['per_rank'][0]['stage_ms'].items()})
    except Exception as e: print(f, 'ERR', e, open(f).read()[-300:])
PY
