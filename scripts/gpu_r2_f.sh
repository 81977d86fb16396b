#!/bin/bash
# round 2, call F: peer-fence tests, then A/B of the k_network experiment builds on C3 (and C2): product vs each variant.
(timeout 600 python -m pytest "tests/test_gpu_parity.py::test_peer_frame_ipc_two_processes" "tests/test_gpu_parity.py::test_peer_frame_one_sided_fence_pipeline" tests/test_gpu_termination.py -q 2>&1 | tail -30) > gpurun_out/r2f_pytest.log
L=nerf_cuda_b200
run() { # name, env...
  name=$1; shift
  for wl in C3 C2; do
    (env "$@" timeout 300 python bench.py --no-cpu --workload $wl --steps 16 --warmup 4 2>&1 | tail -1) > gpurun_out/r2f_ab_${name}_${wl}.json
  done
}
run base A=1
run base_again A=1
run floor_magic NRF_LIB=$PWD/$L/libnerf_b200_floor_magic.so
run evict NRF_LIB=$PWD/$L/libnerf_b200_evict.so
run floor_magic_evict NRF_LIB=$PWD/$L/libnerf_b200_floor_magic_evict.so
run smem_l01 NRF_LIB=$PWD/$L/libnerf_b200_smem_l01.so
run persist32 NRF_L2_PERSIST=32
run persist64 NRF_L2_PERSIST=64
# the variants must stay bit-exact: reference-CUDA golden vectors through each variant
for v in floor_magic evict smem_l01; do
  (NRF_LIB=$PWD/$L/libnerf_b200_$v.so timeout 300 python -m pytest tests/test_ref_cuda_golden.py -m gpu -q 2>&1 | tail -3) > gpurun_out/r2f_golden_$v.log
done
tail -8 gpurun_out/r2f_pytest.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2f_ab_*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('r2f_ab_')[1], round(d['ms_per_step'],3), 'net', round(d['roofline']['ms_per_launch'],3), 'e2e', round(d['e2e']['ms_per_step'],3))
    except Exception as e: print(f, 'ERR', e)
PY
tail -2 gpurun_out/r2f_golden_*.log
