#!/bin/bash
# round 2, call I: the level-layout-specialised k_network against the generic one (NRF_NO_SPECIALISE=1): parity at full size, A/B.
(timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -q -k "full_size" 2>&1 | tail -4) > gpurun_out/r2i_pytest.log
for wl in C3 C2; do
  (timeout 300 python bench.py --no-cpu --workload $wl 2>&1 | tail -1) > gpurun_out/r2i_spec_$wl.json
  (NRF_NO_SPECIALISE=1 timeout 300 python bench.py --no-cpu --workload $wl 2>&1 | tail -1) > gpurun_out/r2i_generic_$wl.json
  (timeout 300 python bench.py --no-cpu --workload $wl 2>&1 | tail -1) > gpurun_out/r2i_spec2_$wl.json
done
cat gpurun_out/r2i_pytest.log
