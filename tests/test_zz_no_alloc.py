"""nrf_eval_samples_noalloc (include/nerf_b200_ext.h; no counterpart in the reference):
sample evaluation with its scratch in a fixed file-scope device array, for callers that cannot cudaMalloc.
Checked through the C ABI against the existing entry point nrf_probe_samples on the same inputs, two calls in a row that
share the scratch (the second smaller, so the first call's leftovers sit behind it), plus the size limit.
Collected last on purpose (file name)."""
import numpy as np
import pytest

from nerf_cuda_b200 import NerfRender, synthetic as syn

NOALLOC_MAX = 65536   # NRF_NOALLOC_MAX_SAMPLES


def rel_err(a, b):   # the network tolerance of tests/test_gpu_parity.py: relative, against max(|ref|, 0.02)
    return np.abs(a - b) / np.maximum(np.abs(b), 2e-2)


@pytest.mark.gpu
def test_no_alloc_entry_point_matches_probe_samples(built_lib):
    import torch
    sc = syn.make_scene("tiny")
    r = NerfRender(1)
    r.load_raw(sc.desc, sc.params, sc.grid, sc.aabb)
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(2024)
    for call, n in enumerate((5000, 777)):     # ragged sizes; the second call reuses the scratch the first one filled
        pos = rng.uniform(0, 1, size=(n, 3)).astype(np.float32)
        d = rng.normal(size=(n, 3))
        dirs01 = (d / np.linalg.norm(d, axis=1, keepdims=True) * 0.5 + 0.5).astype(np.float32)
        ref = r.probe_samples(pos, dirs01)     # existing entry point (allocates its staging buffers)
        p4 = torch.zeros(n, 4, dtype=torch.float32, device=dev)
        d4 = torch.zeros(n, 4, dtype=torch.float32, device=dev)
        p4[:, :3] = torch.from_numpy(pos).to(dev)
        d4[:, :3] = torch.from_numpy(dirs01).to(dev)
        out = torch.full((n, 4), float("nan"), dtype=torch.float16, device=dev)
        torch.cuda.synchronize()               # the inputs were written on torch's stream, the library has its own
        free0 = torch.cuda.mem_get_info(0)[0]
        r.eval_samples_noalloc(n, p4.data_ptr(), d4.data_ptr(), out.data_ptr())
        r.synchronize()
        if call:   # (the first call may make the driver load the two kernels' code lazily; the second finds everything in place)
            assert torch.cuda.mem_get_info(0)[0] >= free0      # nothing was allocated on the device by the call
        got = out.float().cpu().numpy()
        assert np.isfinite(got).all()
        assert rel_err(got, ref).max() <= 1e-2, (n, rel_err(got, ref).max())
        assert np.array_equal(got, ref)        # same kernels' arithmetic on the same fp16 features: in fact identical
    # the limit: one sample more than the fixed scratch holds is refused with a clear error, the maximum itself runs
    big = torch.zeros(NOALLOC_MAX + 1, 4, dtype=torch.float32, device=dev)
    bout = torch.zeros(NOALLOC_MAX + 1, 4, dtype=torch.float16, device=dev)
    torch.cuda.synchronize()
    with pytest.raises(Exception, match="exceed the fixed scratch"):
        r.eval_samples_noalloc(NOALLOC_MAX + 1, big.data_ptr(), big.data_ptr(), bout.data_ptr())
    big[:, :3] = 0.5
    torch.cuda.synchronize()
    r.eval_samples_noalloc(NOALLOC_MAX, big.data_ptr(), big.data_ptr(), bout.data_ptr())
    r.synchronize()
    o = bout[:NOALLOC_MAX].float().cpu().numpy()
    ref1 = r.probe_samples(np.full((1, 3), 0.5, np.float32), np.full((1, 3), 0.5, np.float32))
    assert np.array_equal(o, np.repeat(ref1, NOALLOC_MAX, axis=0))
    del p4, d4, out, big, bout
    r.close()
