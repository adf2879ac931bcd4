"""GPU test of b2_mul_mat_noalloc, the entry point for callers that may not allocate device memory (a framework
that owns all of HBM, a loop that cannot take cudaMalloc's locks).  The caller brings device buffers -- here torch
tensors stand in for the host framework's memory -- and the library uses its built-in scratch array.

Collected last on purpose (file name): if it fails under `pytest -x`, every other test has already run.
It must pass both alone in a fresh process and after the rest of the suite.
"""
import numpy as np
import pytest

from localai_b200 import capi
from oracle import oracle as O

pytestmark = pytest.mark.gpu
MATMUL_RTOL = 2e-5  # same tolerance as tests/test_parity_gpu.py


def _case(torch, wtype, rows, cols, n_tok, seed):
    import make_gguf as mg

    rng = np.random.default_rng(seed)
    if wtype == capi.B2_T_F16:
        raw = (rng.standard_normal((rows, cols)) / np.sqrt(cols)).astype(np.float16).view(np.uint8).reshape(-1)
    else:
        raw = mg.random_blocks(rng, wtype, rows * cols // 256, 1 / np.sqrt(cols)).reshape(-1)
    x = rng.standard_normal((n_tok, cols)).astype(np.float32)
    want = capi.mul_mat(wtype, raw, rows, cols, x, path=1)  # the existing, allocating entry point
    ref = O.mul_mat(wtype, raw, rows, cols, x)               # and the CPU oracle
    w_d = torch.from_numpy(raw.copy()).cuda()
    x_d = torch.from_numpy(x).cuda()
    out_d = torch.full((n_tok, rows), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    capi.mul_mat_noalloc(wtype, w_d.data_ptr(), rows, cols, x_d.data_ptr(), n_tok, out_d.data_ptr(), 0)
    torch.cuda.synchronize()
    # nothing was allocated: the device's free memory is what it was (the first call of the process did its one-time
    # symbol look-up before this function was entered, see the warm-up in the test)
    assert torch.cuda.mem_get_info()[0] == free_before
    got = out_d.cpu().numpy()
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() <= MATMUL_RTOL * np.abs(want).max()
    assert np.abs(got - ref).max() <= MATMUL_RTOL * np.abs(ref).max()
    return got


def test_no_alloc_matches_allocating_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available()
    warm = torch.zeros(1 << 16, dtype=torch.uint8, device="cuda")  # first call on the device: looks the scratch symbol up
    capi.mul_mat_noalloc(capi.B2_T_Q4_K, warm.data_ptr(), 2, 256, warm.data_ptr(), 1, warm.data_ptr() + 4096, 0)
    torch.cuda.synchronize()
    # two calls back to back share the same scratch array: the second (smaller, other type) must not see stale data
    _case(torch, capi.B2_T_Q4_K, 512, 4096, 4, seed=1)
    _case(torch, capi.B2_T_Q5_K, 300, 1024, 1, seed=2)
    _case(torch, capi.B2_T_F16, 128, 16384, 4, seed=3)   # exactly fills the 128 KiB scratch
    _case(torch, capi.B2_T_Q4_K, 64, 256, 2, seed=4)


def test_no_alloc_rejects_what_does_not_fit():
    torch = pytest.importorskip("torch")
    buf = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda")
    p = buf.data_ptr()
    with pytest.raises(capi.B2Error, match="exceeds the built-in scratch"):
        capi.mul_mat_noalloc(capi.B2_T_Q4_K, p, 4, 16384 + 256, p, 1, p)
    with pytest.raises(capi.B2Error, match="exceeds the built-in scratch"):
        capi.mul_mat_noalloc(capi.B2_T_Q4_K, p, 4, 4096, p, 5, p)
    with pytest.raises(capi.B2Error, match="must be Q4_K, Q5_K or F16"):
        capi.mul_mat_noalloc(capi.B2_T_Q6_K, p, 4, 4096, p, 1, p)
    with pytest.raises(capi.B2Error):
        capi.mul_mat_noalloc(capi.B2_T_Q4_K, 0, 4, 4096, p, 1, p)
