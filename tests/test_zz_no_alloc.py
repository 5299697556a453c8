"""f2f_tvl1_gaussian_noalloc: the allocation-free variant of f2f_tvl1_gaussian (scratch in a static device
array, include/f2f_b200.h).  It has no counterpart in the reference; the check is the existing entry point,
bit for bit (the bar tests/test_tvl1_gpu.py holds the Gaussian to), on two calls in a row that share the
static scratch -- the second one smaller, so that the array still holds the first call's values around it.

Collected last on purpose.
"""
import numpy as np
import pytest
import torch

from frame2frame_b200 import _lib, synth

pytestmark = pytest.mark.gpu
MAX_PIXELS = 1920 * 1080          # F2F_NOALLOC_MAX_PIXELS, include/f2f_b200.h


def _pair(lib, cuda, H, W, t, sigma):
    I = torch.from_numpy(np.ascontiguousarray(synth.noisy_frame(H, W, t))).to(cuda)
    want, tmp = I.clone(), torch.empty_like(I)
    _lib.check(lib.f2f_tvl1_gaussian(want.data_ptr(), W, H, sigma, tmp.data_ptr(), None), "gaussian")
    got = I.clone()
    torch.cuda.synchronize(cuda)
    free0 = torch.cuda.mem_get_info(cuda)[0]
    _lib.check(lib.f2f_tvl1_gaussian_noalloc(got.data_ptr(), W, H, sigma, None), "gaussian_noalloc")
    torch.cuda.synchronize(cuda)
    assert torch.cuda.mem_get_info(cuda)[0] == free0, "the call changed the device's free memory"
    return got, want, I


def test_noalloc_gaussian_matches_existing_entry_point_twice(cuda):
    lib = _lib.lib()
    with torch.cuda.device(cuda):
        for (H, W, t, sigma) in [(1080, 1920, 3, 0.8), (135, 240, 5, 0.8), (47, 33, 2, 1.4)]:
            got, want, I = _pair(lib, cuda, H, W, t, sigma)
            assert torch.equal(got, want), (H, W)
            assert not torch.equal(got, I)


def test_noalloc_gaussian_rejects_sizes_beyond_the_static_scratch(cuda):
    lib = _lib.lib()
    with torch.cuda.device(cuda):
        x = torch.zeros(16, device=cuda)
        rc = lib.f2f_tvl1_gaussian_noalloc(x.data_ptr(), 1921, 1080, 0.8, None)
        assert rc == -1 and b"F2F_NOALLOC_MAX_PIXELS" in lib.f2f_last_error()
        assert lib.f2f_tvl1_gaussian_noalloc(x.data_ptr(), 0, 4, 0.8, None) == -1
        assert lib.f2f_tvl1_gaussian_noalloc(None, 4, 4, 0.8, None) == -1
        assert 1921 * 1080 > MAX_PIXELS
