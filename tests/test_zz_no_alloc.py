"""The allocation-free entry point (svtme_cuda_nxm_sad_kernel_noalloc: operand staging and result word in a file-scope
device array, no cudaMalloc) against the existing svtme_cuda_nxm_sad_kernel, through the C ABI.  Collected last on
purpose: a failure here must not hide the parity tests."""
import numpy as np
import pytest

from svtme import api

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    d = api.MeDevice(device=0)
    yield d
    d.close()


def _blocks(seed, w, h, stride):
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, (h + 2, stride), dtype=np.uint8), rng.integers(0, 256, (h + 2, stride), dtype=np.uint8)


def test_noalloc_matches_the_allocating_entry_on_consecutive_calls(dev):
    # two calls in a row that share the static scratch: the second block is smaller than the first (and then narrower
    # than a word), so operands or a result left over from the first call would show up in it
    calls = [(64, 128, 1), (48, 17, 2), (6, 3, 3), (64, 32, 4), (8, 8, 5), (64, 32, 4)]
    for (w, h, seed) in calls:
        stride = w + 19
        src, ref = _blocks(seed, w, h, stride)
        want = dev.nxm_sad_kernel(src, 3, stride, ref, 1, stride, h, w)
        got = dev.nxm_sad_kernel_noalloc(src, 3, stride, ref, 1, stride, h, w)
        assert got == want, (w, h, got, want)  # integer SAD: exact
        assert want == int(np.abs(src.reshape(-1)[3:][:h * stride].reshape(h, stride)[:, :w].astype(np.int64) -
                                  ref.reshape(-1)[1:][:h * stride].reshape(h, stride)[:, :w].astype(np.int64)).sum())
    # identical inputs twice in a row: the cleared result word, not an accumulated one
    src, ref = _blocks(9, 64, 64, 80)
    a = dev.nxm_sad_kernel_noalloc(src, 0, 80, ref, 0, 80, 64, 64)
    b = dev.nxm_sad_kernel_noalloc(src, 0, 80, ref, 0, 80, 64, 64)
    assert a == b == dev.nxm_sad_kernel(src, 0, 80, ref, 0, 80, 64, 64)


def test_noalloc_rejects_sizes_beyond_the_static_scratch(dev):
    src, ref = _blocks(1, 128, 129, 160)
    for (w, h) in [(65, 8), (64, 129), (0, 8), (8, 0)]:
        with pytest.raises(api.SvtMeError) as e:
            dev.nxm_sad_kernel_noalloc(src, 0, 160, ref, 0, 160, h, w)
        if w and h:
            assert "static scratch" in str(e.value)
    # and the entry still works afterwards
    assert dev.nxm_sad_kernel_noalloc(src, 0, 160, ref, 0, 160, 16, 16) == dev.nxm_sad_kernel(src, 0, 160, ref, 0, 160, 16, 16)
