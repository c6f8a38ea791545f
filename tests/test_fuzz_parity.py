"""Bounded randomised parity sweep inside `pytest -m gpu`: tools/fuzz_parity.py with a fixed seed -- random picture sizes,
presets MR..M13, hierarchy depths, temporal layers, clip kinds, reference lists and modes (open-loop ME, temporal filter,
dynamic-GOP detector, picture statistics); every output of the CUDA path, including the pixel-compare count, against the
oracle.  Bit-exact or fail."""
import importlib.util
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [1234, 20261105])
def test_fuzz_parity_bounded(seed):
    spec = importlib.util.spec_from_file_location("fuzz_parity", os.path.join(ROOT, "tools", "fuzz_parity.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    bad = mod.run(iters=100, seed=seed, verbose=False)
    assert not bad, "%d of 100 random cases differ from the oracle; first: %s" % (len(bad), bad[0])
