"""Dynamic-GOP detector HME level 0: dg_detector_hme_level0 (pd_process.c:493-590), the
TASK_DG_DETECTOR_HME branch of the ME kernel thread (me_process.c:327-329).

CPU: the oracle restatement against the compiled reference (C and AVX2 kernels) and against the committed golden
metrics (tests/golden/dg_golden.json, written by the compiled reference).
GPU: svtme_dg_detector_hme_level0 through the C ABI against the oracle (metrics, per-block SAD / vector, algorithmic
work) and against the golden metrics."""
import json
import os

import numpy as np
import pytest

import helpers as H
from svtme import api

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "dg_golden.json")))


def _ids(cases):
    return ["%dx%d-res%d-%s" % tuple(c["case"][:4]) for c in cases]


@pytest.mark.parametrize("g", GOLD, ids=_ids(GOLD))
def test_oracle_reproduces_golden_metrics(oracle, g):
    w, h, res, kind, seed, cur, rf = g["case"]
    _, po = H.make_pictures(oracle, w, h, 3, seed=seed, kind=kind)
    m, sad, mv = oracle.dg_hme_level0(po[cur], po[rf], res)
    assert m == g["metrics"]
    # the metrics are the sums of the per-block outputs
    assert int(sad.sum()) == m["tot_dist"] and int((sad > 16 * 16 * 30).sum()) == m["tot_cplx"]
    assert int((mv != 0).any(axis=1).sum()) == m["tot_active"]
    assert (mv % 4 == 0).all()


@pytest.mark.parametrize("isa", [0, 1], ids=["c", "avx2"])
@pytest.mark.parametrize("case", [(320, 184, 0, "pan"), (320, 184, 1, "noise"), (640, 360, 2, "pan"), (448, 264, 3, "pan"),
                                  (448, 264, 5, "noise"), (456, 272, 3, "pan"), (448, 264, 3, "zero"), (64, 64, 3, "pan"),
                                  (1280, 720, 3, "pan")],
                         ids=lambda c: "%dx%d-res%d-%s" % c)
def test_oracle_matches_reference(reference, oracle, case, isa):
    w, h, res, kind = case
    if reference.set_isa(isa) != isa:
        pytest.skip("AVX2 kernels not available")
    try:
        _, pr = H.make_pictures(reference, w, h, 3, seed=w + res, kind=kind)
        _, po = H.make_pictures(oracle, w, h, 3, seed=w + res, kind=kind)
        for cur, rf in ((2, 0), (0, 1)):
            a, _, _ = reference.dg_hme_level0(pr[cur], pr[rf], res)
            b, _, _ = oracle.dg_hme_level0(po[cur], po[rf], res)
            assert a == b, (case, cur, rf)
    finally:
        reference.set_isa(0)


# ---- GPU ----
@pytest.fixture(scope="module")
def dev():
    d = api.MeDevice(device=0)
    yield d
    d.close()


def _gpu_check(dev, oracle, w, h, res, kind, seed, cur, rf):
    frames, po = H.make_pictures(oracle, w, h, 3, seed=seed, kind=kind)
    want_m, want_sad, want_mv = oracle.dg_hme_level0(po[cur], po[rf], res)
    want_px = oracle.last_px_compares()
    a, b = dev.register(frames[cur], cur), dev.register(frames[rf], rf)
    try:
        n0 = dev.launch_count()
        m, sad, mv, ms, px = dev.dg_hme_level0(a, b, res)
        assert dev.launch_count() == n0 + 1 and ms > 0
    finally:
        a.release()
        b.release()
    assert np.array_equal(sad, want_sad), np.argwhere(sad != want_sad)[:5].tolist()
    assert np.array_equal(mv, want_mv), np.argwhere(mv != want_mv)[:5].tolist()
    assert m == want_m
    assert px == want_px
    return m


@pytest.mark.gpu
@pytest.mark.parametrize("g", GOLD, ids=_ids(GOLD))
def test_gpu_matches_oracle_and_golden(dev, oracle, g):
    m = _gpu_check(dev, oracle, *g["case"])
    assert m == g["metrics"]


@pytest.mark.gpu
@pytest.mark.parametrize("case", [(64, 64, 3, "pan"), (72, 40, 3, "noise"), (128, 64, 0, "pan"), (456, 272, 1, "pan"),
                                  (456, 272, 2, "noise"), (1000, 560, 3, "pan"), (448, 264, 6, "noise"), (3840, 2160, 5, "pan")],
                         ids=lambda c: "%dx%d-res%d-%s" % c)
def test_gpu_matches_oracle_more_shapes(dev, oracle, case):
    """Tiny pictures (windows clipped on every side, widths below 8), sizes that are not multiples of 64, every window
    class, and the benchmark's 4K size."""
    w, h, res, kind = case
    _gpu_check(dev, oracle, w, h, res, kind, seed=w + h, cur=1, rf=0)


@pytest.mark.gpu
def test_gpu_per_block_outputs_are_optional(dev, oracle):
    frames, po = H.make_pictures(oracle, 448, 264, 2, seed=5, kind="pan")
    want, _, _ = oracle.dg_hme_level0(po[1], po[0], 3)
    a, b = dev.register(frames[1], 1), dev.register(frames[0], 0)
    m, sad, mv, _, _ = dev.dg_hme_level0(a, b, 3, per_block=False)
    assert m == want and sad is None and mv is None
