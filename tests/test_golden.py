"""Golden vectors (tests/golden/me_golden.npz, produced by the compiled reference with tests/golden/make_golden.py):
the oracle restatement on CPU and the CUDA path on the GPU must both reproduce them bit for bit.  Nothing here reads
/root/reference."""
import ctypes as C
import os

import numpy as np
import pytest

import helpers as H
import refbind
from svtme import abi, api

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "me_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def _cases(g):
    n = 0
    while "case%d_knobs" % n in g:
        k = abi.Knobs.from_buffer_copy(g["case%d_knobs" % n].tobytes())
        meta = g["case%d_meta" % n].tolist()
        em, tl, cur, n0, n1 = meta[:5]
        l0, l1 = meta[5:5 + n0], meta[5 + n0:5 + n0 + n1]
        want = {name[len("case%d_" % n):]: g[name] for name in g.files
                if name.startswith("case%d_" % n) and not name.endswith(("_knobs", "_meta"))}
        yield n, k, cur, l0, l1, want, (em, tl)
        n += 1


def _tf_cases(g):
    n = 0
    while "tf%d_knobs" % n in g:
        k = abi.Knobs.from_buffer_copy(g["tf%d_knobs" % n].tobytes())
        lvl, th, cur, nb = g["tf%d_meta" % n].tolist()
        want = {name[len("tf%d_" % n):]: g[name] for name in g.files
                if name.startswith("tf%d_" % n) and not name.endswith(("_knobs", "_meta"))}
        yield n, k, cur, nb, want, (lvl, th)
        n += 1


def _check(got, want, what):
    for name, arr in want.items():
        assert name in got.arrays, "%s: missing output %s" % (what, name)
        assert np.array_equal(got.arrays[name], arr), "%s: %s differs from the golden vector" % (what, name)


def test_golden_file_is_current_abi(gold):
    assert gold["case0_knobs"].size == C.sizeof(abi.Knobs), "SvtMeKnobs layout changed: regenerate tests/golden"


def test_oracle_reproduces_golden_vectors(oracle, gold):
    frames = list(gold["frames"])
    pics = [refbind.PaddedPicture(f, i) for i, f in enumerate(frames)]
    for p in pics:
        oracle.build_planes(p, 1)
    for lvl in (1, 2):
        assert np.array_equal(pics[3].planes[lvl], gold["plane%d_pic3" % lvl]), "decimated plane %d" % lvl
    for n, k, cur, l0, l1, want, _ in _cases(gold):
        got = oracle.me_picture(k, pics[cur], H.refs_dict(pics, l0, l1), n_threads=2, extended=True)
        _check(got, want, "oracle case %d" % n)


def test_oracle_reproduces_golden_temporal_filter_vectors(oracle, gold):
    pics = [refbind.PaddedPicture(f, i) for i, f in enumerate(list(gold["frames"]))]
    for p in pics:
        oracle.build_planes(p, 1)
    n_cases = 0
    for n, k, cur, nb, want, _ in _tf_cases(gold):
        got = oracle.me_picture(k, pics[cur], {(0, 0): pics[nb]}, n_threads=2, extended=True)
        _check(got, want, "oracle tf case %d" % n)
        n_cases += 1
    assert n_cases == 4


def test_knob_mirror_reproduces_golden_temporal_filter_knobs(gold):
    for n, k, cur, nb, want, (lvl, th) in _tf_cases(gold):
        mine = api.derive_tf_knobs(hme_me_level=lvl, qp_opt=1, me_exit_th=th, qp=35, width=192, height=136,
                                   picture_number=cur, temporal_layer_index=0, is_ref=1)
        assert bytes(mine) == bytes(k), "tf case %d: derived knobs differ from the reference's" % n


def test_knob_mirror_reproduces_golden_knobs(gold):
    """svtme_derive_knobs (host code of the product) gives the SvtMeKnobs the reference derived."""
    for n, k, cur, l0, l1, want, (em, tl) in _cases(gold):
        mine = api.derive_knobs(enc_mode=em, qp=35, frame_rate_q16=30 << 16, width=192, height=136,
                                hierarchical_levels=3, picture_number=cur, temporal_layer_index=tl,
                                is_ref=1 if tl < 3 else 0, slice_type_b=0 if n == 5 else 1)
        assert bytes(mine) == bytes(k), "case %d: derived knobs differ from the reference's" % n


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", [0], ids=["staged"])
def test_cuda_reproduces_golden_vectors(gold, pipeline):
    dev = api.MeDevice(max_width=256, max_height=256, max_pictures=16, pipeline=pipeline)
    try:
        frames = list(gold["frames"])
        pd = [dev.register(np.ascontiguousarray(f), i) for i, f in enumerate(frames)]
        for lvl in (1, 2):
            pad = refbind.PADS[lvl]
            assert np.array_equal(dev.read_plane(pd[3], lvl, pad), gold["plane%d_pic3" % lvl]), "decimated plane %d" % lvl
        for n, k, cur, l0, l1, want, _ in _cases(gold):
            got = dev.me_picture(k, pd[cur], H.refs_dict(pd, l0, l1), extended=True)
            _check(got, want, "cuda case %d" % n)
        for n, k, cur, nb, want, _ in _tf_cases(gold):
            got = dev.me_picture(k, pd[cur], {(0, 0): pd[nb]}, extended=True)
            _check(got, want, "cuda tf case %d" % n)
    finally:
        dev.close()
