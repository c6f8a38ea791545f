"""Temporal-filter mode (me_type = ME_MCTF) of the 64x64 motion search: the calls produce_temporally_filtered_pic makes
(temporal_filtering.c:3129-3174) -- one neighbouring picture as the only reference, full-SAD HME, no distance scaling
of the integer-search area, no reference pruning, early exit on the HME SAD, no candidates / statistics.

CPU: the knob mirror, the oracle restatement and the host-compiled device control code against the compiled reference.
GPU: both pipelines against the oracle through the C ABI."""
import ctypes as C
import itertools
import os
import subprocess

import numpy as np
import pytest

import helpers as H
import refbind
from svtme import abi, api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, HH = 448, 264  # 7 x 5 blocks, the last row 8 pixels high
TF_ARRAYS = ("sb_best_sad", "sb_best_mv", "hme_sc", "do_ref", "tf_exit_sb", "tf_blk_counts")


def tf_knobs(binding, level, exit_th, w=W, h=HH, qp_opt=1, qp=35, poc=3):
    return binding.derive_tf_knobs(hme_me_level=level, qp_opt=qp_opt, me_exit_th=exit_th, qp=qp, width=w, height=h,
                                   picture_number=poc, temporal_layer_index=0, is_ref=1)


@pytest.mark.parametrize("level,qp_opt", list(itertools.product(range(5), [0, 1])))
def test_tf_knobs_match_reference(reference, level, qp_opt):
    """svtme_derive_tf_knobs (csrc/knobs.cc) against the reference's svt_aom_sig_deriv_me_tf, byte for byte."""
    for (w, h), qp, th in itertools.product([(320, 240), (640, 360), (832, 480), (1280, 720), (3840, 2160)],
                                            [20, 35, 63], [0, 256]):
        a = reference.derive_tf_knobs(hme_me_level=level, qp_opt=qp_opt, me_exit_th=th, qp=qp, width=w, height=h,
                                      picture_number=7, temporal_layer_index=0, is_ref=1)
        b = api.derive_tf_knobs(hme_me_level=level, qp_opt=qp_opt, me_exit_th=th, qp=qp, width=w, height=h,
                                picture_number=7, temporal_layer_index=0, is_ref=1)
        assert bytes(a) == bytes(b), (w, h, qp, th, a.as_dict(), b.as_dict())
        assert a.me_type == abi.ME_MCTF and a.hme_search_method == 1 and a.num_of_ref_pic_to_search[0] == 1


@pytest.fixture(scope="module")
def clips(reference, oracle):
    out = {}
    for kind in ("pan", "static", "zero"):
        out[kind] = (H.make_pictures(reference, W, HH, 7, seed=3, kind=kind)[1],
                     H.make_pictures(oracle, W, HH, 7, seed=3, kind=kind)[1])
    return out


CASES = list(itertools.product(range(5), [0, 16 * 16, 4000], [1, 3]))


@pytest.mark.parametrize("level,exit_th,dist", CASES)
def test_oracle_matches_reference(reference, oracle, clips, level, exit_th, dist):
    pr, po = clips["pan"]
    k = tf_knobs(reference, level, exit_th)
    a = reference.me_picture(k, pr[3], {(0, 0): pr[3 - dist]}, n_threads=4)
    b = oracle.me_picture(k, po[3], {(0, 0): po[3 - dist]}, n_threads=4)
    H.assert_same(a, b, "tf level %d exit %d dist %d" % (level, exit_th, dist))
    assert int(a["tf_blk_counts"].sum()) == a.n_b64
    if exit_th == 0:
        assert not a["tf_exit_sb"].any()


@pytest.mark.parametrize("kind", ["static", "zero"])
@pytest.mark.parametrize("level", [0, 2, 4])
def test_oracle_matches_reference_flat_clips(reference, oracle, clips, kind, level):
    """Identical / all-zero pictures: every block exits after HME once the threshold is above 0."""
    pr, po = clips[kind]
    k = tf_knobs(reference, level, 16 * 16)
    a = reference.me_picture(k, pr[3], {(0, 0): pr[4]}, n_threads=4)
    b = oracle.me_picture(k, po[3], {(0, 0): po[4]}, n_threads=4)
    H.assert_same(a, b, "%s level %d" % (kind, level))
    assert a["tf_exit_sb"].all()
    assert not a["sb_best_mv"].any() and (a["sb_best_sad"] == abi.MAX_SAD_VALUE).all()


class Emul(refbind._Binding):
    def __init__(self, lib):
        self.lib, self.prefix = lib, "svtemu_"
        lib.svtemu_me_picture.argtypes = [C.POINTER(abi.Knobs), C.POINTER(refbind.RefPic), C.POINTER(refbind.RefPic),
                                          C.POINTER(abi.PictureResults), C.c_int]


@pytest.fixture(scope="module")
def emul():
    d = os.path.join(ROOT, "tests", "host_emul")
    so = os.path.join(d, "libsvtme_emul.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so, os.path.join(d, "emul.cc")])
    return Emul(C.CDLL(so))


@pytest.mark.parametrize("level,exit_th,dist", CASES)
def test_device_control_code_matches_oracle(reference, oracle, emul, clips, level, exit_th, dist):
    """csrc/me_control.cuh compiled for the host with the scalar backend."""
    _, po = clips["pan"]
    k = tf_knobs(reference, level, exit_th)
    a = oracle.me_picture(k, po[3], {(0, 0): po[3 - dist]}, n_threads=4)
    b = emul.me_picture(k, po[3], {(0, 0): po[3 - dist]})
    H.assert_same(a, b, "tf level %d exit %d dist %d" % (level, exit_th, dist))


# ---- GPU: the CUDA path through the C ABI ----
GW, GH = 832, 488  # 13 x 8 blocks, partial last row


@pytest.fixture(scope="module")
def gpu_clip(oracle):
    return H.make_pictures(oracle, GW, GH, 7, seed=11, kind="pan")


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", [0], ids=["staged"])
@pytest.mark.parametrize("level,exit_th,dist", [(0, 0, 1), (0, 256, 2), (1, 0, 3), (1, 4000, 1), (2, 256, 1), (2, 0, 2),
                                                (3, 0, 1), (3, 4000, 3), (4, 0, 1), (4, 4000, 2)])
def test_gpu_matches_oracle(oracle, gpu_clip, pipeline, level, exit_th, dist):
    frames, po = gpu_clip
    k = api.derive_tf_knobs(hme_me_level=level, qp_opt=1, me_exit_th=exit_th, qp=35, width=GW, height=GH,
                            picture_number=3, temporal_layer_index=0, is_ref=1)
    want = oracle.me_picture(k, po[3], {(0, 0): po[3 - dist]}, n_threads=8)
    # the planes are the open-loop ME's (pcs->enable_hme_level1_flag, pic_analysis_process.c:1981): sixteenth from quarter
    dev = api.MeDevice(device=0, pipeline=pipeline)
    try:
        cur = dev.register(frames[3], 3)
        ref = dev.register(frames[3 - dist], 3 - dist)
        got = dev.me_picture(k, cur, {(0, 0): ref})
    finally:
        dev.close()
    for n in TF_ARRAYS:
        assert np.array_equal(want[n], got[n]), (n, np.argwhere(want[n] != got[n])[:5].tolist())
    assert int(got["tf_blk_counts"].sum()) == got.n_b64


def test_compact_outputs_oracle_and_reference(reference, oracle, clips):
    """tf_best_sad / tf_best_mv = the [0][0] rows of sb_best_sad / sb_best_mv."""
    pr, po = clips["pan"]
    k = tf_knobs(reference, 2, 256)
    for b, pics in ((reference, pr), (oracle, po)):
        full = b.me_picture(k, pics[3], {(0, 0): pics[2]}, n_threads=2)
        out = abi.ResultBuffers(full.n_b64, k, tf_compact=True)
        arr = (refbind.RefPic * 8)()
        arr[0] = pics[2].struct
        assert b._f("me_picture")(C.byref(k), C.byref(pics[3].struct), arr, C.byref(out.struct), 2) == 0
        assert "sb_best_sad" not in out.arrays
        assert np.array_equal(out["tf_best_sad"], full["sb_best_sad"][:, 0, 0]) and np.array_equal(out["tf_best_mv"], full["sb_best_mv"][:, 0, 0])
        assert np.array_equal(out["tf_exit_sb"], full["tf_exit_sb"]) and np.array_equal(out["hme_sc"], full["hme_sc"])


@pytest.mark.gpu
@pytest.mark.parametrize("pipeline", [0], ids=["staged"])
def test_gpu_compact_outputs(oracle, gpu_clip, pipeline):
    frames, po = gpu_clip
    k = api.derive_tf_knobs(hme_me_level=2, qp_opt=1, me_exit_th=256, qp=35, width=GW, height=GH, picture_number=3,
                            temporal_layer_index=0, is_ref=1)
    want = oracle.me_picture(k, po[3], {(0, 0): po[1]}, n_threads=8)
    dev = api.MeDevice(device=0, pipeline=pipeline)
    try:
        cur, ref = dev.register(frames[3], 3), dev.register(frames[1], 1)
        t, got = dev.search_async(k, cur, {(0, 0): ref}, tf_compact=True)
        dev.wait(t)
    finally:
        dev.close()
    assert "sb_best_sad" not in got.arrays
    assert np.array_equal(got["tf_best_sad"], want["sb_best_sad"][:, 0, 0]) and np.array_equal(got["tf_best_mv"], want["sb_best_mv"][:, 0, 0])
    for n in ("hme_sc", "tf_exit_sb", "tf_blk_counts"):
        assert np.array_equal(got[n], want[n]), n
