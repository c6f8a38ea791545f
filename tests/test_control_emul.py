"""CPU unit test of the product's DEVICE control code (csrc/me_control.cuh) compiled for the host with a scalar
backend (tests/host_emul/emul.cc): the phase-ordered control flow must reproduce the oracle bit-exactly.  This checks
the host-compilable logic only; the CUDA search engines are covered by the -m gpu tests."""
import ctypes as C
import itertools
import os
import subprocess

import pytest

import helpers as H
import refbind
from svtme import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, HH = 448, 264


class Emul(refbind._Binding):
    def __init__(self, lib):
        self.lib, self.prefix = lib, "svtemu_"
        lib.svtemu_me_picture.argtypes = [C.POINTER(abi.Knobs), C.POINTER(refbind.RefPic), C.POINTER(refbind.RefPic),
                                          C.POINTER(abi.PictureResults), C.c_int]


@pytest.fixture(scope="module", params=["lanes-forward", "lanes-reverse"])
def emul(request):
    """The control steps are written for a warp (lane = reference x quadrant); on the host the 32 lanes of a region run
    one after another.  Running them in both orders catches a lane that reads what another lane of the same region
    writes -- which the device, where they run together, would not reproduce."""
    d = os.path.join(ROOT, "tests", "host_emul")
    rev = request.param == "lanes-reverse"
    so = os.path.join(d, "libsvtme_emul_rev.so" if rev else "libsvtme_emul.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared"] + (["-DSVTME_LANES_REVERSE"] if rev else []) +
                          ["-o", so, os.path.join(d, "emul.cc")])
    return Emul(C.CDLL(so))


@pytest.fixture(scope="module")
def clips(oracle):
    return {kind: H.make_pictures(oracle, W, HH, 9, seed=5, kind=kind)[1] for kind in ("pan", "static", "zero")}


CASES = [dict(enc_mode=m, tl=t) for m, t in itertools.product([10, 8, 6, 4, 1], [0, 2, 3])] + [
    dict(enc_mode=9, rtc_tune=1), dict(enc_mode=11, rtc_tune=1), dict(enc_mode=7, rtc_tune=1, sc_class1=1),
    dict(enc_mode=6, sc_class1=1), dict(enc_mode=10, sc_class1=1), dict(enc_mode=6, slice_type_b=0),
    dict(enc_mode=4, ref_list0_count_try=1, ref_list1_count_try=0), dict(enc_mode=12, rate_control_vbr=1),
    dict(enc_mode=7, similar_brightness_refs=1, tl=3),
]


@pytest.mark.parametrize("kind", ["pan", "static", "zero"])
@pytest.mark.parametrize("cfg", CASES, ids=lambda c: "-".join("%s%s" % kv for kv in c.items()))
def test_control_matches_oracle(reference, oracle, emul, clips, cfg, kind):
    if kind != "pan" and cfg.get("tl", 1) != 2 and len(cfg) == 2:
        pytest.skip("adversarial clips: one temporal layer per preset is enough")
    pics = clips[kind]
    cfg = dict(cfg)
    tl = cfg.pop("tl", 1)
    em = cfg.pop("enc_mode")
    k = H.preset_knobs(reference, em, W, HH, 4, tl, **cfg)
    l0, l1 = H.ref_lists(4, 9, k.num_of_ref_pic_to_search[0], k.num_of_ref_pic_to_search[1])
    refs = H.refs_dict(pics, l0, l1)
    a = oracle.me_picture(k, pics[4], refs, n_threads=4)
    b = emul.me_picture(k, pics[4], refs)
    H.assert_same(a, b, "M%d tl%d %s %s" % (em, tl, cfg, kind))
