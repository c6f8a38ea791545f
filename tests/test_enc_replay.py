"""Encoder-level replay (BASELINE.json config 1 as written: "ME results dumped from the reference CPU path").

tests/golden/enc_m10.npz / enc_m4.npz hold what the C-only reference encoder (SvtAv1EncApp --preset 10 / 4, built by
`make -C oracle enc` with the dump hooks of oracle/enc_overlay/) did for every picture its ME kernel searched: the planes, the
knob block taken from the encoder's own MeContext/pcs by the shim, the real reference lists, and the MeSbResults + eight pcs
arrays.  Generator: tests/golden/make_enc_golden.py.

CPU (`-m "not gpu"`): the oracle restatement and, where present, the compiled reference harness replay every picture
bit-exactly; the encoder-side shim (integration/svtme_shim.c, compiled against the reference's headers into
oracle/_ref/libsvtme_shim.so) round-trips knobs and results and reproduces the last-SB global-motion gate.
GPU (`-m gpu`): the CUDA path replays every picture through svtme_picture_search, and once more through the compiled shim
driver (svtme_shim_me_picture: the call me_process.c would make), which scatters into a real PictureParentControlSet.
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np
import pytest

import refbind
from svtme import abi, api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
SHIM_SO = os.path.join(ROOT, "oracle", "_ref", "libsvtme_shim.so")
FIXTURES = ["enc_m10", "enc_m4", "enc_m6"]
FIELDS = (["me_mv", "me_cand", "total_me_candidate_index"] + list(abi.ResultBuffers.SCALARS32) +
          list(abi.ResultBuffers.SCALARS8) + ["picture_sums"])


def parse_res(buf, n, pus, mr, mc):
    """The flat layout of svtme_shim_results_layout (integration/svtme_shim.c): fields in order, each 16-byte aligned."""
    off, out = 0, {}

    def field(name, dt, shape):
        nonlocal off
        off = (off + 15) & ~15
        a = np.frombuffer(buf, dt, int(np.prod(shape)), off).reshape(shape)
        off += a.nbytes
        out[name] = a

    field("me_mv", np.uint32, (n, pus, mr))
    field("me_cand", np.uint8, (n, pus, mc))
    field("total_me_candidate_index", np.uint8, (n, pus))
    for s in abi.ResultBuffers.SCALARS32:
        field(s, np.uint32, (n,))
    for s in abi.ResultBuffers.SCALARS8:
        field(s, np.uint8, (n,))
    field("picture_sums", np.uint64, (4,))
    return out, off


class Fixture:
    def __init__(self, name):
        self.z = np.load(os.path.join(GOLD, name + ".npz"))
        self.meta = json.loads(bytes(self.z["meta"]))
        self.n = self.meta["pictures"]
        self.luma = self.z["luma"]

    def pic(self, i):
        z = self.z
        head = [int(v) for v in z[f"pic_{i}_head"]]
        d = dict(zip(("poc", "W", "H", "n_b64", "pus", "slice_type", "hme_l1", "tl", "knob_bytes"), head))
        assert d["knob_bytes"] == C.sizeof(abi.Knobs), "fixture made with another SvtMeKnobs layout: regenerate"
        d["knobs"] = abi.Knobs.from_buffer_copy(bytes(z[f"pic_{i}_knobs"]))
        d["refs"] = z[f"pic_{i}_refs"]
        d["planes"] = z[f"pic_{i}_planes"]
        d["res_bytes"] = bytes(z[f"pic_{i}_res"])
        d["res"], used = parse_res(d["res_bytes"], d["n_b64"], d["pus"], d["knobs"].max_refs, d["knobs"].max_cand)
        assert used == len(d["res_bytes"])
        d["gmc"] = bytes(z[f"pic_{i}_gmc"])
        d["gate"] = [int(v) for v in z[f"pic_{i}_gate"]]
        d["cur1"], d["cur2"] = z[f"pic_{i}_cur1"], z[f"pic_{i}_cur2"]
        return d


@pytest.fixture(scope="module", params=FIXTURES)
def fx(request):
    return Fixture(request.param)


def test_fixture_is_a_real_encode(fx):
    """The fixture carries what a synthetic harness cannot: the encoder's own prediction structure."""
    pics = [fx.pic(i) for i in range(fx.n)]
    assert fx.n >= 16, "at least two mini-GOPs"
    assert {p["slice_type"] for p in pics} <= {0, 1}  # B_SLICE / P_SLICE (definitions.h): no I picture is searched
    shapes = {(p["knobs"].num_of_list_to_search, tuple(p["knobs"].num_of_ref_pic_to_search)) for p in pics}
    assert len(shapes) >= 2, "reference list shapes vary over the hierarchy: %s" % shapes
    assert len({p["tl"] for p in pics}) >= 3
    for p in pics:  # every reference of every picture is a plane of the fixture
        k = p["knobs"]
        for li in range(k.num_of_list_to_search):
            for ri in range(k.num_of_ref_pic_to_search[li]):
                assert p["planes"][1 + li * 4 + ri] >= 0 and p["refs"][li, ri] >= 0
        assert p["res"]["total_me_candidate_index"].max() >= 1


def _replay_cpu(binding, fx):
    for i in range(fx.n):
        p = fx.pic(i)

        def padded(idx, num):
            pp = refbind.PaddedPicture(fx.luma[idx], num)
            binding.build_planes(pp, p["hme_l1"])
            return pp

        cur = padded(p["planes"][0], p["poc"])
        # the encoder's own decimated planes of this picture == ours (pic_analysis_process.c:1945-2002)
        assert np.array_equal(cur.interior(1), p["cur1"]) and np.array_equal(cur.interior(2), p["cur2"])
        refs = {(li, ri): padded(p["planes"][1 + li * 4 + ri], int(p["refs"][li, ri]))
                for li in range(2) for ri in range(4) if p["planes"][1 + li * 4 + ri] >= 0}
        out = binding.me_picture(p["knobs"], cur, refs, extended=False)
        bad = [f for f in FIELDS if not np.array_equal(p["res"][f], out[f])]
        assert not bad, "picture %d (poc %d): %s differ from the encoder's dump" % (i, p["poc"], bad)


def test_oracle_replays_encoder_dump(oracle, fx):
    _replay_cpu(oracle, fx)


def test_reference_harness_replays_encoder_dump(reference, fx):
    """The harness (oracle/ref_harness.c) around the unmodified motion_estimation.c fed with the dumped knobs gives what the
    encoder gave: the harness' pcs/MeContext filling is complete."""
    _replay_cpu(reference, fx)


# ---------------------------------------------------------------------------------------------------------------------
# TPL source-based search (src_ops_process.c:608-884): TplSrcStats as the encoder's own dispenser wrote them
# ---------------------------------------------------------------------------------------------------------------------
THDR = ("poc W H aligned_w n_b64 pus slice_type tl hl me16 max_cand max_refs max_l0 dispenser_level disable_intra_nref "
        "intra_mode_end use_sad subpel_depth subsample_tx tpl_lad_mg in_loop_ois l0 l1 pad").split()
TREF_DT = np.dtype([("poc", "<u8"), ("group_idx", "<i4"), ("excluded", "<u4"), ("present", "<u4"), ("max_w", "<u4"),
                    ("max_h", "<u4"), ("pad", "<u4")])
TSRC_DT = np.dtype([("mv_row", "<i2"), ("mv_col", "<i2"), ("best_rf_idx", "<i4"), ("ref_frame_poc", "<u8"), ("best_mode", "u1"),
                    ("best_intra_mode", "u1"), ("pad", "u1", (6,))])
FULL_PEL = 3  # SUBPEL_FORCE_STOP (definitions.h): EIGHTH_PEL, QUARTER_PEL, HALF_PEL, FULL_PEL
I_SLICE = 2


def tpl_case(fx, i):
    """One TPL record of the fixture -> (head dict, abi.TplCfg, abi.Knobs, plane indices, abi.ResultBuffers | None, expected)."""
    z = fx.z
    hd = dict(zip(THDR, [int(v) for v in z[f"tpl_{i}_head"]]))
    refs = np.frombuffer(bytes(z[f"tpl_{i}_refs"]), TREF_DT).reshape(2, 4)
    cfg = abi.TplCfg()
    cfg.dispenser_search_level = hd["dispenser_level"]
    cfg.use_sad_in_src_search = hd["use_sad"]
    cfg.subpel_full_pel = int(hd["subpel_depth"] == FULL_PEL)
    cfg.intra_dc_only = int(hd["intra_mode_end"] == 0 and hd["in_loop_ois"] == 1)
    cfg.disable_intra_pred = int(hd["disable_intra_nref"] and hd["tl"] == hd["hl"])
    cfg.slice_type_i = int(hd["slice_type"] == I_SLICE)
    for li in range(2):
        for ri in range(4):
            cfg.ref_excluded[li][ri] = int(refs[li, ri]["excluded"])
            cfg.ref_poc[li][ri] = int(refs[li, ri]["poc"])
            if refs[li, ri]["present"]:  # the MV clamp uses ref_pic_ptr->max_width / max_height: the picture size here
                assert (refs[li, ri]["max_w"], refs[li, ri]["max_h"]) == (hd["W"], hd["H"])
    j = int(z[f"tpl_{i}_me"])
    if j >= 0:
        p = fx.pic(j)
        k = p["knobs"]
        me = abi.ResultBuffers(p["n_b64"], k, extended=False)
        for f in ("me_mv", "me_cand", "total_me_candidate_index"):
            me[f][...] = p["res"][f]
    else:
        k = abi.Knobs()
        k.enable_me_16x16, k.enable_me_8x8 = hd["me16"], int(hd["pus"] == 85)
        k.max_cand, k.max_refs, k.max_l0 = hd["max_cand"], hd["max_refs"], hd["max_l0"]
        me = None
    assert (k.enable_me_16x16, k.max_cand, k.max_refs, k.max_l0) == (hd["me16"], hd["max_cand"], hd["max_refs"], hd["max_l0"])
    gw, gh = (hd["aligned_w"] + 15) >> 4, (hd["H"] + 15) >> 4
    want = np.frombuffer(bytes(z[f"tpl_{i}_src"]), TSRC_DT).reshape(gh, gw)
    return hd, cfg, k, z[f"tpl_{i}_planes"], me, want


def check_tpl(hd, got, want):
    """Every block the dispenser searched: mode, winner, vector and POC equal the encoder's TplSrcStats."""
    live = got["block_size"] > 0
    assert live.any()
    sizes = set(np.unique(got["block_size"][live]).tolist())
    assert sizes <= {16, 32, 64}
    for name, a, b in (("best_mode", got["best_mode"], want["best_mode"]), ("best_rf_idx", got["best_rf_idx"], want["best_rf_idx"]),
                       ("mv_row", got["mv"][..., 0], want["mv_row"]), ("mv_col", got["mv"][..., 1], want["mv_col"]),
                       ("ref_frame_poc", got["ref_frame_poc"], want["ref_frame_poc"])):
        bad = np.argwhere(live & (a.astype(np.int64) != b.astype(np.int64)))
        assert len(bad) == 0, "poc %d: %s differs at cells %s: %s vs %s" % (
            hd["poc"], name, bad[:4].tolist(), [a[tuple(c)] for c in bad[:4]], [b[tuple(c)] for c in bad[:4]])
    return sizes


def test_oracle_tpl_src_search_matches_encoder(oracle, fx):
    n, sizes, newmv = fx.meta["tpl_pictures"], set(), 0
    assert n >= 5
    for i in range(n):
        hd, cfg, k, planes, me, want = tpl_case(fx, i)

        def padded(idx, num):
            pp = refbind.PaddedPicture(fx.luma[idx], num)
            oracle.build_planes(pp, 1)
            return pp

        cur = padded(planes[0], hd["poc"])
        refs = {(li, ri): padded(planes[1 + li * 4 + ri], int(cfg.ref_poc[li][ri]))
                for li in range(2) for ri in range(4) if planes[1 + li * 4 + ri] >= 0}
        got = oracle.tpl_src_search(cfg, k, cur, refs, me)
        sizes |= check_tpl(hd, got, want)
        newmv += int((got["best_mode"] == abi.TPL_NEWMV).sum())
    assert newmv > 0
    if fx.meta["preset"] >= 10:
        assert 32 in sizes  # dispenser_search_level 1 (TPL level 5)
    if fx.meta["width"] % 64:
        assert sizes == {16} or 16 in sizes  # incomplete 64x64 blocks fall back to 16x16


# ---------------------------------------------------------------------------------------------------------------------
# the encoder-side shim, compiled against the reference's headers
# ---------------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def shim():
    if os.path.exists("/root/reference/Source/Lib/Codec/me_process.c"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "shim"], stdout=subprocess.DEVNULL)
    if not os.path.exists(SHIM_SO):
        pytest.skip("oracle/_ref/libsvtme_shim.so not available (built where /root/reference exists)")
    api.load_library()
    C.CDLL(api.LIB_PATH, mode=C.RTLD_GLOBAL)  # the shim driver's svtme_* symbols resolve against the product library
    lib = C.CDLL(SHIM_SO)
    lib.shimtb_results_bytes.restype = C.c_size_t
    lib.shimtb_results_bytes.argtypes = [C.c_uint32] * 4
    lib.shimtb_roundtrip.argtypes = [C.POINTER(abi.Knobs), C.c_char_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_char_p,
                                     C.POINTER(abi.Knobs), C.c_void_p, C.c_void_p]
    lib.shimtb_open.argtypes = [C.c_uint32, C.c_uint32, C.c_int, C.c_uint32, C.c_int, C.POINTER(C.c_void_p)]
    lib.shimtb_me_picture.argtypes = [C.c_void_p, C.POINTER(abi.Knobs), C.c_char_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                      C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.c_void_p, C.c_void_p]
    lib.svtme_shim_close.argtypes = [C.c_void_p]
    lib.svtme_shim_close.restype = None
    lib.svtme_shim_last_error.argtypes = [C.c_void_p]
    lib.svtme_shim_last_error.restype = C.c_char_p
    return lib


def _check_gate(p, gate):
    """gate = (run_gm, level, downsample_level) from the shim; p["gate"] = (ran, level, downsample_level) of the encoder."""
    ran, level, ds = p["gate"][:3]
    assert gate[0] == ran, "me_process.c:276 decision differs"
    if ran:
        assert (gate[1], gate[2]) == (level, ds), "global_me.c:171-205 gate differs: shim %s, encoder %s" % (gate, p["gate"])


def test_shim_roundtrip_and_gm_gate(shim, fx):
    """knobs -> scs/pcs/MeContext -> svtme_shim_knobs_from_me_context gives the knobs back; results scatter into
    MeSbResults + pcs arrays and gather back byte-identical; the gate computed from the gathered sums is the encoder's."""
    ran_any = False
    for i in range(fx.n):
        p = fx.pic(i)
        k = p["knobs"]
        assert shim.shimtb_results_bytes(p["n_b64"], p["pus"], k.max_refs, k.max_cand) == len(p["res_bytes"])
        k_back = abi.Knobs()
        res_back = np.zeros(len(p["res_bytes"]), np.uint8)
        gate = np.zeros(3, np.uint8)
        rc = shim.shimtb_roundtrip(C.byref(k), p["gmc"], p["W"], p["H"], p["slice_type"], p["res_bytes"], C.byref(k_back),
                                   res_back.ctypes.data, gate.ctypes.data)
        assert rc == 0
        assert bytes(k_back) == bytes(k), "knob block does not survive the trip through MeContext/pcs"
        assert res_back.tobytes() == p["res_bytes"]
        _check_gate(p, [int(v) for v in gate])
        ran_any |= bool(p["gate"][0])
    if fx.meta["preset"] <= 4:
        assert ran_any, "global motion runs at this preset: the gate must have been exercised"


# ---------------------------------------------------------------------------------------------------------------------
# GPU
# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
def test_cuda_replays_encoder_dump(fx):
    p0 = fx.pic(0)
    dev = api.MeDevice(device=0, max_inflight=4, sixteenth_from_full=0 if p0["hme_l1"] else 1)
    try:
        resident = {}

        def plane(idx, num):
            if idx not in resident:
                resident[idx] = dev.register(np.ascontiguousarray(fx.luma[idx]), num)
            return resident[idx]

        for i in range(fx.n):
            p = fx.pic(i)
            cur = plane(p["planes"][0], p["poc"])
            assert np.array_equal(dev.read_plane(cur, 1, 0), p["cur1"]) and np.array_equal(dev.read_plane(cur, 2, 0), p["cur2"])
            refs = {(li, ri): plane(p["planes"][1 + li * 4 + ri], int(p["refs"][li, ri]))
                    for li in range(2) for ri in range(4) if p["planes"][1 + li * 4 + ri] >= 0}
            out = dev.me_picture(p["knobs"], cur, refs, extended=False)
            bad = [f for f in FIELDS if not np.array_equal(p["res"][f], out[f])]
            assert not bad, "picture %d (poc %d): %s differ from the encoder's dump" % (i, p["poc"], bad)
    finally:
        dev.close()


@pytest.mark.gpu
@pytest.mark.parametrize("cache", [32, 4])  # 4 is raised to 9 by the shim: every search recycles least-recently-used pictures
def test_cuda_through_compiled_shim(shim, fx, cache):
    """svtme_shim_me_picture -- upload, search, scatter into a PictureParentControlSet, gate -- gives the pcs contents the
    reference encoder had after its own ME of the same picture."""
    p0 = fx.pic(0)
    h = C.c_void_p()
    assert shim.shimtb_open(p0["W"], p0["H"], 0, cache, p0["hme_l1"], C.byref(h)) == 0
    try:
        for i in range(fx.n):
            p = fx.pic(i)
            k = p["knobs"]
            cur = np.ascontiguousarray(fx.luma[p["planes"][0]])
            keep, ptrs, nums = [], (C.c_void_p * 8)(), (C.c_uint64 * 8)()
            for s in range(8):
                idx = p["planes"][1 + s]
                if idx >= 0:
                    a = np.ascontiguousarray(fx.luma[idx])
                    keep.append(a)
                    ptrs[s] = a.ctypes.data
                    nums[s] = int(p["refs"][s // 4, s % 4])
            res = np.zeros(len(p["res_bytes"]), np.uint8)
            gate = np.zeros(3, np.uint8)
            rc = shim.shimtb_me_picture(h, C.byref(k), p["gmc"], p["W"], p["H"], p["slice_type"], cur.ctypes.data, ptrs, nums,
                                        res.ctypes.data, gate.ctypes.data)
            assert rc == 0, shim.svtme_shim_last_error(h)
            got, _ = parse_res(res.tobytes(), p["n_b64"], p["pus"], k.max_refs, k.max_cand)
            bad = [f for f in FIELDS if not np.array_equal(p["res"][f], got[f])]
            assert not bad, "picture %d (poc %d): %s differ from the encoder's pcs" % (i, p["poc"], bad)
            _check_gate(p, [int(v) for v in gate])
    finally:
        shim.svtme_shim_close(h)


@pytest.mark.gpu
def test_cuda_tpl_src_search_matches_encoder(oracle, fx):
    """The TPL source-based decisions of the CUDA path -- standalone over the dumped ME results, and fused behind the picture's
    own ME search -- equal the encoder's TplSrcStats; costs equal the oracle's."""
    dev = api.MeDevice(device=0, max_inflight=4)
    try:
        resident = {}

        def plane(idx, num):
            if idx not in resident:
                resident[idx] = dev.register(np.ascontiguousarray(fx.luma[idx]), num)
            return resident[idx]

        for i in range(fx.meta["tpl_pictures"]):
            hd, cfg, k, planes, me, want = tpl_case(fx, i)
            cur = plane(planes[0], hd["poc"])
            refs = {(li, ri): plane(planes[1 + li * 4 + ri], int(cfg.ref_poc[li][ri]))
                    for li in range(2) for ri in range(4) if planes[1 + li * 4 + ri] >= 0}
            got = dev.tpl_src_search(k, cur, refs, cfg, me)
            check_tpl(hd, got, want)

            def padded(idx, num):
                pp = refbind.PaddedPicture(fx.luma[idx], num)
                oracle.build_planes(pp, 1)
                return pp

            orc = oracle.tpl_src_search(cfg, k, padded(planes[0], hd["poc"]),
                                        {key: padded(planes[1 + key[0] * 4 + key[1]], 0) for key in refs}, me)
            assert not got.diff(orc), "poc %d: %s differ from the oracle" % (hd["poc"], got.diff(orc))
            if me is not None:  # fused: ME on the device, TPL over its results in HBM
                me2, fused = dev.me_tpl_picture(k, cur, refs, cfg)
                assert not me2.diff(me, names=["me_mv", "me_cand", "total_me_candidate_index"])
                assert not fused.diff(got), "poc %d: fused differs from standalone: %s" % (hd["poc"], fused.diff(got))
    finally:
        dev.close()


@pytest.mark.gpu
def test_tpl_refuses_unsupported_controls():
    dev = api.MeDevice(device=0)
    try:
        f = np.zeros((64, 64), np.uint8)
        a, b = dev.register(f, 0), dev.register(f, 1)
        k = api.derive_knobs(enc_mode=6, qp=35, frame_rate_q16=30 << 16, width=64, height=64, hierarchical_levels=3,
                             picture_number=1, temporal_layer_index=0, is_ref=1, slice_type_b=0)
        for field in ("use_sad_in_src_search", "subpel_full_pel", "intra_dc_only"):
            cfg = abi.TplCfg()
            cfg.use_sad_in_src_search = cfg.subpel_full_pel = cfg.intra_dc_only = 1
            setattr(cfg, field, 0)
            with pytest.raises(api.SvtMeError):
                dev.me_tpl_picture(k, a, {(0, 0): b}, cfg)
    finally:
        dev.close()


# ---------------------------------------------------------------------------------------------------------------------
# dynamic-GOP detector and temporal-filter ME against what the encoder's own picture decision / filter searched
# (tests/golden/enc_pairs.npz, generator make_pairs_golden.py)
# ---------------------------------------------------------------------------------------------------------------------
def _pairs():
    z = np.load(os.path.join(GOLD, "enc_pairs.npz"))
    return z, json.loads(bytes(z["meta"]).decode())


def _fnv(words):
    h = 1469598103934665603
    for v in words:
        h = ((h ^ int(v)) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


def _dg_cases(z, meta, binding):
    for i in range(meta["dg_pairs"]):
        hd = [int(v) for v in z["dg_%d_head" % i]]
        luma = z["luma_%dx%d" % (hd[2], hd[3])]
        pics = []
        for idx, num, s16 in ((z["dg_%d_planes" % i][0], hd[0], z["dg_%d_s16" % i]), (z["dg_%d_planes" % i][1], hd[1], z["dg_%d_r16" % i])):
            pp = refbind.PaddedPicture(luma[idx], num)
            binding.build_planes(pp, hd[5])
            assert np.array_equal(pp.interior(2), s16), "sixteenth plane of picture %d differs from the encoder's" % num
            pics.append(pp)
        yield hd, luma, pics, dict(tot_dist=hd[6], tot_cplx=hd[7], tot_active=hd[8], sum_in_vectors=hd[9])


def _tf_cases(z, meta):
    from svtme import abi
    for i in range(meta["tf_pairs"]):
        hd = [int(v) for v in z["tf_%d_head" % i]]
        k = abi.Knobs.from_buffer_copy(bytes(z["tf_%d_knobs" % i]))
        assert k.me_type == 1 and k.num_of_list_to_search == 1 and k.picture_number == hd[0]
        yield hd, z["luma_%dx%d" % (hd[2], hd[3])], z["tf_%d_planes" % i], k, z["tf_%d_blocks" % i], z["tf_%d_hash" % i]


def _check_tf(hd, out, blocks, hashes):
    """what the filter reads after svt_aom_motion_estimation_b64: exit flag, HME centre, the 85 SADs / MVs of [0][0]"""
    n = len(blocks)
    sad = np.asarray(out["sb_best_sad"]).reshape(n, 2, 4, 85)[:, 0, 0]
    mv = np.asarray(out["sb_best_mv"]).reshape(n, 2, 4, 85)[:, 0, 0]
    sc = np.asarray(out["hme_sc"]).reshape(n, 2, 4, 2)[:, 0, 0]
    ex = np.asarray(out["tf_exit_sb"]).reshape(n)
    for b in range(n):
        what = "centre %d neighbour %d block %d" % (hd[0], hd[1], b)
        assert int(ex[b]) == blocks[b][0] and (int(sc[b][0]), int(sc[b][1])) == (blocks[b][1], blocks[b][2]), what
        if not blocks[b][0]:
            assert (_fnv(sad[b]), _fnv(mv[b])) == (int(hashes[b][0]), int(hashes[b][1])), what
    horz = int(np.sum(np.abs(sc[:, 0].astype(np.int64)) > np.abs(sc[:, 1].astype(np.int64))))
    assert [int(v) for v in out["tf_blk_counts"]] == [horz, n - horz]  # motion_estimation.c:2469-2474


def _replay_pairs_cpu(binding):
    z, meta = _pairs()
    assert meta["dg_pairs"] >= 3 and meta["tf_pairs"] >= 15 and meta["tf_early_exits"] >= 100
    for hd, _, (src, ref), want in _dg_cases(z, meta, binding):
        got = binding.dg_hme_level0(src, ref, hd[4])[0]
        assert {k: int(got[k]) for k in want} == want, "detector pair %d / %d" % (hd[0], hd[1])
    for hd, luma, planes, k, blocks, hashes in _tf_cases(z, meta):
        pics = []
        for idx, num in ((planes[0], hd[0]), (planes[1], hd[1])):
            pp = refbind.PaddedPicture(luma[idx], num)
            binding.build_planes(pp, hd[4])
            pics.append(pp)
        _check_tf(hd, binding.me_picture(k, pics[0], {(0, 0): pics[1]}), blocks, hashes)


def test_oracle_dg_and_tf_match_the_encoder(oracle):
    _replay_pairs_cpu(oracle)


def test_reference_harness_dg_and_tf_match_the_encoder(reference):
    _replay_pairs_cpu(reference)


def test_shim_library_exports_its_header(shim):
    """Every function integration/svtme_shim.h declares is defined by the compiled shim (svtme_shim.c + svtme_shim_driver.c)."""
    import re
    text = open(os.path.join(ROOT, "integration", "svtme_shim.h")).read()
    names = sorted(set(re.findall(r"\b(svtme_shim_[a-z0-9_]+)\s*\(", text)))
    assert len(names) >= 15, names
    missing = [n for n in names if not hasattr(shim, n)]
    assert not missing, missing


def test_shim_tf_knobs_and_fill(shim, oracle):
    """The pure-host half of the temporal-filter shim on the reference's own structs (oracle/shim_testbed.c): the knob block of
    every fixture pair survives MeContext / pcs -> svtme_shim_tf_knobs_from_me_context (with pa_me_data NULL, as in picture
    decision), and svtme_shim_tf_fill_b64 leaves in the MeContext what the encoder's filter read after its own ME call."""
    z, meta = _pairs()
    shim.shimtb_tf_roundtrip.argtypes = [C.POINTER(abi.Knobs), C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                         C.c_int, C.c_uint8, C.POINTER(abi.Knobs), C.c_void_p]
    seen = set()
    for hd, luma, planes, k, blocks, hashes in _tf_cases(z, meta):
        if (hd[0], hd[2]) in seen:  # one pair per centre picture is enough here
            continue
        seen.add((hd[0], hd[2]))
        pics = []
        for idx, num in ((planes[0], hd[0]), (planes[1], hd[1])):
            pp = refbind.PaddedPicture(luma[idx], num)
            oracle.build_planes(pp, hd[4])
            pics.append(pp)
        out = oracle.me_picture(k, pics[0], {(0, 0): pics[1]})
        n = len(blocks)
        sad = np.ascontiguousarray(np.asarray(out["sb_best_sad"]).reshape(n, 2, 4, 85)[:, 0, 0], np.uint32)
        mv = np.ascontiguousarray(np.asarray(out["sb_best_mv"]).reshape(n, 2, 4, 85)[:, 0, 0], np.uint32)
        sc = np.asarray(out["hme_sc"]).reshape(n, 2, 4, 2)[:, 0, 0]
        horz = vert = 0
        for b in range(n):
            k_back, got = abi.Knobs(), np.zeros(175, np.int64)
            th = 0 if b % 3 == 0 else 35
            assert shim.shimtb_tf_roundtrip(C.byref(k), hd[2], hd[3], sad[b].ctypes.data, mv[b].ctypes.data, int(sc[b][0]), int(sc[b][1]),
                                            int(blocks[b][0]), th, C.byref(k_back), got.ctypes.data) == 0
            assert bytes(k_back) == bytes(k), (hd, {f: (a, c) for (f, a), (_, c) in zip(k.as_dict().items(), k_back.as_dict().items()) if a != c})
            what = "centre %d block %d" % (hd[0], b)
            assert (int(got[170]), int(got[171])) == (blocks[b][1], blocks[b][2]), what  # the encoder's own HME centre
            assert int(got[172]) == (255 if blocks[b][0] else th), what
            if not blocks[b][0]:  # tier-zero order: 64x64, 4 x 32x32, 16 x 16x16, 64 x 8x8 (me_context.h:52-138)
                assert (_fnv(got[0:85]), _fnv(got[85:170])) == (int(hashes[b][0]), int(hashes[b][1])), what
            horz += int(got[173]) - 100
            vert += int(got[174]) - 200
            assert (int(got[173]) - 100) + (int(got[174]) - 200) == 1
        assert [horz, vert] == [int(v) for v in out["tf_blk_counts"]]
    assert len(seen) >= 6


@pytest.mark.gpu
def test_cuda_dg_and_tf_match_the_encoder():
    from svtme import api
    z, meta = _pairs()
    for hl1 in (0, 1):
        dev = api.MeDevice(device=0, sixteenth_from_full=0 if hl1 else 1)
        try:
            for i in range(meta["dg_pairs"]):
                hd = [int(v) for v in z["dg_%d_head" % i]]
                if hd[5] != hl1:
                    continue
                luma, pl = z["luma_%dx%d" % (hd[2], hd[3])], z["dg_%d_planes" % i]
                src, ref = dev.register(luma[pl[0]], hd[0]), dev.register(luma[pl[1]], hd[1])
                m = dev.dg_hme_level0(src, ref, hd[4])[0]
                assert [int(m[k]) for k in ("tot_dist", "tot_cplx", "tot_active", "sum_in_vectors")] == hd[6:10], hd
                src.release()
                ref.release()
            for hd, luma, planes, k, blocks, hashes in _tf_cases(z, meta):
                if hd[4] != hl1:
                    continue
                cur, nb = dev.register(luma[planes[0]], hd[0]), dev.register(luma[planes[1]], hd[1])
                _check_tf(hd, dev.me_picture(k, cur, {(0, 0): nb}), blocks, hashes)
                cur.release()
                nb.release()
        finally:
            dev.close()


# ---------------------------------------------------------------------------------------------------------------------
# the drop-in, end to end: the reference encoder with its open-loop ME on the GPU writes the same bitstream
# ---------------------------------------------------------------------------------------------------------------------
ENC_BIN = os.path.join(ROOT, "oracle", "_ref", "SvtAv1EncApp_dump")


@pytest.mark.gpu
@pytest.mark.parametrize("preset,w,h,frames,depth", [(10, 320, 192, 17, 8), (6, 200, 136, 17, 8), (4, 192, 128, 17, 8), (10, 1280, 720, 9, 8),
                                                          # 33+ frames below M8: picture decision runs the dynamic-GOP detector
                                                          (6, 200, 136, 40, 8), (7, 640, 496, 34, 8), (7, 896, 744, 34, 8),
                                                          # 10-bit input: ME, TPL and the filter's ME read the 8 MSB planes
                                                          (8, 320, 192, 17, 10), (4, 192, 128, 17, 10)])
def test_encoder_bitstream_identical_with_gpu_me(tmp_path, preset, w, h, frames, depth):
    """oracle/_ref/SvtAv1EncApp_dump = the unmodified C-only reference encoder + one hook in front of me_process.c's 64x64
    block loop (oracle/enc_overlay/).  SVTME_GPU_ME=1 sends every picture's open-loop ME through the compiled shim to
    libsvtme_cuda.so, and every picture's TPL source-based decisions come from tpl_src_kernel (svtme_shim_tpl_picture); mode
    decision, the rest of TPL, rate control and entropy coding then run on those results.  The AV1 bitstream must
    be the one the CPU ME gives, byte for byte, and the GPU run's own ME dump must equal the CPU run's."""
    if not os.path.exists(ENC_BIN):
        pytest.skip("oracle/_ref/SvtAv1EncApp_dump not built (make -C oracle enc, where /root/reference exists)")
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_enc_golden", os.path.join(GOLD, "make_enc_golden.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    clip = str(tmp_path / "clip.yuv")
    gen.synth_clip(clip, w, h, frames, seed=4321 + preset)
    depth_args = []
    if depth == 10:  # the same clip with two random low bits, 16-bit little-endian samples
        a = np.fromfile(clip, np.uint8).astype(np.uint16)
        ((a << 2) | np.random.default_rng(preset).integers(0, 4, size=a.shape, dtype=np.uint16)).astype("<u2").tofile(clip)
        depth_args = ["--input-depth", "10"]
    outs = {}
    for mode in ("cpu", "gpu"):
        ivf, dump = str(tmp_path / (mode + ".ivf")), str(tmp_path / (mode + ".dump"))
        tdump = str(tmp_path / (mode + ".tpl"))
        dgdump = str(tmp_path / (mode + ".dg"))
        tfdump = str(tmp_path / (mode + ".tf"))
        env = dict(os.environ, SVTME_DUMP=dump, SVTME_DUMP_TPL=tdump, SVTME_DUMP_DG=dgdump, SVTME_DUMP_TF=tfdump)
        if mode == "gpu":
            env["SVTME_GPU_ME"] = "1"
            env["SVTME_GPU_STATS"] = str(tmp_path / "gpu_stats.json")
        else:
            env.pop("SVTME_GPU_ME", None)
        r = subprocess.run([ENC_BIN, "-i", clip, "-w", str(w), "-h", str(h), "--fps", "30", "-n", str(frames), "--preset", str(preset),
                            "--lp", "1", "-b", ivf, *depth_args], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        assert r.returncode == 0, "%s encode failed:\n%s" % (mode, r.stdout[-400:])
        outs[mode] = (open(ivf, "rb").read(), os.path.getsize(dump) if os.path.exists(dump) else 0, open(tdump, "rb").read(),
                      open(dgdump).read() if os.path.exists(dgdump) else "",  # written at the first detector call only
                      open(tfdump).read().splitlines() if os.path.exists(tfdump) else [])
    assert len(outs["cpu"][0]) > 1000
    stats = json.load(open(str(tmp_path / "gpu_stats.json")))
    cpu, gpu, problems = outs["cpu"], outs["gpu"], []
    if not (stats["me_pictures"] == frames - 1 and stats["tpl_pictures"] >= 3 and stats["tpl_blocks"] >= 3 * (w // 32) * (h // 32)):
        problems.append("GPU paths not taken: %s" % stats)
    # DGDetectorMetrics of every (source, reference) pair picture decision evaluated
    n_dg = len(cpu[3].splitlines())
    if not (n_dg == (3 * ((frames - 1) // 32) if preset < 8 else 0) and stats["dg_pictures"] == n_dg):
        problems.append("dynamic-GOP detector pairs: %d in the CPU run, stats %s" % (n_dg, stats))
    if gpu[3] != cpu[3]:
        problems.append("dynamic-GOP detector metrics differ:\n%s\nvs\n%s" % (gpu[3], cpu[3]))
    # temporal filter: per (centre, neighbour, block) the early-exit flag, the HME centre and the 85 SADs / MVs the filter reads
    tf = [{tuple(l.split()[:3]): l for l in o[4]} for o in (cpu, gpu)]
    if not (len(tf[0]) >= (w // 64) * (h // 64) and stats["tf_blocks"] == len(cpu[4]) and tf[0].keys() == tf[1].keys()):
        problems.append("temporal-filter blocks: %d in the CPU run, %d in the GPU run, stats %s" % (len(cpu[4]), len(gpu[4]), stats))
    bad = [(tf[0][k], tf[1][k]) for k in sorted(tf[0]) if k in tf[1] and tf[0][k] != tf[1][k]]
    if bad:
        problems.append("temporal-filter ME differs in %d of %d blocks, exited %d; first (cpu, gpu): %s" %
                        (len(bad), len(tf[0]), sum(l.split()[3] == "1" for l in cpu[4]), bad[:3]))
    if not (cpu[1] > 0 and gpu[1] == 0):
        problems.append("the GPU run must not have gone through the CPU block loop (its dump hooks sit there)")
    # TplSrcStats of every picture (+ the planes and ME results TPL read): the GPU run took them from tpl_src_kernel
    if not (len(cpu[2]) > 1000 and gpu[2] == cpu[2]):
        problems.append("TPL source-based decisions differ")
    if gpu[0] != cpu[0]:
        problems.append("bitstreams differ: %d vs %d bytes" % (len(gpu[0]), len(cpu[0])))
    assert not problems, "\n".join(problems)
