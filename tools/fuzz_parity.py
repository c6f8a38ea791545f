"""Randomised parity sweep (run on the GPU box): random picture sizes, presets, temporal layers, clip kinds, reference
lists and modes (open-loop ME, TPL source-based search, temporal filter, dynamic-GOP detector, picture statistics) -- every output of the CUDA
path against the oracle.  python tools/fuzz_parity.py [iterations] [seed] [min_w/8 max_w/8 min_h/8 max_h/8]
(default sizes 64..712 x 64..472; e.g. `60 4242 120 481 68 271` draws 960x544 .. 3840x2160)"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "svt-av1-mirror_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import helpers as H  # noqa: E402
import refbind  # noqa: E402
from svtme import api  # noqa: E402


def run(iters=40, seed=1234, verbose=True, w8=(8, 90), h8=(8, 60), n_threads=8):
    """Returns the list of failure messages (empty = every output of every case bit-exact)."""
    rng = np.random.default_rng(seed)
    oracle = refbind.Oracle()
    bad = []
    for it in range(iters):
        w = int(rng.integers(*w8)) * 8
        h = int(rng.integers(*h8)) * 8
        kind = str(rng.choice(["pan", "pan", "noise", "static", "zero"]))
        n = int(rng.integers(8, 10))  # enough pictures for 4 + 3 references
        seed = int(rng.integers(0, 1 << 30))
        frames, po = H.make_pictures(oracle, w, h, n, seed=seed, kind=kind)
        dev = api.MeDevice(device=0, pipeline=0 * int(rng.integers(0, 2)))  # the draw keeps the case stream of earlier rounds
        try:
            pd = [dev.register(f, i) for i, f in enumerate(frames)]
            cur = int(rng.integers(0, n))
            mode = str(rng.choice(["me", "me", "me", "tf", "dg", "stats", "tpl"]))
            what = "%dx%d %s n=%d seed=%d cur=%d %s" % (w, h, kind, n, seed, cur, mode)
            if mode in ("me", "tpl"):
                em = int(rng.integers(-1, 14))
                hl = int(rng.integers(2, 6))
                tl = int(rng.integers(0, hl + 1))
                extra = {}
                if rng.random() < 0.2:
                    extra["rtc_tune"] = 1
                if rng.random() < 0.2:
                    extra["sc_class1"] = 1
                qp = int(rng.choice([20, 35, 50, 63]))
                k = api.derive_knobs(enc_mode=em, qp=qp, frame_rate_q16=int(rng.choice([24, 30, 60])) << 16, width=w, height=h,
                                     hierarchical_levels=hl, picture_number=cur, temporal_layer_index=tl,
                                     is_ref=1 if tl < hl else 0, slice_type_b=int(rng.random() < 0.85), **extra)
                l0, l1 = H.ref_lists(cur, n, k.num_of_ref_pic_to_search[0], k.num_of_ref_pic_to_search[1])
                assert len(l0) == k.num_of_ref_pic_to_search[0] and len(l1) == k.num_of_ref_pic_to_search[1]
                what += " M%d hl%d tl%d qp%d %s refs %s %s" % (em, hl, tl, qp, extra, l0, l1)
                a = oracle.me_picture(k, po[cur], H.refs_dict(po, l0, l1), n_threads=n_threads)
                px = oracle.last_px_compares()
                b = dev.me_picture(k, pd[cur], H.refs_dict(pd, l0, l1), extended=True)
                H.assert_same(a, b, what)
                assert b.px_compares == px, what
                if mode == "tpl":  # the TPL source-based decisions over those ME results: fused on the device vs the oracle
                    from svtme import abi
                    c = abi.TplCfg()
                    c.use_sad_in_src_search = c.subpel_full_pel = c.intra_dc_only = 1
                    c.dispenser_search_level = int(rng.integers(0, 3)) if k.enable_me_16x16 else int(rng.integers(0, 2))
                    c.disable_intra_pred = int(rng.random() < 0.3)
                    for li, lst in enumerate((l0, l1)):
                        for ri, q in enumerate(lst):
                            c.ref_poc[li][ri] = q
                            c.ref_excluded[li][ri] = int(rng.random() < 0.15)
                    what += " tpl level %d nointra %d" % (c.dispenser_search_level, c.disable_intra_pred)
                    t0 = oracle.tpl_src_search(c, k, po[cur], H.refs_dict(po, l0, l1), a)
                    me1, t1 = dev.me_tpl_picture(k, pd[cur], H.refs_dict(pd, l0, l1), c)
                    assert not t1.diff(t0), (what, t1.diff(t0))
                    assert not me1.diff(a, names=["me_mv", "me_cand", "total_me_candidate_index"]), what
                    t2 = dev.tpl_src_search(k, pd[cur], H.refs_dict(pd, l0, l1), c, a)
                    assert not t2.diff(t0), (what, "standalone", t2.diff(t0))
            elif mode == "tf":
                nb = int(rng.choice([i for i in range(n) if i != cur]))
                k = api.derive_tf_knobs(hme_me_level=int(rng.integers(0, 5)), qp_opt=int(rng.integers(0, 2)),
                                        me_exit_th=int(rng.choice([0, 256, 4000])), qp=35, width=w, height=h, picture_number=cur,
                                        temporal_layer_index=0, is_ref=1)
                a = oracle.me_picture(k, po[cur], {(0, 0): po[nb]}, n_threads=n_threads)
                b = dev.me_picture(k, pd[cur], {(0, 0): pd[nb]})
                for name in ("sb_best_sad", "sb_best_mv", "hme_sc", "do_ref", "tf_exit_sb", "tf_blk_counts"):
                    assert np.array_equal(a[name], b[name]), (what, name)
            elif mode == "dg":
                nb = int(rng.choice([i for i in range(n) if i != cur]))
                res = int(rng.integers(0, 7))
                m0, s0, v0 = oracle.dg_hme_level0(po[cur], po[nb], res)
                m1, s1, v1, _, px = dev.dg_hme_level0(pd[cur], pd[nb], res)
                assert m0 == m1 and np.array_equal(s0, s1) and np.array_equal(v0, v1) and px == oracle.last_px_compares(), what
            else:
                rw, rh, step = int(rng.integers(1, 5)), int(rng.integers(1, 5)), int(rng.choice([1, 4]))
                if (w >> 2) < rw or (h >> 2) < rh:
                    continue
                a, b = oracle.picture_statistics(po[cur], rw, rh, step), dev.picture_statistics(pd[cur], rw, rh, step)
                assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2] == b[2], what
            if verbose:
                print("ok  ", what, flush=True)
        except AssertionError as ex:
            bad.append(str(ex)[:600])
            print("FAIL", str(ex)[:600], flush=True)
        finally:
            dev.close()
    print("fuzz: %d iterations, %d failures" % (iters, len(bad)))
    return bad


if __name__ == "__main__":
    a = [int(v) for v in sys.argv[1:]]
    size = dict(w8=(a[2], a[3]), h8=(a[4], a[5]), n_threads=min(64, os.cpu_count() or 8)) if len(a) >= 6 else {}
    sys.exit(1 if run(a[0] if len(a) > 0 else 40, a[1] if len(a) > 1 else 1234, **size) else 0)
