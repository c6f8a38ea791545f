"""N>1 host logic on CPU: two processes over gloo shard the mini-GOPs of a clip, search their pictures independently
(with the oracle standing in for the device -- TEST ONLY) and the job-level reduction gives the whole-job numbers.
The union of the shards must equal the single-process result: the path has no cross-rank data dependence."""
import os
import pickle
import sys
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers as H
from svtme import api, shard, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, HH, GOP, NGOP = 128, 128, 2, 3


def _search_minigops(oracle, gops):
    """{picture: bytes of its ME results} for the non-base pictures of the given mini-GOPs."""
    n = GOP * NGOP + 1
    frames, pics = H.make_pictures(oracle, W, HH, n, seed=5, kind="pan")
    out = {}
    for g in gops:
        for cur in range(g * GOP + 1, (g + 1) * GOP + 1):
            tl = 0 if cur % GOP == 0 else 1
            k = api.derive_knobs(enc_mode=10, qp=35, frame_rate_q16=30 << 16, width=W, height=HH, hierarchical_levels=3,
                                 picture_number=cur, temporal_layer_index=tl, is_ref=1, slice_type_b=1)  # host code, no GPU needed
            lo, hi = g * GOP, (g + 1) * GOP  # references stay inside the mini-GOP (+ its leading base picture)
            l0 = [p for p in range(cur - 1, lo - 1, -1)][:k.num_of_ref_pic_to_search[0]]
            l1 = [p for p in range(cur + 1, hi + 1)][:k.num_of_ref_pic_to_search[1]] or l0[:k.num_of_ref_pic_to_search[1]]
            res = oracle.me_picture(k, pics[cur], H.refs_dict(pics, l0, l1), n_threads=1, extended=False)
            out[cur] = b"".join(res.arrays[a].tobytes() for a in sorted(res.arrays))
    return out


def _worker(rank, world, port, outdir):
    for p in (os.path.join(ROOT, "svt-av1-mirror_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import refbind
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = shard.shard_minigops(NGOP, rank, world)
        need = shard.pictures_of(mine, GOP)
        res = _search_minigops(refbind.Oracle(), mine)
        assert set(res).issubset(set(need))
        dist.barrier()
        # the picture consecutive mini-GOPs share travels rank r -> r + 1
        frames = synth.make_clip(W, HH, len(mine) * GOP + 1, seed=5, kind="pan", first=mine[0] * GOP)
        last = torch.from_numpy(np.ascontiguousarray(frames[-1]))
        recv = torch.zeros_like(last)
        got = shard.handoff_boundary(dist, last, recv)
        assert got == (rank > 0)
        if got:
            assert torch.equal(recv, torch.from_numpy(np.ascontiguousarray(frames[0]))), "handed-over picture differs"
        pictures, ms = shard.reduce_job(len(res), 10.0 * (rank + 1), dist)
        with open(os.path.join(outdir, "rank%d.pkl" % rank), "wb") as f:
            pickle.dump({"gops": list(mine), "res": res, "pictures": pictures, "ms": ms}, f)
    finally:
        dist.destroy_process_group()


def test_shard_partition_properties():
    for n in range(0, 20):
        for world in (1, 2, 3, 4, 8):
            parts = [list(shard.shard_minigops(n, r, world)) for r in range(world)]
            assert sum(parts, []) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    assert shard.pictures_of([1], 4) == [4, 5, 6, 7, 8]
    with pytest.raises(ValueError):
        shard.shard_minigops(4, 2, 2)
    assert shard.reduce_job(7, 3.5) == (7, 3.5)


def test_two_ranks_over_gloo_equal_one_process(oracle):
    world = 2
    port = 29500 + (os.getpid() % 2000)
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_worker, args=(world, port, d), nprocs=world, join=True)
        got = [pickle.load(open(os.path.join(d, "rank%d.pkl" % r), "rb")) for r in range(world)]
    assert sorted(got[0]["gops"] + got[1]["gops"]) == list(range(NGOP))
    assert not set(got[0]["res"]) & set(got[1]["res"])
    merged = dict(got[0]["res"])
    merged.update(got[1]["res"])
    want = _search_minigops(oracle, range(NGOP))
    assert merged.keys() == want.keys() and all(merged[k] == want[k] for k in want)
    for g in got:  # every rank sees the whole-job numbers: SUM of pictures, MAX of times
        assert g["pictures"] == GOP * NGOP and g["ms"] == 20.0


# ---------------------------------------------------------------------------------------------------------------------
# strong scaling: ONE clip over the ranks, reference planes replicated (bench.py --scaling strong)
# ---------------------------------------------------------------------------------------------------------------------
def _minigop(levels=4, n0=3, n1=2):
    n = (1 << levels) + 1
    return [(e["poc"], None, e["list0"][:n0], e["list1"][:n1]) for e in synth.minigop_structure(n, levels, n0, n1)]


def test_strong_plan_covers_every_reference():
    pics = _minigop()
    for world in (1, 2, 3, 4, 8):
        owner, transfers = shard.strong_plan(pics, world)
        assert set(owner) == {0} | {e[0] for e in pics} and set(owner.values()) <= set(range(world))
        assert transfers == sorted(set(transfers)) and all(s != d and owner[p] == s for p, s, d in transfers)
        for n, (poc, _, l0, l1) in enumerate(pics):
            r = n % world
            assert owner[poc] == r  # a picture is uploaded and decimated by the rank that searches it
            for p in l0 + l1:  # every reference is on the searching rank: its own, or received exactly once
                assert owner[p] == r or (p, owner[p], r) in transfers
        used = {(p, n % world) for n, (_, _, l0, l1) in enumerate(pics) for p in l0 + l1}
        assert all((p, d) in used for p, _, d in transfers), "a plane set moves to a rank that never reads it"
        if world == 1:
            assert not transfers
    with pytest.raises(ValueError):
        shard.strong_plan(pics, 0)


def _strong_worker(rank, world, port, outdir):
    for p in (os.path.join(ROOT, "svt-av1-mirror_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pics = _minigop(levels=3, n0=2, n1=2)
        owner, transfers = shard.strong_plan(pics, world)
        # stand-in for a picture's plane set: bytes that only its owner can produce
        planes = {p: torch.full((64,), (7 * p + 3) % 251, dtype=torch.uint8) for p in owner if owner[p] == rank}
        recv = {t: torch.zeros(64, dtype=torch.uint8) for t in transfers if t[2] == rank}
        ops = []
        for t in transfers:  # the same order on every rank: the operations pair up
            if t[1] == rank:
                ops.append(dist.P2POp(dist.isend, planes[t[0]], t[2]))
            elif t[2] == rank:
                ops.append(dist.P2POp(dist.irecv, recv[t], t[1]))
        for r in (dist.batch_isend_irecv(ops) if ops else []):
            r.wait()
        have = dict(planes)
        have.update({t[0]: b for t, b in recv.items()})
        ok = all(int(have[p][0]) == (7 * p + 3) % 251 and bool((have[p] == have[p][0]).all())
                 for n, (_, _, l0, l1) in enumerate(pics) if n % world == rank for p in l0 + l1)
        with open(os.path.join(outdir, "strong%d.pkl" % rank), "wb") as f:
            pickle.dump({"ok": ok, "searches": [n for n in range(len(pics)) if n % world == rank], "received": len(recv)}, f)
    finally:
        dist.destroy_process_group()


def test_strong_exchange_over_gloo():
    world = 2
    port = 31500 + (os.getpid() % 2000)
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_strong_worker, args=(world, port, d), nprocs=world, join=True)
        got = [pickle.load(open(os.path.join(d, "strong%d.pkl" % r), "rb")) for r in range(world)]
    assert all(g["ok"] for g in got), "a rank searched against planes it does not hold"
    assert sorted(got[0]["searches"] + got[1]["searches"]) == list(range(8))
    assert got[0]["received"] + got[1]["received"] > 0
