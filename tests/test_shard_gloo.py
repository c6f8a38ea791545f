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
