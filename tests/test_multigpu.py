"""Reference planes across handles / GPUs through the C ABI: a picture is decimated once and its
three padded planes are copied -- peer copy between two handles of one process (svtme_picture_replicate), or
export -> caller's transport -> import (what one process per GPU does over NCCL) -- and the searches that read the
copies must give the oracle's results.  With one GPU both handles sit on device 0; with two or more the copies cross
NVLink."""
import numpy as np
import pytest

import helpers as H
import refbind
from svtme import api

pytestmark = pytest.mark.gpu

W, HH = 448, 264


def _case(oracle):
    frames, po = H.make_pictures(oracle, W, HH, 9, seed=21, kind="pan")
    k = api.derive_knobs(enc_mode=6, qp=35, frame_rate_q16=30 << 16, width=W, height=HH, hierarchical_levels=3,
                         picture_number=4, temporal_layer_index=2, is_ref=1, slice_type_b=1)
    l0, l1 = H.ref_lists(4, 9, k.num_of_ref_pic_to_search[0], k.num_of_ref_pic_to_search[1])
    want = oracle.me_picture(k, po[4], H.refs_dict(po, l0, l1), n_threads=4)
    return frames, k, l0, l1, want


def _second_device():
    import torch
    return 1 if torch.cuda.device_count() > 1 else 0


def test_replicated_references_give_the_same_search(oracle):
    frames, k, l0, l1, want = _case(oracle)
    owner = api.MeDevice(device=0)
    other = api.MeDevice(device=_second_device())
    mine = {p: owner.register(frames[p], p) for p in sorted(set(l0 + l1))}   # decimated once, on the owner
    cur = other.register(frames[4], 4)                                       # the searched picture lives on `other`
    copies = {p: owner.replicate_to(pic, other) for p, pic in mine.items()}
    for pic in mine.values():
        pic.release()  # while the copies may still be in flight: the release waits for them
    for lvl, pad in enumerate(refbind.PADS):
        po = refbind.PaddedPicture(frames[l0[0]], l0[0])
        oracle.build_planes(po, 1)
        assert np.array_equal(other.read_plane(copies[l0[0]], lvl, pad), po.planes[lvl]), "copied plane %d differs" % lvl
    got = other.me_picture(k, cur, H.refs_dict(copies, l0, l1), extended=True)
    H.assert_same(want, got, "replicated references")
    owner.close()
    other.close()


def test_export_import_round_trip(oracle):
    import torch
    frames, k, l0, l1, want = _case(oracle)
    owner = api.MeDevice(device=0)
    dst_dev = _second_device()
    other = api.MeDevice(device=dst_dev)
    nbytes = api.MeDevice.planes_bytes(W, HH)
    copies, keep = {}, []
    for p in sorted(set(l0 + l1)):
        pic = owner.register(frames[p], p)
        buf = torch.empty(nbytes, dtype=torch.uint8, device="cuda:0")
        owner.export_planes(pic, buf.data_ptr(), nbytes)
        pic.release()
        moved = buf.to("cuda:%d" % dst_dev)  # stands in for the NCCL send / recv of a one-process-per-GPU job
        torch.cuda.synchronize()
        keep.append(moved)
        copies[p] = other.import_planes(moved.data_ptr(), W, HH, p)
    cur = other.register(frames[4], 4)
    got = other.me_picture(k, cur, H.refs_dict(copies, l0, l1), extended=True)
    H.assert_same(want, got, "imported references")
    with pytest.raises(api.SvtMeError):
        owner.export_planes(cur, keep[0].data_ptr(), nbytes - 1) if dst_dev == 0 else other.export_planes(cur, keep[0].data_ptr(), nbytes - 1)
    owner.close()
    other.close()


def test_planes_device_is_the_export_without_the_copy(oracle):
    """svtme_picture_planes_device hands out the plane allocation itself: the same bytes svtme_picture_export_planes copies."""
    import ctypes as C
    import torch
    frames, _, _, _, _ = _case(oracle)
    dev = api.MeDevice(device=0)
    pic = dev.register(frames[2], 2)
    nbytes = api.MeDevice.planes_bytes(W, HH)
    ptr, n = dev.planes_device(pic)
    assert n == nbytes and ptr
    buf = torch.empty(nbytes, dtype=torch.uint8, device="cuda:0")
    dev.export_planes(pic, buf.data_ptr(), nbytes)

    class Raw:
        __cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}

    view = torch.as_tensor(Raw(), device="cuda:0")
    assert view.data_ptr() == ptr and torch.equal(view, buf)
    pic.release()
    dev.close()
