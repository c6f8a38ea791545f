#!/usr/bin/env python3
"""Generator of the encoder-level golden fixtures tests/golden/enc_*.npz (BASELINE.json config 1 as written:
"ME results dumped from the reference CPU path" of SvtAv1EncApp).

Runs HERE only (needs /root/reference):
    make -C oracle enc                      # oracle/_ref/SvtAv1EncApp_dump: the C-only reference encoder + dump hooks
    python tests/golden/make_enc_golden.py  # writes tests/golden/enc_m10.npz, enc_m4.npz, enc_m6.npz

For each fixture a deterministic synthetic 4:2:0 clip (global pan + a moving patch + noise) is encoded with
`--preset P --lp 1`; the hooks of oracle/enc_overlay/ serialise, for every picture the ME kernel searched, the planes it
read, the knob block the shim extracted from the encoder's own MeContext/pcs, the reference lists and the MeSbResults +
eight pcs arrays the reference wrote.  The fixture keeps every distinct luma plane once.

npz keys:
  meta                 json: preset, size, frames, encoder command
  luma                 [n_planes, H, W] uint8       every distinct full-resolution plane ME read
  pic_<i>_head         uint64[9]: picture_number, W, H, b64_total_count, pu_count, slice_type, hme_level1_planes,
                       temporal_layer_index, sizeof(SvtMeKnobs)
  pic_<i>_knobs        uint8[sizeof(SvtMeKnobs)]
  pic_<i>_refs         int64[2,4] picture numbers (-1 unused);  pic_<i>_planes int32[1+2*4]: index into `luma` of cur, refs (-1)
  pic_<i>_cur1/_cur2   the encoder's own quarter / sixteenth planes of the source picture (decimation parity)
  pic_<i>_res          uint8[...] flat results in svtme_shim_results_layout order
  pic_<i>_gmc / _gate  inputs and outcome of the last-SB global-motion gate
TPL source-based search (src_ops_process.c:608-968), one record per picture whose source path the encoder computed:
  tpl_<i>_head         uint64[24]: picture_number, then the 23 uint32 of the THDR chunk (oracle/enc_overlay/me_dump.c)
  tpl_<i>_refs         uint8 raw TREF chunk: per [2][4] {uint64 poc; int32 group_idx; uint32 excluded, present, max_w, max_h, pad}
  tpl_<i>_planes       int32[9] index into `luma` of the source and of tpl_ref_ds_ptr_array[l][r] (-1)
  tpl_<i>_me           int32: index of the pic_<j> record holding this picture's ME results (-1 for an I picture)
  tpl_<i>_src          uint8 raw TSRC chunk: per 16x16 cell {int16 mv_row, mv_col; int32 best_rf_idx; uint64 ref_frame_poc;
                       uint8 best_mode, best_intra_mode, pad[6]}
"""
import hashlib
import json
import os
import pathlib
import struct
import subprocess
import sys
import tempfile

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parents[2]
ENC = ROOT / "oracle/_ref/SvtAv1EncApp_dump"


def synth_clip(path, w, h, n, seed, pan=(2, 1), noise=2):
    """Textured background panning (2,1) px/frame, a patch moving the other way, a flash at frame 9, +-2 noise.
    pan=(0, 0), noise=0: a still background (the temporal filter's ME leaves early on such blocks)."""
    rng = np.random.default_rng(seed)
    big = rng.integers(0, 256, size=(h // 4 + 64, w // 4 + 64), dtype=np.uint8)
    big = np.kron(big, np.ones((4, 4), dtype=np.uint8)).astype(np.int32)
    # smooth a little so that sub-sampled SADs have gradients
    big = (big + np.roll(big, 1, 0) + np.roll(big, 1, 1) + np.roll(big, (1, 1), (0, 1))) // 4
    patch = rng.integers(0, 256, size=(48, 64)).astype(np.int32)
    with open(path, "wb") as f:
        for t in range(n):
            ox, oy = 40 + pan[0] * t, 30 + pan[1] * t
            y = big[oy:oy + h, ox:ox + w].copy()
            px, py = (w - 80 - 3 * t) % (w - 64), (20 + 2 * t) % (h - 48)
            y[py:py + 48, px:px + 64] = patch
            if t == 9:
                y = y + 12
            y = np.clip(y + rng.integers(-noise, noise + 1, size=y.shape), 0, 255).astype(np.uint8)
            f.write(y.tobytes())
            f.write(np.full((h // 2, w // 2), 128, np.uint8).tobytes() * 2)


def parse_dump(path):
    data = open(path, "rb").read()
    recs, cur, off = [], {}, 0
    while off < len(data):
        tag = data[off:off + 4].decode()
        (n,) = struct.unpack_from("<I", data, off + 4)
        payload = data[off + 8:off + 8 + n]
        off += 8 + n
        if tag == "END ":
            recs.append(cur)
            cur = {}
        else:
            cur[tag.strip()] = payload
    return recs


def build(name, preset, w, h, frames, extra=()):
    with tempfile.TemporaryDirectory() as td:
        clip, dump = os.path.join(td, "clip.yuv"), os.path.join(td, "me.dump")
        synth_clip(clip, w, h, frames, seed=1234 + preset)
        cmd = [str(ENC), "-i", clip, "-w", str(w), "-h", str(h), "--fps", "30", "-n", str(frames), "--preset", str(preset),
               "--lp", "1", "-b", os.path.join(td, "out.ivf"), *extra]
        tdump = os.path.join(td, "tpl.dump")
        env = dict(os.environ, SVTME_DUMP=dump, SVTME_DUMP_TPL=tdump)
        r = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            sys.exit(f"encoder failed:\n{r.stdout[-2000:]}")
        recs = parse_dump(dump)
        trecs = parse_dump(tdump)
    lumas, index, out = [], {}, {}

    def plane_id(buf, W, H):
        key = hashlib.sha1(buf).digest()
        if key not in index:
            index[key] = len(lumas)
            lumas.append(np.frombuffer(buf, np.uint8).reshape(H, W))
        return index[key]

    recs.sort(key=lambda r: struct.unpack_from("<Q", r["HEAD"])[0])
    for i, r in enumerate(recs):
        head = struct.unpack("<Q8I", r["HEAD"])
        W, H = head[1], head[2]
        out[f"pic_{i}_head"] = np.array(head, np.uint64)
        out[f"pic_{i}_knobs"] = np.frombuffer(r["KNOB"], np.uint8)
        refs = np.frombuffer(r["REFS"], np.uint64).astype(np.int64).reshape(2, 4)
        out[f"pic_{i}_refs"] = refs
        planes = np.full(9, -1, np.int32)
        planes[0] = plane_id(r["CUR0"], W, H)
        for li in range(2):
            for ri in range(4):
                if f"R{li}{ri}0" in r:
                    planes[1 + li * 4 + ri] = plane_id(r[f"R{li}{ri}0"], W, H)
        out[f"pic_{i}_planes"] = planes
        out[f"pic_{i}_cur1"] = np.frombuffer(r["CUR1"], np.uint8).reshape(H // 2, W // 2)
        out[f"pic_{i}_cur2"] = np.frombuffer(r["CUR2"], np.uint8).reshape(H // 4, W // 4)
        out[f"pic_{i}_res"] = np.frombuffer(r["RES"], np.uint8)
        out[f"pic_{i}_gmc"] = np.frombuffer(r["GMC"], np.uint8)
        out[f"pic_{i}_gate"] = np.frombuffer(r["GATE"], np.uint8)
    poc_to_pic = {int(out[f"pic_{i}_head"][0]): i for i in range(len(recs))}
    trecs.sort(key=lambda r: struct.unpack_from("<Q", r["THDR"])[0])
    for i, r in enumerate(trecs):
        head = struct.unpack("<Q23I4x", r["THDR"])
        W, H = head[1], head[2]
        out[f"tpl_{i}_head"] = np.array(head, np.uint64)
        out[f"tpl_{i}_refs"] = np.frombuffer(r["TREF"], np.uint8)
        planes = np.full(9, -1, np.int32)
        planes[0] = plane_id(r["TCUR"], W, H)
        for li in range(2):
            for ri in range(4):
                if f"T{li}{ri}0" in r:
                    planes[1 + li * 4 + ri] = plane_id(r[f"T{li}{ri}0"], W, H)
        out[f"tpl_{i}_planes"] = planes
        j = poc_to_pic.get(head[0], -1)
        if "TME" in r:  # the ME results TPL read are the ones ME wrote: keep them once
            assert j >= 0 and bytes(out[f"pic_{j}_res"]) == r["TME"], "TPL saw other ME results than ME dumped"
        out[f"tpl_{i}_me"] = np.array(j if "TME" in r else -1, np.int32)
        out[f"tpl_{i}_src"] = np.frombuffer(r["TSRC"], np.uint8)
    meta = dict(tpl_pictures=len(trecs), preset=preset, width=w, height=h, frames=frames, pictures=len(recs),
                command=" ".join(["SvtAv1EncApp", *cmd[1:]]).replace(os.path.dirname(clip), "<tmp>"),
                reference="SVT-AV1 v3.0.2, C-only build (oracle/Makefile target enc)")
    out["meta"] = np.frombuffer(json.dumps(meta).encode(), np.uint8)
    out["luma"] = np.stack(lumas)
    dst = ROOT / "tests/golden" / f"{name}.npz"
    np.savez_compressed(dst, **out)
    print(f"{dst}: {len(recs)} ME pictures, {len(trecs)} TPL pictures, {len(lumas)} distinct planes, {dst.stat().st_size / 1024:.0f} KiB")


if __name__ == "__main__":
    if not ENC.exists():
        sys.exit("build the dump encoder first: make -C oracle enc")
    build("enc_m10", 10, 320, 192, 17)
    build("enc_m4", 4, 192, 128, 17)
    build("enc_m6", 6, 200, 136, 17)  # incomplete 64x64 blocks on the right and bottom edges
