#!/usr/bin/env python3
"""Generator of tests/golden/enc_pairs.npz: what the reference ENCODER's dynamic-GOP detector and temporal filter searched.

Runs HERE only (needs /root/reference):
    make -C oracle enc
    python tests/golden/make_pairs_golden.py

oracle/_ref/SvtAv1EncApp_dump (the C-only reference encoder + the hooks of oracle/enc_overlay/) encodes two synthetic clips
with SVTME_DUMP_PAIRS / SVTME_DUMP_TF set.  Kept per pair:
  dynamic-GOP detector (pd_process.c:492-634): the two full planes as picture decision saw them (before temporal filtering),
  the encoder's own sixteenth planes, input_resolution and the DGDetectorMetrics early_hme read;
  temporal filter (temporal_filtering.c:3129-3174): the centre plane at the first ME call of the centre (none of its blocks
  filtered yet), the neighbour's plane, the knob block read from the live MeContext, and per 64x64 block what the filter reads
  after svt_aom_motion_estimation_b64: early-exit flag, HME centre, FNV-1a of the 85 SADs and of the 85 MVs.

npz keys: meta (json), luma_<W>x<H> [n, H, W] u8 (distinct planes per size),
  dg_<i>_head int64[10] (src, ref, W, H, input_resolution, hme_level1_planes, tot_dist, tot_cplx, tot_active, sum_in_vectors),
  dg_<i>_planes int32[2] (src, ref), dg_<i>_s16 / dg_<i>_r16 the encoder's sixteenth planes,
  tf_<i>_head int64[5] (centre, ref, W, H, hme_level1_planes), tf_<i>_knobs u8[], tf_<i>_planes int32[2] (centre, ref),
  tf_<i>_blocks int64[n_b64, 3] (exit, hme_sc_x, hme_sc_y), tf_<i>_hash uint64[n_b64, 2] (SADs, MVs; 0 where exited).
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

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent))
import make_enc_golden as g  # noqa: E402

ROOT = pathlib.Path(__file__).resolve().parents[2]
# (preset, w, h, frames, clip kwargs): 40 frames below M8 -> the detector runs; a still clip -> the filter's ME exits early
CASES = [(6, 200, 136, 40, {}), (10, 320, 192, 17, dict(pan=(0, 0), noise=0)), (8, 704, 480, 9, dict(pan=(0, 0), noise=1))]


def main():
    if not g.ENC.exists():
        sys.exit("build the dump encoder first: make -C oracle enc")
    out, lumas, index, meta = {}, {}, {}, []
    n_dg = n_tf = 0

    def plane_id(buf, W, H):
        key = (W, H, hashlib.sha1(buf).digest())
        if key not in index:
            lst = lumas.setdefault((W, H), [])
            index[key] = len(lst)
            lst.append(np.frombuffer(buf, np.uint8).reshape(H, W))
        return index[key]

    with tempfile.TemporaryDirectory() as td:
        for preset, w, h, n, kw in CASES:
            clip, pairs, tf = os.path.join(td, "c.yuv"), os.path.join(td, "pairs.bin"), os.path.join(td, "tf.txt")
            g.synth_clip(clip, w, h, n, seed=555 + preset, **kw)
            cmd = [str(g.ENC), "-i", clip, "-w", str(w), "-h", str(h), "--fps", "30", "-n", str(n), "--preset", str(preset),
                   "--lp", "1", "-b", os.path.join(td, "o.ivf")]
            r = subprocess.run(cmd, env=dict(os.environ, SVTME_DUMP_PAIRS=pairs, SVTME_DUMP_TF=tf), stdout=subprocess.PIPE,
                               stderr=subprocess.STDOUT, text=True)
            if r.returncode != 0:
                sys.exit("encoder failed:\n" + r.stdout[-1500:])
            blocks = {}
            for line in open(tf):
                c, rf, b, ex, sx, sy, hs, hm = line.split()
                blocks.setdefault((int(c), int(rf)), {})[int(b)] = (int(ex), int(sx), int(sy), int(hs, 16), int(hm, 16))
            centre_plane = None
            for rec in g.parse_dump(pairs):
                if "DGHD" in rec:
                    hd = struct.unpack("<10q", rec["DGHD"])
                    W, H = hd[2], hd[3]
                    out[f"dg_{n_dg}_head"] = np.array(hd, np.int64)
                    out[f"dg_{n_dg}_planes"] = np.array([plane_id(rec["DGS0"], W, H), plane_id(rec["DGR0"], W, H)], np.int32)
                    out[f"dg_{n_dg}_s16"] = np.frombuffer(rec["DGS2"], np.uint8).reshape(H // 4, W // 4)
                    out[f"dg_{n_dg}_r16"] = np.frombuffer(rec["DGR2"], np.uint8).reshape(H // 4, W // 4)
                    n_dg += 1
                else:
                    hd = struct.unpack("<6q", rec["TFHD"])
                    W, H = hd[2], hd[3]
                    if "TFC0" in rec:
                        centre_plane = plane_id(rec["TFC0"], W, H)
                    b = blocks[(hd[0], hd[1])]
                    nb = ((W + 63) // 64) * ((H + 63) // 64)
                    assert sorted(b) == list(range(nb)), "the filter did not search every block of the pair"
                    out[f"tf_{n_tf}_head"] = np.array(hd[:5], np.int64)
                    out[f"tf_{n_tf}_knobs"] = np.frombuffer(rec["TFKN"], np.uint8)
                    out[f"tf_{n_tf}_planes"] = np.array([centre_plane, plane_id(rec["TFR0"], W, H)], np.int32)
                    out[f"tf_{n_tf}_blocks"] = np.array([b[i][:3] for i in range(nb)], np.int64)
                    out[f"tf_{n_tf}_hash"] = np.array([(0, 0) if b[i][0] else b[i][3:] for i in range(nb)], np.uint64)
                    n_tf += 1
            meta.append(dict(preset=preset, width=w, height=h, frames=n, clip=kw and {k: list(v) if isinstance(v, tuple) else v for k, v in kw.items()},
                             command=" ".join(["SvtAv1EncApp", *cmd[5:-2]])))
    for (W, H), lst in lumas.items():
        out[f"luma_{W}x{H}"] = np.stack(lst)
    exits = sum(int(out[f"tf_{i}_blocks"][:, 0].sum()) for i in range(n_tf))
    out["meta"] = np.frombuffer(json.dumps(dict(cases=meta, dg_pairs=n_dg, tf_pairs=n_tf, tf_early_exits=exits,
                                                reference="SVT-AV1 v3.0.2, C-only build (oracle/Makefile target enc)")).encode(), np.uint8)
    dst = ROOT / "tests/golden/enc_pairs.npz"
    np.savez_compressed(dst, **out)
    print(f"{dst}: {n_dg} detector pairs, {n_tf} filter pairs ({exits} early exits), {sum(len(v) for v in lumas.values())} planes, "
          f"{dst.stat().st_size / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
