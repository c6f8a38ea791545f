#!/usr/bin/env python3
"""Generator of tests/golden/enc_tf_knobs.json: the knob block of the temporal filter's ME as the reference ENCODER sets it up.

Runs HERE only (needs /root/reference):
    make -C oracle enc
    python tests/golden/make_tf_knobs_golden.py

oracle/_ref/SvtAv1EncApp_dump (the C-only reference encoder + the hooks of oracle/enc_overlay/) is run on short synthetic
clips; at the first block of every (centre, neighbour) pair of produce_temporally_filtered_pic the hook after the ME call
(temporal_filtering.c:3174) writes the SvtMeKnobs block svtme_shim_tf_knobs_from_me_context extracts from the live MeContext,
next to the TfControls / scs fields svtme_derive_tf_knobs takes.  tests/test_knobs.py checks the derivation against them.
Distinct records only (the knobs do not depend on the neighbour).
"""
import json
import os
import pathlib
import struct
import subprocess
import sys
import tempfile

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent))
import make_enc_golden as g  # noqa: E402

ROOT = pathlib.Path(__file__).resolve().parents[2]
HDR = ("picture_number", "ref_picture_number", "width", "height", "hme_me_level", "qp_opt", "me_exit_th", "qp",
       "temporal_layer_index", "is_ref", "enc_mode", "knobs_bytes")
CASES = [(2, 192, 128, 17, ()), (4, 192, 128, 17, ()), (6, 200, 136, 17, ()), (8, 704, 480, 9, ()), (10, 320, 192, 17, ()),
         (12, 320, 192, 17, ()), (7, 1280, 720, 3, ()), (5, 640, 368, 9, ("--crf", "20")), (9, 960, 544, 9, ("--crf", "55"))]


def main():
    if not g.ENC.exists():
        sys.exit("build the dump encoder first: make -C oracle enc")
    recs, seen = [], set()
    with tempfile.TemporaryDirectory() as td:
        for preset, w, h, n, extra in CASES:
            clip, tk = os.path.join(td, "c.yuv"), os.path.join(td, "tk.bin")
            g.synth_clip(clip, w, h, n, seed=99 + preset)
            if os.path.exists(tk):
                os.remove(tk)
            cmd = [str(g.ENC), "-i", clip, "-w", str(w), "-h", str(h), "--fps", "30", "-n", str(n), "--preset", str(preset),
                   "--lp", "1", "-b", os.path.join(td, "o.ivf"), *extra]
            r = subprocess.run(cmd, env=dict(os.environ, SVTME_DUMP_TF_KNOBS=tk), stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
            if r.returncode != 0:
                sys.exit("encoder failed:\n" + r.stdout[-1500:])
            data = open(tk, "rb").read()
            off = 0
            while off < len(data):
                hdr = struct.unpack_from("<12Q", data, off)
                knobs = data[off + 96:off + 96 + hdr[11]]
                off += 96 + hdr[11]
                key = (hdr[2:11], knobs[8:])  # without the picture numbers
                if key in seen:
                    continue
                seen.add(key)
                recs.append(dict(zip(HDR, hdr), knobs=knobs.hex(), command=" ".join(["SvtAv1EncApp", *cmd[5:-2], *extra])))
    dst = ROOT / "tests/golden/enc_tf_knobs.json"
    dst.write_text(json.dumps(dict(reference="SVT-AV1 v3.0.2, C-only build (oracle/Makefile target enc)", records=recs), indent=1) + "\n")
    print(f"{dst}: {len(recs)} records")


if __name__ == "__main__":
    main()
