"""Context only: wall time of the C-only reference encoder (oracle/_ref/SvtAv1EncApp_dump, --lp 1) with its
open-loop ME + TPL source search on the host cores and on the GPU (SVTME_GPU_ME=1), same clip, same bitstream.
python tools/enc_wall.py [preset] [width] [height] [frames]"""
import importlib.util, os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("g", os.path.join(ROOT, "tests", "golden", "make_enc_golden.py"))
g = importlib.util.module_from_spec(spec); spec.loader.exec_module(g)
preset, w, h, n = (int(a) for a in (sys.argv[1:5] + ["10", "1280", "720", "33"][len(sys.argv) - 1:]))
with tempfile.TemporaryDirectory() as td:
    clip = os.path.join(td, "c.yuv"); g.synth_clip(clip, w, h, n, seed=99)
    out = {}
    for mode in ("cpu", "gpu", "cpu", "gpu"):
        env = dict(os.environ); env.pop("SVTME_GPU_ME", None)
        if mode == "gpu": env["SVTME_GPU_ME"] = "1"
        ivf = os.path.join(td, mode + ".ivf"); t0 = time.perf_counter()
        r = subprocess.run([str(g.ENC), "-i", clip, "-w", str(w), "-h", str(h), "--fps", "30", "-n", str(n), "--preset", str(preset), "--lp", "1", "-b", ivf],
                           env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        dt = time.perf_counter() - t0
        assert r.returncode == 0, r.stdout[-500:]
        out.setdefault(mode, []).append(dt); out[mode + "_bytes"] = open(ivf, "rb").read()
    print("preset %d %dx%d %d frames: CPU ME %.2f s, GPU ME %.2f s (best of 2), bitstreams identical: %s" % (
        preset, w, h, n, min(out["cpu"]), min(out["gpu"]), out["cpu_bytes"] == out["gpu_bytes"]))
