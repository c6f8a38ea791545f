#!/usr/bin/env python
"""bench.py -- open-loop ME pictures/s at 4K (BASELINE.json configs[2]: 3840x2160 8-bit, preset 6).

One STEP = the hot path over --gops-per-step mini-GOPs per GPU, streamed without draining the pipeline in between
(about 0.1 s of device time): every picture is decimated (quarter + sixteenth planes + padding) and every non-intra
picture is searched against its reference lists (pre-HME, HME L0/L1/L2, full-pel 64x64 search, pruning, candidates,
per-SB statistics).  Pictures are independent objects: with N GPUs every rank runs its own mini-GOPs (weak scaling, no
data-path collective; the default line).  --scaling strong deals the pictures of ONE clip round-robin over the GPUs and
replicates decimated reference planes to the GPUs that need them (svtme_picture_export_planes -> NCCL ->
svtme_picture_import_planes).  --config 1..5 selects a BASELINE.json configuration (3 = the default).

  value     device-resident: luma planes already in HBM, results stay in HBM
  e2e       through the public C ABI with HOST buffers: H2D of every luma plane and D2H of every result inside the
            timed region
  roofline  the search of one picture (the stream-ordered st_* kernels of csrc/me_staged.cu, launched serialised with
            CUDA events around them): algorithmic pixel-compares / device time against the measured VABSDIFF4 issue
            peak (INT pipe -- SAD is not a tensor-core op), plus the HBM view of the same launches
  cpu_baseline / --impl reference: the UNMODIFIED reference C/AVX2 ME (oracle/_ref/libsvtref.so, compiled from the
            reference's own sources) on the host cores, on a bounded sample of the same workload
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "svt-av1-mirror_b200"))

import numpy as np  # noqa: E402

Q16 = 1 << 16


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--preset", type=int, default=6)
    ap.add_argument("--hier-levels", type=int, default=5, help="mini-GOP = 2^levels pictures")
    ap.add_argument("--inflight", type=int, default=8, help="searches in flight (one stream each)")
    ap.add_argument("--cpu-sample", type=int, default=32, help="pictures of the workload timed on the host cores per step")
    ap.add_argument("--cpu-seconds", type=float, default=8.0, help="CPU-baseline budget of the default run (about this much work)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--mode", default="me", choices=["me", "tf", "dg"],
                    help="me: the headline open-loop ME workload; tf: the temporal-filter (ME_MCTF) searches of the same "
                         "mini-GOP -- a secondary measurement of SURVEY section 8(f) rank 2, not the headline metric")
    ap.add_argument("--tf-window", type=int, default=3, help="tf mode: neighbouring pictures searched on each side of a centre")
    ap.add_argument("--config", type=int, default=0, choices=[0, 1, 2, 3, 4, 5],
                    help="BASELINE.json configs[N-1]: 1 = 1080p M10, 2 = 1080p M4 (7 references), 3 = 4K M6 (the default "
                         "workload), 4 = 4K 10-bit input (ME on the 8-bit MSB plane), 5 = 8K M8; 0 = the flags as given")
    ap.add_argument("--gops-per-step", type=int, default=0, help="mini-GOPs streamed per step and GPU (0: about 0.1 s of device time)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--bit-depth", type=int, default=8, choices=[8, 10])
    ap.add_argument("--tpl", action="store_true",
                    help="fuse the TPL source-based search (src_ops_process.c:608-884) behind every picture's ME: the lookahead "
                         "consumer of the ME results reads them in HBM (BASELINE.json configs[3])")
    a = ap.parse_args()
    if a.config:
        cfg = {1: dict(width=1920, height=1080, preset=10, hier_levels=4), 2: dict(width=1920, height=1080, preset=4, hier_levels=5),
               3: dict(width=3840, height=2160, preset=6, hier_levels=5), 4: dict(width=3840, height=2160, preset=6, hier_levels=5, bit_depth=10, tpl=True),
               5: dict(width=7680, height=4320, preset=8, hier_levels=4)}[a.config]
        for k, v in cfg.items():
            setattr(a, k, v)
    if a.gops_per_step <= 0:  # ~0.1 s of device time per step at the rates of DESIGN.md section 8
        px = a.width * a.height * (1 << a.hier_levels)
        a.gops_per_step = max(1, min(64, int(round(12 * (3840 * 2160 * 32) / px * (0.35 if a.preset <= 4 else 1.0)))))
        if a.mode != "me":
            a.gops_per_step = 1
    return a


SUBMIT_THREADS = int(os.environ.get("SVTME_BENCH_THREADS", "1"))  # host threads submitting searches in the device-resident step


class Workload:
    """A synthetic 8-bit clip of one mini-GOP (+ its leading base picture) and its prediction structure."""

    def __init__(self, args, seed, minigop=0, derive=None, derive_tf=None):
        from svtme import synth  # numpy only
        if derive is None:
            from svtme import api
            derive, derive_tf = api.derive_knobs, api.derive_tf_knobs
        self.w, self.h, self.hl = args.width, args.height, args.hier_levels
        self.gop = 1 << self.hl
        self.n_frames = self.gop + 1
        # mini-GOP `minigop` of the clip: its 2^levels pictures plus the leading base picture it shares with its predecessor
        self.frames = synth.make_clip(self.w, self.h, self.n_frames, seed=seed, kind="pan", first=minigop * self.gop)
        k0 = derive(enc_mode=args.preset, qp=35, frame_rate_q16=60 * Q16, width=self.w, height=self.h,
                              hierarchical_levels=self.hl, picture_number=1, temporal_layer_index=1, is_ref=1, slice_type_b=1)
        self.pics = []  # decode order
        for e in synth.minigop_structure(self.n_frames, self.hl, k0.max_l0, max(1, k0.max_refs - k0.max_l0)):
            k = derive(enc_mode=args.preset, qp=35, frame_rate_q16=60 * Q16, width=self.w, height=self.h,
                                 hierarchical_levels=self.hl, picture_number=e["poc"], temporal_layer_index=e["temporal_layer"],
                                 is_ref=e["is_ref"], slice_type_b=1, ref_list0_count_try=len(e["list0"]),
                                 ref_list1_count_try=len(e["list1"]))
            l0 = e["list0"][:k.num_of_ref_pic_to_search[0]]
            l1 = e["list1"][:k.num_of_ref_pic_to_search[1]]
            self.pics.append((e["poc"], k, l0, l1))
        self.bit_depth = args.bit_depth
        self.name = "%dx%d %d-bit 4:2:0 synthetic%s, preset M%d, mini-GOP %d (hier. levels %d), refs<=%d+%d" % (
            self.w, self.h, args.bit_depth, " (ME on the 8-bit MSB plane)" if args.bit_depth == 10 else "", args.preset, self.gop, self.hl,
            k0.max_l0, k0.max_refs - k0.max_l0)
        self.units, self.unit, self.extended = self.gop, "pictures", False
        self.metric = "ME pictures/sec at 4K"
        self.tpl = bool(getattr(args, "tpl", False)) and args.mode == "me"
        if self.tpl:
            # TPL controls of the preset: level 4 up to M8 (16x16 blocks), level 5 above (32x32 blocks); both SAD / FULL_PEL / DC
            # (get_tpl_params_level + set_tpl_params, initial_rc_process.c:279-374).  Every reference lies in the TPL group.
            from svtme import abi
            self.tpl_cfgs = []
            for poc, k, l0, l1 in self.pics:
                c = abi.TplCfg()
                c.dispenser_search_level = 0 if args.preset <= 8 else 1
                c.use_sad_in_src_search = c.subpel_full_pel = c.intra_dc_only = 1
                c.disable_intra_pred = int(k.temporal_layer_index == self.hl)  # disable_intra_pred_nref, src_ops_process.c:567
                for li, lst in enumerate((l0, l1)):
                    for ri, q in enumerate(lst):
                        c.ref_poc[li][ri] = q
                self.tpl_cfgs.append(c)
            self.name += ", TPL source-based search fused behind each picture's ME"
        self.needed = list(range(self.n_frames))  # pictures a step uploads: the whole mini-GOP + its leading base picture
        if args.mode == "tf":
            # The temporal filter's motion searches in this mini-GOP: the filtered (centre) pictures are the leading
            # picture and the layer-0 / layer-1 pictures; each searches its +-tf_window neighbours one at a time
            # (temporal_filtering.c:3129-3174).  TfControls of presets M5-M7 (tf level 5, enc_handle.c:2951-3023):
            # hme_me_level 2, me_exit_th 0, qp_opt 1.
            self.pics = []
            for c in (0, self.gop // 2, self.gop):
                k = derive_tf(hme_me_level=2, qp_opt=1, me_exit_th=0, qp=35, width=self.w, height=self.h,
                                        picture_number=c, temporal_layer_index=0 if c != self.gop // 2 else 1, is_ref=1)
                for d in range(1, args.tf_window + 1):
                    for nb in (c - d, c + d):
                        if 0 <= nb < self.n_frames:
                            self.pics.append((c, k, [nb], []))
            self.units, self.unit, self.extended = len(self.pics), "picture pairs", True
            self.needed = sorted(set(p for c, _, l0, _ in self.pics for p in [c] + l0))  # the filtered pictures and their neighbours
            self.metric = "temporal-filter ME picture-pair searches/sec at 4K"
            self.name = "%dx%d 8-bit synthetic, temporal-filter ME (ME_MCTF, tf hme_me_level 2 = presets M5-M7), centres 0/%d/%d of a mini-GOP of %d, +-%d neighbours = %d pair searches" % (
                self.w, self.h, self.gop // 2, self.gop, self.gop, args.tf_window, len(self.pics))


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons of one GPU sampled through NVML while the timed region runs."""

    def __init__(self, index, period=0.05):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._halt = index, [], set(), None, threading.Event()
        self.period = period

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {getattr(nv, n): n[len("nvmlClocksThrottleReason"):] for n in dir(nv)
                     if n.startswith("nvmlClocksThrottleReason") and n not in ("nvmlClocksThrottleReasonAll", "nvmlClocksThrottleReasonNone")}
            while not self._halt.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, n in names.items():
                    if bit and (r & bit) and n != "GpuIdle":
                        self.reasons.add(n)
                time.sleep(self.period)
        except Exception as ex:  # NVML missing: report that rather than invent clocks
            self.reasons.add("nvml_unavailable:%s" % type(ex).__name__)

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def run_step_device(dev, wl, d_luma, inflight, gops, handoff=None, tpl_bufs=None):
    """Device-resident step: `gops` mini-GOPs streamed back to back -- every picture registered from HBM-resident luma
    (copy, pad, decimate), searched in decode order with `inflight` searches in flight across the mini-GOP boundaries;
    results stay in HBM.  With several ranks the leading base picture of the step arrives from the previous rank."""
    w, h = wl.w, wl.h
    first = handoff() if handoff else None
    pending, live = [], []

    def submit(items, pics, window, g=0):
        for n, (poc, k, l0, l1) in enumerate(items):
            refs = {(0, i): pics[p] for i, p in enumerate(l0)}
            refs.update({(1, i): pics[p] for i, p in enumerate(l1)})
            if len(pending) >= window:
                dev.wait(pending.pop(0))
            if wl.tpl:  # ME results stay in HBM; the (small) TPL decisions come back: they are what the lookahead reads
                pending.append(dev.search_tpl_async(k, pics[poc], refs, wl.tpl_cfgs[n], want_me=False,
                                                    tpl=tpl_bufs[(g & 1) * len(wl.pics) + n])[0])
            else:
                pending.append(dev.search_device_async(k, pics[poc], refs))

    for g in range(gops):
        pics = {i: dev.register_device((first if (g == 0 and i == 0 and first) else d_luma.data_ptr() + i * w * h), w, w, h, i)
                for i in range(wl.n_frames)}
        live.append(pics)
        nt = SUBMIT_THREADS
        if nt <= 1 or wl.tpl:
            submit(wl.pics, pics, inflight, g)
        else:  # the encoder's ME kernel threads submit concurrently (me_process.c); the library is thread-safe
            th = [threading.Thread(target=submit, args=(wl.pics[i::nt], pics, max(1, inflight // nt))) for i in range(nt)]
            for t in th:
                t.start()
            for t in th:
                t.join()
        if len(live) > 1:  # the mini-GOP before this one: its last searches may still run -- the release waits for them on the device
            for q in live.pop(0).values():
                q.release()
    for t in pending:
        dev.wait(t)
    for pics in live:
        for q in pics.values():
            q.release()


def run_step_host(dev, wl, h_luma, outs, inflight, gops, handoff=None, tpl_bufs=None):
    """End-to-end step through the C ABI with host buffers: H2D of each luma plane, D2H of each picture's results.
    Pictures are registered (uploaded + decimated, asynchronously) in the order the searches need them, so the uploads
    of later pictures overlap the searches of earlier ones; every picture of a mini-GOP is uploaded exactly once, and
    the next mini-GOP's uploads start while the last searches of this one are in flight."""
    w, h = wl.w, wl.h
    base = h_luma.data_ptr()
    bps = 2 if wl.bit_depth == 10 else 1
    first = handoff() if handoff else None  # the shared base picture comes over NVLink, not over the host link
    pending, live = [], []
    for g in range(gops):
        pics = {}
        live.append(pics)
        if g == 0 and first:
            pics[0] = dev.register_device(first, w, w, h, 0)

        def pic(i):
            if i not in pics:
                pics[i] = (dev.register_ptr16 if bps == 2 else dev.register_ptr)(base + i * w * h * bps, w, w, h, i)
            return pics[i]

        for n, (poc, k, l0, l1) in enumerate(wl.pics):
            refs = {(0, i): pic(p) for i, p in enumerate(l0)}
            refs.update({(1, i): pic(p) for i, p in enumerate(l1)})
            cur = pic(poc)
            if len(pending) >= inflight:
                dev.wait(pending.pop(0))
            slot = (g & 1) * len(wl.pics) + n  # two sets of result buffers alternate
            if wl.tpl:
                pending.append(dev.search_tpl_async(k, cur, refs, wl.tpl_cfgs[n], out=outs[slot], tpl=tpl_bufs[slot])[0])
            else:
                pending.append(dev.search_async(k, cur, refs, out=outs[slot])[0])
        for i in wl.needed:  # a picture of the workload that no search touched still counts as uploaded
            pic(i)
        if len(live) > 1:
            for q in live.pop(0).values():
                q.release()
    for t in pending:
        dev.wait(t)
    for pics in live:
        for q in pics.values():
            q.release()


def config_dict(wl, args, world):
    """The `config` object of the JSON line: the same keys and values in both arms."""
    return {"workload": wl.name, "%s_per_step_per_gpu" % wl.unit.replace(" ", "_"): wl.units * args.gops_per_step,
            "mini_gops_per_step_per_gpu": args.gops_per_step, "scaling": args.scaling,
            "parallelism": ("pictures of one clip dealt round-robin over %d rank(s), decimated reference planes replicated by NCCL p2p" % world)
                           if args.scaling == "strong" else "pictures sharded by mini-GOP, %d rank(s)" % world,
            "l2": "inputs larger than L2 (%.0f MB of planes per mini-GOP)" % (wl.n_frames * wl.w * wl.h * 1.55 / 1e6), "inflight": args.inflight}


def reference_arm(wl, steps, warmup, seconds=None, threads=None, one_core=True):
    """The reference's own CPU ME (oracle/_ref/libsvtref.so: its unmodified C sources with the SIMD twins its RTCD would
    bind on this host -- AVX2, and the AVX-512 svt_sad_loop_kernel where the CPU has it) on the host cores, parallelised
    the way the encoder is: one pool of ME threads pulls (picture, 6 x 8 segment) items of the whole mini-GOP, the
    decimation of its pictures runs on the same pool first (oracle/ref_harness.c: svtref_me_minigop).  One step = one
    mini-GOP of the workload (a bounded sample of the GPU arm's step, which streams several).  With `seconds` the steps
    repeat until about that much time has been spent.  Nothing of the product is loaded here."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refbind
    kind = "reference"
    if not refbind.have_reference():
        raise SystemExit("oracle/_ref/libsvtref.so is not built")
    ref = refbind.Reference()
    isa = ref.set_isa(2)
    what = "reference C sources compiled by oracle/Makefile, %s kernels" % {0: "C", 1: "AVX2", 2: "AVX2 + AVX-512 svt_sad_loop_kernel"}[isa]
    cores = threads or os.cpu_count() or 1
    padded = {p: refbind.PaddedPicture(wl.frames[p], p) for p in wl.needed}
    items = []
    for poc, k, l0, l1 in wl.pics:
        refs = {(0, i): padded[p] for i, p in enumerate(l0)}
        refs.update({(1, i): padded[p] for i, p in enumerate(l1)})
        items.append((k, padded[poc], refs))
    build = [padded[p] for p in wl.needed]
    n_b64 = ((wl.w + 63) // 64) * ((wl.h + 63) // 64)
    from svtme import abi  # ctypes mirror of the result structs (pure Python)
    outs = [abi.ResultBuffers(n_b64, k, extended=wl.extended) for k, _, _ in items]

    def step(n=cores, its=items, bld=build, o=outs):
        ref.me_minigop(its, bld, n, outs=o[:len(its)])

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    done = 0
    while done < steps or (seconds is not None and time.perf_counter() - t0 < seconds):
        step()
        done += 1
    steps = done
    dt = time.perf_counter() - t0
    pps = len(items) * steps / dt
    cb = {"value": pps, "unit": wl.unit + "/s", "cores": cores, "kind": kind,
          "sample": "%d mini-GOP(s) of the workload (%d searches + %d decimations each) in %.1f s, %s, %d threads on one work queue" % (
              steps, len(items), len(build), dt, what, cores)}
    if wl.tpl:  # the compiled reference has no stand-alone TPL entry: the CPU figure is ME alone, i.e. LESS work than the GPU arm does
        cb["sample"] += "; ME only -- the TPL source-based search the GPU arm adds is not in this CPU figure"
    if one_core:  # the same code on ONE host thread, on a few searches spread over the temporal layers
        idx = sorted(set(int(round(i * (len(items) - 1) / 3)) for i in range(4)))
        t1 = time.perf_counter()
        step(1, [items[i] for i in idx], [items[i][1] for i in idx])
        cb["one_core"] = {"value": len(idx) / (time.perf_counter() - t1), "unit": wl.unit + "/s", "sample": "searches %s of the mini-GOP" % idx}
    return cb, dt, steps


DG_PAIRS = lambda gop: [(gop, 0), (gop, gop // 2), (gop // 2, 0)]  # noqa: E731  eval_sub_mini_gop (pd_process.c:682-705): (end, start), (end, mid), (mid, start)


def dg_cpu_arm(frames, pairs, res, seconds, steps=None):
    """The reference's own dg_detector_hme_level0 (oracle/_ref, AVX2 kernels) on the host cores: the picture pairs of the
    workload run concurrently, one thread each, repeated over all cores (the reference splits a pair into 48 segments
    over its ME threads instead; the work per pair is the same)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import refbind
    from concurrent.futures import ThreadPoolExecutor
    if refbind.have_reference():
        ref, kind = refbind.Reference(), "reference"
        what = "reference C sources compiled by oracle/Makefile, %s kernels" % ("AVX2" if ref.set_isa(1) == 1 else "C")
    else:
        ref, kind, what = refbind.Oracle(), "port", "oracle restatement (oracle/_ref not built)"
    cores = os.cpu_count() or 1
    need = sorted(set(p for pr in pairs for p in pr))
    padded = {p: refbind.PaddedPicture(frames[p], p) for p in need}
    for p in need:
        ref.build_planes(padded[p], 1)
    jobs = [pairs[i % len(pairs)] for i in range(max(cores, len(pairs)))]
    t0, done = time.perf_counter(), 0
    with ThreadPoolExecutor(cores) as ex:
        while (steps is None and time.perf_counter() - t0 < seconds) or (steps is not None and done < steps * len(jobs)):
            list(ex.map(lambda pr: ref.dg_hme_level0(padded[pr[0]], padded[pr[1]], res), jobs))
            done += len(jobs)
    dt = time.perf_counter() - t0
    return {"value": done / dt, "unit": "picture pairs/s", "cores": cores, "kind": kind,
            "sample": "%d pair searches in %.1f s, %d at a time, %s" % (done, dt, len(jobs), what)}, dt, done


def main_dg(args, rank, local_rank, world):
    """--mode dg: the dynamic-GOP detector's HME level 0 -- per mini-GOP the three picture
    pairs eval_sub_mini_gop compares, each a 128 x 128 full-SAD search of every block's sixteenth-plane 16x16."""
    from svtme import synth
    w, h, gop = args.width, args.height, 1 << args.hier_levels
    res = 0 if w * h <= 352 * 240 else 1 if w * h <= 640 * 360 else 2 if w * h <= 854 * 480 else 3 if w * h <= 1280 * 720 else 4 if w * h <= 1920 * 1080 else 5 if w * h <= 3840 * 2160 else 6
    pairs = DG_PAIRS(gop)
    need = sorted(set(p for pr in pairs for p in pr))
    name = "%dx%d 8-bit synthetic, dynamic-GOP detector HME level 0 (16x16 full SAD, 128x128 window per 64x64 block), pairs (end,start) (end,mid) (mid,start) of a mini-GOP of %d" % (w, h, gop)
    metric, unit = "DG-detector HME-L0 picture-pair searches/sec at 4K", "picture pairs"
    if args.impl == "reference":
        if rank != 0:
            return 0
        clip = synth.make_clip(w, h, gop + 1, seed=100, kind="pan")
        cb, dt, done = dg_cpu_arm(clip, pairs, res, None, steps=max(1, min(args.steps, 4)))
        print(json.dumps({"impl": "reference", "metric": metric, "value": cb["value"], "unit": unit + "/s", "n_gpus": args.gpus,
                          "steps": max(1, min(args.steps, 4)), "warmup": 0, "ms_per_step": 1e3 * dt / max(1, min(args.steps, 4)),
                          "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                          "config": {"workload": name}, "cpu_baseline": cb,
                          "e2e": {"value": cb["value"], "unit": unit + "/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return 0
    import torch
    from svtme import api, shard
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    clip = synth.make_clip(w, h, gop + 1, seed=100, kind="pan", first=rank * gop)
    dev = api.MeDevice(device=local_rank, max_pictures=8)
    h_luma = torch.empty((len(need), h, w), dtype=torch.uint8).pin_memory()
    h_luma.numpy()[:] = np.stack([clip[p] for p in need])
    d_luma = h_luma.cuda()
    slot = {p: i for i, p in enumerate(need)}
    stats = {"px": 0, "ms": 0.0, "n": 0}

    def step(host):
        pics = {p: (dev.register_ptr(h_luma.data_ptr() + slot[p] * w * h, w, w, h, p) if host else
                    dev.register_device(d_luma.data_ptr() + slot[p] * w * h, w, w, h, p)) for p in need}
        for a, b in pairs:
            _, _, _, ms, px = dev.dg_hme_level0(pics[a], pics[b], res, per_block=False)
            stats["px"] += px; stats["ms"] += ms; stats["n"] += 1
        for p in pics.values():
            p.release()

    def timed(host, steps):
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            step(host)
        torch.cuda.synchronize()
        e1.record()
        torch.cuda.synchronize()
        return shard.reduce_job(len(pairs) * steps, e0.elapsed_time(e1), dist, device="cuda")

    for _ in range(max(3, args.warmup)):
        step(False)
    sampler = ClockSampler(local_rank, period=0.002)  # a step is well under a millisecond of device time
    sampler.start()
    time.sleep(0.01)
    l0 = dev.launch_count()
    stats.update(px=0, ms=0.0, n=0)
    units, ms_dev = timed(False, args.steps)
    launches = dev.launch_count() - l0
    kern = dict(stats)
    clocks = sampler.stop()
    step(True)
    _, ms_host = timed(True, args.steps)
    if rank != 0:
        return 0
    sad_peak = dev.measure_sad_peak() * 4 / 1e3
    ach = kern["px"] / (kern["ms"] * 1e-3) / 1e12
    line = {"metric": metric, "value": units / (ms_dev * 1e-3), "unit": unit + "/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": name, "picture_pairs_per_step_per_gpu": len(pairs),
                       "l2": "3 pictures re-uploaded / re-decimated per step; windows are re-read from L2 by design (8.6 G pixel-compares over 0.5 MB planes)"},
            "e2e": {"value": units / (ms_host * 1e-3), "unit": unit + "/s", "h2d_bytes_per_step": len(need) * w * h,
                    "d2h_bytes_per_step": 24 * len(pairs), "ms_per_step": ms_host / args.steps},
            "gpu_launches": launches, "clocks": clocks,
            "roofline": {"bound": "int", "kernel": "dg_hme_kernel (me_dg.cu)", "achieved": ach, "peak": sad_peak, "unit": "Tpx-compares/s",
                         "frac": ach / sad_peak, "traffic": None, "px_compares_per_launch": kern["px"] / kern["n"],
                         "avg_launch_ms": kern["ms"] / kern["n"], "launch": "one picture pair = one kernel, CUDA events on its stream inside the timed region",
                         "peak_source": "VABSDIFF4.U8.ACC issue rate measured in this run (svtme_measure_sad_peak) x 4 px"}}
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"], _, _ = dg_cpu_arm(clip, pairs, res, args.cpu_seconds)
    print(json.dumps(line))
    dev.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


class _DeviceBytes:
    """A device allocation of the library as a __cuda_array_interface__ object (torch.as_tensor wraps it without a copy)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}  # never written through this view


def main_strong(args, rank, local_rank, world, dist, wl):
    """--scaling strong: ONE clip on `world` GPUs.  The pictures of every mini-GOP are dealt round-robin (decode order)
    over the ranks; a picture is uploaded and decimated once, by the rank that searches it (the leading base picture by
    rank 0), and its three padded planes go to the ranks whose searches reference it: svtme_picture_export_planes ->
    NCCL point-to-point over NVLink -> svtme_picture_import_planes (no decimation on the receiver).  No collective runs
    inside the searches.  Before the timed region every rank checks one of its searches against the same search with
    all references decimated locally (the single-GPU path, itself pinned to the oracle by the GPU tests)."""
    import torch
    from svtme import abi, api, shard
    w, h, nf, G = wl.w, wl.h, wl.n_frames, args.gops_per_step
    dev = api.MeDevice(device=local_rank, max_inflight=args.inflight, max_pictures=5 * nf)  # three mini-GOPs alive + one in preparation
    h_luma = torch.empty((nf, h, w), dtype=torch.uint8).pin_memory()
    h_luma.numpy()[:] = np.stack(wl.frames)
    d_luma = h_luma.cuda()
    n_b64 = ((w + 63) // 64) * ((h + 63) // 64)
    keep = []

    def pinned(shape, dt):
        t = torch.zeros(tuple(shape), dtype=getattr(torch, np.dtype(dt).name)).pin_memory()
        keep.append(t)
        return t.numpy()

    # who uploads which picture, and the (picture, source rank, destination rank) plane sets that cross GPUs in a mini-GOP
    owner, transfers = shard.strong_plan(wl.pics, world)
    mine = [(n, e) for n, e in enumerate(wl.pics) if n % world == rank]
    outs = {n: abi.ResultBuffers(n_b64, e[1], extended=False, pinned_alloc=pinned) for n, e in mine}
    nbytes = api.MeDevice.planes_bytes(w, h)
    # two sets of exchange buffers: the planes of mini-GOP g + 1 move while mini-GOP g is searched
    recvbuf = [{t: torch.empty(nbytes, dtype=torch.uint8, device="cuda") for t in transfers if t[2] == rank} for _ in range(2)]
    moved = sum(nbytes for t in transfers if t[2] == rank)
    xstream = torch.cuda.Stream()  # the NCCL point-to-point operations: never on the stream the host synchronises
    imported = [[], []]  # pictures whose planes are still being copied out of recvbuf[slot]

    def prepare(host, slot, exchange=True):
        """Upload + decimate this rank's pictures of a mini-GOP and start moving the plane sets other ranks need (and
        receiving the ones this rank needs).  Returns the state `launch` completes."""
        src = (lambda i: dev.register_ptr(h_luma.data_ptr() + i * w * h, w, w, h, i)) if host else \
              (lambda i: dev.register_device(d_luma.data_ptr() + i * w * h, w, w, h, i))
        pics = {p: src(p) for p in owner if owner[p] == rank}  # uploaded / decimated here
        works, done = [], None
        if exchange and world > 1:
            for q in imported[slot]:  # the previous user of these receive buffers has copied its planes out
                if q.ptr:  # (a released picture has: its release waited for the copy)
                    dev.wait_picture(q)
            imported[slot] = []
            ops, views = [], {}
            for t in transfers:
                if t[1] == rank:  # straight out of the picture's plane allocation (no staging copy); it outlives the send
                    if t[0] not in views:
                        views[t[0]] = torch.as_tensor(_DeviceBytes(*dev.planes_device(pics[t[0]])), device="cuda")
                    ops.append(dist.P2POp(dist.isend, views[t[0]], t[2]))
                elif t[2] == rank:
                    ops.append(dist.P2POp(dist.irecv, recvbuf[slot][t], t[1]))
            if ops:
                with torch.cuda.stream(xstream):
                    works = dist.batch_isend_irecv(ops)
                    for r in works:
                        r.wait()  # orders xstream behind the transfers; the host does not block here
                    done = torch.cuda.Event()
                    done.record(xstream)
        elif not exchange:
            for _, (_, _, l0, l1) in mine:
                for p_ in l0 + l1:
                    if p_ not in pics:
                        pics[p_] = src(p_)
        return pics, done, slot

    def launch(host, state, pending):
        """The received plane sets become pictures (no decimation here), then this rank's searches of the mini-GOP."""
        pics, done, slot = state
        if done is not None:
            done.synchronize()  # the transfers had a whole mini-GOP of searches to complete
            for t in transfers:
                if t[2] == rank and t[0] not in pics:
                    pics[t[0]] = dev.import_planes(recvbuf[slot][t].data_ptr(), w, h, t[0])
                    imported[slot].append(pics[t[0]])
        for n, (poc, k, l0, l1) in mine:
            refs = {(0, i): pics[p_] for i, p_ in enumerate(l0)}
            refs.update({(1, i): pics[p_] for i, p_ in enumerate(l1)})
            if len(pending) >= args.inflight:
                dev.wait(pending.pop(0))
            pending.append(dev.search_async(k, pics[poc], refs, out=outs[n])[0] if host else dev.search_device_async(k, pics[poc], refs))
        return pics

    def gop(host, pending, exchange=True):
        return launch(host, prepare(host, 0, exchange), pending)

    def step(host):
        pending, live = [], []
        state = prepare(host, 0)
        for g in range(G):
            nxt = prepare(host, (g + 1) & 1) if g + 1 < G else None  # mini-GOP g + 1's planes start moving now
            pics = launch(host, state, pending)
            live.append(pics)
            if len(live) > 2:  # its imports have long been copied out of the exchange buffers that are about to be reused
                for q in live.pop(0).values():
                    q.release()
            state = nxt
        for t in pending:
            dev.wait(t)
        for pics in live:
            for q in pics.values():
                q.release()

    # the multi-GPU path against the single-GPU path, on this rank's first picture
    if mine:
        n0 = mine[0][0]
        for exchange in (True, False):
            pend = []
            pics = gop(True, pend, exchange)
            for t in pend:
                dev.wait(t)
            for q in pics.values():
                q.release()
            got = {f: np.array(outs[n0][f]) for f in ("me_mv", "me_cand", "total_me_candidate_index", "me_8x8_distortion")}
            if exchange:
                first = got
            elif any(not np.array_equal(first[f], got[f]) for f in got):
                raise SystemExit("rank %d: the search over replicated reference planes differs from the local one" % rank)

    def timed(host, steps):
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            step(host)
        torch.cuda.synchronize()
        e1.record()
        torch.cuda.synchronize()
        return shard.reduce_job(len(mine) * G * steps, e0.elapsed_time(e1), dist, device="cuda")

    for _ in range(max(3, args.warmup)):
        step(False)
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = dev.launch_count()
    pictures, ms_dev = timed(False, args.steps)
    launches = dev.launch_count() - l0
    clocks = sampler.stop()
    step(True)
    _, ms_host = timed(True, args.steps)
    if rank != 0:
        return 0
    line = {"metric": wl.metric, "value": pictures / (ms_dev * 1e-3), "unit": wl.unit + "/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": config_dict(wl, args, world),
            "e2e": {"value": pictures / (ms_host * 1e-3), "unit": wl.unit + "/s",
                    "h2d_bytes_per_step": sum(1 for p in owner if owner[p] == 0) * w * h * G,
                    "d2h_bytes_per_step": sum(o.nbytes() for o in outs.values()) * G, "ms_per_step": ms_host / args.steps,
                    "nvlink_bytes_received_per_step_rank0": moved * G},
            "gpu_launches": launches, "clocks": clocks,
            "verified": "every rank: one search over replicated planes == the same search with locally decimated references"}
    print(json.dumps(line))
    dev.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.mode == "dg":
        return main_dg(args, rank, local_rank, world)
    if args.impl == "reference":
        if rank != 0:
            return 0
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import refbind
        ref = refbind.Reference()  # the knobs come from the reference's own svt_aom_sig_deriv_me: the product is not loaded
        wl = Workload(args, seed=100, derive=ref.derive_knobs, derive_tf=ref.derive_tf_knobs)
        steps, warmup = max(1, min(args.steps, 40)), max(3, args.warmup)  # a step = one mini-GOP on every host core
        cb, dt, steps = reference_arm(wl, steps, warmup)
        line = {"impl": "reference", "metric": wl.metric, "value": cb["value"], "unit": wl.unit + "/s",
                "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": 1e3 * dt / steps, "higher_is_better": True,
                "scaling": args.scaling, "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": config_dict(wl, args, args.gpus),
                "cpu_baseline": cb, "e2e": {"value": cb["value"], "unit": wl.unit + "/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return 0

    import torch
    from svtme import abi, api, shard
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    # weak scaling: one mini-GOP per rank and step; rank r owns mini-GOP r of the job (svtme/shard.py), its own clip
    my_gops = list(shard.shard_minigops(world, rank, world))
    wl = Workload(args, seed=100, minigop=0 if args.scaling == "strong" else my_gops[0])
    if args.scaling == "strong" and args.mode == "me":
        return main_strong(args, rank, local_rank, world, dist, wl)
    dev = api.MeDevice(device=local_rank, max_inflight=args.inflight, max_pictures=3 * wl.n_frames)
    w, h, nf, G = wl.w, wl.h, wl.n_frames, args.gops_per_step
    h_luma8 = torch.empty((nf, h, w), dtype=torch.uint8).pin_memory()
    h_luma8.numpy()[:] = np.stack(wl.frames)
    d_luma = h_luma8.cuda(non_blocking=False)
    if wl.bit_depth == 10:  # 10-bit samples whose 8-bit MSB plane is the 8-bit clip (svt_unpack_and_2bcompress keeps in16 >> 2)
        h_luma = torch.empty((nf, h, w), dtype=torch.int16).pin_memory()
        rng = np.random.default_rng(7)
        h_luma.numpy().view(np.uint16)[:] = (np.stack(wl.frames).astype(np.uint16) << 2) | rng.integers(0, 4, (nf, h, w), dtype=np.uint16)
    else:
        h_luma = h_luma8
    n_b64 = ((w + 63) // 64) * ((h + 63) // 64)
    keep = []

    def pinned(shape, dt):
        t = torch.zeros(tuple(shape), dtype=getattr(torch, np.dtype(dt).name)).pin_memory()
        keep.append(t)
        return t.numpy()

    # temporal-filter mode: the compact [b64][85] outputs (all the filter reads) instead of the full [2][4][85] tables
    outs = [abi.ResultBuffers(n_b64, k, extended=wl.extended, pinned_alloc=pinned, tf_compact=(args.mode == "tf"))
            for _ in range(2) for _, k, _, _ in wl.pics]  # two sets: consecutive mini-GOPs of a step alternate
    d2h = sum(o.nbytes() for o in outs[:len(wl.pics)]) * G
    tpl_bufs = None
    if wl.tpl:
        tpl_bufs = [abi.TplBuffers(w, h, alloc=pinned) for _ in range(2) for _ in wl.pics]
        d2h += sum(t.nbytes() for t in tpl_bufs[:len(wl.pics)]) * G
    h2d = len(wl.needed) * w * h * (2 if wl.bit_depth == 10 else 1) * G

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()  # every step ends with all of its searches waited for
        torch.cuda.synchronize()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        barrier()
        # whole job: SUM of the pictures of all ranks, MAX of their device times (svtme/shard.py)
        return shard.reduce_job(wl.units * G * steps, ms, dist, device="cuda")

    # the picture consecutive mini-GOPs share (last of rank r = leading base picture of rank r + 1) moves GPU to GPU
    # every step: NCCL point-to-point over NVLink; the receiver decimates it locally
    d_recv = torch.empty((h, w), dtype=torch.uint8, device="cuda") if world > 1 else None

    def handoff():
        got = shard.handoff_boundary(dist, d_luma[nf - 1], d_recv)
        torch.cuda.current_stream().synchronize()  # the library's upload streams read the buffer next
        return d_recv.data_ptr() if got else None

    hand = handoff if world > 1 else None
    step_dev = lambda: run_step_device(dev, wl, d_luma, args.inflight, G, hand, tpl_bufs)  # noqa: E731
    step_host = lambda: run_step_host(dev, wl, h_luma, outs, args.inflight, G, hand, tpl_bufs)  # noqa: E731
    if world > 1:
        handoff()
        if rank > 0 and not torch.equal(d_recv, d_luma[0]):
            raise SystemExit("rank %d: the base picture received from rank %d differs from the clip's" % (rank, rank - 1))

    for _ in range(max(3, args.warmup)):
        step_dev()
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = dev.launch_count()
    pictures, ms_dev = timed(step_dev, args.steps)
    launches = dev.launch_count() - l0
    clocks = sampler.stop()
    step_host()
    _, ms_host = timed(step_host, args.steps)
    # context for the end-to-end number: what the host link alone sustains for one step's uploads (pinned -> HBM)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    scratch = torch.empty_like(h_luma, device="cuda")
    scratch.copy_(h_luma, non_blocking=True)
    torch.cuda.synchronize()
    e0.record()
    scratch.copy_(h_luma, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    h2d_only_ms = e0.elapsed_time(e1) * len(wl.needed) / nf * G  # scaled to the pictures a step uploads
    del scratch

    # per-launch view: one picture's search at a time, CUDA events on its stream around it (svtme_search_elapsed_ms)
    px_total, ms_total, n_launch = 0, 0.0, 0
    pics = {i: dev.register_device(d_luma.data_ptr() + i * w * h, w, w, h, i) for i in range(nf)}
    for rep in range(2):
        for poc, k, l0_, l1_ in wl.pics:
            refs = {(0, i): pics[p] for i, p in enumerate(l0_)}
            refs.update({(1, i): pics[p] for i, p in enumerate(l1_)})
            t = dev.search_device_async(k, pics[poc], refs)
            dev.wait(t)
            if rep == 1:
                px_total += dev.px_compares(t)
                ms_total += dev.elapsed_ms(t)
                n_launch += 1
    for p in pics.values():
        p.release()

    value = pictures / (ms_dev * 1e-3)
    e2e = pictures / (ms_host * 1e-3)
    if rank != 0:
        return 0

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak, hbm_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback")
    sad_peak = dev.measure_sad_peak() * 4 / 1e3  # T pixel-compares/s
    refs_avg = sum(len(l0_) + len(l1_) for _, _, l0_, l1_ in wl.pics) / len(wl.pics)
    # compulsory HBM bytes of one search launch: source + every reference, three planes each (1.3125 B/px) + results
    alg_bytes = (1 + refs_avg) * 1.3125 * w * h + d2h / len(wl.pics)
    traffic = None
    try:
        if args.mode == "me":  # the capture is of the headline workload
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))["search_dram_bytes_per_picture"]
    except Exception:
        pass
    # achieved = algorithmic pixel-compares of the step / device time of the step, over the TIMED region (CUDA events,
    # max over ranks), per GPU: the searches of `inflight` pictures overlap there, so the grid tails and the scalar control
    # kernels of one picture are filled by the engines of another.  "serialised" is the same work over one picture's ~21
    # launches run alone (events on its own stream) -- the latency view, and what the ncu launch list adds up to.
    ach_ser = px_total / (ms_total * 1e-3) / 1e12
    ach = (px_total / n_launch) * (value / world) / 1e12
    roofline = {"bound": "int", "kernel": "picture search = st_init / st_light / st_sad16 / st_sad_rows / st_fp_warp / st_fp_sub kernels (me_staged.cu)", "achieved": ach, "peak": sad_peak, "unit": "Tpx-compares/s",
                "frac": ach / sad_peak, "traffic": traffic,
                "peak_source": "VABSDIFF4.U8.ACC issue rate measured in this run (svtme_measure_sad_peak) x 4 px",
                "px_compares_per_launch": px_total / n_launch, "avg_launch_ms": ms_dev / pictures * world,
                "launch": "one picture's search (14 stream-ordered kernels at preset M6); duration = device time of the timed region / pictures searched in it (%d searches in flight)" % args.inflight,
                "serialised": {"achieved": ach_ser, "frac": ach_ser / sad_peak, "avg_launch_ms": ms_total / n_launch,
                               "launch": "the same search launched alone, CUDA events on its stream"},
                "hbm": {"achieved": alg_bytes / (ms_dev / pictures * world * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": alg_bytes / (ms_dev / pictures * world * 1e-3) / 1e9 / hbm_peak, "algorithmic_bytes_per_launch": alg_bytes,
                        "peak_source": hbm_src}}
    line = {"metric": wl.metric, "value": value, "unit": wl.unit + "/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": config_dict(wl, args, world),
            "e2e": {"value": e2e, "unit": wl.unit + "/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_host / args.steps, "h2d_copy_alone_ms_per_step": h2d_only_ms,
                    "h2d_link_gbs": h2d / h2d_only_ms / 1e6},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline}
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"], _, _ = reference_arm(wl, 1, 1, seconds=args.cpu_seconds)
    print(json.dumps(line))
    dev.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
