"""Seeded synthetic 8-bit luma clips: a smooth-noise canvas panned by a per-frame global
integer MV plus independently moving patches, so that HME levels, pruning, SR adjustment and the 8x8-variance
branches all fire; plus adversarial clips (all-zero, white noise, static)."""
import numpy as np


def _smooth_canvas(rng, h, w):
    # low-pass a coarse random field by bilinear upsampling, then add 8 % white noise
    ch, cw = h // 16 + 2, w // 16 + 2
    coarse = rng.integers(0, 256, (ch, cw)).astype(np.float32)
    ys = np.linspace(0, ch - 1.001, h, dtype=np.float32)
    xs = np.linspace(0, cw - 1.001, w, dtype=np.float32)
    y0, x0 = ys.astype(np.int32), xs.astype(np.int32)
    fy, fx = (ys - y0)[:, None], (xs - x0)[None, :]
    a = coarse[y0][:, x0]
    b = coarse[y0][:, x0 + 1]
    c = coarse[y0 + 1][:, x0]
    d = coarse[y0 + 1][:, x0 + 1]
    img = (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy
    img += rng.normal(0, 0.08 * 255 / 3, (h, w)).astype(np.float32)
    return np.clip(img, 0, 255).astype(np.uint8)


def make_clip(width, height, n_frames, seed=0, kind="pan", max_mv=48, n_patches=16, first=0):
    """Returns a list of (height, width) uint8 arrays: frames first .. first + n_frames - 1 of the clip `seed`
    (a clip is one deterministic sequence; `first` lets every rank of a multi-GPU job cut its own mini-GOPs out of it)."""
    if first:
        return make_clip(width, height, first + n_frames, seed, kind, max_mv, n_patches)[first:]
    rng = np.random.default_rng(seed)
    if kind == "zero":
        return [np.zeros((height, width), np.uint8) for _ in range(n_frames)]
    if kind == "noise":
        return [rng.integers(0, 256, (height, width), dtype=np.uint8) for _ in range(n_frames)]
    margin = max_mv * 2 + 64
    canvas = _smooth_canvas(rng, height + 2 * margin, width + 2 * margin)
    if kind == "static":
        f = canvas[margin:margin + height, margin:margin + width].copy()
        return [f.copy() for _ in range(n_frames)]
    psz = min(256, height // 4 * 2, width // 4 * 2)
    patches = [_smooth_canvas(rng, psz, psz) for _ in range(n_patches)]
    ppos = [(int(rng.integers(0, height - psz)), int(rng.integers(0, width - psz))) for _ in range(n_patches)]
    pvel = [(int(rng.integers(-6, 7)), int(rng.integers(-6, 7))) for _ in range(n_patches)]
    frames = []
    oy = ox = 0
    for t in range(n_frames):
        dy, dx = (int(v) for v in rng.integers(-max_mv // 8, max_mv // 8 + 1, 2))
        oy = int(np.clip(oy + dy, -max_mv, max_mv))
        ox = int(np.clip(ox + dx, -max_mv, max_mv))
        f = canvas[margin + oy:margin + oy + height, margin + ox:margin + ox + width].copy()
        for i, pt in enumerate(patches):
            y = (ppos[i][0] + pvel[i][0] * t) % (height - psz)
            x = (ppos[i][1] + pvel[i][1] * t) % (width - psz)
            f[y:y + psz, x:x + psz] = pt
        frames.append(f)
    return frames


def minigop_structure(n_frames, hierarchical_levels, max_l0, max_l1):
    """A random-access hierarchical-B prediction structure in display order: for every non-intra
    picture its temporal layer and the display-order picture numbers of its list-0 (past) and list-1 (future)
    references.  Returns a list of dicts for pictures 1..n-1 in decode order."""
    gop = 1 << hierarchical_levels
    out = []
    decoded = {0}

    def layer_of(poc):
        if poc % gop == 0:
            return 0
        l = hierarchical_levels
        p = poc
        while p % 2 == 0:
            p //= 2
            l -= 1
        return l

    def visit(lo, hi, depth):
        if hi - lo < 2:
            return
        mid = (lo + hi) // 2
        order.append(mid)
        visit(lo, mid, depth + 1)
        visit(mid, hi, depth + 1)

    base = 0
    while base + gop <= n_frames - 1 or base == 0 and n_frames > 1:
        end = min(base + gop, n_frames - 1)
        order = [end]
        visit(base, end, 1)
        for poc in order:
            tl = layer_of(poc) if end - base == gop else min(hierarchical_levels, layer_of(poc))
            if tl == 0:
                past = [p for p in range(poc - gop, -1, -gop)] or [base]
                fut = list(past)  # base-layer list 1 repeats past pictures (pd_process.c: no future base picture exists yet)
            else:
                # references = reference pictures already decoded (decode order), nearest first
                lo = max(base - gop, 0)
                past = [p for p in range(poc - 1, lo - 1, -1) if p in decoded]
                fut = [p for p in range(poc + 1, end + 1) if p in decoded]
            l0 = past[:max_l0] or [base]
            l1 = fut[:max_l1]
            out.append({"poc": poc, "temporal_layer": tl, "list0": l0, "list1": l1,
                        "is_ref": 1 if tl < hierarchical_levels else 0})
            if tl < hierarchical_levels:
                decoded.add(poc)
        base = end
        if end == n_frames - 1:
            break
    return out
