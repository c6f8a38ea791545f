"""TEST INFRASTRUCTURE ONLY: ctypes binding of oracle/_ref/libsvtref.so (the compiled reference) and of
oracle/libsvtme_oracle.so (our C restatement).  Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
legs may import this module."""
import ctypes as C
import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
_pkg = os.path.join(ROOT, "svt-av1-mirror_b200")
if _pkg not in sys.path:
    sys.path.insert(0, _pkg)
from svtme import abi  # noqa: E402


class RefPlane(C.Structure):
    _fields_ = [("buf", C.c_void_p), ("stride", C.c_uint32), ("width", C.c_uint32), ("height", C.c_uint32),
                ("org_x", C.c_uint32), ("org_y", C.c_uint32)]


class RefPic(C.Structure):
    _fields_ = [("lvl", RefPlane * 3), ("picture_number", C.c_uint64)]


# Reference geometry: full-res pad 72 (enc_handle.c:4057-4070), quarter 32, sixteenth 16 (reference_object.c:267-295)
PADS = (72, 32, 16)


class PaddedPicture:
    """Three padded luma planes laid out as the reference's EbPaReferenceObject has them."""

    def __init__(self, luma, picture_number):
        h, w = luma.shape
        assert w % 8 == 0 and h % 8 == 0
        self.width, self.height, self.picture_number = w, h, picture_number
        self.planes = []
        for lvl, pad in enumerate(PADS):
            pw, ph = w >> lvl, h >> lvl
            self.planes.append(np.zeros((ph + 2 * pad, pw + 2 * pad), np.uint8))
        p = PADS[0]
        self.planes[0][p:p + h, p:p + w] = luma
        self.struct = RefPic()
        self.struct.picture_number = picture_number
        for lvl, pad in enumerate(PADS):
            pl = self.struct.lvl[lvl]
            pl.buf = self.planes[lvl].ctypes.data
            pl.stride = self.planes[lvl].shape[1]
            pl.width, pl.height = w >> lvl, h >> lvl
            pl.org_x = pl.org_y = pad

    def interior(self, lvl):
        pad = PADS[lvl]
        return self.planes[lvl][pad:pad + (self.height >> lvl), pad:pad + (self.width >> lvl)]


def _load(path, what):
    if not os.path.exists(path):
        raise RuntimeError("%s not built: %s (run `make -C oracle`)" % (what, path))
    return C.CDLL(path)


class _Binding:
    """Common call surface of the compiled reference (prefix svtref_) and the oracle restatement (svtorc_)."""

    def __init__(self, lib, prefix):
        self.lib, self.prefix = lib, prefix
        f = self._f
        f("derive_knobs").argtypes = [C.POINTER(abi.PresetCfg), C.POINTER(abi.Knobs)]
        if hasattr(lib, prefix + "derive_tf_knobs"):
            f("derive_tf_knobs").argtypes = [C.POINTER(abi.TfCfg), C.POINTER(abi.Knobs)]
        f("me_picture").argtypes = [C.POINTER(abi.Knobs), C.POINTER(RefPic), C.POINTER(RefPic),
                                    C.POINTER(abi.PictureResults), C.c_int]
        f("build_planes").argtypes = [C.POINTER(RefPic), C.c_int]
        f("build_planes").restype = None
        f("sad_loop_kernel").argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                         C.POINTER(C.c_uint64), C.POINTER(C.c_int16), C.POINTER(C.c_int16),
                                         C.c_uint32, C.c_uint8, C.c_int16, C.c_int16]
        f("sad_loop_kernel").restype = None
        f("nxm_sad_kernel").argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]
        f("nxm_sad_kernel").restype = C.c_uint32
        f("downsample_2d").argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32,
                                       C.c_uint32]
        f("downsample_2d").restype = None
        f("generate_padding").argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        f("generate_padding").restype = None
        f("fullpel_search_sblock").argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_int16, C.c_int16,
                                               C.c_uint32, C.c_uint32, C.c_uint8, C.c_void_p, C.c_void_p]
        f("fullpel_search_sblock").restype = None

    def _f(self, name):
        return getattr(self.lib, self.prefix + name)

    def derive_knobs(self, **kw):
        cfg = abi.PresetCfg()
        for k, v in kw.items():
            setattr(cfg, k, v)
        k = abi.Knobs()
        rc = self._f("derive_knobs")(C.byref(cfg), C.byref(k))
        if rc != 0:
            raise RuntimeError("derive_knobs failed: %d" % rc)
        return k

    def derive_tf_knobs(self, **kw):
        cfg = abi.TfCfg()
        for k, v in kw.items():
            setattr(cfg, k, v)
        k = abi.Knobs()
        rc = self._f("derive_tf_knobs")(C.byref(cfg), C.byref(k))
        if rc != 0:
            raise RuntimeError("derive_tf_knobs failed: %d" % rc)
        return k

    def build_planes(self, pic, hme_level1=1):
        self._f("build_planes")(C.byref(pic.struct), hme_level1)

    def me_picture(self, knobs, cur, refs, n_threads=1, extended=True):
        """refs: {(list, idx): PaddedPicture}. Returns abi.ResultBuffers."""
        n_b64 = ((cur.width + 63) // 64) * ((cur.height + 63) // 64)
        out = abi.ResultBuffers(n_b64, knobs, extended=extended)
        arr = (RefPic * 8)()
        for (li, ri), p in refs.items():
            arr[li * 4 + ri] = p.struct
        rc = self._f("me_picture")(C.byref(knobs), C.byref(cur.struct), arr, C.byref(out.struct), n_threads)
        if rc != 0:
            raise RuntimeError("me_picture failed: %d" % rc)
        out._keep = (arr, refs, cur)
        return out

    def dg_hme_level0(self, src, ref, input_resolution, per_block=True):
        """dg_detector_hme_level0 over the whole picture: (metrics dict, sb_sad, sb_mv) -- the per-block arrays only from
        the oracle restatement (the reference keeps the metrics only)."""
        m = abi.DgMetrics()
        n_b64 = ((src.width + 63) // 64) * ((src.height + 63) // 64)
        if self.prefix == "svtorc_":
            sad = np.zeros(n_b64, np.uint32)
            mv = np.zeros((n_b64, 2), np.int16)
            fn = self._f("dg_hme_level0")
            fn.argtypes = [C.POINTER(RefPic), C.POINTER(RefPic), C.c_uint32, C.POINTER(abi.DgMetrics), C.c_void_p, C.c_void_p]
            rc = fn(C.byref(src.struct), C.byref(ref.struct), input_resolution, C.byref(m), sad.ctypes.data, mv.ctypes.data)
        else:
            sad = mv = None
            fn = self._f("dg_hme_level0")
            fn.argtypes = [C.POINTER(RefPic), C.POINTER(RefPic), C.c_uint32, C.POINTER(abi.DgMetrics)]
            rc = fn(C.byref(src.struct), C.byref(ref.struct), input_resolution, C.byref(m))
        if rc != 0:
            raise RuntimeError("dg_hme_level0 failed: %d" % rc)
        return m.as_dict(), sad, mv

    def picture_statistics(self, pic, regions_w=4, regions_h=4, decim_step=4):
        """(histogram [4][4][256], average_intensity_per_region [4][4], avg_luma) of the sixteenth plane."""
        st = abi.PicStats()
        fn = self._f("picture_statistics")
        fn.argtypes = [C.POINTER(RefPic), C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(abi.PicStats)]
        rc = fn(C.byref(pic.struct), regions_w, regions_h, decim_step, C.byref(st))
        if rc != 0:
            raise RuntimeError("picture_statistics failed: %d" % rc)
        return st.as_arrays()

    def sad_loop_kernel(self, src, src_off, src_stride, ref, ref_off, ref_stride, bh, bw, raw_stride, skip, sa_w, sa_h):
        best = C.c_uint64(0)
        x = C.c_int16(-12345)
        y = C.c_int16(-12345)
        self._f("sad_loop_kernel")(src.ctypes.data + src_off, src_stride, ref.ctypes.data + ref_off, ref_stride, bh, bw,
                                   C.byref(best), C.byref(x), C.byref(y), raw_stride, skip, sa_w, sa_h)
        return best.value, x.value, y.value


class Reference(_Binding):
    def __init__(self):
        super().__init__(_load(os.path.join(HERE, "_ref", "libsvtref.so"), "compiled reference"), "svtref_")
        self.lib.svtref_set_isa.argtypes = [C.c_int]
        if hasattr(self.lib, "svtref_me_minigop"):
            pp = C.POINTER(C.c_void_p)
            self.lib.svtref_me_minigop.argtypes = [C.c_int, pp, pp, pp, pp, C.c_int, pp, C.c_int, C.c_int]

    def set_isa(self, isa):
        """0: C kernels, 1: AVX2 twins, 2: AVX2 + the AVX-512 svt_sad_loop_kernel where the host has it (what the
        encoder's RTCD would bind).  Returns what was bound."""
        return self.lib.svtref_set_isa(isa)

    def me_minigop(self, items, build, n_threads, hme_level1=1, extended=False, outs=None):
        """The encoder's parallelisation of the path (oracle/ref_harness.c: svtref_me_minigop): `build` pictures are
        decimated + padded, then `items` = [(knobs, cur, {(list, idx): PaddedPicture})] are searched, all on one pool of
        `n_threads` ME threads pulling (picture, 6 x 8 segment) items.  Returns the result buffers."""
        n = len(items)
        if outs is None:
            outs = [abi.ResultBuffers(((c.width + 63) // 64) * ((c.height + 63) // 64), k, extended=extended) for k, c, _ in items]
        arrs = []
        for _, _, refs in items:
            arr = (RefPic * 8)()
            for (li, ri), p in refs.items():
                arr[li * 4 + ri] = p.struct
            arrs.append(arr)
        vp = C.c_void_p
        ks = (vp * n)(*[C.addressof(k) for k, _, _ in items])
        cs = (vp * n)(*[C.addressof(c.struct) for _, c, _ in items])
        rs = (vp * n)(*[C.addressof(a) for a in arrs])
        os_ = (vp * n)(*[C.addressof(o.struct) for o in outs])
        bs = (vp * max(1, len(build)))(*[C.addressof(b.struct) for b in build])
        rc = self.lib.svtref_me_minigop(n, ks, cs, rs, os_, len(build), bs, hme_level1, n_threads)
        if rc != 0:
            raise RuntimeError("me_minigop failed: %d" % rc)
        for o, a in zip(outs, arrs):
            o._keep = a
        return outs


class Oracle(_Binding):
    def __init__(self):
        super().__init__(_load(os.path.join(HERE, "libsvtme_oracle.so"), "oracle restatement"), "svtorc_")
        self.lib.svtorc_last_px_compares.restype = C.c_uint64

    def tpl_src_search(self, cfg, knobs, cur, refs, me):
        """cfg: abi.TplCfg, me: abi.ResultBuffers of this picture's ME (None for an I picture).  Returns abi.TplBuffers."""
        f = self.lib.svtorc_tpl_src_search
        f.argtypes = [C.POINTER(abi.TplCfg), C.POINTER(abi.Knobs), C.POINTER(RefPic), C.POINTER(RefPic),
                      C.POINTER(abi.PictureResults), C.POINTER(abi.TplResults)]
        out = abi.TplBuffers(cur.width, cur.height)
        arr = (RefPic * 8)()
        for (li, ri), p in refs.items():
            arr[li * 4 + ri] = p.struct
        if me is None:
            n_b64 = ((cur.width + 63) // 64) * ((cur.height + 63) // 64)
            me = abi.ResultBuffers(n_b64, knobs, extended=False)
        rc = f(C.byref(cfg), C.byref(knobs), C.byref(cur.struct), arr, C.byref(me.struct), C.byref(out.struct))
        if rc != 0:
            raise RuntimeError("tpl_src_search failed: %d" % rc)
        return out

    def last_px_compares(self):
        """Pixel-compares of the last me_picture call (the algorithmic work C_alg of SURVEY.md section 8d)."""
        return int(self.lib.svtorc_last_px_compares())


def have_reference():
    return os.path.exists(os.path.join(HERE, "_ref", "libsvtref.so"))
