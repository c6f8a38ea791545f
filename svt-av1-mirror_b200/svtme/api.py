"""Host-side mirror of the C ABI (include/svtme_cuda.h) over ctypes.

The product is libsvtme_cuda.so; this module only loads it and wraps the calls the way the reference's ME kernel
thread uses its own entry points: register pictures (svt_aom_picture_analysis_kernel's decimation), search a picture
against its reference lists (svt_aom_motion_estimation_kernel's SB loop), read the results.  There is no CPU path:
if the library is missing or no CUDA device is present, construction raises.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SVTME_LIB") or os.path.join(_HERE, "libsvtme_cuda.so")  # SVTME_LIB: tuning variants
_PKG_DIR = os.path.dirname(_HERE)

EXPORTS = [
    "svtme_init", "svtme_destroy", "svtme_derive_knobs", "svtme_derive_tf_knobs", "svtme_picture_register", "svtme_picture_register16",
    "svtme_picture_register_device", "svtme_picture_wait", "svtme_picture_read_plane", "svtme_picture_release",
    "svtme_picture_search", "svtme_search_wait", "svtme_picture_search_device", "svtme_search_elapsed_ms",
    "svtme_search_px_compares", "svtme_measure_sad_peak", "svtme_launch_count", "svtme_last_error", "svtme_cuda_sad_loop_kernel", "svtme_cuda_downsample_2d",
    "svtme_cuda_nxm_sad_kernel", "svtme_cuda_fullpel_search_sblock", "svtme_dg_detector_hme_level0",
    "svtme_picture_statistics", "svtme_cuda_nxm_sad_kernel_noalloc", "svtme_picture_planes_bytes", "svtme_picture_replicate",
    "svtme_picture_export_planes", "svtme_picture_import_planes", "svtme_picture_search_tpl", "svtme_tpl_src_search", "svtme_picture_planes_device",
]


class SvtMeError(RuntimeError):
    pass


def build(force=False):
    """Compile libsvtme_cuda.so in-tree with nvcc for sm_100a (svt-av1-mirror_b200/Makefile)."""
    if force and os.path.exists(LIB_PATH):
        os.remove(LIB_PATH)
    subprocess.check_call(["make", "-C", _PKG_DIR], stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def load_library():
    """dlopen the C-ABI library and declare its prototypes.  Fails loudly when it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise SvtMeError("libsvtme_cuda.so is not built (run `make -C svt-av1-mirror_b200`); there is no CPU fallback")
    lib = C.CDLL(LIB_PATH)
    vp, u8p, u32, u64, i32 = C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_int32
    refs_t = C.POINTER(vp * 8)
    lib.svtme_init.argtypes = [C.POINTER(abi.DeviceCfg), C.POINTER(vp)]
    lib.svtme_destroy.argtypes = [vp]
    lib.svtme_destroy.restype = None
    lib.svtme_derive_knobs.argtypes = [C.POINTER(abi.PresetCfg), C.POINTER(abi.Knobs)]
    lib.svtme_derive_tf_knobs.argtypes = [C.POINTER(abi.TfCfg), C.POINTER(abi.Knobs)]
    lib.svtme_picture_register.argtypes = [vp, u64, u8p, u32, u32, u32, C.POINTER(vp)]
    lib.svtme_picture_register16.argtypes = [vp, u64, u8p, u32, u32, u32, C.POINTER(vp)]
    lib.svtme_picture_register_device.argtypes = [vp, u64, vp, u32, u32, u32, C.POINTER(vp)]
    lib.svtme_picture_wait.argtypes = [vp, vp]
    lib.svtme_picture_planes_bytes.argtypes = [u32, u32, C.POINTER(u64)]
    lib.svtme_picture_replicate.argtypes = [vp, vp, vp, C.POINTER(vp)]
    lib.svtme_picture_export_planes.argtypes = [vp, vp, vp, u64]
    lib.svtme_picture_import_planes.argtypes = [vp, u64, vp, u32, u32, C.POINTER(vp)]
    lib.svtme_picture_planes_device.argtypes = [vp, vp, C.POINTER(vp), C.POINTER(u64)]
    lib.svtme_picture_read_plane.argtypes = [vp, vp, i32, u32, u8p, u32]
    lib.svtme_picture_release.argtypes = [vp, vp]
    lib.svtme_picture_search.argtypes = [vp, vp, refs_t, C.POINTER(abi.Knobs), C.POINTER(abi.PictureResults),
                                         C.POINTER(u64)]
    lib.svtme_picture_search_device.argtypes = [vp, vp, refs_t, C.POINTER(abi.Knobs), C.POINTER(u64)]
    lib.svtme_search_wait.argtypes = [vp, u64]
    lib.svtme_search_elapsed_ms.argtypes = [vp, u64, C.POINTER(C.c_float)]
    lib.svtme_search_px_compares.argtypes = [vp, u64, C.POINTER(u64)]
    lib.svtme_measure_sad_peak.argtypes = [vp, C.POINTER(C.c_double)]
    lib.svtme_launch_count.argtypes = [vp]
    lib.svtme_launch_count.restype = u64
    lib.svtme_last_error.argtypes = [vp]
    lib.svtme_last_error.restype = C.c_char_p
    lib.svtme_cuda_sad_loop_kernel.argtypes = [vp, u8p, u32, u8p, u32, u32, u32, C.POINTER(u64), C.POINTER(C.c_int16),
                                               C.POINTER(C.c_int16), u32, C.c_uint8, C.c_int16, C.c_int16]
    lib.svtme_cuda_sad_loop_kernel.restype = None
    lib.svtme_cuda_downsample_2d.argtypes = [vp, u8p, u32, u32, u32, u8p, u32, u32]
    lib.svtme_cuda_downsample_2d.restype = None
    lib.svtme_cuda_nxm_sad_kernel.argtypes = [vp, u8p, u32, u8p, u32, u32, u32]
    lib.svtme_cuda_nxm_sad_kernel.restype = u32
    lib.svtme_cuda_nxm_sad_kernel_noalloc.argtypes = [vp, u8p, u32, u8p, u32, u32, u32, C.POINTER(u32)]
    lib.svtme_cuda_fullpel_search_sblock.argtypes = [vp, u8p, u32, u8p, u32, C.c_int16, C.c_int16, u32, u32, C.c_uint8,
                                                     vp, vp]
    lib.svtme_cuda_fullpel_search_sblock.restype = None
    lib.svtme_dg_detector_hme_level0.argtypes = [vp, vp, vp, u32, C.POINTER(abi.DgMetrics), vp, vp, C.POINTER(C.c_float),
                                                 C.POINTER(u64)]
    lib.svtme_picture_statistics.argtypes = [vp, vp, u32, u32, u32, C.POINTER(abi.PicStats)]
    lib.svtme_picture_search_tpl.argtypes = [vp, vp, refs_t, C.POINTER(abi.Knobs), C.POINTER(abi.PictureResults),
                                             C.POINTER(abi.TplCfg), C.POINTER(abi.TplResults), C.POINTER(u64)]
    lib.svtme_tpl_src_search.argtypes = [vp, vp, refs_t, C.POINTER(abi.Knobs), C.POINTER(abi.PictureResults),
                                         C.POINTER(abi.TplCfg), C.POINTER(abi.TplResults), C.POINTER(u64)]
    _lib = lib
    return lib


def derive_knobs(**kw):
    """svtme_derive_knobs: preset-level configuration -> SvtMeKnobs (pure host code; works without a GPU)."""
    cfg = abi.PresetCfg()
    for k, v in kw.items():
        setattr(cfg, k, v)
    out = abi.Knobs()
    rc = load_library().svtme_derive_knobs(C.byref(cfg), C.byref(out))
    if rc != 0:
        raise SvtMeError("svtme_derive_knobs: 0x%x" % (rc & 0xffffffff))
    return out


def derive_tf_knobs(**kw):
    """svtme_derive_tf_knobs: TfControls fields -> SvtMeKnobs of one temporal-filter search (pure host code)."""
    cfg = abi.TfCfg()
    for k, v in kw.items():
        setattr(cfg, k, v)
    out = abi.Knobs()
    rc = load_library().svtme_derive_tf_knobs(C.byref(cfg), C.byref(out))
    if rc != 0:
        raise SvtMeError("svtme_derive_tf_knobs: 0x%x" % (rc & 0xffffffff))
    return out


class Picture:
    """Handle of one registered picture (SvtMePic*)."""

    def __init__(self, ctx, ptr, picture_number, width, height, keep=None):
        self.ctx, self.ptr, self.picture_number, self.width, self.height = ctx, ptr, picture_number, width, height
        self._keep = keep  # host buffer of an asynchronous upload

    def release(self):
        if self.ptr:
            self.ctx._check(self.ctx.lib.svtme_picture_release(self.ctx.h, self.ptr), "svtme_picture_release")
            self.ptr = None


class MeDevice:
    """One library handle bound to one CUDA device (SvtMeHandle*)."""

    def __init__(self, device=0, max_width=0, max_height=0, max_pictures=0, max_inflight=0, sixteenth_from_full=0,
                 pipeline=None):
        self.lib = load_library()
        if pipeline is None:
            pipeline = 0  # the staged pipeline is the only one that ships
        cfg = abi.DeviceCfg(device, max_width, max_height, max_pictures, max_inflight, sixteenth_from_full, pipeline)
        h = C.c_void_p()
        rc = self.lib.svtme_init(C.byref(cfg), C.byref(h))
        if rc != 0:
            raise SvtMeError("svtme_init failed (0x%x): no usable CUDA device; there is no CPU fallback" % (rc & 0xffffffff))
        self.h = h
        self._pending = {}

    def close(self):
        if self.h:
            self.lib.svtme_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise SvtMeError("%s failed (0x%x): %s" % (what, rc & 0xffffffff, self.lib.svtme_last_error(self.h).decode()))

    # ---- pictures ----
    def register(self, luma, picture_number):
        """Upload an (H, W) uint8 luma plane (or uint16 10-bit) and build its padded full/quarter/sixteenth planes."""
        a = np.ascontiguousarray(luma) if isinstance(luma, np.ndarray) else luma
        h, w = a.shape
        ptr = C.c_void_p()
        if a.dtype == np.uint16:
            rc = self.lib.svtme_picture_register16(self.h, picture_number, a.ctypes.data, a.strides[0] // 2, w, h, C.byref(ptr))
        else:
            assert a.dtype == np.uint8
            rc = self.lib.svtme_picture_register(self.h, picture_number, a.ctypes.data, a.strides[0], w, h, C.byref(ptr))
        self._check(rc, "svtme_picture_register")
        return Picture(self, ptr, picture_number, w, h, keep=a)

    def register_ptr(self, host_ptr, stride, width, height, picture_number, keep=None):
        ptr = C.c_void_p()
        self._check(self.lib.svtme_picture_register(self.h, picture_number, host_ptr, stride, width, height, C.byref(ptr)),
                    "svtme_picture_register")
        return Picture(self, ptr, picture_number, width, height, keep=keep)

    def register_ptr16(self, host_ptr, stride, width, height, picture_number, keep=None):
        """10-bit samples in uint16 at `host_ptr` (stride in samples): the 8-bit MSB plane is what ME sees."""
        ptr = C.c_void_p()
        self._check(self.lib.svtme_picture_register16(self.h, picture_number, host_ptr, stride, width, height, C.byref(ptr)),
                    "svtme_picture_register16")
        return Picture(self, ptr, picture_number, width, height, keep=keep)

    def register_device(self, dev_ptr, stride, width, height, picture_number):
        ptr = C.c_void_p()
        self._check(self.lib.svtme_picture_register_device(self.h, picture_number, dev_ptr, stride, width, height, C.byref(ptr)),
                    "svtme_picture_register_device")
        return Picture(self, ptr, picture_number, width, height)

    # ---- reference planes across GPUs ----
    @staticmethod
    def planes_bytes(width, height):
        n = C.c_uint64(0)
        if load_library().svtme_picture_planes_bytes(width, height, C.byref(n)) != 0:
            raise SvtMeError("svtme_picture_planes_bytes: bad size")
        return n.value

    def replicate_to(self, pic, dst):
        """Copy `pic` (of this handle) into a new picture of handle `dst` (another GPU of this process): peer copy."""
        ptr = C.c_void_p()
        dst._check(self.lib.svtme_picture_replicate(self.h, pic.ptr, dst.h, C.byref(ptr)), "svtme_picture_replicate")
        return Picture(dst, ptr, pic.picture_number, pic.width, pic.height)

    def export_planes(self, pic, d_dst, nbytes):
        self._check(self.lib.svtme_picture_export_planes(self.h, pic.ptr, d_dst, nbytes), "svtme_picture_export_planes")

    def planes_device(self, pic):
        """(device address, bytes) of the picture's plane allocation, once built: a send buffer without a copy."""
        ptr, n = C.c_void_p(), C.c_uint64()
        self._check(self.lib.svtme_picture_planes_device(self.h, pic.ptr, C.byref(ptr), C.byref(n)), "svtme_picture_planes_device")
        return ptr.value, n.value

    def import_planes(self, d_src, width, height, picture_number, keep=None):
        ptr = C.c_void_p()
        self._check(self.lib.svtme_picture_import_planes(self.h, picture_number, d_src, width, height, C.byref(ptr)),
                    "svtme_picture_import_planes")
        return Picture(self, ptr, picture_number, width, height, keep=keep)

    def wait_picture(self, pic):
        self._check(self.lib.svtme_picture_wait(self.h, pic.ptr), "svtme_picture_wait")

    def read_plane(self, pic, level, pad):
        w, h = pic.width >> level, pic.height >> level
        out = np.zeros((h + 2 * pad, w + 2 * pad), np.uint8)
        self._check(self.lib.svtme_picture_read_plane(self.h, pic.ptr, level, pad, out.ctypes.data, out.strides[0]),
                    "svtme_picture_read_plane")
        return out

    # ---- search ----
    @staticmethod
    def _refs_array(refs):
        arr = (C.c_void_p * 8)()
        for (li, ri), p in refs.items():
            arr[li * 4 + ri] = p.ptr
        return arr

    def search_async(self, knobs, cur, refs, out=None, extended=False, alloc=None, tf_compact=False):
        """refs: {(list, idx): Picture}.  Returns (ticket, ResultBuffers)."""
        n_b64 = ((cur.width + 63) // 64) * ((cur.height + 63) // 64)
        if out is None:
            out = abi.ResultBuffers(n_b64, knobs, extended=extended, pinned_alloc=alloc, tf_compact=tf_compact)
        arr = self._refs_array(refs)
        ticket = C.c_uint64()
        self._check(self.lib.svtme_picture_search(self.h, cur.ptr, C.byref(arr), C.byref(knobs), C.byref(out.struct),
                                                  C.byref(ticket)), "svtme_picture_search")
        self._pending[ticket.value] = (out, arr, knobs)
        return ticket.value, out

    def search_device_async(self, knobs, cur, refs):
        arr = self._refs_array(refs)
        ticket = C.c_uint64()
        self._check(self.lib.svtme_picture_search_device(self.h, cur.ptr, C.byref(arr), C.byref(knobs), C.byref(ticket)),
                    "svtme_picture_search_device")
        self._pending[ticket.value] = (None, arr, knobs)
        return ticket.value

    def wait(self, ticket):
        self._check(self.lib.svtme_search_wait(self.h, ticket), "svtme_search_wait")
        out = self._pending.pop(ticket, (None,))[0]
        return out

    def elapsed_ms(self, ticket):
        ms = C.c_float()
        self._check(self.lib.svtme_search_elapsed_ms(self.h, ticket, C.byref(ms)), "svtme_search_elapsed_ms")
        return ms.value

    def px_compares(self, ticket):
        px = C.c_uint64()
        self._check(self.lib.svtme_search_px_compares(self.h, ticket, C.byref(px)), "svtme_search_px_compares")
        return px.value

    def me_picture(self, knobs, cur, refs, extended=True):
        """Synchronous picture search: the call the reference's ME kernel thread makes per picture."""
        t, out = self.search_async(knobs, cur, refs, extended=extended)
        self.wait(t)
        out.px_compares = self.px_compares(t)
        out.kernel_ms = self.elapsed_ms(t)
        return out

    # ---- TPL source-based search (src_ops_process.c:608-884) ----
    def search_tpl_async(self, knobs, cur, refs, tpl_cfg, out=None, want_me=True, alloc=None, tpl=None):
        """ME of the picture + the TPL source-based decisions over the ME results still in HBM.
        Returns (ticket, ResultBuffers | None, TplBuffers)."""
        n_b64 = ((cur.width + 63) // 64) * ((cur.height + 63) // 64)
        if out is None and want_me:
            out = abi.ResultBuffers(n_b64, knobs, extended=False, pinned_alloc=alloc)
        if tpl is None:
            tpl = abi.TplBuffers(cur.width, cur.height, alloc=alloc)
        arr = self._refs_array(refs)
        ticket = C.c_uint64()
        self._check(self.lib.svtme_picture_search_tpl(self.h, cur.ptr, C.byref(arr), C.byref(knobs),
                                                      C.byref(out.struct) if out is not None else None, C.byref(tpl_cfg),
                                                      C.byref(tpl.struct), C.byref(ticket)), "svtme_picture_search_tpl")
        self._pending[ticket.value] = ((out, tpl), arr, knobs, tpl_cfg)
        return ticket.value, out, tpl

    def me_tpl_picture(self, knobs, cur, refs, tpl_cfg):
        t, out, tpl = self.search_tpl_async(knobs, cur, refs, tpl_cfg)
        self.wait(t)
        return out, tpl

    def tpl_src_search(self, knobs, cur, refs, tpl_cfg, me):
        """The TPL source-based decisions over ME results the caller holds (`me`: ResultBuffers, None for an I picture)."""
        tpl = abi.TplBuffers(cur.width, cur.height)
        arr = self._refs_array(refs)
        ticket = C.c_uint64()
        self._check(self.lib.svtme_tpl_src_search(self.h, cur.ptr, C.byref(arr), C.byref(knobs),
                                                  C.byref(me.struct) if me is not None else None, C.byref(tpl_cfg),
                                                  C.byref(tpl.struct), C.byref(ticket)), "svtme_tpl_src_search")
        self._pending[ticket.value] = (tpl, arr, knobs, tpl_cfg, me)
        self.wait(ticket.value)
        return tpl

    def dg_hme_level0(self, src, ref, input_resolution, per_block=True):
        """Dynamic-GOP detector HME level 0 of picture `src` against `ref` (dg_detector_hme_level0, pd_process.c:493):
        (metrics dict, sb_sad, sb_mv, device ms, pixel-compares)."""
        n_b64 = ((src.width + 63) // 64) * ((src.height + 63) // 64)
        m, ms, px = abi.DgMetrics(), C.c_float(), C.c_uint64()
        sad = np.zeros(n_b64, np.uint32) if per_block else None
        mv = np.zeros((n_b64, 2), np.int16) if per_block else None
        self._check(self.lib.svtme_dg_detector_hme_level0(self.h, src.ptr, ref.ptr, input_resolution, C.byref(m),
                                                          sad.ctypes.data if per_block else None,
                                                          mv.ctypes.data if per_block else None, C.byref(ms), C.byref(px)),
                    "svtme_dg_detector_hme_level0")
        return m.as_dict(), sad, mv, ms.value, px.value

    def picture_statistics(self, pic, regions_w=4, regions_h=4, decim_step=4):
        """Histograms / average intensities of the picture's sixteenth plane (svt_aom_gathering_picture_statistics,
        pic_analysis_process.c:1559): (histogram [4][4][256], average_intensity_per_region [4][4], avg_luma)."""
        st = abi.PicStats()
        self._check(self.lib.svtme_picture_statistics(self.h, pic.ptr, regions_w, regions_h, decim_step, C.byref(st)),
                    "svtme_picture_statistics")
        return st.as_arrays()

    def measure_sad_peak(self):
        """Peak VABSDIFF4 rate of this GPU in 1e9 instructions/s (x4 = pixel-compares)."""
        v = C.c_double()
        self._check(self.lib.svtme_measure_sad_peak(self.h, C.byref(v)), "svtme_measure_sad_peak")
        return v.value

    def launch_count(self):
        return int(self.lib.svtme_launch_count(self.h))

    # ---- RTCD-level entry points (host pointers, synchronous) ----
    def sad_loop_kernel(self, src, src_off, src_stride, ref, ref_off, ref_stride, bh, bw, raw_stride, skip, sa_w, sa_h):
        best = C.c_uint64(0)
        x = C.c_int16(-12345)
        y = C.c_int16(-12345)
        self.lib.svtme_cuda_sad_loop_kernel(self.h, src.ctypes.data + src_off, src_stride, ref.ctypes.data + ref_off,
                                            ref_stride, bh, bw, C.byref(best), C.byref(x), C.byref(y), raw_stride, skip,
                                            sa_w, sa_h)
        return best.value, x.value, y.value

    def nxm_sad_kernel(self, src, src_off, src_stride, ref, ref_off, ref_stride, h, w):
        return self.lib.svtme_cuda_nxm_sad_kernel(self.h, src.ctypes.data + src_off, src_stride, ref.ctypes.data + ref_off,
                                                  ref_stride, h, w)

    def nxm_sad_kernel_noalloc(self, src, src_off, src_stride, ref, ref_off, ref_stride, h, w):
        """The allocation-free variant (static device scratch); raises on sizes beyond 64 x 128."""
        out = C.c_uint32(0xdeadbeef)
        self._check(self.lib.svtme_cuda_nxm_sad_kernel_noalloc(self.h, src.ctypes.data + src_off, src_stride, ref.ctypes.data + ref_off,
                                                               ref_stride, h, w, C.byref(out)), "svtme_cuda_nxm_sad_kernel_noalloc")
        return out.value

    def downsample_2d(self, inp, w, h, step):
        ow, oh = (w - (step >> 1) + step - 1) // step, (h - (step >> 1) + step - 1) // step
        out = np.zeros((oh, ow), np.uint8)
        self.lib.svtme_cuda_downsample_2d(self.h, inp.ctypes.data, inp.strides[0], w, h, out.ctypes.data, out.strides[0], step)
        return out

    def fullpel_search_sblock(self, src, src_off, src_stride, ref, ref_off, ref_stride, xo, yo, sa_w, sa_h, sub_sad,
                              best_sad, best_mv):
        self.lib.svtme_cuda_fullpel_search_sblock(self.h, src.ctypes.data + src_off, src_stride, ref.ctypes.data + ref_off,
                                                  ref_stride, xo, yo, sa_w, sa_h, sub_sad, best_sad.ctypes.data,
                                                  best_mv.ctypes.data)
