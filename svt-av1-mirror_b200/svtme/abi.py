"""ctypes mirror of include/svtme_cuda.h (POD structs only).

Kept byte-compatible with the C header; tests/test_abi.py checks sizeof() against the library.
"""
import ctypes as C

MAX_LISTS = 2
MAX_REFS_PER_LIST = 4
SQUARE_PU_COUNT = 85
MAX_SAD_VALUE = 128 * 128 * 255
ME_OPEN_LOOP = 0  # SVTME_ME_OPEN_LOOP
ME_MCTF = 1       # SVTME_ME_MCTF


class Area(C.Structure):
    _fields_ = [("width", C.c_uint16), ("height", C.c_uint16)]


class AreaMinMax(C.Structure):
    _fields_ = [("sa_min", Area), ("sa_max", Area)]


class Knobs(C.Structure):
    _fields_ = [
        ("picture_number", C.c_uint64),
        ("hierarchical_levels", C.c_uint8),
        ("similar_brightness_refs", C.c_uint8),
        ("input_resolution", C.c_uint8),
        ("enable_me_8x8", C.c_uint8),
        ("enable_me_16x16", C.c_uint8),
        ("max_number_of_pus_per_sb", C.c_uint8),
        ("gm_enabled", C.c_uint8),
        ("gm_use_distance_based_active_th", C.c_uint8),
        ("only_l_bwd", C.c_uint8),
        ("max_cand", C.c_uint8),
        ("max_refs", C.c_uint8),
        ("max_l0", C.c_uint8),
        ("num_of_list_to_search", C.c_uint8),
        ("num_of_ref_pic_to_search", C.c_uint8 * MAX_LISTS),
        ("temporal_layer_index", C.c_uint8),
        ("is_ref", C.c_uint8),
        ("hme_search_method", C.c_uint8),
        ("me_search_method", C.c_uint8),
        ("enable_hme_flag", C.c_uint8),
        ("enable_hme_level0_flag", C.c_uint8),
        ("enable_hme_level1_flag", C.c_uint8),
        ("enable_hme_level2_flag", C.c_uint8),
        ("num_hme_sa_w", C.c_uint16),
        ("num_hme_sa_h", C.c_uint16),
        ("hme_l0_sa", AreaMinMax),
        ("hme_l1_sa", Area),
        ("hme_l2_sa", Area),
        ("me_sa", AreaMinMax),
        ("enable_me_hme_ref_pruning", C.c_uint8),
        ("prune_ref_if_hme_sad_dev_bigger_than_th", C.c_uint16),
        ("prune_ref_if_me_sad_dev_bigger_than_th", C.c_uint16),
        ("zz_sad_th", C.c_uint32),
        ("zz_sad_pct", C.c_uint16),
        ("phme_sad_th", C.c_uint32),
        ("phme_sad_pct", C.c_uint16),
        ("enable_me_sr_adjustment", C.c_uint8),
        ("reduce_me_sr_based_on_mv_length_th", C.c_uint16),
        ("stationary_hme_sad_abs_th", C.c_uint16),
        ("stationary_me_sr_divisor", C.c_uint16),
        ("reduce_me_sr_based_on_hme_sad_abs_th", C.c_uint16),
        ("me_sr_divisor_for_low_hme_sad", C.c_uint16),
        ("distance_based_hme_resizing", C.c_uint8),
        ("me_8x8_var_enabled", C.c_uint8),
        ("me_sr_div4_th", C.c_uint32),
        ("me_sr_div2_th", C.c_uint32),
        ("me_sr_mult2_th", C.c_uint32),
        ("mv_sa_adj_enabled", C.c_uint8),
        ("mv_sa_adj_nearest_ref_only", C.c_uint8),
        ("mv_sa_adj_mv_size_th", C.c_uint16),
        ("mv_sa_adj_sa_multiplier", C.c_uint16),
        ("prehme_enable", C.c_uint8),
        ("prehme_sa_cfg", AreaMinMax * 2),
        ("prehme_skip_search_line", C.c_uint8),
        ("prehme_l1_early_exit", C.c_uint8),
        ("prune_me_candidates_th", C.c_int32),
        ("use_best_unipred_cand_only", C.c_uint8),
        ("reduce_hme_l0_sr_th_min", C.c_uint8),
        ("reduce_hme_l0_sr_th_max", C.c_uint8),
        ("me_early_exit_th", C.c_uint32),
        ("me_safe_limit_zz_th", C.c_uint32),
        ("prev_me_stage_based_exit_th", C.c_uint32),
        ("me_type", C.c_uint8),
        ("tf_me_exit_th", C.c_uint32),
    ]

    def as_dict(self):
        def conv(v):
            if isinstance(v, C.Structure):
                return {n: conv(getattr(v, n)) for n, _ in v._fields_}
            if isinstance(v, C.Array):
                return [conv(x) for x in v]
            return v
        return conv(self)


class PresetCfg(C.Structure):
    _fields_ = [
        ("enc_mode", C.c_int8),
        ("rtc_tune", C.c_uint8),
        ("sc_class1", C.c_uint8),
        ("rate_control_vbr", C.c_uint8),
        ("fast_decode", C.c_uint8),
        ("qp", C.c_uint32),
        ("frame_rate_q16", C.c_uint32),
        ("width", C.c_uint16),
        ("height", C.c_uint16),
        ("hierarchical_levels", C.c_uint8),
        ("picture_number", C.c_uint64),
        ("temporal_layer_index", C.c_uint8),
        ("is_ref", C.c_uint8),
        ("slice_type_b", C.c_uint8),
        ("similar_brightness_refs", C.c_uint8),
        ("ref_list0_count_try", C.c_uint8),
        ("ref_list1_count_try", C.c_uint8),
    ]


class TfCfg(C.Structure):
    """SvtMeTfCfg: the TfControls fields the temporal-filter knob derivation reads."""
    _fields_ = [
        ("hme_me_level", C.c_uint8),
        ("qp_opt", C.c_uint8),
        ("me_exit_th", C.c_uint32),
        ("qp", C.c_uint32),
        ("width", C.c_uint16),
        ("height", C.c_uint16),
        ("picture_number", C.c_uint64),
        ("temporal_layer_index", C.c_uint8),
        ("is_ref", C.c_uint8),
    ]


class PictureResults(C.Structure):
    _fields_ = [
        ("b64_total_count", C.c_uint32),
        ("pu_count", C.c_uint32),
        ("me_mv", C.POINTER(C.c_uint32)),
        ("me_cand", C.POINTER(C.c_uint8)),
        ("total_me_candidate_index", C.POINTER(C.c_uint8)),
        ("me_64x64_distortion", C.POINTER(C.c_uint32)),
        ("me_32x32_distortion", C.POINTER(C.c_uint32)),
        ("me_16x16_distortion", C.POINTER(C.c_uint32)),
        ("me_8x8_distortion", C.POINTER(C.c_uint32)),
        ("me_8x8_cost_variance", C.POINTER(C.c_uint32)),
        ("rc_me_distortion", C.POINTER(C.c_uint32)),
        ("stationary_block_present_sb", C.POINTER(C.c_uint8)),
        ("rc_me_allow_gm", C.POINTER(C.c_uint8)),
        ("sb_best_sad", C.POINTER(C.c_uint32)),
        ("sb_best_mv", C.POINTER(C.c_uint32)),
        ("do_ref", C.POINTER(C.c_uint8)),
        ("hme_sc", C.POINTER(C.c_int16)),
        ("tf_exit_sb", C.POINTER(C.c_uint8)),
        ("tf_blk_counts", C.POINTER(C.c_uint32)),
        ("picture_sums", C.POINTER(C.c_uint64)),
        ("tf_best_sad", C.POINTER(C.c_uint32)),
        ("tf_best_mv", C.POINTER(C.c_uint32)),
    ]


class DgMetrics(C.Structure):
    """SvtMeDgMetrics = DGDetectorMetrics of the reference (pcs.h:731-737) without the segment counter."""
    _fields_ = [("tot_dist", C.c_uint64), ("tot_cplx", C.c_uint32), ("tot_active", C.c_uint32), ("sum_in_vectors", C.c_int32)]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


class PicStats(C.Structure):
    """SvtMePicStats: per-region histograms / average intensities of the sixteenth plane and the picture's avg_luma."""
    _fields_ = [("picture_histogram", ((C.c_uint32 * 256) * 4) * 4), ("average_intensity_per_region", (C.c_uint64 * 4) * 4),
                ("avg_luma", C.c_uint64)]

    def as_arrays(self):
        return (np.ctypeslib.as_array(self.picture_histogram).copy(), np.ctypeslib.as_array(self.average_intensity_per_region).copy(),
                int(self.avg_luma))


class DeviceCfg(C.Structure):
    _fields_ = [
        ("device", C.c_int32),
        ("max_width", C.c_uint32),
        ("max_height", C.c_uint32),
        ("max_pictures", C.c_uint32),
        ("max_inflight", C.c_uint32),
        ("sixteenth_from_full", C.c_uint32),
        ("pipeline", C.c_uint32),
    ]


TPL_DC_PRED = 0   # SVTME_TPL_DC_PRED
TPL_NEWMV = 16    # SVTME_TPL_NEWMV


class TplCfg(C.Structure):
    """SvtMeTplCfg: the controls of the TPL source-based search (src_ops_process.c:519-884)."""
    _fields_ = [("dispenser_search_level", C.c_uint8), ("use_sad_in_src_search", C.c_uint8), ("subpel_full_pel", C.c_uint8),
                ("intra_dc_only", C.c_uint8), ("disable_intra_pred", C.c_uint8), ("slice_type_i", C.c_uint8),
                ("ref_excluded", (C.c_uint8 * MAX_REFS_PER_LIST) * MAX_LISTS), ("ref_poc", (C.c_uint64 * MAX_REFS_PER_LIST) * MAX_LISTS)]


class TplResults(C.Structure):
    _fields_ = [("grid_width", C.c_uint32), ("grid_height", C.c_uint32), ("block_size", C.POINTER(C.c_uint8)),
                ("best_mode", C.POINTER(C.c_uint8)), ("best_rf_idx", C.POINTER(C.c_int8)), ("mv", C.POINTER(C.c_int16)),
                ("ref_frame_poc", C.POINTER(C.c_uint64)), ("best_inter_cost", C.POINTER(C.c_uint32)),
                ("best_intra_cost", C.POINTER(C.c_uint32))]


def pu_count_of(knobs):
    """Number of PUs stored per SB (pcs.c:106-110)."""
    if not knobs.enable_me_16x16:
        return 5
    return 85 if knobs.enable_me_8x8 else 21


import numpy as np  # noqa: E402


class ResultBuffers:
    """Caller-owned host buffers of one picture (SvtMePictureResults) backed by numpy arrays."""

    SCALARS32 = ("me_64x64_distortion", "me_32x32_distortion", "me_16x16_distortion", "me_8x8_distortion",
                 "me_8x8_cost_variance", "rc_me_distortion")
    SCALARS8 = ("stationary_block_present_sb", "rc_me_allow_gm")

    def __init__(self, n_b64, knobs, extended=True, pinned_alloc=None, tf_compact=False):
        self.n_b64 = n_b64
        self.pu_count = pu_count_of(knobs)
        mr, mc = max(1, knobs.max_refs), max(1, knobs.max_cand)
        alloc = pinned_alloc or (lambda shape, dt: np.zeros(shape, dt))
        self.arrays = {}
        if knobs.me_type != ME_MCTF:  # the temporal filter reads no candidates or statistics
            self.arrays = {
                "me_mv": alloc((n_b64, self.pu_count, mr), np.uint32),
                "me_cand": alloc((n_b64, self.pu_count, mc), np.uint8),
                "total_me_candidate_index": alloc((n_b64, self.pu_count), np.uint8),
            }
            for n in self.SCALARS32:
                self.arrays[n] = alloc((n_b64,), np.uint32)
            for n in self.SCALARS8:
                self.arrays[n] = alloc((n_b64,), np.uint8)
            self.arrays["picture_sums"] = alloc((4,), np.uint64)
        if knobs.me_type == ME_MCTF and tf_compact:  # only what the temporal filter reads
            self.arrays["tf_best_sad"] = alloc((n_b64, 85), np.uint32)
            self.arrays["tf_best_mv"] = alloc((n_b64, 85), np.uint32)
        if extended or knobs.me_type == ME_MCTF:
            if not (knobs.me_type == ME_MCTF and tf_compact):
                self.arrays["sb_best_sad"] = alloc((n_b64, 2, 4, 85), np.uint32)
                self.arrays["sb_best_mv"] = alloc((n_b64, 2, 4, 85), np.uint32)
            self.arrays["do_ref"] = alloc((n_b64, 2, 4), np.uint8)
            self.arrays["hme_sc"] = alloc((n_b64, 2, 4, 2), np.int16)
            if knobs.me_type == ME_MCTF:
                self.arrays["tf_exit_sb"] = alloc((n_b64,), np.uint8)
                self.arrays["tf_blk_counts"] = alloc((2,), np.uint32)
        s = PictureResults()
        s.b64_total_count = n_b64
        s.pu_count = self.pu_count
        for name, arr in self.arrays.items():
            ctype = dict(PictureResults._fields_)[name]
            setattr(s, name, arr.ctypes.data_as(ctype))
        self.struct = s

    def __getitem__(self, k):
        return self.arrays[k]

    def nbytes(self):
        return sum(a.nbytes for a in self.arrays.values())

    def diff(self, other, names=None):
        """Names of arrays that differ from `other` (bit-exact compare)."""
        bad = []
        for n in (names or self.arrays.keys()):
            if n in other.arrays and not np.array_equal(self.arrays[n], other.arrays[n]):
                bad.append(n)
        return bad


class TplBuffers:
    """Caller-owned host buffers of one picture's TPL source-based decisions (SvtMeTplResults) backed by numpy arrays."""
    FIELDS = (("block_size", np.uint8, ()), ("best_mode", np.uint8, ()), ("best_rf_idx", np.int8, ()), ("mv", np.int16, (2,)),
              ("ref_frame_poc", np.uint64, ()), ("best_inter_cost", np.uint32, ()), ("best_intra_cost", np.uint32, ()))

    def __init__(self, width, height, alloc=None):
        self.gw, self.gh = (width + 15) >> 4, (height + 15) >> 4
        alloc = alloc or (lambda shape, dt: np.zeros(shape, dt))
        # one arena, the arrays in SvtMeTplResults order, each 16-byte aligned: the library's own layout, so that page-locked
        # buffers receive the whole picture's decisions in one device-to-host copy
        cells, off, offs = self.gw * self.gh, 0, []
        for n, dt, tail in self.FIELDS:
            off = (off + 15) & ~15
            offs.append(off)
            off += cells * np.dtype(dt).itemsize * (tail[0] if tail else 1)
        self.arena = alloc(((off + 15) & ~15,), np.uint8)
        self.arrays = {}
        for (n, dt, tail), o in zip(self.FIELDS, offs):
            cnt = cells * (tail[0] if tail else 1)
            self.arrays[n] = self.arena[o:o + cnt * np.dtype(dt).itemsize].view(dt).reshape((self.gh, self.gw) + tail)
        s = TplResults()
        s.grid_width, s.grid_height = self.gw, self.gh
        for n, a in self.arrays.items():
            setattr(s, n, a.ctypes.data_as(dict(TplResults._fields_)[n]))
        self.struct = s

    def __getitem__(self, k):
        return self.arrays[k]

    def nbytes(self):
        return sum(a.nbytes for a in self.arrays.values())

    def diff(self, other):
        return [n for n in self.arrays if not np.array_equal(self.arrays[n], other.arrays[n])]
