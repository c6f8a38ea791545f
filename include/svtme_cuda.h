/*
 * svtme_cuda.h -- C ABI of the B200-native open-loop motion-estimation library
 * (libsvtme_cuda.so).
 *
 * This is the drop-in boundary for SVT-AV1's open-loop ME hot path.  Every
 * entry point cites the reference interface it replaces (paths relative to the
 * SVT-AV1 v3.0.2 tree).  Only plain C types cross the boundary: pointers,
 * sizes, and the POD structs declared here.  No torch / CUDA types.
 *
 * Return codes reuse the numeric values of the reference's EbErrorType
 * (Source/API/EbSvtAv1.h:121-135).
 */
#ifndef SVTME_CUDA_H
#define SVTME_CUDA_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SVTME_OK                       0            /* EB_ErrorNone */
#define SVTME_ERR_INSUFFICIENT         ((int32_t)0x80001000) /* EB_ErrorInsufficientResources */
#define SVTME_ERR_UNDEFINED            ((int32_t)0x80001001) /* EB_ErrorUndefined (CUDA failure) */
#define SVTME_ERR_BAD_PARAMETER        ((int32_t)0x80001005) /* EB_ErrorBadParameter */

#define SVTME_MAX_LISTS          2   /* MAX_NUM_OF_REF_PIC_LIST, definitions.h:2337 */
#define SVTME_MAX_REFS_PER_LIST  4   /* REF_LIST_MAX_DEPTH / MAX_REF_IDX, EbSvtAv1Enc.h:36, definitions.h:2352 */
#define SVTME_SQUARE_PU_COUNT    85  /* SQUARE_PU_COUNT, me_sb_results.h:24 */
#define SVTME_PU_COUNT_NO_8X8    21  /* MAX_SB64_PU_COUNT_NO_8X8, me_sb_results.h:25 */
#define SVTME_PU_COUNT_WO_16X16  5   /* MAX_SB64_PU_COUNT_WO_16X16, me_sb_results.h:26 */
#define SVTME_MAX_SAD_VALUE      (128u * 128u * 255u) /* motion_estimation.h:85 */

/* EbInputResolution classes, definitions.h:2076-2086 */
enum {
    SVTME_INPUT_SIZE_240p = 0, SVTME_INPUT_SIZE_360p = 1, SVTME_INPUT_SIZE_480p = 2,
    SVTME_INPUT_SIZE_720p = 3, SVTME_INPUT_SIZE_1080p = 4, SVTME_INPUT_SIZE_4K = 5,
    SVTME_INPUT_SIZE_8K = 6
};

/* SearchArea / SearchAreaMinMax, me_context.h:321-328 */
typedef struct SvtMeArea {
    uint16_t width;
    uint16_t height;
} SvtMeArea;
typedef struct SvtMeAreaMinMax {
    SvtMeArea sa_min;
    SvtMeArea sa_max;
} SvtMeAreaMinMax;

/*
 * SvtMeKnobs -- the flattened, derived ME configuration of ONE picture.
 *
 * It is the union of
 *  (a) the MeContext knob fields written by svt_aom_sig_deriv_me
 *      (enc_mode_config.c:671-808; struct at me_context.h:366-509),
 *  (b) the per-picture fields me_process.c copies into MeContext
 *      (me_process.c:216-227), and
 *  (c) the PictureParentControlSet / SequenceControlSet fields that
 *      motion_estimation.c reads directly.
 * The encoder keeps deriving these on the CPU exactly as today; the shim only
 * copies them out of MeContext/pcs.
 */
/* SvtMeKnobs.me_type.  0 keeps a zero-initialised knob block on the open-loop path; the shim maps
 * EbMeType (me_context.h:45-50) ME_OPEN_LOOP(3) -> 0 (me_process.c:216) and ME_MCTF(1) -> 1 (temporal_filtering.c:296). */
#define SVTME_ME_OPEN_LOOP 0
#define SVTME_ME_MCTF      1

typedef struct SvtMeKnobs {
    /* (c) pcs / scs fields */
    uint64_t picture_number;          /* pcs->picture_number */
    uint8_t  hierarchical_levels;     /* pcs->hierarchical_levels   (motion_estimation.c:2419) */
    uint8_t  similar_brightness_refs; /* pcs->similar_brightness_refs (:2420) */
    uint8_t  input_resolution;        /* scs->input_resolution class (:2851, :2998) */
    uint8_t  enable_me_8x8;           /* pcs->enable_me_8x8  (pcs.c:1392) */
    uint8_t  enable_me_16x16;         /* pcs->enable_me_16x16 (pcs.c:1389) */
    uint8_t  max_number_of_pus_per_sb;/* pcs->max_number_of_pus_per_sb, always 85 (resource_coordination_process.c:425) */
    uint8_t  gm_enabled;              /* pcs->gm_ctrls.enabled (:3149) */
    uint8_t  gm_use_distance_based_active_th; /* pcs->gm_ctrls.use_distance_based_active_th (:2879) */
    uint8_t  only_l_bwd;              /* scs->mrp_ctrls.only_l_bwd (:2783) */
    uint8_t  max_cand;                /* pa_me_data->max_cand (pd_process.c:3517) */
    uint8_t  max_refs;                /* pa_me_data->max_refs */
    uint8_t  max_l0;                  /* pa_me_data->max_l0 */

    /* (b) per-picture MeContext fields, me_process.c:219-227 */
    uint8_t  num_of_list_to_search;
    uint8_t  num_of_ref_pic_to_search[SVTME_MAX_LISTS];
    uint8_t  temporal_layer_index;
    uint8_t  is_ref;

    /* (a) derived MeContext knobs */
    uint8_t  hme_search_method;       /* 0 = SUB_SAD_SEARCH, 1 = FULL_SAD_SEARCH (definitions.h) */
    uint8_t  me_search_method;
    uint8_t  enable_hme_flag;
    uint8_t  enable_hme_level0_flag;
    uint8_t  enable_hme_level1_flag;
    uint8_t  enable_hme_level2_flag;
    uint16_t num_hme_sa_w;            /* always 2 */
    uint16_t num_hme_sa_h;            /* always 2 */
    SvtMeAreaMinMax hme_l0_sa;
    SvtMeArea       hme_l1_sa;
    SvtMeArea       hme_l2_sa;
    SvtMeAreaMinMax me_sa;
    /* MeHmeRefPruneCtrls, me_context.h:280-290 */
    uint8_t  enable_me_hme_ref_pruning;
    uint16_t prune_ref_if_hme_sad_dev_bigger_than_th;
    uint16_t prune_ref_if_me_sad_dev_bigger_than_th;
    uint32_t zz_sad_th;
    uint16_t zz_sad_pct;
    uint32_t phme_sad_th;
    uint16_t phme_sad_pct;
    /* MeSrCtrls, me_context.h:292-305 */
    uint8_t  enable_me_sr_adjustment;
    uint16_t reduce_me_sr_based_on_mv_length_th;
    uint16_t stationary_hme_sad_abs_th;
    uint16_t stationary_me_sr_divisor;
    uint16_t reduce_me_sr_based_on_hme_sad_abs_th;
    uint16_t me_sr_divisor_for_low_hme_sad;
    uint8_t  distance_based_hme_resizing;
    /* Me8x8VarCtrls, me_context.h:310-319 */
    uint8_t  me_8x8_var_enabled;
    uint32_t me_sr_div4_th;
    uint32_t me_sr_div2_th;
    uint32_t me_sr_mult2_th;
    /* MvBasedSearchAdj, me_context.h:356-364 */
    uint8_t  mv_sa_adj_enabled;
    uint8_t  mv_sa_adj_nearest_ref_only;
    uint16_t mv_sa_adj_mv_size_th;
    uint16_t mv_sa_adj_sa_multiplier;
    /* PreHmeCtrls, me_context.h:336-341 */
    uint8_t  prehme_enable;
    SvtMeAreaMinMax prehme_sa_cfg[2];
    uint8_t  prehme_skip_search_line;
    uint8_t  prehme_l1_early_exit;
    /* scalars, me_context.h:491-508 */
    int32_t  prune_me_candidates_th;
    uint8_t  use_best_unipred_cand_only;
    uint8_t  reduce_hme_l0_sr_th_min;
    uint8_t  reduce_hme_l0_sr_th_max;
    uint32_t me_early_exit_th;
    uint32_t me_safe_limit_zz_th;
    uint32_t prev_me_stage_based_exit_th;
    /* temporal-filter mode (SVTME_ME_MCTF): the caller is produce_temporally_filtered_pic
     * (temporal_filtering.c:3129-3174); one list, one reference = the neighbouring picture of the window */
    uint8_t  me_type;                 /* SVTME_ME_OPEN_LOOP / SVTME_ME_MCTF, MeContext.me_type (me_context.h EbMeType) */
    uint32_t tf_me_exit_th;           /* MeContext.tf_me_exit_th = pcs->tf_ctrls.me_exit_th (temporal_filtering.c:3158) */
} SvtMeKnobs;

/*
 * Preset-level inputs of svt_aom_sig_deriv_me (enc_mode_config.c:671) and of
 * the pcs/scs fields around it.  svtme_derive_knobs() is a convenience mirror
 * of that derivation for callers that do not run the encoder (tests, bench);
 * an integrated encoder copies its own derived MeContext instead.
 */
typedef struct SvtMePresetCfg {
    int8_t   enc_mode;             /* ENC_MR(-1) .. ENC_M13, EbSvtAv1Enc.h:45-61 */
    uint8_t  rtc_tune;             /* pred_structure == SVT_AV1_PRED_LOW_DELAY_B */
    uint8_t  sc_class1;            /* screen-content class */
    uint8_t  rate_control_vbr;     /* rate_control_mode == VBR (mrp level, enc_handle.c:4163) */
    uint8_t  fast_decode;          /* static_config.fast_decode */
    uint32_t qp;                   /* static_config.qp (default 35, enc_settings.h:22) */
    uint32_t frame_rate_q16;       /* scs->frame_rate (Q16), enc_mode_config.c:335 */
    uint16_t width;                /* luma width  (multiple of 8) */
    uint16_t height;               /* luma height (multiple of 8) */
    uint8_t  hierarchical_levels;  /* pcs->hierarchical_levels */
    /* per picture */
    uint64_t picture_number;
    uint8_t  temporal_layer_index;
    uint8_t  is_ref;
    uint8_t  slice_type_b;         /* 1 = B_SLICE (two lists), 0 = P_SLICE */
    uint8_t  similar_brightness_refs;
    uint8_t  ref_list0_count_try;  /* 0 => take the mrp default for this layer */
    uint8_t  ref_list1_count_try;
} SvtMePresetCfg;

/*
 * Host result buffers of one picture, caller-owned.  They are the flat,
 * picture-level form of the per-SB MeSbResults arrays (me_sb_results.h:44-51,
 * allocated at pcs.c:98-119) plus the eight per-SB pcs statistics written by
 * compute_distortion / perform_gm_detection (motion_estimation.c:2964-3007,
 * 2838-2961).  pu_count = 85 / 21 / 5 as at pcs.c:106-110.
 *
 * Entries the reference leaves unwritten (MVs of pruned references,
 * candidates beyond total_me_candidate_index) are 0 here.
 */
typedef struct SvtMePictureResults {
    uint32_t  b64_total_count;             /* in: capacity check; picture width/height in 64x64 blocks multiplied */
    uint32_t  pu_count;                    /* in: 85, 21 or 5 */
    uint32_t *me_mv;                       /* [b64][pu_count][max_refs]  MvCandidate.as_int = (y<<16)|(x&0xffff), full-pel */
    uint8_t  *me_cand;                     /* [b64][pu_count][max_cand]  MeCandidate bit-fields: dir:2 | ref_idx_l0:2<<2 | ref_idx_l1:2<<4 | ref0_list<<6 | ref1_list<<7 */
    uint8_t  *total_me_candidate_index;    /* [b64][pu_count] */
    uint32_t *me_64x64_distortion;         /* [b64] */
    uint32_t *me_32x32_distortion;
    uint32_t *me_16x16_distortion;
    uint32_t *me_8x8_distortion;
    uint32_t *me_8x8_cost_variance;
    uint32_t *rc_me_distortion;
    uint8_t  *stationary_block_present_sb;
    uint8_t  *rc_me_allow_gm;
    /* Optional (may be NULL): MeContext state at the end of each SB, for parity checks of SADs and selected refs. */
    uint32_t *sb_best_sad;                 /* [b64][2][4][85] p_sb_best_sad, tier-zero PU order; MAX_SAD_VALUE where the ref was not searched */
    uint32_t *sb_best_mv;                  /* [b64][2][4][85] p_sb_best_mv */
    uint8_t  *do_ref;                      /* [b64][2][4]  search_results[l][r].do_ref after me_prune_ref */
    int16_t  *hme_sc;                      /* [b64][2][4][2] search_results[l][r].hme_sc_x / hme_sc_y */
    /* Optional, temporal-filter mode only (NULL otherwise).  In that mode the outputs the filter consumes are
     * sb_best_mv / sb_best_sad of [0][0] (p_best_mv64x64 .. 16x16, temporal_filtering.c:1866-2114, 2685),
     * hme_sc of [0][0], and these two: */
    uint8_t  *tf_exit_sb;                  /* [b64] 1: search_results[0][0].hme_sad < tf_me_exit_th, the integer search
                                            * was skipped and tf_use_pred_64x64_only_th = 0xFF (motion_estimation.c:3109) */
    uint32_t *tf_blk_counts;               /* [2] tf_tot_horz_blks, tf_tot_vert_blks of this picture pair (:2469-2474) */
    /* Optional (may be NULL; not in temporal-filter mode): the picture-level sums the ME kernel thread's consumers
     * loop over the per-block arrays for, accumulated on the device:
     *   [0] sum of me_8x8_distortion        -> Initial RC: norm_me_dist = [0] / b64_total_count (initial_rc_process.c:683-691)
     *   [1] sum of rc_me_distortion         -> global_me.c:162-168: total_me_sad = (uint32_t)[1] (the reference sums in 32 bits)
     *   [2] sum of stationary_block_present_sb -> total_stationary_sb
     *   [3] sum of rc_me_allow_gm           -> total_gm_sbs */
    uint64_t *picture_sums;                /* [4] */
    /* Optional, temporal-filter mode only: the [0][0] rows of sb_best_sad / sb_best_mv alone -- all the filter reads (one
     * reference per search).  A caller that passes these and leaves sb_best_sad / sb_best_mv NULL gets 0.7 KB per block
     * copied back instead of 5.4 KB. */
    uint32_t *tf_best_sad;                 /* [b64][85] p_sb_best_sad[0][0] */
    uint32_t *tf_best_mv;                  /* [b64][85] p_sb_best_mv[0][0] */
} SvtMePictureResults;

typedef struct SvtMeDeviceCfg {
    int32_t  device;            /* CUDA device ordinal */
    uint32_t max_width;         /* largest luma width that will be registered */
    uint32_t max_height;
    uint32_t max_pictures;      /* device picture slots (source + references in flight); 0 => 64 */
    uint32_t max_inflight;      /* concurrent picture searches (streams); 0 => 4 */
    uint32_t sixteenth_from_full; /* 1: HME level 1 is off, the sixteenth plane is decimated from the full plane with
                                   * step 4 (pic_analysis_process.c:1981-1992); 0: from the quarter plane with step 2 */
    uint32_t pipeline;          /* must be 0: staged -- every search engine of the per-SB program is one kernel over the
                                 * whole picture (the round-1 persistent kernel, value 1, is no longer built) */
} SvtMeDeviceCfg;

typedef struct SvtMeHandle SvtMeHandle;   /* opaque */
typedef struct SvtMePic    SvtMePic;      /* opaque: one picture's full / quarter / sixteenth luma planes in HBM */

/* Library/device bring-up.  Replaces: the allocation of PA reference objects
 * (reference_object.c:217-345) and ME contexts (enc_handle.c:2102-2110). */
int32_t svtme_init(const SvtMeDeviceCfg *cfg, SvtMeHandle **out);
void    svtme_destroy(SvtMeHandle *h);

/* Mirror of svt_aom_sig_deriv_me + set_hme/me_search_params (enc_mode_config.c:136-808),
 * HME flags (:1608-1619), enable_me_8x8 (:76-96), gm level (:103-129, 1821-1903),
 * mrp controls (enc_handle.c:3347-3640, 4141-4176), max_cand/max_refs (pcs.c:91-96). Pure host code. */
int32_t svtme_derive_knobs(const SvtMePresetCfg *cfg, SvtMeKnobs *out);

/*
 * Temporal-filter (alt-ref) motion search: the knobs of one (centre picture, neighbouring picture) pair as
 * produce_temporally_filtered_pic sets them up before svt_aom_motion_estimation_b64
 * (temporal_filtering.c:3129-3174): me_type = ME_MCTF (:296), HME always full-SAD (:316), one list / one
 * reference (:3140-3142), hme_l0_sa = hme_l0_sa_default_tf (set_hme_search_params_mctf level 0, :2759-2767),
 * tf_me_exit_th = tf_ctrls.me_exit_th (:3158).  The inputs are the TfControls fields the derivation reads
 * (definitions.h:172-260; per-preset values at enc_handle.c:2560-3160).
 */
typedef struct SvtMeTfCfg {
    uint8_t  hme_me_level;         /* tf_ctrls.hme_me_level 0..4: search areas (enc_mode_config.c:588-653), HME levels (:1620-1643) */
    uint8_t  qp_opt;               /* tf_ctrls.qp_opt: modulate the ME area by qp (:656-663) */
    uint32_t me_exit_th;           /* tf_ctrls.me_exit_th */
    uint32_t qp;                   /* scs->static_config.qp */
    uint16_t width;                /* luma size: input_resolution class for hme_me_level 2 */
    uint16_t height;
    uint64_t picture_number;       /* of the centre picture */
    uint8_t  temporal_layer_index; /* of the centre picture (temporal_filtering.c:3143-3146) */
    uint8_t  is_ref;
} SvtMeTfCfg;

/* Mirror of svt_aom_sig_deriv_me_tf + tf_set_me_hme_params_oq (enc_mode_config.c:588-663, 814-854) and the
 * per-call overrides listed above.  Pure host code. */
int32_t svtme_derive_tf_knobs(const SvtMeTfCfg *cfg, SvtMeKnobs *out);

/* Upload one 8-bit luma plane, then build the padded full / quarter / sixteenth
 * planes on the device.  Replaces: copy_frame_buffer's luma memcpy
 * (enc_handle.c:4920-4925), svt_aom_pad_input_pictures (pic_analysis_process.c:2004),
 * svt_aom_downsample_filtering_input_picture (:1945-2002) incl. RTCD downsample_2d
 * (aom_dsp_rtcd.c:509) and svt_aom_generate_padding (pic_operators.c:338).
 * `luma` points at pixel (0,0); width/height must be multiples of 8.
 * Asynchronous: the host buffer must stay valid until svtme_picture_wait(). */
int32_t svtme_picture_register(SvtMeHandle *h, uint64_t picture_number, const uint8_t *luma, uint32_t stride,
                               uint32_t width, uint32_t height, SvtMePic **out);
/* 10-bit input: ME sees the 8-bit MSB plane (in16 >> 2), as svt_unpack_and_2bcompress
 * does (pack_unpack_c.c:77-117, enc_handle.c:4964-4972).  stride in samples. */
int32_t svtme_picture_register16(SvtMeHandle *h, uint64_t picture_number, const uint16_t *luma16, uint32_t stride,
                                 uint32_t width, uint32_t height, SvtMePic **out);
/* Same, from a luma plane that is already in device memory (device pointer). */
int32_t svtme_picture_register_device(SvtMeHandle *h, uint64_t picture_number, const void *d_luma, uint32_t stride,
                                      uint32_t width, uint32_t height, SvtMePic **out);
/*
 * Reference planes across GPUs (no counterpart in the single-process, single-memory reference:
 * there every ME thread reads the same EbPaReferenceObject, reference_object.c:217-345).  A picture is uploaded and
 * decimated ONCE, on the GPU that owns it; the GPUs whose searches reference it get a copy of its three padded planes
 * (one allocation of svtme_picture_planes_bytes(width, height) bytes, the same layout on every device).
 *  - one process driving several GPUs (the encoder is one process, enc_handle.c:2245-2307): svtme_picture_replicate
 *    copies `pic` of src_h into a new picture of dst_h with cudaMemcpyPeerAsync (NVLink when the devices are peers).
 *    Asynchronous; `pic` may be released right away (its release waits for the copy).
 *  - one process per GPU (torch.distributed / NCCL): svtme_picture_export_planes copies the planes into a device buffer
 *    of the caller (synchronous: ready to send when it returns), the caller moves the bytes (ncclSend / ncclBroadcast),
 *    svtme_picture_import_planes builds a picture from the received bytes (asynchronous: d_src must stay valid until
 *    svtme_picture_wait).  No decimation runs on the receiving GPU.
 */
int32_t svtme_picture_planes_bytes(uint32_t width, uint32_t height, uint64_t *bytes);
int32_t svtme_picture_replicate(SvtMeHandle *src_h, SvtMePic *pic, SvtMeHandle *dst_h, SvtMePic **out);
int32_t svtme_picture_export_planes(SvtMeHandle *h, SvtMePic *pic, void *d_dst, uint64_t dst_bytes);
int32_t svtme_picture_import_planes(SvtMeHandle *h, uint64_t picture_number, const void *d_src, uint32_t width, uint32_t height,
                                    SvtMePic **out);
/* Zero-copy form of the export: the device address of the picture's plane allocation (svtme_picture_planes_bytes bytes, the
 * layout svtme_picture_import_planes expects), for a send straight out of it.  Returns when the planes are built.  Read-only;
 * valid until the picture is released, so release it only after the send has completed. */
int32_t svtme_picture_planes_device(SvtMeHandle *h, SvtMePic *pic, const void **d_planes, uint64_t *bytes);
/* Wait until the planes of `pic` are built. */
int32_t svtme_picture_wait(SvtMeHandle *h, SvtMePic *pic);
/* Copy a padded plane back (level 0 = full, 1 = quarter, 2 = sixteenth) for the
 * CPU users of the decimated planes (temporal filter, DG detector) and for tests.
 * dst receives (width>>level) x (height>>level) pixels plus `pad` replicated
 * pixels on every side (pad <= 64 >> level); dst_stride >= (width>>level)+2*pad. Synchronous. */
int32_t svtme_picture_read_plane(SvtMeHandle *h, SvtMePic *pic, int32_t level, uint32_t pad, uint8_t *dst,
                                 uint32_t dst_stride);
int32_t svtme_picture_release(SvtMeHandle *h, SvtMePic *pic);

/* Open-loop ME of every 64x64 block of `cur` against refs[list][idx].
 * Replaces: the segment/SB loop of svt_aom_motion_estimation_kernel
 * (me_process.c:174-290) and svt_aom_motion_estimation_b64
 * (motion_estimation.c:3076-3153) for all b64 of the picture.
 * Asynchronous; `out` buffers are filled when svtme_search_wait() returns.
 * Unused refs[][] entries may be NULL. */
int32_t svtme_picture_search(SvtMeHandle *h, SvtMePic *cur, SvtMePic *const refs[SVTME_MAX_LISTS][SVTME_MAX_REFS_PER_LIST],
                             const SvtMeKnobs *knobs, SvtMePictureResults *out, uint64_t *ticket);
int32_t svtme_search_wait(SvtMeHandle *h, uint64_t ticket);
/* Device-resident variant for throughput measurement: results stay in HBM
 * (no D2H); used by bench.py for `value`.  Returns device time of the last
 * completed search of this ticket in milliseconds via svtme_search_elapsed_ms. */
int32_t svtme_picture_search_device(SvtMeHandle *h, SvtMePic *cur,
                                    SvtMePic *const refs[SVTME_MAX_LISTS][SVTME_MAX_REFS_PER_LIST],
                                    const SvtMeKnobs *knobs, uint64_t *ticket);
int32_t svtme_search_elapsed_ms(SvtMeHandle *h, uint64_t ticket, float *ms);
/* Algorithmic work of a completed search: pixel-compares (block pixels x candidate positions summed over every SAD
 * evaluation the reference's svt_sad_loop_kernel / svt_nxm_sad_kernel / full-pel scan would perform). */
int32_t svtme_search_px_compares(SvtMeHandle *h, uint64_t ticket, uint64_t *px_compares);
/*
 * Dynamic-GOP detector, HME level 0.  Replaces: the TASK_DG_DETECTOR_HME branch of the ME
 * kernel thread (me_process.c:327-329) = dg_detector_hme_level0 (pd_process.c:493-590) for EVERY segment of the
 * picture, with early_hme_b64 (:392-491): the 16x16 sixteenth-plane block of each 64x64 block of `src` is searched
 * with full SAD over a 16 / 64 / 128-square window (input_resolution <= 360p / <= 480p / larger, :497-498) of `ref`'s
 * sixteenth plane, and the picture-level DGDetectorMetrics (pcs.h:731-737) are accumulated on the device.
 * Synchronous.  Optional outputs (may be NULL): sb_sad [b64] level-0 SAD, sb_mv [b64][2] = (col, row) in full-resolution
 * pixels, elapsed_ms = device time of the search, px_compares = its algorithmic work.
 */
typedef struct SvtMeDgMetrics {
    uint64_t tot_dist;        /* sum of the level-0 SADs */
    uint32_t tot_cplx;        /* blocks with SAD > 16*16*30 */
    uint32_t tot_active;      /* blocks with a non-zero vector */
    int32_t  sum_in_vectors;  /* vectors pointing away from (+) / towards (-) the picture centre, per component */
} SvtMeDgMetrics;
int32_t svtme_dg_detector_hme_level0(SvtMeHandle *h, SvtMePic *src, SvtMePic *ref, uint32_t input_resolution,
                                     SvtMeDgMetrics *metrics, uint32_t *sb_sad, int16_t *sb_mv, float *elapsed_ms,
                                     uint64_t *px_compares);
/*
 * Picture-analysis statistics on the sixteenth plane (second half).  Replaces: the
 * histogram part of svt_aom_gathering_picture_statistics (pic_analysis_process.c:1559-1577) =
 * sub_sample_luma_generate_pixel_intensity_histogram_bins (:1461-1524) with calculate_histogram (:166-184): per region
 * of the regions_per_width x regions_per_height grid (4 x 4 for pictures of at least 64 x 64, enc_handle.c:4221-4224;
 * the last row / column of regions takes the remainder) a 256-bin histogram of every decim_step-th pixel of every
 * decim_step-th row of the sixteenth plane (decim_step 1 with scene-change detection, else 4), bins starting at 1 and
 * scaled by 16 * decim_step^2, the rounded average intensity of the region, and the picture's avg_luma.  The plane is the
 * one svtme_picture_register built, so nothing is uploaded.  Synchronous.
 */
typedef struct SvtMePicStats {
    uint32_t picture_histogram[4][4][256];          /* [region_in_picture_width_index][region_in_picture_height_index][bin] */
    uint64_t average_intensity_per_region[4][4];
    uint64_t avg_luma;
} SvtMePicStats;
int32_t svtme_picture_statistics(SvtMeHandle *h, SvtMePic *pic, uint32_t regions_per_width, uint32_t regions_per_height,
                                 uint32_t decim_step, SvtMePicStats *out);
/*
 * TPL source-based search.  Replaces: the `tpl_src_data_ready == 0` decision of
 * tpl_mc_flow_dispenser_sb_generic (src_ops_process.c:608-884) for every TPL block of a picture -- the direct consumer of
 * me_mv_array / me_candidate_array:
 *   intra: the DC predictor of the block from its source neighbours (intra_dc_sad_path, :623-660: get_neighbor_samples_dc /
 *          svt_aom_update_neighbor_samples_array_open_loop_mb with the 127 / 129 fill outside the picture,
 *          svt_aom_dc_pred[x > 0][y > 0], intra_prediction.c:1034-1073, 2588) and its SAD against the source;
 *   inter: every single-reference ME candidate of the co-located ME block (:759-772), full-pel MV clamped so that the block
 *          stays within TPL_PADX / TPL_PADY = 32 pixels of the reference (:792-802), svt_nxm_sad_kernel against the
 *          reference's source plane (:851-858), first-best-wins (:874-883);
 *   decision: NEWMV when best_inter_cost < best_intra_cost, else DC_PRED (:886-887).
 * Blocks are 16x16 / 32x32 / 64x64 (dispenser_search_level 0 / 1 / 2) inside complete 64x64 blocks and 16x16 inside incomplete
 * ones (:2048-2050); a block less than half inside the picture is skipped (:584-586).  Results are indexed like the
 * reference's tpl_src_stats_buffer: cell (y >> 4) * grid_width + (x >> 4) of the block origin, grid_width =
 * (width + 15) >> 4 (:561, 603-604).
 *
 * Supported controls: use_sad_in_src_search = 1, subpel_depth = FULL_PEL, intra_mode_end = DC_PRED, in_loop_ois = 1 -- TPL
 * levels 4 and 5 of set_tpl_params (initial_rc_process.c:351-374), i.e. presets M3..M13 of v3.0.2 (get_tpl_params_level,
 * :279-292).  Anything else (SATD cost, sub-pel refinement, the PAETH intra sweep of M0-M2) is refused with an error, never
 * approximated.  The transform / quantisation that follows the decision (srcrf_dist, :888-956) stays on the host: one block
 * per decision instead of one per candidate.
 */
#define SVTME_TPL_DC_PRED 0   /* PredictionMode DC_PRED (definitions.h:1127) */
#define SVTME_TPL_NEWMV  16   /* PredictionMode NEWMV   (definitions.h:1143) */
typedef struct SvtMeTplCfg {
    uint8_t  dispenser_search_level;  /* tpl_ctrls.dispenser_search_level */
    uint8_t  use_sad_in_src_search;   /* must be 1 */
    uint8_t  subpel_full_pel;         /* must be 1: tpl_ctrls.subpel_depth == FULL_PEL */
    uint8_t  intra_dc_only;           /* must be 1: tpl_ctrls.intra_mode_end == DC_PRED && scs->in_loop_ois */
    uint8_t  disable_intra_pred;      /* tpl_ctrls.disable_intra_pred_nref && temporal_layer_index == hierarchical_levels (:567) */
    uint8_t  slice_type_i;            /* pcs->slice_type == I_SLICE: no inter candidates (:767) */
    uint8_t  ref_excluded[SVTME_MAX_LISTS][SVTME_MAX_REFS_PER_LIST]; /* ref_tpl_group_idx > 0 && !base_pcs->tpl_valid_pic[idx] (:783-785) */
    uint64_t ref_poc[SVTME_MAX_LISTS][SVTME_MAX_REFS_PER_LIST];      /* tpl_ref_ds_ptr_array[l][r].picture_number (:879) */
} SvtMeTplCfg;
typedef struct SvtMeTplResults {
    uint32_t  grid_width;       /* in: capacity check, (width + 15) >> 4 */
    uint32_t  grid_height;      /* in: (height + 15) >> 4 */
    uint8_t  *block_size;       /* [cell] 16 / 32 / 64 where a searched block starts, else 0 (the other arrays are 0 there) */
    uint8_t  *best_mode;        /* [cell] SVTME_TPL_DC_PRED / SVTME_TPL_NEWMV      -> TplSrcStats.best_mode */
    int8_t   *best_rf_idx;      /* [cell] list * 4 + index of the winner, -1: none -> TplSrcStats.best_rf_idx */
    int16_t  *mv;               /* [cell][2] row, col in 1/8 pel                   -> TplSrcStats.mv */
    uint64_t *ref_frame_poc;    /* [cell]                                          -> TplSrcStats.ref_frame_poc */
    uint32_t *best_inter_cost;  /* [cell] 0xFFFFFFFF: no candidate (INT64_MAX in the reference) */
    uint32_t *best_intra_cost;  /* [cell] 0xFFFFFFFF: intra prediction disabled */
} SvtMeTplResults;
/* Open-loop ME of the picture (exactly svtme_picture_search) followed, on the same stream, by the TPL source-based search over
 * the ME results while they are still in HBM: TPL's references are the ME references (tpl_regular_setup_me_refs,
 * src_ops_process.c:111-144 links the same EbPaReferenceObject.input_padded_pic).  `out` may be NULL when only the TPL
 * decisions are wanted.  Asynchronous; both result sets are filled when svtme_search_wait() returns. */
int32_t svtme_picture_search_tpl(SvtMeHandle *h, SvtMePic *cur, SvtMePic *const refs[SVTME_MAX_LISTS][SVTME_MAX_REFS_PER_LIST],
                                 const SvtMeKnobs *knobs, SvtMePictureResults *out, const SvtMeTplCfg *tpl, SvtMeTplResults *tpl_out,
                                 uint64_t *ticket);
/* The same search over ME results the caller already holds (`me`: me_mv, me_cand, total_me_candidate_index of a finished
 * svtme_picture_search with the same knobs; they are uploaded).  For a TPL pass that runs long after ME.  Asynchronous. */
int32_t svtme_tpl_src_search(SvtMeHandle *h, SvtMePic *cur, SvtMePic *const refs[SVTME_MAX_LISTS][SVTME_MAX_REFS_PER_LIST],
                             const SvtMeKnobs *knobs, const SvtMePictureResults *me, const SvtMeTplCfg *tpl, SvtMeTplResults *tpl_out,
                             uint64_t *ticket);
/* Diagnostic: measured peak issue rate of the packed-byte SAD instruction (VABSDIFF4.U8.ACC, 4 pixel-compares per
 * lane) on this device, in 1e9 instructions per second; the INT-roofline denominator of DESIGN.md. */
int32_t svtme_measure_sad_peak(SvtMeHandle *h, double *gsad4_per_s);
/* Kernel launches issued by this handle so far (bench.py's gpu_launches). */
uint64_t svtme_launch_count(SvtMeHandle *h);
/* Last CUDA error string of this handle (for diagnostics). */
const char *svtme_last_error(SvtMeHandle *h);

/*
 * RTCD-level entry points (host pointers, synchronous).  Signatures are those
 * of the reference's function-pointer table so the reference's kernel unit
 * tests (test/SadTest.cc, test/PictureOperatorTest.cc) can be pointed at them;
 * a per-call offload is NOT the fast path.
 */
/* aom_dsp_rtcd.h:779 svt_sad_loop_kernel */
void svtme_cuda_sad_loop_kernel(SvtMeHandle *h, const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                                uint32_t ref_stride, uint32_t block_height, uint32_t block_width, uint64_t *best_sad,
                                int16_t *x_search_center, int16_t *y_search_center, uint32_t src_stride_raw,
                                uint8_t skip_search_line, int16_t search_area_width, int16_t search_area_height);
/* aom_dsp_rtcd.h:841 downsample_2d */
void svtme_cuda_downsample_2d(SvtMeHandle *h, const uint8_t *input_samples, uint32_t input_stride,
                              uint32_t input_area_width, uint32_t input_area_height, uint8_t *decim_samples,
                              uint32_t decim_stride, uint32_t decim_step);
/* aom_dsp_rtcd.h:856 svt_nxm_sad_kernel */
uint32_t svtme_cuda_nxm_sad_kernel(SvtMeHandle *h, const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                                   uint32_t ref_stride, uint32_t height, uint32_t width);
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * svtme_cuda_nxm_sad_kernel, but the operand staging and the result word live in a file-scope device array of the
 * library (33 024 bytes per GPU, held until exit) instead of the handle's cudaMalloc'ed scratch: the call allocates
 * nothing.  Limit: width <= 64, height <= 128; larger blocks return SVTME_ERR_BAD_PARAMETER (text in
 * svtme_last_error).  All calls share the array: host callers are serialised by a mutex and the call returns after
 * its work has left the GPU; calls made to overlap on the GPU by other means would race. */
int32_t svtme_cuda_nxm_sad_kernel_noalloc(SvtMeHandle *h, const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                                          uint32_t ref_stride, uint32_t height, uint32_t width, uint32_t *sad);
/* open_loop_me_fullpel_search_sblock (motion_estimation.c:781-817) = the fused
 * svt_initialize_buffer_32bits + svt_ext_all_sad_calculation_8x8_16x16 +
 * svt_ext_eight_sad_calculation_32x32_64x64 + svt_ext_sad_calculation_* loop
 * (aom_dsp_rtcd.h:842-853) for one 64x64 block and one search window.
 * best_sad/best_mv: [85] in tier-zero order, initialised by the caller. */
void svtme_cuda_fullpel_search_sblock(SvtMeHandle *h, const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                                      uint32_t ref_stride, int16_t x_search_area_origin, int16_t y_search_area_origin,
                                      uint32_t search_area_width, uint32_t search_area_height, uint8_t sub_sad,
                                      uint32_t *best_sad, uint32_t *best_mv);

#ifdef __cplusplus
}
#endif
#endif /* SVTME_CUDA_H */
