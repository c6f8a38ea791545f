/*
 * ref_harness.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Thin driver around the UNMODIFIED reference sources compiled from
 * /root/reference (see oracle/Makefile): it fills a zero-initialised
 * PictureParentControlSet / SequenceControlSet / MeContext with just the
 * fields motion_estimation.c touches and calls the reference's own
 *   svt_aom_sig_deriv_me            (enc_mode_config.c:671)
 *   svt_aom_motion_estimation_b64   (motion_estimation.c:3076)
 *   svt_aom_downsample_2d_c / svt_aom_generate_padding
 * per SB, mirroring the SB loop of me_process.c:174-290.  The result is
 * written in the flat layout of include/svtme_cuda.h so tests can memcmp it
 * against the oracle restatement and the CUDA path.
 *
 * Built into oracle/_ref/libsvtref.so.  Nothing here is copied from the
 * reference: it only #includes its headers at build time.
 */
#define AOM_RTCD_C /* instantiate the RTCD function-pointer table in this TU (aom_dsp_rtcd.h:25) */
#define RTCD_C     /* same for common_dsp_rtcd.h:42 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

#include "aom_dsp_rtcd.h"
#include "common_dsp_rtcd.h"
#include "pcs.h"
#include "sequence_control_set.h"
#include "motion_estimation.h"
#include "me_context.h"
#include "compute_sad.h"
#include "me_sad_calculation.h"
#include "pic_operators.h"
#include "enc_mode_config.h"

#include "../include/svtme_cuda.h"
#include "ref_harness.h"

/* AVX2 twins, compiled from ASM_AVX2/compute_sad_intrin_avx2.c when the build enables them. */
#ifdef SVTREF_HAVE_AVX2
void svt_sad_loop_kernel_avx2_intrin(uint8_t *src, uint32_t src_stride, uint8_t *ref, uint32_t ref_stride,
                                     uint32_t block_height, uint32_t block_width, uint64_t *best_sad,
                                     int16_t *x_search_center, int16_t *y_search_center, uint32_t src_stride_raw,
                                     uint8_t skip_search_line, int16_t search_area_width, int16_t search_area_height);
#endif

static int g_isa = 0;

/* Bind the eight RTCD rows of the hot path (aom_dsp_rtcd.c:501,509-515). */
void svt_aom_downsample_2d_avx2(uint8_t *input_samples, uint32_t input_stride, uint32_t input_area_width,
                                uint32_t input_area_height, uint8_t *decim_samples, uint32_t decim_stride, uint32_t decim_step);
void svt_sad_loop_kernel_avx512_intrin(uint8_t *src, uint32_t src_stride, uint8_t *ref, uint32_t ref_stride, uint32_t block_height,
                                       uint32_t block_width, uint64_t *best_sad, int16_t *x_search_center, int16_t *y_search_center,
                                       uint32_t src_stride_raw, uint8_t skip_search_line, int16_t search_area_width,
                                       int16_t search_area_height);
int svtref_set_isa(int isa) {
    svt_sad_loop_kernel                       = svt_sad_loop_kernel_c;
    downsample_2d                             = svt_aom_downsample_2d_c;
    svt_ext_sad_calculation_8x8_16x16         = svt_ext_sad_calculation_8x8_16x16_c;
    svt_ext_sad_calculation_32x32_64x64       = svt_ext_sad_calculation_32x32_64x64_c;
    svt_ext_all_sad_calculation_8x8_16x16     = svt_ext_all_sad_calculation_8x8_16x16_c;
    svt_ext_eight_sad_calculation_32x32_64x64 = svt_ext_eight_sad_calculation_32x32_64x64_c;
    svt_initialize_buffer_32bits              = svt_initialize_buffer_32bits_c;
    svt_nxm_sad_kernel                        = svt_nxm_sad_kernel_helper_c;
    svt_memcpy                                = svt_memcpy_c;
    g_isa                                     = 0;
#ifdef SVTREF_HAVE_AVX2
    if (isa >= 1 && __builtin_cpu_supports("avx2")) {
        svt_sad_loop_kernel                       = svt_sad_loop_kernel_avx2_intrin;
        downsample_2d                             = svt_aom_downsample_2d_avx2; /* aom_dsp_rtcd.c:509 */
        /* isa 2: what the RTCD binds on an AVX-512 host (aom_dsp_rtcd.c:501) */
        if (isa == 2 && __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw"))
            svt_sad_loop_kernel                   = svt_sad_loop_kernel_avx512_intrin;
        svt_ext_sad_calculation_8x8_16x16         = svt_ext_sad_calculation_8x8_16x16_avx2_intrin;
        svt_ext_all_sad_calculation_8x8_16x16     = svt_ext_all_sad_calculation_8x8_16x16_avx2;
        svt_ext_eight_sad_calculation_32x32_64x64 = svt_ext_eight_sad_calculation_32x32_64x64_avx2;
        svt_nxm_sad_kernel                        = svt_nxm_sad_kernel_helper_avx2;
        g_isa                                     = svt_sad_loop_kernel == svt_sad_loop_kernel_avx512_intrin ? 2 : 1;
    }
#else
    (void)isa;
#endif
    return g_isa;
}
int svtref_get_isa(void) { return g_isa; }

__attribute__((constructor)) static void svtref_ctor(void) { svtref_set_isa(0); }

/* ---- kernel-level pass-throughs (the reference's own RTCD targets) ---- */
void svtref_sad_loop_kernel(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                            uint32_t block_height, uint32_t block_width, uint64_t *best_sad, int16_t *x, int16_t *y,
                            uint32_t src_stride_raw, uint8_t skip_search_line, int16_t sa_w, int16_t sa_h) {
    svt_sad_loop_kernel((uint8_t *)src, src_stride, (uint8_t *)ref, ref_stride, block_height, block_width, best_sad,
                        x, y, src_stride_raw, skip_search_line, sa_w, sa_h);
}
uint32_t svtref_nxm_sad_kernel(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                               uint32_t height, uint32_t width) {
    return svt_nxm_sad_kernel(src, src_stride, ref, ref_stride, height, width);
}
void svtref_downsample_2d(const uint8_t *in, uint32_t in_stride, uint32_t w, uint32_t h, uint8_t *out,
                          uint32_t out_stride, uint32_t step) {
    downsample_2d((uint8_t *)in, in_stride, w, h, out, out_stride, step);
}
void svtref_generate_padding(uint8_t *pic, uint32_t stride, uint32_t w, uint32_t h, uint32_t pad_w, uint32_t pad_h) {
    svt_aom_generate_padding(pic, stride, w, h, pad_w, pad_h);
}

/* Pull the reference's motion_estimation.c into this TU (unmodified, from where
 * it lies) so that its static functions (open_loop_me_fullpel_search_sblock,
 * hme_level_0, ...) can be driven directly by kernel-level parity tests. */
#include "motion_estimation.c"

void svtref_fullpel_search_sblock(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                                  int16_t x_org, int16_t y_org, uint32_t sa_w, uint32_t sa_h, uint8_t sub_sad,
                                  uint32_t *best_sad, uint32_t *best_mv) {
    /* `ref` points at the top-left candidate; the reference adds the ME_FILTER_TAP/2
     * lead-in back itself (motion_estimation.c:444-446), so rewind by it here. */
    MeContext *ctx = (MeContext *)calloc(1, sizeof(MeContext));
    ctx->me_search_method                = sub_sad ? SUB_SAD_SEARCH : FULL_SAD_SEARCH;
    ctx->b64_src_ptr                     = (uint8_t *)src;
    ctx->b64_src_stride                  = src_stride;
    ctx->interpolated_full_stride[0][0]  = ref_stride;
    ctx->integer_buffer_ptr[0][0]        = (uint8_t *)ref - (ME_FILTER_TAP >> 1) - (ME_FILTER_TAP >> 1) * ref_stride;
    memcpy(ctx->p_sb_best_sad[0][0], best_sad, sizeof(uint32_t) * SQUARE_PU_COUNT);
    memcpy(ctx->p_sb_best_mv[0][0], best_mv, sizeof(uint32_t) * SQUARE_PU_COUNT);
    ctx->p_best_sad_64x64 = &ctx->p_sb_best_sad[0][0][ME_TIER_ZERO_PU_64x64];
    ctx->p_best_sad_32x32 = &ctx->p_sb_best_sad[0][0][ME_TIER_ZERO_PU_32x32_0];
    ctx->p_best_sad_16x16 = &ctx->p_sb_best_sad[0][0][ME_TIER_ZERO_PU_16x16_0];
    ctx->p_best_sad_8x8   = &ctx->p_sb_best_sad[0][0][ME_TIER_ZERO_PU_8x8_0];
    ctx->p_best_mv64x64   = &ctx->p_sb_best_mv[0][0][ME_TIER_ZERO_PU_64x64];
    ctx->p_best_mv32x32   = &ctx->p_sb_best_mv[0][0][ME_TIER_ZERO_PU_32x32_0];
    ctx->p_best_mv16x16   = &ctx->p_sb_best_mv[0][0][ME_TIER_ZERO_PU_16x16_0];
    ctx->p_best_mv8x8     = &ctx->p_sb_best_mv[0][0][ME_TIER_ZERO_PU_8x8_0];
    open_loop_me_fullpel_search_sblock(ctx, 0, 0, x_org, y_org, sa_w, sa_h);
    memcpy(best_sad, ctx->p_sb_best_sad[0][0], sizeof(uint32_t) * SQUARE_PU_COUNT);
    memcpy(best_mv, ctx->p_sb_best_mv[0][0], sizeof(uint32_t) * SQUARE_PU_COUNT);
    free(ctx);
}

/* ---- knob transfer between the flat ABI struct and the reference's MeContext / pcs / scs ---- */
static void area_to(SvtMeArea *d, const SearchArea *s) { d->width = s->width; d->height = s->height; }
static void area_from(SearchArea *d, const SvtMeArea *s) { d->width = s->width; d->height = s->height; }

static void knobs_from_ctx(SvtMeKnobs *k, const MeContext *c) {
    k->hme_search_method      = c->hme_search_method;
    k->me_search_method       = c->me_search_method;
    k->enable_hme_flag        = c->enable_hme_flag;
    k->enable_hme_level0_flag = c->enable_hme_level0_flag;
    k->enable_hme_level1_flag = c->enable_hme_level1_flag;
    k->enable_hme_level2_flag = c->enable_hme_level2_flag;
    k->num_hme_sa_w           = c->num_hme_sa_w;
    k->num_hme_sa_h           = c->num_hme_sa_h;
    area_to(&k->hme_l0_sa.sa_min, &c->hme_l0_sa.sa_min);
    area_to(&k->hme_l0_sa.sa_max, &c->hme_l0_sa.sa_max);
    area_to(&k->hme_l1_sa, &c->hme_l1_sa);
    area_to(&k->hme_l2_sa, &c->hme_l2_sa);
    area_to(&k->me_sa.sa_min, &c->me_sa.sa_min);
    area_to(&k->me_sa.sa_max, &c->me_sa.sa_max);
    k->enable_me_hme_ref_pruning               = c->me_hme_prune_ctrls.enable_me_hme_ref_pruning;
    k->prune_ref_if_hme_sad_dev_bigger_than_th = c->me_hme_prune_ctrls.prune_ref_if_hme_sad_dev_bigger_than_th;
    k->prune_ref_if_me_sad_dev_bigger_than_th  = c->me_hme_prune_ctrls.prune_ref_if_me_sad_dev_bigger_than_th;
    k->zz_sad_th                               = c->me_hme_prune_ctrls.zz_sad_th;
    k->zz_sad_pct                              = c->me_hme_prune_ctrls.zz_sad_pct;
    k->phme_sad_th                             = c->me_hme_prune_ctrls.phme_sad_th;
    k->phme_sad_pct                            = c->me_hme_prune_ctrls.phme_sad_pct;
    k->enable_me_sr_adjustment              = c->me_sr_adjustment_ctrls.enable_me_sr_adjustment;
    k->reduce_me_sr_based_on_mv_length_th   = c->me_sr_adjustment_ctrls.reduce_me_sr_based_on_mv_length_th;
    k->stationary_hme_sad_abs_th            = c->me_sr_adjustment_ctrls.stationary_hme_sad_abs_th;
    k->stationary_me_sr_divisor             = c->me_sr_adjustment_ctrls.stationary_me_sr_divisor;
    k->reduce_me_sr_based_on_hme_sad_abs_th = c->me_sr_adjustment_ctrls.reduce_me_sr_based_on_hme_sad_abs_th;
    k->me_sr_divisor_for_low_hme_sad        = c->me_sr_adjustment_ctrls.me_sr_divisor_for_low_hme_sad;
    k->distance_based_hme_resizing          = c->me_sr_adjustment_ctrls.distance_based_hme_resizing;
    k->me_8x8_var_enabled = c->me_8x8_var_ctrls.enabled;
    k->me_sr_div4_th      = c->me_8x8_var_ctrls.me_sr_div4_th;
    k->me_sr_div2_th      = c->me_8x8_var_ctrls.me_sr_div2_th;
    k->me_sr_mult2_th     = c->me_8x8_var_ctrls.me_sr_mult2_th;
    k->mv_sa_adj_enabled          = c->mv_based_sa_adj.enabled;
    k->mv_sa_adj_nearest_ref_only = c->mv_based_sa_adj.nearest_ref_only;
    k->mv_sa_adj_mv_size_th       = c->mv_based_sa_adj.mv_size_th;
    k->mv_sa_adj_sa_multiplier    = c->mv_based_sa_adj.sa_multiplier;
    k->prehme_enable = c->prehme_ctrl.enable;
    for (int i = 0; i < 2; i++) {
        area_to(&k->prehme_sa_cfg[i].sa_min, &c->prehme_ctrl.prehme_sa_cfg[i].sa_min);
        area_to(&k->prehme_sa_cfg[i].sa_max, &c->prehme_ctrl.prehme_sa_cfg[i].sa_max);
    }
    k->prehme_skip_search_line     = c->prehme_ctrl.skip_search_line;
    k->prehme_l1_early_exit        = c->prehme_ctrl.l1_early_exit;
    k->prune_me_candidates_th      = c->prune_me_candidates_th;
    k->use_best_unipred_cand_only  = c->use_best_unipred_cand_only;
    k->reduce_hme_l0_sr_th_min     = c->reduce_hme_l0_sr_th_min;
    k->reduce_hme_l0_sr_th_max     = c->reduce_hme_l0_sr_th_max;
    k->me_early_exit_th            = c->me_early_exit_th;
    k->me_safe_limit_zz_th         = c->me_safe_limit_zz_th;
    k->prev_me_stage_based_exit_th = c->prev_me_stage_based_exit_th;
    k->me_type                     = c->me_type == ME_MCTF ? SVTME_ME_MCTF : SVTME_ME_OPEN_LOOP;
    k->tf_me_exit_th               = c->tf_me_exit_th;
}

static void ctx_from_knobs(MeContext *c, const SvtMeKnobs *k) {
    c->hme_search_method      = k->hme_search_method;
    c->me_search_method       = k->me_search_method;
    c->enable_hme_flag        = k->enable_hme_flag;
    c->enable_hme_level0_flag = k->enable_hme_level0_flag;
    c->enable_hme_level1_flag = k->enable_hme_level1_flag;
    c->enable_hme_level2_flag = k->enable_hme_level2_flag;
    c->num_hme_sa_w           = k->num_hme_sa_w;
    c->num_hme_sa_h           = k->num_hme_sa_h;
    area_from(&c->hme_l0_sa.sa_min, &k->hme_l0_sa.sa_min);
    area_from(&c->hme_l0_sa.sa_max, &k->hme_l0_sa.sa_max);
    area_from(&c->hme_l1_sa, &k->hme_l1_sa);
    area_from(&c->hme_l2_sa, &k->hme_l2_sa);
    area_from(&c->me_sa.sa_min, &k->me_sa.sa_min);
    area_from(&c->me_sa.sa_max, &k->me_sa.sa_max);
    c->me_hme_prune_ctrls.enable_me_hme_ref_pruning               = k->enable_me_hme_ref_pruning;
    c->me_hme_prune_ctrls.prune_ref_if_hme_sad_dev_bigger_than_th = k->prune_ref_if_hme_sad_dev_bigger_than_th;
    c->me_hme_prune_ctrls.prune_ref_if_me_sad_dev_bigger_than_th  = k->prune_ref_if_me_sad_dev_bigger_than_th;
    c->me_hme_prune_ctrls.zz_sad_th                               = k->zz_sad_th;
    c->me_hme_prune_ctrls.zz_sad_pct                              = k->zz_sad_pct;
    c->me_hme_prune_ctrls.phme_sad_th                             = k->phme_sad_th;
    c->me_hme_prune_ctrls.phme_sad_pct                            = k->phme_sad_pct;
    c->me_sr_adjustment_ctrls.enable_me_sr_adjustment              = k->enable_me_sr_adjustment;
    c->me_sr_adjustment_ctrls.reduce_me_sr_based_on_mv_length_th   = k->reduce_me_sr_based_on_mv_length_th;
    c->me_sr_adjustment_ctrls.stationary_hme_sad_abs_th            = k->stationary_hme_sad_abs_th;
    c->me_sr_adjustment_ctrls.stationary_me_sr_divisor             = k->stationary_me_sr_divisor;
    c->me_sr_adjustment_ctrls.reduce_me_sr_based_on_hme_sad_abs_th = k->reduce_me_sr_based_on_hme_sad_abs_th;
    c->me_sr_adjustment_ctrls.me_sr_divisor_for_low_hme_sad        = k->me_sr_divisor_for_low_hme_sad;
    c->me_sr_adjustment_ctrls.distance_based_hme_resizing          = k->distance_based_hme_resizing;
    c->me_8x8_var_ctrls.enabled        = k->me_8x8_var_enabled;
    c->me_8x8_var_ctrls.me_sr_div4_th  = k->me_sr_div4_th;
    c->me_8x8_var_ctrls.me_sr_div2_th  = k->me_sr_div2_th;
    c->me_8x8_var_ctrls.me_sr_mult2_th = k->me_sr_mult2_th;
    c->mv_based_sa_adj.enabled          = k->mv_sa_adj_enabled;
    c->mv_based_sa_adj.nearest_ref_only = k->mv_sa_adj_nearest_ref_only;
    c->mv_based_sa_adj.mv_size_th       = k->mv_sa_adj_mv_size_th;
    c->mv_based_sa_adj.sa_multiplier    = k->mv_sa_adj_sa_multiplier;
    c->prehme_ctrl.enable = k->prehme_enable;
    for (int i = 0; i < 2; i++) {
        area_from(&c->prehme_ctrl.prehme_sa_cfg[i].sa_min, &k->prehme_sa_cfg[i].sa_min);
        area_from(&c->prehme_ctrl.prehme_sa_cfg[i].sa_max, &k->prehme_sa_cfg[i].sa_max);
    }
    c->prehme_ctrl.skip_search_line = k->prehme_skip_search_line;
    c->prehme_ctrl.l1_early_exit    = k->prehme_l1_early_exit;
    c->prune_me_candidates_th       = k->prune_me_candidates_th;
    c->use_best_unipred_cand_only   = k->use_best_unipred_cand_only;
    c->reduce_hme_l0_sr_th_min      = k->reduce_hme_l0_sr_th_min;
    c->reduce_hme_l0_sr_th_max      = k->reduce_hme_l0_sr_th_max;
    c->me_early_exit_th             = k->me_early_exit_th;
    c->me_safe_limit_zz_th          = k->me_safe_limit_zz_th;
    c->prev_me_stage_based_exit_th  = k->prev_me_stage_based_exit_th;
    c->me_type       = k->me_type == SVTME_ME_MCTF ? ME_MCTF : ME_OPEN_LOOP; /* temporal_filtering.c:296 / me_process.c:216 */
    c->tf_me_exit_th = k->tf_me_exit_th;
    c->num_of_list_to_search        = k->num_of_list_to_search;
    c->num_of_ref_pic_to_search[0]  = k->num_of_ref_pic_to_search[0];
    c->num_of_ref_pic_to_search[1]  = k->num_of_ref_pic_to_search[1];
    c->temporal_layer_index         = k->temporal_layer_index;
    c->is_ref                       = k->is_ref;
}

/* mrp level selection + static set_mrp_ctrl live in Globals/enc_handle.c; see ref_mrp.c */
void svtref_set_mrp(SequenceControlSet *scs, int enc_mode, int rtc_tune, int rate_control_vbr, int fast_decode);

/* Build scs/pcs the way the encoder would for a preset, then let the reference derive the ME knobs. */
int svtref_derive_knobs(const SvtMePresetCfg *cfg, SvtMeKnobs *k) {
    SequenceControlSet      *scs = (SequenceControlSet *)calloc(1, sizeof(*scs));
    PictureParentControlSet *pcs = (PictureParentControlSet *)calloc(1, sizeof(*pcs));
    MeContext               *ctx = (MeContext *)calloc(1, sizeof(*ctx));
    if (!scs || !pcs || !ctx) return -1;
    memset(k, 0, sizeof(*k));
    pcs->scs = scs;
    scs->static_config.pred_structure = cfg->rtc_tune ? SVT_AV1_PRED_LOW_DELAY_B : SVT_AV1_PRED_RANDOM_ACCESS;
    scs->static_config.qp             = cfg->qp;
    scs->static_config.enc_mode       = cfg->enc_mode;
    scs->static_config.fast_decode    = cfg->fast_decode;
    scs->frame_rate                   = cfg->frame_rate_q16;
    svt_aom_derive_input_resolution(&scs->input_resolution, (uint32_t)cfg->width * cfg->height);
    svtref_set_mrp(scs, cfg->enc_mode, cfg->rtc_tune, cfg->rate_control_vbr, cfg->fast_decode);

    pcs->enc_mode             = cfg->enc_mode;
    pcs->sc_class1            = cfg->sc_class1;
    pcs->input_resolution     = scs->input_resolution;
    pcs->hierarchical_levels  = cfg->hierarchical_levels;
    pcs->temporal_layer_index = cfg->temporal_layer_index;
    pcs->slice_type           = cfg->slice_type_b ? B_SLICE : P_SLICE;
    pcs->picture_number       = cfg->picture_number;
    pcs->is_ref               = cfg->is_ref;
    /* HME flags: svt_aom_sig_deriv_multi_processes, enc_mode_config.c:1608-1619 (too entangled to call) */
    pcs->enable_hme_flag        = 1;
    pcs->enable_hme_level0_flag = 1;
    pcs->enable_hme_level1_flag = 1;
    pcs->enable_hme_level2_flag = (cfg->sc_class1 || cfg->enc_mode <= ENC_M6) ? 1 : 0;
    /* enc_mode_config.c:1802-1805 */
    pcs->use_best_me_unipred_cand_only = cfg->enc_mode <= ENC_M3 ? 0 : 1;
    /* pcs.c:1389-1394 */
    pcs->enable_me_16x16 = svt_aom_get_enable_me_16x16(cfg->enc_mode);
    pcs->enable_me_8x8   = pcs->enable_me_16x16
          ? svt_aom_get_enable_me_8x8(cfg->enc_mode, cfg->rtc_tune, scs->input_resolution)
          : 0;
    pcs->max_number_of_pus_per_sb = SQUARE_PU_COUNT; /* resource_coordination_process.c:425 */
    /* enc_mode_config.c:1604 */
    svt_aom_set_gm_controls(pcs, svt_aom_derive_gm_level(pcs, true));

    svt_aom_sig_deriv_me(scs, pcs, ctx);
    knobs_from_ctx(k, ctx);

    const MrpCtrls *m = &scs->mrp_ctrls;
    /* pd_process.c:3503-3519 + pcs.c:91-96 */
    uint8_t l0 = MAX(m->sc_base_ref_list0_count,
                     MAX(m->base_ref_list0_count, MAX(m->sc_non_base_ref_list0_count, m->non_base_ref_list0_count)));
    uint8_t l1 = MAX(m->sc_base_ref_list1_count,
                     MAX(m->base_ref_list1_count, MAX(m->sc_non_base_ref_list1_count, m->non_base_ref_list1_count)));
    uint8_t max_refs, max_cand;
    svt_aom_get_max_allocated_me_refs(l0, l1, &max_refs, &max_cand);
    k->max_cand = max_cand;
    k->max_refs = max_refs;
    k->max_l0   = l0;
    /* update_count_try, pd_process.c:3670-3695 (ref_list{0,1}_count = what the caller offers, 0 = plenty) */
    uint8_t c0, c1;
    if (cfg->sc_class1) {
        c0 = cfg->temporal_layer_index == 0 ? m->sc_base_ref_list0_count : m->sc_non_base_ref_list0_count;
        c1 = cfg->temporal_layer_index == 0 ? m->sc_base_ref_list1_count : m->sc_non_base_ref_list1_count;
    } else {
        c0 = cfg->temporal_layer_index == 0 ? m->base_ref_list0_count : m->non_base_ref_list0_count;
        c1 = cfg->temporal_layer_index == 0 ? m->base_ref_list1_count : m->non_base_ref_list1_count;
    }
    if (cfg->ref_list0_count_try) c0 = MIN(c0, cfg->ref_list0_count_try);
    if (cfg->ref_list1_count_try) c1 = MIN(c1, cfg->ref_list1_count_try);
    k->num_of_list_to_search       = cfg->slice_type_b ? 2 : 1; /* me_process.c:219-221 */
    k->num_of_ref_pic_to_search[0] = c0;
    k->num_of_ref_pic_to_search[1] = cfg->slice_type_b ? c1 : 0;
    k->temporal_layer_index        = cfg->temporal_layer_index;
    k->is_ref                      = cfg->is_ref;
    k->picture_number              = cfg->picture_number;
    k->hierarchical_levels         = cfg->hierarchical_levels;
    k->similar_brightness_refs     = cfg->similar_brightness_refs;
    k->input_resolution            = scs->input_resolution;
    k->enable_me_8x8               = pcs->enable_me_8x8;
    k->enable_me_16x16             = pcs->enable_me_16x16;
    k->max_number_of_pus_per_sb    = pcs->max_number_of_pus_per_sb;
    k->gm_enabled                  = pcs->gm_ctrls.enabled;
    k->gm_use_distance_based_active_th = pcs->gm_ctrls.use_distance_based_active_th;
    k->only_l_bwd                  = m->only_l_bwd;
    free(ctx);
    free(pcs);
    free(scs);
    return 0;
}

/* Temporal-filter knobs: the reference's own svt_aom_sig_deriv_me_tf on a pcs carrying the given TfControls, then the
 * per-block overrides of produce_temporally_filtered_pic (temporal_filtering.c:296, 316, 2759-2767, 3140-3158). */
int svtref_derive_tf_knobs(const SvtMeTfCfg *cfg, SvtMeKnobs *k) {
    SequenceControlSet      *scs = (SequenceControlSet *)calloc(1, sizeof(*scs));
    PictureParentControlSet *pcs = (PictureParentControlSet *)calloc(1, sizeof(*pcs));
    MeContext               *ctx = (MeContext *)calloc(1, sizeof(*ctx));
    if (!scs || !pcs || !ctx) return -1;
    memset(k, 0, sizeof(*k));
    pcs->scs                  = scs;
    scs->static_config.qp     = cfg->qp;
    svt_aom_derive_input_resolution(&scs->input_resolution, (uint32_t)cfg->width * cfg->height);
    pcs->tf_ctrls.hme_me_level = cfg->hme_me_level;
    pcs->tf_ctrls.qp_opt       = cfg->qp_opt;
    pcs->tf_ctrls.me_exit_th   = cfg->me_exit_th;
    /* enc_mode_config.c:1620-1643 (inside svt_aom_sig_deriv_multi_processes, too entangled to call) */
    pcs->tf_enable_hme_flag        = 1;
    pcs->tf_enable_hme_level0_flag = 1;
    pcs->tf_enable_hme_level1_flag = cfg->hme_me_level <= 2;
    pcs->tf_enable_hme_level2_flag = cfg->hme_me_level == 0;
    svt_aom_sig_deriv_me_tf(pcs, ctx);
    ctx->me_type           = ME_MCTF;                  /* temporal_filtering.c:296 */
    ctx->hme_search_method = FULL_SAD_SEARCH;          /* :316 */
    ctx->hme_l0_sa         = ctx->hme_l0_sa_default_tf; /* set_hme_search_params_mctf(ctx, 0), :2759-2767 */
    ctx->tf_me_exit_th     = pcs->tf_ctrls.me_exit_th; /* :3158 */
    knobs_from_ctx(k, ctx);
    k->num_of_list_to_search       = 1; /* :3140-3142 */
    k->num_of_ref_pic_to_search[0] = 1;
    k->temporal_layer_index        = cfg->temporal_layer_index;
    k->is_ref                      = cfg->is_ref;
    k->picture_number              = cfg->picture_number;
    k->input_resolution            = scs->input_resolution;
    k->enable_me_8x8 = 1; k->enable_me_16x16 = 1; k->max_number_of_pus_per_sb = SQUARE_PU_COUNT;
    k->max_cand = 1; k->max_refs = 1; k->max_l0 = 1;
    free(ctx);
    free(pcs);
    free(scs);
    return 0;
}

/* ---- picture-level driver: the SB loop of me_process.c:174-290 over the whole picture ---- */
typedef struct RefJob {
    const SvtMeKnobs        *k;
    const SvtRefPic         *cur;
    const SvtRefPic         *refs; /* [2][4] */
    SvtMePictureResults     *out;
    PictureParentControlSet *pcs;
    uint32_t                 pic_w_b64, pic_h_b64;
    volatile uint32_t        next_sb;
    uint64_t                 px_compares; /* not counted by the reference; left 0 */
} RefJob;

static void desc_from_plane(EbPictureBufferDesc *d, const SvtRefPlane *p) {
    memset(d, 0, sizeof(*d));
    d->buffer_y = (uint8_t *)p->buf;
    d->stride_y = (uint16_t)p->stride;
    d->org_x    = (uint16_t)p->org_x;
    d->org_y    = (uint16_t)p->org_y;
    d->width    = (uint16_t)p->width;
    d->height   = (uint16_t)p->height;
    d->max_width  = d->width;
    d->max_height = d->height;
}

/* one ME thread's view of one picture: its MeContext and the buffer descriptors it points into */
typedef struct RefThread {
    MeContext          *ctx;
    EbPictureBufferDesc cur_d[3], ref_d[2][4][3];
} RefThread;

static void ref_thread_setup(RefThread *t, const RefJob *job) {
    const SvtMeKnobs    *k = job->k;
    MeContext           *ctx = t->ctx;
    EbPictureBufferDesc *cur_d = t->cur_d;
    EbPictureBufferDesc (*ref_d)[4][3] = t->ref_d;
    memset(ctx, 0, sizeof(*ctx));
    ctx_from_knobs(ctx, k);
    for (int l = 0; l < 3; l++) desc_from_plane(&cur_d[l], &job->cur->lvl[l]);
    for (int li = 0; li < k->num_of_list_to_search; li++)
        for (int ri = 0; ri < k->num_of_ref_pic_to_search[li]; ri++) {
            const SvtRefPic *rp = &job->refs[li * 4 + ri];
            for (int l = 0; l < 3; l++) desc_from_plane(&ref_d[li][ri][l], &rp->lvl[l]);
            /* me_process.c:248-262 */
            ctx->me_ds_ref_array[li][ri].picture_ptr           = &ref_d[li][ri][0];
            ctx->me_ds_ref_array[li][ri].quarter_picture_ptr   = &ref_d[li][ri][1];
            ctx->me_ds_ref_array[li][ri].sixteenth_picture_ptr = &ref_d[li][ri][2];
            ctx->me_ds_ref_array[li][ri].picture_number        = rp->picture_number;
        }
}

/* one 64x64 block: the body of the SB loop of me_process.c:174-290 */
static void ref_thread_sb(RefThread *t, RefJob *job, uint32_t sb) {
    const SvtMeKnobs        *k   = job->k;
    PictureParentControlSet *pcs = job->pcs;
    SvtMePictureResults     *out = job->out;
    MeContext               *ctx = t->ctx;
    EbPictureBufferDesc     *cur_d = t->cur_d;
    const uint32_t pus = out->pu_count, mr = k->max_refs, mc = k->max_cand;
    {
        uint32_t x = (sb % job->pic_w_b64) * 64, y = (sb / job->pic_w_b64) * 64;
        /* me_process.c:183-214 */
        ctx->b64_src_ptr    = cur_d[0].buffer_y + (cur_d[0].org_y + y) * cur_d[0].stride_y + cur_d[0].org_x + x;
        ctx->b64_src_stride = cur_d[0].stride_y;
        if (ctx->enable_hme_level1_flag) {
            ctx->quarter_b64_buffer =
                cur_d[1].buffer_y + (cur_d[1].org_y + (y >> 1)) * cur_d[1].stride_y + cur_d[1].org_x + (x >> 1);
            ctx->quarter_b64_buffer_stride = cur_d[1].stride_y;
        }
        if (ctx->enable_hme_level0_flag) {
            ctx->sixteenth_b64_buffer =
                cur_d[2].buffer_y + (cur_d[2].org_y + (y >> 2)) * cur_d[2].stride_y + cur_d[2].org_x + (x >> 2);
            ctx->sixteenth_b64_buffer_stride = cur_d[2].stride_y;
        }
        /* the knob block is re-derived per message in the encoder (me_process.c:121); areas mutated
         * by get_hme_l0_search_area are restored by the reference itself (:1999-2003). */
        ctx->me_type = k->me_type == SVTME_ME_MCTF ? ME_MCTF : ME_OPEN_LOOP;
        ctx->tf_use_pred_64x64_only_th = 1; /* any value but 0xFF: the reference overwrites it per block (temporal_filtering.c:3161) */
        /* harness convention (header of svtme_cuda.h): unsearched refs report MAX_SAD_VALUE, unwritten outputs 0 */
        for (int li = 0; li < 2; li++)
            for (int ri = 0; ri < 4; ri++)
                for (int p = 0; p < SQUARE_PU_COUNT; p++) ctx->p_sb_best_sad[li][ri][p] = MAX_SAD_VALUE;
        if (k->me_type != SVTME_ME_MCTF) {
            MeSbResults *r = pcs->pa_me_data->me_results[sb];
            memset(r->me_mv_array, 0, sizeof(MvCandidate) * pus * mr);
            memset(r->me_candidate_array, 0, sizeof(MeCandidate) * pus * mc);
            memset(r->total_me_candidate_index, 0, pus);
        }

        svt_aom_motion_estimation_b64(pcs, sb, x, y, ctx, &cur_d[0]);

        if (out->tf_exit_sb) out->tf_exit_sb[sb] = ctx->tf_use_pred_64x64_only_th == (uint8_t)~0;
        if (out->sb_best_sad) memcpy(out->sb_best_sad + (size_t)sb * 2 * 4 * 85, ctx->p_sb_best_sad, sizeof(uint32_t) * 2 * 4 * 85);
        if (out->sb_best_mv) memcpy(out->sb_best_mv + (size_t)sb * 2 * 4 * 85, ctx->p_sb_best_mv, sizeof(uint32_t) * 2 * 4 * 85);
        if (out->tf_best_sad) memcpy(out->tf_best_sad + (size_t)sb * 85, ctx->p_sb_best_sad[0][0], sizeof(uint32_t) * 85);
        if (out->tf_best_mv) memcpy(out->tf_best_mv + (size_t)sb * 85, ctx->p_sb_best_mv[0][0], sizeof(uint32_t) * 85);
        for (int li = 0; li < 2; li++)
            for (int ri = 0; ri < 4; ri++) {
                int used = li < k->num_of_list_to_search && ri < k->num_of_ref_pic_to_search[li];
                if (out->do_ref) out->do_ref[(size_t)sb * 8 + li * 4 + ri] = used ? ctx->search_results[li][ri].do_ref : 0;
                if (out->hme_sc) {
                    out->hme_sc[((size_t)sb * 8 + li * 4 + ri) * 2 + 0] = used ? ctx->search_results[li][ri].hme_sc_x : 0;
                    out->hme_sc[((size_t)sb * 8 + li * 4 + ri) * 2 + 1] = used ? ctx->search_results[li][ri].hme_sc_y : 0;
                }
            }
    }
}

static void ref_thread_done(RefThread *t, RefJob *job) {
    SvtMePictureResults *out = job->out;
    MeContext           *ctx = t->ctx;
    if (out->tf_blk_counts) { /* temporal_filtering.c:4224-4227: per-segment counters summed into the centre pcs */
        __sync_fetch_and_add(&out->tf_blk_counts[0], ctx->tf_tot_horz_blks);
        __sync_fetch_and_add(&out->tf_blk_counts[1], ctx->tf_tot_vert_blks);
        ctx->tf_tot_horz_blks = ctx->tf_tot_vert_blks = 0;
    }
}

static void *ref_worker(void *arg) {
    RefJob   *job = (RefJob *)arg;
    RefThread t;
    t.ctx = (MeContext *)malloc(sizeof(MeContext));
    ref_thread_setup(&t, job);
    const uint32_t total = job->pic_w_b64 * job->pic_h_b64;
    for (;;) {
        uint32_t sb = __sync_fetch_and_add(&job->next_sb, 1);
        if (sb >= total) break;
        ref_thread_sb(&t, job, sb);
    }
    ref_thread_done(&t, job);
    free(t.ctx);
    return NULL;
}

/* Picture-level sums (SvtMePictureResults.picture_sums): the loops of Initial RC (initial_rc_process.c:683-691) and of
 * svt_aom_global_motion_estimation (global_me.c:162-168) over the per-block arrays of a completed picture. */
static void picture_sums(const SvtMeKnobs *k, SvtMePictureResults *out, uint32_t n_b64) {
    if (!out->picture_sums || k->me_type == SVTME_ME_MCTF) return;
    uint64_t dist = 0, me_sad = 0, stat = 0, gm = 0;
    for (uint32_t b = 0; b < n_b64; b++) {
        dist += out->me_8x8_distortion[b];
        me_sad += out->rc_me_distortion[b];
        stat += out->stationary_block_present_sb[b];
        gm += out->rc_me_allow_gm[b];
    }
    out->picture_sums[0] = dist; out->picture_sums[1] = me_sad; out->picture_sums[2] = stat; out->picture_sums[3] = gm;
}

static int ref_job_open(RefJob *job, const SvtMeKnobs *k, const SvtRefPic *cur, const SvtRefPic *refs, SvtMePictureResults *out) {
    const uint32_t W = cur->lvl[0].width, H = cur->lvl[0].height;
    const uint32_t wb = (W + 63) / 64, hb = (H + 63) / 64, total = wb * hb;
    if (out->b64_total_count < total) return -2;
    SequenceControlSet      *scs = (SequenceControlSet *)calloc(1, sizeof(*scs));
    PictureParentControlSet *pcs = (PictureParentControlSet *)calloc(1, sizeof(*pcs));
    MotionEstimationData    *med = (MotionEstimationData *)calloc(1, sizeof(*med));
    pcs->scs                       = scs;
    scs->input_resolution          = k->input_resolution;
    scs->mrp_ctrls.only_l_bwd      = k->only_l_bwd;
    pcs->aligned_width             = (uint16_t)((W + 7) & ~7u);
    pcs->aligned_height            = (uint16_t)((H + 7) & ~7u);
    pcs->picture_number            = k->picture_number;
    pcs->hierarchical_levels       = k->hierarchical_levels;
    pcs->temporal_layer_index      = k->temporal_layer_index;
    pcs->similar_brightness_refs   = k->similar_brightness_refs;
    pcs->enable_me_8x8             = k->enable_me_8x8;
    pcs->enable_me_16x16           = k->enable_me_16x16;
    pcs->max_number_of_pus_per_sb  = k->max_number_of_pus_per_sb;
    pcs->gm_ctrls.enabled          = k->gm_enabled;
    pcs->gm_ctrls.use_distance_based_active_th = k->gm_use_distance_based_active_th;
    pcs->pa_me_data                = med;
    med->max_cand                  = k->max_cand;
    med->max_refs                  = k->max_refs;
    med->max_l0                    = k->max_l0;
    med->me_results                = (MeSbResults **)calloc(total, sizeof(MeSbResults *));
    MeSbResults *sbr               = (MeSbResults *)calloc(total, sizeof(MeSbResults));
    const uint32_t pus = out->pu_count;
    for (uint32_t i = 0; i < total; i++) {
        med->me_results[i]              = &sbr[i];
        sbr[i].me_mv_array              = (MvCandidate *)(out->me_mv + (size_t)i * pus * k->max_refs);
        sbr[i].me_candidate_array       = (MeCandidate *)(out->me_cand + (size_t)i * pus * k->max_cand);
        sbr[i].total_me_candidate_index = out->total_me_candidate_index + (size_t)i * pus;
    }
    pcs->b64_geom = (B64Geom *)calloc(total, sizeof(B64Geom));
    for (uint32_t i = 0; i < total; i++) { /* b64 geometry as built in svt_aom_b64_geom_init_pcs: clipped to the aligned picture */
        uint32_t x = (i % wb) * 64, y = (i / wb) * 64;
        pcs->b64_geom[i].org_x  = (uint16_t)x;
        pcs->b64_geom[i].org_y  = (uint16_t)y;
        pcs->b64_geom[i].width  = (uint8_t)MIN(64u, pcs->aligned_width - x);
        pcs->b64_geom[i].height = (uint8_t)MIN(64u, pcs->aligned_height - y);
    }
    pcs->me_64x64_distortion         = out->me_64x64_distortion;
    pcs->me_32x32_distortion         = out->me_32x32_distortion;
    pcs->me_16x16_distortion         = out->me_16x16_distortion;
    pcs->me_8x8_distortion           = out->me_8x8_distortion;
    pcs->me_8x8_cost_variance        = out->me_8x8_cost_variance;
    pcs->rc_me_distortion            = out->rc_me_distortion;
    pcs->stationary_block_present_sb = out->stationary_block_present_sb;
    pcs->rc_me_allow_gm              = out->rc_me_allow_gm;

    if (out->tf_blk_counts) out->tf_blk_counts[0] = out->tf_blk_counts[1] = 0;
    memset(job, 0, sizeof(*job));
    job->k = k; job->cur = cur; job->refs = refs; job->out = out; job->pcs = pcs;
    job->pic_w_b64 = wb; job->pic_h_b64 = hb;
    return 0;
}

static void ref_job_close(RefJob *job) {
    PictureParentControlSet *pcs = job->pcs;
    MotionEstimationData    *med = pcs->pa_me_data;
    picture_sums(job->k, job->out, job->pic_w_b64 * job->pic_h_b64);
    free(pcs->b64_geom);
    free(med->me_results[0]);
    free(med->me_results);
    free(med);
    free(pcs->scs);
    free(pcs);
}

int svtref_me_picture(const SvtMeKnobs *k, const SvtRefPic *cur, const SvtRefPic *refs, SvtMePictureResults *out,
                      int n_threads) {
    RefJob job;
    int    rc = ref_job_open(&job, k, cur, refs, out);
    if (rc) return rc;
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    pthread_t th[256];
    for (int t = 1; t < n_threads; t++) pthread_create(&th[t], NULL, ref_worker, &job);
    ref_worker(&job);
    for (int t = 1; t < n_threads; t++) pthread_join(th[t], NULL);
    ref_job_close(&job);
    return 0;
}

void svtref_build_planes(const SvtRefPic *pic, int hme_level1);

/* ---- a mini-GOP at a time, parallelised the way the encoder is (enc_handle.c:395-398, 706-773; me_process.c:147-157):
 * a pool of ME threads pulls (picture, segment) items -- 6 x 8 segments per picture, several pictures in flight -- from
 * one queue; the decimation of the pictures (the Picture Analysis kernel's share of the path) runs on the same pool
 * first, one picture per item.  Used by bench.py's CPU arms; results are the same bytes as svtref_me_picture's. ---- */
typedef struct RefPool {
    RefJob            *jobs;
    int                n_jobs;
    const SvtRefPic  **build;
    int                n_build, hme_level1;
    volatile uint32_t  next_build, next_item;
    pthread_barrier_t  bar;
} RefPool;
enum { kSegW = 6, kSegH = 8 };

static void *pool_worker(void *arg) {
    RefPool *pl = (RefPool *)arg;
    for (;;) {  /* phase 1: decimation + padding, one picture per item */
        uint32_t i = __sync_fetch_and_add(&pl->next_build, 1);
        if (i >= (uint32_t)pl->n_build) break;
        svtref_build_planes(pl->build[i], pl->hme_level1);
    }
    pthread_barrier_wait(&pl->bar);
    RefThread t;
    t.ctx = (MeContext *)malloc(sizeof(MeContext));
    int cur_job = -1;
    for (;;) {  /* phase 2: (picture, segment) items in decode order */
        uint32_t it = __sync_fetch_and_add(&pl->next_item, 1);
        if (it >= (uint32_t)pl->n_jobs * kSegW * kSegH) break;
        const int ji = (int)(it / (kSegW * kSegH)), seg = (int)(it % (kSegW * kSegH));
        RefJob   *job = &pl->jobs[ji];
        if (ji != cur_job) {
            if (cur_job >= 0) ref_thread_done(&t, &pl->jobs[cur_job]);
            ref_thread_setup(&t, job);  /* the encoder re-derives the knobs per message as well (me_process.c:121) */
            cur_job = ji;
        }
        const uint32_t wb = job->pic_w_b64, hb = job->pic_h_b64, sx = seg % kSegW, sy = seg / kSegW;
        const uint32_t x0 = sx * wb / kSegW, x1 = (sx + 1) * wb / kSegW, y0 = sy * hb / kSegH, y1 = (sy + 1) * hb / kSegH;
        for (uint32_t y = y0; y < y1; y++)
            for (uint32_t x = x0; x < x1; x++) ref_thread_sb(&t, job, y * wb + x);
    }
    if (cur_job >= 0) ref_thread_done(&t, &pl->jobs[cur_job]);
    free(t.ctx);
    return NULL;
}

int svtref_me_minigop(int n_pics, const SvtMeKnobs *const *k, const SvtRefPic *const *cur, const SvtRefPic *const *refs,
                      SvtMePictureResults *const *out, int n_build, const SvtRefPic *const *build, int hme_level1,
                      int n_threads) {
    RefPool pl;
    memset(&pl, 0, sizeof(pl));
    pl.jobs = (RefJob *)calloc((size_t)n_pics, sizeof(RefJob));
    for (int i = 0; i < n_pics; i++) {
        int rc = ref_job_open(&pl.jobs[i], k[i], cur[i], refs[i], out[i]);
        if (rc) return rc;
    }
    pl.n_jobs = n_pics; pl.build = (const SvtRefPic **)build; pl.n_build = n_build; pl.hme_level1 = hme_level1;
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    pthread_barrier_init(&pl.bar, NULL, (unsigned)n_threads);
    pthread_t th[256];
    for (int t = 1; t < n_threads; t++) pthread_create(&th[t], NULL, pool_worker, &pl);
    pool_worker(&pl);
    for (int t = 1; t < n_threads; t++) pthread_join(th[t], NULL);
    pthread_barrier_destroy(&pl.bar);
    for (int i = 0; i < n_pics; i++) ref_job_close(&pl.jobs[i]);
    free(pl.jobs);
    return 0;
}

/* Decimate + pad exactly as svt_aom_downsample_filtering_input_picture (pic_analysis_process.c:1945-2002)
 * and svt_aom_pad_input_pictures (:2004-2014) do, calling the reference's own function. `full` must already
 * hold the picture at its origin; its border is (re)generated here. */
void svtref_build_planes(const SvtRefPic *pic, int hme_level1) {
    EbPictureBufferDesc     d[3];
    PictureParentControlSet pcs;
    for (int l = 0; l < 3; l++) desc_from_plane(&d[l], &pic->lvl[l]);
    svt_aom_generate_padding(d[0].buffer_y, d[0].stride_y, d[0].width, d[0].height, d[0].org_x, d[0].org_y);
    memset(&pcs, 0, sizeof(pcs));
    pcs.enable_hme_flag        = 1;
    pcs.enable_hme_level0_flag = 1;
    pcs.enable_hme_level1_flag = hme_level1 ? 1 : 0;
    svt_aom_downsample_filtering_input_picture(&pcs, &d[0], &d[1], &d[2]);
}

/* ---- dynamic-GOP detector HME level 0: the reference's own dg_detector_hme_level0 (pd_process.c:493-590), compiled from
 * pd_process.c, driven as the TASK_DG_DETECTOR_HME branch of the ME kernel thread does (me_process.c:327-329) with one
 * segment covering the picture.  The mutex / semaphore calls of that function become no-ops here (single thread). ---- */
#include "reference_object.h"
#include "object.h"
EbErrorType svt_block_on_mutex(EbHandle mutex_handle) { (void)mutex_handle; return EB_ErrorNone; }
EbErrorType svt_release_mutex(EbHandle mutex_handle) { (void)mutex_handle; return EB_ErrorNone; }
EbErrorType svt_post_semaphore(EbHandle semaphore_handle) { (void)semaphore_handle; return EB_ErrorNone; }
void        dg_detector_hme_level0(struct PictureParentControlSet *ppcs, uint32_t seg_idx);

int svtref_dg_hme_level0(const SvtRefPic *src, const SvtRefPic *ref, uint32_t input_resolution, SvtMeDgMetrics *m) {
    PictureParentControlSet *pcs[2];
    EbPaReferenceObject     *obj[2];
    EbObjectWrapper         *wr[2];
    EbPictureBufferDesc      d[2];
    SequenceControlSet      *scs = (SequenceControlSet *)calloc(1, sizeof(*scs));
    DGDetectorSeg           *dg  = (DGDetectorSeg *)calloc(1, sizeof(*dg));
    scs->b64_size = 64;
    desc_from_plane(&d[0], &src->lvl[2]);
    desc_from_plane(&d[1], &ref->lvl[2]);
    for (int i = 0; i < 2; i++) {
        pcs[i] = (PictureParentControlSet *)calloc(1, sizeof(PictureParentControlSet));
        obj[i] = (EbPaReferenceObject *)calloc(1, sizeof(EbPaReferenceObject));
        wr[i]  = (EbObjectWrapper *)calloc(1, sizeof(EbObjectWrapper));
        obj[i]->sixteenth_downsampled_picture_ptr = &d[i];
        wr[i]->object_ptr                         = obj[i];
        pcs[i]->pa_ref_pic_wrapper                = wr[i];
        pcs[i]->scs                               = scs;
    }
    pcs[0]->dg_detector              = dg;
    dg->ref_pic                      = pcs[1];
    pcs[0]->input_resolution         = (uint8_t)input_resolution;
    pcs[0]->aligned_width            = (uint16_t)((src->lvl[0].width + 7) & ~7u);
    pcs[0]->aligned_height           = (uint16_t)((src->lvl[0].height + 7) & ~7u);
    pcs[0]->me_segments_column_count = 1;
    pcs[0]->me_segments_row_count    = 1;
    dg_detector_hme_level0(pcs[0], 0);
    m->tot_dist       = dg->metrics.tot_dist;
    m->tot_cplx       = dg->metrics.tot_cplx;
    m->tot_active     = dg->metrics.tot_active;
    m->sum_in_vectors = dg->metrics.sum_in_vectors;
    const int done    = dg->metrics.seg_completed == 1;
    for (int i = 0; i < 2; i++) { free(pcs[i]); free(obj[i]); free(wr[i]); }
    free(scs); free(dg);
    return done ? 0 : -1;
}

/* ---- picture-analysis statistics: the reference's svt_aom_gathering_picture_statistics (pic_analysis_process.c:1559)
 * with calc_hist set and the variance pass off; decim_step 1 <=> scene_change_detection (:1492) ---- */
void svt_aom_gathering_picture_statistics(SequenceControlSet *scs, PictureParentControlSet *pcs,
                                          EbPictureBufferDesc *input_padded_pic,
                                          EbPictureBufferDesc *sixteenth_decimated_picture_ptr);
int svtref_picture_statistics(const SvtRefPic *pic, uint32_t rw_n, uint32_t rh_n, uint32_t step, SvtMePicStats *out) {
    if (rw_n < 1 || rw_n > 4 || rh_n < 1 || rh_n > 4 || (step != 1 && step != 4)) return -1;
    SequenceControlSet      *scs = (SequenceControlSet *)calloc(1, sizeof(*scs));
    PictureParentControlSet *pcs = (PictureParentControlSet *)calloc(1, sizeof(*pcs));
    EbPictureBufferDesc      d[2];
    uint32_t                *rows[4][4], **cols[4];
    desc_from_plane(&d[0], &pic->lvl[0]);
    desc_from_plane(&d[1], &pic->lvl[2]);
    scs->calc_hist                                     = 1;
    scs->calculate_variance                            = 0;
    scs->static_config.scene_change_detection          = step == 1;
    scs->picture_analysis_number_of_regions_per_width  = rw_n;
    scs->picture_analysis_number_of_regions_per_height = rh_n;
    for (int a = 0; a < 4; a++) {
        for (int b = 0; b < 4; b++) rows[a][b] = out->picture_histogram[a][b];
        cols[a] = rows[a];
    }
    memset(out, 0, sizeof(*out));
    pcs->picture_histogram = cols;
    svt_aom_gathering_picture_statistics(scs, pcs, &d[0], &d[1]);
    for (uint32_t a = 0; a < rw_n; a++)
        for (uint32_t b = 0; b < rh_n; b++) out->average_intensity_per_region[a][b] = pcs->average_intensity_per_region[a][b];
    out->avg_luma = pcs->avg_luma;
    free(pcs); free(scs);
    return 0;
}
