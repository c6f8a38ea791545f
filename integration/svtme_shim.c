/*
 * svtme_shim.c -- pure host part of the encoder-side shim (see svtme_shim.h).  C99, compiled against the reference's headers;
 * calls nothing in libsvtme_cuda.so.
 */
#include <string.h>
#include "svtme_shim.h"

static void area_of(SvtMeArea *d, const SearchArea *s) {
    d->width  = s->width;
    d->height = s->height;
}
static void area_mm_of(SvtMeAreaMinMax *d, const SearchAreaMinMax *s) {
    area_of(&d->sa_min, &s->sa_min);
    area_of(&d->sa_max, &s->sa_max);
}

/* the block svt_aom_sig_deriv_me (open loop, enc_mode_config.c:671-808) or svt_aom_sig_deriv_me_tf (temporal filter, :814-854)
 * wrote into the MeContext */
static void knobs_from_ctx(const MeContext *c, SvtMeKnobs *k) {
    k->hme_search_method      = c->hme_search_method;
    k->me_search_method       = c->me_search_method;
    k->enable_hme_flag        = c->enable_hme_flag;
    k->enable_hme_level0_flag = c->enable_hme_level0_flag;
    k->enable_hme_level1_flag = c->enable_hme_level1_flag;
    k->enable_hme_level2_flag = c->enable_hme_level2_flag;
    k->num_hme_sa_w           = c->num_hme_sa_w;
    k->num_hme_sa_h           = c->num_hme_sa_h;
    area_mm_of(&k->hme_l0_sa, &c->hme_l0_sa);
    area_of(&k->hme_l1_sa, &c->hme_l1_sa);
    area_of(&k->hme_l2_sa, &c->hme_l2_sa);
    area_mm_of(&k->me_sa, &c->me_sa);
    k->enable_me_hme_ref_pruning               = c->me_hme_prune_ctrls.enable_me_hme_ref_pruning;
    k->prune_ref_if_hme_sad_dev_bigger_than_th = c->me_hme_prune_ctrls.prune_ref_if_hme_sad_dev_bigger_than_th;
    k->prune_ref_if_me_sad_dev_bigger_than_th  = c->me_hme_prune_ctrls.prune_ref_if_me_sad_dev_bigger_than_th;
    k->zz_sad_th                               = c->me_hme_prune_ctrls.zz_sad_th;
    k->zz_sad_pct                              = c->me_hme_prune_ctrls.zz_sad_pct;
    k->phme_sad_th                             = c->me_hme_prune_ctrls.phme_sad_th;
    k->phme_sad_pct                            = c->me_hme_prune_ctrls.phme_sad_pct;
    k->enable_me_sr_adjustment                 = c->me_sr_adjustment_ctrls.enable_me_sr_adjustment;
    k->reduce_me_sr_based_on_mv_length_th      = c->me_sr_adjustment_ctrls.reduce_me_sr_based_on_mv_length_th;
    k->stationary_hme_sad_abs_th               = c->me_sr_adjustment_ctrls.stationary_hme_sad_abs_th;
    k->stationary_me_sr_divisor                = c->me_sr_adjustment_ctrls.stationary_me_sr_divisor;
    k->reduce_me_sr_based_on_hme_sad_abs_th    = c->me_sr_adjustment_ctrls.reduce_me_sr_based_on_hme_sad_abs_th;
    k->me_sr_divisor_for_low_hme_sad           = c->me_sr_adjustment_ctrls.me_sr_divisor_for_low_hme_sad;
    k->distance_based_hme_resizing             = c->me_sr_adjustment_ctrls.distance_based_hme_resizing;
    k->me_8x8_var_enabled                      = c->me_8x8_var_ctrls.enabled;
    k->me_sr_div4_th                           = c->me_8x8_var_ctrls.me_sr_div4_th;
    k->me_sr_div2_th                           = c->me_8x8_var_ctrls.me_sr_div2_th;
    k->me_sr_mult2_th                          = c->me_8x8_var_ctrls.me_sr_mult2_th;
    k->mv_sa_adj_enabled                       = c->mv_based_sa_adj.enabled;
    k->mv_sa_adj_nearest_ref_only              = c->mv_based_sa_adj.nearest_ref_only;
    k->mv_sa_adj_mv_size_th                    = c->mv_based_sa_adj.mv_size_th;
    k->mv_sa_adj_sa_multiplier                 = c->mv_based_sa_adj.sa_multiplier;
    k->prehme_enable                           = c->prehme_ctrl.enable;
    for (int i = 0; i < 2; i++) area_mm_of(&k->prehme_sa_cfg[i], &c->prehme_ctrl.prehme_sa_cfg[i]);
    k->prehme_skip_search_line     = c->prehme_ctrl.skip_search_line;
    k->prehme_l1_early_exit        = c->prehme_ctrl.l1_early_exit;
    k->prune_me_candidates_th      = c->prune_me_candidates_th;
    k->use_best_unipred_cand_only  = c->use_best_unipred_cand_only;
    k->reduce_hme_l0_sr_th_min     = c->reduce_hme_l0_sr_th_min;
    k->reduce_hme_l0_sr_th_max     = c->reduce_hme_l0_sr_th_max;
    k->me_early_exit_th            = c->me_early_exit_th;
    k->me_safe_limit_zz_th         = c->me_safe_limit_zz_th;
    k->prev_me_stage_based_exit_th = c->prev_me_stage_based_exit_th;
    k->tf_me_exit_th               = 0;
}

void svtme_shim_knobs_from_me_context(const SequenceControlSet *scs, const PictureParentControlSet *pcs, const MeContext *c,
                                      SvtMeKnobs *k) {
    memset(k, 0, sizeof(*k));
    /* (c) what motion_estimation.c reads from pcs / scs directly */
    k->picture_number                  = pcs->picture_number;
    k->hierarchical_levels             = pcs->hierarchical_levels;
    k->similar_brightness_refs         = pcs->similar_brightness_refs;
    k->input_resolution                = (uint8_t)scs->input_resolution;
    k->enable_me_8x8                   = pcs->enable_me_8x8;
    k->enable_me_16x16                 = pcs->enable_me_16x16;
    k->max_number_of_pus_per_sb        = pcs->max_number_of_pus_per_sb;
    k->gm_enabled                      = pcs->gm_ctrls.enabled;
    k->gm_use_distance_based_active_th = pcs->gm_ctrls.use_distance_based_active_th;
    k->only_l_bwd                      = scs->mrp_ctrls.only_l_bwd;
    k->max_cand                        = pcs->pa_me_data->max_cand;
    k->max_refs                        = pcs->pa_me_data->max_refs;
    k->max_l0                          = pcs->pa_me_data->max_l0;
    /* (b) what the SB loop copies into MeContext per message, me_process.c:216-227 */
    k->num_of_list_to_search       = pcs->slice_type == P_SLICE ? 1 : 2;
    k->num_of_ref_pic_to_search[0] = pcs->ref_list0_count_try;
    k->num_of_ref_pic_to_search[1] = pcs->slice_type == B_SLICE ? pcs->ref_list1_count_try : 0;
    k->temporal_layer_index        = pcs->temporal_layer_index;
    k->is_ref                      = pcs->is_ref;
    k->me_type                     = SVTME_ME_OPEN_LOOP;
    knobs_from_ctx(c, k);
}

/* Temporal-filter searches (produce_temporally_filtered_pic, temporal_filtering.c:3129-3174): `c` is the MeContext of the TF
 * segment thread after create_me_context_and_picture_control (:296, 316), the per-pair lines :3140-3164 and
 * set_hme_search_params_mctf (:2759-2767), i.e. as svt_aom_motion_estimation_b64 would see it. */
void svtme_shim_tf_knobs_from_me_context(const SequenceControlSet *scs, const PictureParentControlSet *centre_pcs, const MeContext *c,
                                         SvtMeKnobs *k) {
    memset(k, 0, sizeof(*k));
    k->picture_number          = centre_pcs->picture_number;
    k->hierarchical_levels     = centre_pcs->hierarchical_levels;
    k->similar_brightness_refs = centre_pcs->similar_brightness_refs;
    k->input_resolution        = (uint8_t)scs->input_resolution;
    /* read by the candidate / statistics stage only, which ME_MCTF skips (motion_estimation.c:3126); pa_me_data is not
     * assigned yet when picture decision runs the filter */
    k->enable_me_8x8 = k->enable_me_16x16 = 1;
    k->max_number_of_pus_per_sb = SQUARE_PU_COUNT;
    k->max_cand = k->max_refs = k->max_l0 = 1;
    knobs_from_ctx(c, k);
    k->me_type                     = SVTME_ME_MCTF;
    k->num_of_list_to_search       = 1;
    k->num_of_ref_pic_to_search[0] = 1;
    k->num_of_ref_pic_to_search[1] = 0;
    k->temporal_layer_index        = c->temporal_layer_index;
    k->is_ref                      = c->is_ref;
    k->tf_me_exit_th               = c->tf_me_exit_th;
}

/* What svt_aom_motion_estimation_b64 leaves in the MeContext for the temporal filter (ME_MCTF), from one block's row of a
 * whole-picture search: p_sb_best_sad / p_sb_best_mv [0][0] with the p_best_* pointers on them (motion_estimation.c:1368-1384),
 * the HME centre (:2465), the horizontal / vertical block count (:2469-2474) and the early-exit mark (:3109-3113). */
void svtme_shim_tf_fill_b64(MeContext *ctx, const uint32_t *best_sad, const uint32_t *best_mv, int16_t hme_sc_x, int16_t hme_sc_y,
                            int exited) {
    memcpy(ctx->p_sb_best_sad[0][0], best_sad, SQUARE_PU_COUNT * sizeof(uint32_t));
    memcpy(ctx->p_sb_best_mv[0][0], best_mv, SQUARE_PU_COUNT * sizeof(uint32_t));
    ctx->p_best_sad_64x64 = &ctx->p_sb_best_sad[0][0][ME_TIER_ZERO_PU_64x64];
    ctx->p_best_sad_32x32 = &ctx->p_sb_best_sad[0][0][ME_TIER_ZERO_PU_32x32_0];
    ctx->p_best_sad_16x16 = &ctx->p_sb_best_sad[0][0][ME_TIER_ZERO_PU_16x16_0];
    ctx->p_best_sad_8x8   = &ctx->p_sb_best_sad[0][0][ME_TIER_ZERO_PU_8x8_0];
    ctx->p_best_mv64x64   = &ctx->p_sb_best_mv[0][0][ME_TIER_ZERO_PU_64x64];
    ctx->p_best_mv32x32   = &ctx->p_sb_best_mv[0][0][ME_TIER_ZERO_PU_32x32_0];
    ctx->p_best_mv16x16   = &ctx->p_sb_best_mv[0][0][ME_TIER_ZERO_PU_16x16_0];
    ctx->p_best_mv8x8     = &ctx->p_sb_best_mv[0][0][ME_TIER_ZERO_PU_8x8_0];
    ctx->search_results[0][0].hme_sc_x = hme_sc_x;
    ctx->search_results[0][0].hme_sc_y = hme_sc_y;
    const int ax = hme_sc_x < 0 ? -hme_sc_x : hme_sc_x, ay = hme_sc_y < 0 ? -hme_sc_y : hme_sc_y;
    if (ax > ay) ctx->tf_tot_horz_blks++;
    else ctx->tf_tot_vert_blks++;
    if (exited) ctx->tf_use_pred_64x64_only_th = (uint8_t)~0;
}

uint32_t svtme_shim_pu_count(const PictureParentControlSet *pcs) {
    return pcs->enable_me_16x16 ? (pcs->enable_me_8x8 ? SQUARE_PU_COUNT : MAX_SB64_PU_COUNT_NO_8X8) : MAX_SB64_PU_COUNT_WO_16X16;
}

size_t svtme_shim_results_layout(SvtMePictureResults *res, uint32_t n_b64, uint32_t pus, uint32_t max_refs, uint32_t max_cand,
                                 void *mem) {
    size_t   off  = 0;
    uint8_t *base = (uint8_t *)mem;
#define SHIM_FIELD(field, type, count)                        \
    do {                                                      \
        off = (off + 15) & ~(size_t)15;                       \
        if (base) res->field = (type *)(base + off);          \
        off += sizeof(type) * (size_t)(count);                \
    } while (0)
    if (base) {
        memset(res, 0, sizeof(*res));
        res->b64_total_count = n_b64;
        res->pu_count        = pus;
    }
    SHIM_FIELD(me_mv, uint32_t, (size_t)n_b64 * pus * max_refs);
    SHIM_FIELD(me_cand, uint8_t, (size_t)n_b64 * pus * max_cand);
    SHIM_FIELD(total_me_candidate_index, uint8_t, (size_t)n_b64 * pus);
    SHIM_FIELD(me_64x64_distortion, uint32_t, n_b64);
    SHIM_FIELD(me_32x32_distortion, uint32_t, n_b64);
    SHIM_FIELD(me_16x16_distortion, uint32_t, n_b64);
    SHIM_FIELD(me_8x8_distortion, uint32_t, n_b64);
    SHIM_FIELD(me_8x8_cost_variance, uint32_t, n_b64);
    SHIM_FIELD(rc_me_distortion, uint32_t, n_b64);
    SHIM_FIELD(stationary_block_present_sb, uint8_t, n_b64);
    SHIM_FIELD(rc_me_allow_gm, uint8_t, n_b64);
    SHIM_FIELD(picture_sums, uint64_t, 4);
#undef SHIM_FIELD
    return off;
}

/* MeCandidate is a one-byte bit-field struct; the flat buffer carries it as that byte (svtme_cuda.h, me_cand). */
typedef char shim_me_candidate_is_one_byte[sizeof(MeCandidate) == 1 ? 1 : -1];
typedef char shim_mv_candidate_is_four_bytes[sizeof(MvCandidate) == 4 ? 1 : -1];

void svtme_shim_scatter_results(PictureParentControlSet *pcs, const SvtMePictureResults *res) {
    const MotionEstimationData *med = pcs->pa_me_data;
    const uint32_t              pus = res->pu_count, mr = med->max_refs, mc = med->max_cand;
    const uint32_t              n   = pcs->b64_total_count;
    for (uint32_t b = 0; b < n; b++) {
        MeSbResults *r = med->me_results[b];
        memcpy(r->me_mv_array, res->me_mv + (size_t)b * pus * mr, sizeof(MvCandidate) * pus * mr);
        memcpy(r->me_candidate_array, res->me_cand + (size_t)b * pus * mc, sizeof(MeCandidate) * pus * mc);
        memcpy(r->total_me_candidate_index, res->total_me_candidate_index + (size_t)b * pus, pus);
    }
    memcpy(pcs->me_64x64_distortion, res->me_64x64_distortion, sizeof(uint32_t) * n);
    memcpy(pcs->me_32x32_distortion, res->me_32x32_distortion, sizeof(uint32_t) * n);
    memcpy(pcs->me_16x16_distortion, res->me_16x16_distortion, sizeof(uint32_t) * n);
    memcpy(pcs->me_8x8_distortion, res->me_8x8_distortion, sizeof(uint32_t) * n);
    memcpy(pcs->me_8x8_cost_variance, res->me_8x8_cost_variance, sizeof(uint32_t) * n);
    memcpy(pcs->rc_me_distortion, res->rc_me_distortion, sizeof(uint32_t) * n);
    memcpy(pcs->stationary_block_present_sb, res->stationary_block_present_sb, n);
    memcpy(pcs->rc_me_allow_gm, res->rc_me_allow_gm, n);
}

void svtme_shim_gather_results(const PictureParentControlSet *pcs, SvtMePictureResults *res) {
    const MotionEstimationData *med = pcs->pa_me_data;
    const uint32_t              pus = res->pu_count, mr = med->max_refs, mc = med->max_cand;
    const uint32_t              n   = pcs->b64_total_count;
    uint64_t                    s[4] = {0, 0, 0, 0};
    for (uint32_t b = 0; b < n; b++) {
        const MeSbResults *r = med->me_results[b];
        memcpy(res->me_mv + (size_t)b * pus * mr, r->me_mv_array, sizeof(MvCandidate) * pus * mr);
        memcpy(res->me_cand + (size_t)b * pus * mc, r->me_candidate_array, sizeof(MeCandidate) * pus * mc);
        memcpy(res->total_me_candidate_index + (size_t)b * pus, r->total_me_candidate_index, pus);
        s[0] += pcs->me_8x8_distortion[b];
        s[1] += pcs->rc_me_distortion[b];
        s[2] += pcs->stationary_block_present_sb[b];
        s[3] += pcs->rc_me_allow_gm[b];
    }
    memcpy(res->me_64x64_distortion, pcs->me_64x64_distortion, sizeof(uint32_t) * n);
    memcpy(res->me_32x32_distortion, pcs->me_32x32_distortion, sizeof(uint32_t) * n);
    memcpy(res->me_16x16_distortion, pcs->me_16x16_distortion, sizeof(uint32_t) * n);
    memcpy(res->me_8x8_distortion, pcs->me_8x8_distortion, sizeof(uint32_t) * n);
    memcpy(res->me_8x8_cost_variance, pcs->me_8x8_cost_variance, sizeof(uint32_t) * n);
    memcpy(res->rc_me_distortion, pcs->rc_me_distortion, sizeof(uint32_t) * n);
    memcpy(res->stationary_block_present_sb, pcs->stationary_block_present_sb, n);
    memcpy(res->rc_me_allow_gm, pcs->rc_me_allow_gm, n);
    if (res->picture_sums) memcpy(res->picture_sums, s, sizeof(s));
}

/* thresholds of global_me.c:25-28 (file-local #defines there) */
#define SHIM_GMV_ME_SAD_TH_0 1
#define SHIM_GMV_ME_SAD_TH_1 5
#define SHIM_GMV_ME_SAD_TH_2 10
#define SHIM_GMV_PIC_VAR_TH 750

void svtme_shim_gm_gate(const PictureParentControlSet *pcs, const EbPictureBufferDesc *input_pic, const uint64_t sums[4],
                        SvtMeShimGmGate *g) {
    memset(g, 0, sizeof(*g));
    g->run_gm = pcs->gm_ctrls.enabled && (!pcs->gm_ctrls.pp_enabled || pcs->gm_pp_detected); /* me_process.c:276 */
    /* global_me.c:162-169: the reference accumulates in uint32_t; the device sums are 64-bit, so wrap the same way */
    const uint32_t total_me_sad        = (uint32_t)sums[1];
    const uint32_t total_stationary_sb = (uint32_t)sums[2];
    const uint32_t total_gm_sbs        = (uint32_t)sums[3];
    const uint32_t average_me_sad      = total_me_sad / ((uint32_t)input_pic->width * input_pic->height);
    g->average_me_sad                  = average_me_sad;
    uint8_t level = average_me_sad < SHIM_GMV_ME_SAD_TH_0 ? 0
        : average_me_sad < SHIM_GMV_ME_SAD_TH_1           ? 1
        : average_me_sad < SHIM_GMV_ME_SAD_TH_2           ? 2
                                                          : 3; /* :178-185 */
    if (pcs->gm_ctrls.downsample_level == GM_ADAPT_0)
        g->downsample_level = average_me_sad < SHIM_GMV_ME_SAD_TH_1 ? GM_DOWN : GM_FULL;
    else if (pcs->gm_ctrls.downsample_level == GM_ADAPT_1)
        g->downsample_level = (average_me_sad < SHIM_GMV_ME_SAD_TH_2 &&
                               (!pcs->scs->calculate_variance || pcs->pic_avg_variance < SHIM_GMV_PIC_VAR_TH))
            ? GM_DOWN16
            : GM_DOWN;
    else
        g->downsample_level = pcs->gm_ctrls.downsample_level; /* :186-198 */
    if (pcs->gm_ctrls.bypass_based_on_me) { /* :199-205 */
        if (total_gm_sbs < (uint32_t)(pcs->b64_total_count >> 1) ||
            (pcs->gm_ctrls.use_stationary_block && total_stationary_sb > (uint32_t)((pcs->b64_total_count * 5) / 100)))
            level = 0;
    }
    g->level = level;
}
