/*
 * shim_testbed.c -- TEST INFRASTRUCTURE ONLY.  Builds the reference's own structures (SequenceControlSet,
 * PictureParentControlSet, MeContext, EbPaReferenceObject behind EbObjectWrapper) from plain arrays so that tests can call
 * the encoder-side shim (integration/svtme_shim*.c) exactly as me_process.c would, without running the encoder.
 * Compiled against the reference's headers into oracle/_ref/libsvtme_shim.so (oracle/Makefile target `shim`).
 */
#include <stdlib.h>
#include <string.h>
#include "svtme_shim.h"
#include "reference_object.h"
#include "sys_resource_manager.h"

/* the 12-byte GMC chunk of the encoder dump (oracle/enc_overlay/me_dump.c) */
typedef struct ShimTbGmc {
    uint8_t  enabled, pp_enabled, gm_pp_detected, downsample_level, bypass_based_on_me, use_stationary_block, calculate_variance,
        pad;
    uint32_t pic_avg_variance, input_width, input_height;
} ShimTbGmc;

typedef struct ShimTbPic { /* one PA reference object with a padded luma plane */
    EbObjectWrapper     wrapper;
    EbPaReferenceObject obj;
    EbPictureBufferDesc desc;
    uint8_t            *mem;
} ShimTbPic;

typedef struct ShimTb {
    SequenceControlSet      scs;
    PictureParentControlSet pcs;
    MotionEstimationData    med;
    MeContext               ctx;
    EbPictureBufferDesc     input_pic;
    MeSbResults            *sbr;
    void                   *res_mem;
    ShimTbPic               cur, refs[2][4];
} ShimTb;

static void tb_area(SearchArea *d, const SvtMeArea *s) {
    d->width  = s->width;
    d->height = s->height;
}

/* the inverse of svtme_shim_knobs_from_me_context: put a knob block back where the encoder keeps it */
static void tb_spread_knobs(ShimTb *t, const SvtMeKnobs *k, uint32_t slice_type) {
    SequenceControlSet      *scs = &t->scs;
    PictureParentControlSet *pcs = &t->pcs;
    MeContext               *c   = &t->ctx;
    pcs->scs                     = scs;
    pcs->picture_number          = k->picture_number;
    pcs->hierarchical_levels     = k->hierarchical_levels;
    pcs->similar_brightness_refs = k->similar_brightness_refs;
    scs->input_resolution        = k->input_resolution;
    pcs->enable_me_8x8           = k->enable_me_8x8;
    pcs->enable_me_16x16         = k->enable_me_16x16;
    pcs->max_number_of_pus_per_sb = k->max_number_of_pus_per_sb;
    pcs->gm_ctrls.use_distance_based_active_th = k->gm_use_distance_based_active_th;
    scs->mrp_ctrls.only_l_bwd    = k->only_l_bwd;
    pcs->pa_me_data              = &t->med;
    t->med.max_cand              = k->max_cand;
    t->med.max_refs              = k->max_refs;
    t->med.max_l0                = k->max_l0;
    pcs->slice_type              = (SliceType)slice_type;
    pcs->ref_list0_count_try     = k->num_of_ref_pic_to_search[0];
    pcs->ref_list1_count_try     = k->num_of_ref_pic_to_search[1];
    pcs->temporal_layer_index    = k->temporal_layer_index;
    pcs->is_ref                  = k->is_ref;

    c->hme_search_method      = k->hme_search_method;
    c->me_search_method       = k->me_search_method;
    c->enable_hme_flag        = k->enable_hme_flag;
    c->enable_hme_level0_flag = k->enable_hme_level0_flag;
    c->enable_hme_level1_flag = k->enable_hme_level1_flag;
    c->enable_hme_level2_flag = k->enable_hme_level2_flag;
    c->num_hme_sa_w           = k->num_hme_sa_w;
    c->num_hme_sa_h           = k->num_hme_sa_h;
    tb_area(&c->hme_l0_sa.sa_min, &k->hme_l0_sa.sa_min);
    tb_area(&c->hme_l0_sa.sa_max, &k->hme_l0_sa.sa_max);
    tb_area(&c->hme_l1_sa, &k->hme_l1_sa);
    tb_area(&c->hme_l2_sa, &k->hme_l2_sa);
    tb_area(&c->me_sa.sa_min, &k->me_sa.sa_min);
    tb_area(&c->me_sa.sa_max, &k->me_sa.sa_max);
    MeHmeRefPruneCtrls *pr = &c->me_hme_prune_ctrls;
    pr->enable_me_hme_ref_pruning               = k->enable_me_hme_ref_pruning;
    pr->prune_ref_if_hme_sad_dev_bigger_than_th = k->prune_ref_if_hme_sad_dev_bigger_than_th;
    pr->prune_ref_if_me_sad_dev_bigger_than_th  = k->prune_ref_if_me_sad_dev_bigger_than_th;
    pr->zz_sad_th                               = k->zz_sad_th;
    pr->zz_sad_pct                              = k->zz_sad_pct;
    pr->phme_sad_th                             = k->phme_sad_th;
    pr->phme_sad_pct                            = k->phme_sad_pct;
    MeSrCtrls *sr = &c->me_sr_adjustment_ctrls;
    sr->enable_me_sr_adjustment              = k->enable_me_sr_adjustment;
    sr->reduce_me_sr_based_on_mv_length_th   = k->reduce_me_sr_based_on_mv_length_th;
    sr->stationary_hme_sad_abs_th            = k->stationary_hme_sad_abs_th;
    sr->stationary_me_sr_divisor             = k->stationary_me_sr_divisor;
    sr->reduce_me_sr_based_on_hme_sad_abs_th = k->reduce_me_sr_based_on_hme_sad_abs_th;
    sr->me_sr_divisor_for_low_hme_sad        = k->me_sr_divisor_for_low_hme_sad;
    sr->distance_based_hme_resizing          = k->distance_based_hme_resizing;
    c->me_8x8_var_ctrls.enabled        = k->me_8x8_var_enabled;
    c->me_8x8_var_ctrls.me_sr_div4_th  = k->me_sr_div4_th;
    c->me_8x8_var_ctrls.me_sr_div2_th  = k->me_sr_div2_th;
    c->me_8x8_var_ctrls.me_sr_mult2_th = k->me_sr_mult2_th;
    c->mv_based_sa_adj.enabled          = k->mv_sa_adj_enabled;
    c->mv_based_sa_adj.nearest_ref_only = k->mv_sa_adj_nearest_ref_only;
    c->mv_based_sa_adj.mv_size_th       = k->mv_sa_adj_mv_size_th;
    c->mv_based_sa_adj.sa_multiplier    = k->mv_sa_adj_sa_multiplier;
    c->prehme_ctrl.enable               = k->prehme_enable;
    for (int i = 0; i < 2; i++) {
        tb_area(&c->prehme_ctrl.prehme_sa_cfg[i].sa_min, &k->prehme_sa_cfg[i].sa_min);
        tb_area(&c->prehme_ctrl.prehme_sa_cfg[i].sa_max, &k->prehme_sa_cfg[i].sa_max);
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
    c->me_type                      = ME_OPEN_LOOP;
}

static void tb_gmc(ShimTb *t, const ShimTbGmc *g) {
    t->pcs.gm_ctrls.enabled              = g->enabled;
    t->pcs.gm_ctrls.pp_enabled           = g->pp_enabled;
    t->pcs.gm_pp_detected                = g->gm_pp_detected;
    t->pcs.gm_ctrls.downsample_level     = g->downsample_level;
    t->pcs.gm_ctrls.bypass_based_on_me   = g->bypass_based_on_me;
    t->pcs.gm_ctrls.use_stationary_block = g->use_stationary_block;
    t->scs.calculate_variance            = g->calculate_variance;
    t->pcs.pic_avg_variance              = (uint16_t)g->pic_avg_variance;
    t->input_pic.width                   = (uint16_t)g->input_width;
    t->input_pic.height                  = (uint16_t)g->input_height;
}

/* result storage of the pcs as pcs.c:98-119, 1302-1310 allocate it */
static int tb_results(ShimTb *t, uint32_t W, uint32_t H, uint32_t pus) {
    PictureParentControlSet *pcs = &t->pcs;
    const uint32_t           n   = ((W + 63) / 64) * ((H + 63) / 64);
    pcs->aligned_width           = (uint16_t)W;
    pcs->aligned_height          = (uint16_t)H;
    pcs->b64_total_count         = (uint16_t)n;
    t->med.b64_total_count       = (uint16_t)n;
    SvtMePictureResults r;
    const size_t        bytes = svtme_shim_results_layout(&r, n, pus, t->med.max_refs, t->med.max_cand, NULL);
    t->res_mem                = calloc(1, bytes);
    t->sbr                    = (MeSbResults *)calloc(n, sizeof(MeSbResults));
    t->med.me_results         = (MeSbResults **)calloc(n, sizeof(MeSbResults *));
    if (!t->res_mem || !t->sbr || !t->med.me_results) return -1;
    svtme_shim_results_layout(&r, n, pus, t->med.max_refs, t->med.max_cand, t->res_mem);
    for (uint32_t b = 0; b < n; b++) {
        t->med.me_results[b]               = &t->sbr[b];
        t->sbr[b].me_mv_array              = (MvCandidate *)(r.me_mv + (size_t)b * pus * t->med.max_refs);
        t->sbr[b].me_candidate_array       = (MeCandidate *)(r.me_cand + (size_t)b * pus * t->med.max_cand);
        t->sbr[b].total_me_candidate_index = r.total_me_candidate_index + (size_t)b * pus;
    }
    pcs->me_64x64_distortion         = r.me_64x64_distortion;
    pcs->me_32x32_distortion         = r.me_32x32_distortion;
    pcs->me_16x16_distortion         = r.me_16x16_distortion;
    pcs->me_8x8_distortion           = r.me_8x8_distortion;
    pcs->me_8x8_cost_variance        = r.me_8x8_cost_variance;
    pcs->rc_me_distortion            = r.rc_me_distortion;
    pcs->stationary_block_present_sb = r.stationary_block_present_sb;
    pcs->rc_me_allow_gm              = r.rc_me_allow_gm;
    return 0;
}

static int tb_picture(ShimTbPic *p, const uint8_t *luma, uint32_t W, uint32_t H, uint64_t picture_number) {
    const uint32_t pad = 72; /* enc_handle.c:4057-4070 */
    const uint32_t st  = W + 2 * pad;
    p->mem             = (uint8_t *)calloc((size_t)st, H + 2 * pad);
    if (!p->mem) return -1;
    for (uint32_t y = 0; y < H; y++) memcpy(p->mem + (size_t)(y + pad) * st + pad, luma + (size_t)y * W, W);
    p->desc.buffer_y        = p->mem;
    p->desc.stride_y        = (uint16_t)st;
    p->desc.org_x           = (uint16_t)pad;
    p->desc.org_y           = (uint16_t)pad;
    p->desc.width           = (uint16_t)W;
    p->desc.height          = (uint16_t)H;
    p->obj.input_padded_pic = &p->desc;
    p->obj.picture_number   = picture_number;
    p->wrapper.object_ptr   = &p->obj;
    return 0;
}

static void tb_free(ShimTb *t) {
    if (!t) return;
    free(t->cur.mem);
    for (int i = 0; i < 8; i++) free(t->refs[i / 4][i % 4].mem);
    free(t->res_mem);
    free(t->sbr);
    free(t->med.me_results);
    free(t);
}

static ShimTb *tb_new(const SvtMeKnobs *k, const ShimTbGmc *gmc, uint32_t W, uint32_t H, uint32_t slice_type) {
    ShimTb *t = (ShimTb *)calloc(1, sizeof(*t));
    if (!t) return NULL;
    tb_spread_knobs(t, k, slice_type);
    tb_gmc(t, gmc);
    t->scs.max_input_luma_width  = (uint16_t)W;
    t->scs.max_input_luma_height = (uint16_t)H;
    if (tb_results(t, W, H, svtme_shim_pu_count(&t->pcs)) != 0) {
        tb_free(t);
        return NULL;
    }
    return t;
}

/*
 * CPU-only: put `k` into scs/pcs/MeContext, read it back through svtme_shim_knobs_from_me_context (-> k_back), scatter the flat
 * results `res_in` into the pcs, gather them again (-> res_back, same layout), evaluate the GM gate from the gathered sums.
 * gate_out: run_gm, level, downsample_level.
 */
int shimtb_roundtrip(const SvtMeKnobs *k, const void *gmc, uint32_t W, uint32_t H, uint32_t slice_type, const void *res_in,
                     SvtMeKnobs *k_back, void *res_back, uint8_t gate_out[3]) {
    ShimTb *t = tb_new(k, (const ShimTbGmc *)gmc, W, H, slice_type);
    if (!t) return -1;
    svtme_shim_knobs_from_me_context(&t->scs, &t->pcs, &t->ctx, k_back);
    const uint32_t      n = t->pcs.b64_total_count, pus = svtme_shim_pu_count(&t->pcs);
    SvtMePictureResults in, back;
    const size_t        bytes = svtme_shim_results_layout(&in, n, pus, k->max_refs, k->max_cand, NULL);
    void               *copy  = malloc(bytes);
    memcpy(copy, res_in, bytes);
    svtme_shim_results_layout(&in, n, pus, k->max_refs, k->max_cand, copy);
    svtme_shim_scatter_results(&t->pcs, &in);
    memset(res_back, 0, bytes);
    svtme_shim_results_layout(&back, n, pus, k->max_refs, k->max_cand, res_back);
    svtme_shim_gather_results(&t->pcs, &back);
    SvtMeShimGmGate g;
    svtme_shim_gm_gate(&t->pcs, &t->input_pic, back.picture_sums, &g);
    gate_out[0] = g.run_gm;
    gate_out[1] = g.level;
    gate_out[2] = g.downsample_level;
    free(copy);
    tb_free(t);
    return 0;
}

/*
 * GPU: one picture through svtme_shim_me_picture, exactly the call me_process.c would make.  `lumas[li*4+ri]` are tight W x H
 * planes (NULL where unused).  res_out receives the pcs contents gathered in the flat layout; gate_out as above.
 */
int shimtb_me_picture(SvtMeShim *s, const SvtMeKnobs *k, const void *gmc, uint32_t W, uint32_t H, uint32_t slice_type,
                      const uint8_t *cur_luma, const uint8_t *const *lumas, const uint64_t *ref_numbers, void *res_out,
                      uint8_t gate_out[3]) {
    ShimTb *t = tb_new(k, (const ShimTbGmc *)gmc, W, H, slice_type);
    if (!t) return -1;
    int rc = tb_picture(&t->cur, cur_luma, W, H, k->picture_number);
    t->pcs.pa_ref_pic_wrapper = &t->cur.wrapper;
    for (int li = 0; li < 2 && !rc; li++)
        for (int ri = 0; ri < 4 && !rc; ri++)
            if (lumas[li * 4 + ri]) {
                rc = tb_picture(&t->refs[li][ri], lumas[li * 4 + ri], W, H, ref_numbers[li * 4 + ri]);
                t->pcs.ref_pa_pic_ptr_array[li][ri] = &t->refs[li][ri].wrapper;
            }
    SvtMeShimGmGate g;
    memset(&g, 0, sizeof(g));
    if (!rc) rc = svtme_shim_me_picture(s, &t->scs, &t->pcs, &t->ctx, &t->input_pic, &g);
    if (!rc) {
        const uint32_t      n = t->pcs.b64_total_count, pus = svtme_shim_pu_count(&t->pcs);
        SvtMePictureResults back;
        const size_t        bytes = svtme_shim_results_layout(&back, n, pus, k->max_refs, k->max_cand, NULL);
        memset(res_out, 0, bytes);
        svtme_shim_results_layout(&back, n, pus, k->max_refs, k->max_cand, res_out);
        svtme_shim_gather_results(&t->pcs, &back);
        gate_out[0] = g.run_gm;
        gate_out[1] = g.level;
        gate_out[2] = g.downsample_level;
    }
    tb_free(t);
    return rc;
}

/* open/close pass-throughs that build the scs the driver reads */
int shimtb_open(uint32_t W, uint32_t H, int device, uint32_t pictures, int hme_level1_planes, SvtMeShim **out) {
    SequenceControlSet *scs = (SequenceControlSet *)calloc(1, sizeof(*scs));
    if (!scs) return -1;
    scs->max_input_luma_width  = (uint16_t)W;
    scs->max_input_luma_height = (uint16_t)H;
    int rc                     = svtme_shim_open(scs, device, pictures, hme_level1_planes, out);
    free(scs);
    return rc;
}
size_t shimtb_results_bytes(uint32_t n_b64, uint32_t pus, uint32_t max_refs, uint32_t max_cand) {
    SvtMePictureResults r;
    return svtme_shim_results_layout(&r, n_b64, pus, max_refs, max_cand, NULL);
}

/*
 * CPU-only, temporal filter: put the TF knob block `k` where the filter's MeContext / centre pcs keep it (pa_me_data is not
 * assigned yet when picture decision runs the filter: NULL here), read it back through svtme_shim_tf_knobs_from_me_context
 * (-> k_back), then hand one block's results to svtme_shim_tf_fill_b64 and report what the filter would read:
 * out[0..84] SADs through p_best_sad_64x64 / 32x32 / 16x16 / 8x8, out[85..169] MVs through p_best_mv*, out[170..171] the HME
 * centre, out[172] tf_use_pred_64x64_only_th, out[173..174] tf_tot_horz_blks / tf_tot_vert_blks.
 */
int shimtb_tf_roundtrip(const SvtMeKnobs *k, uint32_t W, uint32_t H, const uint32_t *best_sad, const uint32_t *best_mv, int sc_x,
                        int sc_y, int exited, uint8_t th_before, SvtMeKnobs *k_back, int64_t out[175]) {
    static const ShimTbGmc no_gm;
    ShimTb *t = tb_new(k, &no_gm, W, H, 1);
    if (!t) return -1;
    t->pcs.pa_me_data           = NULL;
    t->ctx.me_type              = ME_MCTF;
    t->ctx.temporal_layer_index = k->temporal_layer_index;
    t->ctx.is_ref               = k->is_ref;
    t->ctx.tf_me_exit_th        = (uint16_t)k->tf_me_exit_th;
    svtme_shim_tf_knobs_from_me_context(&t->scs, &t->pcs, &t->ctx, k_back);
    MeContext *c                 = &t->ctx;
    c->tf_use_pred_64x64_only_th = th_before;
    c->tf_tot_horz_blks          = 100;
    c->tf_tot_vert_blks          = 200;
    svtme_shim_tf_fill_b64(c, best_sad, best_mv, (int16_t)sc_x, (int16_t)sc_y, exited);
    out[0] = c->p_best_sad_64x64[0];
    for (int i = 0; i < 4; i++) out[1 + i] = c->p_best_sad_32x32[i];
    for (int i = 0; i < 16; i++) out[5 + i] = c->p_best_sad_16x16[i];
    for (int i = 0; i < 64; i++) out[21 + i] = c->p_best_sad_8x8[i];
    out[85] = c->p_best_mv64x64[0];
    for (int i = 0; i < 4; i++) out[86 + i] = c->p_best_mv32x32[i];
    for (int i = 0; i < 16; i++) out[90 + i] = c->p_best_mv16x16[i];
    for (int i = 0; i < 64; i++) out[106 + i] = c->p_best_mv8x8[i];
    out[170] = c->search_results[0][0].hme_sc_x;
    out[171] = c->search_results[0][0].hme_sc_y;
    out[172] = c->tf_use_pred_64x64_only_th;
    out[173] = c->tf_tot_horz_blks;
    out[174] = c->tf_tot_vert_blks;
    tb_free(t);
    return 0;
}
