/*
 * me_dump.c -- TEST INFRASTRUCTURE ONLY.  Linked into the C-only reference encoder that oracle/Makefile target `enc`
 * builds into oracle/_ref/SvtAv1EncApp_dump, and called from the three one-line hooks make_overlay.py inserts into a
 * build-time copy of Source/Lib/Codec/me_process.c and global_me.c.  With SVTME_DUMP=<file> set, every open-loop ME
 * picture of the encode is serialised there: the planes ME read, the knob block extracted from the encoder's own
 * MeContext/pcs by the shim (integration/svtme_shim.c), the reference lists, and what the reference's ME wrote.
 *
 * Record = a sequence of chunks { char tag[4]; uint32 bytes; payload }, closed by an "END " chunk:
 *   HEAD  uint64 picture_number, uint32 width, height, b64_total_count, pu_count, slice_type, hme_level1_planes,
 *         temporal_layer_index, sizeof(SvtMeKnobs)
 *   KNOB  SvtMeKnobs (include/svtme_cuda.h)
 *   REFS  uint64 picture_number[2][4] (~0 where unused)
 *   CUR0/CUR1/CUR2  full / quarter / sixteenth luma of the source picture, tight rows (no padding)
 *   R<l><i>0        full luma of reference (list l, index i)
 *   RES   the flat result buffers in svtme_shim_results_layout order, tight
 *   GMC   the pcs/scs fields the last-SB gate reads (gm_ctrls, gm_pp_detected, pic_avg_variance, input size)
 *   GATE  uint8 ran (svt_aom_global_motion_estimation was called), level, downsample_level (0xff when it did not run)
 *         -- `level`/`downsample_level` come from the hook inside global_me.c, i.e. the reference's own locals
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "svtme_shim.h"
#include "reference_object.h"

static FILE *g_f;
static int   g_tried;
static int   g_gm_level = -1, g_gm_ds = -1;

static FILE *dump_file(void) {
    if (!g_tried) {
        g_tried        = 1;
        const char *fn = getenv("SVTME_DUMP");
        if (fn && *fn) g_f = fopen(fn, "wb");
    }
    return g_f;
}

static void chunk(const char *tag, const void *p, uint32_t n) {
    fwrite(tag, 1, 4, g_f);
    fwrite(&n, 4, 1, g_f);
    if (n) fwrite(p, 1, n, g_f);
}

static void plane_chunk(const char *tag, const EbPictureBufferDesc *d, uint32_t w, uint32_t h) {
    uint8_t *tight = (uint8_t *)malloc((size_t)w * h);
    for (uint32_t y = 0; y < h; y++)
        memcpy(tight + (size_t)y * w, d->buffer_y + (size_t)(d->org_y + y) * d->stride_y + d->org_x, w);
    chunk(tag, tight, w * h);
    free(tight);
}

/* ---- $SVTME_DUMP_PAIRS: the planes + knobs of every dynamic-GOP detector pair and temporal-filter pair (fixture enc_pairs.npz) */
#include <pthread.h>
static FILE           *g_pf;
static int             g_ptried;
static pthread_mutex_t g_pmu = PTHREAD_MUTEX_INITIALIZER;
static FILE           *pairs_file(void) {
    if (!g_ptried) {
        g_ptried       = 1;
        const char *fn = getenv("SVTME_DUMP_PAIRS");
        if (fn && *fn) g_pf = fopen(fn, "wb");
    }
    return g_pf;
}
static void pchunk(const char *tag, const void *p, uint32_t n) {
    fwrite(tag, 1, 4, g_pf);
    fwrite(&n, 4, 1, g_pf);
    if (n) fwrite(p, 1, n, g_pf);
}
static void pplane(const char *tag, const EbPictureBufferDesc *d, uint32_t w, uint32_t h) {
    uint8_t *tight = (uint8_t *)malloc((size_t)w * h);
    for (uint32_t y = 0; y < h; y++)
        memcpy(tight + (size_t)y * w, d->buffer_y + (size_t)(d->org_y + y) * d->stride_y + d->org_x, w);
    pchunk(tag, tight, w * h);
    free(tight);
}
static const EbPaReferenceObject *pa_of(const PictureParentControlSet *p) {
    return (const EbPaReferenceObject *)p->pa_ref_pic_wrapper->object_ptr;
}

/* before each svt_aom_motion_estimation_b64: clear what the pooled MeSbResults still holds from an earlier picture, so
 * that entries the search leaves unwritten (MVs of pruned references, candidates past the count) read 0 as in the flat ABI */
void svtme_dump_pre_sb(PictureParentControlSet *pcs, uint32_t b64) {
    if (!dump_file()) return;
    const MotionEstimationData *med = pcs->pa_me_data;
    const uint32_t              pus = svtme_shim_pu_count(pcs);
    MeSbResults                *r   = med->me_results[b64];
    memset(r->me_mv_array, 0, sizeof(MvCandidate) * pus * med->max_refs);
    memset(r->me_candidate_array, 0, sizeof(MeCandidate) * pus * med->max_cand);
    memset(r->total_me_candidate_index, 0, pus);
}

/* inside svt_aom_global_motion_estimation, after the bypass test (global_me.c:205) */
void svtme_dump_gm_level(PictureParentControlSet *pcs, uint8_t level) {
    g_gm_level = level;
    g_gm_ds    = pcs->gm_downsample_level;
}

/* at the end of the last-SB hook (me_process.c:287), still under me_processed_b64_mutex */
void svtme_dump_picture(SequenceControlSet *scs, PictureParentControlSet *pcs, MeContext *ctx, EbPictureBufferDesc *input_pic) {
    if (!dump_file()) return;
    SvtMeKnobs k;
    svtme_shim_knobs_from_me_context(scs, pcs, ctx, &k);
    const uint32_t W = pcs->aligned_width, H = pcs->aligned_height, pus = svtme_shim_pu_count(pcs);
    struct {
        uint64_t picture_number;
        uint32_t width, height, b64_total_count, pu_count, slice_type, hme_level1_planes, temporal_layer_index, knob_bytes;
    } head = {pcs->picture_number, W, H, pcs->b64_total_count, pus, (uint32_t)pcs->slice_type,
              (uint32_t)(pcs->enable_hme_level1_flag || pcs->tf_enable_hme_level1_flag), pcs->temporal_layer_index,
              (uint32_t)sizeof(SvtMeKnobs)};
    chunk("HEAD", &head, sizeof(head));
    chunk("KNOB", &k, sizeof(k));

    uint64_t refs[2][4];
    memset(refs, 0xff, sizeof(refs));
    const EbPaReferenceObject *cur = (const EbPaReferenceObject *)pcs->pa_ref_pic_wrapper->object_ptr;
    plane_chunk("CUR0", cur->input_padded_pic, W, H);
    plane_chunk("CUR1", cur->quarter_downsampled_picture_ptr, W >> 1, H >> 1);
    plane_chunk("CUR2", cur->sixteenth_downsampled_picture_ptr, W >> 2, H >> 2);
    for (int li = 0; li < k.num_of_list_to_search; li++)
        for (int ri = 0; ri < k.num_of_ref_pic_to_search[li]; ri++) {
            const EbPaReferenceObject *ro = (const EbPaReferenceObject *)pcs->ref_pa_pic_ptr_array[li][ri]->object_ptr;
            char                       tag[5];
            snprintf(tag, sizeof(tag), "R%d%d0", li, ri);
            refs[li][ri] = ro->picture_number;
            plane_chunk(tag, ro->input_padded_pic, W, H);
        }
    chunk("REFS", refs, sizeof(refs));

    SvtMePictureResults res;
    const size_t        bytes = svtme_shim_results_layout(&res, pcs->b64_total_count, pus, k.max_refs, k.max_cand, NULL);
    void               *mem   = calloc(1, bytes);
    svtme_shim_results_layout(&res, pcs->b64_total_count, pus, k.max_refs, k.max_cand, mem);
    svtme_shim_gather_results(pcs, &res);
    chunk("RES ", mem, (uint32_t)bytes);

    /* the inputs of the last-SB gate (me_process.c:276, global_me.c:162-205), so that a test can rebuild a pcs for the shim */
    struct {
        uint8_t  enabled, pp_enabled, gm_pp_detected, downsample_level, bypass_based_on_me, use_stationary_block,
            calculate_variance, pad;
        uint32_t pic_avg_variance, input_width, input_height;
    } c = {pcs->gm_ctrls.enabled, pcs->gm_ctrls.pp_enabled, pcs->gm_pp_detected, pcs->gm_ctrls.downsample_level,
           pcs->gm_ctrls.bypass_based_on_me, pcs->gm_ctrls.use_stationary_block, scs->calculate_variance, 0,
           pcs->pic_avg_variance, input_pic->width, input_pic->height};
    chunk("GMC ", &c, sizeof(c));
    /* what the reference did: did svt_aom_global_motion_estimation run, and its two locals after the bypass test */
    struct {
        uint8_t ran, level, downsample_level, pad;
    } g = {(uint8_t)(g_gm_level >= 0), (uint8_t)(g_gm_level < 0 ? 0xff : g_gm_level), (uint8_t)(g_gm_ds < 0 ? 0xff : g_gm_ds), 0};
    chunk("GATE", &g, sizeof(g));
    chunk("END ", NULL, 0);
    fflush(g_f);
    free(mem);
    g_gm_level = g_gm_ds = -1;
}

/* ---------------------------------------------------------------------------------------------------------------------
 * TPL source-based search (src_ops_process.c:608-968): with SVTME_DUMP_TPL=<file>, every picture whose source path was
 * just computed (tpl_src_data_ready == 0) is serialised after its last SB (hook behind the tpl_disp_done_semaphore wait
 * of tpl_mc_flow_dispenser, src_ops_process.c:1396).  Record:
 *   THDR  uint64 picture_number; uint32 width, height (enhanced_pic), aligned_width, b64_total_count, pu_count, slice_type,
 *         temporal_layer_index, hierarchical_levels, enable_me_16x16, max_cand, max_refs, max_l0, dispenser_search_level,
 *         disable_intra_pred_nref, intra_mode_end, use_sad_in_src_search, subpel_depth, subsample_tx, tpl_lad_mg, in_loop_ois,
 *         ref_list0_count_try, ref_list1_count_try
 *   TREF  per [2][4]: uint64 picture_number, int32 ref_tpl_group_idx, uint32 excluded (src_ops_process.c:783-785),
 *         uint32 present, uint32 max_width, max_height
 *   TCUR / T<l><i>0   luma of pcs->enhanced_pic / of tpl_ref_ds_ptr_array[l][i].picture_ptr, tight rows
 *   TME   the picture's ME results, flat (svtme_shim_results_layout)
 *   TSRC  per 16x16 grid cell [(height+15)/16][(aligned_width+15)/16]: int16 mv_row, mv_col; int32 best_rf_idx;
 *         uint64 ref_frame_poc; uint8 best_mode, best_intra_mode, pad[2]   (TplSrcStats, coding_unit.h:323-331)
 */
static FILE *g_tf;
static int   g_ttried;

static void tchunk(const char *tag, const void *p, uint32_t n) {
    fwrite(tag, 1, 4, g_tf);
    fwrite(&n, 4, 1, g_tf);
    if (n) fwrite(p, 1, n, g_tf);
}
static void tplane(const char *tag, const EbPictureBufferDesc *d, uint32_t w, uint32_t h) {
    uint8_t *tight = (uint8_t *)malloc((size_t)w * h);
    for (uint32_t y = 0; y < h; y++)
        memcpy(tight + (size_t)y * w, d->buffer_y + (size_t)(d->org_y + y) * d->stride_y + d->org_x, w);
    tchunk(tag, tight, w * h);
    free(tight);
}

void svtme_dump_tpl(SequenceControlSet *scs, PictureParentControlSet *pcs) {
    if (!g_ttried) {
        g_ttried       = 1;
        const char *fn = getenv("SVTME_DUMP_TPL");
        if (fn && *fn) g_tf = fopen(fn, "wb");
    }
    if (!g_tf || pcs->tpl_src_data_ready || scs->tpl_lad_mg == 0) return;
    const EbPictureBufferDesc *in  = pcs->enhanced_pic;
    const uint32_t             pus = svtme_shim_pu_count(pcs);
    struct {
        uint64_t picture_number;
        uint32_t v[24]; /* 23 used; the 24th keeps the record free of uninitialised padding */
    } h = {pcs->picture_number,
           {in->width, in->height, pcs->aligned_width, pcs->b64_total_count, pus, (uint32_t)pcs->slice_type,
            pcs->temporal_layer_index, pcs->hierarchical_levels, pcs->enable_me_16x16, pcs->pa_me_data->max_cand,
            pcs->pa_me_data->max_refs, pcs->pa_me_data->max_l0, pcs->tpl_ctrls.dispenser_search_level,
            pcs->tpl_ctrls.disable_intra_pred_nref, (uint32_t)pcs->tpl_ctrls.intra_mode_end, pcs->tpl_ctrls.use_sad_in_src_search,
            (uint32_t)pcs->tpl_ctrls.subpel_depth, pcs->tpl_ctrls.subsample_tx, scs->tpl_lad_mg, scs->in_loop_ois,
            pcs->ref_list0_count_try, pcs->ref_list1_count_try, 0, 0}};
    tchunk("THDR", &h, sizeof(h));
    struct {
        uint64_t picture_number;
        int32_t  group_idx;
        uint32_t excluded, present, max_width, max_height, pad;
    } r[2][4];
    memset(r, 0, sizeof(r));
    tplane("TCUR", in, in->width, in->height);
    const uint32_t n_list = pcs->slice_type == I_SLICE ? 0 : 2;
    for (uint32_t li = 0; li < n_list; li++) {
        const uint32_t cnt = li == 0 ? pcs->ref_list0_count_try : (pcs->slice_type == B_SLICE ? pcs->ref_list1_count_try : 0);
        for (uint32_t ri = 0; ri < cnt && ri < 4; ri++) {
            const EbPictureBufferDesc *d = (const EbPictureBufferDesc *)pcs->tpl_data.tpl_ref_ds_ptr_array[li][ri].picture_ptr;
            const int32_t              g = pcs->tpl_data.ref_tpl_group_idx[li][ri];
            r[li][ri].picture_number     = pcs->tpl_data.tpl_ref_ds_ptr_array[li][ri].picture_number;
            r[li][ri].group_idx          = g;
            r[li][ri].excluded           = g > 0 && pcs->tpl_data.base_pcs->tpl_valid_pic[g] == 0;
            r[li][ri].present            = d != NULL;
            if (d) {
                char tag[5];
                snprintf(tag, sizeof(tag), "T%u%u0", li, ri);
                r[li][ri].max_width  = d->max_width;
                r[li][ri].max_height = d->max_height;
                tplane(tag, d, in->width, in->height);
            }
        }
    }
    tchunk("TREF", r, sizeof(r));
    if (pcs->slice_type != I_SLICE) {
        SvtMePictureResults res;
        const size_t bytes = svtme_shim_results_layout(&res, pcs->b64_total_count, pus, pcs->pa_me_data->max_refs,
                                                       pcs->pa_me_data->max_cand, NULL);
        void        *mem   = calloc(1, bytes);
        svtme_shim_results_layout(&res, pcs->b64_total_count, pus, pcs->pa_me_data->max_refs, pcs->pa_me_data->max_cand, mem);
        svtme_shim_gather_results(pcs, &res);
        tchunk("TME ", mem, (uint32_t)bytes);
        free(mem);
    }
    const uint32_t gw = (pcs->aligned_width + 15) >> 4, gh = (in->height + 15) >> 4;
    struct Cell {
        int16_t  mv_row, mv_col;
        int32_t  best_rf_idx;
        uint64_t ref_frame_poc;
        uint8_t  best_mode, best_intra_mode, pad[6];
    } *cells = (struct Cell *)calloc((size_t)gw * gh, sizeof(struct Cell));
    for (uint32_t i = 0; i < gw * gh; i++) {
        const TplSrcStats *s   = &pcs->pa_me_data->tpl_src_stats_buffer[i];
        cells[i].mv_row        = s->mv.row;
        cells[i].mv_col        = s->mv.col;
        cells[i].best_rf_idx   = s->best_rf_idx;
        cells[i].ref_frame_poc = s->ref_frame_poc;
        cells[i].best_mode     = s->best_mode;
        cells[i].best_intra_mode = (uint8_t)s->best_intra_mode;
    }
    tchunk("TSRC", cells, (uint32_t)((size_t)gw * gh * sizeof(struct Cell)));
    tchunk("END ", NULL, 0);
    fflush(g_tf);
    free(cells);
}

/* ---- dynamic-GOP detector: one text line per early_hme (pd_process.c:592-634) pair, after all its segments ---- */
void svtme_dump_dg(PictureParentControlSet *src_pcs) {
    static FILE *f;
    static int   tried;
    if (!tried) {
        tried          = 1;
        const char *fn = getenv("SVTME_DUMP_DG");
        if (fn && *fn) f = fopen(fn, "w");
    }
    if (pairs_file()) {
        pthread_mutex_lock(&g_pmu);
        const PictureParentControlSet *rp = src_pcs->dg_detector->ref_pic;
        const DGDetectorMetrics       *dm = &src_pcs->dg_detector->metrics;
        const uint32_t W = src_pcs->aligned_width, H = src_pcs->aligned_height;
        const uint64_t hd[10] = {src_pcs->picture_number, rp->picture_number, W, H, src_pcs->input_resolution,
                                 (uint64_t)(src_pcs->enable_hme_level1_flag || src_pcs->tf_enable_hme_level1_flag), dm->tot_dist,
                                 dm->tot_cplx, dm->tot_active, (uint64_t)(int64_t)dm->sum_in_vectors};
        pchunk("DGHD", hd, sizeof(hd));
        pplane("DGS0", pa_of(src_pcs)->input_padded_pic, W, H);
        pplane("DGR0", pa_of(rp)->input_padded_pic, W, H);
        pplane("DGS2", pa_of(src_pcs)->sixteenth_downsampled_picture_ptr, W >> 2, H >> 2);
        pplane("DGR2", pa_of(rp)->sixteenth_downsampled_picture_ptr, W >> 2, H >> 2);
        pchunk("END ", NULL, 0);
        fflush(g_pf);
        pthread_mutex_unlock(&g_pmu);
    }
    if (!f) return;
    const DGDetectorMetrics *m = &src_pcs->dg_detector->metrics;
    fprintf(f, "%llu %llu %llu %u %u %d\n", (unsigned long long)src_pcs->picture_number,
            (unsigned long long)src_pcs->dg_detector->ref_pic->picture_number, (unsigned long long)m->tot_dist, m->tot_cplx,
            m->tot_active, m->sum_in_vectors);
    fflush(f);
}

/* ---- temporal-filter ME: one text line per (centre, neighbour, block) with what the filter reads from the MeContext ---- */
void svtme_dump_tf_b64(PictureParentControlSet *centre_pcs, PictureParentControlSet *ref_pcs, MeContext *ctx, unsigned b64_index) {
    static FILE *f;
    static int   tried;
    if (!tried) {
        tried          = 1;
        const char *fn = getenv("SVTME_DUMP_TF");
        if (fn && *fn) f = fopen(fn, "w");
    }
    if (pairs_file()) {
        static uint64_t seen_centre = ~0ull, seen_refs[ALTREF_MAX_NFRAMES];
        static int      n_seen;
        pthread_mutex_lock(&g_pmu); /* a second TF thread waits here until the centre plane is on file, then may filter */
        const int new_centre = seen_centre != centre_pcs->picture_number;
        if (new_centre) {
            seen_centre = centre_pcs->picture_number;
            n_seen      = 0;
        }
        int known = 0;
        for (int i = 0; i < n_seen; i++) known |= seen_refs[i] == ref_pcs->picture_number;
        if (!known && n_seen < ALTREF_MAX_NFRAMES) {
            seen_refs[n_seen++] = ref_pcs->picture_number;
            SvtMeKnobs k;
            svtme_shim_tf_knobs_from_me_context(centre_pcs->scs, centre_pcs, ctx, &k);
            const uint32_t W = centre_pcs->aligned_width, H = centre_pcs->aligned_height;
            const uint64_t hd[6] = {centre_pcs->picture_number, ref_pcs->picture_number, W, H,
                                    (uint64_t)(centre_pcs->enable_hme_level1_flag || centre_pcs->tf_enable_hme_level1_flag), sizeof(k)};
            pchunk("TFHD", hd, sizeof(hd));
            pchunk("TFKN", &k, sizeof(k));
            if (new_centre) pplane("TFC0", pa_of(centre_pcs)->input_padded_pic, W, H);
            pplane("TFR0", pa_of(ref_pcs)->input_padded_pic, W, H);
            pchunk("END ", NULL, 0);
            fflush(g_pf);
        }
        pthread_mutex_unlock(&g_pmu);
    }
    /* the knob block of the pair (first block only) next to the TfControls fields svtme_derive_tf_knobs takes */
    static FILE *fk;
    static int   ktried;
    if (!ktried) {
        ktried         = 1;
        const char *fn = getenv("SVTME_DUMP_TF_KNOBS");
        if (fn && *fn) fk = fopen(fn, "wb");
    }
    if (fk && b64_index == 0) {
        SvtMeKnobs k;
        svtme_shim_tf_knobs_from_me_context(centre_pcs->scs, centre_pcs, ctx, &k);
        const EbPictureBufferDesc *d = ((const EbPaReferenceObject *)centre_pcs->pa_ref_pic_wrapper->object_ptr)->input_padded_pic;
        const uint64_t hdr[12] = {centre_pcs->picture_number, ref_pcs->picture_number, (d->width + 7u) & ~7u, (d->height + 7u) & ~7u,
                                  centre_pcs->tf_ctrls.hme_me_level, centre_pcs->tf_ctrls.qp_opt, centre_pcs->tf_ctrls.me_exit_th,
                                  centre_pcs->scs->static_config.qp, centre_pcs->temporal_layer_index, centre_pcs->is_ref,
                                  centre_pcs->scs->static_config.enc_mode, sizeof(k)};
        fwrite(hdr, sizeof(hdr), 1, fk);
        fwrite(&k, sizeof(k), 1, fk);
        fflush(fk);
    }
    if (!f) return;
    const int exited = ctx->tf_use_pred_64x64_only_th == (uint8_t)~0;
    uint64_t  hs = 1469598103934665603ull, hm = hs; /* FNV-1a over the 85 SADs / 85 MVs; stale when the search exited early */
    if (!exited)
        for (int i = 0; i < SQUARE_PU_COUNT; i++) {
            hs = (hs ^ ctx->p_sb_best_sad[0][0][i]) * 1099511628211ull;
            hm = (hm ^ ctx->p_sb_best_mv[0][0][i]) * 1099511628211ull;
        }
    fprintf(f, "%llu %llu %u %d %d %d %016llx %016llx\n", (unsigned long long)centre_pcs->picture_number,
            (unsigned long long)ref_pcs->picture_number, b64_index, exited, ctx->search_results[0][0].hme_sc_x,
            ctx->search_results[0][0].hme_sc_y, (unsigned long long)hs, (unsigned long long)hm);
}
