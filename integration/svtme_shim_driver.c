/*
 * svtme_shim_driver.c -- the part of the encoder-side shim that calls libsvtme_cuda.so (see svtme_shim.h).
 * C99, compiled against the reference's headers.  One call per TASK_PAME picture replaces the SB loop of
 * svt_aom_motion_estimation_kernel (Source/Lib/Codec/me_process.c:174-290).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "svtme_shim.h"
#include "reference_object.h"

typedef struct ShimPic {
    uint64_t  picture_number;
    uint64_t  last_use;
    SvtMePic *pic; /* NULL: free slot */
} ShimPic;

/* temporal filter: the results of one (centre, neighbour) pair, whole picture */
#define SHIM_TF_PAIRS ALTREF_MAX_NFRAMES /* definitions.h: 33 */
typedef struct ShimTfPair {
    uint64_t  ref_picture_number;
    int       valid;
    void     *mem;
    size_t    bytes;
    uint32_t *best_sad, *best_mv; /* [b64][85] */
    int16_t  *hme_sc;             /* [b64][2][4][2] */
    uint8_t  *exit_sb;            /* [b64] */
    uint32_t  blk_counts[2];
} ShimTfPair;

struct SvtMeShim {
    SvtMeHandle *h;
    uint64_t     tf_centre_number;
    SvtMePic    *tf_centre;
    ShimTfPair   tf_pairs[SHIM_TF_PAIRS];
    ShimPic     *pics;
    uint32_t     n_pics;
    uint64_t     clock;
    void        *res_mem; /* flat result buffers of one picture, grown on demand */
    size_t       res_bytes;
    char         err[256];
};

const char *svtme_shim_last_error(SvtMeShim *s) { return s ? s->err : "svtme shim: no instance"; }

static int32_t shim_fail(SvtMeShim *s, const char *what) {
    snprintf(s->err, sizeof(s->err), "%s: %s", what, svtme_last_error(s->h));
    return -1;
}

int32_t svtme_shim_open(const SequenceControlSet *scs, int32_t device, uint32_t pictures, int hme_level1_planes, SvtMeShim **out) {
    SvtMeShim *s = (SvtMeShim *)calloc(1, sizeof(*s));
    if (!s) return -1;
    SvtMeDeviceCfg cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.device       = device;
    cfg.max_width    = scs->max_input_luma_width;
    cfg.max_height   = scs->max_input_luma_height;
    /* + 2: the uncached pictures of svtme_shim_dg_picture (2) / svtme_shim_tf_b64 (centre + 1) */
    cfg.max_pictures = pictures ? (pictures < 9 ? 9 : pictures) + 3 : 0;
    cfg.max_inflight = 4;
    /* pic_analysis_process.c:1968-1992: the sixteenth plane is decimated from the quarter plane when the HME level 1
     * planes exist (step 2), from the full plane otherwise (step 4) */
    cfg.sixteenth_from_full = !hme_level1_planes;
    if (svtme_init(&cfg, &s->h) != 0) {
        free(s);
        return -1;
    }
    /* at least one search's worth (source + 2 x 4 references): the least recently used picture is recycled when the cache is
     * full, and that must never be one this very search has just collected */
    s->n_pics = pictures ? (pictures < 9 ? 9 : pictures) : 64;
    s->pics   = (ShimPic *)calloc(s->n_pics, sizeof(ShimPic));
    *out      = s;
    return 0;
}

void svtme_shim_close(SvtMeShim *s) {
    if (!s) return;
    for (uint32_t i = 0; i < s->n_pics; i++)
        if (s->pics[i].pic) svtme_picture_release(s->h, s->pics[i].pic);
    if (s->tf_centre) svtme_picture_release(s->h, s->tf_centre);
    for (int i = 0; i < SHIM_TF_PAIRS; i++) free(s->tf_pairs[i].mem);
    svtme_destroy(s->h);
    free(s->pics);
    free(s->res_mem);
    free(s);
}

void svtme_shim_forget(SvtMeShim *s, uint64_t picture_number) {
    for (uint32_t i = 0; i < s->n_pics; i++)
        if (s->pics[i].pic && s->pics[i].picture_number == picture_number) {
            svtme_picture_release(s->h, s->pics[i].pic);
            s->pics[i].pic = NULL;
        }
}

/* The device planes of one PA reference object: resident already, or uploaded from its padded 8-bit luma
 * (EbPaReferenceObject.input_padded_pic, reference_object.h; the planes ME reads at me_process.c:139-143, 248-262). */
static SvtMePic *shim_resident_desc(SvtMeShim *s, const EbPictureBufferDesc *d, uint64_t picture_number) {
    ShimPic *free_slot = NULL, *oldest = NULL;
    for (uint32_t i = 0; i < s->n_pics; i++) {
        ShimPic *p = &s->pics[i];
        if (p->pic && p->picture_number == picture_number) {
            p->last_use = ++s->clock;
            return p->pic;
        }
        if (!p->pic) {
            if (!free_slot) free_slot = p;
        } else if (!oldest || p->last_use < oldest->last_use)
            oldest = p;
    }
    if (!free_slot) { /* the encoder did not call svtme_shim_forget: recycle the least recently used picture */
        svtme_picture_release(s->h, oldest->pic);
        oldest->pic = NULL;
        free_slot   = oldest;
    }
    const uint8_t             *luma = d->buffer_y + (size_t)d->org_y * d->stride_y + d->org_x;
    /* the ME planes are multiples of 8 wide/high (pcs->aligned_width/height): the padded picture holds them */
    const uint32_t w = (d->width + 7u) & ~7u, h = (d->height + 7u) & ~7u;
    SvtMePic      *pic = NULL;
    if (svtme_picture_register(s->h, picture_number, luma, d->stride_y, w, h, &pic) != 0) return NULL;
    free_slot->pic            = pic;
    free_slot->picture_number = picture_number;
    free_slot->last_use       = ++s->clock;
    return pic;
}

static SvtMePic *shim_resident(SvtMeShim *s, const EbPaReferenceObject *obj, uint64_t picture_number) {
    return shim_resident_desc(s, obj->input_padded_pic, picture_number);
}

int32_t svtme_shim_me_picture(SvtMeShim *s, const SequenceControlSet *scs, PictureParentControlSet *pcs, const MeContext *ctx,
                              const EbPictureBufferDesc *input_pic, SvtMeShimGmGate *gate) {
    SvtMeKnobs k;
    svtme_shim_knobs_from_me_context(scs, pcs, ctx, &k);

    const EbPaReferenceObject *cur_obj = (const EbPaReferenceObject *)pcs->pa_ref_pic_wrapper->object_ptr;
    SvtMePic                  *cur     = shim_resident(s, cur_obj, pcs->picture_number);
    if (!cur) return shim_fail(s, "register source picture");
    SvtMePic *refs[SVTME_MAX_LISTS][SVTME_MAX_REFS_PER_LIST];
    memset(refs, 0, sizeof(refs));
    for (int li = 0; li < k.num_of_list_to_search; li++)
        for (int ri = 0; ri < k.num_of_ref_pic_to_search[li]; ri++) {
            const EbPaReferenceObject *ro = (const EbPaReferenceObject *)pcs->ref_pa_pic_ptr_array[li][ri]->object_ptr;
            refs[li][ri]                  = shim_resident(s, ro, ro->picture_number);
            if (!refs[li][ri]) return shim_fail(s, "register reference picture");
        }

    SvtMePictureResults res;
    const uint32_t      pus  = svtme_shim_pu_count(pcs);
    const size_t        need = svtme_shim_results_layout(&res, pcs->b64_total_count, pus, k.max_refs, k.max_cand, NULL);
    if (need > s->res_bytes) {
        free(s->res_mem);
        s->res_mem   = malloc(need);
        s->res_bytes = s->res_mem ? need : 0;
        if (!s->res_mem) {
            snprintf(s->err, sizeof(s->err), "out of host memory for %zu result bytes", need);
            return -1;
        }
    }
    svtme_shim_results_layout(&res, pcs->b64_total_count, pus, k.max_refs, k.max_cand, s->res_mem);

    uint64_t ticket = 0;
    if (svtme_picture_search(s->h, cur, (SvtMePic * const(*)[SVTME_MAX_REFS_PER_LIST]) refs, &k, &res, &ticket) != 0)
        return shim_fail(s, "svtme_picture_search");
    if (svtme_search_wait(s->h, ticket) != 0) return shim_fail(s, "svtme_search_wait");

    svtme_shim_scatter_results(pcs, &res);
    pcs->me_processed_b64_count = pcs->b64_total_count; /* me_process.c:273 */
    svtme_shim_gm_gate(pcs, input_pic, res.picture_sums, gate);
    return 0;
}

/* ---- dynamic-GOP detector, HME level 0 (pd_process.c:492-590) ---- */

static SvtMePic *shim_upload_once(SvtMeShim *s, const PictureParentControlSet *p) {
    const EbPictureBufferDesc *d    = ((const EbPaReferenceObject *)p->pa_ref_pic_wrapper->object_ptr)->input_padded_pic;
    const uint8_t             *luma = d->buffer_y + (size_t)d->org_y * d->stride_y + d->org_x;
    SvtMePic                  *pic  = NULL;
    if (svtme_picture_register(s->h, p->picture_number, luma, d->stride_y, (d->width + 7u) & ~7u, (d->height + 7u) & ~7u, &pic) != 0)
        return NULL;
    return pic;
}

int32_t svtme_shim_dg_picture(SvtMeShim *s, PictureParentControlSet *ppcs) {
    SvtMePic *src = shim_upload_once(s, ppcs);
    SvtMePic *ref = src ? shim_upload_once(s, ppcs->dg_detector->ref_pic) : NULL;
    int32_t   rc  = -1;
    if (src && ref) {
        SvtMeDgMetrics m;
        rc = svtme_dg_detector_hme_level0(s->h, src, ref, ppcs->input_resolution, &m, NULL, NULL, NULL, NULL);
        if (rc == 0) { /* pd_process.c:543-580, summed over the picture */
            DGDetectorMetrics *dst = &ppcs->dg_detector->metrics;
            dst->tot_dist += m.tot_dist;
            dst->tot_cplx += m.tot_cplx;
            dst->tot_active += m.tot_active;
            dst->sum_in_vectors += m.sum_in_vectors;
        }
    }
    if (rc != 0) shim_fail(s, "svtme_dg_detector_hme_level0");
    if (src) svtme_picture_release(s->h, src);
    if (ref) svtme_picture_release(s->h, ref);
    return rc;
}

/* ---- temporal-filter ME (temporal_filtering.c:3129-3174) ---- */

static ShimTfPair *shim_tf_pair(SvtMeShim *s, const SequenceControlSet *scs, PictureParentControlSet *centre,
                                PictureParentControlSet *ref_pcs, const MeContext *ctx) {
    if (!s->tf_centre || s->tf_centre_number != centre->picture_number) { /* a new centre picture: drop the previous one */
        if (s->tf_centre) svtme_picture_release(s->h, s->tf_centre);
        for (int i = 0; i < SHIM_TF_PAIRS; i++) s->tf_pairs[i].valid = 0;
        s->tf_centre = shim_upload_once(s, centre);
        if (!s->tf_centre) {
            shim_fail(s, "register the temporal filter's centre picture");
            return NULL;
        }
        s->tf_centre_number = centre->picture_number;
    }
    ShimTfPair *p = NULL;
    for (int i = 0; i < SHIM_TF_PAIRS; i++) {
        if (s->tf_pairs[i].valid && s->tf_pairs[i].ref_picture_number == ref_pcs->picture_number) return &s->tf_pairs[i];
        if (!p && !s->tf_pairs[i].valid) p = &s->tf_pairs[i];
    }
    if (!p) {
        snprintf(s->err, sizeof(s->err), "more than %d neighbours of one temporal-filter centre picture", SHIM_TF_PAIRS);
        return NULL;
    }
    SvtMeKnobs k;
    svtme_shim_tf_knobs_from_me_context(scs, centre, ctx, &k);
    const EbPictureBufferDesc *d     = ((const EbPaReferenceObject *)centre->pa_ref_pic_wrapper->object_ptr)->input_padded_pic;
    const uint32_t             w = (d->width + 7u) & ~7u, h = (d->height + 7u) & ~7u;
    const size_t               n_b64 = (size_t)((w + 63) / 64) * ((h + 63) / 64);
    const size_t               at_mv = n_b64 * 85 * 4, at_sc = 2 * at_mv, at_exit = at_sc + n_b64 * 16 * 2, need = at_exit + n_b64;
    if (need > p->bytes) {
        free(p->mem);
        p->mem   = malloc(need);
        p->bytes = p->mem ? need : 0;
        if (!p->mem) {
            snprintf(s->err, sizeof(s->err), "out of host memory for %zu temporal-filter result bytes", need);
            return NULL;
        }
    }
    p->best_sad = (uint32_t *)p->mem;
    p->best_mv  = (uint32_t *)((uint8_t *)p->mem + at_mv);
    p->hme_sc   = (int16_t *)((uint8_t *)p->mem + at_sc);
    p->exit_sb  = (uint8_t *)p->mem + at_exit;

    SvtMePic *ref = shim_upload_once(s, ref_pcs);
    if (!ref) {
        shim_fail(s, "register a temporal-filter neighbour");
        return NULL;
    }
    SvtMePic *refs[SVTME_MAX_LISTS][SVTME_MAX_REFS_PER_LIST];
    memset(refs, 0, sizeof(refs));
    refs[0][0] = ref;
    SvtMePictureResults res;
    memset(&res, 0, sizeof(res));
    res.b64_total_count = (uint32_t)n_b64;
    res.pu_count        = SQUARE_PU_COUNT;
    res.tf_best_sad     = p->best_sad;
    res.tf_best_mv      = p->best_mv;
    res.hme_sc          = p->hme_sc;
    res.tf_exit_sb      = p->exit_sb;
    res.tf_blk_counts   = p->blk_counts;
    uint64_t ticket     = 0;
    int32_t  rc = svtme_picture_search(s->h, s->tf_centre, (SvtMePic * const(*)[SVTME_MAX_REFS_PER_LIST]) refs, &k, &res, &ticket);
    if (rc == 0) rc = svtme_search_wait(s->h, ticket);
    if (rc != 0) shim_fail(s, "temporal-filter search");
    svtme_picture_release(s->h, ref);
    if (rc != 0) return NULL;
    p->ref_picture_number = ref_pcs->picture_number;
    p->valid              = 1;
    return p;
}

int32_t svtme_shim_tf_b64(SvtMeShim *s, const SequenceControlSet *scs, PictureParentControlSet *centre_pcs,
                          PictureParentControlSet *ref_pcs, MeContext *ctx, uint32_t b64_index) {
    const ShimTfPair *p = shim_tf_pair(s, scs, centre_pcs, ref_pcs, ctx);
    if (!p) return -1;
    svtme_shim_tf_fill_b64(ctx, p->best_sad + (size_t)b64_index * SQUARE_PU_COUNT, p->best_mv + (size_t)b64_index * SQUARE_PU_COUNT,
                           p->hme_sc[(size_t)b64_index * 16], p->hme_sc[(size_t)b64_index * 16 + 1], p->exit_sb[b64_index]);
    return 0;
}

/* ---- TPL source-based search (src_ops_process.c:608-887) over the ME results the pcs holds ---- */

size_t svtme_shim_tpl_layout(SvtMeTplResults *res, uint32_t width, uint32_t height, void *mem) {
    const uint32_t gw = (width + 15) >> 4, gh = (height + 15) >> 4;
    const size_t   cells = (size_t)gw * gh, elem[7] = {1, 1, 1, 4, 8, 4, 4};
    size_t         off = 0, at[7];
    for (int i = 0; i < 7; i++) {
        off   = (off + 15) & ~(size_t)15;
        at[i] = off;
        off += elem[i] * cells;
    }
    if (mem) {
        uint8_t *b = (uint8_t *)mem;
        memset(res, 0, sizeof(*res));
        res->grid_width = gw; res->grid_height = gh;
        res->block_size = b + at[0]; res->best_mode = b + at[1]; res->best_rf_idx = (int8_t *)(b + at[2]);
        res->mv = (int16_t *)(b + at[3]); res->ref_frame_poc = (uint64_t *)(b + at[4]);
        res->best_inter_cost = (uint32_t *)(b + at[5]); res->best_intra_cost = (uint32_t *)(b + at[6]);
    }
    return (off + 15) & ~(size_t)15;
}

/* 1: the picture's TPL controls are outside what the library restates (keep the CPU path); 0: `out` is filled; < 0: error */
int32_t svtme_shim_tpl_picture(SvtMeShim *s, const SequenceControlSet *scs, PictureParentControlSet *pcs, SvtMeTplResults *out) {
    const TplControls *tc = &pcs->tpl_ctrls;
    if (!tc->use_sad_in_src_search || tc->subpel_depth != FULL_PEL || tc->intra_mode_end != DC_PRED || scs->in_loop_ois == 0) return 1;
    SvtMeTplCfg c;
    memset(&c, 0, sizeof(c));
    c.dispenser_search_level = tc->dispenser_search_level;
    c.use_sad_in_src_search = c.subpel_full_pel = c.intra_dc_only = 1;
    c.disable_intra_pred = tc->disable_intra_pred_nref && pcs->temporal_layer_index == pcs->hierarchical_levels; /* :567 */
    c.slice_type_i       = pcs->slice_type == I_SLICE;                                                           /* :767 */
    /* the pcs / scs part of the knob block (the search kernels do not run: only list shapes and array strides are read) */
    SvtMeKnobs k;
    memset(&k, 0, sizeof(k));
    k.picture_number           = pcs->picture_number;
    k.enable_me_8x8            = pcs->enable_me_8x8;
    k.enable_me_16x16          = pcs->enable_me_16x16;
    k.max_number_of_pus_per_sb = pcs->max_number_of_pus_per_sb;
    k.max_cand                 = pcs->pa_me_data->max_cand;
    k.max_refs                 = pcs->pa_me_data->max_refs;
    k.max_l0                   = pcs->pa_me_data->max_l0;
    k.num_of_list_to_search    = pcs->slice_type == P_SLICE ? 1 : 2;
    k.num_of_ref_pic_to_search[0] = pcs->ref_list0_count_try;
    k.num_of_ref_pic_to_search[1] = pcs->slice_type == B_SLICE ? pcs->ref_list1_count_try : 0;
    k.num_hme_sa_w = k.num_hme_sa_h = 2;
    k.me_type      = SVTME_ME_OPEN_LOOP;

    SvtMePic *cur = shim_resident_desc(s, pcs->enhanced_pic, pcs->picture_number);
    if (!cur) return shim_fail(s, "register source picture (TPL)");
    SvtMePic *refs[SVTME_MAX_LISTS][SVTME_MAX_REFS_PER_LIST];
    memset(refs, 0, sizeof(refs));
    if (!c.slice_type_i)
        for (int li = 0; li < k.num_of_list_to_search; li++)
            for (int ri = 0; ri < k.num_of_ref_pic_to_search[li]; ri++) {
                const EbDownScaledBufDescPtrArray *r = &pcs->tpl_data.tpl_ref_ds_ptr_array[li][ri]; /* :111-144 */
                const int32_t                      g = pcs->tpl_data.ref_tpl_group_idx[li][ri];
                c.ref_poc[li][ri]      = r->picture_number;
                c.ref_excluded[li][ri] = g > 0 && pcs->tpl_data.base_pcs->tpl_valid_pic[g] == 0;   /* :783-785 */
                refs[li][ri]           = shim_resident_desc(s, (const EbPictureBufferDesc *)r->picture_ptr, r->picture_number);
                if (!refs[li][ri]) return shim_fail(s, "register reference picture (TPL)");
            }
    SvtMePictureResults me;
    memset(&me, 0, sizeof(me));
    if (!c.slice_type_i) {
        const uint32_t pus  = svtme_shim_pu_count(pcs);
        const size_t   need = svtme_shim_results_layout(&me, pcs->b64_total_count, pus, k.max_refs, k.max_cand, NULL);
        if (need > s->res_bytes) {
            free(s->res_mem);
            s->res_mem   = malloc(need);
            s->res_bytes = s->res_mem ? need : 0;
            if (!s->res_mem) return -1;
        }
        svtme_shim_results_layout(&me, pcs->b64_total_count, pus, k.max_refs, k.max_cand, s->res_mem);
        svtme_shim_gather_results(pcs, &me);
    }
    uint64_t ticket = 0;
    if (svtme_tpl_src_search(s->h, cur, (SvtMePic * const(*)[SVTME_MAX_REFS_PER_LIST]) refs, &k, c.slice_type_i ? NULL : &me, &c, out, &ticket) != 0)
        return shim_fail(s, "svtme_tpl_src_search");
    if (svtme_search_wait(s->h, ticket) != 0) return shim_fail(s, "svtme_search_wait (TPL)");
    return 0;
}
