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

struct SvtMeShim {
    SvtMeHandle *h;
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
    cfg.max_pictures = pictures;
    cfg.max_inflight = 4;
    /* pic_analysis_process.c:1968-1992: the sixteenth plane is decimated from the quarter plane when the HME level 1
     * planes exist (step 2), from the full plane otherwise (step 4) */
    cfg.sixteenth_from_full = !hme_level1_planes;
    if (svtme_init(&cfg, &s->h) != 0) {
        free(s);
        return -1;
    }
    s->n_pics = pictures ? pictures : 64;
    s->pics   = (ShimPic *)calloc(s->n_pics, sizeof(ShimPic));
    *out      = s;
    return 0;
}

void svtme_shim_close(SvtMeShim *s) {
    if (!s) return;
    for (uint32_t i = 0; i < s->n_pics; i++)
        if (s->pics[i].pic) svtme_picture_release(s->h, s->pics[i].pic);
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
static SvtMePic *shim_resident(SvtMeShim *s, const EbPaReferenceObject *obj, uint64_t picture_number) {
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
    const EbPictureBufferDesc *d    = obj->input_padded_pic;
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
