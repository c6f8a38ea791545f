/*
 * svtme_shim.h -- the encoder-side glue between SVT-AV1's ME kernel thread and libsvtme_cuda.so.
 *
 * This file is compiled AGAINST THE REFERENCE'S OWN HEADERS (pcs.h, me_context.h, sequence_control_set.h,
 * me_sb_results.h): it is what a maintainer drops into Source/Lib/Codec next to me_process.c.  Nothing in the
 * product library depends on it; tests build it into oracle/_ref/libsvtme_shim.so (oracle/Makefile target `shim`)
 * and call it from tests/test_enc_replay.py.
 *
 * It replaces, for TASK_PAME messages, the body of svt_aom_motion_estimation_kernel between
 * "if (pcs->slice_type != I_SLICE)" and the open-loop intra search (Source/Lib/Codec/me_process.c:163-290):
 *   - the per-message MeContext/pcs fields the SB loop copies             -> svtme_shim_knobs_from_me_context
 *   - the SB loop + svt_aom_motion_estimation_b64 (motion_estimation.c:3076) -> svtme_shim_me_picture
 *   - the writes into pcs->pa_me_data->me_results[] and the eight pcs arrays  -> svtme_shim_scatter_results
 *   - the last-SB hook (me_process.c:272-287) and the picture-level sums that open
 *     svt_aom_global_motion_estimation (global_me.c:162-215)               -> svtme_shim_gm_gate
 */
#ifndef SVTME_SHIM_H
#define SVTME_SHIM_H

#include "pcs.h"
#include "sequence_control_set.h"
#include "me_context.h"
#include "me_sb_results.h"
#include "svtme_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- pure host code, no CUDA symbol needed (svtme_shim.c) ---- */

/* The flat knob block of one picture from the encoder's own derived state: `ctx` after svt_aom_sig_deriv_me
 * (me_process.c:121), `pcs` / `scs` as the SB loop reads them (me_process.c:216-227, motion_estimation.c passim). */
void svtme_shim_knobs_from_me_context(const SequenceControlSet *scs, const PictureParentControlSet *pcs, const MeContext *ctx,
                                      SvtMeKnobs *k);
/* The same for one temporal-filter search (ME_MCTF): `ctx` as svt_aom_motion_estimation_b64 sees it at temporal_filtering.c:3169 */
void svtme_shim_tf_knobs_from_me_context(const SequenceControlSet *scs, const PictureParentControlSet *centre_pcs, const MeContext *ctx,
                                         SvtMeKnobs *k);

/* 85 / 21 / 5 square PUs per 64x64 block, as MeSbResults is allocated (pcs.c:106-110). */
uint32_t svtme_shim_pu_count(const PictureParentControlSet *pcs);

/* Flat picture results -> pcs->pa_me_data->me_results[b64] (MeSbResults, me_sb_results.h:44-51) and the eight per-block
 * pcs arrays (pcs.h:871-881).  `res` must have been produced with this picture's knobs. */
void svtme_shim_scatter_results(PictureParentControlSet *pcs, const SvtMePictureResults *res);
/* The reverse copy (used by the dump hook of the encoder overlay and by tests). */
void svtme_shim_gather_results(const PictureParentControlSet *pcs, SvtMePictureResults *res);

/* Decision of the last-SB hook.  `sums` = SvtMePictureResults.picture_sums of the picture. */
typedef struct SvtMeShimGmGate {
    uint8_t run_gm;           /* me_process.c:276: gm_ctrls.enabled && (!gm_ctrls.pp_enabled || gm_pp_detected);
                               * 0 => the caller clears pcs->is_global_motion (:280) */
    uint8_t level;            /* global_motion_estimation_level 0..3 after the bypass test (global_me.c:171-205) */
    uint8_t downsample_level; /* pcs->gm_downsample_level (global_me.c:184-196) */
    uint32_t average_me_sad;  /* global_me.c:169 */
} SvtMeShimGmGate;
void svtme_shim_gm_gate(const PictureParentControlSet *pcs, const EbPictureBufferDesc *input_pic, const uint64_t sums[4],
                        SvtMeShimGmGate *gate);

/* Bytes of one picture's flat result buffers for (b64 count, pu count, max_refs, max_cand); fills the pointers of `res`
 * into `mem` when it is not NULL.  The optional parity fields are left NULL except picture_sums. */
size_t svtme_shim_results_layout(SvtMePictureResults *res, uint32_t b64_total_count, uint32_t pu_count, uint32_t max_refs,
                                 uint32_t max_cand, void *mem);

/* ---- the driver: calls libsvtme_cuda.so (svtme_shim_driver.c) ---- */

typedef struct SvtMeShim SvtMeShim;
/* One per encoder instance (enc_handle.c:2102-2110 builds the ME contexts there). `pictures` >= size of the PA reference pool.
 * hme_level1_planes = pcs->enable_hme_level1_flag || pcs->tf_enable_hme_level1_flag, constant over an encode
 * (pic_analysis_process.c:1949-1992 picks the decimation chain of the sixteenth plane with it). */
int32_t svtme_shim_open(const SequenceControlSet *scs, int32_t device, uint32_t pictures, int hme_level1_planes, SvtMeShim **out);
void    svtme_shim_close(SvtMeShim *s);
/* The whole picture's open-loop ME: to be called ONCE per TASK_PAME picture (the encoder posts one segment message per
 * picture when the shim is active: pcs->me_segments_total_count = 1, pd_process.c:3545-3557).  Uploads the luma of `pcs` and
 * of every reference that is not resident yet, searches, scatters the results into `pcs`, evaluates the GM gate. */
int32_t svtme_shim_me_picture(SvtMeShim *s, const SequenceControlSet *scs, PictureParentControlSet *pcs, const MeContext *ctx,
                              const EbPictureBufferDesc *input_pic, SvtMeShimGmGate *gate);
/* TPL source-based search of one picture (the `tpl_src_data_ready == 0` decision of tpl_mc_flow_dispenser_sb_generic,
 * src_ops_process.c:608-887) over the ME results `pcs` holds; to be called once per picture before its dispenser tasks are
 * posted (src_ops_process.c:1378).  `out`: arrays laid out by svtme_shim_tpl_layout(out, pcs->enhanced_pic->width, height, mem).
 * Returns 1 when the picture's TPL controls are outside what the library restates (the CPU path stays), 0 on success. */
size_t  svtme_shim_tpl_layout(SvtMeTplResults *res, uint32_t width, uint32_t height, void *mem);
int32_t svtme_shim_tpl_picture(SvtMeShim *s, const SequenceControlSet *scs, PictureParentControlSet *pcs, SvtMeTplResults *out);
/* Dynamic-GOP detector (TASK_DG_DETECTOR_HME, me_process.c:327-329): the level-0 search of dg_detector_hme_level0
 * (pd_process.c:492-590) for ALL segments of `ppcs` against ppcs->dg_detector->ref_pic, added into
 * ppcs->dg_detector->metrics (the caller holds metrics_mutex or is the only segment doing work; seg_completed and the
 * frame_done semaphore stay with the caller).  Both pictures are uploaded for this call only and not cached: the detector
 * runs in picture decision, BEFORE temporal filtering rewrites the source planes ME will read later. */
int32_t svtme_shim_dg_picture(SvtMeShim *s, PictureParentControlSet *ppcs);
/* Temporal-filter ME (produce_temporally_filtered_pic, temporal_filtering.c:3129-3174).  Replaces the call of
 * svt_aom_motion_estimation_b64 at :3169 for block `b64_index` of `centre_pcs` against `ref_pcs`: the first request of a
 * (centre, reference) pair searches the WHOLE picture on the GPU (ME_MCTF mode of svtme_picture_search) and keeps the pair's
 * results until the next centre picture; every request then puts the block's outputs where the filter reads them:
 * ctx->p_sb_best_sad / p_sb_best_mv [0][0] with the p_best_* pointers on them, search_results[0][0].hme_sc_x / y,
 * tf_use_pred_64x64_only_th = 0xFF when the HME SAD was under tf_me_exit_th (motion_estimation.c:3109-3113), and
 * tf_tot_horz_blks / tf_tot_vert_blks (:2469-2474).  The centre picture is uploaded at its first request -- before any block
 * of it has been filtered in place -- and the neighbours per pair; neither goes into the open-loop picture cache.
 * Not thread-safe: TF segment threads call it under one lock. */
/* the pure-host half of svtme_shim_tf_b64: one block's row of the pair's results into the MeContext */
void svtme_shim_tf_fill_b64(MeContext *ctx, const uint32_t *best_sad, const uint32_t *best_mv, int16_t hme_sc_x, int16_t hme_sc_y,
                            int exited);
int32_t svtme_shim_tf_b64(SvtMeShim *s, const SequenceControlSet *scs, PictureParentControlSet *centre_pcs,
                          PictureParentControlSet *ref_pcs, MeContext *ctx, uint32_t b64_index);
/* To be called where the encoder releases a PA reference object (pic_mgr / pd_process release of pa_ref_pic_wrapper). */
void svtme_shim_forget(SvtMeShim *s, uint64_t picture_number);
const char *svtme_shim_last_error(SvtMeShim *s);

#ifdef __cplusplus
}
#endif
#endif
