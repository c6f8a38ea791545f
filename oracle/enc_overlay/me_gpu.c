/*
 * me_gpu.c -- TEST INFRASTRUCTURE ONLY.  The drop-in demonstrated end to end: linked into the C-only reference encoder that
 * oracle/Makefile target `enc` builds, and called from the hook make_overlay.py inserts in front of the 64x64 block loop of
 * me_process.c.  With SVTME_GPU_ME=1 every TASK_PAME picture's open-loop ME is done by libsvtme_cuda.so through the encoder-side
 * shim (integration/svtme_shim_driver.c: upload, search, scatter into MeSbResults + the eight pcs arrays, global-motion gate) --
 * exactly the call INTEGRATION.md section 3 describes -- and the rest of the encoder runs unchanged on those results.
 * tests/test_enc_replay.py encodes the same clip with and without the variable and compares the two bitstreams byte for byte.
 */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "svtme_shim.h"
#include "global_me.h"

static SvtMeShim      *g_shim;
static int             g_state; /* 0: not looked at, 1: CPU ME, 2: GPU ME */
static pthread_mutex_t g_mu = PTHREAD_MUTEX_INITIALIZER;
/* what went through the GPU, written to $SVTME_GPU_STATS at exit (the test checks that the GPU paths were really taken) */
static unsigned long long g_n_me, g_n_tpl_pics, g_n_tpl_blocks, g_n_dg, g_n_tf_blocks;
static void               write_stats(void) {
    const char *fn = getenv("SVTME_GPU_STATS");
    FILE       *f  = fn ? fopen(fn, "w") : NULL;
    if (f) {
        fprintf(f, "{\"me_pictures\": %llu, \"tpl_pictures\": %llu, \"tpl_blocks\": %llu, \"dg_pictures\": %llu, \"tf_blocks\": %llu}\n", g_n_me,
                g_n_tpl_pics, g_n_tpl_blocks, g_n_dg, g_n_tf_blocks);
        fclose(f);
    }
}

/* first call: reads SVTME_GPU_ME and brings the library up.  Returns with g_mu HELD when the GPU path is on. */
static int gpu_me_enabled(SequenceControlSet *scs, PictureParentControlSet *pcs) {
    pthread_mutex_lock(&g_mu);
    if (g_state == 0) {
        const char *v = getenv("SVTME_GPU_ME");
        g_state       = (v && v[0] == '1') ? 2 : 1;
        atexit(write_stats);
        /* the two flags are set by picture analysis (pic_analysis_process.c:1968), so they are valid in every later kernel */
        if (g_state == 2 &&
            svtme_shim_open(scs, 0, 64, pcs->enable_hme_level1_flag || pcs->tf_enable_hme_level1_flag, &g_shim) != 0) {
            fprintf(stderr, "svtme: SVTME_GPU_ME=1 but the CUDA library could not be brought up -- no CPU fallback\n");
            abort();
        }
    }
    if (g_state != 2) {
        pthread_mutex_unlock(&g_mu);
        return 0;
    }
    return 1;
}

/* returns 1 when the picture's ME is (being) done on the GPU: the caller skips its 64x64 block loop */
int svtme_gpu_me_picture(SequenceControlSet *scs, PictureParentControlSet *pcs, MeContext *ctx, EbPictureBufferDesc *input_pic,
                         unsigned segment_index) {
    if (!gpu_me_enabled(scs, pcs)) return 0;
    if (segment_index == 0) { /* one call per picture; the shim's picture cache is not thread-safe: under the lock */
        SvtMeShimGmGate gate;
        if (svtme_shim_me_picture(g_shim, scs, pcs, ctx, input_pic, &gate) != 0) {
            fprintf(stderr, "svtme: %s\n", svtme_shim_last_error(g_shim));
            abort();
        }
        /* the last-SB hook, me_process.c:272-287 (svt_aom_global_motion_estimation re-derives the same level from the arrays) */
        g_n_me++;
        if (gate.run_gm) svt_aom_global_motion_estimation(pcs, input_pic);
        else memset(pcs->is_global_motion, false, MAX_NUM_OF_REF_PIC_LIST * REF_LIST_MAX_DEPTH);
    }
    pthread_mutex_unlock(&g_mu);
    return 1;
}

/* TASK_DG_DETECTOR_HME (hook in front of the block loop of dg_detector_hme_level0, pd_process.c:517): segment 0 searches the
 * whole picture on the GPU and adds the picture's metrics; every segment then falls through to the reference's own
 * seg_completed / frame_done_sem bookkeeping with an empty block range. */
int svtme_gpu_dg_picture(PictureParentControlSet *ppcs, unsigned seg_idx) {
    if (!gpu_me_enabled(ppcs->scs, ppcs)) return 0;
    if (seg_idx == 0) {
        svt_block_on_mutex(ppcs->dg_detector->metrics_mutex);
        if (svtme_shim_dg_picture(g_shim, ppcs) != 0) {
            fprintf(stderr, "svtme: %s\n", svtme_shim_last_error(g_shim));
            abort();
        }
        svt_release_mutex(ppcs->dg_detector->metrics_mutex);
        g_n_dg++;
    }
    pthread_mutex_unlock(&g_mu);
    return 1;
}

/* Temporal-filter ME (hook in front of the svt_aom_motion_estimation_b64 call of produce_temporally_filtered_pic,
 * temporal_filtering.c:3169): 1 = the block's ME outputs are in `ctx`, taken from the whole-picture GPU search of the pair. */
int svtme_gpu_tf_b64(PictureParentControlSet *centre_pcs, PictureParentControlSet *ref_pcs, MeContext *ctx, unsigned b64_index) {
    if (!gpu_me_enabled(centre_pcs->scs, centre_pcs)) return 0;
    if (svtme_shim_tf_b64(g_shim, centre_pcs->scs, centre_pcs, ref_pcs, ctx, b64_index) != 0) {
        fprintf(stderr, "svtme: %s\n", svtme_shim_last_error(g_shim));
        abort();
    }
    g_n_tf_blocks++;
    pthread_mutex_unlock(&g_mu);
    return 1;
}

/* ---- the TPL source-based decisions of one picture from the GPU (hooks in src_ops_process.c, see make_overlay.py) ----
 * tpl_mc_flow_dispenser handles one picture at a time (it blocks on tpl_disp_done_semaphore), so one result set is enough;
 * the per-block hook only reads it (several dispenser threads). */
static uint64_t        g_tpl_poc = ~0ull;
static SvtMeTplResults g_tpl;
static void           *g_tpl_mem;
static size_t          g_tpl_bytes;

void svtme_gpu_tpl_prepare(SequenceControlSet *scs, PictureParentControlSet *pcs) {
    g_tpl_poc = ~0ull;
    if (g_state != 2 || pcs->tpl_src_data_ready) return; /* CPU ME, or the source path of this picture is already known */
    pthread_mutex_lock(&g_mu);
    const size_t need = svtme_shim_tpl_layout(&g_tpl, pcs->enhanced_pic->width, pcs->enhanced_pic->height, NULL);
    if (need > g_tpl_bytes) {
        free(g_tpl_mem);
        g_tpl_mem   = malloc(need);
        g_tpl_bytes = need;
    }
    svtme_shim_tpl_layout(&g_tpl, pcs->enhanced_pic->width, pcs->enhanced_pic->height, g_tpl_mem);
    const int32_t rc = svtme_shim_tpl_picture(g_shim, scs, pcs, &g_tpl);
    if (rc < 0) {
        fprintf(stderr, "svtme: %s\n", svtme_shim_last_error(g_shim));
        abort();
    }
    if (rc == 0) g_n_tpl_pics++;
    if (rc == 0) g_tpl_poc = pcs->picture_number; /* rc == 1: controls the library does not restate -> the CPU search runs */
    pthread_mutex_unlock(&g_mu);
}

/* 1: the decision of the block at (x, y) is filled in and the caller skips its intra + inter source search */
int svtme_gpu_tpl_block(PictureParentControlSet *pcs, unsigned x, unsigned y, unsigned char *best_mode, int *best_rf_idx,
                        void *final_best_mv, unsigned long long *best_ref_poc) {
    if (g_tpl_poc != pcs->picture_number) return 0;
    const uint32_t cell = (y >> 4) * g_tpl.grid_width + (x >> 4);
    if (!g_tpl.block_size[cell]) {
        fprintf(stderr, "svtme: TPL block (%u, %u) of picture %llu has no decision\n", x, y, (unsigned long long)pcs->picture_number);
        abort();
    }
    __sync_fetch_and_add(&g_n_tpl_blocks, 1ull);
    MV *mv        = (MV *)final_best_mv;
    *best_mode    = g_tpl.best_mode[cell];
    *best_rf_idx  = g_tpl.best_rf_idx[cell];
    mv->row       = g_tpl.mv[2 * cell];
    mv->col       = g_tpl.mv[2 * cell + 1];
    *best_ref_poc = g_tpl.ref_frame_poc[cell];
    return 1;
}
