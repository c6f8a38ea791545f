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

/* returns 1 when the picture's ME is (being) done on the GPU: the caller skips its 64x64 block loop */
int svtme_gpu_me_picture(SequenceControlSet *scs, PictureParentControlSet *pcs, MeContext *ctx, EbPictureBufferDesc *input_pic,
                         unsigned segment_index) {
    pthread_mutex_lock(&g_mu);
    if (g_state == 0) {
        const char *v = getenv("SVTME_GPU_ME");
        g_state       = (v && v[0] == '1') ? 2 : 1;
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
    if (segment_index == 0) { /* one call per picture; the shim's picture cache is not thread-safe: under the lock */
        SvtMeShimGmGate gate;
        if (svtme_shim_me_picture(g_shim, scs, pcs, ctx, input_pic, &gate) != 0) {
            fprintf(stderr, "svtme: %s\n", svtme_shim_last_error(g_shim));
            abort();
        }
        /* the last-SB hook, me_process.c:272-287 (svt_aom_global_motion_estimation re-derives the same level from the arrays) */
        if (gate.run_gm) svt_aom_global_motion_estimation(pcs, input_pic);
        else memset(pcs->is_global_motion, false, MAX_NUM_OF_REF_PIC_LIST * REF_LIST_MAX_DEPTH);
    }
    pthread_mutex_unlock(&g_mu);
    return 1;
}
