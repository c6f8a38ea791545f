// emul.cc -- TEST INFRASTRUCTURE ONLY.
//
// Compiles the DEVICE control code of the product (svt-av1-mirror_b200/csrc/me_control.cuh) for the host with a
// scalar single-thread backend, so that the phase-ordered control flow can be unit-tested against the oracle on
// machines without a GPU.  The search engines here are deliberately naive loops written for this test; they are
// NOT a CPU fallback of the product: nothing in libsvtme_cuda.so links or calls this file.
#include <cstdlib>
#include <cstring>
#include "../../svt-av1-mirror_b200/csrc/me_control.cuh"
#include "../../oracle/planes.h"

using namespace svtme;

namespace {
struct ScalarBackend {
    bool     leader() const { return true; }
    void     sync() const {}
    uint32_t tid() const { return 0; }
    uint32_t nthreads() const { return 1; }
    void     stage_src(const SbState &) const {}

    static uint32_t sad(const uint8_t *a, int as, const uint8_t *b, int bs, int rows, int w) {
        uint32_t s = 0;
        for (int y = 0; y < rows; y++)
            for (int x = 0; x < w; x++) s += (uint32_t)abs((int)a[y * as + x] - (int)b[y * bs + x]);
        return s;
    }
    void run_nxm(SbState &s) const {
        for (int i = 0; i < s.n_nxm; i++) {
            NxmJob &j = s.nxm_jobs[i];
            j.sad     = sad(j.src, j.src_stride, j.ref, j.ref_stride, j.rows, j.width);
        }
    }
    void run_sad(SbState &s) const {
        for (int i = 0; i < s.n_sad; i++) {
            SadJob &j = s.sad_jobs[i];
            const bool skip = j.skip_line && j.block_w == 16 && j.rows <= 16;
            for (int y = 0; y < j.sa_h; y++) {
                if (skip && (y & 1) == 0) continue;
                for (int x = 0; x < j.sa_w; x++) {
                    uint32_t v = sad(j.blk, j.blk_stride, j.ref + (int64_t)y * j.raw_stride + x, j.ref_stride, j.rows, j.block_w);
                    if (v < j.best_sad) { j.best_sad = v; j.bx = (int16_t)x; j.by = (int16_t)y; j.found = 1; }
                }
            }
        }
    }
    const MeParams *p;
    void run_fullpel(SbState &s) const {
        static const uint8_t z16[16] = {0, 1, 4, 5, 2, 3, 6, 7, 8, 9, 12, 13, 10, 11, 14, 15};
        const bool sub = p->k.me_search_method == 0;
        for (int i = 0; i < s.n_fp; i++) {
            const FpJob &j = s.fp_jobs[i];
            uint32_t *bs = s.best_sad[j.li][j.ri], *bm = s.best_mv[j.li][j.ri];
            for (int y = 0; y < j.h; y++)
                for (int x = 0; x < j.w; x++) {
                    const uint32_t mv = ((uint32_t)(y + j.yo) << 16) | (uint16_t)(x + j.xo);
                    const uint8_t *r  = j.ref + (int64_t)y * j.ref_stride + x;
                    uint32_t s16[16], s64 = 0;
                    for (int b = 0; b < 16; b++) {
                        const int by = b >> 2, bx = b & 3, z = z16[b];
                        uint32_t  t  = 0;
                        for (int q = 0; q < 4; q++) {
                            const uint8_t *sp = s.src + (int64_t)(16 * by + 8 * (q >> 1)) * s.src_stride + 16 * bx + 8 * (q & 1);
                            const uint8_t *rp = r + (int64_t)(16 * by + 8 * (q >> 1)) * j.ref_stride + 16 * bx + 8 * (q & 1);
                            uint32_t v = sub ? sad(sp, 2 * s.src_stride, rp, 2 * j.ref_stride, 4, 8) << 1
                                             : sad(sp, s.src_stride, rp, j.ref_stride, 8, 8);
                            const int pu = 21 + 4 * z + q;
                            if (v < bs[pu]) { bs[pu] = v; bm[pu] = mv; }
                            t += v;
                        }
                        if (t < bs[5 + z]) { bs[5 + z] = t; bm[5 + z] = mv; }
                        s16[z] = t;
                    }
                    for (int q = 0; q < 4; q++) {
                        uint32_t t = s16[4 * q] + s16[4 * q + 1] + s16[4 * q + 2] + s16[4 * q + 3];
                        if (t < bs[1 + q]) { bs[1 + q] = t; bm[1 + q] = mv; }
                        s64 += t;
                    }
                    if (s64 < bs[0]) { bs[0] = s64; bm[0] = mv; }
                }
        }
    }
};

DevPic to_dev(const SvtRefPic &r) {
    DevPic d;
    for (int l = 0; l < 3; l++) {
        const SvtRefPlane &pl = r.lvl[l];
        d.lvl[l].org     = pl.buf + (size_t)pl.org_y * pl.stride + pl.org_x;
        d.lvl[l].stride  = (int32_t)pl.stride;
        d.lvl[l].width   = (int32_t)pl.width;
        d.lvl[l].height  = (int32_t)pl.height;
        d.lvl[l].ref_org = (int32_t)pl.org_x;
    }
    d.picture_number = r.picture_number;
    return d;
}
}  // namespace

extern "C" int svtemu_me_picture(const SvtMeKnobs *k, const SvtRefPic *cur, const SvtRefPic *refs,
                                 SvtMePictureResults *out, int n_threads) {
    (void)n_threads;
    MeParams p;
    memset(&p, 0, sizeof(p));
    p.k   = *k;
    p.cur = to_dev(*cur);
    for (int li = 0; li < k->num_of_list_to_search; li++)
        for (int ri = 0; ri < k->num_of_ref_pic_to_search[li]; ri++) p.refs[li][ri] = to_dev(refs[li * 4 + ri]);
    p.wb = (cur->lvl[0].width + 63) / 64;
    p.hb = (cur->lvl[0].height + 63) / 64;
    if (out->b64_total_count < p.wb * p.hb) return -2;
    p.pu_count = out->pu_count;
    p.out.me_mv = out->me_mv; p.out.me_cand = out->me_cand; p.out.total_me_candidate_index = out->total_me_candidate_index;
    p.out.me_64x64_distortion = out->me_64x64_distortion; p.out.me_32x32_distortion = out->me_32x32_distortion;
    p.out.me_16x16_distortion = out->me_16x16_distortion; p.out.me_8x8_distortion = out->me_8x8_distortion;
    p.out.me_8x8_cost_variance = out->me_8x8_cost_variance; p.out.rc_me_distortion = out->rc_me_distortion;
    p.out.stationary_block_present_sb = out->stationary_block_present_sb; p.out.rc_me_allow_gm = out->rc_me_allow_gm;
    p.out.sb_best_sad = out->sb_best_sad; p.out.sb_best_mv = out->sb_best_mv; p.out.do_ref = out->do_ref; p.out.hme_sc = out->hme_sc;
    p.out.tf_exit_sb = out->tf_exit_sb; p.out.tf_blk_counts = out->tf_blk_counts;
    p.out.tf_best_sad = out->tf_best_sad; p.out.tf_best_mv = out->tf_best_mv;
    if (out->tf_blk_counts) out->tf_blk_counts[0] = out->tf_blk_counts[1] = 0;
    p.out.picture_sums = k->me_type == SVTME_ME_MCTF ? nullptr : (unsigned long long *)out->picture_sums;
    if (p.out.picture_sums) memset(p.out.picture_sums, 0, 32);
    p.n_steps = build_program(p.k, p.prog);
    ScalarBackend be;
    be.p = &p;
    SbState *s = (SbState *)malloc(sizeof(SbState));
    for (uint32_t sb = 0; sb < p.wb * p.hb; sb++) me_sb(be, p, *s, sb);
    free(s);
    return 0;
}
