// me_control.cuh -- per-SB control flow of the open-loop ME path, written once for the device.
//
// The reference runs one 64x64 block at a time on one CPU thread and interleaves scalar decisions
// with SAD searches (svt_aom_motion_estimation_b64, motion_estimation.c:3076-3153).  Here one CTA
// owns one 64x64 block and the same decisions are organised in PHASES: a leader thread turns the
// state of the block into a list of independent search jobs (SadJob / NxmJob / FpJob), the whole
// CTA executes the list, the leader folds the results back.  A phase boundary is placed wherever
// the reference has a data dependence between references:
//   * pre-HME list-1 early exit reads the list-0 result of the same index   (:1706-1717)
//   * HME-L0 area of refs other than (0,0) reads the (0,0) level-0 centre    (:1810-1825, rtc only)
//   * full-pel SR halving reads p_sb_best_sad[0][0][0]                        (:1355-1364, sr adj 2)
// Everything else inside a phase is independent and runs concurrently.
//
// The code is templated on a Backend that provides leader()/sync()/tid()/nthreads() and the three
// collective engines run_nxm / run_sad / run_fullpel; the CUDA backend is in me_kernel.cu.  (A
// scalar backend used only by the CPU unit tests of this control code lives under tests/host_emul.)
//
// All integer widths, truncations and unsigned wrap-arounds follow the reference literally; the
// file:line next to each block is the reference statement it implements.
#pragma once
#include "me_types.cuh"

namespace svtme {

#define SVTME_ABS(x) ((x) < 0 ? -(x) : (x))
#define SVTME_MIN(a, b) ((a) < (b) ? (a) : (b))
#define SVTME_MAX(a, b) ((a) > (b) ? (a) : (b))

// ---- lanes ----
// A control step is executed by ONE WARP: lane l = (list << 4) | (reference << 2) | quadrant owns reference (li, ri)
// and one of its (up to) four search jobs -- quadrant (qw, qh) of an HME level, or pre-HME region sr = q < 2.  Lanes
// exchange data only through the block state (shared memory on the device) between SVTME_WSYNC() points, so the same
// code runs on the host as 32 sequential iterations (tests/host_emul).  Search jobs live in FIXED slots: SadJob slot =
// lane, NxmJob slot = 2 * (li * 4 + ri) + {0, 1}, FpJob slot = li * 4 + ri; an unused slot has an empty window.
#if defined(__CUDA_ARCH__)
#define SVTME_LANES(l) for (int l = (int)(threadIdx.x & 31u), l##_it = 0; l##_it < 1; ++l##_it)
#define SVTME_WSYNC() __syncwarp()
#elif defined(SVTME_LANES_REVERSE)  // host test build: the lanes of a region must not depend on each other's order
#define SVTME_LANES(l) for (int l = 31; l >= 0; --l)
#define SVTME_WSYNC() ((void)0)
#else
#define SVTME_LANES(l) for (int l = 0; l < 32; ++l)
#define SVTME_WSYNC() ((void)0)
#endif

struct LaneRef {
    int  li, ri, q, qw, qh;
    bool used;  // (li, ri) is a reference of this picture
};
SVTME_HD LaneRef lane_ref(const SvtMeKnobs &k, int l) {
    LaneRef r;
    r.li = l >> 4; r.ri = (l >> 2) & 3; r.q = l & 3; r.qw = r.q & 1; r.qh = r.q >> 1;
    r.used = r.li < k.num_of_list_to_search && r.ri < k.num_of_ref_pic_to_search[r.li];
    return r;
}
constexpr uint8_t kNoJob = 0xff;  // SadJob::level of an unused slot
SVTME_HD void clear_sad_slot(SbState &s, int slot) {
    SadJob &j = s.sad_jobs[slot];
    j.level = kNoJob; j.sa_w = 0; j.sa_h = 0; j.found = 0;
}
SVTME_HD int nxm_slot(int li, int ri, int second) { return 2 * (li * 4 + ri) + second; }

SVTME_HD bool hme_active(const SvtMeKnobs &k, int li) { return k.temporal_layer_index > 0 || li == 0; }

// get_me_reference's distance (:1232-1234) and svt_aom_get_scaled_picture_distance (:1239-1243)
SVTME_HD uint16_t ref_dist(const MeParams &p, int li, int ri) {
    int64_t d = (int64_t)p.k.picture_number - (int64_t)p.refs[li][ri].picture_number;
    if (d < 0) d = -d;
    return (uint16_t)(int16_t)d;
}
SVTME_HD uint16_t scaled_dist(uint16_t dist) { return (uint16_t)(((dist * 5) / 8) + ((dist % 8) == 0 ? 0 : 1)); }

// Window clipping shared by hme_level_0/1/2 (:847-880, 948-982, 1049-1078) and prehme_core (:1584-1629).
// The left/top correction only moves the origin (the size correction there is a no-op in the reference because
// the origin has already been overwritten); right/bottom corrections shrink with MAX(1, .).
SVTME_HD void clip_window(int16_t org_x, int16_t org_y, int16_t pad_w, int16_t pad_h, int16_t pic_w, int16_t pic_h,
                          int16_t &ox, int16_t &oy, int16_t &w, int16_t &h, bool round8) {
    if ((org_x + ox) < -pad_w) ox = (int16_t)(-pad_w - org_x);
    if ((org_x + ox) > pic_w - 1) ox = (int16_t)(ox - ((org_x + ox) - (pic_w - 1)));
    if ((org_x + ox + w) > pic_w) w = (int16_t)SVTME_MAX(1, w - ((org_x + ox + w) - pic_w));
    if (round8) w = (int16_t)((w < 8) ? w : (w & ~0x07));
    if ((org_y + oy) < -pad_h) oy = (int16_t)(-pad_h - org_y);
    if ((org_y + oy) > pic_h - 1) oy = (int16_t)(oy - ((org_y + oy) - (pic_h - 1)));
    if ((org_y + oy + h) > pic_h) h = (int16_t)SVTME_MAX(1, h - ((org_y + oy + h) - pic_h));
}

// Source block of this SB on plane `lvl` (me_process.c:183-214)
SVTME_HD const uint8_t *cur_block(const MeParams &p, const SbState &s, int lvl) {
    const DevPlane &pl = p.cur.lvl[lvl];
    return pl.org + (int64_t)(s.org_y >> lvl) * pl.stride + (s.org_x >> lvl);
}

// Fill SadJob `slot` for an HME search at `lvl` (0 full, 1 quarter, 2 sixteenth) of reference (li,ri):
// the tail of hme_level_0/1/2 / prehme_core (:882-909, 984-1011, 1080-1104, 1631-1655).
SVTME_HD void make_sad_job(const MeParams &p, SbState &s, int slot, int lvl, int li, int ri, int16_t ox, int16_t oy, int16_t w,
                           int16_t h, uint8_t skip, uint8_t dst_level, uint8_t qw, uint8_t qh) {
    const DevPlane &rp   = p.refs[li][ri].lvl[lvl];
    const DevPlane &cp   = p.cur.lvl[lvl];
    const bool      full = p.k.hme_search_method == 1;  // FULL_SAD_SEARCH
    SadJob         &j    = s.sad_jobs[slot];
    const int16_t org_x = (int16_t)(((int16_t)s.org_x) >> lvl), org_y = (int16_t)(((int16_t)s.org_y) >> lvl);
    j.blk        = cur_block(p, s, lvl);
    j.blk_stride = full ? cp.stride : cp.stride * 2;
    // x_top_left = (org + block) + origin as int16; index = x + y*stride (:883-888)
    const int16_t x_tl = (int16_t)(((int16_t)rp.ref_org + org_x) + ox), y_tl = (int16_t)(((int16_t)rp.ref_org + org_y) + oy);
    j.ref        = rp.org + (int64_t)(y_tl - rp.ref_org) * rp.stride + (x_tl - rp.ref_org);
    j.ref_stride = full ? rp.stride : rp.stride * 2;
    j.raw_stride = rp.stride;
    const uint32_t bh = s.b64_h >> lvl;
    j.rows    = (int16_t)(full ? bh : bh >> 1);
    j.block_w = (int16_t)(s.b64_w >> lvl);
    j.sa_w = w; j.sa_h = h; j.sa_ox = ox; j.sa_oy = oy;
    j.skip_line = skip;
    j.mv_scale  = (uint8_t)(1 << lvl);
    j.level = dst_level; j.li = (uint8_t)li; j.ri = (uint8_t)ri; j.qw = qw; j.qh = qh;
    j.found = 0; j.best_sad = 0xffffff; j.bx = 0; j.by = 0;
}

// Fold one finished SadJob back: *best_sad *= 2 in sub-SAD mode, MV = (argmin + origin) * scale in int16
// (:911-917, 1013-1019, 1106-1110, 1657-1664).  When nothing was found (empty window) the reference leaves the
// previous centre in place; here that previous value is the SB-initial 0.
SVTME_HD void fold_sad_job(const MeParams &p, SbState &s, const SadJob &j) {
    const bool full = p.k.hme_search_method == 1;
    uint64_t   sad  = full ? (uint64_t)j.best_sad : (uint64_t)j.best_sad * 2;
    if (j.level < 3) {
        int16_t &x = s.lx[j.level][j.li][j.ri][j.qw][j.qh], &y = s.ly[j.level][j.li][j.ri][j.qw][j.qh];
        if (j.found) { x = j.bx; y = j.by; }
        x = (int16_t)(x + j.sa_ox); x = (int16_t)(x * j.mv_scale);
        y = (int16_t)(y + j.sa_oy); y = (int16_t)(y * j.mv_scale);
        s.lsad[j.level][j.li][j.ri][j.qw][j.qh] = sad;
    } else {
        PreHme &q = s.prehme[j.li][j.ri][j.qw];
        if (j.found) { q.mv_x = j.bx; q.mv_y = j.by; }
        q.mv_x = (int16_t)(q.mv_x + j.sa_ox); q.mv_x = (int16_t)(q.mv_x * j.mv_scale);
        q.mv_y = (int16_t)(q.mv_y + j.sa_oy); q.mv_y = (int16_t)(q.mv_y * j.mv_scale);
        q.sad   = sad;
        q.valid = 1;
    }
}

// set_quadrants of the reference code paths, one quadrant per lane
SVTME_HD void set_quadrant(const SvtMeKnobs &k, SbState &s, int lvl, const LaneRef &r, int16_t x, int16_t y, uint64_t sad) {
    if (r.qh < k.num_hme_sa_h && r.qw < k.num_hme_sa_w) {
        s.lx[lvl][r.li][r.ri][r.qw][r.qh] = x; s.ly[lvl][r.li][r.ri][r.qw][r.qh] = y; s.lsad[lvl][r.li][r.ri][r.qw][r.qh] = sad;
    }
}

// init_me_hme_data (:3010-3070) + the block-size derivation of svt_aom_motion_estimation_b64 (:3093-3100): the CONTROL
// part of the state (`s`, zeroed by the caller) -- the result tables are initialised by sb_init_tables
SVTME_HD void sb_init_head(const MeParams &p, SbState &s, SbState *home, uint32_t b64_index) {
    const DevPlane &f  = p.cur.lvl[0];
    const uint16_t  aw = (uint16_t)((f.width + 7) & ~7), ah = (uint16_t)((f.height + 7) & ~7);
    s.home      = home;
    s.b64_index = b64_index;
    s.org_x     = (b64_index % p.wb) * 64;
    s.org_y     = (b64_index / p.wb) * 64;
    s.b64_w     = (aw - s.org_x) < 64 ? aw - s.org_x : 64;
    s.b64_h     = (ah - s.org_y) < 64 ? ah - s.org_y : 64;
    s.src        = f.org + (int64_t)s.org_y * f.stride + s.org_x;
    s.src_stride = f.stride;
    s.hme_l0_sa  = p.k.hme_l0_sa;
    for (int li = 0; li < kMaxLists; li++)
        for (int ri = 0; ri < kMaxRefs; ri++) {
            s.do_ref[li][ri] = 1; s.hme_sad[li][ri] = kMaxU32; s.divisor[li][ri] = 1; s.zz_sad[li][ri] = kMaxU32;
        }
    for (int i = 0; i < kMaxSadJobs; i++) clear_sad_slot(s, i);
    s.n_sad = kMaxSadJobs; s.n_nxm = kMaxNxmJobs; s.n_fp = kMaxFpJobs;  // fixed slots; rows / w == 0 marks the unused ones
}

// host emulation: zero the whole SB state, then the control part and the tables
template <class B>
SVTME_HD void sb_init(B &be, const MeParams &p, SbState &s, uint32_t b64_index) {
    uint32_t *w = reinterpret_cast<uint32_t *>(&s);
    for (uint32_t i = be.tid(); i < sizeof(SbState) / 4; i += be.nthreads()) w[i] = 0;
    be.sync();
    if (be.leader()) sb_init_head(p, s, &s, b64_index);
    // unsearched references report MAX_SAD_VALUE (convention of include/svtme_cuda.h)
    uint32_t *bs = &s.best_sad[0][0][0];
    be.sync();
    for (uint32_t i = be.tid(); i < kMaxLists * kMaxRefs * kPus; i += be.nthreads()) bs[i] = SVTME_MAX_SAD_VALUE;
    be.sync();
}

#define SVTME_FOR_REFS(k, li, ri)                                \
    for (int li = 0; li < (k).num_of_list_to_search; ++li)       \
        for (int ri = 0; ri < (k).num_of_ref_pic_to_search[li]; ++ri)

// Zero-MV (or any fixed-MV) 64x32-subsampled SAD job against the full-resolution plane (get_zz_sad :1667-1689,
// check_00_center :1153-1163, 1189-1197).
SVTME_HD void make_nxm_job(const MeParams &p, SbState &s, int slot, int li, int ri, int16_t mvx, int16_t mvy, uint8_t kind) {
    const DevPlane &rp = p.refs[li][ri].lvl[0];
    NxmJob         &j  = s.nxm_jobs[slot];
    j.src        = s.src;
    j.src_stride = s.src_stride << 1;
    j.ref        = rp.org + (int64_t)((int16_t)s.org_y + mvy) * rp.stride + ((int16_t)s.org_x + mvx);
    j.ref_stride = rp.stride << 1;
    j.rows       = (int16_t)(s.b64_h >> 1);
    j.width      = (int16_t)s.b64_w;
    j.li = (uint8_t)li; j.ri = (uint8_t)ri; j.kind = kind; j.sad = 0;
}
SVTME_HD void clear_nxm_slots(SbState &s) {
    SVTME_LANES(l) if (l < kMaxNxmJobs) { s.nxm_jobs[l].rows = 0; s.nxm_jobs[l].sad = 0; }
    SVTME_WSYNC();
}
SVTME_HD void clear_fp_slots(SbState &s) {
    SVTME_LANES(l) if (l < kMaxFpJobs) { s.fp_jobs[l].w = 0; s.fp_jobs[l].h = 0; }
    SVTME_WSYNC();
}

// ---- init_zz_sad (:2382-2437): make the zero-MV SAD jobs / fold them and prune ----
SVTME_HD void zz_make(const MeParams &p, SbState &s) {
    const SvtMeKnobs &k = p.k;
    clear_nxm_slots(s);
    SVTME_LANES(l) {
        const LaneRef r = lane_ref(k, l);
        if (r.used && r.q == 0 && hme_active(k, r.li)) make_nxm_job(p, s, nxm_slot(r.li, r.ri, 0), r.li, r.ri, 0, 0, 0);
    }
    SVTME_WSYNC();
}
SVTME_HD void zz_fold(const MeParams &p, SbState &s) {
    const SvtMeKnobs &k = p.k;
    SVTME_LANES(l) {
        const LaneRef r = lane_ref(k, l);
        if (r.q != 0) continue;
        const NxmJob &j = s.nxm_jobs[nxm_slot(r.li, r.ri, 0)];
        if (j.rows <= 0) continue;
        uint32_t zz = j.sad << 1;
        zz = (zz * 64 * 64) / (s.b64_w * s.b64_h);  // normalise incomplete b64 (:2395)
        s.zz_sad[r.li][r.ri] = zz;
    }
    SVTME_WSYNC();
    SVTME_LANES(l) {
        const LaneRef r = lane_ref(k, l);
        if (!r.used || r.q != 0) continue;
        uint32_t best = kMaxU32;  // over the references whose zero-MV SAD was computed
        for (int i = 0; i < kMaxLists * kMaxRefs; i++)
            if (s.nxm_jobs[2 * i].rows > 0) best = SVTME_MIN(best, s.zz_sad[i >> 2][i & 3]);
        if (k.temporal_layer_index > 0 && best < k.zz_sad_th && r.ri != 0) {
            if ((s.zz_sad[r.li][r.ri] - best) * 100 > ((uint32_t)k.zz_sad_pct * best)) s.do_ref[r.li][r.ri] = 0;  // u32 wrap (:2409)
        }
        if (k.me_safe_limit_zz_th) {
            const bool safe = k.hierarchical_levels > 0 && k.num_of_list_to_search == 2 &&
                k.temporal_layer_index >= k.hierarchical_levels && k.similar_brightness_refs &&
                s.zz_sad[0][0] < k.me_safe_limit_zz_th && s.zz_sad[1][0] < k.me_safe_limit_zz_th;
            if (safe && r.ri > 0) s.do_ref[r.li][r.ri] = 0;
        }
    }
    SVTME_WSYNC();
}

// ---- prehme_b64 (:1722-1796) with check_prehme_early_exit (:1693-1719) and prehme_core (:1568-1666) ----
// Returns true when the early exit applies (the region's result is then already final).
SVTME_HD bool prehme_early_exit(const SvtMeKnobs &k, SbState &s, int li, int ri, int sr) {
    PreHme &q = s.prehme[li][ri][sr];
    if (k.me_early_exit_th && s.zz_sad[li][ri] < k.me_early_exit_th) {
        q.mv_x = 0; q.mv_y = 0; q.sad = 0; q.valid = 1;
        return true;
    }
    if (k.prehme_l1_early_exit) {
        const PreHme &a = s.prehme[0][ri][sr];
        if (li == 1 && a.valid && ((a.sad < (32 * 32)) || ((SVTME_ABS(a.mv_x) < 16) && (SVTME_ABS(a.mv_y) < 16)))) {
            q.mv_x = (int16_t)-a.mv_x; q.mv_y = (int16_t)-a.mv_y; q.sad = a.sad; q.valid = 1;
            return true;
        }
    }
    return false;
}

// `arg`: 0 = list 0 only, 1 = list 1 only (its early exit reads the list-0 results, :1706-1717), 2 = both lists.
// List 1 of a base-layer picture mirrors list 0 (:1769-1777) and has no searches of its own.  Lane = (li, ri, region).
SVTME_HD void prehme_make(const MeParams &p, SbState &s, int arg) {
    const SvtMeKnobs &k = p.k;
    const bool l1_active = k.num_of_list_to_search > 1 && hme_active(k, 1);
    SVTME_LANES(l) {
        clear_sad_slot(s, l);
        const LaneRef r = lane_ref(k, l);
        if (!r.used || r.q >= 2) continue;
        if (r.li == 0 ? !(arg == 0 || arg == 2) : !(l1_active && (arg == 1 || arg == 2))) continue;
        const int li = r.li, ri = r.ri, sr = r.q;
        const uint32_t f = scaled_dist(ref_dist(p, li, ri));
        if (prehme_early_exit(k, s, li, ri, sr)) continue;
        PreHme &q = s.prehme[li][ri][sr];
        if (!s.do_ref[li][ri]) { q.mv_x = 0; q.mv_y = 0; q.sad = kMaxU32; continue; }
        q.sa_w = (uint16_t)SVTME_MIN((k.prehme_sa_cfg[sr].sa_min.width * f), (uint32_t)k.prehme_sa_cfg[sr].sa_max.width);
        q.sa_h = (uint16_t)SVTME_MIN((k.prehme_sa_cfg[sr].sa_min.height * f), (uint32_t)k.prehme_sa_cfg[sr].sa_max.height);
        const DevPlane &rp = p.refs[li][ri].lvl[2];
        int16_t w = (int16_t)q.sa_w, h = (int16_t)q.sa_h;
        int16_t ox = (int16_t)(-(int16_t)(w >> 1)), oy = (int16_t)(-(int16_t)(h >> 1));
        clip_window((int16_t)(((int16_t)s.org_x) >> 2), (int16_t)(((int16_t)s.org_y) >> 2), (int16_t)(rp.ref_org - 1),
                    (int16_t)(rp.ref_org - 1), (int16_t)rp.width, (int16_t)rp.height, ox, oy, w, h, false);
        make_sad_job(p, s, l, 2, li, ri, ox, oy, w, h, k.prehme_skip_search_line, 3, (uint8_t)sr, 0);
        s.performed_phme[li][ri][sr] = 1;
    }
    SVTME_WSYNC();
}
SVTME_HD void sad_fold(const MeParams &p, SbState &s) {
    SVTME_LANES(l) if (s.sad_jobs[l].level != kNoJob) fold_sad_job(p, s, s.sad_jobs[l]);
    SVTME_WSYNC();
}
SVTME_HD void prehme_final(const MeParams &p, SbState &s) {
    const SvtMeKnobs &k = p.k;
    SVTME_LANES(l) {  // a list without HME of its own mirrors list 0 (:1769-1777)
        const LaneRef r = lane_ref(k, l);
        if (r.used && r.q < 2 && !hme_active(k, r.li)) {
            s.prehme[1][r.ri][r.q].mv_x = (int16_t)-s.prehme[0][r.ri][r.q].mv_x;
            s.prehme[1][r.ri][r.q].mv_y = (int16_t)-s.prehme[0][r.ri][r.q].mv_y;
            s.prehme[1][r.ri][r.q].sad  = s.prehme[0][r.ri][r.q].sad;
        }
    }
    SVTME_WSYNC();
    SVTME_LANES(l) {
        const LaneRef r = lane_ref(k, l);
        if (!r.used || r.q != 0 || r.ri == 0 || !s.do_ref[r.li][r.ri]) continue;
        uint32_t best_sad = kMaxU32;
        SVTME_FOR_REFS(k, li, ri) if (hme_active(k, li)) {
            const uint32_t m = (uint32_t)SVTME_MIN(s.prehme[li][ri][0].sad, s.prehme[li][ri][1].sad);
            best_sad         = SVTME_MIN(best_sad, m);
        }
        if (k.temporal_layer_index > 0 && best_sad < k.phme_sad_th) {
            const uint32_t m = (uint32_t)SVTME_MIN(s.prehme[r.li][r.ri][0].sad, s.prehme[r.li][r.ri][1].sad);
            if ((m - best_sad) * 100 > ((uint32_t)k.phme_sad_pct * best_sad)) s.do_ref[r.li][r.ri] = 0;  // u32 wrap (:1790)
        }
    }
    SVTME_WSYNC();
}

// ---- get_hme_l0_search_area (:1800-1867); works on a private copy of hme_l0_sa, which is what the reference's
// mutate-then-restore (:1843-1850, 1999-2003) amounts to ----
SVTME_HD void hme_l0_area(const MeParams &p, const SbState &s, int li, int ri, int16_t &sa_w, int16_t &sa_h) {
    const SvtMeKnobs &k  = p.k;
    SvtMeAreaMinMax   sa = s.hme_l0_sa;
    if (k.enable_me_sr_adjustment && k.distance_based_hme_resizing) {
        uint8_t is_hor = 1, is_ver = 1, is_still = 0;
        if (k.reduce_hme_l0_sr_th_min && k.reduce_hme_l0_sr_th_max && (li || ri)) {
            const int16_t mx = s.lx[0][0][0][0][0], my = s.ly[0][0][0][0][0];
            is_ver   = (SVTME_ABS(mx) < k.reduce_hme_l0_sr_th_min) && (SVTME_ABS(my) > k.reduce_hme_l0_sr_th_max);
            is_hor   = (SVTME_ABS(mx) > k.reduce_hme_l0_sr_th_max) && (SVTME_ABS(my) < k.reduce_hme_l0_sr_th_min);
            is_still = (SVTME_ABS(mx) < (k.reduce_hme_l0_sr_th_min * 3)) && (SVTME_ABS(my) < (k.reduce_hme_l0_sr_th_min * 3));
        }
        uint8_t xo = 1, yo = 1;
        if (!is_ver) yo = 2;
        if (!is_hor) xo = 2;
        if (k.enable_me_sr_adjustment == 2 && is_still) { xo = 4; yo = 4; }
        sa.sa_min.width  = (uint16_t)(sa.sa_min.width / (xo + ri));
        sa.sa_min.height = (uint16_t)(sa.sa_min.height / (yo + ri));
        sa.sa_max.width  = (uint16_t)(sa.sa_max.width / (xo + ri));
        sa.sa_max.height = (uint16_t)(sa.sa_max.height / (yo + ri));
    }
    const int32_t f = scaled_dist(ref_dist(p, li, ri));
    int16_t w = (int16_t)(sa.sa_min.width / k.num_hme_sa_w);
    w         = (int16_t)SVTME_MIN((((w * f) + 15) & ~0x0F), (((sa.sa_max.width / k.num_hme_sa_w) + 15) & ~0x0F));
    int16_t h = (int16_t)(sa.sa_min.height / k.num_hme_sa_h);
    h         = (int16_t)SVTME_MIN((h * f), sa.sa_max.height / k.num_hme_sa_h);
    sa_w = w; sa_h = h;
}

// One quadrant of one reference's share of hme_level0_b64 (:1921-2033) up to the job creation.  Returns true if the
// reference's quadrants are searched.
SVTME_HD bool hme_l0_make_job(const MeParams &p, SbState &s, int slot, const LaneRef &r) {
    const SvtMeKnobs &k = p.k;
    const int li = r.li, ri = r.ri;
    if (k.me_early_exit_th && s.zz_sad[li][ri] < (k.me_early_exit_th >> 2)) { set_quadrant(k, s, 0, r, 0, 0, 0); return false; }
    if (k.prev_me_stage_based_exit_th) {
        const int sr = s.prehme[li][ri][0].sad <= s.prehme[li][ri][1].sad ? 0 : 1;
        if (s.performed_phme[li][ri][sr] && s.prehme[li][ri][sr].sad < (k.prev_me_stage_based_exit_th >> 4)) {
            set_quadrant(k, s, 0, r, s.prehme[li][ri][sr].mv_x, s.prehme[li][ri][sr].mv_y, s.prehme[li][ri][sr].sad);
            return false;
        }
    }
    if (!s.do_ref[li][ri]) { set_quadrant(k, s, 0, r, 0, 0, kMaxU32); return false; }
    if (!hme_active(k, li)) return false;
    if (r.qh >= k.num_hme_sa_h || r.qw >= k.num_hme_sa_w) return true;
    int16_t sa_w = 0, sa_h = 0;
    hme_l0_area(p, s, li, ri, sa_w, sa_h);
    const DevPlane &rp = p.refs[li][ri].lvl[2];
    {   // hme_level_0 (:820-920)
        const int qw = r.qw, qh = r.qh;
        int16_t w = (int16_t)((sa_w + 7) & ~0x07), h = sa_h;
        const int16_t xd = (int16_t)(w * qw), yd = (int16_t)(h * qh);
        int16_t ox = (int16_t)(-(int16_t)((w * k.num_hme_sa_w) >> 1) + xd);
        int16_t oy = (int16_t)(-(int16_t)((h * k.num_hme_sa_h) >> 1) + yd);
        clip_window((int16_t)(((int16_t)s.org_x) >> 2), (int16_t)(((int16_t)s.org_y) >> 2), (int16_t)(rp.ref_org - 1),
                    (int16_t)(rp.ref_org - 1), (int16_t)rp.width, (int16_t)rp.height, ox, oy, w, h, true);
        make_sad_job(p, s, slot, 2, li, ri, ox, oy, w, h, 0, 0, (uint8_t)qw, (uint8_t)qh);
    }
    return true;
}

// Tail of the per-reference loop of hme_level0_b64: replace the worst quadrant by the better pre-HME result
// (:2005-2032) with get_worst_quadrant (:1872-1901; the running max is not updated in its last branch).
SVTME_HD void hme_l0_replace_worst(const SvtMeKnobs &k, SbState &s, int li, int ri) {
    if (!k.prehme_enable) return;
    uint8_t bw = 0, bh = 0;
    if (k.num_hme_sa_w == 2 && k.num_hme_sa_h == 2) {
        uint64_t m = 0;
        if (s.lsad[0][li][ri][0][0] > m) { m = s.lsad[0][li][ri][0][0]; bw = 0; bh = 0; }
        if (s.lsad[0][li][ri][1][0] > m) { m = s.lsad[0][li][ri][1][0]; bw = 1; bh = 0; }
        if (s.lsad[0][li][ri][0][1] > m) { m = s.lsad[0][li][ri][0][1]; bw = 0; bh = 1; }
        if (s.lsad[0][li][ri][1][1] > m) { bw = 1; bh = 1; }
    }
    const int sr = s.prehme[li][ri][0].sad <= s.prehme[li][ri][1].sad ? 0 : 1;
    if (s.prehme[li][ri][sr].sad < s.lsad[0][li][ri][bw][bh]) {
        s.lsad[0][li][ri][bw][bh] = s.prehme[li][ri][sr].sad;
        s.lx[0][li][ri][bw][bh]   = s.prehme[li][ri][sr].mv_x;
        s.ly[0][li][ri][bw][bh]   = s.prehme[li][ri][sr].mv_y;
    }
}

// `arg`: 0 = reference (0,0) only, 1 = all others (they read the (0,0) level-0 centre, :1804-1825), 2 = all.
SVTME_HD void l0_make(const MeParams &p, SbState &s, int arg) {
    const SvtMeKnobs &k = p.k;
    SVTME_LANES(l) {
        clear_sad_slot(s, l);
        const LaneRef r = lane_ref(k, l);
        bool searched = false;
        const bool first = (r.li == 0 && r.ri == 0);
        if (r.used && (arg == 2 || (first == (arg == 0)))) searched = hme_l0_make_job(p, s, l, r);
        if (r.q == 0) s.l0_done[r.li][r.ri] = searched ? 1 : 0;  // references whose quadrants are searched in THIS pass
    }
    SVTME_WSYNC();
}
SVTME_HD void l0_fold(const MeParams &p, SbState &s) {
    const SvtMeKnobs &k = p.k;
    sad_fold(p, s);
    SVTME_LANES(l) {
        const LaneRef r = lane_ref(k, l);
        if (r.used && r.q == 0 && s.l0_done[r.li][r.ri]) hme_l0_replace_worst(k, s, r.li, r.ri);
    }
    SVTME_WSYNC();
}

// ---- hme_level1_b64 (:2041-2122) + hme_level_1 (:923-1022) ----
SVTME_HD void l1_make(const MeParams &p, SbState &s) {
    const SvtMeKnobs &k = p.k;
    SVTME_LANES(l) {
        clear_sad_slot(s, l);
        const LaneRef r = lane_ref(k, l);
        const int li = r.li, ri = r.ri, qw = r.qw, qh = r.qh;
        if (!r.used || !hme_active(k, li)) continue;
        if (k.me_early_exit_th && s.zz_sad[li][ri] < (k.me_early_exit_th >> 2)) { set_quadrant(k, s, 1, r, 0, 0, 0); continue; }
        if (!s.do_ref[li][ri]) { set_quadrant(k, s, 1, r, 0, 0, kMaxU32); continue; }
        if (qh >= k.num_hme_sa_h || qw >= k.num_hme_sa_w) continue;
        const DevPlane &rp = p.refs[li][ri].lvl[1];
        if (k.prev_me_stage_based_exit_th && s.lsad[0][li][ri][qw][qh] < (k.prev_me_stage_based_exit_th >> 5)) {
            s.lx[1][li][ri][qw][qh] = s.lx[0][li][ri][qw][qh]; s.ly[1][li][ri][qw][qh] = s.ly[0][li][ri][qw][qh];
            s.lsad[1][li][ri][qw][qh] = s.lsad[0][li][ri][qw][qh];
            continue;
        }
        int16_t w = (int16_t)(((int16_t)k.hme_l1_sa.width + 7) & ~0x07), h = (int16_t)k.hme_l1_sa.height;
        const int16_t cx = (int16_t)(s.lx[0][li][ri][qw][qh] >> 1), cy = (int16_t)(s.ly[0][li][ri][qw][qh] >> 1);
        int16_t ox = (int16_t)(-(w >> 1) + cx), oy = (int16_t)(-(h >> 1) + cy);
        clip_window((int16_t)(((int16_t)s.org_x) >> 1), (int16_t)(((int16_t)s.org_y) >> 1), (int16_t)(rp.ref_org - 1),
                    (int16_t)(rp.ref_org - 1), (int16_t)rp.width, (int16_t)rp.height, ox, oy, w, h, true);
        make_sad_job(p, s, l, 1, li, ri, ox, oy, w, h, 0, 1, (uint8_t)qw, (uint8_t)qh);
    }
    SVTME_WSYNC();
}

// ---- hme_level2_b64 (:2127-2177) + hme_level_2 (:1025-1113; pad is BLOCK_SIZE_64-1 regardless of the buffer) ----
SVTME_HD void l2_make(const MeParams &p, SbState &s) {
    const SvtMeKnobs &k = p.k;
    SVTME_LANES(l) {
        clear_sad_slot(s, l);
        const LaneRef r = lane_ref(k, l);
        const int li = r.li, ri = r.ri, qw = r.qw, qh = r.qh;
        if (!r.used || !hme_active(k, li) || qh >= k.num_hme_sa_h || qw >= k.num_hme_sa_w) continue;
        const DevPlane &rp = p.refs[li][ri].lvl[0];
        if (k.prev_me_stage_based_exit_th && s.lsad[1][li][ri][qw][qh] < (k.prev_me_stage_based_exit_th >> 2)) {
            s.lx[2][li][ri][qw][qh] = s.lx[1][li][ri][qw][qh]; s.ly[2][li][ri][qw][qh] = s.ly[1][li][ri][qw][qh];
            s.lsad[2][li][ri][qw][qh] = s.lsad[1][li][ri][qw][qh];
            continue;
        }
        int16_t w = (int16_t)(((int16_t)k.hme_l2_sa.width + 7) & ~0x07), h = (int16_t)k.hme_l2_sa.height;
        int16_t ox = (int16_t)(-(w >> 1) + s.lx[1][li][ri][qw][qh]), oy = (int16_t)(-(h >> 1) + s.ly[1][li][ri][qw][qh]);
        clip_window((int16_t)s.org_x, (int16_t)s.org_y, 63, 63, (int16_t)rp.width, (int16_t)rp.height, ox, oy, w, h, true);
        make_sad_job(p, s, l, 0, li, ri, ox, oy, w, h, 0, 2, (uint8_t)qw, (uint8_t)qh);
    }
    SVTME_WSYNC();
}

// ---- set_final_seach_centre_sb (:2182-2380) followed by hme_prune_ref_and_adjust_sr (:2477-2518) ----
SVTME_HD void final_centre_and_prune(const MeParams &p, SbState &s) {
    const SvtMeKnobs &k = p.k;
    int L = -1;  // the three per-level blocks of the reference are mutually exclusive
    if (k.enable_hme_level0_flag && !k.enable_hme_level1_flag && !k.enable_hme_level2_flag) L = 0;
    if (k.enable_hme_level1_flag && !k.enable_hme_level2_flag) L = 1;
    if (k.enable_hme_level2_flag) L = 2;
    // a reference with HME results of its own: the minimum over its quadrants, lane = reference
    SVTME_LANES(l) {
        const LaneRef r = lane_ref(k, l);
        if (!r.used || r.q != 0 || !hme_active(k, r.li) || !k.enable_hme_flag || L < 0) continue;
        const int li = r.li, ri = r.ri;
        int16_t  xc = s.lx[L][li][ri][0][0], yc = s.ly[L][li][ri][0][0];
        uint64_t sad = s.lsad[L][li][ri][0][0];
        uint32_t w = 1, h = 0;  // scan starts at [1][0]: ties keep the earlier quadrant (:2226-2261)
        while (h < k.num_hme_sa_h) {
            while (w < k.num_hme_sa_w) {
                if (s.lsad[L][li][ri][w][h] < sad) { xc = s.lx[L][li][ri][w][h]; yc = s.ly[L][li][ri][w][h]; sad = s.lsad[L][li][ri][w][h]; }
                w++;
            }
            w = 0; h++;
        }
        s.hme_sc_x[li][ri] = xc; s.hme_sc_y[li][ri] = yc; s.hme_sad[li][ri] = sad;
    }
    SVTME_WSYNC();
    // the reference's locals carry over from one reference to the next (:2193, 2370): one lane walks the lists
    SVTME_LANES(l) {
        if (l != 0) continue;
        int16_t  xc = 0, yc = 0, sx = 0, sy = 0;
        uint64_t hme_mv_sad = 0;
        SVTME_FOR_REFS(k, li, ri) {
            if (hme_active(k, li)) {
                if (k.enable_hme_flag) {
                    if (L >= 0) { xc = s.hme_sc_x[li][ri]; yc = s.hme_sc_y[li][ri]; hme_mv_sad = s.hme_sad[li][ri]; }
                    sx = xc; sy = yc;
                }
            } else {
                sx = 0; sy = 0;
            }
            s.hme_sc_x[li][ri] = sx; s.hme_sc_y[li][ri] = sy; s.hme_sad[li][ri] = hme_mv_sad;
        }
        // temporal-filter mode: no reference pruning or SR adjustment (:3103); a block whose HME SAD is already under
        // the threshold skips the integer search altogether (:3109-3113)
        if (k.me_type == SVTME_ME_MCTF) s.tf_exit = s.hme_sad[0][0] < k.tf_me_exit_th;
    }
    SVTME_WSYNC();
    if (k.me_type == SVTME_ME_MCTF) return;
    if (!k.enable_hme_flag) return;  // prune_ref = enable_hme_flag && me_type != ME_MCTF (:3103, 3115)
    const uint16_t th = k.prune_ref_if_hme_sad_dev_bigger_than_th;
    SVTME_LANES(l) {  // every [2][4] slot, used or not
        const int li = l >> 4, ri = (l >> 2) & 3;
        if ((l & 3) != 0) continue;
        if (k.enable_me_hme_ref_pruning && th != (uint16_t)~0 && ri >= 1) {
            uint64_t best = ~0ull;  // over ALL [2][4] slots; unused ones hold the init value MAX_U32
            for (int i = 0; i < kMaxLists; i++)
                for (int j = 0; j < kMaxRefs; j++)
                    if (s.hme_sad[i][j] < best) best = s.hme_sad[i][j];
            if ((s.hme_sad[li][ri] - best) * 100 > ((uint64_t)th * best)) s.do_ref[li][ri] = 0;  // u64 wrap (:2492)
        }
        if (k.enable_me_sr_adjustment) {
            if (SVTME_ABS(s.hme_sc_x[li][ri]) <= k.reduce_me_sr_based_on_mv_length_th &&
                SVTME_ABS(s.hme_sc_y[li][ri]) <= k.reduce_me_sr_based_on_mv_length_th &&
                s.hme_sad[li][ri] < k.stationary_hme_sad_abs_th)
                s.divisor[li][ri] = k.stationary_me_sr_divisor;
            else if (s.hme_sad[li][ri] < k.reduce_me_sr_based_on_hme_sad_abs_th)
                s.divisor[li][ri] = k.me_sr_divisor_for_low_hme_sad;
        }
    }
    SVTME_WSYNC();
}

// ---- integer_search_b64 (:1249-1516), split into its data-dependent steps ----
// Step A (:1294-1320): area from distance / MV size / divisor, before any SAD is needed.
SVTME_HD void isearch_area(const MeParams &p, SbState &s, int li, int ri) {
    const SvtMeKnobs &k = p.k;
    int16_t xc = s.hme_sc_x[li][ri], yc = s.hme_sc_y[li][ri];
    int16_t w = (int16_t)k.me_sa.sa_min.width, h = (int16_t)k.me_sa.sa_min.height;
    const uint16_t dist = k.me_type == SVTME_ME_MCTF ? ref_dist(p, li, ri) : scaled_dist(ref_dist(p, li, ri));  // :1300
    w = (int16_t)SVTME_MIN((w * dist), (int)k.me_sa.sa_max.width);
    h = (int16_t)SVTME_MIN((h * dist), (int)k.me_sa.sa_max.height);
    if (k.mv_sa_adj_enabled && (!k.mv_sa_adj_nearest_ref_only || ri == 0)) {
        if (SVTME_ABS(xc) > k.mv_sa_adj_mv_size_th) w = (int16_t)(w * k.mv_sa_adj_sa_multiplier);
        if (SVTME_ABS(yc) > k.mv_sa_adj_mv_size_th) h = (int16_t)(h * k.mv_sa_adj_sa_multiplier);
    }
    // int16 / uint32 is an UNSIGNED division in the reference (:1314-1318)
    w = (int16_t)((SVTME_MAX(1u, ((uint32_t)(int32_t)w / s.divisor[li][ri])) + 7) & ~0x07u);
    h = (int16_t)SVTME_MAX(3u, ((uint32_t)(int32_t)h / s.divisor[li][ri]));
    s.isc_x[li][ri] = xc; s.isc_y[li][ri] = yc;
    s.isa_w[li][ri] = w; s.isa_h[li][ri] = h; s.isa_w0[li][ri] = w; s.isa_h0[li][ri] = h;
    s.c00_done[li][ri] = 0; s.c00_sad[li][ri] = 0; s.probe[li][ri] = 0;
}

// check_00_center (:1139-1206), first half: clamp the HME centre into the picture (+63 px) and queue the SAD(s).
SVTME_HD void isearch_c00_jobs(const MeParams &p, SbState &s, int li, int ri) {
    const SvtMeKnobs &k  = p.k;
    const DevPlane   &rp = p.refs[li][ri].lvl[0];
    const int16_t org_x = (int16_t)s.org_x, org_y = (int16_t)s.org_y, pad = 63;
    int16_t xc = s.isc_x[li][ri], yc = s.isc_y[li][ri];
    xc = ((org_x + xc) < -pad) ? (int16_t)(-pad - org_x) : xc;
    xc = ((org_x + xc) > (int16_t)rp.width - 1) ? (int16_t)(xc - ((org_x + xc) - ((int16_t)rp.width - 1))) : xc;
    yc = ((org_y + yc) < -pad) ? (int16_t)(-pad - org_y) : yc;
    yc = ((org_y + yc) > (int16_t)rp.height - 1) ? (int16_t)(yc - ((org_y + yc) - ((int16_t)rp.height - 1))) : yc;
    s.isc_x[li][ri] = xc; s.isc_y[li][ri] = yc;
    s.c00_done[li][ri] = 1;
    if (!k.me_early_exit_th) make_nxm_job(p, s, nxm_slot(li, ri, 0), li, ri, 0, 0, 1);  // zero-MV SAD (else zz_sad is reused, :1155-1156)
    make_nxm_job(p, s, nxm_slot(li, ri, 1), li, ri, xc, yc, 2);                         // HME-centre SAD
}

// Step B (:1321-1365 + 1391-1412): early exit / SR adjustment, then decide whether the 8x8-variance probe runs.
SVTME_HD void isearch_adjust(const MeParams &p, SbState &s, int li, int ri, uint32_t zero_sad_raw, uint32_t hme_sad_raw) {
    const SvtMeKnobs &k = p.k;
    int16_t  w = s.isa_w[li][ri], h = s.isa_h[li][ri];
    uint64_t best_hme_sad = ~0ull;
    if (k.me_early_exit_th) {
        if (s.zz_sad[li][ri] < (k.me_early_exit_th / 6)) { w = 1; h = 1; }
    } else {
        uint8_t accurate = 1;
        if (s.c00_done[li][ri]) {
            const uint32_t zero_mv_sad = zero_sad_raw << 1, hme_mv_sad = hme_sad_raw << 1;
            const uint64_t zc = (uint64_t)(uint32_t)(zero_mv_sad << 8), hc = (uint64_t)(uint32_t)(hme_mv_sad << 8);  // COST_PRECISION
            if (SVTME_MIN(zc, hc) == zc) { s.isc_x[li][ri] = 0; s.isc_y[li][ri] = 0; }
            best_hme_sad = hme_mv_sad;
            if (s.isc_x[li][ri] == 0 && s.isc_y[li][ri] == 0) accurate = 0;
        }
        if (k.enable_me_sr_adjustment == 2) {
            if ((accurate && (best_hme_sad < (24 * 24))) || (k.is_ref && s.hme_sad[li][ri] < (24 * 24))) h = (int16_t)(h / 2);
            if ((li || ri) && s.home->best_sad[0][0][0] < 5000 && h == s.isa_h0[li][ri] && w == s.isa_w0[li][ri]) {
                h = (int16_t)(h >> 1); w = (int16_t)(w >> 1);
            }
        }
    }
    s.isa_w[li][ri] = w; s.isa_h[li][ri] = h;
    s.probe[li][ri] = (k.me_8x8_var_enabled && (w * h > 24)) ? 1 : 0;
}

SVTME_HD void make_fp_job(const MeParams &p, SbState &s, int li, int ri, int16_t xo, int16_t yo, int16_t w, int16_t h) {
    const DevPlane &rp = p.refs[li][ri].lvl[0];
    FpJob          &j  = s.fp_jobs[li * kMaxRefs + ri];
    j.ref        = rp.org + (int64_t)((int32_t)s.org_y + yo) * rp.stride + ((int32_t)s.org_x + xo);
    j.ref_stride = rp.stride;
    j.xo = xo; j.yo = yo; j.w = w; j.h = h; j.li = (uint8_t)li; j.ri = (uint8_t)ri;
}

// Step C (:1414-1491): 8x8-variance scaling of the area (`ssq` = sum of squared deviations of the 64 8x8 SADs of the
// probe, wrapped to 32 bits as in the reference), then clip to the picture (+63 px) and queue the main scan.
SVTME_HD void isearch_final_job(const MeParams &p, SbState &s, int li, int ri, uint32_t ssq) {
    const SvtMeKnobs &k = p.k;
    int16_t w = s.isa_w[li][ri], h = s.isa_h[li][ri];
    const int16_t xc = s.isc_x[li][ri], yc = s.isc_y[li][ri];
    if (s.probe[li][ri]) {
        const uint32_t var = ssq / 64;
        if (var > k.me_sr_mult2_th) { w = (int16_t)((SVTME_MAX(1, w * 3 / 2) + 7) & ~0x7); h = (int16_t)SVTME_MAX(1, h * 3 / 2); }
        if (var < k.me_sr_div4_th) {
            w = (int16_t)((SVTME_MAX(1, w >> 2) + 7) & ~0x7);
            h = (int16_t)SVTME_MAX(1, h >> 2);
            h = (int16_t)SVTME_MAX(3, h);
        } else if (var < k.me_sr_div2_th) {
            w = (int16_t)((SVTME_MIN(w, w >> 1) + 7) & ~0x7);
            h = (int16_t)SVTME_MIN(h, h >> 1);
            h = (int16_t)SVTME_MAX(3, h);
        }
    }
    const DevPlane &f  = p.cur.lvl[0];
    const int16_t pic_w = (int16_t)((f.width + 7) & ~7), pic_h = (int16_t)((f.height + 7) & ~7);  // pcs->aligned_*
    const int16_t org_x = (int16_t)s.org_x, org_y = (int16_t)s.org_y, pad = 63;
    int16_t xo = (int16_t)(xc - (w >> 1)), yo = (int16_t)(yc - (h >> 1));
    xo = ((org_x + xo) < -pad) ? (int16_t)(-pad - org_x) : xo;  // left: origin only (:1445-1450)
    xo = ((org_x + xo) > pic_w - 1) ? (int16_t)(xo - ((org_x + xo) - (pic_w - 1))) : xo;
    w  = ((org_x + xo + w) > pic_w) ? (int16_t)SVTME_MAX(1, w - ((org_x + xo + w) - pic_w)) : w;
    w  = (w < 8) ? w : (int16_t)(w & ~0x07);
    yo = ((org_y + yo) < -pad) ? (int16_t)(-pad - org_y) : yo;
    yo = ((org_y + yo) > pic_h - 1) ? (int16_t)(yo - ((org_y + yo) - (pic_h - 1))) : yo;
    h  = (org_y + yo + h > pic_h) ? (int16_t)SVTME_MAX(1, h - ((org_y + yo + h) - pic_h)) : h;
    make_fp_job(p, s, li, ri, xo, yo, w, h);
}

// `arg`: 0 = reference (0,0) only, 1 = the others (SR adjustment level 2 lets p_sb_best_sad[0][0][0] of THIS SB steer
// them, :1355-1364), 2 = all references.
SVTME_HD bool is_in_pass(const SbState &s, int li, int ri, int arg) {
    return s.do_ref[li][ri] && !s.tf_exit && (arg == 2 || ((li == 0 && ri == 0) == (arg == 0)));
}
// A: area from distance / MV / divisor, and the zero-MV / HME-centre SAD jobs of check_00_center
SVTME_HD void is_a(const MeParams &p, SbState &s, int arg) {
    const SvtMeKnobs &k = p.k;
    clear_nxm_slots(s);
    SVTME_LANES(l) {
        const LaneRef r = lane_ref(k, l);
        if (!r.used || r.q != 0 || !is_in_pass(s, r.li, r.ri, arg)) continue;
        isearch_area(p, s, r.li, r.ri);
        if (!k.me_early_exit_th && (s.isc_x[r.li][r.ri] != 0 || s.isc_y[r.li][r.ri] != 0) && k.is_ref) isearch_c00_jobs(p, s, r.li, r.ri);
    }
    SVTME_WSYNC();
}
// B: early exit / SR adjustment; queue the 1x1 probe at the search centre, which runs BEFORE the main scan on the
// same best arrays (:1391-1412)
SVTME_HD void is_b(const MeParams &p, SbState &s, int arg) {
    const SvtMeKnobs &k = p.k;
    clear_fp_slots(s);
    SVTME_LANES(l) {
        const LaneRef r = lane_ref(k, l);
        if (!r.used || r.q != 0 || !is_in_pass(s, r.li, r.ri, arg)) continue;
        const int li = r.li, ri = r.ri;
        const NxmJob &jz = s.nxm_jobs[nxm_slot(li, ri, 0)], &jh = s.nxm_jobs[nxm_slot(li, ri, 1)];
        // zz_sad stands in for the zero-MV SAD when early exit is on
        const uint32_t zero_raw = (jz.rows > 0 && jz.kind == 1) ? jz.sad : s.zz_sad[li][ri];
        const uint32_t hme_raw  = (jh.rows > 0 && jh.kind == 2) ? jh.sad : 0;
        isearch_adjust(p, s, li, ri, zero_raw, hme_raw);
        if (s.probe[li][ri]) make_fp_job(p, s, li, ri, s.isc_x[li][ri], s.isc_y[li][ri], 1, 1);
    }
    SVTME_WSYNC();
}
// C: 8x8-variance scaling, clipping, the main scan.  The four lanes of a reference share the 64 8x8 SADs of its probe.
SVTME_HD void is_c(const MeParams &p, SbState &s, int arg) {
    const SvtMeKnobs &k = p.k;
    clear_fp_slots(s);
    SVTME_LANES(l) {
        const LaneRef r = lane_ref(k, l);
        uint32_t      part = 0;
        if (r.used && is_in_pass(s, r.li, r.ri, arg) && s.probe[r.li][r.ri]) {
            const uint32_t *bs = s.home->best_sad[r.li][r.ri];
            const uint32_t  mean = bs[0] / 64;
            for (int i = 16 * r.q; i < 16 * r.q + 16; i++) {
                const int32_t d = (int32_t)bs[21 + i] - (int32_t)mean;
                part += (uint32_t)(d * d);
            }
        }
        s.scratch[l] = part;
    }
    SVTME_WSYNC();
    SVTME_LANES(l) {
        const LaneRef r = lane_ref(k, l);
        if (!r.used || r.q != 0 || !is_in_pass(s, r.li, r.ri, arg)) continue;
        isearch_final_job(p, s, r.li, r.ri, s.scratch[l] + s.scratch[l + 1] + s.scratch[l + 2] + s.scratch[l + 3]);
    }
    SVTME_WSYNC();
}

// ---- me_prune_ref (:1522-1565) ----
SVTME_HD void me_prune_ref(const MeParams &p, SbState &s) {
    const SvtMeKnobs &k = p.k;
    SVTME_LANES(l) {
        const LaneRef r = lane_ref(k, l);
        uint64_t      part = 0;
        if (r.used && s.do_ref[r.li][r.ri])
            for (int i = 16 * r.q; i < 16 * r.q + 16; i++) part += s.home->best_sad[r.li][r.ri][21 + i];
        s.scratch64[l] = part;
    }
    SVTME_WSYNC();
    SVTME_LANES(l) {
        const LaneRef r = lane_ref(k, l);
        if (!r.used || r.q != 0) continue;
        s.hme_sad[r.li][r.ri] = s.do_ref[r.li][r.ri] == 0 ? (uint64_t)SVTME_MAX_SAD_VALUE * 64
                                                           : s.scratch64[l] + s.scratch64[l + 1] + s.scratch64[l + 2] + s.scratch64[l + 3];
    }
    SVTME_WSYNC();
    const uint16_t th = k.prune_ref_if_me_sad_dev_bigger_than_th;
    if (k.enable_me_hme_ref_pruning && th != (uint16_t)~0) {
        SVTME_LANES(l) {
            const int li = l >> 4, ri = (l >> 2) & 3;
            if ((l & 3) != 0 || ri < 1) continue;
            uint64_t best = ~0ull;
            for (int i = 0; i < kMaxLists; i++)
                for (int j = 0; j < kMaxRefs; j++)
                    if (s.hme_sad[i][j] < best) best = s.hme_sad[i][j];
            if ((s.hme_sad[li][ri] - best) * 100 > ((uint64_t)th * best)) s.do_ref[li][ri] = 0;
        }
        SVTME_WSYNC();
    }
}

// ---- ME candidate construction (:2532-2835), one PU (n_idx, tier-zero order) per call so PUs run in parallel ----
struct SbOut {
    uint32_t *mv;     // [pu_count][max_refs]
    uint8_t  *cand;   // [pu_count][max_cand]
    uint8_t  *count;  // [pu_count]
};

// z_to_raster (:2520-2531), computed: tier-zero index -> raster index inside its level.
// 16x16: n = 5 + z16 with z16 = (quad32 << 2) | sub; 8x8: n = 21 + 4*z16 + q (q = row*2 + col inside the 16x16).
SVTME_HD uint8_t z_to_raster(unsigned n) {
    if (n < 5) return (uint8_t)n;
    const unsigned z16 = n < 21 ? n - 5 : (n - 21) >> 2;
    const unsigned r16 = ((z16 >> 3) << 1) | ((z16 >> 1) & 1), c16 = (((z16 >> 2) & 1) << 1) | (z16 & 1);
    if (n < 21) return (uint8_t)(5 + r16 * 4 + c16);
    const unsigned q = (n - 21) & 3;
    return (uint8_t)(21 + (2 * r16 + (q >> 1)) * 8 + 2 * c16 + (q & 1));
}
// MeCandidate bit-fields under GCC (me_sb_results.h:28-34), LSB first, values truncated to the field width
SVTME_HD uint8_t pack_cand(unsigned dir, unsigned i0, unsigned i1, unsigned l0, unsigned l1) {
    return (uint8_t)((dir & 3) | ((i0 & 3) << 2) | ((i1 & 3) << 4) | ((l0 & 1) << 6) | ((l1 & 1) << 7));
}
SVTME_HD bool use_me_pu(const SvtMeKnobs &k, unsigned n) { return k.enable_me_16x16 ? (k.enable_me_8x8 || n < 21) : n < 5; }
SVTME_HD unsigned number_of_pus(const SvtMeKnobs &k) { return k.enable_me_16x16 ? (k.enable_me_8x8 ? k.max_number_of_pus_per_sb : 21) : 5; }

SVTME_HD void cand_one_pu(const MeParams &p, SbState &s, const SbOut &o, unsigned n) {
    const SvtMeKnobs &k   = p.k;
    SbState          &t   = *s.home;  // the result tables live in the block's own state (HBM/L2)
    const uint8_t    *nr  = k.num_of_ref_pic_to_search;
    const bool        use = use_me_pu(k, n);
    if (nr[0] == 1 && nr[1] == 0) {  // construct_me_candidate_array_single_ref (:2646-2697)
        const unsigned pu = z_to_raster(n);
        t.me_distortion[pu] = t.best_sad[0][0][n];
        if (s.do_ref[0][0] == 0) return;
        if (use) { o.cand[pu * k.max_cand] = pack_cand(0, 0, 0, 0, 0); o.mv[pu * k.max_refs] = t.best_mv[0][0][n]; }
        return;
    }
    if (nr[0] == 1 && nr[1] == 1) {  // construct_me_candidate_array_mrp_off (:2532-2645)
        const unsigned pu = z_to_raster(n);
        uint32_t num_lists = k.num_of_list_to_search;
        uint8_t org0 = s.do_ref[0][0], org1 = (num_lists == 1) ? 0 : s.do_ref[1][0];
        if (num_lists < 2 || !s.do_ref[1][0]) num_lists = 1;
        const uint32_t prune_th = (org0 && org1) ? (uint32_t)k.prune_me_candidates_th : 0;
        uint8_t blk[2] = {org0, org1};
        uint8_t off = 0;
        const uint32_t s0 = t.best_sad[0][0][n], s1 = t.best_sad[1][0][n];
        const uint32_t best = (org0 && org1) ? SVTME_MIN(s0, s1) : org0 ? s0 : s1;
        t.me_distortion[pu] = best;
        int min_list = -1;
        if (k.use_best_unipred_cand_only && blk[0] && blk[1]) min_list = s0 < s1 ? 0 : 1;
        for (int li = 0; (uint32_t)li < num_lists && (use || off == 0); ++li) {
            if (blk[li] == 0) continue;
            if (prune_th > 0) {
                const uint32_t d = (t.best_sad[li][0][n] - best) * 100;  // u32 (:2597-2600)
                if (d > (best * prune_th)) { blk[li] = 0; continue; }
            }
            if (min_list != -1 && min_list != li) {
                if (use) o.mv[pu * k.max_refs + (li ? k.max_l0 : 0)] = t.best_mv[li][0][n];
                continue;
            }
            if (use) {
                o.cand[pu * k.max_cand + off]                 = pack_cand((unsigned)li, 0, 0, li == 0 ? 0 : 24, li == 1 ? 1 : 24);
                o.mv[pu * k.max_refs + (li ? k.max_l0 : 0)] = t.best_mv[li][0][n];
            }
            off++;
        }
        if (blk[0] && blk[1] && use) {
            o.cand[pu * k.max_cand + off] = pack_cand(2, 0, 0, 0, 1);  // BI_PRED (LAST, BWD)
            o.count[pu]                   = (uint8_t)(off + 1);
        }
        return;
    }
    // construct_me_candidate_array (:2698-2835)
    const uint32_t num_lists = k.num_of_list_to_search;
    const unsigned pu = (n > 4) ? z_to_raster(n) : n;
    uint8_t off = 0;
    uint8_t blk[kMaxLists][kMaxRefs] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
    const uint32_t prune_th = (uint32_t)k.prune_me_candidates_th;
    uint32_t best = ~0u;
    for (uint32_t li = 0; li < num_lists; li++)
        for (uint32_t ri = 0; ri < nr[li]; ri++) {
            blk[li][ri] = s.do_ref[li][ri];
            if (!blk[li][ri]) continue;
            if (t.best_sad[li][ri][n] < best) best = t.best_sad[li][ri][n];
        }
    t.me_distortion[pu] = best;
    for (uint32_t li = 0; li < num_lists && (use || off == 0); ++li)
        for (uint32_t ri = 0; ri < nr[li] && (use || off == 0); ++ri) {
            if (!blk[li][ri]) continue;
            if (prune_th > 0) {
                const uint32_t d = (t.best_sad[li][ri][n] - best) * 100;  // u32 (:2744-2748)
                if (d > (best * prune_th)) { blk[li][ri] = 0; continue; }
            }
            if (use) {
                o.cand[pu * k.max_cand + off] = pack_cand(li, ri, ri, li == 0 ? 0 : 24, li == 1 ? 1 : 24);
                o.mv[pu * k.max_refs + (li ? k.max_l0 : 0) + ri] = t.best_mv[li][ri][n];
            }
            off++;
        }
    if (num_lists == 2 && use) {
        for (uint32_t a = 0; a < nr[0]; a++)
            for (uint32_t b = 0; b < nr[1]; b++) {
                if (k.only_l_bwd && (a > 0 || b > 0)) continue;
                if (blk[0][a] && blk[1][b]) o.cand[pu * k.max_cand + off++] = pack_cand(2, a, b, 0, 1);
            }
        if (!k.only_l_bwd) {
            for (uint32_t a = 1; a < nr[0]; a++)
                if (blk[0][0] && blk[0][a]) o.cand[pu * k.max_cand + off++] = pack_cand(2, 0, a, 0, 0);
            if (nr[1] == 3 && blk[1][0] && blk[1][2]) o.cand[pu * k.max_cand + off++] = pack_cand(2, 0, 2, 1, 1);
        }
    }
    if (use) o.count[pu] = off;
}

// ---- compute_distortion (:2964-3007); leader only ----
SVTME_HD void compute_distortion(const MeParams &p, const SbState &s, uint32_t sums[4]) {
    const SbState &t = *s.home;
    const uint32_t i = s.b64_index;
    uint32_t d64 = t.me_distortion[0], d32 = 0, d16 = 0, d8 = 0;
    for (int j = 0; j < 4; j++) d32 += t.me_distortion[1 + j];
    for (int j = 0; j < 16; j++) d16 += t.me_distortion[5 + j];
    for (int j = 0; j < 64; j++) d8 += t.me_distortion[21 + j];
    const uint64_t mean = d8 / 64;
    uint64_t       ssq  = 0;
    for (int j = 0; j < 64; j++) {
        const int64_t diff = (int64_t)t.me_distortion[21 + j] - (int64_t)mean;
        ssq += (uint64_t)(diff * diff);
    }
    p.out.me_8x8_cost_variance[i] = (uint32_t)(ssq / 64);
    p.out.rc_me_distortion[i]     = (p.k.input_resolution <= SVTME_INPUT_SIZE_480p) ? d8 : d16;
    const uint32_t pix            = s.b64_w * s.b64_h;  // b64_geom width * height
    p.out.me_64x64_distortion[i]  = (d64 * 4096u) / pix;  // u32 products as in the reference (:3003-3006)
    p.out.me_32x32_distortion[i]  = (d32 * 4096u) / pix;
    p.out.me_16x16_distortion[i]  = (d16 * 4096u) / pix;
    p.out.me_8x8_distortion[i]    = (d8 * 4096u) / pix;
    sums[0] = (d8 * 4096u) / pix;
    sums[1] = (p.k.input_resolution <= SVTME_INPUT_SIZE_480p) ? d8 : d16;
}

// ---- perform_gm_detection (:2838-2961); leader only.  n indexes BOTH the raster-ordered candidate array and the
// tier-zero-ordered p_sb_best_mv, exactly as the reference does. ----
SVTME_HD void gm_detection(const MeParams &p, const SbState &s, const SbOut &o, uint32_t sums[4]) {
    const SvtMeKnobs &k = p.k;
    const SbState    &t = *s.home;
    uint32_t stationary = 0, tot = 0;
    uint8_t  cnt[kMaxLists][kMaxRefs][2][2];
    for (int l = 0; l < kMaxLists; l++)
        for (int r = 0; r < kMaxRefs; r++) { cnt[l][r][0][0] = 0; cnt[l][r][0][1] = 0; cnt[l][r][1][0] = 0; cnt[l][r][1][1] = 0; }
    const bool     low_res = k.input_resolution <= SVTME_INPUT_SIZE_480p;
    const unsigned n_blk   = low_res ? 64 : 16;
    for (unsigned i = 0; i < n_blk; i++) {
        unsigned n = (low_res ? 21 : 5) + i;
        if (low_res) {
            if (!k.enable_me_8x8) {
                // me_idx_85_8x8_to_16x16_conversion (definitions.h:2613): raster 8x8 -> raster 16x16
                if (n >= 21) { const unsigned r = (n - 21) >> 3, c = (n - 21) & 7; n = 5 + (r >> 1) * 4 + (c >> 1); }
                // me_idx_16x16_to_parent_32x32_conversion (definitions.h:2626)
                if (!k.enable_me_16x16 && n >= 5) { const unsigned r = (n - 5) >> 2, c = (n - 5) & 3; n = 1 + (r >> 1) * 2 + (c >> 1); }
            }
        } else if (!k.enable_me_16x16 && n >= 5) {
            const unsigned r = (n - 5) >> 2, c = (n - 5) & 3; n = 1 + (r >> 1) * 2 + (c >> 1);
        }
        const uint8_t  c   = o.cand[n * k.max_cand];
        const unsigned dir = c & 3;
        const unsigned li  = (dir == 0 || dir == 2) ? ((c >> 6) & 1) : ((c >> 7) & 1);
        const unsigned ri  = (dir == 0 || dir == 2) ? ((c >> 2) & 3) : ((c >> 4) & 3);
        const uint64_t pp = k.picture_number, rp = p.refs[li][ri].picture_number;
        int active_th;
        if (low_res) {
            const int16_t  d16  = (int16_t)(SVTME_MAX(pp, rp) - SVTME_MIN(pp, rp));
            const uint16_t dist = (uint16_t)SVTME_ABS(d16);
            active_th           = k.gm_use_distance_based_active_th ? SVTME_MAX(dist >> 1, 4) : 4;
        } else {
            const int16_t  d16  = (int16_t)(pp - rp);
            const uint16_t dist = (uint16_t)SVTME_ABS(d16);
            active_th           = k.gm_use_distance_based_active_th ? SVTME_MAX(dist * 16, 32) : 32;
        }
        const uint32_t mv = t.best_mv[li][ri][n];
        const int mx = (int)(int16_t)(mv & 0xFFFF) * 4, my = (int)(int16_t)(mv >> 16) * 4;
        if (mx < -active_th) cnt[li][ri][0][0]++; else if (mx > active_th) cnt[li][ri][0][1]++;
        if (my < -active_th) cnt[li][ri][1][0]++; else if (my > active_th) cnt[li][ri][1][1]++;
        const int st = low_res ? 0 : 4;
        if (SVTME_ABS(mx) <= st && SVTME_ABS(my) <= st) stationary++;
        tot++;
    }
    if (stationary > ((tot * 5) / 100)) { p.out.stationary_block_present_sb[s.b64_index] = 1; sums[2] = 1; }
    for (int l = 0; l < kMaxLists; l++)
        for (int r = 0; r < kMaxRefs; r++)
            for (int c = 0; c < 2; c++)
                for (int q = 0; q < 2; q++)
                    if (cnt[l][r][c][q] > (tot / 2)) { p.out.rc_me_allow_gm[s.b64_index] = 1; sums[3] = 1; }
}

// ---- the step program: svt_aom_motion_estimation_b64 (:3076-3153) for ME_OPEN_LOOP with hme_b64 (:2441-2475) and
// integer_search_b64 (:1249-1516) unrolled into control steps (scalar, per SB) and engine steps (data parallel) ----
// STEP_PROBE: the 1x1 full-pel "search" at the search centre that feeds the 8x8-variance test -- an FpJob list like
// STEP_FP, but light enough to run beside the control steps instead of in a search-engine kernel
// STEP_PLAN16: cut the sixteenth-plane jobs just made into sub-windows and batches for st_sad16_kernel (device only: the
// host emulation's searches need no plan)
// STEP_KEYS16: read the minima the engine's warps folded into that plan back into the job table
enum StepKind : uint8_t { STEP_CTL = 0, STEP_NXM = 1, STEP_SAD = 2, STEP_FP = 3, STEP_PROBE = 4, STEP_PLAN16 = 5, STEP_KEYS16 = 6 };
enum CtlOp : uint8_t {
    C_ZZ_MAKE, C_ZZ_FOLD, C_PREHME_MAKE, C_SAD_FOLD, C_PREHME_FINAL, C_L0_MAKE, C_L0_FOLD, C_L1_MAKE, C_L2_MAKE,
    C_FINAL_CENTRE, C_IS_A, C_IS_B, C_IS_C, C_PRUNE
};

SVTME_HDX int build_program(const SvtMeKnobs &k, Step *st) {
    int n = 0;
    auto ctl = [&](uint8_t op, uint8_t arg) { st[n].kind = STEP_CTL; st[n].op = op; st[n].arg = arg; n++; };
    auto eng = [&](uint8_t kind, uint8_t arg = 0) { st[n].kind = kind; st[n].op = 0; st[n].arg = arg; n++; };  // SAD steps: arg = plane level
    const bool l1_hme = k.temporal_layer_index > 0;  // hme_active(k, 1)
    if (k.me_early_exit_th || k.me_safe_limit_zz_th) { ctl(C_ZZ_MAKE, 0); eng(STEP_NXM); ctl(C_ZZ_FOLD, 0); }
    if (k.prehme_enable) {
        const bool split = k.num_of_list_to_search > 1 && l1_hme && k.prehme_l1_early_exit;
        for (int pass = 0; pass < (split ? 2 : 1); pass++) {
            ctl(C_PREHME_MAKE, (uint8_t)(split ? pass : 2));
            if (k.hme_search_method == 0) eng(STEP_PLAN16);
            eng(STEP_SAD, 2);
            if (k.hme_search_method == 0) eng(STEP_KEYS16);
            ctl(C_SAD_FOLD, 0);
        }
        ctl(C_PREHME_FINAL, 0);
    }
    if (k.enable_hme_flag) {
        if (k.enable_hme_level0_flag) {
            const bool dep = k.enable_me_sr_adjustment && k.distance_based_hme_resizing && k.reduce_hme_l0_sr_th_min &&
                k.reduce_hme_l0_sr_th_max;
            for (int pass = 0; pass < (dep ? 2 : 1); pass++) {
                ctl(C_L0_MAKE, (uint8_t)(dep ? pass : 2));
                if (k.hme_search_method == 0) eng(STEP_PLAN16);
                eng(STEP_SAD, 2);
                if (k.hme_search_method == 0) eng(STEP_KEYS16);
                ctl(C_L0_FOLD, 0);
            }
        }
        if (k.enable_hme_level1_flag) { ctl(C_L1_MAKE, 0); eng(STEP_SAD, 1); ctl(C_SAD_FOLD, 0); }
        if (k.enable_hme_level2_flag) { ctl(C_L2_MAKE, 0); eng(STEP_SAD, 0); ctl(C_SAD_FOLD, 0); }
    }
    ctl(C_FINAL_CENTRE, 0);
    {
        const bool dep = (k.enable_me_sr_adjustment == 2) && !k.me_early_exit_th;
        for (int pass = 0; pass < (dep ? 2 : 1); pass++) {
            const uint8_t a = (uint8_t)(dep ? pass : 2);
            ctl(C_IS_A, a);
            if (!k.me_early_exit_th && k.is_ref) eng(STEP_NXM);  // check_00_center's SADs: is_a queues them only then
            ctl(C_IS_B, a);
            if (k.me_8x8_var_enabled) eng(STEP_PROBE);  // the 1x1 probe of the 8x8-variance test
            ctl(C_IS_C, a); eng(STEP_FP);
        }
    }
    if (k.enable_hme_flag && k.enable_me_hme_ref_pruning && k.me_type != SVTME_ME_MCTF) ctl(C_PRUNE, 0);
    return n;
}

SVTME_HD void ctl_step(const MeParams &p, SbState &s, uint8_t op, uint8_t arg) {
    switch (op) {
    case C_ZZ_MAKE: zz_make(p, s); break;
    case C_ZZ_FOLD: zz_fold(p, s); break;
    case C_PREHME_MAKE: prehme_make(p, s, arg); break;
    case C_SAD_FOLD: sad_fold(p, s); break;
    case C_PREHME_FINAL: prehme_final(p, s); break;
    case C_L0_MAKE: l0_make(p, s, arg); break;
    case C_L0_FOLD: l0_fold(p, s); break;
    case C_L1_MAKE: l1_make(p, s); break;
    case C_L2_MAKE: l2_make(p, s); break;
    case C_FINAL_CENTRE: final_centre_and_prune(p, s); break;
    case C_IS_A: is_a(p, s, arg); break;
    case C_IS_B: is_b(p, s, arg); break;
    case C_IS_C: is_c(p, s, arg); break;
    case C_PRUNE: me_prune_ref(p, s); break;
    default: break;
    }
}

// candidates (one PU per thread), distortion statistics and GM detection of one SB (:3126-3151)
template <class B>
SVTME_HD void sb_results(B &be, const MeParams &p, SbState &s, uint32_t b64_index) {
    const SvtMeKnobs &k = p.k;
    SbOut o;
    o.mv    = p.out.me_mv + (size_t)b64_index * p.pu_count * k.max_refs;
    o.cand  = p.out.me_cand + (size_t)b64_index * p.pu_count * k.max_cand;
    o.count = p.out.total_me_candidate_index + (size_t)b64_index * p.pu_count;
    const uint8_t *nr     = k.num_of_ref_pic_to_search;
    const bool     simple = nr[0] == 1 && nr[1] <= 1;  // single_ref / mrp_off start from count = 1 (:2555, 2664)
    const unsigned npus   = number_of_pus(k);
    for (uint32_t i = be.tid(); i < p.pu_count * k.max_refs; i += be.nthreads()) o.mv[i] = 0;
    for (uint32_t i = be.tid(); i < p.pu_count * k.max_cand; i += be.nthreads()) o.cand[i] = 0;
    for (uint32_t i = be.tid(); i < p.pu_count; i += be.nthreads()) o.count[i] = (simple && i < npus) ? 1 : 0;
    be.sync();
    for (unsigned n = be.tid(); n < k.max_number_of_pus_per_sb; n += be.nthreads()) cand_one_pu(p, s, o, n);
    be.sync();
    if (be.leader()) {
        uint32_t sums[4] = {0, 0, 0, 0};
        compute_distortion(p, s, sums);
        p.out.stationary_block_present_sb[b64_index] = 0;
        p.out.rc_me_allow_gm[b64_index]              = 0;
        if (k.gm_enabled) gm_detection(p, s, o, sums);
        if (p.out.picture_sums) {
            // what Initial RC (initial_rc_process.c:683-691) and svt_aom_global_motion_estimation (global_me.c:162-168)
            // sum over the per-block arrays once the picture is complete
            unsigned long long *ps = p.out.picture_sums;
#ifdef __CUDA_ARCH__
            ps += (b64_index % kSumSlots) * kSumSlotWords;
            if (sums[0]) atomicAdd(ps + 0, (unsigned long long)sums[0]);
            if (sums[1]) atomicAdd(ps + 1, (unsigned long long)sums[1]);
            if (sums[2]) atomicAdd(ps + 2, 1ull);
            if (sums[3]) atomicAdd(ps + 3, 1ull);
#else
            for (int i = 0; i < 4; i++) ps[i] += sums[i];
#endif
        }
    }
}

// ---- results of one SB: candidates (one PU per thread), distortion statistics, GM detection, optional exports ----
template <class B>
SVTME_HD void sb_finish(B &be, const MeParams &p, SbState &s, uint32_t b64_index) {
    const SvtMeKnobs &k = p.k;
    const SbState    &t = *s.home;
    const uint8_t *nr     = k.num_of_ref_pic_to_search;
    if (k.me_type == SVTME_ME_MCTF) {
        // no candidates, distortions or GM for the temporal filter (:3126); it reads the best MVs / SADs below
        if (be.leader()) {
            p.out.tf_exit_sb[b64_index] = s.tf_exit;
            uint32_t *cnt = p.out.tf_blk_counts + (SVTME_ABS(s.hme_sc_x[0][0]) > SVTME_ABS(s.hme_sc_y[0][0]) ? 0 : 1);  // :2469-2474
#ifdef __CUDA_ARCH__
            atomicAdd(cnt, 1u);
#else
            ++*cnt;
#endif
        }
    } else {
        sb_results(be, p, s, b64_index);
    }
    // optional MeContext state for parity checks
    if (p.out.sb_best_sad)
        for (uint32_t i = be.tid(); i < kMaxLists * kMaxRefs * kPus; i += be.nthreads()) p.out.sb_best_sad[(size_t)b64_index * 680 + i] = (&t.best_sad[0][0][0])[i];
    if (p.out.sb_best_mv)
        for (uint32_t i = be.tid(); i < kMaxLists * kMaxRefs * kPus; i += be.nthreads()) p.out.sb_best_mv[(size_t)b64_index * 680 + i] = (&t.best_mv[0][0][0])[i];
    if (p.out.tf_best_sad)
        for (uint32_t i = be.tid(); i < kPus; i += be.nthreads()) p.out.tf_best_sad[(size_t)b64_index * kPus + i] = t.best_sad[0][0][i];
    if (p.out.tf_best_mv)
        for (uint32_t i = be.tid(); i < kPus; i += be.nthreads()) p.out.tf_best_mv[(size_t)b64_index * kPus + i] = t.best_mv[0][0][i];
    if (be.leader()) {
        for (int li = 0; li < kMaxLists; li++)
            for (int ri = 0; ri < kMaxRefs; ri++) {
                const bool used = li < k.num_of_list_to_search && ri < nr[li];
                const size_t o8 = (size_t)b64_index * 8 + li * 4 + ri;
                if (p.out.do_ref) p.out.do_ref[o8] = used ? s.do_ref[li][ri] : 0;
                if (p.out.hme_sc) { p.out.hme_sc[o8 * 2] = used ? s.hme_sc_x[li][ri] : 0; p.out.hme_sc[o8 * 2 + 1] = used ? s.hme_sc_y[li][ri] : 0; }
            }
    }
    be.sync();
}

// ---- one CTA (or the scalar test backend) runs the whole program for one SB ----
template <class B>
SVTME_HD void me_sb(B &be, const MeParams &p, SbState &s, uint32_t b64_index) {
    sb_init(be, p, s, b64_index);
    be.stage_src(s);  // engines keep the 64x64 source block on chip for the whole SB
    for (int i = 0; i < p.n_steps; i++) {
        const Step st = p.prog[i];
        switch (st.kind) {
        case STEP_CTL: ctl_step(p, s, st.op, st.arg); break;  // a warp on the device, 32 sequential lanes on the host
        case STEP_NXM: be.run_nxm(s); break;
        case STEP_SAD: be.run_sad(s); break;
        case STEP_PLAN16: case STEP_KEYS16: break;
        default: be.run_fullpel(s); break;
        }
        be.sync();
    }
    sb_finish(be, p, s, b64_index);
}

}  // namespace svtme
