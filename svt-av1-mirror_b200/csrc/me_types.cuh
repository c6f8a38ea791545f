// me_types.cuh -- device-side data structures of the open-loop ME path.
//
// Names follow the reference's domain: a picture has three luma planes (full, quarter, sixteenth);
// a 64x64 block ("b64"/SB) is searched against up to 2x4 reference pictures; results are per
// square PU (85 per SB, "tier-zero" order, me_context.h:52-138).
#pragma once
#include <stdint.h>
#include "../../include/svtme_cuda.h"

#if defined(__CUDACC__)
#define SVTME_HD __device__ __forceinline__
#define SVTME_HDX __host__ __device__ inline
#define SVTME_D __device__ __forceinline__
#else
#define SVTME_HD inline
#define SVTME_HDX inline
#define SVTME_D inline
#endif

namespace svtme {

constexpr int kMaxLists = SVTME_MAX_LISTS;
constexpr int kMaxRefs  = SVTME_MAX_REFS_PER_LIST;
constexpr int kPus      = SVTME_SQUARE_PU_COUNT;
constexpr uint32_t kMaxU32 = 0xFFFFFFFFu;

// One padded luma plane in HBM.  `org` points at pixel (0,0); at least `pad_x`/`pad_y` replicated
// pixels exist on every side.  `ref_org` is the origin (= padding) the REFERENCE's buffer of this
// level has (72 / 32 / 16: enc_handle.c:4057-4070, reference_object.c:267-295); the HME window
// clipping depends on it (pad_width = org_x - 1, motion_estimation.c:838-839), independent of how
// much padding the device buffer really has.
struct DevPlane {
    const uint8_t *org;
    int32_t        stride;
    int32_t        width, height;
    int32_t        ref_org;
};

struct DevPic {
    DevPlane lvl[3];  // 0 full, 1 quarter, 2 sixteenth
    uint64_t picture_number;
};

// Parameters of the TPL source-based search kernel (me_tpl.cu), passed by value.
struct TplParams {
    DevPic          cur;
    DevPic          refs[kMaxLists][kMaxRefs];
    const uint32_t *me_mv;     // [b64][pus][max_refs] of the picture's ME results, in HBM
    const uint8_t  *me_cand;   // [b64][pus][max_cand]
    const uint8_t  *me_count;  // [b64][pus]
    uint32_t        pus, max_refs, max_cand, max_l0;
    uint32_t        wb, hb, grid_width;
    uint8_t         level, disable_intra_pred, slice_type_i, enable_me_16x16;
    uint32_t        ref_excluded;  // bit list * 4 + index
    uint64_t        ref_poc[kMaxLists * kMaxRefs];
    uint8_t        *block_size, *best_mode;
    int8_t         *best_rf_idx;
    int16_t        *mv;
    uint64_t       *ref_frame_poc;
    uint32_t       *best_inter_cost, *best_intra_cost;
};

// Device result buffers of one picture (same layout as SvtMePictureResults).
struct DevResults {
    uint32_t *me_mv;
    uint8_t  *me_cand;
    uint8_t  *total_me_candidate_index;
    uint32_t *me_64x64_distortion, *me_32x32_distortion, *me_16x16_distortion, *me_8x8_distortion;
    uint32_t *me_8x8_cost_variance, *rc_me_distortion;
    uint8_t  *stationary_block_present_sb, *rc_me_allow_gm;
    uint32_t *sb_best_sad, *sb_best_mv;  // optional
    uint8_t  *do_ref;                    // optional
    int16_t  *hme_sc;                    // optional
    uint8_t  *tf_exit_sb;                // temporal-filter mode only
    uint32_t *tf_blk_counts;             // temporal-filter mode only: [horz, vert]
    // optional: sums of me_8x8_distortion, rc_me_distortion, stationary blocks, GM votes.  On the device kSumSlots partial
    // sums, 64 bytes apart (block b adds into slot b % kSumSlots): 2040 blocks adding into ONE address serialise in L2 for
    // ~16 us; the host adds the slots up.  The host emulation uses slot 0 only.
    unsigned long long *picture_sums;
    uint32_t *tf_best_sad, *tf_best_mv;  // optional, temporal-filter mode: [b64][85] = the [0][0] rows of sb_best_sad / sb_best_mv
};
constexpr int kSumSlots = 64, kSumSlotWords = 8;

// device padding of the three planes (>= the reference's 72 / 32 / 16, rounded so that rows stay 16-byte aligned)
constexpr int kDevPad[3] = {96, 48, 32};
// TMA boxes of the sixteenth plane: kS16Classes widths in bytes (pitch = 4 x odd words: rows 8 banks apart) x
// kS16HClasses heights (8 / 32 / 128 rows: the TMA unit's cost is per box, so tall windows take tall boxes);
// descriptor index = width class * kS16HClasses + height class
constexpr int kS16Classes = 5, kS16HClasses = 3;
constexpr int kS16BoxBytes[kS16Classes] = {48, 80, 112, 144, 176};
constexpr int kS16BoxRows[kS16HClasses] = {8, 32, 128};
constexpr int kTmapBytes = 128;

// One step of the per-SB program (me_control.cuh: build_program)
struct Step {
    uint8_t kind, op, arg;
};
constexpr int kMaxSteps = 48;

struct SbState;

struct MeParams {
    SvtMeKnobs k;
    DevPic     cur;
    DevPic     refs[kMaxLists][kMaxRefs];
    DevResults out;
    uint32_t   wb, hb;      // picture size in 64x64 blocks
    uint32_t   pu_count;    // 85 / 21 / 5
    uint32_t  *work_counter;  // persistent-CTA SB queue
    unsigned long long *px_counter;  // optional: algorithmic pixel-compares of the picture (roofline numerator)
    SbState   *states;        // staged pipeline: one SbState per 64x64 block in HBM/L2
    void      *plans16;       // one S16Plan per block (me_staged.cu): the sixteenth-plane searches' sub-windows
    uint32_t  *q_units;       // unit queue of the sixteenth-plane engine: one entry per sub-window of the picture
    uint32_t  *q_ctr;         // [count, head] per program step (zeroed when the search is submitted)
    int32_t    n_steps;
    Step       prog[kMaxSteps];
};

// TMA descriptors of every reference picture's padded sixteenth plane: kS16Classes x kS16HClasses CUtensorMaps (128 bytes
// each), box = kS16BoxBytes[class] x kS16BoxRows[height class]; st_sad16_kernel stages its search windows with them.
// Passed to the kernels BY VALUE as a __grid_constant__ parameter (17 KB of the 32 KB a launch may carry): descriptors
// in parameter space need no proxy fence before the TMA unit fetches them, unlike descriptors read from global memory
// (a fence.proxy.tensormap per CTA was measured at -8 % of the picture rate).
// Two more descriptors per picture for the HME level-2 windows (st_sad_rows_kernel<16, 32>): the even and the odd rows of the
// padded full plane as two tensors of doubled row pitch; box = one parity plane of a window (112 bytes x 36 rows).  Index
// kS16Classes * kS16HClasses + parity.  (The same for the quarter plane / HME level 1 -- 80 bytes x 20 rows -- was measured
// 0.5 % slower than that kernel's 16-byte cp.async copies and is not built.)
constexpr int kRowsTmaMaps = 2;
constexpr int kMapsPerPic  = kS16Classes * kS16HClasses + kRowsTmaMaps;
constexpr int kRowsTmaBoxW = 112, kRowsTmaBoxH = 36;
struct alignas(64) S16Maps {
    unsigned char m[kMaxLists * kMaxRefs][kS16Classes * kS16HClasses][kTmapBytes];
};
// the parity-view descriptors of the references' full planes, for st_sad_rows_kernel<16, 32> (2 KB: a launch copies its
// parameters, and at 1080p M10 the pipeline is bound by launches)
struct alignas(64) RowsMaps {
    unsigned char m[kMaxLists * kMaxRefs][kRowsTmaMaps][kTmapBytes];
};

// ---- work descriptors produced by the per-SB control code and consumed by the search engines ----

// One svt_sad_loop_kernel call (compute_sad_c.c:58): block (block_w x rows, rows at blk_stride) against every
// position of an sa_w x sa_h window whose rows are raw_stride apart; block rows sample the reference at ref_stride.
struct SadJob {
    const uint8_t *blk;
    const uint8_t *ref;
    int32_t        blk_stride, ref_stride, raw_stride;
    int16_t        rows, block_w;
    int16_t        sa_w, sa_h;
    int16_t        sa_ox, sa_oy;  // window origin relative to the block position (added to the result)
    uint8_t        skip_line;     // search odd candidate rows only (16-wide blocks of <= 16 rows)
    uint8_t        mv_scale;      // 4 / 2 / 1
    uint8_t        level, li, ri, qw, qh;  // destination: HME level (3 = pre-HME, qw = search region)
    uint8_t        found;
    uint32_t       best_sad;  // raw (not doubled)
    int16_t        bx, by;    // raw argmin inside the window
};

// One svt_nxm_sad_kernel call (compute_sad_c.c:209)
struct NxmJob {
    const uint8_t *src;
    const uint8_t *ref;
    int32_t        src_stride, ref_stride;
    int16_t        rows, width;
    uint8_t        li, ri, kind;
    uint32_t       sad;
};

// One open_loop_me_fullpel_search_sblock call (motion_estimation.c:781)
struct FpJob {
    const uint8_t *ref;  // reference pixel under the block's top-left for MV (xo, yo)
    int32_t        ref_stride;
    int16_t        xo, yo;
    int16_t        w, h;
    uint8_t        li, ri;
};

struct PreHme {
    uint64_t sad;
    uint16_t sa_w, sa_h;
    int16_t  mv_x, mv_y;
    uint8_t  valid;
};

constexpr int kMaxSadJobs = 32;
constexpr int kMaxNxmJobs = 16;
constexpr int kMaxFpJobs  = 8;

// SB-scoped part of the reference's MeContext (me_context.h:366-509).  The control part (everything up to `home`) is
// worked on in shared memory by every kernel of the staged pipeline and kept in HBM/L2 between them; the result tables
// behind it are reached through `home`.
struct alignas(16) SbState {
    uint64_t lsad[3][kMaxLists][kMaxRefs][2][2];  // hme_level{0,1,2}_sad[l][r][w][h]
    uint64_t hme_sad[kMaxLists][kMaxRefs];
    PreHme   prehme[kMaxLists][kMaxRefs][2];
    int16_t  lx[3][kMaxLists][kMaxRefs][2][2], ly[3][kMaxLists][kMaxRefs][2][2];
    int16_t  hme_sc_x[kMaxLists][kMaxRefs], hme_sc_y[kMaxLists][kMaxRefs];
    uint32_t divisor[kMaxLists][kMaxRefs];
    uint32_t zz_sad[kMaxLists][kMaxRefs];
    uint8_t  do_ref[kMaxLists][kMaxRefs];
    uint8_t  performed_phme[kMaxLists][kMaxRefs][2];
    uint8_t  tf_exit;  // temporal-filter mode: HME SAD under tf_me_exit_th, no integer search (motion_estimation.c:3109)
    SvtMeAreaMinMax hme_l0_sa;
    uint32_t org_x, org_y, b64_w, b64_h, b64_index;
    const uint8_t *src;  // b64_src_ptr
    int32_t  src_stride;
    // integer-search scratch per reference
    int16_t  isc_x[kMaxLists][kMaxRefs], isc_y[kMaxLists][kMaxRefs];    // search centre after check_00_center
    int16_t  isa_w[kMaxLists][kMaxRefs], isa_h[kMaxLists][kMaxRefs];    // search area
    int16_t  isa_w0[kMaxLists][kMaxRefs], isa_h0[kMaxLists][kMaxRefs];  // before SR reduction
    uint32_t c00_sad[kMaxLists][kMaxRefs];
    uint8_t  c00_done[kMaxLists][kMaxRefs];
    uint8_t  probe[kMaxLists][kMaxRefs];
    // job lists
    alignas(16) SadJob sad_jobs[kMaxSadJobs];  // fetched by the engines with 16-byte copies
    alignas(16) NxmJob nxm_jobs[kMaxNxmJobs];
    alignas(16) FpJob  fp_jobs[kMaxFpJobs];
    int32_t  n_sad, n_nxm, n_fp;
    unsigned long long px;  // algorithmic pixel-compares of this SB so far (staged pipeline)
    uint8_t  l0_done[kMaxLists][kMaxRefs];  // references whose HME-L0 quadrants were searched in the current pass
    uint32_t scratch[32];                    // lane-to-lane exchange of the control steps (me_control.cuh)
    uint64_t scratch64[32];
    // Everything above is the CONTROL part: the staged pipeline's control kernel works on a shared-memory copy of it
    // (kSbHeadBytes) and reaches the search results below through `home`, the block's own state in HBM/L2.
    SbState *home;
    alignas(16) uint32_t best_sad[kMaxLists][kMaxRefs][kPus];
    uint32_t best_mv[kMaxLists][kMaxRefs][kPus];
    uint32_t me_distortion[kPus];
};

}  // namespace svtme
