// knobs.cc -- svtme_derive_knobs: host mirror of the reference's per-picture ME signal derivation.
//
// An integrated encoder does NOT need this: it copies its own derived MeContext.  Tests, the
// bench and stand-alone users do, so the preset tables are restated here from the reference:
//   svt_aom_sig_deriv_me            enc_mode_config.c:671-808
//   set_me_search_params            enc_mode_config.c:213-340
//   set_hme_search_params           enc_mode_config.c:136-208
//   prune / sr-adjust / mv-sa / 8x8-var / pre-HME level tables   enc_mode_config.c:341-583
//   HME level flags                 enc_mode_config.c:1608-1619
//   enable_me_8x8 / gm level        enc_mode_config.c:76-129, 1821-1903
//   MRP controls                    enc_handle.c:3347-3610, 4141-4176
//   max_cand / max_refs             pcs.c:91-96, pd_process.c:3503-3519, 3670-3695
//   input resolution class          sequence_control_set.c:113-131, definitions.h:2051-2057
// tests/test_knobs.py pins every field against the compiled reference over the preset grid.
#include <cstring>
#include <algorithm>
#include "../../include/svtme_cuda.h"

namespace {

struct Mrp {
    uint8_t sc_base0, sc_base1, sc_nb0, sc_nb1, base0, base1, nb0, nb1;
    uint8_t safe_limit_nref;
    uint32_t safe_limit_zz_th;
    uint8_t only_l_bwd;
};

// set_mrp_ctrl, enc_handle.c:3347-3596 (only the fields the ME path reads)
Mrp mrp_table(int level) {
    switch (level) {
    case 0: return {1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0};
    case 1: return {2, 2, 2, 2, 4, 3, 4, 3, 0, 0, 0};
    case 2:
    case 3: return {2, 2, 2, 2, 4, 3, 4, 3, 0, 0, 1};
    case 4:
    case 5: return {2, 2, 1, 1, 4, 3, 4, 3, 0, 0, 1};
    case 6: return {2, 2, 1, 1, 4, 3, 4, 3, 1, 60000, 1};
    case 7: return {2, 2, 1, 1, 3, 2, 3, 2, 1, 60000, 1};
    case 8:
    case 9: return {2, 2, 1, 1, 3, 2, 3, 2, 2, 0, 1};
    case 10: return {2, 2, 1, 1, 3, 2, 2, 2, 2, 0, 1};
    case 11: return {2, 2, 1, 1, 3, 2, 1, 1, 2, 0, 1};
    default: return {2, 2, 1, 1, 2, 2, 1, 1, 2, 0, 1};  // 12
    }
}

int input_resolution_class(uint32_t px) {
    if (px < 0x28500) return SVTME_INPUT_SIZE_240p;
    if (px < 0x4CE00) return SVTME_INPUT_SIZE_360p;
    if (px < 0xA1400) return SVTME_INPUT_SIZE_480p;
    if (px < 0x16DA00) return SVTME_INPUT_SIZE_720p;
    if (px < 0x535200) return SVTME_INPUT_SIZE_1080p;
    if (px < 0x140A000) return SVTME_INPUT_SIZE_4K;
    return SVTME_INPUT_SIZE_8K;
}

inline SvtMeArea A(uint16_t w, uint16_t h) { SvtMeArea a; a.width = w; a.height = h; return a; }
inline int clip3(int lo, int hi, int v) { return v < lo ? lo : (v > hi ? hi : v); }

// set_me_search_params, enc_mode_config.c:213-340
void me_search_params(const SvtMePresetCfg &c, int res, SvtMeKnobs &k) {
    const int m = c.enc_mode;
    int q_mult = 0;
    SvtMeArea &mn = k.me_sa.sa_min, &mx = k.me_sa.sa_max;
    if (c.rtc_tune) {
        if (c.sc_class1) {
            if (m <= 7) { mn = A(32, 32); mx = A(96, 96); }
            else if (res < SVTME_INPUT_SIZE_1080p) { mn = A(16, 16); mx = A(32, 16); }
            else if (m <= 9) { mn = A(24, 24); mx = A(24, 24); }
            else { mn = A(16, 6); mx = A(16, 9); }
        } else if (m <= 8) {
            if (res < SVTME_INPUT_SIZE_1080p) { mn = A(16, 16); mx = A(32, 16); }
            else { mn = A(16, 6); mx = A(16, 9); }
        } else if (res < SVTME_INPUT_SIZE_720p) { mn = A(8, 3); mx = A(16, 9); }
        else if (res < SVTME_INPUT_SIZE_1080p) { mn = A(8, 1); mx = A(16, 7); }
        else if (res < SVTME_INPUT_SIZE_4K) { mn = A(8, 1); mx = A(8, 7); }
        else { mn = A(8, 1); mx = A(8, 1); }
    } else if (c.sc_class1) {
        if (m <= 1) { mn = A(175, 175); mx = A(750, 750); }
        else if (m <= 6) { mn = A(48, 48); mx = A(224, 224); }
        else if (m <= 7) { mn = A(32, 32); mx = A(164, 164); }
        else if (m <= 8) { mn = A(32, 32); mx = A(96, 96); }
        else if (m <= 9) { mn = A(16, 16); mx = A(96, 96); }
        else { mn = A(8, 8); mx = A(32, 32); }
    } else if (m <= 1) { mn = A(64, 64); mx = A(256, 256); }
    else if (m <= 2) { mn = A(32, 32); mx = A(128, 128); }
    else if (m <= 4) { mn = A(24, 24); mx = A(104, 104); }
    else if (m <= 6) { mn = A(16, 16); mx = A(64, 32); q_mult = 7; }
    else if (m <= 9) {
        if (c.hierarchical_levels <= 3) {
            if (res < SVTME_INPUT_SIZE_4K) { mn = A(8, 5); mx = A(16, 9); }
            else { mn = A(8, 1); mx = A(8, 1); }
        } else if (res < SVTME_INPUT_SIZE_1080p) { mn = A(16, 16); mx = A(32, 16); }
        else { mn = A(16, 6); mx = A(16, 9); }
        q_mult = 7;
    } else { mn = A(16, 6); mx = A(16, 6); q_mult = 6; }
    if (q_mult) {
        const int qw = clip3(500, 1000, (int)(q_mult * ((31 * (int)c.qp) - 700)) >> 3);
        mn.width  = (uint16_t)std::max(8, (mn.width * qw) / 1000);
        mn.height = (uint16_t)std::max(3, (mn.height * qw) / 1000);
        mx.width  = (uint16_t)std::max(8, (mx.width * qw) / 1000);
        mx.height = (uint16_t)std::max(3, (mx.height * qw) / 1000);
    }
    if (c.frame_rate_q16 >> 16) {  // "low frame rate" flag is the integer fps (enc_mode_config.c:335)
        mn.width  = (uint16_t)((mn.width * 3) >> 1);
        mn.height = (uint16_t)((mn.height * 3) >> 1);
    }
}

// set_hme_search_params, enc_mode_config.c:136-208
void hme_search_params(const SvtMePresetCfg &c, int res, SvtMeKnobs &k) {
    const int m = c.enc_mode;
    k.num_hme_sa_w = 2; k.num_hme_sa_h = 2;
    int q_mult = 0;
    SvtMeArea &mn = k.hme_l0_sa.sa_min, &mx = k.hme_l0_sa.sa_max;
    if (m <= 1) {
        if (res < SVTME_INPUT_SIZE_4K) { mn = A(32, 32); mx = A(192, 192); }
        else { mn = A(240, 240); mx = A(480, 480); }
    } else if (m <= 3) { mn = A(32, 32); mx = A(192, 192); }
    else if (m <= 6) { mn = A(32, 32); mx = A(192, 192); q_mult = 3; }
    else if (!c.rtc_tune && m <= 7) {
        if (c.sc_class1 || res >= SVTME_INPUT_SIZE_4K) { mn = A(32, 32); mx = A(192, 192); }
        else { mn = A(16, 16); mx = A(192, 192); }
        q_mult = 3;
    } else if ((!c.rtc_tune && m <= 9) || (c.rtc_tune && m <= 7)) {
        if (c.sc_class1) { mn = A(32, 32); mx = A(192, 192); }
        else { mn = A(16, 16); mx = A(192, 192); }
        q_mult = 3;
    } else {
        if (c.sc_class1) { mn = A(32, 32); mx = A(192, 192); }
        else if (res < SVTME_INPUT_SIZE_4K) { mn = A(8, 8); mx = A(96, 96); }
        else { mn = A(16, 16); mx = A(96, 96); }
        q_mult = 3;
    }
    if (q_mult) {
        const int qw = clip3(500, 1000, (int)(q_mult * ((8 * (int)c.qp) - 125)));
        mn.width  = (uint16_t)std::max(8, (mn.width * qw) / 1000);
        mn.height = (uint16_t)std::max(8, (mn.height * qw) / 1000);
        mx.width  = (uint16_t)std::max(96, (mx.width * qw) / 1000);
        mx.height = (uint16_t)std::max(96, (mx.height * qw) / 1000);
    }
    if (m <= -1) { k.hme_l1_sa = A(16, 16); k.hme_l2_sa = A(16, 16); }
    else { k.hme_l1_sa = A(8, 3); k.hme_l2_sa = A(8, 3); }
}

}  // namespace

namespace {

// svt_aom_set_me_hme_ref_prune_ctrls, enc_mode_config.c:341-420
void prune_ctrls(int level, SvtMeKnobs &k) {
    static const uint16_t hme_th[7] = {0xffff, 80, 50, 30, 15, 5, 5};
    k.enable_me_hme_ref_pruning               = level ? 1 : 0;
    k.prune_ref_if_hme_sad_dev_bigger_than_th = hme_th[level];
    k.prune_ref_if_me_sad_dev_bigger_than_th  = level >= 4 ? 60 : 0xffff;
    k.zz_sad_th    = level == 6 ? 20 * 64 * 64 : 0;
    k.zz_sad_pct   = level == 6 ? 5 : 0;
    k.phme_sad_th  = level == 6 ? 10 * 64 * 64 : 0;
    k.phme_sad_pct = level == 6 ? 5 : 0;
}

// svt_aom_set_me_sr_adjustment_ctrls, enc_mode_config.c:444-520 (needs the HME level flags set first)
void sr_adjust_ctrls(int level, SvtMeKnobs &k) {
    if (level == 0) {
        k.enable_me_sr_adjustment = 0;
    } else {
        static const uint16_t mvth[6] = {0, 4, 4, 4, 16, 20};
        static const uint16_t stat[6] = {0, 12000, 12000, 12000, 20000, 24000};
        static const uint16_t low[6]  = {0, 6000, 6000, 12000, 20000, 24000};
        k.enable_me_sr_adjustment              = level >= 4 ? 2 : 1;
        k.reduce_me_sr_based_on_mv_length_th   = mvth[level];
        k.stationary_hme_sad_abs_th            = stat[level];
        k.stationary_me_sr_divisor             = 8;
        k.reduce_me_sr_based_on_hme_sad_abs_th = low[level];
        k.me_sr_divisor_for_low_hme_sad        = 8;
        k.distance_based_hme_resizing          = level >= 2 ? 1 : 0;
    }
    if (k.enable_hme_level2_flag == 0) {  // thresholds scale with the resolution HME stopped at (:507-519)
        const int d = k.enable_hme_level1_flag == 1 ? 4 : 16;
        k.stationary_hme_sad_abs_th            = (uint16_t)(k.stationary_hme_sad_abs_th / d);
        k.reduce_me_sr_based_on_hme_sad_abs_th = (uint16_t)(k.reduce_me_sr_based_on_hme_sad_abs_th / d);
    }
}

// svt_aom_set_prehme_ctrls, enc_mode_config.c:539-583
void prehme_ctrls(int level, SvtMeKnobs &k) {
    k.prehme_enable = level ? 1 : 0;
    if (!level) return;
    static const uint16_t v_min_h[5] = {0, 144, 100, 100, 100}, v_max_h[5] = {0, 496, 400, 350, 350};
    static const uint16_t h_min_w[5] = {0, 144, 96, 32, 32}, h_max_w[5] = {0, 496, 384, 200, 128};
    const uint16_t hh = level >= 3 ? 7 : 3;
    k.prehme_sa_cfg[0].sa_min = A(8, v_min_h[level]);
    k.prehme_sa_cfg[0].sa_max = A(8, v_max_h[level]);
    k.prehme_sa_cfg[1].sa_min = A(h_min_w[level], hh);
    k.prehme_sa_cfg[1].sa_max = A(h_max_w[level], hh);
    k.prehme_skip_search_line = level >= 3 ? 1 : 0;
    k.prehme_l1_early_exit    = level == 4 ? 1 : 0;
}

}  // namespace

extern "C" int32_t svtme_derive_knobs(const SvtMePresetCfg *cfg, SvtMeKnobs *out) {
    if (!cfg || !out || cfg->enc_mode < -1 || cfg->enc_mode > 13 || cfg->width == 0 || cfg->height == 0)
        return SVTME_ERR_BAD_PARAMETER;
    const SvtMePresetCfg &c = *cfg;
    SvtMeKnobs           &k = *out;
    memset(&k, 0, sizeof(k));
    const int  m       = c.enc_mode;
    const int  res     = input_resolution_class((uint32_t)c.width * c.height);
    const bool is_base = c.temporal_layer_index == 0;

    // MRP level, enc_handle.c:4141-4176; low-delay drops list 1 (:3597-3609)
    int mrp_level;
    if (c.rtc_tune) mrp_level = m <= 9 ? 10 : 11;
    else if (m <= 2) mrp_level = (!(res <= SVTME_INPUT_SIZE_360p) && !(c.fast_decode <= 1)) ? 9 : 2;
    else if (m <= 4) mrp_level = (!(res <= SVTME_INPUT_SIZE_360p) && !(c.fast_decode <= 1)) ? 9 : 5;
    else if (!c.rate_control_vbr) mrp_level = m <= 7 ? 9 : m <= 9 ? 10 : 0;
    else mrp_level = 12;
    Mrp mrp = mrp_table(mrp_level);
    if (c.rtc_tune) { mrp.sc_base1 = mrp.sc_nb1 = mrp.base1 = mrp.nb1 = 0; mrp.safe_limit_nref = 0; mrp.only_l_bwd = 0; }

    me_search_params(c, res, k);
    hme_search_params(c, res, k);
    // HME flags, enc_mode_config.c:1608-1619
    k.enable_hme_flag = k.enable_hme_level0_flag = k.enable_hme_level1_flag = 1;
    k.enable_hme_level2_flag = (c.sc_class1 || m <= 6) ? 1 : 0;
    k.hme_search_method = 0;  // SUB_SAD_SEARCH (:689-690)
    k.me_search_method  = 0;
    if (c.rtc_tune) {
        if (c.sc_class1) { k.reduce_hme_l0_sr_th_min = 8; k.reduce_hme_l0_sr_th_max = 100; }
        else if (m > 7) { k.reduce_hme_l0_sr_th_min = 8; k.reduce_hme_l0_sr_th_max = 200; }
    }
    int prehme_level;
    if (c.sc_class1) prehme_level = c.rtc_tune ? 1 : 2;
    else if (c.rtc_tune) prehme_level = m <= 9 ? 4 : 0;
    else prehme_level = m <= 7 ? 2 : 4;
    if (!k.enable_hme_level1_flag) prehme_level = 0;
    prehme_ctrls(prehme_level, k);

    int prune_level;
    if (c.sc_class1) prune_level = m <= 2 ? 1 : m <= 7 ? 3 : 6;
    else if (m <= 0) prune_level = is_base ? 1 : 2;
    else if (m <= 1) prune_level = is_base ? 1 : 4;
    else if (m <= 3) prune_level = is_base ? 1 : 5;
    else if (m <= 9) prune_level = is_base ? 1 : 6;
    else prune_level = 6;
    prune_ctrls(prune_level, k);

    int sr_level;
    if (c.sc_class1) sr_level = m <= 7 ? 4 : 5;
    else if (m <= -1) sr_level = 0;
    else if (m <= 0) sr_level = 1;
    else sr_level = 3;
    sr_adjust_ctrls(sr_level, k);

    // svt_aom_set_mv_based_sa_ctrls level 2 / 0 (:422-442)
    if (m <= 3) { k.mv_sa_adj_enabled = 1; k.mv_sa_adj_nearest_ref_only = 1; k.mv_sa_adj_mv_size_th = 25; k.mv_sa_adj_sa_multiplier = 2; }
    // svt_aom_set_me_8x8_var_ctrls level 2 (:521-538)
    k.me_8x8_var_enabled = 1; k.me_sr_div4_th = 80000; k.me_sr_div2_th = 150000; k.me_sr_mult2_th = 0xffffffffu;
    k.prune_me_candidates_th     = m <= 6 ? 0 : 65;
    k.use_best_unipred_cand_only = m <= 3 ? 0 : 1;  // pcs->use_best_me_unipred_cand_only, enc_mode_config.c:1802-1805
    if (c.rtc_tune) k.me_early_exit_th = c.sc_class1 ? 64 * 64 : (m <= 9 ? 64 * 64 * 8 : 64 * 64 * 9);
    else k.me_early_exit_th = m <= 4 ? 0 : 64 * 64 * 8;
    k.me_safe_limit_zz_th         = mrp.safe_limit_nref == 1 ? mrp.safe_limit_zz_th : 0;
    k.prev_me_stage_based_exit_th = (c.rtc_tune && c.sc_class1) ? 64 * 64 * 4 : 0;

    // result geometry: pd_process.c:3503-3519, pcs.c:91-96
    const uint8_t l0 = std::max(std::max(mrp.sc_base0, mrp.base0), std::max(mrp.sc_nb0, mrp.nb0));
    const uint8_t l1 = std::max(std::max(mrp.sc_base1, mrp.base1), std::max(mrp.sc_nb1, mrp.nb1));
    k.max_refs = (uint8_t)(l0 + l1);
    k.max_cand = (uint8_t)(l0 + l1 + l0 * l1 + (l0 - 1) + (l1 == 3 ? 1 : 0));
    k.max_l0   = l0;
    // reference counts of this picture: update_count_try, pd_process.c:3670-3695
    uint8_t c0, c1;
    if (c.sc_class1) { c0 = is_base ? mrp.sc_base0 : mrp.sc_nb0; c1 = is_base ? mrp.sc_base1 : mrp.sc_nb1; }
    else { c0 = is_base ? mrp.base0 : mrp.nb0; c1 = is_base ? mrp.base1 : mrp.nb1; }
    if (c.ref_list0_count_try) c0 = std::min(c0, c.ref_list0_count_try);
    if (c.ref_list1_count_try) c1 = std::min(c1, c.ref_list1_count_try);
    k.num_of_list_to_search       = c.slice_type_b ? 2 : 1;  // me_process.c:219-221
    k.num_of_ref_pic_to_search[0] = c0;
    k.num_of_ref_pic_to_search[1] = c.slice_type_b ? c1 : 0;
    k.temporal_layer_index    = c.temporal_layer_index;
    k.is_ref                  = c.is_ref;
    k.picture_number          = c.picture_number;
    k.hierarchical_levels     = c.hierarchical_levels;
    k.similar_brightness_refs = c.similar_brightness_refs;
    k.input_resolution        = (uint8_t)res;
    // enc_mode_config.c:76-101 / pcs.c:1389-1394
    k.enable_me_16x16 = 1;
    if (c.rtc_tune) k.enable_me_8x8 = m <= 7;
    else k.enable_me_8x8 = m <= 5 ? 1 : (m <= 8 ? (res <= SVTME_INPUT_SIZE_720p) : 0);
    k.max_number_of_pus_per_sb = SVTME_SQUARE_PU_COUNT;
    // gm level (super-res off): MR -> 2, <= M4 -> 3, else off (:103-116); B/P slices only
    k.gm_enabled = m <= 4 ? 1 : 0;
    k.gm_use_distance_based_active_th = 0;
    k.only_l_bwd = mrp.only_l_bwd;
    return SVTME_OK;
}

// Temporal-filter knobs: svt_aom_sig_deriv_me_tf (enc_mode_config.c:814-854), tf_set_me_hme_params_oq (:588-663),
// the TF HME level flags (:1620-1643) and the per-call overrides of produce_temporally_filtered_pic
// (temporal_filtering.c:296, 316, 2759-2767, 3140-3158).
extern "C" int32_t svtme_derive_tf_knobs(const SvtMeTfCfg *cfg, SvtMeKnobs *out) {
    if (!cfg || !out || cfg->hme_me_level > 4 || cfg->width == 0 || cfg->height == 0) return SVTME_ERR_BAD_PARAMETER;
    SvtMeKnobs &k = *out;
    memset(&k, 0, sizeof(k));
    const int lvl = cfg->hme_me_level;
    const int res = input_resolution_class((uint32_t)cfg->width * cfg->height);
    k.num_hme_sa_w = 2;
    k.num_hme_sa_h = 2;
    switch (lvl) {
    case 0:
        k.hme_l0_sa.sa_min = A(30, 30); k.hme_l0_sa.sa_max = A(60, 60);
        k.hme_l1_sa = A(16, 16); k.hme_l2_sa = A(16, 16);
        k.me_sa.sa_min = A(60, 60); k.me_sa.sa_max = A(120, 120);
        break;
    case 1:
        k.hme_l0_sa.sa_min = A(16, 16); k.hme_l0_sa.sa_max = A(32, 32);
        k.hme_l1_sa = A(16, 16); k.hme_l2_sa = A(16, 16);
        k.me_sa.sa_min = A(16, 16); k.me_sa.sa_max = A(32, 32);
        break;
    case 2:
        if (res <= SVTME_INPUT_SIZE_360p) { k.hme_l0_sa.sa_min = A(8, 8); k.hme_l0_sa.sa_max = A(8, 8); k.hme_l1_sa = A(8, 8); }
        else if (res <= SVTME_INPUT_SIZE_480p) { k.hme_l0_sa.sa_min = A(8, 8); k.hme_l0_sa.sa_max = A(16, 16); k.hme_l1_sa = A(8, 8); }
        else { k.hme_l0_sa.sa_min = A(16, 16); k.hme_l0_sa.sa_max = A(32, 32); k.hme_l1_sa = A(16, 16); }
        k.hme_l2_sa = A(16, 16);
        k.me_sa.sa_min = A(8, 8); k.me_sa.sa_max = A(8, 8);
        break;
    case 3:
        k.hme_l0_sa.sa_min = A(8, 8); k.hme_l0_sa.sa_max = A(8, 8);
        k.hme_l1_sa = A(8, 8); k.hme_l2_sa = A(8, 8);
        k.me_sa.sa_min = A(8, 8); k.me_sa.sa_max = A(8, 8);
        break;
    default:
        k.hme_l0_sa.sa_min = A(4, 4); k.hme_l0_sa.sa_max = A(4, 4);
        k.hme_l1_sa = A(8, 8); k.hme_l2_sa = A(8, 8);
        k.me_sa.sa_min = A(8, 8); k.me_sa.sa_max = A(8, 8);
        break;
    }
    if (cfg->qp_opt) {  // :656-663
        const int q = clip3(250, 1000, (int)(8 * cfg->qp) - 125);
        auto mod = [q](uint16_t v) { const int r = (v * q) / 1000; return (uint16_t)(r > 8 ? r : 8); };
        k.me_sa.sa_min.width = mod(k.me_sa.sa_min.width); k.me_sa.sa_min.height = mod(k.me_sa.sa_min.height);
        k.me_sa.sa_max.width = mod(k.me_sa.sa_max.width); k.me_sa.sa_max.height = mod(k.me_sa.sa_max.height);
    }
    k.enable_hme_flag        = 1;
    k.enable_hme_level0_flag = 1;
    k.enable_hme_level1_flag = lvl <= 2 ? 1 : 0;
    k.enable_hme_level2_flag = lvl == 0 ? 1 : 0;
    k.hme_search_method = 1;                 // FULL_SAD_SEARCH, forced per block (temporal_filtering.c:316)
    k.me_search_method  = lvl <= 2 ? 1 : 0;  // :822-834
    prehme_ctrls(0, k);
    prune_ctrls(0, k);
    sr_adjust_ctrls(0, k);
    // mv-based SA adjustment and 8x8-variance control are level 0 = off (:844-846)
    k.me_early_exit_th            = lvl <= 1 ? 0 : 64 * 64 * 4;
    k.prev_me_stage_based_exit_th = lvl <= 1 ? 0 : 64 * 64 * 4;
    k.me_type       = SVTME_ME_MCTF;
    k.tf_me_exit_th = cfg->me_exit_th;
    k.num_of_list_to_search       = 1;
    k.num_of_ref_pic_to_search[0] = 1;
    k.temporal_layer_index        = cfg->temporal_layer_index;
    k.is_ref                      = cfg->is_ref;
    k.picture_number              = cfg->picture_number;
    k.input_resolution            = (uint8_t)res;
    // only read by the candidate / statistics stage, which this mode skips; kept at the encoder's constants
    k.enable_me_8x8 = 1; k.enable_me_16x16 = 1; k.max_number_of_pus_per_sb = SVTME_SQUARE_PU_COUNT;
    k.max_cand = 1; k.max_refs = 1; k.max_l0 = 1;
    return SVTME_OK;
}
