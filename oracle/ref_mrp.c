/*
 * ref_mrp.c -- TEST INFRASTRUCTURE ONLY.
 * The reference's MRP (multi-reference-picture) control table is a static
 * function of Globals/enc_handle.c.  Pull that file into this TU unmodified
 * (its other undefined symbols are satisfied by the auto-generated stubs, see
 * oracle/Makefile) and expose set_mrp_ctrl.  The mrp_level choice restates
 * enc_handle.c:4141-4176.
 */
#include "enc_handle.c"

void svtref_set_mrp(SequenceControlSet *scs, int enc_mode, int rtc_tune, int rate_control_vbr, int fast_decode) {
    uint8_t mrp_level;
    if (rtc_tune) {
        mrp_level = enc_mode <= ENC_M9 ? 10 : 11;
    } else if (enc_mode <= ENC_M2) {
        mrp_level = (!(scs->input_resolution <= INPUT_SIZE_360p_RANGE) && !(fast_decode <= 1)) ? 9 : 2;
    } else if (enc_mode <= ENC_M4) {
        mrp_level = (!(scs->input_resolution <= INPUT_SIZE_360p_RANGE) && !(fast_decode <= 1)) ? 9 : 5;
    } else if (!rate_control_vbr) {
        mrp_level = enc_mode <= ENC_M7 ? 9 : enc_mode <= ENC_M9 ? 10 : 0;
    } else {
        mrp_level = 12;
    }
    set_mrp_ctrl(scs, mrp_level);
}
