/* ref_harness.h -- TEST INFRASTRUCTURE ONLY: C interface of oracle/_ref/libsvtref.so (the compiled reference). */
#ifndef SVTREF_HARNESS_H
#define SVTREF_HARNESS_H
#include <stdint.h>
#include "../include/svtme_cuda.h"
#include "planes.h"
#ifdef __cplusplus
extern "C" {
#endif
int      svtref_set_isa(int isa); /* 0 = C kernels, 1 = AVX2 twins (if built and supported); returns what was bound */
int      svtref_get_isa(void);
int      svtref_derive_knobs(const SvtMePresetCfg *cfg, SvtMeKnobs *out);
int      svtref_derive_tf_knobs(const SvtMeTfCfg *cfg, SvtMeKnobs *out);
int      svtref_me_picture(const SvtMeKnobs *k, const SvtRefPic *cur, const SvtRefPic *refs /*[2][4]*/,
                           SvtMePictureResults *out, int n_threads);
void     svtref_build_planes(const SvtRefPic *pic, int hme_level1);
void     svtref_sad_loop_kernel(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                                uint32_t block_height, uint32_t block_width, uint64_t *best_sad, int16_t *x, int16_t *y,
                                uint32_t src_stride_raw, uint8_t skip_search_line, int16_t sa_w, int16_t sa_h);
uint32_t svtref_nxm_sad_kernel(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                               uint32_t height, uint32_t width);
void     svtref_downsample_2d(const uint8_t *in, uint32_t in_stride, uint32_t w, uint32_t h, uint8_t *out,
                              uint32_t out_stride, uint32_t step);
void     svtref_generate_padding(uint8_t *pic, uint32_t stride, uint32_t w, uint32_t h, uint32_t pad_w, uint32_t pad_h);
void     svtref_fullpel_search_sblock(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                                      int16_t x_org, int16_t y_org, uint32_t sa_w, uint32_t sa_h, uint8_t sub_sad,
                                      uint32_t *best_sad, uint32_t *best_mv);
/* the reference's dg_detector_hme_level0 (pd_process.c:493) over the whole picture */
int      svtref_dg_hme_level0(const SvtRefPic *src, const SvtRefPic *ref, uint32_t input_resolution, SvtMeDgMetrics *m);
/* the reference's svt_aom_gathering_picture_statistics (pic_analysis_process.c:1559), histogram part */
int      svtref_picture_statistics(const SvtRefPic *pic, uint32_t regions_w, uint32_t regions_h, uint32_t decim_step, SvtMePicStats *out);
#ifdef __cplusplus
}
#endif
#endif
