// me_launch.h -- host-callable launch wrappers of the device kernels (me_kernel.cu).
#pragma once
#include <cuda_runtime.h>
#include "me_types.cuh"

namespace svtme {
cudaError_t rtcd_set_attributes();  // per device, with the device current
cudaError_t launch_pad_plane(uint8_t *org, int stride, int w, int h, int pad, cudaStream_t st);
cudaError_t launch_decimate(const DevPic &pic, int pad_q, int pad_s, int sixteenth_from_full, cudaStream_t st);
cudaError_t launch_unpack16(const uint16_t *in, int in_stride, uint8_t *out, int out_stride, int w, int h, cudaStream_t st);
cudaError_t launch_downsample_generic(const uint8_t *in, int in_stride, int w, int h, uint8_t *out, int out_stride, int step,
                                      cudaStream_t st);
cudaError_t launch_rtcd_sad_loop(const SadJob &job, uint32_t *d_out, cudaStream_t st);
cudaError_t launch_rtcd_nxm(const NxmJob &job, uint32_t *d_out, cudaStream_t st);
cudaError_t launch_rtcd_fullpel(const uint8_t *src, int src_stride, const FpJob &job, int sub_sad, uint32_t *d_best_sad,
                                uint32_t *d_best_mv, cudaStream_t st);
cudaError_t launch_sad_peak(uint32_t *d_out, int grid, int iters, cudaStream_t st);
}  // namespace svtme
namespace svtme {
// staged (picture-wide kernels) execution of the step program; p.states must hold staged_state_bytes(n_sb) bytes
cudaError_t launch_me_picture_staged(const MeParams &p, const S16Maps &maps, const RowsMaps &rmaps, int sm_count, cudaStream_t st, int *n_launches);
cudaError_t staged_set_attributes();  // per device, with the device current
void        staged_print_phases();    // -DSVTME_PHASES builds only
size_t      staged_state_bytes(uint32_t n_sb);
size_t      staged_plan_offset(uint32_t n_sb);
size_t      staged_queue_offset(uint32_t n_sb);
constexpr int kQueueCounterBytes = 2 * 4 * kMaxSteps;  // [count, head] per program step
// TPL source-based search over the ME results in HBM (me_tpl.cu); the output arrays must be zeroed first
cudaError_t launch_tpl_src(const TplParams &p, cudaStream_t st);
// dynamic-GOP detector HME level 0 (me_dg.cu): picture-level sums accumulated by the kernel (zero them first)
struct DgSums {
    unsigned long long tot_dist;
    uint32_t           tot_cplx, tot_active;
    int32_t            sum_in_vectors;
    uint32_t           pad_;
    unsigned long long px_compares;
};
// picture-analysis histograms of the sixteenth plane (me_dg.cu): raw counts / sums per region [wi * 4 + hi] (zero them first)
struct StatAcc {
    uint32_t           hist[16][256];
    unsigned long long sum[16];
};
cudaError_t launch_hist16(const DevPlane &p16, int rw_n, int rh_n, int step, StatAcc *d_acc, cudaStream_t st);
cudaError_t launch_dg_hme(const DevPlane &src16, const DevPlane &ref16, int sa, uint32_t wb, uint32_t hb, uint32_t *d_sb_sad,
                          int16_t *d_sb_mv, DgSums *d_out, cudaStream_t st);
}  // namespace svtme
