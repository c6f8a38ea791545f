// me_dg.cu -- dynamic-GOP detector, HME level 0.
//
// Reference: dg_detector_hme_level0 (pd_process.c:493-590) with early_hme_b64 (:392-491), run by the ME kernel
// thread for TASK_DG_DETECTOR_HME messages (me_process.c:327-329): for every 64x64 block of the source picture, its
// 16x16 block of the sixteenth plane is searched with FULL SAD over a 16 / 64 / 128-square window of the reference
// picture's sixteenth plane (svt_sad_loop_kernel, compute_sad_c.c:58-101: first candidate in raster order wins on
// strict '<'), and four picture-level sums (DGDetectorMetrics, pcs.h:731-737) are accumulated under a mutex.
//
// Here: one kernel over the picture, one CTA per 64x64 block.  Unlike the per-SB ME searches this one is large and
// regular (128 x 128 candidates x 256 pixels = 4.2 M pixel-compares per block), so the kernel is built around the
// VABSDIFF4 issue rate alone:
//   * the window (<= 143 x 143 bytes) is staged once, byte-aligned on the way in (candidate 0 at byte 0 of a word);
//   * an item = 4 x-adjacent candidates x R candidate rows.  The R rows share R + 15 window rows: each window row is
//     loaded once (5 LDS), shifted once (12 SHF) and used by up to R block rows -> 64 R SADs per 5 (R + 15) LDS and
//     12 (R + 15) SHF; R = 8: 2048 SADs, 115 LDS, 276 SHF (SAD = 84 % of the ALU-pipe instructions);
//   * the 16x16 source block lives in 64 registers, the R x 4 accumulators in 32;
//   * minima are kept as (sad << 14 | raster index): the reference's tie-break independent of the item order.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include "me_engines.cuh"
#include "me_launch.h"

namespace svtme {

constexpr int kDgThreads = 128;
constexpr int kDgPitch   = 36;            // words per staged row: 32 groups of 4 candidates + the 4 words of the block width
constexpr int kDgStrip   = 8;             // candidate rows per item
constexpr int kDgRows    = 128 + 15 + kDgStrip;

__global__ void __launch_bounds__(kDgThreads, 4)
dg_hme_kernel(const DevPlane sp, const DevPlane rp, int sa0, uint32_t wb, uint32_t hb, uint32_t *sb_sad, int16_t *sb_mv, DgSums *out) {
    extern __shared__ __align__(16) uint32_t dg_win[];  // kDgRows x kDgPitch
    __shared__ uint32_t wbest[kDgThreads / 32];
    const uint32_t xb = blockIdx.x % wb, yb = blockIdx.x / wb;
    // early_hme_b64 (pd_process.c:406-452): the window is centred on the block; left / top clipping moves it without
    // shrinking it (:415-418, 437-440), right / bottom clipping shrinks it, the width is then rounded down to 8
    const int org_x = (int)(xb * 16), org_y = (int)(yb * 16);
    int       sa_w = (sa0 + 7) & ~7, sa_h = sa0;
    const int pad_w = rp.ref_org - 1, pad_h = rp.ref_org - 1;
    int       ox = -(sa_w >> 1), oy = -(sa_h >> 1);
    if (org_x + ox < -pad_w) ox = -pad_w - org_x;
    if (org_x + ox > rp.width - 1) ox -= (org_x + ox) - (rp.width - 1);
    if (org_x + ox + sa_w > rp.width) sa_w = max(1, sa_w - ((org_x + ox + sa_w) - rp.width));
    sa_w = sa_w < 8 ? sa_w : sa_w & ~7;
    if (org_y + oy < -pad_h) oy = -pad_h - org_y;
    if (org_y + oy > rp.height - 1) oy -= (org_y + oy) - (rp.height - 1);
    if (org_y + oy + sa_h > rp.height) sa_h = max(1, sa_h - ((org_y + oy + sa_h) - rp.height));

    // the 16x16 source block (always 16x16: blocks cut by the picture edge read the padding, :524-536)
    uint32_t       sreg[16][4];
    const uint8_t *b0 = sp.org + (int64_t)org_y * sp.stride + org_x;
#pragma unroll
    for (int r = 0; r < 16; r++) {
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(b0 + (int64_t)r * sp.stride));
        sreg[r][0] = v.x; sreg[r][1] = v.y; sreg[r][2] = v.z; sreg[r][3] = v.w;
    }
    // stage the window, byte-aligned: word (r, c) = window bytes 4c .. 4c + 3 of row r
    const int ng = (sa_w + 3) >> 2, nwords = ng + 4, nrows = sa_h + 15;
    {
        const uint8_t *g  = rp.org + (int64_t)(org_y + oy) * rp.stride + (org_x + ox);
        const int      a  = (int)((uintptr_t)g & 3);
        const uint8_t *ga = g - a;
        // flat over (row, word) with the constant pitch as the divisor (a full window has nwords == kDgPitch)
        for (int idx = threadIdx.x; idx < nrows * kDgPitch; idx += kDgThreads) {
            const int r = idx / kDgPitch, c = idx - r * kDgPitch;
            if (c < nwords) {
                const uint32_t *q = reinterpret_cast<const uint32_t *>(ga + (int64_t)r * rp.stride) + c;
#ifdef SVTME_BOUNDS  // diagnostics build: both words must lie inside the plane's physical padding (32 pixels at this level)
                {
                    const int64_t off = reinterpret_cast<const uint8_t *>(q) - (rp.org - 32 * (int64_t)rp.stride - 32);  // from the padded origin
                    const int64_t y = off / rp.stride - 32, x = off % rp.stride - 32;
                    if (off < 0 || y >= rp.height + 32 || x + 8 > rp.width + 32) {
                        printf("svtme bounds: dg window word at (%lld, %lld) outside the padded %d x %d plane\n", (long long)x, (long long)y, rp.width, rp.height);
                        __trap();
                    }
                }
#endif
                dg_win[idx] = __funnelshift_r(__ldg(q), __ldg(q + 1), 8 * a);
            }
        }
    }
    __syncthreads();
    const int nstrips = (sa_h + kDgStrip - 1) / kDgStrip, items = ng * nstrips;
    uint32_t  best = ~0u;
    for (int it = threadIdx.x; it < items; it += kDgThreads) {
        const int       s = it / ng, g = it - s * ng;
        const uint32_t *rowp = dg_win + (s * kDgStrip) * kDgPitch + g;
        uint32_t        acc[kDgStrip][4];
#pragma unroll
        for (int y2 = 0; y2 < kDgStrip; y2++) acc[y2][0] = acc[y2][1] = acc[y2][2] = acc[y2][3] = 0;
#pragma unroll
        for (int rr = 0; rr < kDgStrip + 15; rr++) {
            uint32_t w[5];
#pragma unroll
            for (int c = 0; c < 5; c++) w[c] = rowp[rr * kDgPitch + c];
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const uint32_t s0 = w[c], s1 = fsr(w[c], w[c + 1], 1), s2 = fsr(w[c], w[c + 1], 2), s3 = fsr(w[c], w[c + 1], 3);
#pragma unroll
                for (int y2 = 0; y2 < kDgStrip; y2++) {
                    const int i = rr - y2;  // block row that meets window row rr for candidate row y2 of the strip
                    if (i >= 0 && i < 16) {
                        const uint32_t sw = sreg[i][c];
                        acc[y2][0] = sad4(s0, sw, acc[y2][0]); acc[y2][1] = sad4(s1, sw, acc[y2][1]);
                        acc[y2][2] = sad4(s2, sw, acc[y2][2]); acc[y2][3] = sad4(s3, sw, acc[y2][3]);
                    }
                }
            }
        }
        // keys (sad << 14 | raster index); items that lie wholly inside the window (all of them for an interior block)
        // skip the per-candidate bounds tests
        const int ybase = s * kDgStrip, xbase = 4 * g;
        if (ybase + kDgStrip <= sa_h && xbase + 4 <= sa_w) {
#pragma unroll
            for (int y2 = 0; y2 < kDgStrip; y2++) {
                const uint32_t o = (uint32_t)((ybase + y2) * sa_w + xbase);
                const uint32_t k0 = acc[y2][0] * 16384u + o, k1 = acc[y2][1] * 16384u + (o + 1), k2 = acc[y2][2] * 16384u + (o + 2),
                               k3 = acc[y2][3] * 16384u + (o + 3);
                best = min(min(best, k0), min(min(k1, k2), k3));
            }
        } else {
#pragma unroll
            for (int y2 = 0; y2 < kDgStrip; y2++) {
                const int y = ybase + y2;
#pragma unroll
                for (int k = 0; k < 4; k++)
                    if (y < sa_h && xbase + k < sa_w) best = min(best, (acc[y2][k] << 14) | (uint32_t)(y * sa_w + xbase + k));
            }
        }
    }
    best = __reduce_min_sync(0xffffffffu, best);
    if ((threadIdx.x & 31) == 0) wbest[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t b = wbest[0];
#pragma unroll
        for (int i = 1; i < kDgThreads / 32; i++) b = min(b, wbest[i]);
        const uint32_t sad = b >> 14, idx = b & 0x3fffu;
        const int      y = (int)idx / sa_w, x = (int)idx - y * sa_w;
        const int16_t  col = (int16_t)((x + ox) * 4), row = (int16_t)((y + oy) * 4);  // :484-487
        if (sb_sad) sb_sad[blockIdx.x] = sad;
        if (sb_mv) { sb_mv[2 * blockIdx.x] = col; sb_mv[2 * blockIdx.x + 1] = row; }
        // the picture-level metrics (:541-581)
        int in = 0;
        if (yb < hb / 2) in += row > 0 ? -1 : row < 0 ? 1 : 0;
        else if (yb > hb / 2) in += row > 0 ? 1 : row < 0 ? -1 : 0;
        if (xb < wb / 2) in += col > 0 ? -1 : col < 0 ? 1 : 0;
        else if (xb > wb / 2) in += col > 0 ? 1 : col < 0 ? -1 : 0;
        atomicAdd(&out->tot_dist, (unsigned long long)sad);
        if (sad > 16 * 16 * 30) atomicAdd(&out->tot_cplx, 1u);
        if (col != 0 || row != 0) atomicAdd(&out->tot_active, 1u);
        if (in) atomicAdd(&out->sum_in_vectors, in);
        atomicAdd(&out->px_compares, (unsigned long long)sa_w * (unsigned long long)sa_h * 256ull);
    }
}

cudaError_t launch_dg_hme(const DevPlane &src16, const DevPlane &ref16, int sa, uint32_t wb, uint32_t hb, uint32_t *d_sb_sad,
                          int16_t *d_sb_mv, DgSums *d_out, cudaStream_t st) {
    dg_hme_kernel<<<wb * hb, kDgThreads, kDgRows * kDgPitch * 4, st>>>(src16, ref16, sa, wb, hb, d_sb_sad, d_sb_mv, d_out);
    return cudaGetLastError();
}

}  // namespace svtme

// ---- picture-analysis statistics on the sixteenth plane (second half) ----
// sub_sample_luma_generate_pixel_intensity_histogram_bins (pic_analysis_process.c:1461-1524) with calculate_histogram
// (:166-184): per region a 256-bin count of every step-th pixel of every step-th row, and their sum.  One CTA per
// (region, row slice): counts in shared memory, one global atomic per non-empty bin at the end.  The host applies the
// reference's "+1, x 16 step^2" scaling and the rounded averages (svtme_picture_statistics, abi.cu).  HBM-bound and tiny
// (0.5 MB at 4K).
namespace svtme {

__global__ void __launch_bounds__(256) hist16_kernel(const DevPlane p, int rw_n, int rh_n, int step, StatAcc *acc) {
    __shared__ uint32_t           h[256];
    __shared__ unsigned long long ssum;
    const int wi = blockIdx.x / rh_n, hi = blockIdx.x - wi * rh_n;
    const int region_w = p.width / rw_n, region_h = p.height / rh_n;
    const int w = region_w + (wi == rw_n - 1 ? p.width - rw_n * region_w : 0);
    const int hh = region_h + (hi == rh_n - 1 ? p.height - rh_n * region_h : 0);
    h[threadIdx.x] = 0;
    if (threadIdx.x == 0) ssum = 0;
    __syncthreads();
    const uint8_t *s = p.org + (int64_t)(hi * region_h) * p.stride + wi * region_w;
    const int      nsx = (w + step - 1) / step, nsy = (hh + step - 1) / step;
    uint32_t       sum = 0;  // <= 255 x samples of this thread
    for (int sy = blockIdx.y; sy < nsy; sy += gridDim.y) {
        const uint8_t *row = s + (int64_t)(sy * step) * p.stride;
        for (int sx = threadIdx.x; sx < nsx; sx += blockDim.x) {
            const uint32_t v = row[sx * step];
            atomicAdd(&h[v], 1u);
            sum += v;
        }
    }
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((threadIdx.x & 31) == 0 && sum) atomicAdd(&ssum, (unsigned long long)sum);
    __syncthreads();
    const int r = wi * 4 + hi;
    if (h[threadIdx.x]) atomicAdd(&acc->hist[r][threadIdx.x], h[threadIdx.x]);
    if (threadIdx.x == 0 && ssum) atomicAdd(&acc->sum[r], ssum);
}

cudaError_t launch_hist16(const DevPlane &p16, int rw_n, int rh_n, int step, StatAcc *d_acc, cudaStream_t st) {
    const int nsy = (p16.height / rh_n + step - 1) / step;
    dim3      grid((unsigned)(rw_n * rh_n), (unsigned)max(1, min(32, nsy / 4)));
    hist16_kernel<<<grid, 256, 0, st>>>(p16, rw_n, rh_n, step, d_acc);
    return cudaGetLastError();
}

}  // namespace svtme
