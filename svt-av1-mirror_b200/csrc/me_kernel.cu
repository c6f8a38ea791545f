// me_kernel.cu -- device kernels of libsvtme_cuda.so (sm_100a) and their launch wrappers.
//
//   pad_plane_kernel       svt_aom_generate_padding (pic_operators.c:338) on the full-resolution plane
//   decimate_kernel        svt_aom_downsample_filtering_input_picture (pic_analysis_process.c:1945-2002):
//                          quarter + sixteenth planes and their padding in one pass over the full plane
//   unpack16_kernel        8-bit MSB plane of a 10-bit input (pack_unpack_c.c:77-117)
//   rtcd_*_kernel          single-CTA wrappers that run one engine job (RTCD-level entry points)
#include <cuda_runtime.h>
#include "me_engines.cuh"
#include "me_launch.h"

namespace svtme {

struct alignas(16) CtaShared {
    SbState      sb;
    EngineShared eng;
};
constexpr size_t kSmemBytes = ((sizeof(CtaShared) + 15) & ~size_t(15)) + kWinBytes;

__device__ __forceinline__ CudaBackend make_backend(const MeParams *p, unsigned char *smem) {
    CtaShared  *cs = reinterpret_cast<CtaShared *>(smem);
    CudaBackend be;
    be.p   = p;
    be.sh  = &cs->eng;
    be.win = smem + ((sizeof(CtaShared) + 15) & ~size_t(15));
    return be;
}

// ---- planes ----
__global__ void pad_plane_kernel(uint8_t *org, int stride, int w, int h, int pad) {
    // one thread per 16-byte vector of the padded domain; vectors entirely inside the picture are left alone
    const int vec_per_row = (w + 2 * pad + 15) >> 4;  // the row pitch leaves room for the last partial vector
    const int total       = vec_per_row * (h + 2 * pad);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int y = i / vec_per_row - pad, x = (i % vec_per_row) * 16 - pad;
        if (y >= 0 && y < h && x >= 0 && x + 16 <= w) continue;
        const int      cy  = min(max(y, 0), h - 1);
        const uint8_t *src = org + (int64_t)cy * stride;
        uint32_t       v[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            uint32_t a = 0;
#pragma unroll
            for (int q = 0; q < 4; q++) a |= (uint32_t)src[min(max(x + 4 * k + q, 0), w - 1)] << (8 * q);
            v[k] = a;
        }
        // interior bytes of a straddling vector are rewritten with their own values
        *reinterpret_cast<uint4 *>(org + (int64_t)y * stride + x) = make_uint4(v[0], v[1], v[2], v[3]);
    }
}

__device__ __forceinline__ uint32_t quarter_px(const uint8_t *f, int stride, int xq, int yq) {
    const uint8_t *r0 = f + (int64_t)(2 * yq) * stride + 2 * xq;
    return ((uint32_t)r0[0] + r0[1] + r0[stride] + r0[stride + 1] + 2) >> 2;
}

// Thread = one pixel of the padded sixteenth plane; it also writes the 2x2 quarter pixels it covers.
__global__ void decimate_kernel(DevPic pic, int pad_q, int pad_s, int sixteenth_from_full) {
    const DevPlane &F = pic.lvl[0], &Q = pic.lvl[1], &S = pic.lvl[2];
    const int dom_w = S.width + 2 * pad_s, dom_h = S.height + 2 * pad_s;
    const int xs = (int)(blockIdx.x * blockDim.x + threadIdx.x) - pad_s;
    const int ys = (int)(blockIdx.y * blockDim.y + threadIdx.y) - pad_s;
    if (xs >= S.width + pad_s || ys >= S.height + pad_s) return;
    (void)dom_w; (void)dom_h;
    uint8_t *q_out = const_cast<uint8_t *>(Q.org), *s_out = const_cast<uint8_t *>(S.org);
    const bool interior = xs >= 0 && xs < S.width && ys >= 0 && ys < S.height;
    uint32_t   q00, q01, q10, q11, sv;
    if (interior) {
        uint32_t r[4];
#pragma unroll
        for (int k = 0; k < 4; k++) r[k] = *reinterpret_cast<const uint32_t *>(F.org + (int64_t)(4 * ys + k) * F.stride + 4 * xs);
        auto px = [](uint32_t w, int i) { return (w >> (8 * i)) & 0xffu; };
        q00 = (px(r[0], 0) + px(r[0], 1) + px(r[1], 0) + px(r[1], 1) + 2) >> 2;
        q01 = (px(r[0], 2) + px(r[0], 3) + px(r[1], 2) + px(r[1], 3) + 2) >> 2;
        q10 = (px(r[2], 0) + px(r[2], 1) + px(r[3], 0) + px(r[3], 1) + 2) >> 2;
        q11 = (px(r[2], 2) + px(r[2], 3) + px(r[3], 2) + px(r[3], 3) + 2) >> 2;
        if (sixteenth_from_full) sv = (px(r[1], 1) + px(r[1], 2) + px(r[2], 1) + px(r[2], 2) + 2) >> 2;  // step 4: rows/cols 1,2
        else sv = (q00 + q01 + q10 + q11 + 2) >> 2;
        *reinterpret_cast<uchar2 *>(q_out + (int64_t)(2 * ys) * Q.stride + 2 * xs)     = make_uchar2((uint8_t)q00, (uint8_t)q01);
        *reinterpret_cast<uchar2 *>(q_out + (int64_t)(2 * ys + 1) * Q.stride + 2 * xs) = make_uchar2((uint8_t)q10, (uint8_t)q11);
        s_out[(int64_t)ys * S.stride + xs] = (uint8_t)sv;
        return;
    }
    // padding: replicate the edge pixel of the respective plane (values recomputed from the full plane)
    const int cx = min(max(xs, 0), S.width - 1), cy = min(max(ys, 0), S.height - 1);
    if (sixteenth_from_full) {
        const uint8_t *r1 = F.org + (int64_t)(4 * cy + 1) * F.stride + 4 * cx + 1;
        sv = ((uint32_t)r1[0] + r1[1] + r1[F.stride] + r1[F.stride + 1] + 2) >> 2;
    } else {
        sv = (quarter_px(F.org, F.stride, 2 * cx, 2 * cy) + quarter_px(F.org, F.stride, 2 * cx + 1, 2 * cy) +
              quarter_px(F.org, F.stride, 2 * cx, 2 * cy + 1) + quarter_px(F.org, F.stride, 2 * cx + 1, 2 * cy + 1) + 2) >> 2;
    }
    s_out[(int64_t)ys * S.stride + xs] = (uint8_t)sv;
#pragma unroll
    for (int j = 0; j < 2; j++)
#pragma unroll
        for (int i = 0; i < 2; i++) {
            const int xq = 2 * xs + i, yq = 2 * ys + j;
            if (xq < -pad_q || xq >= Q.width + pad_q || yq < -pad_q || yq >= Q.height + pad_q) continue;
            q_out[(int64_t)yq * Q.stride + xq] =
                (uint8_t)quarter_px(F.org, F.stride, min(max(xq, 0), Q.width - 1), min(max(yq, 0), Q.height - 1));
        }
}

__global__ void unpack16_kernel(const uint16_t *in, int in_stride, uint8_t *out, int out_stride, int w, int h) {
    const int x = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y;
    if (x >= w || y >= h) return;
    const uint16_t *ip = in + (int64_t)y * in_stride + x;
    uint32_t        v  = 0;
#pragma unroll
    for (int q = 0; q < 4; q++)
        if (x + q < w) v |= (uint32_t)((ip[q] >> 2) & 0xff) << (8 * q);
    *reinterpret_cast<uint32_t *>(out + (int64_t)y * out_stride + x) = v;  // rows are 16-byte aligned and w % 8 == 0
}

// Generic strided 2-D average used by the RTCD-level downsample_2d entry (any decim_step).
__global__ void downsample_generic_kernel(const uint8_t *in, int in_stride, int w, int h, uint8_t *out, int out_stride, int step) {
    const int half = step >> 1;
    const int ox = blockIdx.x * blockDim.x + threadIdx.x, oy = blockIdx.y * blockDim.y + threadIdx.y;
    const int x = half + ox * step, y = half + oy * step;
    if (x >= w || y >= h) return;
    const uint8_t *r0 = in + (int64_t)(y - 1) * in_stride, *r1 = in + (int64_t)y * in_stride;
    out[(int64_t)oy * out_stride + ox] = (uint8_t)(((uint32_t)r0[x - 1] + r0[x] + r1[x - 1] + r1[x] + 2) >> 2);
}

// ---- RTCD-level single-job kernels ----
__global__ void __launch_bounds__(kThreads) rtcd_sad_loop_kernel(SadJob job, uint32_t *out /* [4]: found, sad, x, y */) {
    extern __shared__ __align__(16) unsigned char smem[];
    CtaShared  *cs = reinterpret_cast<CtaShared *>(smem);
    CudaBackend be = make_backend(nullptr, smem);
    if (threadIdx.x == 0) { cs->sb.sad_jobs[0] = job; cs->sb.n_sad = 1; }
    __syncthreads();
    be.run_sad(cs->sb);
    __syncthreads();
    if (threadIdx.x == 0) {
        const SadJob &j = cs->sb.sad_jobs[0];
        out[0] = j.found; out[1] = j.best_sad; out[2] = (uint32_t)(int32_t)j.bx; out[3] = (uint32_t)(int32_t)j.by;
    }
}

__global__ void __launch_bounds__(kThreads) rtcd_nxm_kernel(NxmJob job, uint32_t *out) {
    extern __shared__ __align__(16) unsigned char smem[];
    CtaShared  *cs = reinterpret_cast<CtaShared *>(smem);
    CudaBackend be = make_backend(nullptr, smem);
    if (threadIdx.x == 0) { cs->sb.nxm_jobs[0] = job; cs->sb.n_nxm = 1; }
    __syncthreads();
    be.run_nxm(cs->sb);
    __syncthreads();
    if (threadIdx.x == 0) out[0] = cs->sb.nxm_jobs[0].sad;
}

__global__ void __launch_bounds__(kThreads) rtcd_fullpel_kernel(const uint8_t *src, int src_stride, FpJob job, int sub_sad,
                                                               uint32_t *best_sad, uint32_t *best_mv) {
    extern __shared__ __align__(16) unsigned char smem[];
    CtaShared  *cs = reinterpret_cast<CtaShared *>(smem);
    CudaBackend be = make_backend(nullptr, smem);
    SbState    &s  = cs->sb;
    if (threadIdx.x == 0) { s.fp_jobs[0] = job; s.n_fp = 1; s.src = src; s.src_stride = src_stride; }
    for (int i = threadIdx.x; i < kPus; i += blockDim.x) { s.best_sad[0][0][i] = best_sad[i]; s.best_mv[0][0][i] = best_mv[i]; }
    __syncthreads();
    be.stage_src(s);
    if (sub_sad) be.run_fullpel_t<true>(s);
    else be.run_fullpel_t<false>(s);
    __syncthreads();
    for (int i = threadIdx.x; i < kPus; i += blockDim.x) { best_sad[i] = s.best_sad[0][0][i]; best_mv[i] = s.best_mv[0][0][i]; }
}

// ---- INT-pipe peak probe: 8 independent VABSDIFF4.U8.ACC chains per thread, nothing else in the loop ----
__global__ void __launch_bounds__(256) sad_peak_kernel(uint32_t *out, int iters, uint32_t seed) {
    uint32_t a[8], acc[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = seed * (i + 1) + threadIdx.x; acc[i] = 0; }
    uint32_t b = seed ^ threadIdx.x;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 4; u++) {
#pragma unroll
            for (int i = 0; i < 8; i++) acc[i] = sad4(a[i], b, acc[i]);
            b += acc[0] & 1;
        }
    }
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) r += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

cudaError_t launch_sad_peak(uint32_t *d_out, int grid, int iters, cudaStream_t st) {
    sad_peak_kernel<<<grid, 256, 0, st>>>(d_out, iters, 12345u);
    return cudaGetLastError();
}

// ---- launch wrappers (host) ----
// the opt-in to more than 48 KB of dynamic shared memory is per device: svtme_init calls this with its device current
cudaError_t rtcd_set_attributes() {
    cudaError_t e = cudaFuncSetAttribute(rtcd_sad_loop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(rtcd_nxm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(rtcd_fullpel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    return e;
}

cudaError_t launch_pad_plane(uint8_t *org, int stride, int w, int h, int pad, cudaStream_t st) {
    const int total = ((w + 2 * pad + 15) >> 4) * (h + 2 * pad);
    pad_plane_kernel<<<(total + 255) / 256, 256, 0, st>>>(org, stride, w, h, pad);
    return cudaGetLastError();
}

cudaError_t launch_decimate(const DevPic &pic, int pad_q, int pad_s, int sixteenth_from_full, cudaStream_t st) {
    const dim3 blk(64, 4);
    const dim3 grd((pic.lvl[2].width + 2 * pad_s + blk.x - 1) / blk.x, (pic.lvl[2].height + 2 * pad_s + blk.y - 1) / blk.y);
    decimate_kernel<<<grd, blk, 0, st>>>(pic, pad_q, pad_s, sixteenth_from_full);
    return cudaGetLastError();
}

cudaError_t launch_unpack16(const uint16_t *in, int in_stride, uint8_t *out, int out_stride, int w, int h, cudaStream_t st) {
    const dim3 blk(128), grd(((w + 3) / 4 + 127) / 128, h);
    unpack16_kernel<<<grd, blk, 0, st>>>(in, in_stride, out, out_stride, w, h);
    return cudaGetLastError();
}

cudaError_t launch_downsample_generic(const uint8_t *in, int in_stride, int w, int h, uint8_t *out, int out_stride, int step,
                                      cudaStream_t st) {
    const dim3 blk(32, 8), grd((w / step + 32) / 32, (h / step + 8) / 8);
    downsample_generic_kernel<<<grd, blk, 0, st>>>(in, in_stride, w, h, out, out_stride, step);
    return cudaGetLastError();
}

cudaError_t launch_rtcd_sad_loop(const SadJob &job, uint32_t *d_out, cudaStream_t st) {
    rtcd_sad_loop_kernel<<<1, kThreads, kSmemBytes, st>>>(job, d_out);
    return cudaGetLastError();
}
cudaError_t launch_rtcd_nxm(const NxmJob &job, uint32_t *d_out, cudaStream_t st) {
    rtcd_nxm_kernel<<<1, kThreads, kSmemBytes, st>>>(job, d_out);
    return cudaGetLastError();
}
cudaError_t launch_rtcd_fullpel(const uint8_t *src, int src_stride, const FpJob &job, int sub_sad, uint32_t *d_best_sad,
                                uint32_t *d_best_mv, cudaStream_t st) {
    rtcd_fullpel_kernel<<<1, kThreads, kSmemBytes, st>>>(src, src_stride, job, sub_sad, d_best_sad, d_best_mv);
    return cudaGetLastError();
}

}  // namespace svtme
