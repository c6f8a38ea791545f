// me_tpl.cu -- TPL source-based search on the device.
//
// The decision the reference takes per TPL block while `tpl_src_data_ready == 0`
// (tpl_mc_flow_dispenser_sb_generic, Source/Lib/Codec/src_ops_process.c:519-884) for the controls v3.0.2 uses at presets
// M3..M13: DC-predictor SAD from the source neighbours (:623-660) against the SAD of every single-reference ME candidate at
// its clamped full-pel vector (:759-883), first-best-wins, NEWMV when the inter cost is strictly smaller (:886-887).
// It reads the picture's ME results where the search kernels left them in HBM, so the candidates never cross PCIe.
//
// One warp per TPL block (16 / 4 / 1 per 64x64 block), in small CTAs (four warps: a quarter of a 64x64 block) so that the
// kernel slots in beside the search kernels of the pictures in flight.  A lane holds one source row of the block in registers
// for all candidates (16-pixel blocks: each half-warp holds the 16 rows and the two halves cost two candidates per pass), reads
// the matching reference row as aligned words funnel-shifted into place, packed-byte SAD (VABSDIFF4) per word, a (half-)warp
// add-reduction per candidate, the running best in lane-uniform registers.  The kernel is bound by L1 wavefronts of the reference
// rows (ncu: l1tex throughput 87 % of peak with word loads), not by arithmetic: ~5 candidates x 256 bytes per 16x16 block.
#include <cuda_runtime.h>
#include <stdint.h>

#include "me_types.cuh"
#include "me_launch.h"

namespace svtme {

namespace {

constexpr int kTplPad = 32;  // TPL_PADX / TPL_PADY, encode_context.h:43-44
#ifndef SVTME_TPL_THREADS
#define SVTME_TPL_THREADS 128
#endif
constexpr int kTplThreads = SVTME_TPL_THREADS;    // a 64x64 block has 16 block slots (one warp each); a CTA takes kTplThreads / 32 of them
constexpr int kTplCtasPerSb = 512 / kTplThreads;

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) { return __reduce_add_sync(0xffffffffu, v); }

// One row of the block per lane.  RW = words per row (4 / 8 / 16 for 16 / 32 / 64-pixel blocks).
template <int RW>
struct Row {
    uint32_t w[RW];
    __device__ __forceinline__ void load_aligned(const uint8_t *p) {  // source rows: x is a multiple of 16, planes are 16-byte aligned
#pragma unroll
        for (int i = 0; i < RW; i += 4) {
            const uint4 v = *reinterpret_cast<const uint4 *>(p + 4 * i);
            w[i] = v.x; w[i + 1] = v.y; w[i + 2] = v.z; w[i + 3] = v.w;
        }
    }
    __device__ __forceinline__ uint32_t sad_const(uint32_t dc4) const {
        uint32_t a = 0;
#pragma unroll
        for (int i = 0; i < RW; i++) a = __vsadu4(w[i], dc4) + a;
        return a;
    }
    // against RW * 4 bytes at any alignment: the RW / 4 + 1 aligned 16-byte chunks that cover them (128-bit loads: the kernel is
    // bound by L1 wavefronts -- one row per lane means one cache line per lane and load instruction -- so a row costs 2 / 3 / 5
    // load instructions instead of 5 / 9 / 17 word loads), rotated by the word offset with two select stages, then
    // funnel-shifted by the byte offset.  The bytes read beyond the row are inside the plane's physical padding.
    __device__ __forceinline__ uint32_t sad_ref(const uint8_t *ref) const {
        constexpr int   NV = RW / 4 + 1;
        const uint32_t  m = (uint32_t)(reinterpret_cast<uintptr_t>(ref) & 15), wi = m >> 2, sh = (m & 3) * 8;
        const uint4    *q4 = reinterpret_cast<const uint4 *>(ref - m);
        uint32_t        q[4 * NV];
#pragma unroll
        for (int v = 0; v < NV; v++) {
            const uint4 t = q4[v];
            q[4 * v] = t.x; q[4 * v + 1] = t.y; q[4 * v + 2] = t.z; q[4 * v + 3] = t.w;
        }
#pragma unroll
        for (int i = 0; i + 1 < 4 * NV; i++) q[i] = (wi & 1) ? q[i + 1] : q[i];
#pragma unroll
        for (int i = 0; i + 2 < 4 * NV; i++) q[i] = (wi & 2) ? q[i + 2] : q[i];
        uint32_t a = 0;
#pragma unroll
        for (int i = 0; i < RW; i++) a = __vsadu4(w[i], __funnelshift_r(q[i], q[i + 1], sh)) + a;
        return a;
    }
};

// DC predictor value of the block at (x, y): svt_aom_dc_pred[x > 0][y > 0] over neighbour arrays that hold 127 (above) /
// 129 (left) beyond the picture's right / bottom edge (enc_intra_prediction.c:1153-1155, 1175, 1198)
__device__ uint32_t dc_value(const DevPlane &p, int x, int y, int size, int lane) {
    const uint8_t *src = p.org + (size_t)y * p.stride + x;
    uint32_t above = 0, left = 0;
    for (int i = lane; i < size; i += 32) {
        if (y > 0) above += (x + i < p.width) ? src[i - p.stride] : 127u;
        if (x > 0) left += (y + i < p.height) ? src[(size_t)i * p.stride - 1] : 129u;
    }
    above = warp_sum(above);
    left  = warp_sum(left);
    if (x > 0 && y > 0) return (above + left + size) / (2u * size);
    if (y > 0) return (above + (size >> 1)) / size;
    if (x > 0) return (left + (size >> 1)) / size;
    return 128u;
}

// One TPL block of SIZE x SIZE pixels at (x, y) by one warp.  16-pixel blocks: the two half-warps each hold the 16 source rows
// and cost two candidates per pass; 32-pixel blocks: one row per lane; 64-pixel blocks: two rows per lane.
template <int SIZE>
__device__ void tpl_block(const TplParams &p, int sb, int x, int y, uint32_t me_off, int lane) {
    constexpr int RW = SIZE / 4, ROWS_PER_LANE = SIZE == 64 ? 2 : 1, PAR = SIZE == 16 ? 2 : 1;  // PAR candidates per pass
    const DevPlane &sp  = p.cur.lvl[0];
    const uint8_t  *src = sp.org + (size_t)y * sp.stride + x;
    const int       row = SIZE == 16 ? (lane & 15) : lane, half = SIZE == 16 ? (lane >> 4) : 0;
    Row<RW> s[ROWS_PER_LANE];
#pragma unroll
    for (int r = 0; r < ROWS_PER_LANE; r++) s[r].load_aligned(src + (size_t)(row + 32 * r) * sp.stride);
    // lane ci fetches candidate ci, lane r the vector of reference slot r of this PU: independent loads, in flight together with
    // the source rows and the DC neighbours.  max_cand <= 23 and max_refs <= 7 (MAX_PA_ME_CAND / MAX_PA_ME_MV, me_sb_results.h:21-22)
    const size_t   pu    = (size_t)sb * p.pus + me_off;
    const uint32_t total = p.slice_type_i ? 0 : p.me_count[pu];
    uint32_t my_ref = 0xffu, my_mv = 0;
    if (!p.slice_type_i) {
        if ((uint32_t)lane < p.max_refs) my_mv = p.me_mv[pu * p.max_refs + lane];
        if ((uint32_t)lane < p.max_cand) {
            const uint32_t cand = p.me_cand[pu * p.max_cand + lane];
            const uint32_t dir = cand & 3, li = dir, ri = li == 0 ? (cand >> 2) & 3 : (cand >> 4) & 3;
            if ((uint32_t)lane < total && dir <= 1 && !(p.ref_excluded & (1u << (li * 4 + ri))))  // :775-776, :783-785
                my_ref = li * 4 + ri;
        }
    }
    auto mv_slot = [&](uint32_t rf) { return ((rf >> 2) ? p.max_l0 : 0u) + (rf & 3); };  // :787-788
    uint32_t best_intra = 0xFFFFFFFFu, best_inter = 0xFFFFFFFFu;
    int      best_rf = -1, best_row = 0, best_col = 0;
    if (!p.disable_intra_pred) {
        const uint32_t dc4 = dc_value(sp, x, y, SIZE, lane) * 0x01010101u;
        uint32_t a = 0;
#pragma unroll
        for (int r = 0; r < ROWS_PER_LANE; r++) a += s[r].sad_const(dc4);
        if (SIZE == 16 && half) a = 0;  // both half-warps hold the same rows
        best_intra = warp_sum(a);
    }
    // candidate c0 + half of the block: its reference index and clamped full-pel vector, and this lane's row SAD against it
    auto cost_of = [&](uint32_t ci, uint32_t &rf, int &mvx, int &mvy) -> uint32_t {
        rf = __shfl_sync(0xffffffffu, my_ref, ci & 31);  // lanes >= total hold 0xff: no candidate
        const uint32_t mvw = __shfl_sync(0xffffffffu, my_mv, (rf == 0xffu ? 0u : mv_slot(rf)) & 31);
        mvx = (int16_t)(mvw & 0xffff);
        mvy = (int16_t)(mvw >> 16);
        if (rf == 0xffu) return 0;
        const DevPlane &rp = p.refs[rf >> 2][rf & 3].lvl[0];
        // :794-804, in full-pel units (the reference clamps the <<3 value with the same comparisons on >>3)
        if (x + mvx < -kTplPad) mvx = -kTplPad - x;
        if (x + SIZE + mvx > kTplPad + rp.width - 1) mvx = (kTplPad + rp.width - 1) - (x + SIZE);
        if (y + mvy < -kTplPad) mvy = -kTplPad - y;
        if (y + SIZE + mvy > kTplPad + rp.height - 1) mvy = (kTplPad + rp.height - 1) - (y + SIZE);
        const uint8_t *ref = rp.org + (ptrdiff_t)(y + mvy + row) * rp.stride + (x + mvx);
        uint32_t a = 0;
#pragma unroll
        for (int r = 0; r < ROWS_PER_LANE; r++) a += s[r].sad_ref(ref + (ptrdiff_t)(32 * r) * rp.stride);
        return a;
    };
    auto reduce = [&](uint32_t a) -> uint32_t {
        if (SIZE == 16) {  // per half-warp sums
#pragma unroll
            for (int o = 8; o; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
            return a;
        }
        return warp_sum(a);
    };
    auto take = [&](uint32_t a, uint32_t rf, int mvx, int mvy) {  // the PAR candidates of one pass, in candidate order
#pragma unroll
        for (int h = 0; h < PAR; h++) {
            const int      from = h * 16;
            const uint32_t cost = __shfl_sync(0xffffffffu, a, from), crf = __shfl_sync(0xffffffffu, rf, from);
            const int      cx = __shfl_sync(0xffffffffu, mvx, from), cy = __shfl_sync(0xffffffffu, mvy, from);
            if (crf != 0xffu && cost < best_inter) {  // :874-883: first best wins
                best_inter = cost;
                best_rf    = (int)crf;
                best_row   = cy * 8;
                best_col   = cx * 8;
            }
        }
    };
    // two passes per iteration: the reference rows of both are requested before either is reduced
    for (uint32_t c0 = 0; c0 < total; c0 += 2 * PAR) {
        uint32_t rf0, rf1;
        int      x0, y0, x1, y1;
        uint32_t a0 = cost_of(c0 + half, rf0, x0, y0);
        uint32_t a1 = cost_of(c0 + PAR + half, rf1, x1, y1);
        a0 = reduce(a0);
        a1 = reduce(a1);
        take(a0, rf0, x0, y0);
        take(a1, rf1, x1, y1);
    }
    if (lane == 0) {
        const uint32_t cell = (uint32_t)(y >> 4) * p.grid_width + (uint32_t)(x >> 4);
        p.block_size[cell]      = (uint8_t)SIZE;
        p.best_mode[cell]       = (best_rf >= 0 && best_inter < best_intra) ? 16 : 0;  // NEWMV : DC_PRED (:886-887)
        p.best_rf_idx[cell]     = (int8_t)best_rf;
        p.mv[cell * 2 + 0]      = (int16_t)best_row;
        p.mv[cell * 2 + 1]      = (int16_t)best_col;
        p.ref_frame_poc[cell]   = best_rf >= 0 ? p.ref_poc[best_rf] : 0ull;
        p.best_inter_cost[cell] = best_inter;
        p.best_intra_cost[cell] = best_intra;
    }
}

__global__ void __launch_bounds__(kTplThreads) tpl_src_kernel(const TplParams p) {
    const int sb = blockIdx.x / kTplCtasPerSb, lane = threadIdx.x & 31;
    const int warp = (blockIdx.x % kTplCtasPerSb) * (kTplThreads / 32) + (threadIdx.x >> 5);  // block slot 0..15 of this 64x64 block
    const int sx = (sb % p.wb) * 64, sy = (sb / p.wb) * 64;
    const int W = p.cur.lvl[0].width, H = p.cur.lvl[0].height;
    const bool complete = sx + 64 <= W && sy + 64 <= H;  // b64_geom width == 64 && height == 64 (:2048-2050)
    const int  level = complete ? p.level : 0, size = 16 << level, n = 64 / size;
    for (int b = warp; b < n * n; b += 16) {
        const int by = b / n, bx = b - by * n;
        const int x = sx + bx * size, y = sy + by * size;
        if (x + (size >> 1) > W || y + (size >> 1) > H) continue;  // at least half of the block inside (:584-586)
        // the ME block co-located with this TPL block (tpl_blk_idx_tab[1], :353-355: raster order within its level)
        uint32_t me_off = level == 0 ? 5 + by * 4 + bx : level == 1 ? 1 + by * 2 + bx : 0;
        if (!p.enable_me_16x16) me_off = (me_off - 1) / 4;  // :761-762
        if (level == 0) tpl_block<16>(p, sb, x, y, me_off, lane);
        else if (level == 1) tpl_block<32>(p, sb, x, y, me_off, lane);
        else tpl_block<64>(p, sb, x, y, me_off, lane);
    }
}

}  // namespace

cudaError_t launch_tpl_src(const TplParams &p, cudaStream_t st) {
    tpl_src_kernel<<<p.wb * p.hb * kTplCtasPerSb, kTplThreads, 0, st>>>(p);
    return cudaGetLastError();
}

}  // namespace svtme
