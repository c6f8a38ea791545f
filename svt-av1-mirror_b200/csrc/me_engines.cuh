// me_engines.cuh -- the CUDA search engines of the open-loop ME path (sm_100a).
//
// One CTA owns one 64x64 block (see me_control.cuh).  The control code hands the engines lists of
// independent search jobs; the whole CTA executes a list:
//
//   run_nxm     svt_nxm_sad_kernel            (compute_sad_c.c:209)   one warp per job
//   run_sad     svt_sad_loop_kernel           (compute_sad_c.c:58)    HME / pre-HME windows
//   run_fullpel open_loop_me_fullpel_search_sblock (motion_estimation.c:781) = the fused
//               svt_ext_all_sad_calculation_8x8_16x16 + svt_ext_eight_sad_calculation_32x32_64x64
//               + svt_ext_sad_calculation_* loop
//
// Common machinery
//   * Search windows are staged into shared memory in TILES (a rectangle of candidate positions of
//     one job plus the block extent), 128-bit coalesced loads from the 16-byte aligned plane rows.
//     A batch of tiles fills the window buffer; jobs larger than the buffer are split.
//   * A thread evaluates 4 x-adjacent candidates that share a shared-memory word: the four byte
//     alignments come from funnel shifts (SHF issues beside the SAD pipe), the SADs are
//     VABSDIFF4.U8.ACC (4 pixels per lane per instruction).
//   * Tie-breaking: every candidate carries its raster index inside the job's window; minima are
//     taken over the key (SAD, raster index), which reproduces "first in raster order wins on a
//     strict <" of the reference exactly, independent of thread or tile order.
#pragma once
#include "me_control.cuh"

namespace svtme {

#ifndef SVTME_THREADS
#define SVTME_THREADS 256
#endif
#ifndef SVTME_WIN_KB
#define SVTME_WIN_KB 64
#endif
#ifndef SVTME_MIN_CTAS
#define SVTME_MIN_CTAS 2
#endif
constexpr int      kThreads   = SVTME_THREADS;
constexpr int      kMaxTiles  = 48;
constexpr uint32_t kWinBytes  = SVTME_WIN_KB * 1024;
constexpr int      kTileMaxNx = 128;  // candidates per tile row

struct Tile {
    uint32_t off;      // byte offset of the tile's first row in the window buffer (16-byte aligned)
    uint32_t item0;    // first flattened work item of this tile
    uint16_t job;
    uint16_t pitch;    // bytes per staged row (multiple of 16)
    int16_t  x0, y0;   // first candidate of the tile, relative to the job's window origin
    int16_t  nx, ny;   // candidates per row / candidate rows
    uint16_t ngroups;  // groups of 4 candidates (on shared-memory word boundaries) per candidate row
    uint16_t mis;      // byte offset of candidate x0 inside a staged row (global address & 15)
    uint16_t nrows;    // staged rows
    uint16_t nyeff;    // candidate rows that are actually searched (skip_line halves them)
};

struct EngineShared {
    Tile               tiles[kMaxTiles];
    int32_t            n_tiles;
    uint32_t           n_items;
    int32_t            cur_job, cur_x, cur_y;  // planning cursor
    uint32_t           next_sb;                // persistent-CTA work queue broadcast
    unsigned long long px;                     // pixel-compares of this SB (algorithmic work, same definition as the oracle's)
    unsigned long long sad_key[kMaxSadJobs];
    unsigned long long fp_key[kMaxFpJobs][kPus];
    alignas(16) uint32_t blk[64 * 16];   // HME source block of the current phase: [row][word], zero past block_w
    alignas(16) uint32_t src64[64 * 16]; // 64x64 source block of this SB: [row][word]
};

__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t fsr(uint32_t lo, uint32_t hi, int bytes) { return __funnelshift_r(lo, hi, bytes * 8); }

struct CudaBackend {
    const MeParams *p;
    EngineShared   *sh;
    uint8_t        *win;  // kWinBytes, 16-byte aligned

    __device__ __forceinline__ bool     leader() const { return threadIdx.x == 0; }
    __device__ __forceinline__ void     sync() const { __syncthreads(); }
    __device__ __forceinline__ uint32_t tid() const { return threadIdx.x; }
    __device__ __forceinline__ uint32_t nthreads() const { return blockDim.x; }

    // ---------------------------------------------------------------------------------------------
    // svt_nxm_sad_kernel: one warp per job, lane = (row parity group, word column).
    // ---------------------------------------------------------------------------------------------
    __device__ void run_nxm(SbState &s) const {
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
        if (threadIdx.x == 0)
            for (int ji = 0; ji < s.n_nxm; ji++) sh->px += (unsigned long long)s.nxm_jobs[ji].rows * s.nxm_jobs[ji].width;
        for (int ji = warp; ji < s.n_nxm; ji += nwarps) {
            NxmJob        &j  = s.nxm_jobs[ji];
            const int      nw = (j.width + 3) >> 2;
            uint32_t       acc = 0;
            for (int i = lane; i < j.rows * nw; i += 32) {
                const int      r = i / nw, c = i - r * nw;
                const uint8_t *sp = j.src + (int64_t)r * j.src_stride + 4 * c;
                const uint8_t *rp = j.ref + (int64_t)r * j.ref_stride + 4 * c;
                const int      n  = min(4, j.width - 4 * c);
                uint32_t       a = 0, b = 0;
                for (int q = 0; q < n; q++) { a |= (uint32_t)sp[q] << (8 * q); b |= (uint32_t)rp[q] << (8 * q); }
                acc = sad4(a, b, acc);
            }
            for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (lane == 0) j.sad = acc;
        }
    }

    // ---------------------------------------------------------------------------------------------
    // Tile planning (leader): fill the window buffer with tiles starting at the cursor.
    //   W,H      candidate grid of the job;  ext_x / ext_y: block extent beyond a candidate (bytes / rows)
    // ---------------------------------------------------------------------------------------------
    struct JobGeom {
        const uint8_t *ref;
        int32_t        raw_stride;
        int32_t        W, H, ext_x, ext_y;
        bool           skip;
    };
    template <class GeomFn>
    __device__ void plan_batch(int n_jobs, GeomFn geom) const {
        EngineShared &e    = *sh;
        uint32_t      used = 0, items = 0;
        int           nt   = 0;
        while (e.cur_job < n_jobs && nt < kMaxTiles) {
            const JobGeom g = geom(e.cur_job);
            if (g.W <= 0 || g.H <= 0) { e.cur_job++; e.cur_x = 0; e.cur_y = 0; continue; }
            const int      nx  = min(g.W - e.cur_x, kTileMaxNx);
            const uint32_t mis = (uint32_t)((uintptr_t)(g.ref + (int64_t)e.cur_y * g.raw_stride + e.cur_x) & 15);
            uint32_t pitch = (mis + nx + g.ext_x + 8 + 15) & ~15u;  // +8: the shifted loads touch one word past the last candidate
            if (((pitch >> 4) & 1) == 0) pitch += 16;  // odd number of 16-byte units: rows spread over the banks
            const int rows_fit = (int)((kWinBytes - used) / pitch) - g.ext_y;
            int       ny       = min(min(g.H - e.cur_y, rows_fit), 512);  // ny * nx <= 65536: 16-bit tile-local raster index
            if (ny < 1 || (ny < g.H - e.cur_y && ny < 8 && used != 0)) break;  // batch full
            Tile &t = e.tiles[nt++];
            t.off = used; t.item0 = items; t.job = (uint16_t)e.cur_job; t.pitch = (uint16_t)pitch;
            t.x0 = (int16_t)e.cur_x; t.y0 = (int16_t)e.cur_y; t.nx = (int16_t)nx; t.ny = (int16_t)ny;
            t.mis = (uint16_t)mis; t.ngroups = (uint16_t)(((mis & 3) + nx + 3) >> 2);
            t.nrows = (uint16_t)(ny + g.ext_y);
            t.nyeff = (uint16_t)(g.skip ? ((ny + (e.cur_y & 1)) >> 1) : ny);  // odd candidate rows only
            used += pitch * t.nrows;
            items += (uint32_t)t.nyeff * t.ngroups;
            e.cur_y += ny;
            if (e.cur_y >= g.H) {
                e.cur_y = 0; e.cur_x += nx;
                if (e.cur_x >= g.W) { e.cur_x = 0; e.cur_job++; }
            }
        }
        e.n_tiles = nt;
        e.n_items = items;
    }

    // Stage the rows of every tile of the batch: 16-byte vectors from 16-byte aligned addresses.
    template <class GeomFn>
    __device__ void load_batch(GeomFn geom) const {
        const EngineShared &e = *sh;
        for (int ti = 0; ti < e.n_tiles; ti++) {
            const Tile    &t   = e.tiles[ti];
            const JobGeom  g   = geom(t.job);
            const uint8_t *row0 = g.ref + (int64_t)t.y0 * g.raw_stride + t.x0 - t.mis;  // 16-byte aligned
            const uint32_t nvec = t.pitch >> 4, total = nvec * t.nrows;
            for (uint32_t i = threadIdx.x; i < total; i += blockDim.x) {
                const uint32_t r = i / nvec, v = i - r * nvec;
                const uint4    d = __ldg(reinterpret_cast<const uint4 *>(row0 + (int64_t)r * g.raw_stride) + v);
                *reinterpret_cast<uint4 *>(win + t.off + r * t.pitch + (v << 4)) = d;
            }
        }
    }

    // ---------------------------------------------------------------------------------------------
    // svt_sad_loop_kernel jobs.  All jobs of one call share the source block (same SB, same plane).
    // ---------------------------------------------------------------------------------------------
    __device__ void run_sad(SbState &s) const {
        EngineShared &e = *sh;
        if (s.n_sad == 0) return;
        const SadJob &j0    = s.sad_jobs[0];
        const int     rows  = j0.rows, bw = j0.block_w;
        const int     nw    = (bw + 3) >> 2;
        const int     rstep = j0.ref_stride / j0.raw_stride;  // 1 (full SAD) or 2 (sub-sampled SAD)
        // stage the source block, zero beyond block_w; the same mask is applied to the reference words
        const uint32_t tailmask = (bw & 3) ? ((1u << (8 * (bw & 3))) - 1u) : 0xffffffffu;
        for (int i = threadIdx.x; i < rows * nw; i += blockDim.x) {
            const int      r = i / nw, c = i - r * nw;
            const uint8_t *sp = j0.blk + (int64_t)r * j0.blk_stride + 4 * c;
            const int      n  = min(4, bw - 4 * c);
            uint32_t       a  = 0;
            for (int q = 0; q < n; q++) a |= (uint32_t)sp[q] << (8 * q);
            e.blk[r * 16 + c] = a;
        }
        for (int i = threadIdx.x; i < s.n_sad; i += blockDim.x) e.sad_key[i] = ~0ull;
        if (threadIdx.x == 0) {
            e.cur_job = 0; e.cur_x = 0; e.cur_y = 0;
            for (int ji = 0; ji < s.n_sad; ji++) {
                const SadJob &j = s.sad_jobs[ji];
                if (j.sa_w <= 0 || j.sa_h <= 0) continue;
                const bool skip = j.skip_line && j.block_w == 16 && j.rows <= 16;
                e.px += (unsigned long long)j.sa_w * (skip ? (j.sa_h >> 1) : j.sa_h) * j.rows * j.block_w;
            }
        }
        __syncthreads();

        auto geom = [&](int ji) {
            const SadJob &j = s.sad_jobs[ji];
            JobGeom       g;
            g.ref = j.ref; g.raw_stride = j.raw_stride; g.W = j.sa_w; g.H = j.sa_h;
            g.ext_x = bw - 1; g.ext_y = (rows - 1) * rstep;
            g.skip = j.skip_line && j.block_w == 16 && j.rows <= 16;
            return g;
        };
        const int nwfull = (bw & 3) ? nw - 1 : nw;

        for (;;) {
            if (threadIdx.x == 0) plan_batch(s.n_sad, geom);
            __syncthreads();
            if (e.n_tiles == 0) break;
            load_batch(geom);
            __syncthreads();

            unsigned long long best = ~0ull;
            int                bjob = -1, ti = 0;
            for (uint32_t i = threadIdx.x; i < e.n_items; i += blockDim.x) {
                while (ti + 1 < e.n_tiles && i >= e.tiles[ti + 1].item0) ti++;
                const Tile &t = e.tiles[ti];
                if ((int)t.job != bjob) {
                    if (bjob >= 0 && best != ~0ull) atomicMin(&e.sad_key[bjob], best);
                    bjob = t.job; best = ~0ull;
                }
                const uint32_t local = i - t.item0;
                const uint32_t yi = local / t.ngroups, g = local - yi * t.ngroups;
                const bool     skip = s.sad_jobs[t.job].skip_line && bw == 16 && rows <= 16;
                const int      yy   = skip ? (int)(2 * yi + ((t.y0 & 1) ? 0 : 1)) : (int)yi;
                const uint32_t pitchw = t.pitch >> 2;
                const uint32_t *rowp = reinterpret_cast<const uint32_t *>(win + t.off) + (uint32_t)yy * pitchw + (t.mis >> 2) + g;
                uint32_t a0 = 0, a1 = 0, a2 = 0, a3 = 0;
                for (int r = 0; r < rows; r++) {
                    const uint32_t *rp = rowp + (uint32_t)(r * rstep) * pitchw;
                    const uint32_t *bp = &e.blk[r * 16];
                    uint32_t        w0 = rp[0];
#pragma unroll 4
                    for (int c = 0; c < nwfull; c++) {
                        const uint32_t w1 = rp[c + 1], sw = bp[c];
                        a0 = sad4(w0, sw, a0);
                        a1 = sad4(fsr(w0, w1, 1), sw, a1);
                        a2 = sad4(fsr(w0, w1, 2), sw, a2);
                        a3 = sad4(fsr(w0, w1, 3), sw, a3);
                        w0 = w1;
                    }
                    if (nwfull != nw) {
                        const uint32_t w1 = rp[nw], sw = bp[nwfull];
                        a0 = sad4(w0 & tailmask, sw, a0);
                        a1 = sad4(fsr(w0, w1, 1) & tailmask, sw, a1);
                        a2 = sad4(fsr(w0, w1, 2) & tailmask, sw, a2);
                        a3 = sad4(fsr(w0, w1, 3) & tailmask, sw, a3);
                    }
                }
                // candidates x = x0 + 4g - (mis & 3) + {0,1,2,3}
                const int xb = t.x0 + 4 * (int)g - (int)(t.mis & 3);
                const int y  = t.y0 + yy;
                const uint32_t acc[4] = {a0, a1, a2, a3};
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const int x = xb + c;
                    if (x >= t.x0 && x < t.x0 + t.nx) {
                        const unsigned long long key = ((unsigned long long)acc[c] << 32) | ((uint32_t)y << 16) | (uint32_t)x;
                        if (key < best) best = key;
                    }
                }
            }
            if (bjob >= 0 && best != ~0ull) atomicMin(&e.sad_key[bjob], best);
            __syncthreads();
        }
        for (int i = threadIdx.x; i < s.n_sad; i += blockDim.x) {
            SadJob                  &j   = s.sad_jobs[i];
            const unsigned long long key = e.sad_key[i];
            if (key != ~0ull && (uint32_t)(key >> 32) < 0xffffffu) {  // strict '<' against the 0xffffff start value
                j.found = 1; j.best_sad = (uint32_t)(key >> 32);
                j.by = (int16_t)((key >> 16) & 0xffff); j.bx = (int16_t)(key & 0xffff);
            }
        }
    }

    // ---------------------------------------------------------------------------------------------
    // Full-pel search of the 64x64 block: lane = (16x16 block in z-order, candidate group); 32x32 and
    // 64x64 SADs are butterfly sums over the 16 lanes of a candidate group.
    // ---------------------------------------------------------------------------------------------
    template <bool SUB>
    __device__ void run_fullpel_t(SbState &s) const {
        EngineShared &e = *sh;
        constexpr int RSTEP = SUB ? 2 : 1;
        for (int i = threadIdx.x; i < s.n_fp * kPus; i += blockDim.x) (&e.fp_key[0][0])[i] = ~0ull;
        if (threadIdx.x == 0) {
            e.cur_job = 0; e.cur_x = 0; e.cur_y = 0;
            for (int ji = 0; ji < s.n_fp; ji++)
                if (s.fp_jobs[ji].w > 0 && s.fp_jobs[ji].h > 0) e.px += (unsigned long long)s.fp_jobs[ji].w * s.fp_jobs[ji].h * (SUB ? 64 * 32 : 64 * 64);
        }
        __syncthreads();
        auto geom = [&](int ji) {
            const FpJob &j = s.fp_jobs[ji];
            JobGeom      g;
            g.ref = j.ref; g.raw_stride = j.ref_stride; g.W = j.w; g.H = j.h; g.ext_x = 63; g.ext_y = 63; g.skip = false;
            return g;
        };
        const int b = threadIdx.x & 15, sub = threadIdx.x >> 4, nsub = blockDim.x >> 4;
        const int bx = ((b >> 2) & 1) * 2 + (b & 1), by = (b >> 3) * 2 + ((b >> 1) & 1);

        for (;;) {
            if (threadIdx.x == 0) plan_batch(s.n_fp, geom);
            __syncthreads();
            if (e.n_tiles == 0) break;
            load_batch(geom);
            __syncthreads();

            for (int half = 0; half < (SUB ? 1 : 2); half++) {
                // source rows of this lane's 16x16 block: SUB: rows 0,2,..,14; FULL: rows 8*half .. 8*half+7
                uint32_t sreg[8][4];
#pragma unroll
                for (int r = 0; r < 8; r++)
#pragma unroll
                    for (int c = 0; c < 4; c++) sreg[r][c] = e.src64[(16 * by + (SUB ? 2 * r : 8 * half + r)) * 16 + 4 * bx + c];

                uint32_t           best8[4] = {~0u, ~0u, ~0u, ~0u}, best16 = ~0u;
                unsigned long long best32 = ~0ull, best64 = ~0ull;
                int                ti = 0, btile = -1;
                const uint32_t     rounds = (e.n_items + nsub - 1) / nsub;
                for (uint32_t it = 0; it <= rounds; it++) {
                    const uint32_t i      = it * nsub + sub;
                    const bool     active = it < rounds && i < e.n_items;
                    if (active)
                        while (ti + 1 < e.n_tiles && i >= e.tiles[ti + 1].item0) ti++;
                    const int want = active ? ti : -1;
                    if (want != btile) {  // flush this lane's running minima of the finished tile
                        if (btile >= 0) {
                            const Tile  &t  = e.tiles[btile];
                            const FpJob &j  = s.fp_jobs[t.job];
                            auto         gk = [&](uint32_t sad, uint32_t ord) {
                                const uint32_t yy = ord / (uint32_t)t.nx, xx = ord - yy * (uint32_t)t.nx;
                                return ((unsigned long long)(SUB ? sad << 1 : sad) << 32) | (uint32_t)((t.y0 + yy) * j.w + t.x0 + xx);
                            };
                            if (SUB) {
#pragma unroll
                                for (int q = 0; q < 4; q++)
                                    if (best8[q] != ~0u) atomicMin(&e.fp_key[t.job][21 + 4 * b + q], gk(best8[q] >> 16, best8[q] & 0xffff));
                                if (best16 != ~0u) atomicMin(&e.fp_key[t.job][5 + b], gk(best16 >> 16, best16 & 0xffff));
                                if ((b & 3) == 0 && best32 != ~0ull) atomicMin(&e.fp_key[t.job][1 + (b >> 2)], gk((uint32_t)(best32 >> 32), (uint32_t)best32));
                                if (b == 0 && best64 != ~0ull) atomicMin(&e.fp_key[t.job][0], gk((uint32_t)(best64 >> 32), (uint32_t)best64));
                            } else {
                                // FULL: this pass covered the top (half 0) or bottom (half 1) pair of 8x8 blocks
#pragma unroll
                                for (int q = 0; q < 2; q++)
                                    if (best8[q] != ~0u) atomicMin(&e.fp_key[t.job][21 + 4 * b + 2 * half + q], gk(best8[q] >> 16, best8[q] & 0xffff));
                            }
                        }
                        btile = want;
                        best8[0] = best8[1] = best8[2] = best8[3] = best16 = ~0u;
                        best32 = best64 = ~0ull;
                    }
                    if (it == rounds) break;
                    const Tile    &t      = e.tiles[active ? ti : 0];
                    const uint32_t local  = active ? i - t.item0 : 0;
                    const uint32_t yy     = local / t.ngroups, g = local - yy * t.ngroups;
                    const uint32_t pitchw = t.pitch >> 2;
                    uint32_t       acc[4][4];
#pragma unroll
                    for (int c = 0; c < 4; c++) acc[c][0] = acc[c][1] = acc[c][2] = acc[c][3] = 0;
                    if (active) {
                        const uint32_t *rowp = reinterpret_cast<const uint32_t *>(win + t.off) +
                            (yy + 16 * by + (SUB ? 0 : 8 * half)) * pitchw + (t.mis >> 2) + g + 4 * bx;
#pragma unroll
                        for (int r = 0; r < 8; r++) {
                            const uint32_t *rp = rowp + (uint32_t)(r * RSTEP) * pitchw;
                            uint32_t        w[5];
#pragma unroll
                            for (int c = 0; c < 5; c++) w[c] = rp[c];
#pragma unroll
                            for (int c = 0; c < 4; c++) {
                                const int      q  = SUB ? ((r >> 2) * 2 + (c >> 1)) : (c >> 1);
                                const uint32_t sw = sreg[r][c];
                                acc[0][q] = sad4(w[c], sw, acc[0][q]);
                                acc[1][q] = sad4(fsr(w[c], w[c + 1], 1), sw, acc[1][q]);
                                acc[2][q] = sad4(fsr(w[c], w[c + 1], 2), sw, acc[2][q]);
                                acc[3][q] = sad4(fsr(w[c], w[c + 1], 3), sw, acc[3][q]);
                            }
                        }
                    }
                    const int xb = 4 * (int)g - (int)(t.mis & 3);  // candidate x relative to the tile
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        const int      xx    = xb + c;
                        const bool     valid = active && xx >= 0 && xx < t.nx;
                        const uint32_t ord   = yy * (uint32_t)t.nx + (uint32_t)xx;  // < 65536 by construction of the tiles
                        if (SUB) {
                            uint32_t s16 = 0;
#pragma unroll
                            for (int q = 0; q < 4; q++) {
                                if (valid) best8[q] = min(best8[q], (acc[c][q] << 16) | ord);
                                s16 += acc[c][q];
                            }
                            if (valid) best16 = min(best16, (s16 << 16) | ord);
                            uint32_t s32 = s16 + __shfl_xor_sync(0xffffffffu, s16, 1);
                            s32 += __shfl_xor_sync(0xffffffffu, s32, 2);
                            uint32_t s64 = s32 + __shfl_xor_sync(0xffffffffu, s32, 4);
                            s64 += __shfl_xor_sync(0xffffffffu, s64, 8);
                            if (valid) {
                                best32 = min(best32, ((unsigned long long)s32 << 32) | ord);
                                best64 = min(best64, ((unsigned long long)s64 << 32) | ord);
                            }
                        } else if (valid) {
                            best8[0] = min(best8[0], (acc[c][0] << 16) | ord);
                            best8[1] = min(best8[1], (acc[c][1] << 16) | ord);
                        }
                    }
                }
            }
            __syncthreads();
        }
        if (!SUB) {
            // FULL SAD: the 16x16 / 32x32 / 64x64 minima need both halves of every candidate; they are rebuilt here
            // from per-candidate sums in a second, simple pass (the full-SAD mode is off the open-loop hot path).
            fullpel_full_upper_levels(s);
        }
        __syncthreads();
        // fold the window minima into the running best of the reference (strict '<': earlier results win ties)
        for (int i = threadIdx.x; i < s.n_fp * kPus; i += blockDim.x) {
            const int                ji = i / kPus, pu = i - ji * kPus;
            const FpJob             &j  = s.fp_jobs[ji];
            const unsigned long long key = e.fp_key[ji][pu];
            if (key == ~0ull) continue;
            const uint32_t sad = (uint32_t)(key >> 32), ord = (uint32_t)key;
            if (sad < s.best_sad[j.li][j.ri][pu]) {
                const int32_t y = (int32_t)(ord / (uint32_t)j.w), x = (int32_t)(ord - (uint32_t)y * (uint32_t)j.w);
                s.best_sad[j.li][j.ri][pu] = sad;
                s.best_mv[j.li][j.ri][pu]  = ((uint32_t)(y + j.yo) << 16) | (uint16_t)(x + j.xo);
            }
        }
        __syncthreads();
    }

    // 16x16 and larger PUs in FULL-SAD mode: thread = (candidate, 16x16 block) straight from global memory.
    __device__ void fullpel_full_upper_levels(SbState &s) const {
        EngineShared &e = *sh;
        const int b = threadIdx.x & 15, sub = threadIdx.x >> 4, nsub = blockDim.x >> 4;
        const int bx = ((b >> 2) & 1) * 2 + (b & 1), by = (b >> 3) * 2 + ((b >> 1) & 1);
        for (int ji = 0; ji < s.n_fp; ji++) {
            const FpJob   &j      = s.fp_jobs[ji];
            const uint32_t ncand  = (uint32_t)j.w * (uint32_t)j.h;
            const uint32_t rounds = (ncand + nsub - 1) / nsub;
            unsigned long long best16 = ~0ull, best32 = ~0ull, best64 = ~0ull;
            for (uint32_t it = 0; it < rounds; it++) {
                const uint32_t i = it * nsub + sub;
                const bool     active = i < ncand;
                uint32_t       s16 = 0;
                if (active) {
                    const uint32_t y = i / (uint32_t)j.w, x = i - y * (uint32_t)j.w;
                    const uint8_t *rp = j.ref + (int64_t)(y + 16 * by) * j.ref_stride + x + 16 * bx;
                    for (int r = 0; r < 16; r++)
                        for (int c = 0; c < 4; c++) {
                            uint32_t v = 0;
                            for (int q = 0; q < 4; q++) v |= (uint32_t)rp[(int64_t)r * j.ref_stride + 4 * c + q] << (8 * q);
                            s16 = sad4(v, e.src64[(16 * by + r) * 16 + 4 * bx + c], s16);
                        }
                }
                uint32_t s32 = s16 + __shfl_xor_sync(0xffffffffu, s16, 1);
                s32 += __shfl_xor_sync(0xffffffffu, s32, 2);
                uint32_t s64 = s32 + __shfl_xor_sync(0xffffffffu, s32, 4);
                s64 += __shfl_xor_sync(0xffffffffu, s64, 8);
                if (active) {
                    best16 = min(best16, ((unsigned long long)s16 << 32) | i);
                    best32 = min(best32, ((unsigned long long)s32 << 32) | i);
                    best64 = min(best64, ((unsigned long long)s64 << 32) | i);
                }
            }
            if (best16 != ~0ull) atomicMin(&e.fp_key[ji][5 + b], best16);
            if ((b & 3) == 0 && best32 != ~0ull) atomicMin(&e.fp_key[ji][1 + (b >> 2)], best32);
            if (b == 0 && best64 != ~0ull) atomicMin(&e.fp_key[ji][0], best64);
        }
    }

    __device__ void run_fullpel(SbState &s) const {
        if (s.n_fp == 0) return;
        if (p->k.me_search_method == 0) run_fullpel_t<true>(s);
        else run_fullpel_t<false>(s);
    }

    // Stage the 64x64 source block of the SB (rows beyond the picture come from the replicated padding,
    // exactly what the reference reads: motion_estimation.c:452-471).
    __device__ void stage_src(const SbState &s) const {
        for (int i = threadIdx.x; i < 64 * 4; i += blockDim.x) {
            const int   r = i >> 2, v = i & 3;
            const uint4 d = __ldg(reinterpret_cast<const uint4 *>(s.src + (int64_t)r * s.src_stride) + v);
            *reinterpret_cast<uint4 *>(&sh->src64[r * 16 + 4 * v]) = d;
        }
        __syncthreads();
    }
};

}  // namespace svtme
