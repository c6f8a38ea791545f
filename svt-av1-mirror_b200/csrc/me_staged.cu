// me_staged.cu -- the picture-wide ("staged") execution of the per-SB step program.
//
// Every ENGINE step of the program (me_control.cuh: build_program) is ONE KERNEL OVER THE WHOLE PICTURE, one CTA of 128
// threads per 64x64 block:
//     pre-HME / HME level 0 (sixteenth plane)   st_sad16_kernel      CTA-wide TMA-staged windows, flat item list
//     HME level 1 / 2 (quarter / full plane)     st_sad_rows_kernel   warp per 8x3 refinement window, lane = block row
//     1x1 probe + full-pel scan                  st_fp_sub_kernel     CTA-wide windows, rounds dealt to the warps
//     anything unusual (full SAD, picture edge)  st_sad_kernel        generic warp-per-job path
// The control steps between two engines (fold the search results, prune references, derive the next search windows:
// the scalar code of motion_estimation.c around its SAD loops) and the light svt_nxm_sad_kernel probes run in the TAIL
// of the kernel in front of them, on the shared-memory copy of the block's control state that the kernel already holds;
// the last kernel's tail also builds the results.  Block state (SbState) lives in HBM/L2 between kernels.  Kernel
// launches of one picture are stream ordered (programmatic dependent launch); several pictures are in flight on
// different streams.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "me_engines.cuh"
#include "me_launch.h"

namespace svtme {

// -DSVTME_PHASES: thread 0 of every CTA accumulates the clocks it spends in each phase of the engine kernels
// (diagnostics; printed per search by SVTME_TRACE=1)
#ifdef SVTME_PHASES
__device__ unsigned long long g_phase[4][12];  // 0 full-pel scan, 1 pre-HME, 2 HME level 0, 3 HME level 1 / 2 rows kernels
#define PH_DECL unsigned long long ph_t = clock64();
#define PH(k, i) if (threadIdx.x == 0) { const unsigned long long ph_n = clock64(); atomicAdd(&g_phase[k][i], ph_n - ph_t); ph_t = ph_n; }
#else
#define PH_DECL
#define PH(k, i)
#endif

// -DSVTME_BOUNDS: every staged copy checks that its source bytes lie inside the reference plane's allocation
// (physical padding 96 / 48 / 32 + 256 bytes of slack, csrc/abi.cu plane_geometry); a violation prints and traps.
#ifdef SVTME_BOUNDS
__device__ __forceinline__ void check_src(const MeParams &p, int lvl, int li, int ri, const uint8_t *g, int bytes, int site) {
    const DevPlane &pl = p.refs[li][ri].lvl[lvl];
    const int       pad = lvl == 0 ? 96 : lvl == 1 ? 48 : 32;
    const uint8_t  *lo = pl.org - (int64_t)pad * pl.stride - pad;
    const uint8_t  *hi = lo + (int64_t)pl.stride * (pl.height + 2 * pad) + 256;
    if (g < lo || g + bytes > hi) {
        printf("svtme bounds: site %d lvl %d ref %d/%d: %lld bytes before / %lld past the plane\n", site, lvl, li, ri,
               (long long)(lo - g), (long long)(g + bytes - hi));
        __trap();
    }
}
#define CHECK_SRC(p, lvl, li, ri, g, bytes, site) check_src(p, lvl, li, ri, reinterpret_cast<const uint8_t *>(g), bytes, site)
#else
#define CHECK_SRC(p, lvl, li, ri, g, bytes, site)
#endif

#ifndef SVTME_S16_CTAS
#define SVTME_S16_CTAS 6  // CTAs per SM of the persistent sixteenth-plane engine
#endif
#ifndef SVTME_ENGINE_CTAS
#define SVTME_ENGINE_CTAS 0  // 0: one CTA per block
#endif
constexpr int kStThreads  = 128;             // 4 warps per block CTA
constexpr int kStWarps    = kStThreads / 32;
constexpr int kSadWarpWin = 8 * 1024;        // HME window bytes per warp

// Programmatic dependent launch (default; SVTME_PDL=0 turns it off): every kernel of a picture's chain carries the
// programmatic-stream-serialization attribute and waits here, first thing, until its predecessor's grid is complete and
// its memory visible.  No kernel signals `griddepcontrol.launch_dependents` itself: the implicit signal at CTA exit lets
// the successor's CTAs move into the slots the predecessor's last wave frees (launch latency and tail overlap).  An
// explicit signal at the top of the kernel was measured worse (0.418 vs 0.401 ms for one search alone, -2 % end to
// end): the successor's CTAs become resident as soon as the predecessor's last CTA has STARTED, and spin in the slots
// that the predecessor's later waves -- a 2040-CTA grid is 2.3 waves -- and the other searches' kernels need.
// Without the launch attribute the instruction is a no-op.
#define PDL_ENTRY() asm volatile("griddepcontrol.wait;" ::: "memory")

// cooperative helpers of a block CTA
struct StBackend {
    __device__ __forceinline__ bool     leader() const { return threadIdx.x == 0; }
    __device__ __forceinline__ void     sync() const { __syncthreads(); }
    __device__ __forceinline__ uint32_t tid() const { return threadIdx.x; }
    __device__ __forceinline__ uint32_t nthreads() const { return blockDim.x; }
};

// ---- the block's control state on chip ----
// Every kernel of a picture's chain works on a shared-memory copy of the CONTROL part of its block's state (kSbHeadBytes:
// HME centres / SADs, pruning flags, the three job tables; everything but the best-SAD / best-MV tables, which stay in
// HBM/L2 behind SbState::home).  The copy is requested first thing with 16-byte cp.async (it lands while the kernel fetches
// its source block), the engine reads its jobs from it and leaves their results in it, and the control steps that follow
// the engine in the program (fold the results, prune, make the next kernel's jobs) run on it in the kernel's TAIL --
// one warp, lane = reference x quadrant (me_control.cuh) -- before it is written back.  There are no separate control
// launches and no round trip of the state between an engine and its control code.
constexpr int kSbHeadBytes = (int)offsetof(SbState, best_sad);
static_assert(kSbHeadBytes % 16 == 0 && sizeof(SbState) % 16 == 0, "the control part is copied as 16-byte words");
constexpr int kHeadVecs = kSbHeadBytes / 16;

__device__ __forceinline__ void head_load_async(unsigned char *head, const SbState *g) {
    const uint32_t sb = (uint32_t)__cvta_generic_to_shared(head);
    const uint4   *gp = reinterpret_cast<const uint4 *>(g);
    for (int i = threadIdx.x; i < kHeadVecs; i += blockDim.x)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sb + 16u * (uint32_t)i), "l"(gp + i));
    asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void head_wait() {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
}
// the job table has landed; true when any of its slots holds work (a block whose references all took an early exit has none)
template <class F>
__device__ __forceinline__ bool jobs_wait_any(int n, F used) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    return __syncthreads_or((int)threadIdx.x < n && used((int)threadIdx.x)) != 0;
}
__device__ __forceinline__ void head_store(const unsigned char *head, SbState *g) {
    const uint4 *l = reinterpret_cast<const uint4 *>(head);
    uint4       *gp = reinterpret_cast<uint4 *>(g);
    for (int i = threadIdx.x; i < kHeadVecs; i += blockDim.x) gp[i] = l[i];
}

// The search engines only need their job table: it is fetched into shared memory with 16-byte cp.async while the kernel
// loads its source block; results go back into the block's state in HBM/L2.
template <class J, int N>
__device__ __forceinline__ void jobs_load_async(J *dst, const J *src) {
    static_assert(sizeof(J) * N % 16 == 0, "job tables are copied as 16-byte words");
    const uint32_t sb = (uint32_t)__cvta_generic_to_shared(dst);
    const uint4   *gp = reinterpret_cast<const uint4 *>(src);
    for (int i = threadIdx.x; i < (int)(sizeof(J) * N / 16); i += kStThreads)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sb + 16u * (uint32_t)i), "l"(gp + i));
    asm volatile("cp.async.commit_group;" ::: "memory");
}

// ---- svt_nxm_sad_kernel jobs: one warp per job, straight from global memory (2 KB per operand) ----
__device__ void nxm_engine(SbState &s) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    // the used slots of the table, dealt round-robin to the warps
    const unsigned used = __ballot_sync(0xffffffffu, lane < kMaxNxmJobs && s.nxm_jobs[lane & (kMaxNxmJobs - 1)].rows > 0);
    if (threadIdx.x == 0) {
        unsigned long long px = 0;
        for (unsigned m = used; m; m &= m - 1) { const NxmJob &j = s.nxm_jobs[__ffs(m) - 1]; px += (unsigned long long)j.rows * j.width; }
        if (px) s.px += px;  // the only writer at this point (barriers on both sides of the step)
    }
    int idx = 0;
    for (unsigned m = used; m; m &= m - 1, idx++) {
        if (idx % nwarps != warp) continue;
        NxmJob        &j   = s.nxm_jobs[__ffs(m) - 1];
        const int      nw  = (j.width + 3) >> 2, rows = j.rows, width = j.width;
        const uint8_t *src = j.src, *ref = j.ref;
        const int      sst = j.src_stride, rst = j.ref_stride;
        const uint32_t tail = (width & 3) ? ((1u << (8 * (width & 3))) - 1u) : 0xffffffffu;
        uint32_t       acc = 0;
        if (width == 64 && (((uintptr_t)src | (uint32_t)sst | (uint32_t)rst) & 3) == 0) {
            // the common case (64 x 32 sub-sampled rows): lane = (column word, row parity); the loads of 8 rows are
            // independent and issued together.  The second reference word is read even when the row is word aligned --
            // it lies inside the padded plane -- so that the loop has no data-dependent branch.
            const int       c  = lane & 15, sh = 8 * (int)((uintptr_t)ref & 3);
            const uint8_t  *sp = src + (int64_t)(lane >> 4) * sst + 4 * c;
            const uint32_t *rp = reinterpret_cast<const uint32_t *>(ref - ((uintptr_t)ref & 3) + (int64_t)(lane >> 4) * rst) + c;
#pragma unroll 8
            for (int r = lane >> 4; r < rows; r += 2, sp += 2 * (int64_t)sst, rp += (int64_t)rst >> 1) {
                const uint32_t a = *reinterpret_cast<const uint32_t *>(sp), lo = rp[0], hi = rp[1];
                acc = sad4(a, __funnelshift_r(lo, hi, sh), acc);
            }
        } else if ((((uintptr_t)src | (uint32_t)sst) & 3) == 0) {
            // source rows are word aligned (always so on the ME path); reference words come from two aligned loads
            for (int i = lane; i < rows * nw; i += 32) {
                const int       r = i / nw, c = i - r * nw;
                const uint32_t  a = *reinterpret_cast<const uint32_t *>(src + (int64_t)r * sst + 4 * c);
                const uint8_t  *rp = ref + (int64_t)r * rst + 4 * c;
                const uint32_t *ra = reinterpret_cast<const uint32_t *>(rp - ((uintptr_t)rp & 3));
                const uint32_t  lo = ra[0], hi = (((uintptr_t)rp & 3) ? ra[1] : 0u);
                uint32_t        b  = __funnelshift_r(lo, hi, 8 * (int)((uintptr_t)rp & 3));
                const uint32_t m2 = (c == nw - 1) ? tail : 0xffffffffu;
                acc = sad4(a & m2, b & m2, acc);
            }
        } else {
            for (int i = lane; i < rows * nw; i += 32) {
                const int      r = i / nw, c = i - r * nw;
                const uint8_t *sp = src + (int64_t)r * sst + 4 * c, *rp = ref + (int64_t)r * rst + 4 * c;
                const int      nb = min(4, width - 4 * c);
                uint32_t       a = 0, b = 0;
                for (int q = 0; q < nb; q++) { a |= (uint32_t)sp[q] << (8 * q); b |= (uint32_t)rp[q] << (8 * q); }
                acc = sad4(a, b, acc);
            }
        }
        for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) j.sad = acc;
    }
}

// ---- the 1x1 probes of the 8x8-variance test (motion_estimation.c:1391-1412): one open_loop_me_fullpel_search_sblock call
// with a single candidate per reference, sub-sampled SAD.  One warp: lane = (16x16 block b in z-order, upper / lower pair
// of its 8x8 blocks); the operands come straight from global memory. ----
__device__ void probe_engine(const MeParams &p, SbState &s) {
    const int lane = threadIdx.x & 31;
    if (threadIdx.x >= 32) return;
    const int b = lane & 15, h = lane >> 4;
    const int bx = ((b >> 2) & 1) * 2 + (b & 1), by = (b >> 3) * 2 + ((b >> 1) & 1);
    SbState  &tab = *s.home;
    unsigned long long px = 0;
    for (int ji = 0; ji < kMaxFpJobs; ji++) {
        const FpJob &j = s.fp_jobs[ji];
        if (j.w <= 0 || j.h <= 0) continue;
        if (j.w != 1 || j.h != 1) continue;  // (not produced by is_b)
        px += 64ull * 32ull;
        // rows 8h, 8h + 2, 8h + 4, 8h + 6 of the lane's 16x16 block: 16 source bytes (aligned) against 16 reference bytes
        const uint8_t *sp = s.src + (int64_t)(16 * by + 8 * h) * s.src_stride + 16 * bx;
        const uint8_t *rp = j.ref + (int64_t)(16 * by + 8 * h) * j.ref_stride + 16 * bx;
        const int      a  = (int)((uintptr_t)rp & 3), sh = 8 * a;
        const uint32_t *ra = reinterpret_cast<const uint32_t *>(rp - a);
        uint32_t sad8[2] = {0u, 0u};
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const uint4     sv = *reinterpret_cast<const uint4 *>(sp + (int64_t)(2 * r) * s.src_stride);
            const uint32_t *rr = ra + (int64_t)(2 * r) * (j.ref_stride >> 2);
            const uint32_t  w0 = rr[0], w1 = rr[1], w2 = rr[2], w3 = rr[3], w4 = rr[4];  // the fifth word lies inside the padded plane
            sad8[0] = sad4(sv.x, __funnelshift_r(w0, w1, sh), sad8[0]);
            sad8[0] = sad4(sv.y, __funnelshift_r(w1, w2, sh), sad8[0]);
            sad8[1] = sad4(sv.z, __funnelshift_r(w2, w3, sh), sad8[1]);
            sad8[1] = sad4(sv.w, __funnelshift_r(w3, w4, sh), sad8[1]);
        }
        sad8[0] <<= 1; sad8[1] <<= 1;  // sub-sampled SAD: every second row, doubled
        uint32_t s16 = sad8[0] + sad8[1];
        s16 += __shfl_xor_sync(0xffffffffu, s16, 16);
        uint32_t s32 = s16 + __shfl_xor_sync(0xffffffffu, s16, 1);
        s32 += __shfl_xor_sync(0xffffffffu, s32, 2);
        uint32_t s64 = s32 + __shfl_xor_sync(0xffffffffu, s32, 4);
        s64 += __shfl_xor_sync(0xffffffffu, s64, 8);
        // running bests of the reference (strict '<': earlier results win ties)
        uint32_t      *bs = tab.best_sad[j.li][j.ri], *bm = tab.best_mv[j.li][j.ri];
        const uint32_t mv = ((uint32_t)(int32_t)j.yo << 16) | (uint16_t)j.xo;
        auto upd = [&](int pu, uint32_t sad) { if (sad < bs[pu]) { bs[pu] = sad; bm[pu] = mv; } };
        upd(21 + 4 * b + 2 * h, sad8[0]);
        upd(21 + 4 * b + 2 * h + 1, sad8[1]);
        if (h == 0) upd(5 + b, s16);
        if (lane < 16 && (b & 3) == 0) upd(1 + (b >> 2), s32);
        if (lane == 0) upd(0, s64);
    }
    if (lane == 0 && px) s.px += px;
    __threadfence_block();
}


// ---- svt_sad_loop_kernel jobs (pre-HME, HME level 0/1/2): one warp per job ----
// A lane evaluates 4 x-adjacent candidates (they share shared-memory words; alignments 1..3 by funnel shift).  When a
// job has fewer than 32 such items, the block rows of an item are split over L = 2^k lanes and summed by xor shuffles,
// so that 8x3 refinement windows still fill the warp.
template <int NW>
__device__ __forceinline__ void sad_rows(const uint32_t *rowp, uint32_t row_step_w, const uint32_t *blk, int r0, int rstride, int rows,
                                         int nw, uint32_t tailmask, uint32_t &a0, uint32_t &a1, uint32_t &a2, uint32_t &a3) {
    for (int r = r0; r < rows; r += rstride) {
        const uint32_t *rp = rowp + (uint32_t)r * row_step_w;
        const uint32_t *bp = blk + r * 16;
        uint32_t        w0 = rp[0];
        if (NW > 0) {
#pragma unroll
            for (int c = 0; c < NW; c++) {
                const uint32_t w1 = rp[c + 1], sw = bp[c];
                a0 = sad4(w0, sw, a0);
                a1 = sad4(fsr(w0, w1, 1), sw, a1);
                a2 = sad4(fsr(w0, w1, 2), sw, a2);
                a3 = sad4(fsr(w0, w1, 3), sw, a3);
                w0 = w1;
            }
        } else {
            const int nwfull = (tailmask == 0xffffffffu) ? nw : nw - 1;
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
    }
}

template <int NW>
__device__ void warp_sad_job(const SadJob &j, SadJob &jout, const uint32_t *blk, uint8_t *win, int lane, int rows, int bw, int rstep,
                             int win_bytes = kSadWarpWin) {
    const int W = j.sa_w, H = j.sa_h;
    if (W <= 0 || H <= 0) return;
    const bool     skip     = j.skip_line && bw == 16 && rows <= 16;
    const int      nw       = (bw + 3) >> 2;
    const uint32_t tailmask = (bw & 3) ? ((1u << (8 * (bw & 3))) - 1u) : 0xffffffffu;
    const int      ext_y    = (rows - 1) * rstep;
    const uint8_t *ref      = j.ref;
    const int      raw      = j.raw_stride;
    unsigned long long best = ~0ull;
    // widest tile whose rows (a few candidate rows + the block extent) still fit the warp's window buffer
    int nx_cap = (((win_bytes / (ext_y + min(H, 8))) & ~15) - 16) - 15 - (bw - 1) - 8;
    nx_cap     = max(4, min(nx_cap, kTileMaxNx));
    for (int x0 = 0; x0 < W; x0 += nx_cap) {
        const int      nx    = min(W - x0, nx_cap);
        const uint32_t mis   = (uint32_t)((uintptr_t)(ref + x0) & 15);
        uint32_t       pitch = (mis + nx + bw - 1 + 8 + 15) & ~15u;
        if (((pitch >> 4) & 1) == 0) pitch += 16;
        const int rows_fit = max(1, win_bytes / (int)pitch - ext_y);  // >= 1 by the choice of nx_cap
        const int ngroups  = (int)(((mis & 3) + nx + 3) >> 2);
        const uint32_t pitchw = pitch >> 2, nvec = pitch >> 4;
        for (int y0 = 0; y0 < H; y0 += rows_fit) {
            const int ny = min(H - y0, rows_fit);
            // stage rows y0 .. y0 + ny - 1 + ext_y of the window: 16-byte vectors from 16-byte aligned addresses
            const uint8_t *row0  = ref + (int64_t)y0 * raw + x0 - mis;
            const uint32_t total = nvec * (uint32_t)(ny + ext_y);
            __syncwarp();
            for (uint32_t i = lane; i < total; i += 32) {
                const uint32_t r = i / nvec, v = i - r * nvec;
                *reinterpret_cast<uint4 *>(win + r * pitch + (v << 4)) = __ldg(reinterpret_cast<const uint4 *>(row0 + (int64_t)r * raw) + v);
            }
            __syncwarp();
            const int nyeff = skip ? ((ny + (y0 & 1)) >> 1) : ny;  // odd candidate rows only
            const int items = nyeff * ngroups;
            int       L     = 1;
            while (L < 32 && items * (L * 2) <= 32 && L * 2 <= rows) L *= 2;
            const int per_round = 32 / L;
            const int slice = lane & (L - 1), sub = lane / L;
            for (int base = 0; base < items; base += per_round) {
                const int  item   = base + sub;
                const bool active = item < items;
                const int  yi = active ? item / ngroups : 0, g = active ? item - yi * ngroups : 0;
                const int  yy = skip ? (2 * yi + ((y0 & 1) ? 0 : 1)) : yi;
                uint32_t   a0 = 0, a1 = 0, a2 = 0, a3 = 0;
                if (active) {
                    const uint32_t *rowp = reinterpret_cast<const uint32_t *>(win) + (uint32_t)yy * pitchw + (mis >> 2) + g;
                    sad_rows<NW>(rowp, (uint32_t)rstep * pitchw, blk, slice, L, rows, nw, tailmask, a0, a1, a2, a3);
                }
                for (int o = L >> 1; o; o >>= 1) {
                    a0 += __shfl_xor_sync(0xffffffffu, a0, o); a1 += __shfl_xor_sync(0xffffffffu, a1, o);
                    a2 += __shfl_xor_sync(0xffffffffu, a2, o); a3 += __shfl_xor_sync(0xffffffffu, a3, o);
                }
                if (active && slice == 0) {
                    const int      xb = x0 + 4 * g - (int)(mis & 3), y = y0 + yy;
                    const uint32_t acc[4] = {a0, a1, a2, a3};
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        const int x = xb + c;
                        if (x >= x0 && x < x0 + nx) {
                            const unsigned long long key = ((unsigned long long)acc[c] << 32) | ((uint32_t)y << 16) | (uint32_t)x;
                            best = key < best ? key : best;
                        }
                    }
                }
            }
        }
    }
    for (int o = 16; o; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
        best = other < best ? other : best;
    }
    if (lane == 0 && best != ~0ull && (uint32_t)(best >> 32) < 0xffffffu) {  // strict '<' against the 0xffffff start value
        jout.found = 1; jout.best_sad = (uint32_t)(best >> 32);
        jout.by = (int16_t)((best >> 16) & 0xffff); jout.bx = (int16_t)(best & 0xffff);
    }
}

// Window staging for the specialised paths.  Rows are staged word by word (4-byte cp.async, all copies of the warp in
// flight together) with an ODD row pitch in words, and -- when the block samples every second reference row
// (sub-sampled SAD) -- with even and odd window rows in two separate planes, so that the rows a candidate reads are
// consecutive in shared memory.  Lanes that read the same column of different block rows then hit 32 different banks
// (a 16-byte aligned pitch gave 4- to 8-way conflicts: ncu l1tex__data_bank_conflicts of the first staged version).
struct WinLayout {
    uint32_t S;            // row pitch in words (odd)
    uint32_t plane_words;  // offset of the odd-row plane (0 when rows are not split)
    int      a;            // byte offset of candidate x0 inside the first staged word
};
__device__ __forceinline__ uint32_t win_row_word(const WinLayout &wl, int R, bool split) {
    return split ? (uint32_t)(R & 1) * wl.plane_words + (uint32_t)(R >> 1) * wl.S : (uint32_t)R * wl.S;
}
__device__ __forceinline__ void warp_stage_words(uint32_t *win, const uint8_t *start, int raw_stride, int nrows, int nwords,
                                                 const WinLayout &wl, bool split, int lane) {
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(win);
    // 2^q lanes per row (smallest power of two >= nwords, at most 32): a copy instruction covers 32 >> q rows, and the
    // per-copy address arithmetic is a handful of adds (no divisions, no 64-bit multiplies)
    int q = 5;
    while (q > 0 && (1 << (q - 1)) >= nwords) q--;
    const int rpi = 32 >> q, lr = lane >> q, lw = lane & ((1 << q) - 1);
    const int64_t gstep = (int64_t)rpi * raw_stride;
    for (int wd = lw; wd < nwords; wd += 32) {  // a second pass only for rows wider than 32 words
        const uint8_t *g = start + (int64_t)lr * raw_stride + 4 * wd;
        if (!split) {
            uint32_t sa = sbase + 4u * ((uint32_t)lr * wl.S + (uint32_t)wd);
            for (int R = lr; R < nrows; R += rpi, g += gstep, sa += 4u * (uint32_t)rpi * wl.S)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(g));
        } else if (rpi >= 2) {  // the parity of the lane's rows is fixed
            uint32_t sa = sbase + 4u * ((uint32_t)(lr & 1) * wl.plane_words + (uint32_t)(lr >> 1) * wl.S + (uint32_t)wd);
            for (int R = lr; R < nrows; R += rpi, g += gstep, sa += 4u * (uint32_t)(rpi >> 1) * wl.S)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(g));
        } else {
            uint32_t sa = sbase + 4u * (uint32_t)wd;
            for (int R = 0; R < nrows; R += 2, g += 2 * gstep, sa += 4u * wl.S) {
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(g));
                if (R + 1 < nrows) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa + 4u * wl.plane_words), "l"(g + gstep));
            }
        }
    }
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
    __syncwarp();
}

// Full-size blocks (every SB that does not touch the right/bottom picture edge): block words of the lane live in
// registers, the row loop is fully unrolled.  L lanes share an item (4 x-adjacent candidates):
//   sixteenth plane 16x8 rows : L = 1  (windows are large: >= 32 items)         32 source words per lane
//   quarter plane  32x16 rows : L = 16 (8x3 refinement windows: 6..9 items)      8 source words per lane (one row)
//   full plane     64x32 rows : L = 32                                          16 source words per lane (one row)
template <int NW, int ROWS, int L>
__device__ void warp_sad_job_fast(const SadJob &j, SadJob &jout, const uint32_t *blk, uint32_t *win, int lane, int rstep) {
    constexpr int RPL = ROWS / L, PER_ROUND = 32 / L;
    const int W = j.sa_w, H = j.sa_h;
    if (W <= 0 || H <= 0) return;
    constexpr int bw = NW * 4;
    const bool     skip  = j.skip_line && bw == 16 && ROWS <= 16;
    const bool     split = rstep == 2;
    const int      ext_y = (ROWS - 1) * rstep;
    const uint8_t *ref   = j.ref;
    const int      raw   = j.raw_stride;
    const int      slice = lane & (L - 1), sub = lane / L;
    uint32_t       sreg[RPL][NW];
#pragma unroll
    for (int k = 0; k < RPL; k++)
#pragma unroll
        for (int c = 0; c < NW; c++) sreg[k][c] = blk[(slice + L * k) * 16 + c];
    unsigned long long best = ~0ull;
    constexpr int kWinWords = kSadWarpWin / 4;
    // widest tile whose rows (a few candidate rows + the block extent) still fit the warp's window buffer
    int nx_cap = (kWinWords / (ext_y + min(H, 8) + 1) - 3) * 4 - bw - 8;
    nx_cap     = max(4, min(nx_cap, kTileMaxNx));
    for (int x0 = 0; x0 < W; x0 += nx_cap) {
        const int nx = min(W - x0, nx_cap);
        WinLayout wl;
        wl.a = (int)((uintptr_t)(ref + x0) & 3);
        const int nwords  = (wl.a + nx + bw - 1 + 3) / 4 + 1;  // +1: the shifted loads touch one word past the last candidate
        wl.S              = (uint32_t)nwords | 1u;
        const int rows_fit = max(1, (kWinWords / (int)wl.S) - ext_y - 2);
        const int ngroups  = (wl.a + nx + 3) >> 2;
        const float inv_ng = 1.0f / (float)ngroups;
        for (int y0 = 0; y0 < H; y0 += rows_fit) {
            const int ny = min(H - y0, rows_fit), nrows = ny + ext_y;
            wl.plane_words = split ? (uint32_t)((nrows + 1) >> 1) * wl.S : 0u;
            __syncwarp();
            warp_stage_words(win, ref + (int64_t)y0 * raw + x0 - wl.a, raw, nrows, nwords, wl, split, lane);
            const int nyeff = skip ? ((ny + (y0 & 1)) >> 1) : ny;
            const int items = nyeff * ngroups;
            for (int base = 0; base < items; base += PER_ROUND) {
                const int  item   = base + sub;
                const bool active = item < items;
                const int  yi = active ? (int)(((float)item + 0.5f) * inv_ng) : 0, g = active ? item - yi * ngroups : 0;
                const int  yy = skip ? (2 * yi + ((y0 & 1) ? 0 : 1)) : yi;
                uint32_t   a0 = 0, a1 = 0, a2 = 0, a3 = 0;
                if (active) {
                    // block row r of candidate row yy is window row yy + r * rstep: consecutive rows of one plane
                    const uint32_t row_step = split ? wl.S : (uint32_t)rstep * wl.S;
                    const uint32_t *rowp = win + win_row_word(wl, yy, split) + (uint32_t)slice * row_step + g;
#pragma unroll
                    for (int k = 0; k < RPL; k++) {
                        const uint32_t *rp = rowp + (uint32_t)(L * k) * row_step;
                        uint32_t        w[NW + 1];
#pragma unroll
                        for (int c = 0; c <= NW; c++) w[c] = rp[c];
#pragma unroll
                        for (int c = 0; c < NW; c++) {
                            const uint32_t sw = sreg[k][c];
                            a0 = sad4(w[c], sw, a0);
                            a1 = sad4(fsr(w[c], w[c + 1], 1), sw, a1);
                            a2 = sad4(fsr(w[c], w[c + 1], 2), sw, a2);
                            a3 = sad4(fsr(w[c], w[c + 1], 3), sw, a3);
                        }
                    }
                }
#pragma unroll
                for (int o = L >> 1; o; o >>= 1) {
                    a0 += __shfl_xor_sync(0xffffffffu, a0, o); a1 += __shfl_xor_sync(0xffffffffu, a1, o);
                    a2 += __shfl_xor_sync(0xffffffffu, a2, o); a3 += __shfl_xor_sync(0xffffffffu, a3, o);
                }
                if (active && slice == 0) {
                    const int      xb = x0 + 4 * g - wl.a;
                    const uint32_t yk = (uint32_t)(y0 + yy) << 16;
                    const uint32_t acc[4] = {a0, a1, a2, a3};
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        const int x = xb + c;
                        if (x >= x0 && x < x0 + nx) {
                            const unsigned long long key = ((unsigned long long)acc[c] << 32) | yk | (uint32_t)x;
                            best = key < best ? key : best;
                        }
                    }
                }
            }
        }
    }
    for (int o = 16; o; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
        best = other < best ? other : best;
    }
    if (lane == 0 && best != ~0ull && (uint32_t)(best >> 32) < 0xffffffu) {
        jout.found = 1; jout.best_sad = (uint32_t)(best >> 32);
        jout.by = (int16_t)((best >> 16) & 0xffff); jout.bx = (int16_t)(best & 0xffff);
    }
}

// mbarrier / TMA primitives (PTX ISA: mbarrier, cp.async.bulk.tensor)
SVTME_D void mbar_init(uint32_t mbar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\nfence.mbarrier_init.release.cluster;" ::"r"(mbar), "r"(count) : "memory");
}
SVTME_D void mbar_expect_tx(uint32_t mbar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
SVTME_D void mbar_wait(uint32_t mbar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(mbar), "r"(parity) : "memory");
    } while (!done);
}
// one box of a 2-D tiled tensor map -> shared memory (dst 128-byte aligned), completion counted on `mbar`
SVTME_D void tma_load_2d(uint32_t dst, const void *tmap, int x, int y, uint32_t mbar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(tmap), "r"(x), "r"(y), "r"(mbar) : "memory");
}

// ---- HME level 1 / 2 refinement windows (8x3 candidates on 32x16 / 64x32 sub-sampled blocks): lane = one block row ----
// The search job is tiny (24 candidates) but every candidate touches the whole block, and each window row is used by
// at most two candidate rows -- so shared memory is only the transposer between coalesced global rows and "one row
// per lane".  The lane pulls its window row into registers ONCE per candidate row, normalises the byte alignment with
// one funnel shift per word and evaluates 8 x-candidates from registers; the 8 partial SADs are summed over the rows
// with a transposing butterfly (9 shuffles instead of 40).  ROWS lanes form a group; 32 / ROWS candidate rows run at
// once.  Sub-sampled SAD only (block row r = window row 2r): even and odd window rows live in two planes with an odd
// word pitch, so the 32 rows read by a warp fall into 32 different banks.
template <int NW, int ROWS>
struct RowsCfg {
    // a window row = 16-byte chunks from the 16-byte aligned address below the first candidate:
    // misalignment (<= 15) + 7 more candidates + block width, rounded up
    static constexpr int kChunks  = (15 + 7 + 4 * NW + 15) / 16;  // 6 (64-wide blocks) / 4 (32-wide)
    static constexpr int kSlots   = kChunks <= 4 ? 4 : 8;         // lanes per staged row
    static constexpr int kRpi     = 32 / kSlots;                  // rows per copy instruction (even: a lane keeps its row parity)
    static constexpr int kS       = 4 * (kChunks | 1);            // pitch in words, an ODD number of chunks: 8 consecutive rows
                                                                  // of a plane start in 8 different 16-byte bank groups
    static constexpr int kExtY    = (ROWS - 1) * 2;
    static constexpr int kNy      = 4;                            // candidate rows per staged chunk
    static constexpr int kRowsPad = (kNy + kExtY + 7) & ~7;       // staged in groups of 8 rows
    static constexpr int kPlaneW  = ((kRowsPad / 2) * kS + 31) & ~31;  // words per parity plane, a multiple of 128 bytes: the odd
                                                                  // plane is a legal TMA destination
    static constexpr int kWinW    = 2 * kPlaneW;
    static_assert(kRowsPad % kRpi == 0 && kRpi % 2 == 0, "copy groups");
};

// WO = (address of candidate 0) / 4 mod 4: which word of the 16-byte aligned row holds candidate 0 (static register indexing)
template <int NW, int ROWS, bool FULLW, int WO>
__device__ __forceinline__ void rows_unit(const uint32_t (&sreg)[NW], const uint32_t *win, int lane, int rows, int nw_eff, int a,
                                          int x0, int y0, int wx, int ny, unsigned long long &best) {
    using C = RowsCfg<NW, ROWS>;
    constexpr int SUBS = 32 / ROWS, MAXW = NW + 3;
    static_assert(WO + MAXW < 4 * C::kChunks, "the shifted-in word is inside the staged chunks");
    const int slice = lane & (ROWS - 1), sub = lane / ROWS, shift = 8 * a;
    // which of the 8 candidates this lane owns after the transposing reduction
    const int own = ROWS == 32 ? (((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1))
                               : (((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1));
    for (int yb = 0; yb < ny; yb += SUBS) {
        const int  yy     = yb + sub;
        const bool active = yy < ny && slice < rows;
        uint32_t   acc[8];
#pragma unroll
        for (int i = 0; i < 8; i++) acc[i] = 0;
        if (active) {
            const uint4 *rp = reinterpret_cast<const uint4 *>(win + (yy & 1) * C::kPlaneW + ((yy >> 1) + slice) * C::kS);
            uint32_t     v[4 * C::kChunks];
#pragma unroll
            for (int c = 0; c < C::kChunks; c++) {
                const uint4 q = rp[c];
                v[4 * c] = q.x; v[4 * c + 1] = q.y; v[4 * c + 2] = q.z; v[4 * c + 3] = q.w;
            }
            uint32_t w[MAXW + 1];
#pragma unroll
            for (int c = 0; c < MAXW; c++) w[c] = __funnelshift_r(v[WO + c], v[WO + c + 1], shift);  // candidate 0 now starts at byte 0
            w[MAXW] = 0;
#pragma unroll
            for (int g = 0; g < 2; g++)
#pragma unroll
                for (int c = 0; c < NW; c++) {
                    if (FULLW || c < nw_eff) {
                        const uint32_t sw = sreg[c], lo = w[g + c], hi = w[g + c + 1];
                        acc[4 * g + 0] = sad4(lo, sw, acc[4 * g + 0]);
                        acc[4 * g + 1] = sad4(fsr(lo, hi, 1), sw, acc[4 * g + 1]);
                        acc[4 * g + 2] = sad4(fsr(lo, hi, 2), sw, acc[4 * g + 2]);
                        acc[4 * g + 3] = sad4(fsr(lo, hi, 3), sw, acc[4 * g + 3]);
                    }
                }
        }
        // transposing butterfly over the ROWS lanes of the group: 8 -> 4 -> 2 -> 1 values per lane, then plain sums
        uint32_t v4[4], v2[2], v1;
        {
            const bool up = lane & (ROWS / 2);
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const uint32_t send = up ? acc[i] : acc[4 + i], keep = up ? acc[4 + i] : acc[i];
                v4[i] = keep + __shfl_xor_sync(0xffffffffu, send, ROWS / 2);
            }
        }
        {
            const bool up = lane & (ROWS / 4);
#pragma unroll
            for (int i = 0; i < 2; i++) {
                const uint32_t send = up ? v4[i] : v4[2 + i], keep = up ? v4[2 + i] : v4[i];
                v2[i] = keep + __shfl_xor_sync(0xffffffffu, send, ROWS / 4);
            }
        }
        {
            const bool     up   = lane & (ROWS / 8);
            const uint32_t send = up ? v2[0] : v2[1], keep = up ? v2[1] : v2[0];
            v1 = keep + __shfl_xor_sync(0xffffffffu, send, ROWS / 8);
        }
#pragma unroll
        for (int o = ROWS / 16; o; o >>= 1) v1 += __shfl_xor_sync(0xffffffffu, v1, o);
        if (yy < ny && own < wx && (lane & (ROWS / 8 - 1)) == 0) {
            const unsigned long long key = ((unsigned long long)v1 << 32) | ((uint32_t)(y0 + yy) << 16) | (uint32_t)(x0 + own);
            best = key < best ? key : best;
        }
    }
}

template <int NW, int ROWS, bool FULLW>
__device__ __forceinline__ void rows_job(const MeParams &p, int lvl, const SadJob &j, SadJob &jout, const uint32_t (&sreg)[NW],
                                         uint32_t *win, int lane, int rows, int nw_eff, const RowsMaps *maps = nullptr, uint32_t mbar = 0,
                                         uint32_t *phase = nullptr) {
    using C = RowsCfg<NW, ROWS>;
    const int W = j.sa_w, H = j.sa_h;
    if (W <= 0 || H <= 0) return;
    const uint8_t *ref   = j.ref;
    const uint32_t raw   = (uint32_t)j.raw_stride;
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(win);
    // staging map of this lane: row lr of every group of kRpi rows, chunk lc (16-byte copies)
    const int      lr = lane / C::kSlots, lc = lane % C::kSlots;
    const uint32_t sofs = (uint32_t)(((lr & 1) * C::kPlaneW + (lr >> 1) * C::kS) * 4 + 16 * lc);
    unsigned long long best = ~0ull;
    for (int x0 = 0; x0 < W; x0 += 8) {
        const int a16 = (int)((uintptr_t)(ref + x0) & 15), a = a16 & 3, wo = a16 >> 2, wx = min(8, W - x0);
        for (int y0 = 0; y0 < H; y0 += C::kNy) {
            const int      ny = min(H - y0, C::kNy);
            const uint8_t *g0 = ref + (int64_t)y0 * (int64_t)raw + x0 - a16 + 16 * lc;
            __syncwarp();
            if (maps) {
                // the two parity planes of the window as two TMA boxes (kS * 4 bytes x kRowsPad / 2 rows of the same-parity rows of
                // the padded plane, from the picture's parity-view descriptors); completion on this warp's mbarrier.  Rows or bytes
                // beyond the plane's allocation are zero-filled by the TMA unit; they belong to no candidate.
                constexpr int L = 0;  // the full plane: only the HME level-2 kernel is given descriptors
                static_assert(NW != 16 || (C::kS * 4 == kRowsTmaBoxW && C::kRowsPad / 2 == kRowsTmaBoxH), "box = one parity plane");
                const DevPlane &rpl = p.refs[j.li][j.ri].lvl[L];
                const uint8_t  *row0 = ref + (int64_t)y0 * (int64_t)raw + x0 - a16;  // 16-byte aligned
                constexpr int   kPadL = kDevPad[L], kBoxBytes = C::kS * 4 * (C::kRowsPad / 2);
                const int       offb = (int)(row0 - (rpl.org - (int64_t)kPadL * rpl.stride - kPadL));
                const int       ty = offb / rpl.stride, tx = offb - ty * rpl.stride;
                if (lane == 0) mbar_expect_tx(mbar, 2u * (uint32_t)kBoxBytes);
                __syncwarp();
                if (lane < 2) {  // window rows of parity `lane` = plane rows of parity (ty + lane) & 1, from row (ty + lane) >> 1 of that view
                    const int pr = ty + lane;
                    tma_load_2d(sbase + (uint32_t)lane * (uint32_t)(C::kPlaneW * 4),
                                maps->m[j.li * kMaxRefs + j.ri][pr & 1], tx, pr >> 1, mbar);
                }
                mbar_wait(mbar, *phase & 1u);
                (*phase)++;
            } else
            // always the full kRowsPad rows: the few rows below a short window are inside the plane's physical padding
            // (96 / 48 rows against the 72 / 32 the search may touch), and the copy loop unrolls to immediates
            if (lc < C::kChunks) {
#pragma unroll
                for (int G = 0; G < C::kRowsPad / C::kRpi; G++) {
                    CHECK_SRC(p, lvl, j.li, j.ri, g0 + (uint64_t)(uint32_t)(G * C::kRpi + lr) * (uint64_t)raw, 16, 1);
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sbase + sofs + (uint32_t)(G * (C::kRpi / 2) * C::kS * 4)),
                                 "l"(g0 + (uint64_t)(uint32_t)(G * C::kRpi + lr) * (uint64_t)raw));
                }
            }
            asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
            __syncwarp();
            switch (wo) {
            case 0: rows_unit<NW, ROWS, FULLW, 0>(sreg, win, lane, rows, nw_eff, a, x0, y0, wx, ny, best); break;
            case 1: rows_unit<NW, ROWS, FULLW, 1>(sreg, win, lane, rows, nw_eff, a, x0, y0, wx, ny, best); break;
            case 2: rows_unit<NW, ROWS, FULLW, 2>(sreg, win, lane, rows, nw_eff, a, x0, y0, wx, ny, best); break;
            default: rows_unit<NW, ROWS, FULLW, 3>(sreg, win, lane, rows, nw_eff, a, x0, y0, wx, ny, best); break;
            }
        }
    }
    for (int o = 16; o; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
        best = other < best ? other : best;
    }
    if (lane == 0 && best != ~0ull && (uint32_t)(best >> 32) < 0xffffffu) {  // strict '<' against the 0xffffff start value
        jout.found = 1; jout.best_sad = (uint32_t)(best >> 32);
        jout.by = (int16_t)((best >> 16) & 0xffff); jout.bx = (int16_t)(best & 0xffff);
    }
}

// `lvl` 1: quarter plane, <8, 16>; `lvl` 0: full plane, <16, 32>.  Launched only when hme_search_method is sub-sampled.
#ifndef SVTME_ROWS8_CTAS
#define SVTME_ROWS8_CTAS 6
#endif
template <int NW, int ROWS>
__device__ __forceinline__ void sad_rows_block(const MeParams &p, int lvl, uint32_t sbi, const RowsMaps *maps = nullptr, uint32_t mbar = 0,
                                               uint32_t *phase = nullptr) {
    using C = RowsCfg<NW, ROWS>;
    constexpr int kBlkPitch = 17;  // odd: lane = row reads are conflict free
    __shared__ uint32_t                    blk[ROWS * kBlkPitch];
    __shared__ __align__(128) uint32_t     win[kStWarps][C::kWinW];
    __shared__ __align__(16) SadJob        jobs[kMaxSadJobs];
    SbState  &s = p.states[sbi];
    PH_DECL
    jobs_load_async<SadJob, kMaxSadJobs>(jobs, s.sad_jobs);
    constexpr int n = kMaxSadJobs;  // fixed slots, unused ones have empty windows
    // block geometry as in make_sad_job (me_control.cuh)
    const DevPlane &cp = p.cur.lvl[lvl];
    const uint32_t  aw = (uint32_t)((p.cur.lvl[0].width + 7) & ~7), ah = (uint32_t)((p.cur.lvl[0].height + 7) & ~7);
    const uint32_t  org_x = (sbi % p.wb) * 64, org_y = (sbi / p.wb) * 64;
    const uint32_t  b64_w = min(64u, aw - org_x), b64_h = min(64u, ah - org_y);
    const int       bw = (int)(b64_w >> lvl), rows = (int)((b64_h >> lvl) >> 1), nw_eff = bw >> 2;  // bw is a multiple of 4
    {   // every second row of the source block (same for every job of the SB)
        const uint8_t *b0 = cp.org + (int64_t)(org_y >> lvl) * cp.stride + (org_x >> lvl);
        for (int i = threadIdx.x; i < ROWS * NW; i += blockDim.x) {
            const int r = i / NW, c = i - r * NW;
            blk[r * kBlkPitch + c] = (r < rows && c < nw_eff) ? __ldg(reinterpret_cast<const uint32_t *>(b0 + (int64_t)(2 * r) * cp.stride) + c) : 0u;
        }
    }
    if (!jobs_wait_any(n, [&](int i) { return jobs[i].sa_w > 0 && jobs[i].sa_h > 0; })) return;
    PH(3, 0)  // job fetch + source block
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t  sreg[NW];
#pragma unroll
    for (int c = 0; c < NW; c++) sreg[c] = blk[(lane & (ROWS - 1)) * kBlkPitch + c];
    uint32_t px = 0;
    for (int ji = warp; ji < n; ji += kStWarps) {
        const SadJob &j = jobs[ji];
        if (j.sa_w > 0 && j.sa_h > 0) px += (uint32_t)(j.sa_w * j.sa_h) * (uint32_t)(rows * bw);
        if (nw_eff == NW) rows_job<NW, ROWS, true>(p, lvl, j, s.sad_jobs[ji], sreg, win[warp], lane, rows, nw_eff, maps, mbar, phase);
        else rows_job<NW, ROWS, false>(p, lvl, j, s.sad_jobs[ji], sreg, win[warp], lane, rows, nw_eff, maps, mbar, phase);
    }
    if (lane == 0 && px) atomicAdd(&s.px, (unsigned long long)px);
    PH(3, 1)  // this warp's windows
}
// The engine kernels are PERSISTENT: a CTA walks the picture's blocks with a grid stride.  The grid is sized to a share
// of the GPU's CTA slots (launch_me_picture_staged), so that the engines of two pictures in flight on different
// streams are resident side by side instead of one after the other.
template <int NW, int ROWS>
__global__ void __launch_bounds__(kStThreads, NW == 8 ? SVTME_ROWS8_CTAS : 6) st_sad_rows_kernel(const __grid_constant__ MeParams p, const __grid_constant__ RowsMaps maps, int lvl) {
    PDL_ENTRY();
    __shared__ __align__(8) unsigned long long mbar_store[kStWarps];  // one mbarrier per warp: a warp stages and searches its own windows
    const int      warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(&mbar_store[warp]);
    if (lane == 0) mbar_init(mbar, 1);
    __syncthreads();
    uint32_t       phase = 0;
    const uint32_t n_sb = p.wb * p.hb;
    for (uint32_t sbi = blockIdx.x; sbi < n_sb; sbi += gridDim.x) {
        sad_rows_block<NW, ROWS>(p, lvl, sbi, NW == 16 ? &maps : nullptr, mbar, &phase);  // HME level 1 keeps its cp.async copies
        __syncthreads();  // the shared-memory tables are reused by the next block
    }
}

// `lvl`: plane of this phase (2 sixteenth: pre-HME / level 0, 1 quarter: level 1, 0 full: level 2).  The source block is
// derived from the block index alone, so its staging overlaps the fetch of the job descriptors.
__device__ __forceinline__ void sad_block(const MeParams &p, int lvl, uint32_t sbi) {
    __shared__ __align__(16) uint32_t blk[64 * 16];
    __shared__ __align__(16) uint8_t  win[kStWarps][kSadWarpWin];
    __shared__ __align__(16) SadJob   jobs[kMaxSadJobs];
    SbState  &s = p.states[sbi];
    jobs_load_async<SadJob, kMaxSadJobs>(jobs, s.sad_jobs);
    constexpr int n = kMaxSadJobs;  // fixed slots, unused ones have empty windows
    // block geometry as in make_sad_job (me_control.cuh)
    const DevPlane &cp   = p.cur.lvl[lvl];
    const bool      full = p.k.hme_search_method == 1;
    const uint32_t  aw = (uint32_t)((p.cur.lvl[0].width + 7) & ~7), ah = (uint32_t)((p.cur.lvl[0].height + 7) & ~7);
    const uint32_t  org_x = (sbi % p.wb) * 64, org_y = (sbi / p.wb) * 64;
    const uint32_t  b64_w = min(64u, aw - org_x), b64_h = min(64u, ah - org_y);
    const int       bw = (int)(b64_w >> lvl), rows = (int)(full ? (b64_h >> lvl) : ((b64_h >> lvl) >> 1)), nw = (bw + 3) >> 2;
    const int       rstep = full ? 1 : 2;
    {   // the source block of this phase (same for every job of the SB), zero beyond block_w
        const uint8_t *b0 = cp.org + (int64_t)(org_y >> lvl) * cp.stride + (org_x >> lvl);
        const int      bs = full ? cp.stride : cp.stride * 2;
        for (int i = threadIdx.x; i < rows * nw; i += blockDim.x) {
            const int      r = i / nw, c = i - r * nw;
            const uint8_t *sp = b0 + (int64_t)r * bs + 4 * c;
            uint32_t       a;
            if (4 * c + 4 <= bw && (((uintptr_t)sp) & 3) == 0) a = *reinterpret_cast<const uint32_t *>(sp);
            else {
                a = 0;
                for (int q = 0; q < min(4, bw - 4 * c); q++) a |= (uint32_t)sp[q] << (8 * q);
            }
            blk[r * 16 + c] = a;
        }
    }
    if (!jobs_wait_any(n, [&](int i) { return jobs[i].sa_w > 0 && jobs[i].sa_h > 0; })) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned long long px = 0;
    for (int ji = warp; ji < n; ji += kStWarps) {
        const SadJob &j  = jobs[ji];
        SadJob       &jo = s.sad_jobs[ji];
        uint32_t     *w32 = reinterpret_cast<uint32_t *>(win[warp]);
        if (j.sa_w > 0 && j.sa_h > 0) {
            const bool skip = j.skip_line && bw == 16 && rows <= 16;
            px += (unsigned long long)j.sa_w * (skip ? (j.sa_h >> 1) : j.sa_h) * rows * bw;
        }
        if (bw == 16 && rows == 8) warp_sad_job_fast<4, 8, 1>(j, jo, blk, w32, lane, rstep);
        else if (bw == 32 && rows == 16 && j.sa_w * j.sa_h <= 64) warp_sad_job_fast<8, 16, 16>(j, jo, blk, w32, lane, rstep);
        else if (bw == 64 && rows == 32 && j.sa_w * j.sa_h <= 64) warp_sad_job_fast<16, 32, 32>(j, jo, blk, w32, lane, rstep);
        else if (bw == 16) warp_sad_job<4>(j, jo, blk, win[warp], lane, rows, bw, rstep);
        else if (bw == 32) warp_sad_job<8>(j, jo, blk, win[warp], lane, rows, bw, rstep);
        else if (bw == 64) warp_sad_job<16>(j, jo, blk, win[warp], lane, rows, bw, rstep);
        else warp_sad_job<0>(j, jo, blk, win[warp], lane, rows, bw, rstep);
    }
    if (lane == 0 && px) atomicAdd(&s.px, px);
}
__global__ void __launch_bounds__(kStThreads) st_sad_kernel(const __grid_constant__ MeParams p, int lvl) {
    PDL_ENTRY();
    const uint32_t n_sb = p.wb * p.hb;
    for (uint32_t sbi = blockIdx.x; sbi < n_sb; sbi += gridDim.x) {
        sad_block(p, lvl, sbi);
        __syncthreads();
    }
}

// ---- pre-HME / HME level 0 on the sixteenth plane (16x16 blocks, 8 sub-sampled rows): WARP-autonomous, lane = item ----
// The job windows of a block are cut into SUB-WINDOWS (a column tile of <= 128 candidates x a row strip) that fit a
// warp's private window buffer; the cut (the "plan") is made by the one-warp light kernel in front of this engine, which
// also appends one UNIT per sub-window to the picture's unit queue.  Here every warp of a persistent grid pulls units
// from that queue: it stages the sub-window by TMA (2-D tiled copies of the reference's padded sixteenth plane: boxes of
// 8 or 32 rows x 48 .. 176 bytes, one instruction per box, completion on the warp's own mbarrier -- nothing of the copy
// is executed by the SM's own pipes), walks its items, reduces the minimum over its lanes and folds it into the job's
// 64-bit key (sad, y, x) in HBM/L2 with one atomic.  No CTA-wide barrier, plan or batch exists in this kernel: a warp
// that waits for its copy leaves the SM to the other warps, and blocks with much work and blocks with little are dealt
// to the warps by the queue.  (The CTA-per-block version of this kernel spent 40-50 % of a CTA's life outside the
// search: plan, copy issue and wait, batch barriers -- with the registers and shared memory of four warps held.)
// A box must start on a 16-byte boundary of the plane (the TMA unit raises an illegal-instruction error otherwise:
// tools/microbench/tma_u8.cu), so candidate 0 sits at byte 4 * wo + a of a staged row: groups start at word wo,
// candidates are shifted by a bytes.  An item = 4 x-adjacent candidates on the two candidate rows y and y + 2, which
// share 7 of their 8 reference rows: 9 window rows (45 LDS, 108 SHF) per 256 SADs.  With prehme_skip_search_line only odd
// candidate rows exist (y = 4k + 1 and 4k + 3).  The row pitch is 4 x odd words (8 consecutive rows fall on 8 different
// bank quads).  Blocks cut by the picture edge take the generic per-warp path, one unit per job.
#ifndef SVTME_S16_NYCAP
#define SVTME_S16_NYCAP 448  // candidate rows of a sub-window at most (a unit is searched by ONE warp)
#endif
constexpr int kS16BufWords = 2304;  // per warp: 9 KB
constexpr int kS16MaxSub   = 96;
struct alignas(16) S16Sub {
    int16_t job, x0, y0, nx, ny, ngroups, P, nbox;  // P: row pitch in words = box bytes / 4; nbox: boxes of `brows` rows
    int16_t a, wo;                                  // candidate 0 at byte 4 * wo + a of a staged row
    int32_t items;
    int32_t tx, ty;                                 // the sub-window's first byte in the padded plane (TMA coordinates)
    float   inv_ng, inv_nx;
    int16_t hsel, cls, brows, ref;                  // hsel 1: odd rows only; cls: descriptor index; brows: rows per box; ref: list * 4 + index
};
static_assert(sizeof(S16Sub) == 48, "a sub-window descriptor is fetched as three 16-byte words");

constexpr int kPad16 = kDevPad[2];

// The plan of one block's pre-HME / HME level-0 searches, made by the light kernel in front of the engine (s16_plan) and
// handed over through HBM/L2; the engine's warps fold their minima into job_key, the light kernel behind it reads them.
struct alignas(16) S16Plan {
    unsigned long long job_key[kMaxSadJobs];  // (sad << 32 | y << 16 | x) minimum of each job's window, ~0: nothing found
    int32_t n_sub_all, generic, pad_[2];      // generic: the block takes the generic path (picture edge, too many sub-windows)
    S16Sub  subs[kS16MaxSub];
};
static_assert(sizeof(S16Plan) % 16 == 0 && offsetof(S16Plan, subs) % 16 == 0, "plans are copied as 16-byte words");

// unit of the queue: block index << 8 | generic << 7 | sub-window index (generic: job slot)
// one warp (threads 0 .. 31 of the CTA): lane = job.  Sub-windows = (column tile of <= 128 candidates) x (row strip)
__device__ void s16_plan(const MeParams &p, SbState &s, S16Plan &pl) {
    const int     lane = threadIdx.x & 31;
    constexpr int n    = kMaxSadJobs;
    static_assert(kMaxSadJobs == 32, "lane = job");
    SadJob       *jobs = s.sad_jobs;
    const int     bw = (int)(s.b64_w >> 2), rows = (int)((s.b64_h >> 2) >> 1);
    pl.job_key[lane] = ~0ull;
    if (lane == 0) { pl.n_sub_all = 0; pl.generic = 0; }
    const int  ji    = lane;
    const bool valid = jobs[ji].sa_w > 0 && jobs[ji].sa_h > 0;
    int        cnt = 0, nst = 1, ny_cap = 16;
    uint32_t   px = 0;
    if (valid) {
        const int W = jobs[ji].sa_w, H = jobs[ji].sa_h;
        const int a16  = (int)((uintptr_t)jobs[ji].ref & 15);                          // the same for every tile (x0 = 128 t)
        const int need = 4 * ((a16 >> 2) + (((a16 & 3) + min(128, W) + 3) >> 2)) + 16;  // bytes of the widest tile's rows
        const int pmax = need <= 48 ? 12 : need <= 80 ? 20 : need <= 112 ? 28 : need <= 144 ? 36 : 44;  // its row pitch in words
        // rows of a strip: whole boxes of 32 or of 8 rows that fit the buffer, minus the 14 rows below the last candidate
        const int fit = kS16BufWords / pmax;
        ny_cap = max((fit / 32) * 32, (fit / 8) * 8) - 14;
        ny_cap = min(SVTME_S16_NYCAP, ny_cap) & ~3;
        nst    = H <= ny_cap ? 1 : (H + ny_cap - 1) / ny_cap;
        cnt    = ((W + 127) >> 7) * nst;
        px     = (uint32_t)W * (uint32_t)(jobs[ji].skip_line ? (H >> 1) : H) * (8 * 16);
    }
    for (int o = 16; o; o >>= 1) px += __shfl_xor_sync(0xffffffffu, px, o);  // < 2^32: 32 jobs x 400 x 400 x 128
    if (lane == 0 && px) s.px += (unsigned long long)px;  // the only writer of the shared-memory copy at this point
    int incl = cnt;
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += v;
    }
    const int S = __shfl_sync(0xffffffffu, incl, 31);
    if (S <= kS16MaxSub && cnt > 0) {
        const SadJob &j = jobs[ji];
        for (int local = 0, xt = 0, st = 0; local < cnt; local++) {
            const int x0 = xt * 128, y0 = st * ny_cap, nx = min(128, j.sa_w - x0), ny = min(ny_cap, j.sa_h - y0);
            S16Sub   &u = pl.subs[incl - cnt + local];
            const int a16 = (int)((uintptr_t)(j.ref + x0) & 15), a = a16 & 3, wo = a16 >> 2;  // planes and rows are 16-byte aligned
            const int ngroups = (a + nx + 3) >> 2, need = 4 * (wo + ngroups) + 16;  // a group reads 5 words: 4 candidates + 15 more bytes
            const int cls = need <= 48 ? 0 : need <= 80 ? 1 : need <= 112 ? 2 : need <= 144 ? 3 : 4, P = 12 + 8 * cls;  // kS16BoxBytes[cls] / 4
            const int nyp = (ny + 3) & ~3, nrows = nyp + 14;
            // boxes of 32 rows when they fit the buffer (fewer copies), else of 8
            const int hc = (nrows > 24 && ((nrows + 31) & ~31) * P <= kS16BufWords) ? 1 : 0;
            const int brows = 8 << (2 * hc), nbox = (nrows + brows - 1) / brows;  // kS16BoxRows[hc]
            // the sub-window's first byte inside the reference's padded sixteenth plane (rows of j.raw_stride bytes)
            const DevPlane &rpl = p.refs[j.li][j.ri].lvl[2];
            const int       offb = (int)((j.ref + (int64_t)y0 * j.raw_stride + x0) - (rpl.org - (int64_t)kPad16 * rpl.stride - kPad16));
            const int       ty = offb / rpl.stride;
            u.job = (int16_t)ji; u.x0 = (int16_t)x0; u.y0 = (int16_t)y0; u.nx = (int16_t)nx; u.ny = (int16_t)ny;
            u.ngroups = (int16_t)ngroups; u.P = (int16_t)P; u.nbox = (int16_t)nbox; u.cls = (int16_t)(cls * kS16HClasses + hc); u.brows = (int16_t)brows;
            u.a = (int16_t)a; u.wo = (int16_t)wo;
            u.tx = offb - ty * rpl.stride - a16; u.ty = ty;
            u.hsel  = j.skip_line ? 1 : 0;
            u.ref   = (int16_t)(j.li * kMaxRefs + j.ri);
            u.items = (nyp >> 2) * (j.skip_line ? 1 : 2) * ngroups;
            u.inv_ng = __fdividef(1.0f, (float)ngroups) * 1.0000001f; u.inv_nx = __fdividef(1.0f, (float)nx) * 1.0000001f;
            if (++st == nst) { st = 0; xt++; }
        }
    }
    if (lane == 0) {
        pl.n_sub_all = S;
        pl.generic   = (bw != 16 || rows != 8 || S > kS16MaxSub) ? 1 : 0;
        if (pl.generic) {
            // the work counter of partial blocks: rows x width of THIS block (the sum above counted 8 x 16)
            long long d = 0;
            for (int q = 0; q < n; q++) {
                const SadJob &j = jobs[q];
                if (j.sa_w > 0 && j.sa_h > 0) {
                    const bool skip = j.skip_line && bw == 16 && rows <= 16;
                    d += (long long)j.sa_w * (skip ? (j.sa_h >> 1) : j.sa_h) * rows * bw;
                    d -= (long long)j.sa_w * (j.skip_line ? (j.sa_h >> 1) : j.sa_h) * (8 * 16);
                }
            }
            s.px += (unsigned long long)d;
        }
    }
    __syncwarp();
}

// The plan goes to HBM/L2 and its units into the picture's queue (counter pair `qi`: [count, head]) for the engine kernel
// that follows.  One warp.
__device__ __forceinline__ void s16_plan_store(const MeParams &p, const SbState &s, const S16Plan &pl, uint32_t sbi, int qi) {
    const int    lane = threadIdx.x & 31;
    const uint4 *l = reinterpret_cast<const uint4 *>(&pl);
    uint4       *g = reinterpret_cast<uint4 *>(static_cast<S16Plan *>(p.plans16) + sbi);
    const int    n_used = pl.generic ? 0 : pl.n_sub_all;
    const int    vecs = (int)(offsetof(S16Plan, subs) + (size_t)n_used * sizeof(S16Sub)) / 16;
    for (int i = lane; i < vecs; i += 32) g[i] = l[i];
    // units: sub-windows, or (generic blocks) job slots with a non-empty window
    const bool     gjob = pl.generic && s.sad_jobs[lane].sa_w > 0 && s.sad_jobs[lane].sa_h > 0;
    const unsigned gm   = __ballot_sync(0xffffffffu, gjob);
    const int      nu   = pl.generic ? __popc(gm) : pl.n_sub_all;
    uint32_t       base = 0;
    if (lane == 0 && nu) base = atomicAdd(&p.q_ctr[2 * qi], (uint32_t)nu);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (pl.generic) {
        if (gjob) p.q_units[base + __popc(gm & ((1u << lane) - 1u))] = (sbi << 8) | 128u | (uint32_t)lane;
    } else {
        for (int i = lane; i < nu; i += 32) p.q_units[base + i] = (sbi << 8) | (uint32_t)i;
    }
}

// the minima of a block's jobs, as the engine's warps left them in the plan -> the job table (control part of the
// state, in shared memory): what st_sad_kernel's warps write directly.  One warp, lane = job.
__device__ __forceinline__ void s16_keys(const MeParams &p, SbState &s, uint32_t sbi) {
    const int      lane = threadIdx.x & 31;
    const S16Plan &pl = static_cast<const S16Plan *>(p.plans16)[sbi];
    if (pl.generic) return;  // the generic path has written the job results itself
    const unsigned long long key = pl.job_key[lane];
    if (key != ~0ull && (uint32_t)(key >> 32) < 0xffffffu) {  // strict '<' against the 0xffffff start value
        SadJob &jo = s.sad_jobs[lane];
        jo.found = 1; jo.best_sad = (uint32_t)(key >> 32);
        jo.by = (int16_t)((key >> 16) & 0xffff); jo.bx = (int16_t)(key & 0xffff);
    }
    __syncwarp();
}

__global__ void __launch_bounds__(kStThreads, 6) st_sad16_kernel(const __grid_constant__ MeParams p, const __grid_constant__ S16Maps maps, int qi, int tag) {
    PDL_ENTRY();
    extern __shared__ __align__(128) uint32_t s16_win_all[];  // kStWarps x kS16BufWords
    __shared__ __align__(8) unsigned long long mbar_store[kStWarps];
    const int      warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t      *win = s16_win_all + warp * kS16BufWords;
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(win);
    const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(&mbar_store[warp]);
    if (lane == 0) mbar_init(mbar, 1);
    __syncwarp();
    uint32_t        phase = 0;
    uint32_t       *ctr = p.q_ctr + 2 * qi;
    const uint32_t  count = ctr[0];
    const DevPlane &cp = p.cur.lvl[2];
    S16Plan        *plans = static_cast<S16Plan *>(p.plans16);
    uint32_t        last_sb = 0xffffffffu;
    uint32_t        sreg[8][4];  // rows 0, 2, .. 14 of the block's 16x16 sixteenth-plane block (the same for every lane)
    (void)tag;
    for (;;) {
        uint32_t ui = 0;
        if (lane == 0) ui = atomicAdd(&ctr[1], 1u);
        ui = __shfl_sync(0xffffffffu, ui, 0);
        if (ui >= count) break;
        const uint32_t unit = p.q_units[ui], sbi = unit >> 8, idx = unit & 127u;
        SbState       &s = p.states[sbi];
        const uint32_t org_x = (sbi % p.wb) * 64, org_y = (sbi / p.wb) * 64;
        const uint8_t *b0 = cp.org + (int64_t)(org_y >> 2) * cp.stride + (org_x >> 2);
        if (unit & 128u) {
            // generic path (blocks cut by the picture edge, or with more sub-windows than a plan holds): one job
            const uint32_t aw = (uint32_t)((p.cur.lvl[0].width + 7) & ~7), ah = (uint32_t)((p.cur.lvl[0].height + 7) & ~7);
            const uint32_t b64_w = min(64u, aw - org_x), b64_h = min(64u, ah - org_y);
            const int      bw = (int)(b64_w >> 2), rows = (int)((b64_h >> 2) >> 1), nw = (bw + 3) >> 2;
            uint32_t      *blk = win;  // block words, rows of 16
            for (int i = lane; i < rows * nw; i += 32) {
                const int      r = i / nw, c = i - r * nw;
                const uint8_t *sp = b0 + (int64_t)r * (cp.stride * 2) + 4 * c;
                uint32_t       a = 0;
                for (int q = 0; q < min(4, bw - 4 * c); q++) a |= (uint32_t)sp[q] << (8 * q);
                blk[r * 16 + c] = a;
            }
            __syncwarp();
            const SadJob  j = s.sad_jobs[idx];
            constexpr int kWin = (kS16BufWords - 128) * 4;  // bytes behind the block words
            uint8_t      *w8 = reinterpret_cast<uint8_t *>(win + 128);
            if (bw == 16) warp_sad_job<4>(j, s.sad_jobs[idx], blk, w8, lane, rows, bw, 2, kWin);
            else warp_sad_job<0>(j, s.sad_jobs[idx], blk, w8, lane, rows, bw, 2, kWin);
            __syncwarp();
            last_sb = 0xffffffffu;
            continue;
        }
        // the sub-window's descriptor: three broadcast loads
        S16Sub u;
        {
            const uint4 *up = reinterpret_cast<const uint4 *>(&plans[sbi].subs[idx]);
            uint4       *ud = reinterpret_cast<uint4 *>(&u);
            ud[0] = up[0]; ud[1] = up[1]; ud[2] = up[2];
        }
        if (lane == 0) mbar_expect_tx(mbar, 4u * (uint32_t)(u.nbox * u.brows * u.P));
        __syncwarp();
        {
            const unsigned char *tm = maps.m[u.ref][u.cls];
            const uint32_t       box_bytes = 4u * (uint32_t)u.brows * (uint32_t)u.P;
            if (lane < u.nbox) tma_load_2d(sbase + (uint32_t)lane * box_bytes, tm, u.tx, u.ty + u.brows * lane, mbar);  // <= 9 boxes
        }
        if (sbi != last_sb) {
#pragma unroll
            for (int r = 0; r < 8; r++) {
                const uint4 v = __ldg(reinterpret_cast<const uint4 *>(b0 + (int64_t)(2 * r) * cp.stride));
                sreg[r][0] = v.x; sreg[r][1] = v.y; sreg[r][2] = v.z; sreg[r][3] = v.w;
            }
            last_sb = sbi;
        }
        mbar_wait(mbar, phase & 1u);
        phase++;
        uint32_t        best = ~0u;
        const int       u_ng = u.ngroups, u_nx = u.nx, u_ny = u.ny, u_P = u.P, u_a = u.a, u_hsel = u.hsel;
        const float     u_inv_ng = u.inv_ng;
        const uint32_t *u_win = win + u.wo;
        for (int item = lane; item < u.items; item += 32) {
            const int kh = (int)(((float)item + 0.5f) * u_inv_ng), g = item - kh * u_ng;
            const int k = u_hsel ? kh : (kh >> 1), h = u_hsel ? 1 : (kh & 1);
            const int yy = 4 * k + h;
            uint32_t  acc[2][4];
#pragma unroll
            for (int c = 0; c < 4; c++) acc[0][c] = acc[1][c] = 0;
            const uint32_t *rowp = u_win + yy * u_P + g;  // window rows yy, yy + 2, .. yy + 16
#pragma unroll
            for (int rr = 0; rr < 9; rr++) {
                const uint32_t *rp = rowp + rr * (2 * u_P);
                uint32_t        w[5];
#pragma unroll
                for (int c = 0; c < 5; c++) w[c] = rp[c];
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const uint32_t s0 = w[c], s1 = fsr(w[c], w[c + 1], 1), s2 = fsr(w[c], w[c + 1], 2), s3 = fsr(w[c], w[c + 1], 3);
                    if (rr < 8) {
                        const uint32_t sw = sreg[rr][c];
                        acc[0][0] = sad4(s0, sw, acc[0][0]); acc[0][1] = sad4(s1, sw, acc[0][1]);
                        acc[0][2] = sad4(s2, sw, acc[0][2]); acc[0][3] = sad4(s3, sw, acc[0][3]);
                    }
                    if (rr >= 1) {
                        const uint32_t sw = sreg[rr - 1][c];
                        acc[1][0] = sad4(s0, sw, acc[1][0]); acc[1][1] = sad4(s1, sw, acc[1][1]);
                        acc[1][2] = sad4(s2, sw, acc[1][2]); acc[1][3] = sad4(s3, sw, acc[1][3]);
                    }
                }
            }
            const int xb = 4 * g - u_a;
#pragma unroll
            for (int y2 = 0; y2 < 2; y2++) {
                const int  y = yy + 2 * y2;
                const bool yok = y < u_ny;
                const uint32_t ybase = (uint32_t)(y * u_nx + xb);
#pragma unroll
                for (int c = 0; c < 4; c++)
                    if (yok && (uint32_t)(xb + c) < (uint32_t)u_nx) best = min(best, acc[y2][c] * 65536u + ybase + (uint32_t)c);
            }
        }
        // the sub-window's minimum (sad << 16 | raster index) -> the job's key (sad, y, x): raster order inside the job
        // decides ties between sub-windows
#pragma unroll
        for (int o = 16; o; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
        if (lane == 0 && best != ~0u) {
            const uint32_t sad = best >> 16, ord = best & 0xffffu;
            const uint32_t y = (uint32_t)(((float)ord + 0.5f) * u.inv_nx), x = ord - y * (uint32_t)u.nx;
            atomicMin(&plans[sbi].job_key[u.job], ((unsigned long long)sad << 32) | ((uint32_t)(u.y0 + y) << 16) | (uint32_t)(u.x0 + x));
        }
        __syncwarp();  // every lane is done with the window before the next unit's copy lands in it
    }
}

// ---- sub-sampled full-pel search (the open-loop hot path): CTA-wide windows, balanced rounds ----
// All windows of the block's jobs are staged TOGETHER into shared memory with asynchronous 4-byte copies (one exposed
// memory latency per CTA); the candidates of all windows then form one flat list of ROUNDS dealt to the 4 warps.
// A round = one group of 4 x-adjacent candidates for FOUR candidate rows 4k .. 4k+3: half-warp `half` takes rows
// 4k + half and 4k + half + 2.  Those two rows share 7 of their 8 sub-sampled reference rows, so 9 window rows
// (loads + funnel shifts) serve 8 candidates -- LDS and SHF per SAD are nearly halved against one row per item.
// lane & 15 = a 16x16 block of the 64x64 (z-order), its 8 source rows (32 words) live in registers.
// Layout of a window: even and odd rows in two planes (offset = 16 mod 32 words), row i of a plane at
// i * P + (i >> 3) with P = 0 mod 4: the 32 lanes of a load hit banks 16 * half + by + 4 * bx + const -- no conflicts.
// Minima: 32-bit keys (sad << 12 | candidate index inside the <= 64 x 64 sub-window); 32x32 / 64x64 sums by a
// transposing butterfly (8 shuffles for 8 candidates).
constexpr int kFpBufWords = 8 * 1024;  // 44 KB of windows per CTA
__device__ __forceinline__ bool fp_job_small(const FpJob &j, int &a, int &ngroups, int &P, int &plane);
constexpr int kFpMaxSub   = 12;
struct FpSub {
    int16_t job, x0, y0, nx, ny, a, ngroups, P;
    int32_t off, plane_words, rbase, rounds;
};

// FULL (all 64 rows, the temporal filter's ME at tf hme_me_level <= 2): SAD_full(y) = [even source rows against window
// rows y + 2r] + [odd source rows against window rows y + 1 + 2r] -- two sub-sampled sums.  Half-warp `half` keeps the
// source rows of parity `half`; both halves walk the same candidates (rows 4k + e and 4k + e + 2, e = 0, 1 in turn),
// each reading the plane of parity (e + half) & 1, and one exchange over lane ^ 16 completes the sums: half h ends up
// with the four candidates of row 4k + e + 2h.
template <bool FULL>
__device__ __forceinline__ void fp_sub_block(const MeParams &p, uint32_t sbi, int big_only) {
    extern __shared__ __align__(16) uint32_t fp_win[];  // kFpBufWords
    __shared__ uint32_t               sub_key[kFpMaxSub][kPus];  // (sad << 12 | candidate) minima of the batch's sub-windows
    __shared__ unsigned long long     fp_key[kMaxFpJobs][kPus];
    __shared__ __align__(16) FpJob    jobs[kMaxFpJobs];
    __shared__ FpSub                  subs[kFpMaxSub];
    __shared__ int                    n_sub, cur_job, cur_xt, cur_st;
    PH_DECL
    __shared__ int                    b_lo[kFpMaxSub + 1], fast_nb;
    SbState  &s = p.states[sbi], &tab = s;  // the block's state (HBM/L2): job table in, best-SAD / best-MV tables out
    jobs_load_async<FpJob, kMaxFpJobs>(jobs, s.fp_jobs);
    {
        for (int i = threadIdx.x; i < kMaxFpJobs * kPus; i += blockDim.x) (&fp_key[0][0])[i] = ~0ull;
        if (threadIdx.x == 0) { cur_job = 0; cur_xt = 0; cur_st = 0; }
    }
    constexpr int n = kMaxFpJobs;  // fixed slots, unused ones have empty windows
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = lane & 15, half = lane >> 4;
    const int bx = ((b >> 2) & 1) * 2 + (b & 1), by = (b >> 3) * 2 + ((b >> 1) & 1);
    uint32_t  sreg[8][4];  // every second row of the lane's 16x16 block, straight from the source picture (sb_init geometry)
    {
        const DevPlane &f   = p.cur.lvl[0];
        const uint8_t  *sp0 = f.org + (int64_t)((sbi / p.wb) * 64 + 16 * by) * f.stride + (sbi % p.wb) * 64 + 16 * bx;
#pragma unroll
        for (int r = 0; r < 8; r++) {
            const uint4 v = __ldg(reinterpret_cast<const uint4 *>(sp0 + (int64_t)(2 * r + (FULL ? half : 0)) * f.stride));
            sreg[r][0] = v.x; sreg[r][1] = v.y; sreg[r][2] = v.z; sreg[r][3] = v.w;
        }
    }
    constexpr uint32_t kPxPerCand = FULL ? 64u * 64u : 64u * 32u;
    if (big_only) {
        // the jobs st_fp_warp_kernel has searched (small windows) are not this kernel's
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        if (threadIdx.x < kMaxFpJobs && jobs[threadIdx.x].w > 0 && jobs[threadIdx.x].h > 0) {
            int a, ng, P, plane;
            if (fp_job_small(jobs[threadIdx.x], a, ng, P, plane)) { jobs[threadIdx.x].w = 0; jobs[threadIdx.x].h = 0; }
        }
        __syncthreads();
    }
    if (!jobs_wait_any(n, [&](int i) { return jobs[i].w > 0 && jobs[i].h > 0; })) return;
    // the running best SADs this thread will compare against at the end: requested now, needed last
    constexpr int kFoldPer = (kMaxFpJobs * kPus + kStThreads - 1) / kStThreads;
    uint32_t      prev_best[kFoldPer];
#pragma unroll
    for (int t = 0; t < kFoldPer; t++) {
        const int i = threadIdx.x + t * kStThreads, ji = i / kPus, pu = i - ji * kPus;
        prev_best[t] = (i < n * kPus && jobs[ji].w > 0 && jobs[ji].h > 0) ? tab.best_sad[jobs[ji].li][jobs[ji].ri][pu] : 0u;
    }
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(fp_win);
    PH(0, 0)  // job fetch + source rows
    // ---- common case: every window is one sub-window (<= 64 x 64 candidates): all batches planned at once by warp 0,
    //      lane = job; otherwise thread 0 walks a cursor over (job, column, strip) batch by batch ----
    auto sub_geometry = [&](const FpJob &j, int x0, int y0, int nx, int ny, FpSub &u) {
        const int a = (int)((uintptr_t)(j.ref + x0) & 3);
        const int ngroups = (a + nx + 3) >> 2, P = (ngroups + 16 + 3) & ~3;
        const int nyp = (ny + 3) & ~3, PR = ((nyp + 62 + 7) & ~7) >> 1;  // staged in groups of 8 rows
        int       plane = PR * P + (PR >> 3) + 1;
        plane           = ((plane + 15) & ~31) + 16;  // = 16 mod 32
        u.x0 = (int16_t)x0; u.y0 = (int16_t)y0; u.nx = (int16_t)nx; u.ny = (int16_t)ny;
        u.a = (int16_t)a; u.ngroups = (int16_t)ngroups; u.P = (int16_t)P;
        u.plane_words = plane; u.rounds = (nyp >> 2) * ngroups * (FULL ? 2 : 1);
    };
    if (warp == 0) {
        const bool valid = lane < n && jobs[lane].w > 0 && jobs[lane].h > 0;
        const bool fast  = __all_sync(0xffffffffu, !valid || (jobs[lane].w <= 64 && jobs[lane].h <= 64));
        if (fast) {
            const unsigned vmask = __ballot_sync(0xffffffffu, valid);
            const int      S = __popc(vmask), idx = __popc(vmask & ((1u << lane) - 1u));
            int            words = 0, rounds = 0;
            uint32_t       px = 0;
            if (valid) {
                FpSub &u = subs[idx];
                sub_geometry(jobs[lane], 0, 0, jobs[lane].w, jobs[lane].h, u);
                u.job = (int16_t)lane;
                words = 2 * u.plane_words; rounds = u.rounds;
                px    = (uint32_t)(jobs[lane].w * jobs[lane].h) * kPxPerCand;  // <= 64 * 64 * 4096
            }
            int cw = words, cr = rounds;
            for (int o = 1; o < 8; o <<= 1) {  // jobs live in lanes 0..7
                const int vw = __shfl_up_sync(0xffffffffu, cw, o), vr = __shfl_up_sync(0xffffffffu, cr, o);
                if (lane >= o) { cw += vw; cr += vr; }
            }
            for (int o = 4; o; o >>= 1) px += __shfl_xor_sync(0xffffffffu, px, o);
            if (lane == 0 && px) atomicAdd(&s.px, (unsigned long long)px);
            int base_w = 0, nb = 0;
            if (lane == 0) b_lo[0] = 0;
            for (;;) {
                const unsigned over = __ballot_sync(0xffffffffu, valid && cw - base_w > kFpBufWords);
                if (!over) break;
                const int first = __ffs(over) - 1;
                const int fw = __shfl_sync(0xffffffffu, cw - words, first), fidx = __shfl_sync(0xffffffffu, idx, first);
                nb++;
                if (lane == 0) b_lo[nb] = fidx;
                if (lane >= first) base_w = fw;
            }
            if (valid) { subs[idx].off = cw - words - base_w; subs[idx].rbase = cr - rounds; }
            if (lane == 0) { b_lo[nb + 1] = S; fast_nb = S ? nb + 1 : 0; }
        } else if (lane == 0) fast_nb = -1;
    }
    __syncthreads();
    const int nb_fast = fast_nb;

    for (int bi = 0;; bi++) {
        int s_lo = 0, s_hi = 0;
        if (nb_fast >= 0) {
            if (bi >= nb_fast) break;
            s_lo = b_lo[bi]; s_hi = b_lo[bi + 1];
        } else {
            if (threadIdx.x == 0) {
                // next batch of sub-windows (job, 64-wide column, 64-high strip) that fits the buffer
                int                ns = 0, used = 0, rb = 0;
                unsigned long long px = 0;
                while (cur_job < n && ns < kFpMaxSub) {
                    const FpJob &j = jobs[cur_job];
                    if (j.w <= 0 || j.h <= 0) { cur_job++; continue; }
                    const int nxt = (j.w + 63) >> 6, nst = (j.h + 63) >> 6;
                    const int x0 = cur_xt * 64, y0 = cur_st * 64;
                    FpSub     u;
                    sub_geometry(j, x0, y0, min(64, j.w - x0), min(64, j.h - y0), u);
                    if (used + 2 * u.plane_words > kFpBufWords) break;  // a single sub-window always fits (<= 4.8 K words)
                    u.job = (int16_t)cur_job; u.off = used; u.rbase = rb;
                    subs[ns++] = u;
                    used += 2 * u.plane_words; rb += u.rounds;
                    if (cur_xt == 0 && cur_st == 0) px += (unsigned long long)j.w * j.h * kPxPerCand;
                    if (++cur_st == nst) { cur_st = 0; if (++cur_xt == nxt) { cur_xt = 0; cur_job++; } }
                }
                n_sub = ns;
                if (px) atomicAdd(&s.px, px);
            }
            __syncthreads();
            s_hi = n_sub;
            if (s_hi == 0) break;
        }
        PH(0, 1)  // batch planning
        const int ns = s_hi - s_lo;
        for (int i = threadIdx.x; i < ns * kPus; i += blockDim.x) (&sub_key[s_lo][0])[i] = ~0u;
        // stage: warp w copies rows w, w + 4, ... (fixed parity), lane = word of the row.  Rows go in groups of 8
        // (2 per warp): inside a group the skew (i >> 3) is constant, so the addresses are plain strides.  The rows
        // past the window (< 8 + 3) are inside the plane's physical padding (96 rows against the 72 a search uses).
        for (int si = s_lo; si < s_hi; si++) {
            const FpSub   &u = subs[si];
            const FpJob   &j = jobs[u.job];
            const int      ngrp = (((u.ny + 3) & ~3) + 62 + 7) >> 3, nwords = u.ngroups + 16, raw = j.ref_stride;
            const uint8_t *g0 = j.ref + (int64_t)u.y0 * raw + u.x0 - u.a + (int64_t)warp * raw;
            const uint32_t P4 = 4u * (uint32_t)u.P;
            const uint32_t pbase = sbase + 4u * (uint32_t)(u.off + (warp & 1) * u.plane_words) + (uint32_t)(warp >> 1) * P4;
            for (int wd = lane; wd < nwords; wd += 32) {
                const uint8_t *g  = g0 + 4 * wd;
                uint32_t       sa = pbase + 4u * (uint32_t)wd;
                for (int G = 0; G < ngrp; G++, g += 8 * (int64_t)raw, sa += 4u * P4 + ((G & 1) ? 0u : 4u)) {
                    // rows warp + 8G and warp + 8G + 4: plane rows i = (warp >> 1) + 4G + {0, 2}, skew i >> 3 = G >> 1
                    CHECK_SRC(p, 0, j.li, j.ri, g, 4, 3);
                    CHECK_SRC(p, 0, j.li, j.ri, g + 4 * (int64_t)raw, 4, 3);
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(g));
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa + 2u * P4), "l"(g + 4 * (int64_t)raw));
                }
            }
        }
        PH(0, 2)  // copy issue
        asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        PH(0, 3)  // copy wait
        for (int si = s_lo; si < s_hi; si++) {
            const FpSub &u = subs[si];
            const int    first = (warp - u.rbase) & (kStWarps - 1);
            if (first >= u.rounds) continue;
            const int ngroups = u.ngroups, nx = u.nx, ny = u.ny, P = u.P;
            int       k = first / ngroups, g = first - k * ngroups;
            uint32_t  best8[4] = {~0u, ~0u, ~0u, ~0u}, best16 = ~0u, best32 = ~0u, best64 = ~0u;
            const uint32_t *wbase0 = fp_win + u.off + 4 * bx + (8 * by) * P + by;
            for (int local = first; local < u.rounds; local += kStWarps) {
                // SUB: k counts groups of 4 candidate rows.  FULL: k counts pairs (rows 4kk + e, 4kk + e + 2)
                const int       kk = FULL ? (k >> 1) : k, e = FULL ? (k & 1) : 0;
                const int       yy = 4 * kk + (FULL ? e : half);
                const int       i0 = 2 * kk + (FULL ? (e & half) : 0);
                const uint32_t *wbase = wbase0 + (FULL ? (e ^ half) : half) * u.plane_words;
                const bool      active = yy < ny;
                uint32_t   acc[2][4][4];
#pragma unroll
                for (int y2 = 0; y2 < 2; y2++)
#pragma unroll
                    for (int c = 0; c < 4; c++) acc[y2][c][0] = acc[y2][c][1] = acc[y2][c][2] = acc[y2][c][3] = 0;
                if (active) {
                    const uint32_t *rowp = wbase + g;
#pragma unroll
                    for (int rr = 0; rr < 9; rr++) {
                        const int       i  = i0 + rr;  // + 8 * by is in wbase, and (i + 8 by) >> 3 = by + (i >> 3)
                        const uint32_t *rp = rowp + i * P + (i >> 3);
                        uint32_t        w[5];
#pragma unroll
                        for (int c = 0; c < 5; c++) w[c] = rp[c];
#pragma unroll
                        for (int c = 0; c < 4; c++) {
                            const uint32_t s0 = w[c], s1 = fsr(w[c], w[c + 1], 1), s2 = fsr(w[c], w[c + 1], 2), s3 = fsr(w[c], w[c + 1], 3);
                            if (rr < 8) {
                                const int      q  = (rr >> 2) * 2 + (c >> 1);
                                const uint32_t sw = sreg[rr][c];
                                acc[0][0][q] = sad4(s0, sw, acc[0][0][q]); acc[0][1][q] = sad4(s1, sw, acc[0][1][q]);
                                acc[0][2][q] = sad4(s2, sw, acc[0][2][q]); acc[0][3][q] = sad4(s3, sw, acc[0][3][q]);
                            }
                            if (rr >= 1) {
                                const int      q  = ((rr - 1) >> 2) * 2 + (c >> 1);
                                const uint32_t sw = sreg[rr - 1][c];
                                acc[1][0][q] = sad4(s0, sw, acc[1][0][q]); acc[1][1][q] = sad4(s1, sw, acc[1][1][q]);
                                acc[1][2][q] = sad4(s2, sw, acc[1][2][q]); acc[1][3][q] = sad4(s3, sw, acc[1][3][q]);
                            }
                        }
                    }
                }
                const int xb = 4 * g - u.a;
                if constexpr (FULL) {
                    // complete the sums: this half keeps candidate row yy + 2 * half and gets the other parity's share of it
                    uint32_t s16[4];
                    const int  y = yy + 2 * half;
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        const int      xx = xb + c;
                        const bool     valid = y < ny && xx >= 0 && xx < nx;
                        const uint32_t ord = (uint32_t)(y * nx + xx);  // < 4096
                        uint32_t       t = 0;
#pragma unroll
                        for (int q = 0; q < 4; q++) {
                            const uint32_t send = half ? acc[0][c][q] : acc[1][c][q], keep = half ? acc[1][c][q] : acc[0][c][q];
                            const uint32_t v = keep + __shfl_xor_sync(0xffffffffu, send, 16);
                            if (valid) best8[q] = min(best8[q], v * 4096u + ord);
                            t += v;
                        }
                        if (valid) best16 = min(best16, t * 4096u + ord);
                        s16[c] = t;
                    }
                    // transposing butterfly over the 4 lanes of a 32x32 block: candidate c = 2 * b0 + b1 stays here
                    uint32_t t2[2], t1;
                    {
                        const bool up = b & 1;
#pragma unroll
                        for (int i = 0; i < 2; i++) {
                            const uint32_t send = up ? s16[i] : s16[2 + i], keep = up ? s16[2 + i] : s16[i];
                            t2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
                        }
                    }
                    {
                        const bool     up = b & 2;
                        const uint32_t send = up ? t2[0] : t2[1], keep = up ? t2[1] : t2[0];
                        t1 = keep + __shfl_xor_sync(0xffffffffu, send, 2);
                    }
                    const int  xx = xb + 2 * (b & 1) + ((b >> 1) & 1);
                    const bool ok = y < ny && xx >= 0 && xx < nx;
                    if (ok) best32 = min(best32, t1 * 4096u + (uint32_t)(y * nx + xx));
                    t1 += __shfl_xor_sync(0xffffffffu, t1, 4);
                    t1 += __shfl_xor_sync(0xffffffffu, t1, 8);
                    if (ok) best64 = min(best64, t1 * 4096u + (uint32_t)(y * nx + xx));
                } else {
                    uint32_t  s16[8];
    #pragma unroll
                    for (int y2 = 0; y2 < 2; y2++)
    #pragma unroll
                        for (int c = 0; c < 4; c++) {
                            const int      xx = xb + c, y = yy + 2 * y2;
                            const bool     valid = y < ny && xx >= 0 && xx < nx;
                            const uint32_t ord = (uint32_t)(y * nx + xx);  // < 4096
                            uint32_t       t = 0;
    #pragma unroll
                            for (int q = 0; q < 4; q++) {
                                if (valid) best8[q] = min(best8[q], acc[y2][c][q] * 4096u + ord);
                                t += acc[y2][c][q];
                            }
                            if (valid) best16 = min(best16, t * 4096u + ord);
                            s16[y2 * 4 + c] = t;
                        }
                    // transposing butterfly over the 4 lanes of a 32x32 block, then over the 4 blocks: candidate e = y2 * 4 + c
                    uint32_t t4[4], t2[2], t1;
                    {
                        const bool up = b & 1;  // keeps y2 = b0
    #pragma unroll
                        for (int i = 0; i < 4; i++) {
                            const uint32_t send = up ? s16[i] : s16[4 + i], keep = up ? s16[4 + i] : s16[i];
                            t4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
                        }
                    }
                    {
                        const bool up = b & 2;  // keeps c >> 1 = b1
    #pragma unroll
                        for (int i = 0; i < 2; i++) {
                            const uint32_t send = up ? t4[i] : t4[2 + i], keep = up ? t4[2 + i] : t4[i];
                            t2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
                        }
                    }
                    const int  y32 = yy + 2 * (b & 1), x32 = xb + (b & 2);
                    const bool yok = y32 < ny;
    #pragma unroll
                    for (int i = 0; i < 2; i++) {
                        const int xx = x32 + i;
                        if (yok && xx >= 0 && xx < nx) best32 = min(best32, t2[i] * 4096u + (uint32_t)(y32 * nx + xx));
                    }
                    {
                        const bool     up   = b & 4;  // keeps c & 1 = b2
                        const uint32_t send = up ? t2[0] : t2[1], keep = up ? t2[1] : t2[0];
                        t1 = keep + __shfl_xor_sync(0xffffffffu, send, 4);
                    }
                    t1 += __shfl_xor_sync(0xffffffffu, t1, 8);
                    {
                        const int xx = x32 + ((b >> 2) & 1);
                        if (yok && xx >= 0 && xx < nx) best64 = min(best64, t1 * 4096u + (uint32_t)(y32 * nx + xx));
                    }
                }
                g += kStWarps;
                while (g >= ngroups) { g -= ngroups; k++; }
            }
            // this lane's minima -> the sub-window's keys (native 32-bit shared atomics)
#pragma unroll
            for (int q = 0; q < 4; q++)
                if (best8[q] != ~0u) atomicMin(&sub_key[si][21 + 4 * b + q], best8[q]);
            if (best16 != ~0u) atomicMin(&sub_key[si][5 + b], best16);
            if (best32 != ~0u) atomicMin(&sub_key[si][1 + (b >> 2)], best32);
            if (b < 8 && best64 != ~0u) atomicMin(&sub_key[si][0], best64);
        }
        PH(0, 4)  // search (this warp)
        __syncthreads();
        PH(0, 5)  // waiting for the other warps
        // sub-window keys -> job keys: (SAD -- doubled in sub-sampled mode --, raster index inside the job's window)
        for (int i = threadIdx.x; i < ns * kPus; i += blockDim.x) {
            const int      si = s_lo + i / kPus, pu = i - (i / kPus) * kPus;
            const uint32_t key = sub_key[si][pu];
            if (key == ~0u) continue;
            const FpSub   &u = subs[si];
            const uint32_t sad = key >> 12, ord = key & 4095u, y = ord / (uint32_t)u.nx, x = ord - y * (uint32_t)u.nx;
            atomicMin(&fp_key[u.job][pu], ((unsigned long long)(FULL ? sad : sad << 1) << 32) | (uint32_t)((u.y0 + y) * jobs[u.job].w + u.x0 + x));
        }
        __syncthreads();  // the windows are reused by the next batch
        PH(0, 6)  // key fold
    }
    // fold the window minima into the running best of the reference (strict '<': earlier results win ties)
#pragma unroll
    for (int t = 0; t < kFoldPer; t++) {
        const int i = threadIdx.x + t * kStThreads;
        if (i >= n * kPus) break;
        const int                ji = i / kPus, pu = i - ji * kPus;
        const FpJob             &j  = jobs[ji];
        const unsigned long long key = fp_key[ji][pu];
        if (key == ~0ull) continue;
        const uint32_t sad = (uint32_t)(key >> 32), ord = (uint32_t)key;
        if (sad < prev_best[t]) {
            const int32_t y = (int32_t)(ord / (uint32_t)j.w), x = (int32_t)(ord - (uint32_t)y * (uint32_t)j.w);
            tab.best_sad[j.li][j.ri][pu] = sad;
            tab.best_mv[j.li][j.ri][pu]  = ((uint32_t)(y + j.yo) << 16) | (uint16_t)(x + j.xo);
        }
    }
    PH(0, 7)  // results to the block state
}
// big_only: behind st_fp_warp_kernel, for the jobs that kernel left (counter word `qi` != 0 when there are any)
template <bool FULL>
__global__ void __launch_bounds__(kStThreads, 5) st_fp_sub_kernel(const __grid_constant__ MeParams p, int big_only, int qi) {
    PDL_ENTRY();
    if (big_only && p.q_ctr[2 * qi] == 0) return;
    const uint32_t n_sb = p.wb * p.hb;
    for (uint32_t sbi = blockIdx.x; sbi < n_sb; sbi += gridDim.x) {
        fp_sub_block<FULL>(p, sbi, big_only);
        __syncthreads();
    }
}

// ---- the same search for SMALL windows (what preset M6 and faster produce: a few dozen candidates per reference), one
// WARP per 64x64 block.  The CTA-wide kernel above spends two thirds of a CTA's life outside the search when the windows
// are small (job fetch, plan, copy issue, key fold, results: phases of one warp with the others at a barrier); here a
// warp owns its block -- jobs, window buffer, minima, results -- and never meets a barrier, so the four warps of a CTA
// (four blocks) overlap each other's copies and searches.  A job whose window does not fit the warp's buffer is left to
// the CTA-wide kernel, which then runs behind this one for those jobs only (`big` flag in the step's counter pair).
#ifndef SVTME_FPW_BUF_WORDS
#define SVTME_FPW_BUF_WORDS 3584
#endif
constexpr int kFpWBufWords = SVTME_FPW_BUF_WORDS;  // 14 KB per warp: windows up to 64 x 32 candidates
#ifndef SVTME_FPW_SPLIT
#define SVTME_FPW_SPLIT 2
#endif
__device__ __forceinline__ bool fp_job_small(const FpJob &j, int &a, int &ngroups, int &P, int &plane) {
    a       = (int)((uintptr_t)j.ref & 3);
    ngroups = (a + j.w + 3) >> 2;
    P       = (ngroups + 16 + 3) & ~3;
    const int nyp = (j.h + 3) & ~3, PR = ((nyp + 62 + 7) & ~7) >> 1;  // staged in groups of 8 rows
    plane   = PR * P + (PR >> 3) + 1;
    plane   = ((plane + 15) & ~31) + 16;  // = 16 mod 32
    return j.w <= 64 && j.h <= 64 && 2 * plane <= kFpWBufWords;
}

#ifndef SVTME_FPW_CTAS
#define SVTME_FPW_CTAS 4  // resident CTAs per SM the register budget is set for (128 registers)
#endif
__global__ void __launch_bounds__(kStThreads, SVTME_FPW_CTAS) st_fp_warp_kernel(const __grid_constant__ MeParams p, int qi) {
    PDL_ENTRY();
    extern __shared__ __align__(16) uint32_t fpw_all[];  // kStWarps x kFpWBufWords
    const int      warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // SVTME_FPW_SPLIT warps share a block: warp `part` takes every SVTME_FPW_SPLIT-th job of it
    constexpr int  kSplit = SVTME_FPW_SPLIT;
    const uint32_t sbi = blockIdx.x * (kStWarps / kSplit) + warp / kSplit;
    const int      part = warp % kSplit;
    if (sbi >= p.wb * p.hb) return;
    uint32_t      *win = fpw_all + warp * kFpWBufWords;
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(win);
    SbState       &s = p.states[sbi];
    const int      b = lane & 15, half = lane >> 4;
    const int      bx = ((b >> 2) & 1) * 2 + (b & 1), by = (b >> 3) * 2 + ((b >> 1) & 1);
    uint32_t       sreg[8][4];  // every second row of the lane's 16x16 block, straight from the source picture
    {
        const DevPlane &f   = p.cur.lvl[0];
        const uint8_t  *sp0 = f.org + (int64_t)((sbi / p.wb) * 64 + 16 * by) * f.stride + (sbi % p.wb) * 64 + 16 * bx;
#pragma unroll
        for (int r = 0; r < 8; r++) {
            const uint4 v = __ldg(reinterpret_cast<const uint4 *>(sp0 + (int64_t)(2 * r) * f.stride));
            sreg[r][0] = v.x; sreg[r][1] = v.y; sreg[r][2] = v.z; sreg[r][3] = v.w;
        }
    }
    uint32_t px = 0;
    bool     any_big = false;
    int      ordinal = 0;
    // the block's job table (8 x 24 bytes) in ONE round trip: lane = word of the table, a job's fields come by shuffle
    static_assert(sizeof(FpJob) == 24 && kMaxFpJobs * 6 <= 64, "job table = 48 words");
    const uint32_t *jt  = reinterpret_cast<const uint32_t *>(s.fp_jobs);
    const uint32_t  jt0 = jt[lane], jt1 = lane < kMaxFpJobs * 6 - 32 ? jt[32 + lane] : 0u;
    for (int ji = 0; ji < kMaxFpJobs; ji++) {
        FpJob j;
        {
            uint32_t f[6];
#pragma unroll
            for (int q = 0; q < 6; q++) {
                const int      idx = 6 * ji + q;
                const uint32_t v0 = __shfl_sync(0xffffffffu, jt0, idx & 31), v1 = __shfl_sync(0xffffffffu, jt1, idx & 31);
                f[q] = idx < 32 ? v0 : v1;
            }
            j.ref = reinterpret_cast<const uint8_t *>(((unsigned long long)f[1] << 32) | f[0]);
            j.ref_stride = (int32_t)f[2];
            j.xo = (int16_t)(f[3] & 0xffff); j.yo = (int16_t)(f[3] >> 16);
            j.w = (int16_t)(f[4] & 0xffff); j.h = (int16_t)(f[4] >> 16);
            j.li = (uint8_t)(f[5] & 0xff); j.ri = (uint8_t)((f[5] >> 8) & 0xff);
        }
        if (j.w <= 0 || j.h <= 0) continue;
        int a, ngroups, P, plane;
        if (!fp_job_small(j, a, ngroups, P, plane)) { any_big = true; continue; }
        if ((ordinal++) % kSplit != part) continue;
        px += (uint32_t)(j.w * j.h) * (64u * 32u);
        const int nx = j.w, ny = j.h, nyp = (ny + 3) & ~3;
        uint32_t *bs = s.best_sad[j.li][j.ri], *bm = s.best_mv[j.li][j.ri];
        // the running bests this lane will compare against at the end: requested now, needed last
        const int      pu8 = 21 + 4 * b + 2 * half;
        const uint32_t prev8a = bs[pu8], prev8b = bs[pu8 + 1], prev16 = bs[5 + b], prev32 = bs[1 + (b >> 2)], prev64 = bs[0];
        // stage: lane = word of the row; rows go in groups of 8 (inside a group the skew (i >> 3) is constant).  The rows
        // past the window (< 8 + 3) are inside the plane's physical padding (96 rows against the 72 a search uses).
        {
            const int      ngrp = (nyp + 62 + 7) >> 3, nwords = ngroups + 16, raw = j.ref_stride;
            const uint8_t *g0 = j.ref - a;
            __syncwarp();  // the previous job's window is no longer read
            for (int wd = lane; wd < nwords; wd += 32) {
                const uint8_t *g = g0 + 4 * wd;
                for (int G = 0; G < ngrp; G++) {
                    // rows 8G + r8: plane r8 & 1, plane row i = 4G + (r8 >> 1), skew i >> 3 = G >> 1
                    const uint32_t sa0 = sbase + 4u * (uint32_t)((4 * G) * P + (G >> 1) + wd);
#pragma unroll
                    for (int r8 = 0; r8 < 8; r8++, g += raw) {
                        CHECK_SRC(p, 0, j.li, j.ri, g, 4, 4);
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa0 + 4u * (uint32_t)((r8 & 1) * plane + (r8 >> 1) * P)), "l"(g));
                    }
                }
            }
            asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
            __syncwarp();
        }
        // search: a round = one group of 4 x-adjacent candidates for the candidate rows 4k .. 4k + 3 (half-warp `half` takes
        // rows 4k + half and 4k + half + 2); the rounds of the window one after the other
        uint32_t        best8[4] = {~0u, ~0u, ~0u, ~0u}, best16 = ~0u, best32 = ~0u, best64 = ~0u;
        const uint32_t *wbase = win + 4 * bx + (8 * by) * P + by + half * plane;
        const int       rounds = (nyp >> 2) * ngroups;
        for (int local = 0, k = 0, g = 0; local < rounds; local++) {
            const int  yy = 4 * k + half, i0 = 2 * k;
            const bool active = yy < ny;
            uint32_t   acc[2][4][4];
#pragma unroll
            for (int y2 = 0; y2 < 2; y2++)
#pragma unroll
                for (int c = 0; c < 4; c++) acc[y2][c][0] = acc[y2][c][1] = acc[y2][c][2] = acc[y2][c][3] = 0;
            if (active) {
                const uint32_t *rowp = wbase + g;
#pragma unroll
                for (int rr = 0; rr < 9; rr++) {
                    const int       i  = i0 + rr;  // + 8 * by is in wbase, and (i + 8 by) >> 3 = by + (i >> 3)
                    const uint32_t *rp = rowp + i * P + (i >> 3);
                    uint32_t        w[5];
#pragma unroll
                    for (int c = 0; c < 5; c++) w[c] = rp[c];
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        const uint32_t s0 = w[c], s1 = fsr(w[c], w[c + 1], 1), s2 = fsr(w[c], w[c + 1], 2), s3 = fsr(w[c], w[c + 1], 3);
                        if (rr < 8) {
                            const int      q  = (rr >> 2) * 2 + (c >> 1);
                            const uint32_t sw = sreg[rr][c];
                            acc[0][0][q] = sad4(s0, sw, acc[0][0][q]); acc[0][1][q] = sad4(s1, sw, acc[0][1][q]);
                            acc[0][2][q] = sad4(s2, sw, acc[0][2][q]); acc[0][3][q] = sad4(s3, sw, acc[0][3][q]);
                        }
                        if (rr >= 1) {
                            const int      q  = ((rr - 1) >> 2) * 2 + (c >> 1);
                            const uint32_t sw = sreg[rr - 1][c];
                            acc[1][0][q] = sad4(s0, sw, acc[1][0][q]); acc[1][1][q] = sad4(s1, sw, acc[1][1][q]);
                            acc[1][2][q] = sad4(s2, sw, acc[1][2][q]); acc[1][3][q] = sad4(s3, sw, acc[1][3][q]);
                        }
                    }
                }
            }
            const int xb = 4 * g - a;
            uint32_t  s16[8];
#pragma unroll
            for (int y2 = 0; y2 < 2; y2++)
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const int      xx = xb + c, y = yy + 2 * y2;
                    const bool     valid = y < ny && xx >= 0 && xx < nx;
                    const uint32_t ord = (uint32_t)(y * 64 + xx);  // < 4096; the same order as y * nx + xx (nx <= 64)
                    uint32_t       t = 0;
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        if (valid) best8[q] = min(best8[q], acc[y2][c][q] * 4096u + ord);
                        t += acc[y2][c][q];
                    }
                    if (valid) best16 = min(best16, t * 4096u + ord);
                    s16[y2 * 4 + c] = t;
                }
            // transposing butterfly over the 4 lanes of a 32x32 block, then over the 4 blocks: candidate e = y2 * 4 + c
            uint32_t t4[4], t2[2], t1;
            {
                const bool up = b & 1;  // keeps y2 = b0
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const uint32_t send = up ? s16[i] : s16[4 + i], keep = up ? s16[4 + i] : s16[i];
                    t4[i] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
                }
            }
            {
                const bool up = b & 2;  // keeps c >> 1 = b1
#pragma unroll
                for (int i = 0; i < 2; i++) {
                    const uint32_t send = up ? t4[i] : t4[2 + i], keep = up ? t4[2 + i] : t4[i];
                    t2[i] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
                }
            }
            const int  y32 = yy + 2 * (b & 1), x32 = xb + (b & 2);
            const bool yok = y32 < ny;
#pragma unroll
            for (int i = 0; i < 2; i++) {
                const int xx = x32 + i;
                if (yok && xx >= 0 && xx < nx) best32 = min(best32, t2[i] * 4096u + (uint32_t)(y32 * 64 + xx));
            }
            {
                const bool     up   = b & 4;  // keeps c & 1 = b2
                const uint32_t send = up ? t2[0] : t2[1], keep = up ? t2[1] : t2[0];
                t1 = keep + __shfl_xor_sync(0xffffffffu, send, 4);
            }
            t1 += __shfl_xor_sync(0xffffffffu, t1, 8);
            {
                const int xx = x32 + ((b >> 2) & 1);
                if (yok && xx >= 0 && xx < nx) best64 = min(best64, t1 * 4096u + (uint32_t)(y32 * 64 + xx));
            }
            if (++g == ngroups) { g = 0; k++; }
        }
        // the window's minima: the two half-warps hold different candidate rows; the 32x32 / 64x64 sums of different
        // candidates sit in different lanes of the block group.  Keys are (sad << 12 | raster index): min = first in raster order
#pragma unroll
        for (int q = 0; q < 4; q++) best8[q] = min(best8[q], __shfl_xor_sync(0xffffffffu, best8[q], 16));
        best16 = min(best16, __shfl_xor_sync(0xffffffffu, best16, 16));
        best32 = min(best32, __shfl_xor_sync(0xffffffffu, best32, 1));
        best32 = min(best32, __shfl_xor_sync(0xffffffffu, best32, 2));
        best32 = min(best32, __shfl_xor_sync(0xffffffffu, best32, 16));
        best64 = min(best64, __shfl_xor_sync(0xffffffffu, best64, 1));
        best64 = min(best64, __shfl_xor_sync(0xffffffffu, best64, 2));
        best64 = min(best64, __shfl_xor_sync(0xffffffffu, best64, 4));
        best64 = min(best64, __shfl_xor_sync(0xffffffffu, best64, 16));
        // fold into the running best of the reference (strict '<': earlier results win ties); SAD doubled (sub-sampled rows)
        auto fold = [&](int pu, uint32_t key, uint32_t prev) {
            if (key == ~0u) return;
            const uint32_t sad = (key >> 12) << 1, ord = key & 4095u;
            if (sad < prev) {
                const uint32_t y = ord >> 6, x = ord & 63u;
                bs[pu] = sad;
                bm[pu] = ((uint32_t)((int32_t)y + j.yo) << 16) | (uint16_t)((int32_t)x + j.xo);
            }
        };
        fold(pu8, half ? best8[2] : best8[0], prev8a);
        fold(pu8 + 1, half ? best8[3] : best8[1], prev8b);
        if (half == 0) fold(5 + b, best16, prev16);
        if (half == 0 && (b & 3) == 0) fold(1 + (b >> 2), best32, prev32);
        if (lane == 0) fold(0, best64, prev64);
    }
    for (int o = 16; o; o >>= 1) any_big |= (bool)__shfl_xor_sync(0xffffffffu, (int)any_big, o);  // (uniform anyway)
    if (lane == 0) {
        if (px) atomicAdd(&s.px, (unsigned long long)px);
        if (any_big) atomicOr(&p.q_ctr[2 * qi], 1u);  // the CTA-wide kernel behind this one has work
    }
}

// The tail of a kernel: steps [first, first + count) of the program -- control steps (warp 0) and the light
// svt_nxm_sad_kernel probes (all warps) -- then, in the picture's last kernel, the results (sb_finish), and the state goes
// back to HBM/L2.  Called by every thread of the CTA after a barrier; one copy of this code serves all kernels.
__device__ __forceinline__ void run_tail(const MeParams &p, unsigned char *head, S16Plan &plan, int first, int count, int finish) {
    SbState &s = *reinterpret_cast<SbState *>(head);
    for (int i = first; i < first + count; i++) {
        const Step st = p.prog[i];
        if (st.kind == STEP_CTL) {
            if (threadIdx.x < 32) ctl_step(p, s, st.op, st.arg);
        } else if (st.kind == STEP_PLAN16) {
            // the sixteenth-plane jobs just made -> sub-windows; the plan and its units go to HBM/L2 for the engine
            if (threadIdx.x < 32) {
                s16_plan(p, s, plan);
                s16_plan_store(p, s, plan, blockIdx.x, i);
            }
        } else if (st.kind == STEP_KEYS16) {
            if (threadIdx.x < 32) s16_keys(p, s, blockIdx.x);
        } else if (st.kind == STEP_PROBE) {
            probe_engine(p, s);
        } else {
            nxm_engine(s);
        }
        __syncthreads();
    }
    if (finish) {
        StBackend be;
        sb_finish(be, p, s, blockIdx.x);
        if (threadIdx.x == 0 && p.px_counter) atomicAdd(p.px_counter, s.px);
    }
    head_store(head, s.home);
}

// The steps between two search engines (and the results behind the last one): a light kernel, ONE WARP per block --
// 32 threads, ~2 K registers and 4.7 KB of shared memory per CTA, so that it slips in beside the six resident CTAs of
// another picture's search engine instead of displacing one (a 128-thread version cost 5 % of the picture rate)
constexpr int kLightThreads = 32;
__global__ void __launch_bounds__(kLightThreads) st_light_kernel(const __grid_constant__ MeParams p, int first, int count, int finish) {
    PDL_ENTRY();
    __shared__ __align__(16) unsigned char head[kSbHeadBytes];
    __shared__ S16Plan plan;
    head_load_async(head, &p.states[blockIdx.x]);
    head_wait();
    run_tail(p, head, plan, first, count, finish);
}

// First kernel of a picture: init_me_hme_data for every block, then the steps in front of the first search engine
#ifndef SVTME_INIT_THREADS
#define SVTME_INIT_THREADS 32
#endif
__global__ void __launch_bounds__(kStThreads) st_init_kernel(const __grid_constant__ MeParams p, int count, int finish) {
    PDL_ENTRY();
    __shared__ __align__(16) unsigned char head[kSbHeadBytes];
    __shared__ S16Plan plan;
    SbState *home = &p.states[blockIdx.x];
    for (int i = threadIdx.x; i < kHeadVecs; i += blockDim.x) reinterpret_cast<uint4 *>(head)[i] = make_uint4(0u, 0u, 0u, 0u);
    // the result tables: unsearched references report MAX_SAD_VALUE (convention of include/svtme_cuda.h), p_sb_best_mv = 0
    for (int i = threadIdx.x; i < kMaxLists * kMaxRefs * kPus; i += blockDim.x) {
        (&home->best_sad[0][0][0])[i] = SVTME_MAX_SAD_VALUE;
        (&home->best_mv[0][0][0])[i]  = 0u;
    }
    for (int i = threadIdx.x; i < kPus; i += blockDim.x) home->me_distortion[i] = 0u;
    __syncthreads();
    if (threadIdx.x == 0) sb_init_head(p, *reinterpret_cast<SbState *>(head), home, blockIdx.x);
    __syncthreads();
    run_tail(p, head, plan, 0, count, finish);
}

// ---- host: run the program of one picture as a stream-ordered sequence of picture-wide kernels ----
// One kernel per ENGINE step (pre-HME / HME level 0 / 1 / 2 searches, the 1x1 probe, the full-pel scan) plus the first
// one; the control and svt_nxm_sad_kernel steps between two engines run in the tail of the kernel in front of them, the
// results (candidates, statistics) in the tail of the last kernel: 7 launches per picture at preset M6 where the first
// version of this pipeline needed 21.
cudaError_t staged_set_attributes() {  // per device: call with the device current (svtme_init)
    cudaError_t e = cudaFuncSetAttribute(st_sad16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kStWarps * kS16BufWords * 4);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(st_fp_sub_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFpBufWords * 4);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(st_fp_warp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kStWarps * kFpWBufWords * 4);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(st_fp_sub_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFpBufWords * 4);
    return e;
}

cudaError_t launch_me_picture_staged(const MeParams &p, const S16Maps &maps, const RowsMaps &rmaps, int sm_count, cudaStream_t st, int *n_launches) {
    const int n_sb = (int)(p.wb * p.hb);
#ifdef SVTME_DIAG
    static const int engine_ctas = getenv("SVTME_ENGINE_CTAS") ? atoi(getenv("SVTME_ENGINE_CTAS")) : SVTME_ENGINE_CTAS;
#else
    constexpr int engine_ctas = SVTME_ENGINE_CTAS;
#endif
    int       nl   = 0;
#ifdef SVTME_DIAG
    // diagnostics build only: SVTME_TRACE=1 prints per-kernel device times of every picture on stderr (serialises the
    // stream); SVTME_PDL=0 launches every kernel fully stream-ordered
    static const bool trace = getenv("SVTME_TRACE") != nullptr;
    static const bool pdl = !(getenv("SVTME_PDL") != nullptr && atoi(getenv("SVTME_PDL")) == 0) && !trace;
#else
    constexpr bool trace = false, pdl = true;
#endif
    cudaEvent_t ev[kMaxSteps + 4];
    const char *nm[kMaxSteps + 4];
    int         ne = 0;
    auto mark = [&](const char *name) {
        if (!trace) return;
        cudaEventCreate(&ev[ne]); cudaEventRecord(ev[ne], st); nm[ne] = name; ne++;
    };
    mark("start");
    cudaError_t err = cudaSuccess;
    int threads = kStThreads, grid = n_sb;
    // persistent engine kernels: a share of the CTA slots of the GPU (engine_ctas per SM), see st_sad_rows_kernel
    const int engine_grid = engine_ctas > 0 ? (n_sb < sm_count * engine_ctas ? n_sb : sm_count * engine_ctas) : n_sb;
    auto launch = [&](auto kern, size_t smem, auto... args) {
        if (err != cudaSuccess) return;  // the chain stops at the first failed launch
        cudaLaunchConfig_t   cfg = {};
        cudaLaunchAttribute  at[1];
        cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3((unsigned)threads); cfg.dynamicSmemBytes = smem; cfg.stream = st;
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = pdl ? 1 : 0;
        cfg.attrs = at; cfg.numAttrs = 1;
        err = cudaLaunchKernelEx(&cfg, kern, args...);
        nl++;
    };
    auto is_engine = [&](int i) { return p.prog[i].kind == STEP_SAD || p.prog[i].kind == STEP_FP; };
    int i = 0;
    while (i < p.n_steps && !is_engine(i)) i++;
    threads = SVTME_INIT_THREADS;
    launch(st_init_kernel, 0, p, i, i == p.n_steps ? 1 : 0); mark("init");
    threads = kStThreads;
    while (i < p.n_steps) {
        int j = i + 1;
        while (j < p.n_steps && !is_engine(j)) j++;
        const int tf = i + 1, tc = j - i - 1, fin = j == p.n_steps ? 1 : 0;
        grid = engine_grid;
        if (p.prog[i].kind == STEP_SAD) {
            const int lvl = (int)p.prog[i].arg;
            if (lvl == 1 && p.k.hme_search_method == 0) { launch(st_sad_rows_kernel<8, 16>, 0, p, rmaps, lvl); mark("rows8"); }
            else if (lvl == 0 && p.k.hme_search_method == 0) { launch(st_sad_rows_kernel<16, 32>, 0, p, rmaps, lvl); mark("rows16"); }
            else if (lvl == 2 && p.k.hme_search_method == 0) {
                // (diagnostics) 1: pre-HME windows, 2: HME level-0 quadrants -- by the control step that made the jobs
                int tag = 2, qi = -1;
                for (int q = i - 1; q >= 0; q--) {
                    if (qi < 0 && p.prog[q].kind == STEP_PLAN16) qi = q;  // the plan step in front of this engine names the unit queue
                    if (p.prog[q].kind == STEP_CTL && (p.prog[q].op == C_PREHME_MAKE || p.prog[q].op == C_L0_MAKE)) { tag = p.prog[q].op == C_PREHME_MAKE ? 1 : 2; break; }
                }
                // persistent warps that pull units from the queue: every CTA slot of the GPU, whatever the picture size
#ifdef SVTME_DIAG
                static const int s16_ctas = getenv("SVTME_S16_CTAS") ? atoi(getenv("SVTME_S16_CTAS")) : SVTME_S16_CTAS;
#else
                constexpr int s16_ctas = SVTME_S16_CTAS;
#endif
                grid = sm_count * s16_ctas;
                launch(st_sad16_kernel, (size_t)kStWarps * kS16BufWords * 4, p, maps, qi, tag); mark(tag == 1 ? "prehme" : "hme0");
                grid = engine_grid;
            }
            else { launch(st_sad_kernel, 0, p, lvl); mark("sadg"); }
        } else {
            if (p.k.me_search_method == 0) {
                // small windows: one warp per block; the CTA-wide kernel behind it takes the jobs that did not fit a warp
                grid = (n_sb + kStWarps / SVTME_FPW_SPLIT - 1) / (kStWarps / SVTME_FPW_SPLIT);
                launch(st_fp_warp_kernel, (size_t)kStWarps * kFpWBufWords * 4, p, i); mark("fpw");
                grid = engine_grid;
                launch(st_fp_sub_kernel<false>, (size_t)kFpBufWords * 4, p, 1, i); mark("fp");
            } else {
                launch(st_fp_sub_kernel<true>, (size_t)kFpBufWords * 4, p, 0, i); mark("fp");
            }
        }
        grid = n_sb;
        if (tc > 0 || fin) { threads = kLightThreads; launch(st_light_kernel, 0, p, tf, tc, fin); threads = kStThreads; mark("light"); }
        i = j;
    }
    if (trace && err == cudaSuccess) {
        cudaStreamSynchronize(st);
        float tot = 0.f;
        fprintf(stderr, "[svtme trace] tl=%d refs=%d+%d:", p.k.temporal_layer_index, p.k.num_of_ref_pic_to_search[0], p.k.num_of_ref_pic_to_search[1]);
        for (int q = 1; q < ne; q++) {
            float ms = 0.f;
            cudaEventElapsedTime(&ms, ev[q - 1], ev[q]);
            tot += ms;
            fprintf(stderr, " %s %.1f", nm[q], ms * 1e3f);
        }
        fprintf(stderr, " | total %.1f us\n", tot * 1e3f);
    }
    for (int q = 0; q < ne; q++) cudaEventDestroy(ev[q]);
    if (n_launches) *n_launches = nl;
    return err != cudaSuccess ? err : cudaGetLastError();
}

// -DSVTME_PHASES: where thread 0 of the engine CTAs spent its clocks since the last call (svtme_destroy prints it)
void staged_print_phases() {
#ifdef SVTME_PHASES
    unsigned long long ph[4][12];
    if (cudaMemcpyFromSymbol(ph, g_phase, sizeof(ph)) != cudaSuccess) return;
    static const char *nm[4] = {"fp_sub", "sad16 pre-HME", "sad16 HME-L0", "sad_rows"};
    for (int k = 0; k < 4; k++) {
        unsigned long long sum = 0;
        for (int q = 0; q < 12; q++) sum += ph[k][q];
        if (!sum) continue;
        fprintf(stderr, "[svtme phases] %-14s:", nm[k]);
        for (int q = 0; q < 12; q++) if (ph[k][q]) fprintf(stderr, " p%d %.1f%%", q, 100.0 * (double)ph[k][q] / (double)sum);
        fprintf(stderr, " | %.1f Mclk\n", (double)sum / 1e6);
    }
    memset(ph, 0, sizeof(ph));
    cudaMemcpyToSymbol(g_phase, ph, sizeof(ph));
#endif
}

// block states, then the sixteenth-plane plans, then the unit queue of the sixteenth-plane engine
size_t staged_plan_offset(uint32_t n_sb) { return ((size_t)n_sb * sizeof(SbState) + 255) & ~size_t(255); }
size_t staged_queue_offset(uint32_t n_sb) { return (staged_plan_offset(n_sb) + (size_t)n_sb * sizeof(S16Plan) + 255) & ~size_t(255); }
size_t staged_state_bytes(uint32_t n_sb) { return staged_queue_offset(n_sb) + (size_t)n_sb * kS16MaxSub * sizeof(uint32_t); }

}  // namespace svtme
