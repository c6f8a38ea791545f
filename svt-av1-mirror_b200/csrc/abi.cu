// abi.cu -- the extern "C" boundary of libsvtme_cuda.so (include/svtme_cuda.h).
//
// Host-side runtime of the library: device picture pool (three padded luma planes per picture in HBM),
// per-search result arenas (device + pinned host mirror), a stream per in-flight picture, and the RTCD-level
// single-call entry points.  There is NO CPU implementation behind any entry point: every call either runs
// the CUDA kernels or returns an error code.
#include <cuda.h>
#include <cuda_runtime.h>
#include <algorithm>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/svtme_cuda.h"
#include "me_launch.h"
#include "me_control.cuh"

using namespace svtme;

namespace {

// Device plane geometry: padding >= the reference's (72 / 32 / 16: enc_handle.c:4057-4070, reference_object.c:267-295)
// rounded so that pixel (0,0) and every row start are 16-byte aligned and aligned vector loads of a search window
// never leave the allocation.
constexpr const int *kPad = svtme::kDevPad;
constexpr int kRefOrg[3] = {72, 32, 16};

struct PlaneGeom {
    int    w, h, pad, stride;
    size_t offset, bytes;
};

// One pooled allocation: three padded planes.  A buffer is only ever reused for a picture of the SAME size, so the TMA
// descriptors of its sixteenth plane are encoded once, on the host, when it is allocated; a search hands them to
// st_sad16_kernel by value (S16Maps, me_types.cuh).
struct PicBuffer {
    uint8_t    *base = nullptr;
    size_t      bytes = 0;
    uint32_t    w = 0, h = 0;
    cudaEvent_t free_ev = nullptr;  // recorded at release, after every reader of the previous owner: the next owner waits on it
    uint16_t   *staging16 = nullptr;  // 10-bit uploads land here before the unpack kernel; allocated on the first such upload
                                      // and kept with the buffer (a cudaMalloc / cudaFree pair per picture serialises the device)
    bool        recycled = false;
    unsigned char (*maps)[kTmapBytes] = nullptr;  // host: the kS16Classes x kS16HClasses CUtensorMaps of the sixteenth plane (64-byte aligned)
};

}  // namespace

struct SvtMePic {
    uint64_t    picture_number;
    uint32_t    width, height;
    PicBuffer   buf;
    DevPic      dev;
    PlaneGeom   geom[3];
    cudaEvent_t ready;
    cudaStream_t stream;
    uint16_t   *staging16;
    uint64_t    reader_slots;  // search slots whose streams have read this picture (bit = slot index; guarded by the handle's mutex)
    bool        dg_reader;     // the dynamic-GOP detector / statistics stream has read it
    std::vector<cudaEvent_t> peer_reads;  // copies of this picture to other handles / devices still in flight (svtme_picture_replicate)
};

namespace {

struct ResultLayout {
    size_t   off[20];
    size_t   bytes[20];
    size_t   total;
};
enum { R_MV, R_CAND, R_COUNT, R_D64, R_D32, R_D16, R_D8, R_VAR, R_RC, R_STAT, R_GM, R_SAD, R_BMV, R_DOREF, R_HMESC, R_TFEXIT, R_TFCNT, R_PSUM, R_TFSAD, R_TFMV, R_N };

struct Inflight {
    cudaStream_t        stream = nullptr;
    cudaEvent_t         ev_start = nullptr, ev_kernel = nullptr, ev_done = nullptr;
    uint8_t            *d_res = nullptr, *h_res = nullptr;
    size_t              cap = 0;
    uint32_t           *d_counter = nullptr;
    uint8_t            *d_states = nullptr;
    size_t              states_cap = 0;
    ResultLayout        lay{};
    SvtMePictureResults user{};
    bool                to_host = false, busy = false;
    std::thread::id     owner;  // the thread that submitted the search in flight
    bool                direct[32] = {};  // result array i went straight into the caller's (page-locked) buffer
    S16Maps             maps;             // launch parameter of st_sad16_kernel: the references' TMA descriptors
    svtme::RowsMaps     rmaps;            // launch parameter of st_sad_rows_kernel<16, 32>
    uint64_t            ticket = 0;
    float               ms = 0.f;
    unsigned long long *h_px = nullptr;  // pinned
    unsigned long long  px = 0;
    // TPL source-based search of the same picture (svtme_picture_search_tpl / svtme_tpl_src_search)
    uint8_t            *d_tpl = nullptr, *h_tpl = nullptr;
    size_t              tpl_cap = 0, tpl_cells = 0;
    SvtMeTplResults     tpl_user{};
    bool                tpl_to_host = false;
};

// device / pinned arena of one picture's TPL decisions: the seven SvtMeTplResults arrays, each 16-byte aligned
struct TplLayout {
    size_t off[7], bytes[7], total;
};
inline void tpl_layout(size_t cells, TplLayout *L) {
    const size_t elem[7] = {1, 1, 1, 4, 8, 4, 4};  // block_size, best_mode, best_rf_idx, mv[2], ref_frame_poc, inter, intra
    size_t off = 0;
    for (int i = 0; i < 7; i++) {
        off = (off + 15) & ~size_t(15);
        L->off[i] = off; L->bytes[i] = elem[i] * cells;
        off += L->bytes[i];
    }
    L->total = (off + 15) & ~size_t(15);
}

}  // namespace

struct SvtMeHandle {
    SvtMeDeviceCfg         cfg;
    std::mutex             mu;
    std::condition_variable slot_cv;  // a search slot was freed
    std::vector<PicBuffer> free_bufs;
    std::vector<Inflight>  slots;
    std::vector<cudaStream_t> up_streams;
    size_t                 next_up = 0;
    uint64_t               next_ticket = 1;
    uint64_t               launches = 0;
    int                    sm_count = 0;
    // RTCD-level scratch
    uint8_t               *d_scratch = nullptr;
    size_t                 scratch_bytes = 0;
    uint32_t              *d_small = nullptr;
    float                  last_ms = 0.f;
    void                  *encode_tiled = nullptr;  // cuTensorMapEncodeTiled (driver entry point)
    // dynamic-GOP detector (svtme_dg_detector_hme_level0): its own stream, events and result arena
    std::mutex             dg_mu;
    cudaStream_t           dg_stream = nullptr;
    cudaEvent_t            dg_e0 = nullptr, dg_e1 = nullptr, dg_done = nullptr;
    uint8_t               *d_dg = nullptr, *h_dg = nullptr;
    size_t                 dg_cap = 0;
};

namespace {

// Error text of the calling thread's last failed call (svtme_last_error): thread-local, so concurrent callers of one
// handle never write the same string.
thread_local std::string t_err;
int fail(SvtMeHandle *, int code, const char *what, cudaError_t e = cudaSuccess) {
    t_err = what;
    if (e != cudaSuccess) { t_err += ": "; t_err += cudaGetErrorString(e); }
    return code;
}
#define CK(call)                                                                  \
    do {                                                                          \
        cudaError_t e_ = (call);                                                  \
        if (e_ != cudaSuccess) return fail(h, SVTME_ERR_UNDEFINED, #call, e_);    \
    } while (0)

void plane_geometry(uint32_t w, uint32_t h, PlaneGeom g[3], size_t *total) {
    size_t off = 0;
    for (int l = 0; l < 3; l++) {
        g[l].w = (int)(w >> l); g[l].h = (int)(h >> l); g[l].pad = kPad[l];
        g[l].stride = (g[l].w + 2 * g[l].pad + 127) & ~127;
        g[l].offset = off;
        g[l].bytes  = (size_t)g[l].stride * (g[l].h + 2 * g[l].pad) + 256;  // slack for aligned over-reads of the last row
        off += (g[l].bytes + 255) & ~size_t(255);
    }
    *total = off;
}

inline size_t al4k(size_t v) { return (v + 4095) & ~size_t(4095); }

int set_device(SvtMeHandle *h) {
    CK(cudaSetDevice(h->cfg.device));
    return SVTME_OK;
}

int acquire_buffer(SvtMeHandle *h, uint32_t w, uint32_t hgt, size_t bytes, PicBuffer *out) {
    for (size_t i = 0; i < h->free_bufs.size(); i++)
        if (h->free_bufs[i].w == w && h->free_bufs[i].h == hgt && h->free_bufs[i].bytes == bytes) {
            *out = h->free_bufs[i];
            h->free_bufs.erase(h->free_bufs.begin() + i);
            return SVTME_OK;
        }
    PicBuffer b;
    cudaError_t e = cudaMalloc(&b.base, bytes);
    if (e != cudaSuccess) return fail(h, SVTME_ERR_INSUFFICIENT, "cudaMalloc(picture planes)", e);
    b.bytes = bytes; b.w = w; b.h = hgt;
    e = cudaEventCreateWithFlags(&b.free_ev, cudaEventDisableTiming);
    if (e != cudaSuccess) { cudaFree(b.base); return fail(h, SVTME_ERR_UNDEFINED, "cudaEventCreate", e); }
    *out = b;
    return SVTME_OK;
}

int make_pic(SvtMeHandle *h, uint64_t picture_number, uint32_t width, uint32_t height, SvtMePic **out) {
    if (!h || !out || width == 0 || height == 0 || (width & 7) || (height & 7)) return fail(h, SVTME_ERR_BAD_PARAMETER, "width/height must be non-zero multiples of 8");
    if (width > 16384 || height > 16384) return fail(h, SVTME_ERR_BAD_PARAMETER, "picture too large (int16 MV arithmetic)");
    int rc = set_device(h);
    if (rc) return rc;
    SvtMePic *p = new SvtMePic();
    p->picture_number = picture_number; p->width = width; p->height = height; p->staging16 = nullptr;
    p->reader_slots = 0; p->dg_reader = false;
    size_t total;
    plane_geometry(width, height, p->geom, &total);
    {
        std::lock_guard<std::mutex> lk(h->mu);
        rc = acquire_buffer(h, width, height, total, &p->buf);
        p->stream = h->up_streams[h->next_up++ % h->up_streams.size()];
    }
    if (rc) { delete p; return rc; }
    for (int l = 0; l < 3; l++) {
        const PlaneGeom &g = p->geom[l];
        p->dev.lvl[l].org     = p->buf.base + g.offset + (size_t)g.pad * g.stride + g.pad;
        p->dev.lvl[l].stride  = g.stride;
        p->dev.lvl[l].width   = g.w;
        p->dev.lvl[l].height  = g.h;
        p->dev.lvl[l].ref_org = kRefOrg[l];
    }
    p->dev.picture_number = picture_number;
    auto give_back = [&](int code) {  // error paths: the buffer returns to the pool
        std::lock_guard<std::mutex> lk(h->mu);
        h->free_bufs.push_back(p->buf);
        delete p;
        return code;
    };
    cudaError_t e = cudaEventCreateWithFlags(&p->ready, cudaEventDisableTiming);
    if (e != cudaSuccess) return give_back(fail(h, SVTME_ERR_UNDEFINED, "cudaEventCreate", e));
    // a recycled buffer: everything its previous owner's readers (searches, detector, its own upload) had in flight at
    // release time precedes `free_ev`
    if (p->buf.recycled) cudaStreamWaitEvent(p->stream, p->buf.free_ev, 0);
    // TMA descriptors of the padded sixteenth plane (st_sad16_kernel): encoded once per buffer, kept on the host
    if (!p->buf.maps) {
        typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                     const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
        const PlaneGeom &g = p->geom[2];
        static_assert(sizeof(CUtensorMap) == kTmapBytes, "CUtensorMap is 128 bytes");
        void *mem = nullptr;
        if (posix_memalign(&mem, 64, (size_t)svtme::kMapsPerPic * kTmapBytes) != 0) { cudaEventDestroy(p->ready); return give_back(fail(h, SVTME_ERR_INSUFFICIENT, "tensor maps")); }
        p->buf.maps = reinterpret_cast<unsigned char (*)[kTmapBytes]>(mem);
        for (int c = 0; c < kS16Classes * kS16HClasses; c++) {
            const cuuint64_t gdim[2] = {(cuuint64_t)g.stride, (cuuint64_t)(g.h + 2 * g.pad)}, gstr[1] = {(cuuint64_t)g.stride};
            const cuuint32_t box[2] = {(cuuint32_t)kS16BoxBytes[c / kS16HClasses], (cuuint32_t)kS16BoxRows[c % kS16HClasses]}, estr[2] = {1, 1};
            const CUresult   r = reinterpret_cast<EncodeFn>(h->encode_tiled)(
                reinterpret_cast<CUtensorMap *>(p->buf.maps[c]), CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, p->buf.base + g.offset, gdim, gstr, box, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) { cudaEventDestroy(p->ready); return give_back(fail(h, SVTME_ERR_UNDEFINED, "cuTensorMapEncodeTiled")); }
        }
        for (int q = 0; q < svtme::kRowsTmaMaps; q++) {  // rows of parity q of the padded full plane (HME level-2 windows)
            const int        par = q;
            const PlaneGeom &f = p->geom[0];
            const cuuint64_t rows = (cuuint64_t)(f.h + 2 * f.pad);
            const cuuint64_t gdim[2] = {(cuuint64_t)f.stride, (rows - par + 1) / 2}, gstr[1] = {(cuuint64_t)f.stride * 2};
            const cuuint32_t box[2] = {(cuuint32_t)svtme::kRowsTmaBoxW, (cuuint32_t)svtme::kRowsTmaBoxH}, estr[2] = {1, 1};
            const CUresult   r = reinterpret_cast<EncodeFn>(h->encode_tiled)(
                reinterpret_cast<CUtensorMap *>(p->buf.maps[kS16Classes * kS16HClasses + q]), CU_TENSOR_MAP_DATA_TYPE_UINT8, 2,
                p->buf.base + f.offset + (size_t)par * f.stride, gdim, gstr, box, estr,
                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) { cudaEventDestroy(p->ready); return give_back(fail(h, SVTME_ERR_UNDEFINED, "cuTensorMapEncodeTiled (parity planes)")); }
        }
    }
    *out = p;
    return SVTME_OK;
}

// pad the full plane, build quarter + sixteenth (+ padding), signal `ready`
int build_planes(SvtMeHandle *h, SvtMePic *p) {
    uint8_t *org = const_cast<uint8_t *>(p->dev.lvl[0].org);
    CK(launch_pad_plane(org, p->geom[0].stride, p->geom[0].w, p->geom[0].h, p->geom[0].pad, p->stream));
    CK(launch_decimate(p->dev, p->geom[1].pad, p->geom[2].pad, h->cfg.sixteenth_from_full ? 1 : 0, p->stream));
    CK(cudaEventRecord(p->ready, p->stream));
    std::lock_guard<std::mutex> lk(h->mu);
    h->launches += 2;
    return SVTME_OK;
}

void result_layout(uint32_t n_b64, uint32_t pu_count, const SvtMeKnobs &k, bool extended, ResultLayout *L) {
    const size_t n = n_b64;
    size_t sz[R_N] = {n * pu_count * k.max_refs * 4, n * pu_count * k.max_cand, n * pu_count, n * 4, n * 4, n * 4, n * 4,
                      n * 4, n * 4, n, n, extended ? n * 680 * 4 : 0, extended ? n * 680 * 4 : 0, extended ? n * 8 : 0,
                      extended ? n * 16 * 2 : 0, 0, 0, (size_t)svtme::kSumSlots * svtme::kSumSlotWords * 8, 0, 0};
    if (k.me_type == SVTME_ME_MCTF) {
        // the temporal filter reads no candidates or statistics (motion_estimation.c:3126): only MVs / SADs / centres
        for (int i = R_MV; i <= R_GM; i++) sz[i] = 0;
        sz[R_TFEXIT] = n; sz[R_TFCNT] = 8; sz[R_PSUM] = 0; sz[R_TFSAD] = sz[R_TFMV] = n * 85 * 4;
    }
    size_t off = 0;
    for (int i = 0; i < R_N; i++) {
        L->off[i] = off; L->bytes[i] = sz[i];
        off += (sz[i] + 255) & ~size_t(255);
    }
    L->total = off;
}

int validate_knobs(SvtMeHandle *h, const SvtMeKnobs *k, uint32_t pu_count) {
    if (k->num_of_list_to_search < 1 || k->num_of_list_to_search > SVTME_MAX_LISTS) return fail(h, SVTME_ERR_BAD_PARAMETER, "num_of_list_to_search");
    for (int l = 0; l < SVTME_MAX_LISTS; l++)
        if (k->num_of_ref_pic_to_search[l] > SVTME_MAX_REFS_PER_LIST) return fail(h, SVTME_ERR_BAD_PARAMETER, "num_of_ref_pic_to_search");
    if (k->num_of_ref_pic_to_search[0] < 1) return fail(h, SVTME_ERR_BAD_PARAMETER, "list 0 needs at least one reference");
    if (k->num_hme_sa_w < 1 || k->num_hme_sa_w > 2 || k->num_hme_sa_h < 1 || k->num_hme_sa_h > 2) return fail(h, SVTME_ERR_BAD_PARAMETER, "num_hme_sa_w/h");
    if (k->max_number_of_pus_per_sb != SVTME_SQUARE_PU_COUNT) return fail(h, SVTME_ERR_BAD_PARAMETER, "max_number_of_pus_per_sb must be 85");
    const uint32_t want = k->enable_me_16x16 ? (k->enable_me_8x8 ? 85u : 21u) : 5u;
    if (pu_count != want) return fail(h, SVTME_ERR_BAD_PARAMETER, "pu_count does not match enable_me_8x8/16x16");
    if (k->me_type == SVTME_ME_MCTF) {  // one neighbouring picture per call, no candidate arrays (temporal_filtering.c:3140-3142)
        if (k->num_of_list_to_search != 1 || k->num_of_ref_pic_to_search[0] != 1)
            return fail(h, SVTME_ERR_BAD_PARAMETER, "temporal-filter mode searches exactly one reference");
        return SVTME_OK;
    }
    if (k->me_type != SVTME_ME_OPEN_LOOP) return fail(h, SVTME_ERR_BAD_PARAMETER, "me_type");
    if (k->max_refs == 0 || k->max_cand == 0) return fail(h, SVTME_ERR_BAD_PARAMETER, "max_refs/max_cand");
    const uint32_t n0 = k->num_of_ref_pic_to_search[0], n1 = k->num_of_list_to_search > 1 ? k->num_of_ref_pic_to_search[1] : 0;
    if (k->max_l0 < n0 || (uint32_t)k->max_l0 + n1 > k->max_refs) return fail(h, SVTME_ERR_BAD_PARAMETER, "max_l0/max_refs too small for the reference lists");
    const uint32_t need_cand = (n0 == 1 && n1 <= 1) ? (n1 ? 3u : 1u) : n0 + n1 + n0 * n1 + (n0 ? n0 - 1 : 0) + (n1 == 3 ? 1 : 0);
    if (k->max_cand < need_cand) return fail(h, SVTME_ERR_BAD_PARAMETER, "max_cand too small for the reference lists");
    return SVTME_OK;
}

int search_impl(SvtMeHandle *h, SvtMePic *cur, SvtMePic *const refs[SVTME_MAX_LISTS][SVTME_MAX_REFS_PER_LIST],
                const SvtMeKnobs *knobs, SvtMePictureResults *out, uint32_t pu_count_dev, bool extended, uint64_t *ticket,
                const SvtMeTplCfg *tpl = nullptr, SvtMeTplResults *tpl_out = nullptr, const SvtMePictureResults *me_in = nullptr) {
    if (!h || !cur || !refs || !knobs || !ticket) return fail(h, SVTME_ERR_BAD_PARAMETER, "null argument");
    if (tpl) {  // the controls this kernel restates (include/svtme_cuda.h); anything else is refused, not approximated
        if (!tpl_out) return fail(h, SVTME_ERR_BAD_PARAMETER, "tpl_out is NULL");
        if (knobs->me_type == SVTME_ME_MCTF) return fail(h, SVTME_ERR_BAD_PARAMETER, "TPL search needs open-loop ME results");
        if (!tpl->use_sad_in_src_search || !tpl->subpel_full_pel || !tpl->intra_dc_only)
            return fail(h, SVTME_ERR_BAD_PARAMETER, "TPL source search supports use_sad_in_src_search = 1, subpel_depth = FULL_PEL, intra_mode_end = DC_PRED only");
        if (tpl->dispenser_search_level > 2 || (tpl->dispenser_search_level == 2 && !knobs->enable_me_16x16))
            return fail(h, SVTME_ERR_BAD_PARAMETER, "TPL dispenser_search_level");
        if (tpl_out->grid_width < ((cur->width + 15) >> 4) || tpl_out->grid_height < ((cur->height + 15) >> 4))
            return fail(h, SVTME_ERR_BAD_PARAMETER, "TPL result grid too small");
        if (!tpl_out->block_size || !tpl_out->best_mode || !tpl_out->best_rf_idx || !tpl_out->mv || !tpl_out->ref_frame_poc ||
            !tpl_out->best_inter_cost || !tpl_out->best_intra_cost)
            return fail(h, SVTME_ERR_BAD_PARAMETER, "TPL result arrays");
        if (me_in && (!me_in->me_mv || !me_in->me_cand || !me_in->total_me_candidate_index))
            return fail(h, SVTME_ERR_BAD_PARAMETER, "ME results to search over");
    }
    const uint32_t wb = (cur->width + 63) / 64, hb = (cur->height + 63) / 64, n_b64 = wb * hb;
    const uint32_t pu_count = me_in ? me_in->pu_count : out ? out->pu_count : pu_count_dev;
    const bool intra_only_tpl = tpl && me_in && tpl->slice_type_i;  // an I picture: no lists, no candidates, nothing of ME is read
    int rc = intra_only_tpl ? SVTME_OK : validate_knobs(h, knobs, pu_count);
    if (rc) return rc;
    if (out && out->b64_total_count < n_b64) return fail(h, SVTME_ERR_BAD_PARAMETER, "b64_total_count too small");
    const int n_lists = intra_only_tpl ? 0 : knobs->num_of_list_to_search;
    for (int l = 0; l < n_lists; l++)
        for (int r = 0; r < knobs->num_of_ref_pic_to_search[l]; r++) {
            if (!refs[l][r]) return fail(h, SVTME_ERR_BAD_PARAMETER, "missing reference picture");
            if (refs[l][r]->width != cur->width || refs[l][r]->height != cur->height) return fail(h, SVTME_ERR_BAD_PARAMETER, "reference size differs");
        }
    rc = set_device(h);
    if (rc) return rc;

    // pick a slot; when all are in flight, block until another thread's svtme_search_wait frees one (the ME kernel
    // threads of an encoder each submit and wait for one picture).  If every slot is held by THIS thread nobody else can
    // free one: that is a caller error, reported instead of deadlocking.
    Inflight *sl = nullptr;
    {
        std::unique_lock<std::mutex> lk(h->mu);
        for (;;) {
            bool all_mine = true;
            for (auto &s : h->slots) {
                if (!s.busy) { sl = &s; break; }
                if (s.owner != std::this_thread::get_id()) all_mine = false;
            }
            if (sl) break;
            if (all_mine) return fail(h, SVTME_ERR_INSUFFICIENT, "all search slots are held by this thread: call svtme_search_wait first");
            h->slot_cv.wait(lk);
        }
        sl->busy = true;
        sl->owner = std::this_thread::get_id();
        sl->ticket = h->next_ticket++;
    }
    auto release_slot = [&](int code) {
        { std::lock_guard<std::mutex> lk(h->mu); sl->busy = false; }
        h->slot_cv.notify_one();
        return code;
    };
    result_layout(n_b64, pu_count, *knobs, extended, &sl->lay);
    if (sl->cap < sl->lay.total) {
        if (sl->d_res) cudaFree(sl->d_res);
        if (sl->h_res) cudaFreeHost(sl->h_res);
        sl->d_res = nullptr; sl->h_res = nullptr; sl->cap = 0;
        cudaError_t e = cudaMalloc(&sl->d_res, sl->lay.total);
        if (e == cudaSuccess) e = cudaMallocHost(&sl->h_res, sl->lay.total);
        if (e != cudaSuccess) return release_slot(fail(h, SVTME_ERR_INSUFFICIENT, "result arena", e));
        sl->cap = sl->lay.total;
    }
    if (sl->states_cap < staged_state_bytes(n_b64)) {
        if (sl->d_states) cudaFree(sl->d_states);
        sl->d_states = nullptr; sl->states_cap = 0;
        cudaError_t e = cudaMalloc(&sl->d_states, staged_state_bytes(n_b64));
        if (e != cudaSuccess) return release_slot(fail(h, SVTME_ERR_INSUFFICIENT, "block state arena", e));
        sl->states_cap = staged_state_bytes(n_b64);
    }
    TplLayout TL{};
    if (tpl) {
        sl->tpl_cells = (size_t)tpl_out->grid_width * tpl_out->grid_height;
        tpl_layout(sl->tpl_cells, &TL);
        if (sl->tpl_cap < TL.total) {
            if (sl->d_tpl) cudaFree(sl->d_tpl);
            if (sl->h_tpl) cudaFreeHost(sl->h_tpl);
            sl->d_tpl = nullptr; sl->h_tpl = nullptr; sl->tpl_cap = 0;
            cudaError_t e = cudaMalloc(&sl->d_tpl, TL.total);
            if (e == cudaSuccess) e = cudaMallocHost(&sl->h_tpl, TL.total);
            if (e != cudaSuccess) return release_slot(fail(h, SVTME_ERR_INSUFFICIENT, "TPL result arena", e));
            sl->tpl_cap = TL.total;
        }
    }
    MeParams p;
    memset(&p, 0, sizeof(p));
    S16Maps *maps = &sl->maps;  // unused entries keep whatever they held: the kernel only reads the maps of the picture's references
    p.k   = *knobs;
    p.cur = cur->dev;
    for (int l = 0; l < n_lists; l++)
        for (int r = 0; r < knobs->num_of_ref_pic_to_search[l]; r++) {
            p.refs[l][r] = refs[l][r]->dev;
            memcpy(maps->m[l * kMaxRefs + r], refs[l][r]->buf.maps, sizeof(maps->m[0]));
            memcpy(sl->rmaps.m[l * kMaxRefs + r], refs[l][r]->buf.maps[kS16Classes * kS16HClasses], sizeof(sl->rmaps.m[0]));
        }
    p.n_steps = build_program(p.k, p.prog);
    p.wb = wb; p.hb = hb; p.pu_count = pu_count; p.work_counter = sl->d_counter;
    p.px_counter = reinterpret_cast<unsigned long long *>(sl->d_counter + 2);
    p.states = reinterpret_cast<SbState *>(sl->d_states);
    p.plans16 = sl->d_states + staged_plan_offset(n_b64);
    p.q_units = reinterpret_cast<uint32_t *>(sl->d_states + staged_queue_offset(n_b64));
    p.q_ctr   = sl->d_counter + 16;  // behind the work / pixel counters
    uint8_t *d = sl->d_res;
    const ResultLayout &L = sl->lay;
    p.out.me_mv = (uint32_t *)(d + L.off[R_MV]); p.out.me_cand = d + L.off[R_CAND]; p.out.total_me_candidate_index = d + L.off[R_COUNT];
    p.out.me_64x64_distortion = (uint32_t *)(d + L.off[R_D64]); p.out.me_32x32_distortion = (uint32_t *)(d + L.off[R_D32]);
    p.out.me_16x16_distortion = (uint32_t *)(d + L.off[R_D16]); p.out.me_8x8_distortion = (uint32_t *)(d + L.off[R_D8]);
    p.out.me_8x8_cost_variance = (uint32_t *)(d + L.off[R_VAR]); p.out.rc_me_distortion = (uint32_t *)(d + L.off[R_RC]);
    p.out.stationary_block_present_sb = d + L.off[R_STAT]; p.out.rc_me_allow_gm = d + L.off[R_GM];
    if (extended) {
        p.out.sb_best_sad = (uint32_t *)(d + L.off[R_SAD]); p.out.sb_best_mv = (uint32_t *)(d + L.off[R_BMV]);
        p.out.do_ref = d + L.off[R_DOREF]; p.out.hme_sc = (int16_t *)(d + L.off[R_HMESC]);
    }
    if (knobs->me_type == SVTME_ME_MCTF) {
        p.out.tf_exit_sb = d + L.off[R_TFEXIT]; p.out.tf_blk_counts = (uint32_t *)(d + L.off[R_TFCNT]);
        p.out.tf_best_sad = (uint32_t *)(d + L.off[R_TFSAD]); p.out.tf_best_mv = (uint32_t *)(d + L.off[R_TFMV]);
    } else {
        p.out.picture_sums = (unsigned long long *)(d + L.off[R_PSUM]);
    }
    cudaStream_t st = sl->stream;
    cudaError_t  e  = cudaStreamWaitEvent(st, cur->ready, 0);
    for (int l = 0; l < n_lists && e == cudaSuccess; l++)
        for (int r = 0; r < knobs->num_of_ref_pic_to_search[l] && e == cudaSuccess; r++) e = cudaStreamWaitEvent(st, refs[l][r]->ready, 0);
    if (e == cudaSuccess) e = cudaMemsetAsync(sl->d_counter, 0, 64 + kQueueCounterBytes, st);
    if (e == cudaSuccess && p.out.tf_blk_counts) e = cudaMemsetAsync(p.out.tf_blk_counts, 0, 8, st);
    if (e == cudaSuccess && p.out.picture_sums) e = cudaMemsetAsync(p.out.picture_sums, 0, L.bytes[R_PSUM], st);
    if (e == cudaSuccess) e = cudaEventRecord(sl->ev_start, st);
    int n_launch = 0;
    if (intra_only_tpl) {
    } else if (me_in) {  // TPL over ME results the caller holds: they take the place the search kernels would have written
        const size_t nb = n_b64;
        if (e == cudaSuccess) e = cudaMemcpyAsync(p.out.me_mv, me_in->me_mv, nb * pu_count * knobs->max_refs * 4, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(p.out.me_cand, me_in->me_cand, nb * pu_count * knobs->max_cand, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(p.out.total_me_candidate_index, me_in->total_me_candidate_index, nb * pu_count, cudaMemcpyHostToDevice, st);
    } else if (e == cudaSuccess) e = launch_me_picture_staged(p, *maps, sl->rmaps, h->sm_count, st, &n_launch);
    sl->tpl_to_host = false;
    if (tpl && e == cudaSuccess) {
        svtme::TplParams t;
        memset(&t, 0, sizeof(t));
        t.cur = cur->dev;
        for (int l = 0; l < n_lists; l++)
            for (int r = 0; r < knobs->num_of_ref_pic_to_search[l]; r++) {
                t.refs[l][r] = refs[l][r]->dev;
                if (tpl->ref_excluded[l][r]) t.ref_excluded |= 1u << (l * 4 + r);
                t.ref_poc[l * 4 + r] = tpl->ref_poc[l][r];
            }
        t.me_mv = p.out.me_mv; t.me_cand = p.out.me_cand; t.me_count = p.out.total_me_candidate_index;
        t.pus = pu_count; t.max_refs = knobs->max_refs; t.max_cand = knobs->max_cand; t.max_l0 = knobs->max_l0;
        t.wb = wb; t.hb = hb; t.grid_width = tpl_out->grid_width;
        t.level = tpl->dispenser_search_level; t.disable_intra_pred = tpl->disable_intra_pred;
        t.slice_type_i = tpl->slice_type_i; t.enable_me_16x16 = knobs->enable_me_16x16;
        uint8_t *b = sl->d_tpl;
        t.block_size = b + TL.off[0]; t.best_mode = b + TL.off[1]; t.best_rf_idx = (int8_t *)(b + TL.off[2]);
        t.mv = (int16_t *)(b + TL.off[3]); t.ref_frame_poc = (uint64_t *)(b + TL.off[4]);
        t.best_inter_cost = (uint32_t *)(b + TL.off[5]); t.best_intra_cost = (uint32_t *)(b + TL.off[6]);
        e = cudaMemsetAsync(sl->d_tpl, 0, TL.total, st);
        if (e == cudaSuccess) e = svtme::launch_tpl_src(t, st);
        n_launch += 1;
        sl->tpl_user = *tpl_out;
    }
    if (e == cudaSuccess) e = cudaEventRecord(sl->ev_kernel, st);
    if (tpl && e == cudaSuccess) {
        // caller buffers laid out like the arena (one block, fields 16-byte aligned in order) and page-locked: one copy
        // straight into them; otherwise through the slot's pinned arena, scattered by svtme_search_wait
        uint8_t *base = tpl_out->block_size;
        const void *dst[7] = {tpl_out->block_size, tpl_out->best_mode, tpl_out->best_rf_idx, tpl_out->mv, tpl_out->ref_frame_poc,
                              tpl_out->best_inter_cost, tpl_out->best_intra_cost};
        bool direct = true;
        for (int i = 0; i < 7; i++) direct = direct && dst[i] == base + TL.off[i];
        if (direct) {
            cudaPointerAttributes pa;
            if (!(cudaPointerGetAttributes(&pa, base) == cudaSuccess && pa.type == cudaMemoryTypeHost)) { (void)cudaGetLastError(); direct = false; }
        }
        sl->tpl_to_host = !direct;
        e = cudaMemcpyAsync(direct ? base : sl->h_tpl, sl->d_tpl, direct ? TL.off[6] + TL.bytes[6] : TL.total, cudaMemcpyDeviceToHost, st);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(sl->h_px, sl->d_counter + 2, 8, cudaMemcpyDeviceToHost, st);
    sl->to_host = out != nullptr;
    if (out) {
        sl->user = *out;
        // Large arrays whose destination is page-locked host memory are copied straight into it; the rest goes
        // through the slot's pinned arena (one copy of the covering range) and is scattered by svtme_search_wait.
        void *dst[R_N] = {out->me_mv, out->me_cand, out->total_me_candidate_index, out->me_64x64_distortion, out->me_32x32_distortion,
                          out->me_16x16_distortion, out->me_8x8_distortion, out->me_8x8_cost_variance, out->rc_me_distortion,
                          out->stationary_block_present_sb, out->rc_me_allow_gm, out->sb_best_sad, out->sb_best_mv, out->do_ref, out->hme_sc,
                          out->tf_exit_sb, out->tf_blk_counts, out->picture_sums, out->tf_best_sad, out->tf_best_mv};
        size_t lo = L.total, hi = 0;
        for (int i = 0; i < R_N && e == cudaSuccess; i++) {
            sl->direct[i] = false;
            if (!dst[i] || !L.bytes[i]) continue;
            if (L.bytes[i] >= (64u << 10)) {
                cudaPointerAttributes pa;
                if (cudaPointerGetAttributes(&pa, dst[i]) == cudaSuccess && pa.type == cudaMemoryTypeHost) sl->direct[i] = true;
                else (void)cudaGetLastError();
            }
            if (sl->direct[i]) e = cudaMemcpyAsync(dst[i], sl->d_res + L.off[i], L.bytes[i], cudaMemcpyDeviceToHost, st);
            else { lo = std::min(lo, L.off[i]); hi = std::max(hi, L.off[i] + L.bytes[i]); }
        }
        if (e == cudaSuccess && hi > lo) e = cudaMemcpyAsync(sl->h_res + lo, sl->d_res + lo, hi - lo, cudaMemcpyDeviceToHost, st);
    }
    if (e == cudaSuccess) e = cudaEventRecord(sl->ev_done, st);
    {
        // the planes must outlive this search: every picture remembers the slots whose streams read it, and its release
        // waits for the last thing recorded on each of them (`ev_done`, in stream order behind this search)
        std::lock_guard<std::mutex> lk(h->mu);
        const uint64_t bit = 1ull << (sl - h->slots.data());
        cur->reader_slots |= bit;
        for (int l = 0; l < n_lists; l++)
            for (int r = 0; r < knobs->num_of_ref_pic_to_search[l]; r++) refs[l][r]->reader_slots |= bit;
        h->launches += (uint64_t)n_launch;
    }
    if (e != cudaSuccess) return release_slot(fail(h, SVTME_ERR_UNDEFINED, "search launch", e));
    *ticket = sl->ticket;
    return SVTME_OK;
}

}  // namespace

extern "C" {

int32_t svtme_init(const SvtMeDeviceCfg *cfg, SvtMeHandle **out) {
    if (!cfg || !out) return SVTME_ERR_BAD_PARAMETER;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) return SVTME_ERR_UNDEFINED;  // no CUDA device: no fallback
    if (cfg->device < 0 || cfg->device >= n) return SVTME_ERR_BAD_PARAMETER;
    SvtMeHandle *h = new SvtMeHandle();
    h->cfg = *cfg;
    if (h->cfg.max_pictures == 0) h->cfg.max_pictures = 64;
    if (h->cfg.max_inflight == 0) h->cfg.max_inflight = 4;
    if (h->cfg.max_inflight > 64) h->cfg.max_inflight = 64;  // one bit per slot in SvtMePic::reader_slots
    if (h->cfg.pipeline != 0) { delete h; return SVTME_ERR_BAD_PARAMETER; }  // only the staged pipeline ships
    const uint32_t inflight = h->cfg.max_inflight;
    cudaError_t e = cudaSetDevice(cfg->device);
    cudaDeviceProp prop;
    if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, cfg->device);
    if (e == cudaSuccess && prop.major < 10) { delete h; return SVTME_ERR_UNDEFINED; }  // built for sm_100a only
    if (e == cudaSuccess) h->sm_count = prop.multiProcessorCount;
    // opt-ins to more than 48 KB of dynamic shared memory are per device: set them for THIS handle's device
    if (e == cudaSuccess) e = staged_set_attributes();
    if (e == cudaSuccess) e = rtcd_set_attributes();
    h->slots.resize(inflight);
    for (auto &s : h->slots) {
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreate(&s.ev_start);
        if (e == cudaSuccess) e = cudaEventCreate(&s.ev_kernel);
        if (e == cudaSuccess) e = cudaEventCreate(&s.ev_done);
        if (e == cudaSuccess) e = cudaMalloc(&s.d_counter, 64 + kQueueCounterBytes);
        if (e == cudaSuccess) e = cudaMallocHost(&s.h_px, 64);
    }
    h->up_streams.resize(std::max<uint32_t>(2, inflight));
    for (auto &s : h->up_streams)
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_small, 4096);
    if (e == cudaSuccess) {
        cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
        e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &h->encode_tiled, cudaEnableDefault, &q);
        if (e == cudaSuccess && (q != cudaDriverEntryPointSuccess || !h->encode_tiled)) e = cudaErrorNotSupported;
    }
    if (e != cudaSuccess) { fprintf(stderr, "svtme_init: %s\n", cudaGetErrorString(e)); svtme_destroy(h); return SVTME_ERR_UNDEFINED; }
    *out = h;
    return SVTME_OK;
}

void svtme_destroy(SvtMeHandle *h) {
    if (!h) return;
    cudaSetDevice(h->cfg.device);
    cudaDeviceSynchronize();
    staged_print_phases();
    for (auto &s : h->slots) {
        if (s.stream) cudaStreamDestroy(s.stream);
        if (s.ev_start) cudaEventDestroy(s.ev_start);
        if (s.ev_kernel) cudaEventDestroy(s.ev_kernel);
        if (s.ev_done) cudaEventDestroy(s.ev_done);
        if (s.d_res) cudaFree(s.d_res);
        if (s.h_res) cudaFreeHost(s.h_res);
        if (s.d_tpl) cudaFree(s.d_tpl);
        if (s.h_tpl) cudaFreeHost(s.h_tpl);
        if (s.d_counter) cudaFree(s.d_counter);
        if (s.d_states) cudaFree(s.d_states);
        if (s.h_px) cudaFreeHost(s.h_px);
    }
    for (auto &s : h->up_streams)
        if (s) cudaStreamDestroy(s);
    for (auto &b : h->free_bufs) { cudaFree(b.base); cudaFree(b.staging16); cudaEventDestroy(b.free_ev); free(b.maps); }
    if (h->d_scratch) cudaFree(h->d_scratch);
    if (h->d_small) cudaFree(h->d_small);
    if (h->dg_stream) { cudaStreamDestroy(h->dg_stream); cudaEventDestroy(h->dg_e0); cudaEventDestroy(h->dg_e1); cudaEventDestroy(h->dg_done); }
    if (h->d_dg) cudaFree(h->d_dg);
    if (h->h_dg) cudaFreeHost(h->h_dg);
    delete h;
}

int32_t svtme_picture_register(SvtMeHandle *h, uint64_t picture_number, const uint8_t *luma, uint32_t stride, uint32_t width,
                               uint32_t height, SvtMePic **out) {
    if (!luma || stride < width) return fail(h, SVTME_ERR_BAD_PARAMETER, "luma/stride");
    SvtMePic *p = nullptr;
    int rc = make_pic(h, picture_number, width, height, &p);
    if (rc) return rc;
    cudaError_t e = cudaMemcpy2DAsync(const_cast<uint8_t *>(p->dev.lvl[0].org), p->geom[0].stride, luma, stride, width, height,
                                      cudaMemcpyHostToDevice, p->stream);
    if (e != cudaSuccess) { svtme_picture_release(h, p); return fail(h, SVTME_ERR_UNDEFINED, "cudaMemcpy2DAsync(luma)", e); }
    rc = build_planes(h, p);
    if (rc) { svtme_picture_release(h, p); return rc; }
    *out = p;
    return SVTME_OK;
}

int32_t svtme_picture_register_device(SvtMeHandle *h, uint64_t picture_number, const void *d_luma, uint32_t stride, uint32_t width,
                                      uint32_t height, SvtMePic **out) {
    if (!d_luma || stride < width) return fail(h, SVTME_ERR_BAD_PARAMETER, "luma/stride");
    SvtMePic *p = nullptr;
    int rc = make_pic(h, picture_number, width, height, &p);
    if (rc) return rc;
    cudaError_t e = cudaMemcpy2DAsync(const_cast<uint8_t *>(p->dev.lvl[0].org), p->geom[0].stride, d_luma, stride, width, height,
                                      cudaMemcpyDeviceToDevice, p->stream);
    if (e != cudaSuccess) { svtme_picture_release(h, p); return fail(h, SVTME_ERR_UNDEFINED, "cudaMemcpy2DAsync(device luma)", e); }
    rc = build_planes(h, p);
    if (rc) { svtme_picture_release(h, p); return rc; }
    *out = p;
    return SVTME_OK;
}

int32_t svtme_picture_register16(SvtMeHandle *h, uint64_t picture_number, const uint16_t *luma16, uint32_t stride, uint32_t width,
                                 uint32_t height, SvtMePic **out) {
    if (!luma16 || stride < width) return fail(h, SVTME_ERR_BAD_PARAMETER, "luma/stride");
    SvtMePic *p = nullptr;
    int rc = make_pic(h, picture_number, width, height, &p);
    if (rc) return rc;
    cudaError_t e = cudaSuccess;
    if (!p->buf.staging16) e = cudaMalloc(&p->buf.staging16, (size_t)width * height * 2);  // buffers are pooled per (width, height)
    p->staging16 = p->buf.staging16;
    if (e == cudaSuccess)
        e = cudaMemcpy2DAsync(p->staging16, (size_t)width * 2, luma16, (size_t)stride * 2, (size_t)width * 2, height, cudaMemcpyHostToDevice, p->stream);
    if (e == cudaSuccess)
        e = launch_unpack16(p->staging16, (int)width, const_cast<uint8_t *>(p->dev.lvl[0].org), p->geom[0].stride, (int)width, (int)height, p->stream);
    if (e != cudaSuccess) { svtme_picture_release(h, p); return fail(h, SVTME_ERR_UNDEFINED, "10-bit upload", e); }
    { std::lock_guard<std::mutex> lk(h->mu); h->launches += 1; }
    rc = build_planes(h, p);
    if (rc) { svtme_picture_release(h, p); return rc; }
    *out = p;
    return SVTME_OK;
}

// ---- reference planes across GPUs: a picture is decimated once, on the GPU that owns it, and
// its three padded planes -- one allocation, the same layout for a given picture size on every device -- are copied to
// the GPUs whose searches reference it ----
int32_t svtme_picture_planes_bytes(uint32_t width, uint32_t height, uint64_t *bytes) {
    if (!bytes || width == 0 || height == 0 || (width & 7) || (height & 7)) return SVTME_ERR_BAD_PARAMETER;
    PlaneGeom g[3];
    size_t    total;
    plane_geometry(width, height, g, &total);
    *bytes = total;
    return SVTME_OK;
}

int32_t svtme_picture_replicate(SvtMeHandle *src_h, SvtMePic *pic, SvtMeHandle *dst_h, SvtMePic **out) {
    if (!src_h || !pic || !dst_h || !out) return fail(dst_h, SVTME_ERR_BAD_PARAMETER, "null argument");
    SvtMePic *p = nullptr;
    int rc = make_pic(dst_h, pic->picture_number, pic->width, pic->height, &p);  // makes dst_h's device current
    if (rc) return rc;
    size_t total;
    PlaneGeom g[3];
    plane_geometry(pic->width, pic->height, g, &total);
    const int sd = src_h->cfg.device, dd = dst_h->cfg.device;
    cudaError_t e = cudaSuccess;
    if (sd != dd) {  // direct NVLink path when the devices are peers (an "already enabled" answer is fine)
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, dd, sd) == cudaSuccess && can) {
            e = cudaDeviceEnablePeerAccess(sd, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) { (void)cudaGetLastError(); e = cudaSuccess; }
        }
    }
    // the copy runs on the new picture's stream, behind the source's decimation
    if (e == cudaSuccess) e = cudaStreamWaitEvent(p->stream, pic->ready, 0);
    if (e == cudaSuccess) e = cudaMemcpyPeerAsync(p->buf.base, dd, pic->buf.base, sd, total, p->stream);
    cudaEvent_t done = nullptr;
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&done, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventRecord(done, p->stream);
    if (e == cudaSuccess) e = cudaEventRecord(p->ready, p->stream);
    if (e != cudaSuccess) { if (done) cudaEventDestroy(done); svtme_picture_release(dst_h, p); return fail(dst_h, SVTME_ERR_UNDEFINED, "svtme_picture_replicate", e); }
    { std::lock_guard<std::mutex> lk(src_h->mu); pic->peer_reads.push_back(done); }  // the source's release waits for the copy
    *out = p;
    return SVTME_OK;
}

int32_t svtme_picture_export_planes(SvtMeHandle *h, SvtMePic *pic, void *d_dst, uint64_t dst_bytes) {
    if (!h || !pic || !d_dst) return fail(h, SVTME_ERR_BAD_PARAMETER, "null argument");
    size_t total;
    PlaneGeom g[3];
    plane_geometry(pic->width, pic->height, g, &total);
    if (dst_bytes < total) return fail(h, SVTME_ERR_BAD_PARAMETER, "destination smaller than svtme_picture_planes_bytes");
    int rc = set_device(h);
    if (rc) return rc;
    CK(cudaStreamWaitEvent(pic->stream, pic->ready, 0));
    CK(cudaMemcpyAsync(d_dst, pic->buf.base, total, cudaMemcpyDeviceToDevice, pic->stream));
    CK(cudaStreamSynchronize(pic->stream));  // the bytes are in d_dst when the call returns (the caller's NCCL send comes next)
    return SVTME_OK;
}

int32_t svtme_picture_planes_device(SvtMeHandle *h, SvtMePic *pic, const void **d_planes, uint64_t *bytes) {
    if (!h || !pic || !d_planes) return fail(h, SVTME_ERR_BAD_PARAMETER, "null argument");
    size_t total;
    PlaneGeom g[3];
    plane_geometry(pic->width, pic->height, g, &total);
    int rc = set_device(h);
    if (rc) return rc;
    CK(cudaEventSynchronize(pic->ready));  // the planes are built when the call returns
    *d_planes = pic->buf.base;
    if (bytes) *bytes = total;
    return SVTME_OK;
}

int32_t svtme_picture_import_planes(SvtMeHandle *h, uint64_t picture_number, const void *d_src, uint32_t width, uint32_t height,
                                    SvtMePic **out) {
    if (!d_src) return fail(h, SVTME_ERR_BAD_PARAMETER, "null argument");
    SvtMePic *p = nullptr;
    int rc = make_pic(h, picture_number, width, height, &p);
    if (rc) return rc;
    size_t total;
    PlaneGeom g[3];
    plane_geometry(width, height, g, &total);
    cudaError_t e = cudaMemcpyAsync(p->buf.base, d_src, total, cudaMemcpyDeviceToDevice, p->stream);
    if (e == cudaSuccess) e = cudaEventRecord(p->ready, p->stream);
    if (e != cudaSuccess) { svtme_picture_release(h, p); return fail(h, SVTME_ERR_UNDEFINED, "svtme_picture_import_planes", e); }
    *out = p;
    return SVTME_OK;
}

int32_t svtme_picture_wait(SvtMeHandle *h, SvtMePic *pic) {
    if (!h || !pic) return fail(h, SVTME_ERR_BAD_PARAMETER, "null argument");
    CK(cudaEventSynchronize(pic->ready));
    return SVTME_OK;
}

int32_t svtme_picture_read_plane(SvtMeHandle *h, SvtMePic *pic, int32_t level, uint32_t pad, uint8_t *dst, uint32_t dst_stride) {
    if (!h || !pic || !dst || level < 0 || level > 2) return fail(h, SVTME_ERR_BAD_PARAMETER, "argument");
    const PlaneGeom &g = pic->geom[level];
    if ((int)pad > g.pad || dst_stride < (uint32_t)g.w + 2 * pad) return fail(h, SVTME_ERR_BAD_PARAMETER, "pad/dst_stride");
    int rc = set_device(h);
    if (rc) return rc;
    CK(cudaEventSynchronize(pic->ready));
    const uint8_t *src = pic->dev.lvl[level].org - (size_t)pad * g.stride - pad;
    CK(cudaMemcpy2D(dst, dst_stride, src, g.stride, g.w + 2 * pad, g.h + 2 * pad, cudaMemcpyDeviceToHost));
    return SVTME_OK;
}

int32_t svtme_picture_release(SvtMeHandle *h, SvtMePic *pic) {
    if (!h || !pic) return fail(h, SVTME_ERR_BAD_PARAMETER, "null argument");
    cudaSetDevice(h->cfg.device);
    // The buffer goes back to the pool while work that reads it may still be in flight.  The picture's own stream (its
    // upload and decimation are already in it) waits for the last thing recorded on the stream of every search slot that
    // read the picture and on the detector stream, then `free_ev` is recorded behind all of that; the buffer's next owner
    // waits on `free_ev` before it writes.
    {
        std::lock_guard<std::mutex> lk(h->mu);
        for (size_t i = 0; i < h->slots.size(); i++)
            if (pic->reader_slots & (1ull << i)) cudaStreamWaitEvent(pic->stream, h->slots[i].ev_done, 0);
        if (pic->dg_reader && h->dg_done) cudaStreamWaitEvent(pic->stream, h->dg_done, 0);
        for (cudaEvent_t ev : pic->peer_reads) { cudaStreamWaitEvent(pic->stream, ev, 0); cudaEventDestroy(ev); }
        cudaEventRecord(pic->buf.free_ev, pic->stream);
        pic->buf.recycled = true;
        if (h->free_bufs.size() < h->cfg.max_pictures) h->free_bufs.push_back(pic->buf);
        else { cudaEventSynchronize(pic->buf.free_ev); cudaFree(pic->buf.base); cudaFree(pic->buf.staging16); cudaEventDestroy(pic->buf.free_ev); free(pic->buf.maps); }
    }
    cudaEventDestroy(pic->ready);
    delete pic;
    return SVTME_OK;
}

int32_t svtme_picture_search(SvtMeHandle *h, SvtMePic *cur, SvtMePic *const refs[SVTME_MAX_LISTS][SVTME_MAX_REFS_PER_LIST],
                             const SvtMeKnobs *knobs, SvtMePictureResults *out, uint64_t *ticket) {
    if (out && knobs && knobs->me_type == SVTME_ME_MCTF) {
        const bool full = out->sb_best_mv && out->sb_best_sad, compact = out->tf_best_mv && out->tf_best_sad;
        if (!(full || compact) || !out->hme_sc || !out->tf_exit_sb || !out->tf_blk_counts)
            return fail(h, SVTME_ERR_BAD_PARAMETER, "temporal-filter mode needs sb_best_mv + sb_best_sad (or tf_best_mv + tf_best_sad), hme_sc, tf_exit_sb and tf_blk_counts");
        return search_impl(h, cur, refs, knobs, out, 0, true, ticket);
    }
    if (!out || !out->me_mv || !out->me_cand || !out->total_me_candidate_index || !out->me_64x64_distortion ||
        !out->me_32x32_distortion || !out->me_16x16_distortion || !out->me_8x8_distortion || !out->me_8x8_cost_variance ||
        !out->rc_me_distortion || !out->stationary_block_present_sb || !out->rc_me_allow_gm)
        return fail(h, SVTME_ERR_BAD_PARAMETER, "result buffers");
    const bool extended = out->sb_best_sad || out->sb_best_mv || out->do_ref || out->hme_sc;
    return search_impl(h, cur, refs, knobs, out, 0, extended, ticket);
}

int32_t svtme_picture_search_device(SvtMeHandle *h, SvtMePic *cur, SvtMePic *const refs[SVTME_MAX_LISTS][SVTME_MAX_REFS_PER_LIST],
                                    const SvtMeKnobs *knobs, uint64_t *ticket) {
    if (!knobs) return fail(h, SVTME_ERR_BAD_PARAMETER, "null argument");
    const uint32_t pu = knobs->enable_me_16x16 ? (knobs->enable_me_8x8 ? 85u : 21u) : 5u;
    return search_impl(h, cur, refs, knobs, nullptr, pu, knobs->me_type == SVTME_ME_MCTF, ticket);
}

static Inflight *find_slot(SvtMeHandle *h, uint64_t ticket) {
    for (auto &s : h->slots)
        if (s.ticket == ticket) return &s;
    return nullptr;
}

int32_t svtme_picture_search_tpl(SvtMeHandle *h, SvtMePic *cur, SvtMePic *const refs[SVTME_MAX_LISTS][SVTME_MAX_REFS_PER_LIST],
                                 const SvtMeKnobs *knobs, SvtMePictureResults *out, const SvtMeTplCfg *tpl, SvtMeTplResults *tpl_out,
                                 uint64_t *ticket) {
    if (!tpl || !tpl_out || !knobs) return fail(h, SVTME_ERR_BAD_PARAMETER, "null argument");
    const bool extended = out && (out->sb_best_sad || out->sb_best_mv || out->do_ref || out->hme_sc);
    const uint32_t pus = knobs->enable_me_16x16 ? (knobs->enable_me_8x8 ? 85u : 21u) : 5u;  // pcs.c:106-110
    return search_impl(h, cur, refs, knobs, out, pus, extended, ticket, tpl, tpl_out, nullptr);
}

int32_t svtme_tpl_src_search(SvtMeHandle *h, SvtMePic *cur, SvtMePic *const refs[SVTME_MAX_LISTS][SVTME_MAX_REFS_PER_LIST],
                             const SvtMeKnobs *knobs, const SvtMePictureResults *me, const SvtMeTplCfg *tpl, SvtMeTplResults *tpl_out,
                             uint64_t *ticket) {
    if (!tpl || !tpl_out || !knobs) return fail(h, SVTME_ERR_BAD_PARAMETER, "null argument");
    if (!me && !tpl->slice_type_i) return fail(h, SVTME_ERR_BAD_PARAMETER, "ME results are needed unless the picture is an I slice");
    SvtMePictureResults none;
    memset(&none, 0, sizeof(none));
    static const uint8_t zero = 0;  // an I picture has no candidates: nothing of `me` is read by the kernel
    if (!me) {
        none.pu_count = knobs->enable_me_16x16 ? (knobs->enable_me_8x8 ? 85u : 21u) : 5u;
        none.me_mv = (uint32_t *)&zero; none.me_cand = (uint8_t *)&zero; none.total_me_candidate_index = (uint8_t *)&zero;
    }
    return search_impl(h, cur, refs, knobs, nullptr, 0, false, ticket, tpl, tpl_out, me ? me : &none);
}

int32_t svtme_search_wait(SvtMeHandle *h, uint64_t ticket) {
    if (!h) return SVTME_ERR_BAD_PARAMETER;
    Inflight *sl;
    {
        std::lock_guard<std::mutex> lk(h->mu);
        sl = find_slot(h, ticket);
        if (!sl || !sl->busy) return fail(h, SVTME_ERR_BAD_PARAMETER, "unknown ticket");
    }
    cudaError_t we = cudaSetDevice(h->cfg.device);
    if (we == cudaSuccess) we = cudaEventSynchronize(sl->ev_done);
    if (we != cudaSuccess) {  // the slot is not left busy forever
        { std::lock_guard<std::mutex> lk(h->mu); sl->busy = false; }
        h->slot_cv.notify_one();
        return fail(h, SVTME_ERR_UNDEFINED, "svtme_search_wait", we);
    }
    cudaEventElapsedTime(&sl->ms, sl->ev_start, sl->ev_kernel);
    sl->px = *sl->h_px;
    if (sl->to_host) {
        const ResultLayout &L = sl->lay;
        const SvtMePictureResults &u = sl->user;
        void *dst[R_N] = {u.me_mv, u.me_cand, u.total_me_candidate_index, u.me_64x64_distortion, u.me_32x32_distortion,
                          u.me_16x16_distortion, u.me_8x8_distortion, u.me_8x8_cost_variance, u.rc_me_distortion,
                          u.stationary_block_present_sb, u.rc_me_allow_gm, u.sb_best_sad, u.sb_best_mv, u.do_ref, u.hme_sc,
                          u.tf_exit_sb, u.tf_blk_counts, u.picture_sums, u.tf_best_sad, u.tf_best_mv};
        for (int i = 0; i < R_N; i++)
            if (dst[i] && L.bytes[i] && !sl->direct[i]) {
                if (i == R_PSUM) {  // the device keeps kSumSlots partial sums (me_types.cuh)
                    const uint64_t *part = reinterpret_cast<const uint64_t *>(sl->h_res + L.off[i]);
                    for (int q = 0; q < 4; q++) {
                        uint64_t t = 0;
                        for (int s2 = 0; s2 < svtme::kSumSlots; s2++) t += part[s2 * svtme::kSumSlotWords + q];
                        u.picture_sums[q] = t;
                    }
                } else memcpy(dst[i], sl->h_res + L.off[i], L.bytes[i]);
            }
    }
    if (sl->tpl_to_host) {
        TplLayout TL;
        tpl_layout(sl->tpl_cells, &TL);
        const SvtMeTplResults &u = sl->tpl_user;
        void *dst[7] = {u.block_size, u.best_mode, u.best_rf_idx, u.mv, u.ref_frame_poc, u.best_inter_cost, u.best_intra_cost};
        for (int i = 0; i < 7; i++) memcpy(dst[i], sl->h_tpl + TL.off[i], TL.bytes[i]);
        sl->tpl_to_host = false;
    }
    {
        std::lock_guard<std::mutex> lk(h->mu);
        h->last_ms = sl->ms;
        sl->busy = false;
    }
    h->slot_cv.notify_one();
    return SVTME_OK;
}

int32_t svtme_search_elapsed_ms(SvtMeHandle *h, uint64_t ticket, float *ms) {
    if (!h || !ms) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> lk(h->mu);
    Inflight *sl = find_slot(h, ticket);
    if (!sl || sl->busy) return fail(h, SVTME_ERR_BAD_PARAMETER, "ticket not completed (call svtme_search_wait first)");
    *ms = sl->ms;
    return SVTME_OK;
}

int32_t svtme_search_px_compares(SvtMeHandle *h, uint64_t ticket, uint64_t *px) {
    if (!h || !px) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> lk(h->mu);
    Inflight *sl = find_slot(h, ticket);
    if (!sl || sl->busy) return fail(h, SVTME_ERR_BAD_PARAMETER, "ticket not completed (call svtme_search_wait first)");
    *px = sl->px;
    return SVTME_OK;
}

uint64_t svtme_launch_count(SvtMeHandle *h) {
    if (!h) return 0;
    std::lock_guard<std::mutex> lk(h->mu);
    return h->launches;
}

int32_t svtme_measure_sad_peak(SvtMeHandle *h, double *gsad4_per_s) {
    if (!h || !gsad4_per_s) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> lk(h->mu);
    CK(cudaSetDevice(h->cfg.device));
    const int grid = h->sm_count * 4, iters = 4000;
    uint32_t *d = nullptr;
    CK(cudaMalloc(&d, (size_t)grid * 256 * 4));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    cudaError_t e = cudaSuccess;
    for (int r = 0; r < 6 && e == cudaSuccess; r++) {  // first pass is the warm-up
        cudaEventRecord(e0, 0);
        e = launch_sad_peak(d, grid, iters, 0);
        cudaEventRecord(e1, 0);
        if (e == cudaSuccess) e = cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (r > 0 && ms < best) best = ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
    h->launches += 6;
    if (e != cudaSuccess) return fail(h, SVTME_ERR_UNDEFINED, "sad peak probe", e);
    *gsad4_per_s = (double)grid * 256 * iters * 32.0 / (best * 1e-3) / 1e9;
    return SVTME_OK;
}

int32_t svtme_dg_detector_hme_level0(SvtMeHandle *h, SvtMePic *src, SvtMePic *ref, uint32_t input_resolution,
                                     SvtMeDgMetrics *metrics, uint32_t *sb_sad, int16_t *sb_mv, float *elapsed_ms,
                                     uint64_t *px_compares) {
    if (!h || !src || !ref || !metrics) return fail(h, SVTME_ERR_BAD_PARAMETER, "null argument");
    if (src->width != ref->width || src->height != ref->height) return fail(h, SVTME_ERR_BAD_PARAMETER, "reference size differs");
    int rc = set_device(h);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(h->dg_mu);
    const uint32_t aw = (src->width + 7) & ~7u, ah = (src->height + 7) & ~7u, wb = (aw + 63) / 64, hb = (ah + 63) / 64, n = wb * hb;
    const size_t   off_sad = 256, off_mv = off_sad + al4k((size_t)n * 4), total = off_mv + al4k((size_t)n * 4);
    if (!h->dg_stream) {
        CK(cudaStreamCreateWithFlags(&h->dg_stream, cudaStreamNonBlocking));
        CK(cudaEventCreate(&h->dg_e0));
        CK(cudaEventCreate(&h->dg_e1));
        CK(cudaEventCreateWithFlags(&h->dg_done, cudaEventDisableTiming));
    }
    if (h->dg_cap < total) {
        if (h->d_dg) cudaFree(h->d_dg);
        if (h->h_dg) cudaFreeHost(h->h_dg);
        h->d_dg = h->h_dg = nullptr; h->dg_cap = 0;
        CK(cudaMalloc(&h->d_dg, total));
        CK(cudaMallocHost(&h->h_dg, total));
        h->dg_cap = total;
    }
    cudaStream_t st = h->dg_stream;
    CK(cudaStreamWaitEvent(st, src->ready, 0));
    CK(cudaStreamWaitEvent(st, ref->ready, 0));
    CK(cudaMemsetAsync(h->d_dg, 0, sizeof(svtme::DgSums), st));
    // search area of early_hme (pd_process.c:497-498)
    const int sa = input_resolution <= SVTME_INPUT_SIZE_360p ? 16 : input_resolution <= SVTME_INPUT_SIZE_480p ? 64 : 128;
    CK(cudaEventRecord(h->dg_e0, st));
    CK(svtme::launch_dg_hme(src->dev.lvl[2], ref->dev.lvl[2], sa, wb, hb, reinterpret_cast<uint32_t *>(h->d_dg + off_sad),
                            reinterpret_cast<int16_t *>(h->d_dg + off_mv), reinterpret_cast<svtme::DgSums *>(h->d_dg), st));
    CK(cudaEventRecord(h->dg_e1, st));
    CK(cudaMemcpyAsync(h->h_dg, h->d_dg, (sb_sad || sb_mv) ? total : 256, cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(h->dg_done, st));
    { std::lock_guard<std::mutex> lk2(h->mu); src->dg_reader = ref->dg_reader = true; }
    CK(cudaStreamSynchronize(st));
    const svtme::DgSums *r = reinterpret_cast<const svtme::DgSums *>(h->h_dg);
    metrics->tot_dist = r->tot_dist; metrics->tot_cplx = r->tot_cplx; metrics->tot_active = r->tot_active;
    metrics->sum_in_vectors = r->sum_in_vectors;
    if (sb_sad) memcpy(sb_sad, h->h_dg + off_sad, (size_t)n * 4);
    if (sb_mv) memcpy(sb_mv, h->h_dg + off_mv, (size_t)n * 4);
    if (px_compares) *px_compares = r->px_compares;
    if (elapsed_ms) cudaEventElapsedTime(elapsed_ms, h->dg_e0, h->dg_e1);
    {
        std::lock_guard<std::mutex> lk2(h->mu);
        h->launches += 1;
    }
    return SVTME_OK;
}

int32_t svtme_picture_statistics(SvtMeHandle *h, SvtMePic *pic, uint32_t regions_per_width, uint32_t regions_per_height,
                                 uint32_t decim_step, SvtMePicStats *out) {
    if (!h || !pic || !out) return fail(h, SVTME_ERR_BAD_PARAMETER, "null argument");
    if (regions_per_width < 1 || regions_per_width > 4 || regions_per_height < 1 || regions_per_height > 4 || decim_step < 1 ||
        (pic->width >> 2) < regions_per_width || (pic->height >> 2) < regions_per_height)
        return fail(h, SVTME_ERR_BAD_PARAMETER, "regions / decim_step");
    int rc = set_device(h);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(h->dg_mu);
    const size_t total = al4k(sizeof(svtme::StatAcc));
    if (!h->dg_stream) {
        CK(cudaStreamCreateWithFlags(&h->dg_stream, cudaStreamNonBlocking));
        CK(cudaEventCreate(&h->dg_e0));
        CK(cudaEventCreate(&h->dg_e1));
        CK(cudaEventCreateWithFlags(&h->dg_done, cudaEventDisableTiming));
    }
    if (h->dg_cap < total) {
        if (h->d_dg) cudaFree(h->d_dg);
        if (h->h_dg) cudaFreeHost(h->h_dg);
        h->d_dg = h->h_dg = nullptr; h->dg_cap = 0;
        CK(cudaMalloc(&h->d_dg, total));
        CK(cudaMallocHost(&h->h_dg, total));
        h->dg_cap = total;
    }
    cudaStream_t st = h->dg_stream;
    CK(cudaStreamWaitEvent(st, pic->ready, 0));
    CK(cudaMemsetAsync(h->d_dg, 0, sizeof(svtme::StatAcc), st));
    CK(svtme::launch_hist16(pic->dev.lvl[2], (int)regions_per_width, (int)regions_per_height, (int)decim_step,
                            reinterpret_cast<svtme::StatAcc *>(h->d_dg), st));
    CK(cudaMemcpyAsync(h->h_dg, h->d_dg, sizeof(svtme::StatAcc), cudaMemcpyDeviceToHost, st));
    CK(cudaEventRecord(h->dg_done, st));
    { std::lock_guard<std::mutex> lk2(h->mu); pic->dg_reader = true; }
    CK(cudaStreamSynchronize(st));
    // the reference's scaling and rounding (pic_analysis_process.c:1480, 1504-1521)
    const svtme::StatAcc *a = reinterpret_cast<const svtme::StatAcc *>(h->h_dg);
    const uint32_t pw = pic->width >> 2, ph = pic->height >> 2, region_w = pw / regions_per_width, region_h = ph / regions_per_height;
    const uint32_t scale = 4 * 4 * decim_step * decim_step;
    memset(out, 0, sizeof(*out));
    uint64_t total_sum = 0;
    for (uint32_t wi = 0; wi < regions_per_width; wi++)
        for (uint32_t hi = 0; hi < regions_per_height; hi++) {
            const uint32_t w = region_w + (wi == regions_per_width - 1 ? pw - regions_per_width * region_w : 0);
            const uint32_t hh = region_h + (hi == regions_per_height - 1 ? ph - regions_per_height * region_h : 0);
            for (int b = 0; b < 256; b++) out->picture_histogram[wi][hi][b] = (1u + a->hist[wi * 4 + hi][b]) * scale;
            const uint64_t sum = a->sum[wi * 4 + hi] * (uint64_t)decim_step * decim_step;
            out->average_intensity_per_region[wi][hi] = (uint8_t)((sum + ((w * hh) >> 1)) / (w * hh));
            total_sum += sum;
        }
    out->avg_luma = total_sum / ((uint64_t)pw * ph);
    {
        std::lock_guard<std::mutex> lk2(h->mu);
        h->launches += 1;
    }
    return SVTME_OK;
}

const char *svtme_last_error(SvtMeHandle *h) { return h ? t_err.c_str() : "null handle"; }

// ---- RTCD-level entry points: host pointers, synchronous.  Inputs are copied into an aligned device scratch. ----
static uint8_t *scratch(SvtMeHandle *h, size_t bytes) {
    if (h->scratch_bytes < bytes) {
        if (h->d_scratch) cudaFree(h->d_scratch);
        h->d_scratch = nullptr; h->scratch_bytes = 0;
        if (cudaMalloc(&h->d_scratch, bytes) != cudaSuccess) return nullptr;
        h->scratch_bytes = bytes;
    }
    return h->d_scratch;
}
static inline size_t al(size_t v, size_t a) { return (v + a - 1) / a * a; }

void svtme_cuda_sad_loop_kernel(SvtMeHandle *h, const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                                uint32_t block_height, uint32_t block_width, uint64_t *best_sad, int16_t *x_search_center,
                                int16_t *y_search_center, uint32_t src_stride_raw, uint8_t skip_search_line, int16_t search_area_width,
                                int16_t search_area_height) {
    *best_sad = 0xffffff;  // compute_sad_c.c:71
    if (!h || block_width == 0 || block_height == 0 || block_width > 64 || block_height > 64 || search_area_width <= 0 || search_area_height <= 0) return;
    if (src_stride_raw == 0 || ref_stride % src_stride_raw) { t_err = "ref_stride must be a multiple of src_stride_raw"; return; }
    std::lock_guard<std::mutex> lk(h->mu);
    cudaSetDevice(h->cfg.device);
    // reference rows touched: (sa_h - 1) * raw + (bh - 1) * ref_stride + sa_w + bw - 1 bytes from `ref`
    const size_t ref_rows  = (size_t)(search_area_height - 1) + (size_t)(block_height - 1) * (ref_stride / src_stride_raw) + 1;
    const size_t ref_w     = (size_t)search_area_width + block_width - 1;
    const size_t d_rstride = al(ref_w + 64, 128), d_sstride = 128;
    const size_t need      = d_rstride * (ref_rows + 1) + d_sstride * block_height + 512;
    uint8_t     *d         = scratch(h, need);
    if (!d) { t_err = "scratch alloc"; return; }
    uint8_t *d_ref = d, *d_src = d + al(d_rstride * (ref_rows + 1), 256);
    cudaMemsetAsync(d, 0, need, 0);
    cudaMemcpy2DAsync(d_ref, d_rstride, ref, src_stride_raw, ref_w, ref_rows, cudaMemcpyHostToDevice, 0);
    cudaMemcpy2DAsync(d_src, d_sstride, src, src_stride, block_width, block_height, cudaMemcpyHostToDevice, 0);
    SadJob j;
    memset(&j, 0, sizeof(j));
    j.blk = d_src; j.blk_stride = (int32_t)d_sstride;
    j.ref = d_ref; j.raw_stride = (int32_t)d_rstride; j.ref_stride = (int32_t)(d_rstride * (ref_stride / src_stride_raw));
    j.rows = (int16_t)block_height; j.block_w = (int16_t)block_width;
    j.sa_w = search_area_width; j.sa_h = search_area_height; j.skip_line = skip_search_line; j.mv_scale = 1;
    j.best_sad = 0xffffff;
    uint32_t res[4] = {0, 0, 0, 0};
    cudaError_t e = launch_rtcd_sad_loop(j, h->d_small, 0);
    if (e == cudaSuccess) e = cudaMemcpy(res, h->d_small, sizeof(res), cudaMemcpyDeviceToHost);
    h->launches++;
    if (e != cudaSuccess) { t_err = cudaGetErrorString(e); return; }
    if (res[0]) { *best_sad = res[1]; *x_search_center = (int16_t)(int32_t)res[2]; *y_search_center = (int16_t)(int32_t)res[3]; }
}

void svtme_cuda_downsample_2d(SvtMeHandle *h, const uint8_t *input_samples, uint32_t input_stride, uint32_t input_area_width,
                              uint32_t input_area_height, uint8_t *decim_samples, uint32_t decim_stride, uint32_t decim_step) {
    if (!h || decim_step < 2 || !input_area_width || !input_area_height) return;
    std::lock_guard<std::mutex> lk(h->mu);
    cudaSetDevice(h->cfg.device);
    const size_t in_stride = al(input_area_width, 128);
    const uint32_t ow = (input_area_width - (decim_step >> 1) + decim_step - 1) / decim_step;
    const uint32_t oh = (input_area_height - (decim_step >> 1) + decim_step - 1) / decim_step;
    const size_t out_stride = al(ow, 128);
    uint8_t *d = scratch(h, in_stride * input_area_height + out_stride * oh + 512);
    if (!d) { t_err = "scratch alloc"; return; }
    uint8_t *d_in = d, *d_out = d + al(in_stride * input_area_height, 256);
    cudaMemcpy2DAsync(d_in, in_stride, input_samples, input_stride, input_area_width, input_area_height, cudaMemcpyHostToDevice, 0);
    cudaError_t e = launch_downsample_generic(d_in, (int)in_stride, (int)input_area_width, (int)input_area_height, d_out, (int)out_stride,
                                              (int)decim_step, 0);
    if (e == cudaSuccess) e = cudaMemcpy2D(decim_samples, decim_stride, d_out, out_stride, ow, oh, cudaMemcpyDeviceToHost);
    h->launches++;
    if (e != cudaSuccess) t_err = cudaGetErrorString(e);
}

uint32_t svtme_cuda_nxm_sad_kernel(SvtMeHandle *h, const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                                   uint32_t height, uint32_t width) {
    if (!h || !width || !height) return 0;
    std::lock_guard<std::mutex> lk(h->mu);
    cudaSetDevice(h->cfg.device);
    const size_t st = al(width, 128);
    uint8_t *d = scratch(h, 2 * st * height + 512);
    if (!d) { t_err = "scratch alloc"; return 0; }
    uint8_t *d_src = d, *d_ref = d + al(st * height, 256);
    cudaMemcpy2DAsync(d_src, st, src, src_stride, width, height, cudaMemcpyHostToDevice, 0);
    cudaMemcpy2DAsync(d_ref, st, ref, ref_stride, width, height, cudaMemcpyHostToDevice, 0);
    NxmJob j;
    memset(&j, 0, sizeof(j));
    j.src = d_src; j.ref = d_ref; j.src_stride = (int32_t)st; j.ref_stride = (int32_t)st; j.rows = (int16_t)height; j.width = (int16_t)width;
    uint32_t res = 0;
    cudaError_t e = launch_rtcd_nxm(j, h->d_small, 0);
    if (e == cudaSuccess) e = cudaMemcpy(&res, h->d_small, 4, cudaMemcpyDeviceToHost);
    h->launches++;
    if (e != cudaSuccess) t_err = cudaGetErrorString(e);
    return res;
}

void svtme_cuda_fullpel_search_sblock(SvtMeHandle *h, const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                                      int16_t x_search_area_origin, int16_t y_search_area_origin, uint32_t search_area_width,
                                      uint32_t search_area_height, uint8_t sub_sad, uint32_t *best_sad, uint32_t *best_mv) {
    if (!h || !search_area_width || !search_area_height || search_area_width > 32767 || search_area_height > 32767) return;
    std::lock_guard<std::mutex> lk(h->mu);
    cudaSetDevice(h->cfg.device);
    const size_t ref_w = search_area_width + 63, ref_rows = search_area_height + 63;
    const size_t d_rstride = al(ref_w + 64, 128);
    const size_t need = d_rstride * (ref_rows + 1) + 128 * 64 + 2 * 85 * 4 + 1024;
    uint8_t *d = scratch(h, need);
    if (!d) { t_err = "scratch alloc"; return; }
    uint8_t  *d_ref = d, *d_src = d + al(d_rstride * (ref_rows + 1), 256);
    uint32_t *d_bs = (uint32_t *)(d_src + 128 * 64), *d_bm = d_bs + 88;
    cudaMemsetAsync(d, 0, need, 0);
    cudaMemcpy2DAsync(d_ref, d_rstride, ref, ref_stride, ref_w, ref_rows, cudaMemcpyHostToDevice, 0);
    cudaMemcpy2DAsync(d_src, 128, src, src_stride, 64, 64, cudaMemcpyHostToDevice, 0);
    cudaMemcpyAsync(d_bs, best_sad, 85 * 4, cudaMemcpyHostToDevice, 0);
    cudaMemcpyAsync(d_bm, best_mv, 85 * 4, cudaMemcpyHostToDevice, 0);
    FpJob j;
    memset(&j, 0, sizeof(j));
    j.ref = d_ref; j.ref_stride = (int32_t)d_rstride; j.xo = x_search_area_origin; j.yo = y_search_area_origin;
    j.w = (int16_t)search_area_width; j.h = (int16_t)search_area_height;
    cudaError_t e = launch_rtcd_fullpel(d_src, 128, j, sub_sad ? 1 : 0, d_bs, d_bm, 0);
    if (e == cudaSuccess) e = cudaMemcpy(best_sad, d_bs, 85 * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess) e = cudaMemcpy(best_mv, d_bm, 85 * 4, cudaMemcpyDeviceToHost);
    h->launches++;
    if (e != cudaSuccess) t_err = cudaGetErrorString(e);
}

}  // extern "C"

// ---- allocation-free entry point (no counterpart in the reference) ----
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it):
// svt_nxm_sad_kernel with its operand staging and its result word in a file-scope device array instead of the handle's
// cudaMalloc'ed scratch.  Every GPU that loads the library holds the array (kNoAllocBytes) until exit.  All calls share
// it and it keeps the last call's contents: the result word is cleared on the stream before every launch; calls that
// overlap on the GPU race (the entry serialises host callers with a mutex and synchronises before it returns).
namespace {

constexpr uint32_t kNoAllocMaxWidth = 64, kNoAllocMaxHeight = 128, kNoAllocPitch = 128;
constexpr size_t   kNoAllocResult = 256;  // bytes reserved in front of the operands (result word at offset 0)
constexpr size_t   kNoAllocPlane  = (size_t)kNoAllocPitch * kNoAllocMaxHeight;
constexpr size_t   kNoAllocBytes  = kNoAllocResult + 2 * kNoAllocPlane;  // 33 024 bytes
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

// one warp per group of rows; the partial sums meet in the (cleared) result word
__global__ void __launch_bounds__(128) noalloc_nxm_kernel(unsigned char *scratch, uint32_t rows, uint32_t width) {
    const unsigned char *src = scratch + kNoAllocResult, *ref = src + kNoAllocPlane;
    uint32_t            *out = reinterpret_cast<uint32_t *>(scratch);
    const uint32_t       nw = (width + 3) >> 2, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t             acc = 0;
    for (uint32_t i = warp * 32 + lane; i < rows * nw; i += 128) {
        const uint32_t r = i / nw, c = i - r * nw, nb = min(4u, width - 4 * c);
        uint32_t       a = *reinterpret_cast<const uint32_t *>(src + r * kNoAllocPitch + 4 * c);
        uint32_t       b = *reinterpret_cast<const uint32_t *>(ref + r * kNoAllocPitch + 4 * c);
        if (nb < 4) { const uint32_t m = (1u << (8 * nb)) - 1u; a &= m; b &= m; }  // bytes past the row hold an earlier call's data
        acc = __vsadu4(a, b) + acc;
    }
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0 && acc) atomicAdd(out, acc);
}

std::mutex     g_noalloc_mu;
unsigned char *g_noalloc_ptr[64] = {};  // device address of g_noalloc_scratch, per device ordinal

}  // namespace

extern "C" int32_t svtme_cuda_nxm_sad_kernel_noalloc(SvtMeHandle *h, const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                                                     uint32_t ref_stride, uint32_t height, uint32_t width, uint32_t *sad) {
    if (!h || !src || !ref || !sad) return fail(h, SVTME_ERR_BAD_PARAMETER, "null argument");
    if (width == 0 || height == 0 || width > kNoAllocMaxWidth || height > kNoAllocMaxHeight)
        return fail(h, SVTME_ERR_BAD_PARAMETER, "svtme_cuda_nxm_sad_kernel_noalloc: block larger than the static scratch (at most 64 wide x 128 rows)");
    if (src_stride < width || ref_stride < width) return fail(h, SVTME_ERR_BAD_PARAMETER, "stride smaller than the block width");
    if (h->cfg.device < 0 || h->cfg.device >= 64) return fail(h, SVTME_ERR_BAD_PARAMETER, "device ordinal beyond the per-device table");
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    CK(cudaSetDevice(h->cfg.device));
    unsigned char *&d = g_noalloc_ptr[h->cfg.device];
    if (!d) {  // once per device, with that device current: the array's name is only a host-side stand-in
        void *p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        d = static_cast<unsigned char *>(p);
    }
    CK(cudaMemsetAsync(d, 0, kNoAllocResult, 0));  // the result word must start at zero
    CK(cudaMemcpy2DAsync(d + kNoAllocResult, kNoAllocPitch, src, src_stride, width, height, cudaMemcpyHostToDevice, 0));
    CK(cudaMemcpy2DAsync(d + kNoAllocResult + kNoAllocPlane, kNoAllocPitch, ref, ref_stride, width, height, cudaMemcpyHostToDevice, 0));
    noalloc_nxm_kernel<<<1, 128, 0, 0>>>(d, height, width);
    CK(cudaGetLastError());
    CK(cudaMemcpy(sad, d, sizeof(uint32_t), cudaMemcpyDeviceToHost));  // synchronous: the call has left the GPU when it returns
    {
        std::lock_guard<std::mutex> lk2(h->mu);
        h->launches += 1;
    }
    return SVTME_OK;
}
