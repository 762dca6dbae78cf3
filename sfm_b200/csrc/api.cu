// C ABI of libsfm_match.so (see include/sfm_match.h): context, metadata
// staging and kernel sequencing.  No torch types; device memory that crosses
// the boundary is owned by the caller.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "kernels.h"

namespace {

std::string g_create_error;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        const size_t want = n + n / 4 + 64;
        cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&p), want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

// Host -> device copies of per-call metadata (pair lists, offsets).  cudaMemcpyAsync from pageable
// memory waits for the stream's earlier work before it returns, which serialises the host with the
// previous kernel (1 ms of launch gaps per step on a slow host).  A small ring of pinned staging
// slots makes the copies truly asynchronous; a slot is reused only after its copy has executed.
struct PinnedRing {
    static constexpr int kSlots = 32;
    struct Slot {
        void* p = nullptr;
        size_t cap = 0;
        cudaEvent_t ev = nullptr;
        bool busy = false;
    } slot[kSlots];
    int next = 0;
    cudaError_t copy(void* dst, const void* src, size_t bytes, cudaStream_t st) {
        if (bytes == 0) return cudaSuccess;
        Slot& s = slot[next];
        next = (next + 1) % kSlots;
        cudaError_t e;
        if (!s.ev && (e = cudaEventCreateWithFlags(&s.ev, cudaEventDisableTiming)) != cudaSuccess) return e;
        if (s.busy && (e = cudaEventSynchronize(s.ev)) != cudaSuccess) return e;
        s.busy = false;
        if (s.cap < bytes) {
            if (s.p) cudaFreeHost(s.p);
            s.p = nullptr;
            s.cap = 0;
            const size_t want = bytes + bytes / 2 + 4096;
            if ((e = cudaMallocHost(&s.p, want)) != cudaSuccess) return e;
            s.cap = want;
        }
        std::memcpy(s.p, src, bytes);
        if ((e = cudaMemcpyAsync(dst, s.p, bytes, cudaMemcpyHostToDevice, st)) != cudaSuccess) return e;
        if ((e = cudaEventRecord(s.ev, st)) != cudaSuccess) return e;
        s.busy = true;
        return cudaSuccess;
    }
    void release() {
        for (Slot& s : slot) {
            if (s.busy) cudaEventSynchronize(s.ev);
            if (s.p) cudaFreeHost(s.p);
            if (s.ev) cudaEventDestroy(s.ev);
            s = Slot{};
        }
    }
};

}  // namespace

struct sfm_ctx {
    int device = 0;
    int num_sms = 0;
    std::string err;
    int64_t launches = 0;
    EncodeTiledFn encode = nullptr;

    // registered image set
    const uint8_t* d_desc = nullptr;
    const float* d_kps = nullptr;
    int32_t n_images = 0;
    int64_t total_rows = 0;
    int64_t pad_total = 0;
    std::vector<int64_t> h_off, h_pad;
    CUtensorMap tmap;
    CUtensorMap tmap_x32, tmap_ax32;   // wide kNN variant: fold operands as 32-byte rows (knn_wide.cu)
    CUtensorMap tmap_b64, tmap_x64;    // pair kNN variant: 64-row boxes of the descriptor / fold operand rows (knn_pair.cu)
    DevBuf<uint8_t> d_bias32;
    DevBuf<int32_t> d_fix_list;  // (work item, row) of the rows the wide kNN kernel leaves to its fix-up pass
    DevBuf<int32_t> d_tie_list;  // ping-pong kNN kernel: rows with a dropped unit to merge in afterwards (8 ints each)
    static constexpr int32_t kFixCap = 1 << 18;
    bool have_tmap = false;
    DevBuf<int64_t> d_img_off, d_pad_off;
    DevBuf<int32_t> d_norms, d_norms_pad, d_flags;
    DevBuf<uint32_t> d_par_bits;   // norm parities of the padded column space, 32 per word (knn_ts.cu)
    // sfm_mutual_filter scratch: gathered train descriptors / norms of the surviving matches, their
    // reverse kNN rows, the per-pair match offsets and the reverse pass's work items
    DevBuf<uint8_t> d_qdesc;
    DevBuf<int32_t> d_qnorms, d_knn_rev;
    DevBuf<int64_t> d_moff;
    DevBuf<sfm::WorkItem> d_work_rev;
    std::vector<int64_t> h_moff;
    std::vector<sfm::WorkItem> h_work_rev;
    int fold_flag_host = -1;     // flags[0] read back once per image set (-1: not yet)
    int knn_mode = 0;            // sfm_debug_set_knn_mode
    int knn_auto = 5;            // what mode 0 resolves to: 5 = wide folded kernel + generic fallback
    sfm::KnnTuning knn_tuning;   // sfm_debug_set_knn_waits

    // per-call metadata scratch
    DevBuf<int32_t> d_pairs;
    DevBuf<int64_t> d_meta;      // blk_prefix | row_off | row_off_rev
    DevBuf<sfm::WorkItem> d_work;
    std::vector<int64_t> h_meta;
    PinnedRing ring;             // staging of host metadata

    // sfm_radius_match scratch
    DevBuf<sfm::RadiusBlock> d_rblocks;
    DevBuf<long long> d_rcnt, d_roff;
    DevBuf<int32_t> d_rfirst;
    std::vector<sfm::RadiusBlock> h_rblocks;

    int fail(const char* fmt, ...) {
        char buf[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof(buf), fmt, ap);
        va_end(ap);
        err = buf;
        return 1;
    }
};

#define SFM_CUDA(ctx, call)                                                                      \
    do {                                                                                         \
        cudaError_t e__ = (call);                                                                \
        if (e__ != cudaSuccess)                                                                  \
            return (ctx)->fail("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, \
                               __LINE__);                                                        \
    } while (0)

namespace {

int check_pairs(sfm_ctx* ctx, const int32_t* h_pairs, int32_t n_pairs) {
    if (n_pairs < 0) return ctx->fail("n_pairs < 0");
    if (n_pairs > 0 && !h_pairs) return ctx->fail("h_pairs is NULL");
    for (int32_t p = 0; p < n_pairs; ++p) {
        const int32_t a = h_pairs[2 * p], b = h_pairs[2 * p + 1];
        if (a < 0 || b < 0 || a >= ctx->n_images || b >= ctx->n_images)
            return ctx->fail("pair %d = (%d, %d) out of range for %d images", p, a, b, ctx->n_images);
    }
    return 0;
}

// Copy the pair list and up to three i64 arrays of n_pairs+1 entries to the device.
int stage_meta(sfm_ctx* ctx, const int32_t* h_pairs, int32_t n_pairs, const int64_t* a,
               const int64_t* b, const int64_t* c, cudaStream_t st, const int64_t** da,
               const int64_t** db, const int64_t** dc) {
    const size_t n1 = static_cast<size_t>(n_pairs) + 1;
    SFM_CUDA(ctx, ctx->d_pairs.reserve(static_cast<size_t>(n_pairs) * 2 + 2));
    SFM_CUDA(ctx, ctx->d_meta.reserve(3 * n1));
    if (n_pairs > 0)
        SFM_CUDA(ctx, ctx->ring.copy(ctx->d_pairs.p, h_pairs, sizeof(int32_t) * 2 * n_pairs, st));
    const int64_t* src[3] = {a, b, c};
    const int64_t** dst[3] = {da, db, dc};
    for (int i = 0; i < 3; ++i) {
        if (!dst[i]) continue;
        *dst[i] = nullptr;
        if (!src[i]) continue;
        SFM_CUDA(ctx, ctx->ring.copy(ctx->d_meta.p + i * n1, src[i], sizeof(int64_t) * n1, st));
        *dst[i] = ctx->d_meta.p + i * n1;
    }
    return 0;
}

}  // namespace

extern "C" {

int sfm_abi_version(void) { return 1; }

void sfm_default_params(sfm_verify_params* p) {
    if (!p) return;
    p->ransac_thresh = 3.0;
    p->confidence = 0.99;
    p->max_iters = 1000;
    p->min_matches = 5;
    p->seed = 0x5F3759DF2024ull;
    p->verify_thresh = 3.0;
    p->min_inliers = 15;
    p->lo_iters = 2;
    p->min_ratio = 0.3;
    p->max_error = 2.0;
    p->min_spread = 20.0;
    p->score_mode = 3;
    p->min_work = 1 << 16;
}

const char* sfm_last_error(sfm_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int sfm_create(int device_ordinal, sfm_ctx** out) {
    if (!out) return 1;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e);
        return 1;
    }
    if (device_ordinal < 0 || device_ordinal >= n) {
        g_create_error = "device ordinal out of range";
        return 1;
    }
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device_ordinal);
    if (e != cudaSuccess) {
        g_create_error = cudaGetErrorString(e);
        return 1;
    }
    if (prop.major != 10) {
        char buf[160];
        snprintf(buf, sizeof(buf), "device %d is sm_%d%d; this library is sm_100a only (tcgen05/TMEM)",
                 device_ordinal, prop.major, prop.minor);
        g_create_error = buf;
        return 1;
    }
    e = cudaSetDevice(device_ordinal);
    if (e != cudaSuccess) {
        g_create_error = cudaGetErrorString(e);
        return 1;
    }
    sfm_ctx* ctx = new sfm_ctx();
    ctx->device = device_ordinal;
    ctx->num_sms = prop.multiProcessorCount;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
        g_create_error = "cuTensorMapEncodeTiled not available from the driver";
        delete ctx;
        return 1;
    }
    ctx->encode = reinterpret_cast<EncodeTiledFn>(fn);
    *out = ctx;
    return 0;
}

void sfm_destroy(sfm_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    ctx->d_img_off.release();
    ctx->d_pad_off.release();
    ctx->d_norms.release();
    ctx->d_norms_pad.release();
    ctx->d_flags.release();
    ctx->d_bias32.release();
    ctx->d_qdesc.release();
    ctx->d_qnorms.release();
    ctx->d_knn_rev.release();
    ctx->d_moff.release();
    ctx->d_work_rev.release();
    ctx->d_par_bits.release();
    ctx->d_fix_list.release();
    ctx->d_tie_list.release();
    ctx->d_pairs.release();
    ctx->d_meta.release();
    ctx->d_work.release();
    ctx->d_rblocks.release();
    ctx->d_rcnt.release();
    ctx->d_roff.release();
    ctx->d_rfirst.release();
    ctx->ring.release();
    delete ctx;
}

int64_t sfm_launch_count(sfm_ctx* ctx) { return ctx ? ctx->launches : 0; }

int sfm_set_images(sfm_ctx* ctx, const uint8_t* d_desc, const float* d_kps,
                   const int64_t* h_offsets, int32_t n_images, void* stream) {
    if (!ctx) return 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (n_images < 0 || !h_offsets) return ctx->fail("bad image set arguments");
    if (h_offsets[0] != 0) return ctx->fail("h_offsets[0] must be 0");
    for (int i = 0; i < n_images; ++i)
        if (h_offsets[i + 1] < h_offsets[i]) return ctx->fail("h_offsets must be non-decreasing");
    const int64_t total = h_offsets[n_images];
    if (total >= (int64_t(1) << 31) - 512) return ctx->fail("too many descriptor rows");
    if (total > 0 && !d_desc) return ctx->fail("d_desc is NULL");
    if (reinterpret_cast<uintptr_t>(d_desc) & 15) return ctx->fail("d_desc must be 16-byte aligned");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));

    ctx->d_desc = d_desc;
    ctx->d_kps = d_kps;
    ctx->fold_flag_host = -1;
    ctx->n_images = n_images;
    ctx->total_rows = total;
    ctx->h_off.assign(h_offsets, h_offsets + n_images + 1);
    ctx->h_pad.resize(n_images + 1);
    ctx->h_pad[0] = 0;
    for (int i = 0; i < n_images; ++i) {
        const int64_t n = h_offsets[i + 1] - h_offsets[i];
        ctx->h_pad[i + 1] = ctx->h_pad[i] + (n + SFM_TTILE - 1) / SFM_TTILE * SFM_TTILE;
    }
    ctx->pad_total = ctx->h_pad[n_images];
    if (ctx->pad_total >= (int64_t(1) << 31)) return ctx->fail("padded column space too large");
    SFM_CUDA(ctx, ctx->d_img_off.reserve(n_images + 1));
    SFM_CUDA(ctx, ctx->d_pad_off.reserve(n_images + 1));
    SFM_CUDA(ctx, ctx->d_norms.reserve(static_cast<size_t>(total) + 1));
    SFM_CUDA(ctx, ctx->d_norms_pad.reserve(static_cast<size_t>(ctx->pad_total) + 4));
    SFM_CUDA(ctx, ctx->d_flags.reserve(4));
    SFM_CUDA(ctx, ctx->d_bias32.reserve(static_cast<size_t>(total + SFM_TTILE) * 32));
    SFM_CUDA(ctx, ctx->d_fix_list.reserve(2 * static_cast<size_t>(sfm_ctx::kFixCap)));
    SFM_CUDA(ctx, ctx->d_par_bits.reserve(static_cast<size_t>(ctx->pad_total) / 32 + 4));
    SFM_CUDA(ctx, cudaMemcpyAsync(ctx->d_img_off.p, ctx->h_off.data(), sizeof(int64_t) * (n_images + 1),
                                  cudaMemcpyHostToDevice, st));
    SFM_CUDA(ctx, cudaMemcpyAsync(ctx->d_pad_off.p, ctx->h_pad.data(), sizeof(int64_t) * (n_images + 1),
                                  cudaMemcpyHostToDevice, st));

    ctx->have_tmap = false;
    if (total > 0) {
        const cuuint64_t gdim[2] = {SFM_DESC_DIM, static_cast<cuuint64_t>(total)};
        const cuuint64_t gstride[1] = {SFM_DESC_DIM};
        const cuuint32_t box[2] = {SFM_DESC_DIM, SFM_TTILE};
        const cuuint32_t estr[2] = {1, 1};
        CUresult r = ctx->encode(&ctx->tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2,
                                 const_cast<uint8_t*>(d_desc), gdim, gstride, box, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return ctx->fail("cuTensorMapEncodeTiled failed: CUresult %d", int(r));
        // wide variant: u8 [128 + total, 32], 256-row boxes (train tile) and a 128-row box (constant rows)
        const cuuint64_t wdim[2] = {32, static_cast<cuuint64_t>(total + SFM_TTILE)};
        const cuuint64_t wstride[1] = {32};
        const cuuint32_t wbox[2] = {32, 256};
        const cuuint32_t abox[2] = {32, 128};
        r = ctx->encode(&ctx->tmap_x32, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, ctx->d_bias32.p, wdim, wstride, wbox,
                        estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_32B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r == CUDA_SUCCESS)
            r = ctx->encode(&ctx->tmap_ax32, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, ctx->d_bias32.p, wdim, wstride, abox,
                            estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_32B,
                            CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return ctx->fail("cuTensorMapEncodeTiled (bias32) failed: CUresult %d", int(r));
        const cuuint32_t hbox[2] = {SFM_DESC_DIM, 64};
        const cuuint32_t hxbox[2] = {32, 64};
        r = ctx->encode(&ctx->tmap_b64, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<uint8_t*>(d_desc), gdim, gstride,
                        hbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r == CUDA_SUCCESS)
            r = ctx->encode(&ctx->tmap_x64, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, ctx->d_bias32.p, wdim, wstride, hxbox,
                            estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_32B,
                            CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return ctx->fail("cuTensorMapEncodeTiled (64-row boxes) failed: CUresult %d", int(r));
        ctx->have_tmap = true;
    }
    sfm::launch_norms(d_desc, total, ctx->d_img_off.p, ctx->d_pad_off.p, n_images, ctx->d_norms.p,
                      ctx->d_norms_pad.p, ctx->d_flags.p, st, &ctx->launches);
    sfm::launch_bias32(ctx->d_norms.p, total, ctx->d_bias32.p, ctx->d_flags.p, st, &ctx->launches);
    sfm::launch_parity_bits(ctx->d_norms_pad.p, ctx->pad_total, ctx->d_par_bits.p, st, &ctx->launches);
    SFM_CUDA(ctx, cudaGetLastError());
    return 0;
}

int sfm_knn2(sfm_ctx* ctx, const int32_t* h_pairs, int32_t n_pairs, const int64_t* h_row_off,
             int32_t* d_knn, void* stream) {
    if (!ctx) return 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (check_pairs(ctx, h_pairs, n_pairs)) return 1;
    if (n_pairs == 0) return 0;
    if (!h_row_off) return ctx->fail("NULL argument");
    // an image set whose images are all empty has no rows and no tensor map: every pair is then a
    // "too few features" skip of the reference (utils/find_matches.py:268), not an error
    if (ctx->n_images > 0 && ctx->total_rows == 0) return 0;
    if (!ctx->have_tmap) return ctx->fail("no descriptors registered (sfm_set_images)");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    // work-item prefix: ceil(n_query / 256) blocks per pair
    ctx->h_meta.resize(static_cast<size_t>(n_pairs) + 1);
    ctx->h_meta[0] = 0;
    for (int32_t p = 0; p < n_pairs; ++p) {
        const int32_t qi = h_pairs[2 * p];
        const int64_t nq = ctx->h_off[qi + 1] - ctx->h_off[qi];
        if (h_row_off[p + 1] - h_row_off[p] < nq)
            return ctx->fail("h_row_off leaves pair %d fewer than its %lld query rows", p, (long long)nq);
        ctx->h_meta[p + 1] = ctx->h_meta[p] + (nq + SFM_QBLOCK - 1) / SFM_QBLOCK;
    }
    const int64_t n_work = ctx->h_meta[n_pairs];
    if (n_work == 0) return 0;   // every query image of the list is empty
    if (!d_knn) return ctx->fail("NULL argument");
    const int64_t *d_prefix, *d_row;
    if (stage_meta(ctx, h_pairs, n_pairs, ctx->h_meta.data(), h_row_off, nullptr, st, &d_prefix, &d_row,
                   nullptr))
        return 1;
    SFM_CUDA(ctx, ctx->d_work.reserve(static_cast<size_t>(n_work) + 1));
    sfm::launch_build_work(ctx->d_pairs.p, n_pairs, d_prefix, d_row, ctx->d_img_off.p, ctx->d_pad_off.p,
                           ctx->d_work.p, st, &ctx->launches);
    // knn_mode 0: the wide folded kernel (knn_wide.cu), or the generic one when a norm is beyond the fold range
    // (both are launched; the one that does not apply exits at its first instruction).  1: generic only,
    // 5: wide only, 6/7: its timing probes, 12: CTA-pair kernel (knn_pair.cu), 13/14: its timing probes,
    // 15: wide with the warp-uniform skip of the unit bookkeeping, 16: wide with the lean tracker (A/B variants
    // of the default tracker), 17: ping-pong kernel (accumulators alternate over train tiles), 18 / 19: its
    // timing probes.
    const int km = ctx->knn_mode == 0 ? ctx->knn_auto : ctx->knn_mode;
    int r = 0;
    const int32_t fix_cap = (ctx->knn_tuning.probe_flags & 8u) ? 4 : sfm_ctx::kFixCap;
    // tie list of the wide and ping-pong kernels (rows with a dropped unit to merge in afterwards): one entry per 8
    // query rows of the launch (1 % duplicated rows need about one per 25); a full list sends rows to the fix-up pass
    int32_t tie_cap = 0;
    if (km >= 5 && !(km >= 12 && km <= 14)) {
        tie_cap = static_cast<int32_t>(std::min<int64_t>(n_work * (SFM_QBLOCK / 8) + 4096, int64_t(1) << 27));
        if (ctx->knn_tuning.probe_flags & 512u) tie_cap = 2;   // test switch: the overflow path (rows go to the fix-up pass)
        SFM_CUDA(ctx, ctx->d_tie_list.reserve(8 * static_cast<size_t>(tie_cap)));
    }
    if (km >= 15)   // tracker variants of the wide kernel (A/B against mode 5), ping-pong kernel and its probes
        r = sfm::launch_knn2_wide(ctx->tmap, ctx->tmap_x32, ctx->tmap_ax32, ctx->d_desc, ctx->d_work.p, n_work, ctx->d_norms.p,
                                  ctx->d_norms_pad.p, ctx->d_flags.p, km - 12, ctx->knn_mode != 0, ctx->knn_tuning,
                                  ctx->d_fix_list.p, fix_cap, d_knn, ctx->num_sms, st, &ctx->launches, nullptr,
                                  ctx->d_tie_list.p, tie_cap);
    else if (km >= 12)
        r = sfm::launch_knn2_pair(ctx->tmap, ctx->tmap_b64, ctx->tmap_x64, ctx->tmap_ax32, ctx->d_desc, ctx->d_work.p,
                                  n_work, ctx->d_norms.p, ctx->d_norms_pad.p, ctx->d_par_bits.p, ctx->d_flags.p,
                                  km - 12, ctx->knn_mode != 0, ctx->knn_tuning, ctx->d_fix_list.p, fix_cap, d_knn,
                                  ctx->num_sms, st, &ctx->launches);
    else if (km >= 5)
        r = sfm::launch_knn2_wide(ctx->tmap, ctx->tmap_x32, ctx->tmap_ax32, ctx->d_desc, ctx->d_work.p, n_work, ctx->d_norms.p,
                                  ctx->d_norms_pad.p, ctx->d_flags.p, km >= 6 ? km - 5 : 0, ctx->knn_mode != 0,
                                  ctx->knn_tuning, ctx->d_fix_list.p, fix_cap, d_knn,
                                  ctx->num_sms, st, &ctx->launches, nullptr, ctx->d_tie_list.p, tie_cap);
    const bool folded_auto = (km == 5 || km == 12) && ctx->knn_mode == 0;
    if (!r && (km == 1 || folded_auto))
        r = sfm::launch_knn2(ctx->tmap, ctx->d_work.p, n_work, ctx->d_norms.p, ctx->d_norms_pad.p, ctx->d_flags.p,
                             folded_auto ? 1 : 0, ctx->knn_tuning, d_knn, ctx->num_sms, st, &ctx->launches);
    if (r) return ctx->fail("knn2 launch failed: %s", cudaGetErrorString(static_cast<cudaError_t>(r)));
    return 0;
}

int sfm_filter_matches(sfm_ctx* ctx, const int32_t* h_pairs, int32_t n_pairs,
                       const int64_t* h_row_off, const int32_t* d_knn, const int64_t* h_row_off_rev,
                       const int32_t* d_knn_rev, double ratio, int32_t* d_mq, int32_t* d_mt,
                       int32_t* d_md2, int32_t* d_count, void* stream) {
    if (!ctx) return 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (check_pairs(ctx, h_pairs, n_pairs)) return 1;
    if (n_pairs == 0) return 0;
    if (!h_row_off || !d_knn || !d_mq || !d_mt || !d_md2 || !d_count) return ctx->fail("NULL argument");
    if ((d_knn_rev != nullptr) != (h_row_off_rev != nullptr))
        return ctx->fail("d_knn_rev and h_row_off_rev must be given together");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    const int64_t *d_row, *d_row_rev;
    if (stage_meta(ctx, h_pairs, n_pairs, nullptr, h_row_off, h_row_off_rev, st, nullptr, &d_row,
                   &d_row_rev))
        return 1;
    sfm::launch_filter(ctx->d_pairs.p, n_pairs, ctx->d_img_off.p, d_row,
                       reinterpret_cast<const int4*>(d_knn), d_row_rev,
                       reinterpret_cast<const int4*>(d_knn_rev), ratio, d_mq, d_mt, d_md2, d_count, st,
                       &ctx->launches);
    SFM_CUDA(ctx, cudaGetLastError());
    return 0;
}

int sfm_filter_matches_static(sfm_ctx* ctx, const int32_t* h_pairs, int32_t n_pairs, const int64_t* h_row_off,
                              const int32_t* d_knn, double ratio, int32_t* d_mq, int32_t* d_mt, int32_t* d_md2,
                              int32_t* d_count, void* stream) {
    if (!ctx) return 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (check_pairs(ctx, h_pairs, n_pairs)) return 1;
    if (n_pairs > SFM_FILTER_STATIC_MAX_PAIRS)
        return ctx->fail("n_pairs = %d exceeds SFM_FILTER_STATIC_MAX_PAIRS = %d (static scratch, no allocation): "
                         "split the batch or call sfm_filter_matches",
                         n_pairs, SFM_FILTER_STATIC_MAX_PAIRS);
    if (n_pairs > 0 && (!h_row_off || !d_knn || !d_mq || !d_mt || !d_md2 || !d_count))
        return ctx->fail("NULL argument");
    if (n_pairs > 0 && ctx->h_off.size() != static_cast<size_t>(ctx->n_images) + 1)
        return ctx->fail("no image set registered");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    SFM_CUDA(ctx, sfm::launch_filter_static(ctx->device, h_pairs, n_pairs, ctx->h_off.data(), h_row_off,
                                            reinterpret_cast<const int4*>(d_knn), ratio, d_mq, d_mt, d_md2,
                                            d_count, st, &ctx->launches));
    return 0;
}

int sfm_mutual_filter(sfm_ctx* ctx, const int32_t* h_pairs, int32_t n_pairs, const int64_t* h_row_off,
                      const int32_t* h_count, int32_t* d_mq, int32_t* d_mt, int32_t* d_md2, int32_t* d_count,
                      void* stream) {
    if (!ctx) return 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (check_pairs(ctx, h_pairs, n_pairs)) return 1;
    if (n_pairs == 0) return 0;
    if (!h_row_off || !h_count || !d_mq || !d_mt || !d_md2 || !d_count) return ctx->fail("NULL argument");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    // per-pair offsets of the surviving matches and the work items of the reverse pass: blocks of up to
    // 256 gathered train descriptors (queries) against every row of the pair's FIRST image
    ctx->h_moff.resize(static_cast<size_t>(n_pairs) + 1);
    ctx->h_work_rev.clear();
    ctx->h_moff[0] = 0;
    for (int32_t p = 0; p < n_pairs; ++p) {
        const int32_t qi = h_pairs[2 * p];
        const int64_t nq = ctx->h_off[qi + 1] - ctx->h_off[qi];
        const int32_t m = h_count[p];
        if (m < 0 || m > h_row_off[p + 1] - h_row_off[p])
            return ctx->fail("h_count[%d] = %d does not fit the pair's rows", p, m);
        ctx->h_moff[p + 1] = ctx->h_moff[p] + m;
        for (int32_t b = 0; b < m; b += SFM_QBLOCK) {
            sfm::WorkItem w;
            w.q_row = static_cast<int32_t>(ctx->h_moff[p] + b);
            w.q_cnt = m - b < SFM_QBLOCK ? m - b : SFM_QBLOCK;
            w.t_row = static_cast<int32_t>(ctx->h_off[qi]);
            w.t_cnt = static_cast<int32_t>(nq);
            w.t_pad = static_cast<int32_t>(ctx->h_pad[qi]);
            w.x_row = 0;
            w.out_row = ctx->h_moff[p] + b;
            ctx->h_work_rev.push_back(w);
        }
    }
    const int64_t total = ctx->h_moff[n_pairs];
    if (total == 0) return 0;
    if (total >= (int64_t(1) << 31) - 512) return ctx->fail("too many matches for one mutual pass");
    if (!ctx->have_tmap) return ctx->fail("no descriptors registered (sfm_set_images)");
    if (ctx->fold_flag_host < 0) {   // once per image set (the caller has just synchronised for h_count anyway)
        int32_t f = 0;
        SFM_CUDA(ctx, cudaMemcpyAsync(&f, ctx->d_flags.p, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        SFM_CUDA(ctx, cudaStreamSynchronize(st));
        ctx->fold_flag_host = f != 0;
    }
    if (ctx->fold_flag_host)
        return ctx->fail("a descriptor norm is beyond the fold range: use the reversed kNN pass for the mutual check");
    const int64_t q_rows = total + SFM_QBLOCK;   // a block's tensor-memory load may run past the last match
    SFM_CUDA(ctx, ctx->d_qdesc.reserve(static_cast<size_t>(q_rows) * SFM_DESC_DIM));
    SFM_CUDA(ctx, ctx->d_qnorms.reserve(static_cast<size_t>(q_rows)));
    SFM_CUDA(ctx, ctx->d_knn_rev.reserve(static_cast<size_t>(total) * 4));
    SFM_CUDA(ctx, ctx->d_moff.reserve(static_cast<size_t>(n_pairs) + 1));
    SFM_CUDA(ctx, ctx->d_work_rev.reserve(ctx->h_work_rev.size()));
    const int64_t* d_row;
    if (stage_meta(ctx, h_pairs, n_pairs, nullptr, h_row_off, nullptr, st, nullptr, &d_row, nullptr)) return 1;
    SFM_CUDA(ctx, ctx->ring.copy(ctx->d_moff.p, ctx->h_moff.data(), sizeof(int64_t) * (n_pairs + 1), st));
    SFM_CUDA(ctx, ctx->ring.copy(ctx->d_work_rev.p, ctx->h_work_rev.data(),
                                 sizeof(sfm::WorkItem) * ctx->h_work_rev.size(), st));
    CUtensorMap qmap;
    {
        const cuuint64_t gdim[2] = {SFM_DESC_DIM, static_cast<cuuint64_t>(q_rows)};
        const cuuint64_t gstride[1] = {SFM_DESC_DIM};
        const cuuint32_t box[2] = {SFM_DESC_DIM, SFM_TTILE};
        const cuuint32_t estr[2] = {1, 1};
        CUresult r = ctx->encode(&qmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, ctx->d_qdesc.p, gdim, gstride, box, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return ctx->fail("cuTensorMapEncodeTiled (mutual queries) failed: CUresult %d", int(r));
    }
    sfm::launch_mutual_gather(ctx->d_pairs.p, n_pairs, ctx->d_img_off.p, d_row, ctx->d_moff.p, d_mt, ctx->d_desc,
                              ctx->d_norms.p, ctx->d_qdesc.p, ctx->d_qnorms.p, st, &ctx->launches);
    const sfm::KnnQuery query{&qmap, ctx->d_qdesc.p, ctx->d_qnorms.p};
    const int r = sfm::launch_knn2_wide(ctx->tmap, ctx->tmap_x32, ctx->tmap_ax32, ctx->d_desc, ctx->d_work_rev.p,
                                        static_cast<int64_t>(ctx->h_work_rev.size()), ctx->d_norms.p,
                                        ctx->d_norms_pad.p, ctx->d_flags.p, 0, 0, ctx->knn_tuning, ctx->d_fix_list.p,
                                        sfm_ctx::kFixCap, ctx->d_knn_rev.p, ctx->num_sms, st, &ctx->launches, &query);
    if (r) return ctx->fail("mutual kNN launch failed: %s", cudaGetErrorString(static_cast<cudaError_t>(r)));
    sfm::launch_mutual_compact(n_pairs, d_row, ctx->d_moff.p, reinterpret_cast<const int4*>(ctx->d_knn_rev.p), d_mq,
                               d_mt, d_md2, d_count, st, &ctx->launches);
    SFM_CUDA(ctx, cudaGetLastError());
    return 0;
}

int sfm_verify_pairs(sfm_ctx* ctx, const int32_t* h_pairs, int32_t n_pairs, const int64_t* h_row_off,
                     const int32_t* d_mq, const int32_t* d_mt, const int32_t* d_count,
                     const sfm_verify_params* params, float* d_pts1, float* d_pts2, double* d_F,
                     uint8_t* d_mask, float* d_sym, sfm_pair_metrics* d_metrics, int32_t* d_status,
                     void* stream) {
    if (!ctx) return 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (check_pairs(ctx, h_pairs, n_pairs)) return 1;
    if (n_pairs == 0) return 0;
    if (!h_row_off || !d_mq || !d_mt || !d_count || !d_pts1 || !d_pts2 || !d_F || !d_mask || !d_sym ||
        !d_metrics || !d_status)
        return ctx->fail("NULL argument");
    if (!ctx->d_kps) return ctx->fail("no keypoints registered (sfm_set_images d_kps)");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    sfm_verify_params prm;
    if (params) prm = *params; else sfm_default_params(&prm);
    const int64_t* d_row;
    if (stage_meta(ctx, h_pairs, n_pairs, nullptr, h_row_off, nullptr, st, nullptr, &d_row, nullptr))
        return 1;
    sfm::launch_gather_points(ctx->d_pairs.p, n_pairs, ctx->d_img_off.p, d_row, d_mq, d_mt, d_count,
                              ctx->d_kps, d_pts1, d_pts2, st, &ctx->launches);
    int r = sfm::launch_ransac_verify(n_pairs, d_row, d_count, ctx->d_pairs.p, d_pts1, d_pts2, nullptr,
                                      prm, d_F, d_mask, d_sym, d_metrics, d_status, st, &ctx->launches);
    if (r) return ctx->fail("ransac launch failed: %s", cudaGetErrorString(static_cast<cudaError_t>(r)));
    return 0;
}

static int points_entry(sfm_ctx* ctx, int32_t n_pairs, const int64_t* h_off, const float* d_pts1,
                        const float* d_pts2, const double* d_F_in, const sfm_verify_params* params,
                        double* d_F, uint8_t* d_mask, float* d_sym, sfm_pair_metrics* d_metrics,
                        int32_t* d_status, cudaStream_t st) {
    if (n_pairs < 0) return ctx->fail("n_pairs < 0");
    if (n_pairs == 0) return 0;
    if (!h_off || !d_pts1 || !d_pts2 || !d_mask || !d_sym || !d_metrics || !d_status)
        return ctx->fail("NULL argument");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    sfm_verify_params prm;
    if (params) prm = *params; else sfm_default_params(&prm);
    const size_t n1 = static_cast<size_t>(n_pairs) + 1;
    SFM_CUDA(ctx, ctx->d_meta.reserve(3 * n1));
    SFM_CUDA(ctx, ctx->ring.copy(ctx->d_meta.p, h_off, sizeof(int64_t) * n1, st));
    int r = sfm::launch_ransac_verify(n_pairs, ctx->d_meta.p, nullptr, nullptr, d_pts1, d_pts2, d_F_in, prm,
                                      d_F, d_mask, d_sym, d_metrics, d_status, st, &ctx->launches);
    if (r) return ctx->fail("ransac launch failed: %s", cudaGetErrorString(static_cast<cudaError_t>(r)));
    return 0;
}

int sfm_verify_given_F(sfm_ctx* ctx, int32_t n_pairs, const int64_t* h_off, const float* d_pts1,
                       const float* d_pts2, const double* d_F, const sfm_verify_params* params,
                       uint8_t* d_mask, float* d_sym, sfm_pair_metrics* d_metrics, int32_t* d_status,
                       void* stream) {
    if (!ctx) return 1;
    if (!d_F && n_pairs > 0) return ctx->fail("d_F is NULL");
    return points_entry(ctx, n_pairs, h_off, d_pts1, d_pts2, d_F, params, nullptr, d_mask, d_sym,
                        d_metrics, d_status, static_cast<cudaStream_t>(stream));
}

int sfm_ransac_points(sfm_ctx* ctx, int32_t n_pairs, const int64_t* h_off, const float* d_pts1,
                      const float* d_pts2, const sfm_verify_params* params, double* d_F,
                      uint8_t* d_mask, float* d_sym, sfm_pair_metrics* d_metrics, int32_t* d_status,
                      void* stream) {
    if (!ctx) return 1;
    if (!d_F && n_pairs > 0) return ctx->fail("d_F is NULL");
    return points_entry(ctx, n_pairs, h_off, d_pts1, d_pts2, nullptr, params, d_F, d_mask, d_sym,
                        d_metrics, d_status, static_cast<cudaStream_t>(stream));
}

int sfm_pack_results(sfm_ctx* ctx, int32_t n_pairs, const int64_t* h_row_off, const int32_t* d_count,
                     const int32_t* d_status, int32_t only_ok, const int32_t* d_mq,
                     const int32_t* d_mt, const int32_t* d_md2, const uint8_t* d_mask,
                     const float* d_sym, const float* d_pts1, const float* d_pts2,
                     int64_t* d_pack_off, int32_t* o_mq, int32_t* o_mt, int32_t* o_md2,
                     uint8_t* o_mask, float* o_sym, float* o_pts1, float* o_pts2, void* stream) {
    if (!ctx) return 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (n_pairs < 0) return ctx->fail("n_pairs < 0");
    if (n_pairs == 0) return 0;
    if (!h_row_off || !d_count || !d_pack_off) return ctx->fail("NULL argument");
    if (only_ok && !d_status) return ctx->fail("only_ok needs d_status");
    if ((d_mq && !o_mq) || (d_mt && !o_mt) || (d_md2 && !o_md2) || (d_mask && !o_mask) ||
        (d_sym && !o_sym) || (d_pts1 && !o_pts1) || (d_pts2 && !o_pts2))
        return ctx->fail("a source array has no destination");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t n1 = static_cast<size_t>(n_pairs) + 1;
    SFM_CUDA(ctx, ctx->d_meta.reserve(3 * n1));
    SFM_CUDA(ctx, ctx->ring.copy(ctx->d_meta.p + n1, h_row_off, sizeof(int64_t) * n1, st));
    sfm::PackArrays a{d_mq, d_mt, d_md2, d_mask, d_sym, d_pts1, d_pts2,
                      o_mq, o_mt, o_md2, o_mask, o_sym, o_pts1, o_pts2};
    sfm::launch_pack(n_pairs, ctx->d_meta.p + n1, d_count, d_status, only_ok ? 1 : 0, d_pack_off, a, st,
                     &ctx->launches);
    SFM_CUDA(ctx, cudaGetLastError());
    return 0;
}

// Smallest s with sqrt_rn(s) >= thr: sqrt(s) < thr <=> s < radius_limit(thr) for every s >= 0
// (sqrt_rn is monotone; positive doubles order like their bit patterns).
static double radius_limit(double thr) {
    if (!(thr > 0.0)) return 0.0;                       // sqrt(s) < thr never holds
    if (std::isinf(thr)) return thr;
    uint64_t lo = 0, hi = 0x7ff0000000000000ull;        // sqrt(+0) = 0 < thr, sqrt(inf) = inf >= thr
    while (hi - lo > 1) {
        const uint64_t mid = lo + (hi - lo) / 2;
        double s;
        std::memcpy(&s, &mid, sizeof(s));
        if (std::sqrt(s) >= thr) hi = mid; else lo = mid;
    }
    double s;
    std::memcpy(&s, &hi, sizeof(s));
    return s;
}

int sfm_radius_match(sfm_ctx* ctx, int32_t n_seg, const int64_t* h_track_off, const double* d_tracks,
                     const int64_t* h_pt_off, const double* d_pts, double threshold, int64_t* h_seg_off,
                     int32_t* d_out_track, int32_t* d_out_pt, int64_t cap, void* stream) {
    if (!ctx) return 1;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (n_seg < 0) return ctx->fail("n_seg < 0");
    if (!h_seg_off) return ctx->fail("h_seg_off is NULL");
    if (n_seg == 0) {
        h_seg_off[0] = 0;
        return 0;
    }
    if (!h_track_off || !h_pt_off) return ctx->fail("NULL offset array");
    if ((d_out_track == nullptr) != (d_out_pt == nullptr)) return ctx->fail("pass both output arrays or neither");
    if (std::isnan(threshold)) return ctx->fail("threshold is NaN");
    auto& blocks = ctx->h_rblocks;
    blocks.clear();
    ctx->h_meta.assign(static_cast<size_t>(n_seg) + 1, 0);       // blocks before each segment
    for (int32_t s = 0; s < n_seg; ++s) {
        ctx->h_meta[s] = static_cast<int64_t>(blocks.size());
        const int64_t nt = h_track_off[s + 1] - h_track_off[s], np = h_pt_off[s + 1] - h_pt_off[s];
        if (nt < 0 || np < 0 || h_track_off[0] != 0 || h_pt_off[0] != 0)
            return ctx->fail("segment %d: offsets must start at 0 and not decrease", s);
        if (nt > INT32_MAX || np > INT32_MAX) return ctx->fail("segment %d too large", s);
        if (np == 0) continue;                           // counts stay 0
        for (int64_t r = 0; r < nt; r += sfm::kRadiusRows)
            blocks.push_back({h_track_off[s] + r, h_pt_off[s],
                              static_cast<int32_t>(std::min<int64_t>(sfm::kRadiusRows, nt - r)),
                              static_cast<int32_t>(np), static_cast<int32_t>(r), 0});
    }
    ctx->h_meta[n_seg] = static_cast<int64_t>(blocks.size());
    const int64_t rows = h_track_off[n_seg];
    if (rows > 0 && !d_tracks) return ctx->fail("d_tracks is NULL");
    if (h_pt_off[n_seg] > 0 && !d_pts) return ctx->fail("d_pts is NULL");
    if (blocks.size() > static_cast<size_t>(INT32_MAX)) return ctx->fail("too many track blocks");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    const int32_t nb = static_cast<int32_t>(blocks.size());
    SFM_CUDA(ctx, ctx->d_rblocks.reserve(blocks.size() + 1));
    SFM_CUDA(ctx, ctx->d_rcnt.reserve(static_cast<size_t>(rows) + 1));
    SFM_CUDA(ctx, ctx->d_roff.reserve(2 * (blocks.size() + 1)));          // block offsets | block totals
    SFM_CUDA(ctx, ctx->d_rfirst.reserve(static_cast<size_t>(rows) * sfm::kRadiusKeep + 1));
    if (nb > 0)
        SFM_CUDA(ctx, ctx->ring.copy(ctx->d_rblocks.p, blocks.data(), sizeof(sfm::RadiusBlock) * blocks.size(), st));
    const double s_lim = radius_limit(threshold);
    long long* d_block_off = ctx->d_roff.p;
    long long* d_block_cnt = ctx->d_roff.p + nb + 1;
    sfm::launch_radius_count(ctx->d_rblocks.p, nb, d_tracks, d_pts, s_lim, ctx->d_rcnt.p, ctx->d_rfirst.p,
                             d_block_cnt, d_block_off, st, &ctx->launches);
    SFM_CUDA(ctx, cudaGetLastError());
    // match offsets at the segment boundaries (a segment starts at a block boundary): gathered on
    // the device, n_seg + 1 words come back
    const size_t n1 = static_cast<size_t>(n_seg) + 1;
    SFM_CUDA(ctx, ctx->d_meta.reserve(3 * n1));
    static_assert(sizeof(long long) == sizeof(int64_t), "offset width");
    SFM_CUDA(ctx, ctx->ring.copy(ctx->d_meta.p, ctx->h_meta.data(), sizeof(int64_t) * n1, st));
    sfm::launch_gather_offsets(d_block_off, reinterpret_cast<const long long*>(ctx->d_meta.p),
                               static_cast<int32_t>(n1), reinterpret_cast<long long*>(ctx->d_meta.p + n1), st,
                               &ctx->launches);
    SFM_CUDA(ctx, cudaGetLastError());
    SFM_CUDA(ctx, cudaMemcpyAsync(h_seg_off, ctx->d_meta.p + n1, sizeof(int64_t) * n1, cudaMemcpyDeviceToHost, st));
    SFM_CUDA(ctx, cudaStreamSynchronize(st));
    if (!d_out_track) return 0;
    const int64_t total = h_seg_off[n_seg];
    if (total > cap) return 0;   // caller sees h_seg_off[n_seg] > cap and calls again with room
    // the fill pass writes global positions; segment-local positions follow from h_seg_off
    sfm::launch_radius_fill(ctx->d_rblocks.p, nb, d_tracks, d_pts, s_lim, ctx->d_rcnt.p, ctx->d_rfirst.p,
                            d_block_off, cap, d_out_track, d_out_pt, st, &ctx->launches);
    SFM_CUDA(ctx, cudaGetLastError());
    SFM_CUDA(ctx, cudaStreamSynchronize(st));
    return 0;
}

int sfm_probe_int8_peak(sfm_ctx* ctx, int32_t iters, float* ms, double* ops, void* stream) {
    return sfm_probe_int8_rate(ctx, iters, 256, ms, ops, stream);
}

int sfm_probe_int8_rate(sfm_ctx* ctx, int32_t iters, int32_t tile_n, float* ms, double* ops, void* stream) {
    if (!ctx) return 1;
    if (iters <= 0 || !ms || !ops) return ctx->fail("bad probe arguments");
    if (tile_n != 64 && tile_n != 128 && tile_n != 256) return ctx->fail("tile_n must be 64, 128 or 256");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaEvent_t e0, e1;
    SFM_CUDA(ctx, cudaEventCreate(&e0));
    SFM_CUDA(ctx, cudaEventCreate(&e1));
    int r = sfm::launch_mma_probe(iters, tile_n, ctx->num_sms, st, &ctx->launches);   // warm-up
    if (!r) {
        cudaEventRecord(e0, st);
        r = sfm::launch_mma_probe(iters, tile_n, ctx->num_sms, st, &ctx->launches);
        cudaEventRecord(e1, st);
    }
    cudaError_t e = cudaEventSynchronize(e1);
    if (r || e != cudaSuccess) {
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        return ctx->fail("probe failed: %s", cudaGetErrorString(r ? static_cast<cudaError_t>(r) : e));
    }
    cudaEventElapsedTime(ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *ops = 2.0 * 128.0 * tile_n * 128.0 * static_cast<double>(iters) * ctx->num_sms;
    return 0;
}

int sfm_probe_fp64_rate(sfm_ctx* ctx, int32_t iters, float* ms, double* instr, void* stream) {
    if (!ctx) return 1;
    if (iters <= 0 || !ms || !instr) return ctx->fail("bad probe arguments");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    SFM_CUDA(ctx, ctx->d_meta.reserve(8));
    cudaEvent_t e0, e1;
    SFM_CUDA(ctx, cudaEventCreate(&e0));
    SFM_CUDA(ctx, cudaEventCreate(&e1));
    double* scratch = reinterpret_cast<double*>(ctx->d_meta.p);
    int r = sfm::launch_fp64_probe(iters, ctx->num_sms, scratch, st, &ctx->launches);   // warm-up
    if (!r) {
        cudaEventRecord(e0, st);
        r = sfm::launch_fp64_probe(iters, ctx->num_sms, scratch, st, &ctx->launches);
        cudaEventRecord(e1, st);
    }
    cudaError_t e = cudaEventSynchronize(e1);
    if (!r && e == cudaSuccess) cudaEventElapsedTime(ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (r || e != cudaSuccess)
        return ctx->fail("probe failed: %s", cudaGetErrorString(r ? static_cast<cudaError_t>(r) : e));
    *instr = 8.0 * 512.0 * 4.0 * ctx->num_sms * static_cast<double>(iters);
    return 0;
}

int sfm_debug_dot_tile(sfm_ctx* ctx, int64_t q_row, int64_t t_row, int32_t* d_out, void* stream) {
    if (!ctx) return 1;
    if (!ctx->have_tmap) return ctx->fail("no descriptors registered");
    if (!d_out) return ctx->fail("d_out is NULL");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    int r = sfm::launch_debug_dot(ctx->tmap, static_cast<int32_t>(q_row), static_cast<int32_t>(t_row), d_out,
                                  static_cast<cudaStream_t>(stream), &ctx->launches);
    if (r) return ctx->fail("debug launch failed: %s", cudaGetErrorString(static_cast<cudaError_t>(r)));
    return 0;
}

int sfm_debug_wide_tile(sfm_ctx* ctx, int64_t q_row, int64_t t_row, int32_t* d_out, void* stream) {
    if (!ctx) return 1;
    if (!ctx->have_tmap) return ctx->fail("no descriptors registered");
    if (!d_out) return ctx->fail("d_out is NULL");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    int r = sfm::launch_debug_wide(ctx->tmap, ctx->tmap_x32, ctx->tmap_ax32, static_cast<int32_t>(q_row),
                                   static_cast<int32_t>(t_row), d_out, static_cast<cudaStream_t>(stream),
                                   &ctx->launches);
    if (r) return ctx->fail("debug launch failed: %s", cudaGetErrorString(static_cast<cudaError_t>(r)));
    return 0;
}

int sfm_debug_ts_tile(sfm_ctx* ctx, int64_t q_row, int64_t t_row, int32_t* d_out, void* stream) {
    if (!ctx) return 1;
    if (!ctx->have_tmap) return ctx->fail("no descriptors registered");
    if (!d_out) return ctx->fail("d_out is NULL");
    if (q_row < 0 || t_row < 0 || q_row + 128 > ctx->total_rows || t_row >= ctx->total_rows)
        return ctx->fail("tile rows out of range");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    int r = sfm::launch_debug_ts(ctx->tmap, ctx->tmap_ax32, ctx->d_desc, ctx->d_bias32.p, static_cast<int32_t>(q_row),
                                 static_cast<int32_t>(t_row), d_out, static_cast<cudaStream_t>(stream),
                                 &ctx->launches);
    if (r) return ctx->fail("debug launch failed: %s", cudaGetErrorString(static_cast<cudaError_t>(r)));
    return 0;
}

int sfm_debug_pair_tile(sfm_ctx* ctx, int64_t q_row, int64_t t_row, int32_t* d_out, void* stream) {
    if (!ctx) return 1;
    if (!ctx->have_tmap) return ctx->fail("no descriptors registered");
    if (!d_out) return ctx->fail("d_out is NULL");
    if (q_row < 0 || t_row < 0 || q_row >= ctx->total_rows || t_row >= ctx->total_rows)
        return ctx->fail("tile rows out of range");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    int r = sfm::launch_debug_pair(ctx->tmap, ctx->tmap_b64, ctx->tmap_x64, ctx->tmap_ax32, static_cast<int32_t>(q_row),
                                   static_cast<int32_t>(t_row), d_out, static_cast<cudaStream_t>(stream),
                                   &ctx->launches);
    if (r) return ctx->fail("debug launch failed: %s", cudaGetErrorString(static_cast<cudaError_t>(r)));
    return 0;
}

int sfm_debug_pair_counters(sfm_ctx* ctx, int64_t* h_out8, void* stream) {
    if (!ctx || !h_out8) return 1;
    if (!ctx->d_fix_list.p) return ctx->fail("no descriptors registered");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    SFM_CUDA(ctx, cudaMemcpyAsync(h_out8, ctx->d_fix_list.p + sfm_ctx::kFixCap, 8 * sizeof(int64_t),
                                  cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(stream)));
    SFM_CUDA(ctx, cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
    return 0;
}

int sfm_debug_ts_rate_probe(sfm_ctx* ctx, int32_t groups, int32_t tile_n, int32_t ts, int32_t commit, float* ms,
                            double* ops, void* stream) {
    if (!ctx) return 1;
    if (groups < 1 || !ms || !ops) return ctx->fail("bad probe arguments");
    if (tile_n != 64 && tile_n != 96 && tile_n != 128 && tile_n != 192 && tile_n != 256)
        return ctx->fail("tile_n must be 64, 96, 128, 192 or 256");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaEvent_t e0, e1;
    SFM_CUDA(ctx, cudaEventCreate(&e0));
    SFM_CUDA(ctx, cudaEventCreate(&e1));
    int r = sfm::launch_ts_rate_probe(groups, tile_n, ts, commit, ctx->num_sms, st, &ctx->launches);   // warm-up
    if (!r) {
        cudaEventRecord(e0, st);
        r = sfm::launch_ts_rate_probe(groups, tile_n, ts, commit, ctx->num_sms, st, &ctx->launches);
        cudaEventRecord(e1, st);
    }
    cudaError_t e = cudaEventSynchronize(e1);
    if (!r && e == cudaSuccess) cudaEventElapsedTime(ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (r || e != cudaSuccess)
        return ctx->fail("probe failed: %s", cudaGetErrorString(r ? static_cast<cudaError_t>(r) : e));
    *ops = 2.0 * 128.0 * tile_n * 32.0 * 5.0 * static_cast<double>(groups) * ctx->num_sms;
    return 0;
}

int sfm_debug_tmem_ld_probe(sfm_ctx* ctx, int32_t warps, int32_t variant, int32_t iters, int64_t* h_cycles,
                            void* stream) {
    if (!ctx) return 1;
    if (warps < 1 || warps > 16 || variant < 0 || variant > 6 || iters < 1 || !h_cycles)
        return ctx->fail("bad probe arguments");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t n = static_cast<size_t>(ctx->num_sms) * warps;
    SFM_CUDA(ctx, ctx->d_meta.reserve(n));
    int r = sfm::launch_tmem_ld_probe(warps, variant, iters, ctx->num_sms,
                                      reinterpret_cast<long long*>(ctx->d_meta.p), st, &ctx->launches);
    if (r) return ctx->fail("probe launch failed: %s", cudaGetErrorString(static_cast<cudaError_t>(r)));
    std::vector<int64_t> h(n);
    SFM_CUDA(ctx, cudaMemcpyAsync(h.data(), ctx->d_meta.p, n * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    SFM_CUDA(ctx, cudaStreamSynchronize(st));
    int64_t mx = 0;
    for (int64_t c : h) mx = c > mx ? c : mx;
    *h_cycles = mx;
    return 0;
}

int sfm_debug_mma_handshake_probe(sfm_ctx* ctx, int32_t steps, int32_t k_per_acc, int32_t flags,
                                  int32_t consumers, float* ms, void* stream) {
    if (!ctx) return 1;
    if (steps < 1 || k_per_acc < 1 || k_per_acc > 8 || consumers < 0 || consumers > 9 || !ms)
        return ctx->fail("bad probe arguments");
    if ((flags & 2) && (!(flags & 1) || consumers < 1)) return ctx->fail("hand-shake needs commits and consumers");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaEvent_t e0, e1;
    SFM_CUDA(ctx, cudaEventCreate(&e0));
    SFM_CUDA(ctx, cudaEventCreate(&e1));
    int r = sfm::launch_mma_handshake_probe(steps, k_per_acc, flags, consumers, ctx->num_sms, st, &ctx->launches);
    if (!r) {
        cudaEventRecord(e0, st);
        r = sfm::launch_mma_handshake_probe(steps, k_per_acc, flags, consumers, ctx->num_sms, st, &ctx->launches);
        cudaEventRecord(e1, st);
    }
    cudaError_t e = cudaEventSynchronize(e1);
    if (!r && e == cudaSuccess) cudaEventElapsedTime(ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (r || e != cudaSuccess)
        return ctx->fail("probe failed: %s", cudaGetErrorString(r ? static_cast<cudaError_t>(r) : e));
    return 0;
}

int sfm_debug_set_knn_mode(sfm_ctx* ctx, int32_t mode) {
    if (!ctx) return 1;
    if (!(mode == 0 || mode == 1 || (mode >= 5 && mode <= 7) || (mode >= 12 && mode <= 19)))
        return ctx->fail("knn mode must be 0, 1, 5..7 or 12..19");
    ctx->knn_mode = mode;
    return 0;
}

int sfm_debug_set_knn_probe_flags(sfm_ctx* ctx, int32_t flags) {
    if (!ctx) return 1;
    ctx->knn_tuning.probe_flags = static_cast<uint32_t>(flags) & 0x7FFu;
    return 0;
}

int sfm_debug_set_knn_waits(sfm_ctx* ctx, int32_t wait_ns, int32_t epi_wait_ns) {
    if (!ctx) return 1;
    if (wait_ns < 0 || epi_wait_ns < 0) return ctx->fail("wait hints must be >= 0");
    ctx->knn_tuning.wait_ns = static_cast<uint32_t>(wait_ns);
    ctx->knn_tuning.epi_wait_ns = static_cast<uint32_t>(epi_wait_ns);
    return 0;
}

int sfm_debug_handover_probe(sfm_ctx* ctx, int32_t iters, int32_t n_mma, int32_t tile_n, int32_t trail, int64_t* h_out4,
                             void* stream) {
    if (!ctx || !h_out4) return 1;
    if (iters < 1 || n_mma < 0 || n_mma > 64 || trail < 0 || trail > 64 || (tile_n != 128 && tile_n != 256))
        return ctx->fail("bad probe arguments");
    if (trail > 0 && tile_n != 128) return ctx->fail("trailing MMAs need tile_n = 128 (they use columns 128..255)");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    SFM_CUDA(ctx, ctx->d_meta.reserve(8));
    int r = sfm::launch_handover_probe(iters, n_mma, tile_n, trail, reinterpret_cast<long long*>(ctx->d_meta.p), st, &ctx->launches);
    if (r) return ctx->fail("probe launch failed: %s", cudaGetErrorString(static_cast<cudaError_t>(r)));
    SFM_CUDA(ctx, cudaMemcpyAsync(h_out4, ctx->d_meta.p, 4 * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
    SFM_CUDA(ctx, cudaStreamSynchronize(st));
    return 0;
}

int sfm_debug_fold_flag(sfm_ctx* ctx, int32_t* h_flag, void* stream) {
    if (!ctx || !h_flag) return 1;
    if (!ctx->d_flags.p) return ctx->fail("no descriptors registered");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    SFM_CUDA(ctx, cudaMemcpyAsync(h_flag, ctx->d_flags.p, sizeof(int32_t), cudaMemcpyDeviceToHost,
                                  static_cast<cudaStream_t>(stream)));
    SFM_CUDA(ctx, cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
    return 0;
}

int sfm_debug_brute_rows(sfm_ctx* ctx, int32_t* h_rows, void* stream) {
    if (!ctx || !h_rows) return 1;
    if (!ctx->d_flags.p) return ctx->fail("no descriptors registered");
    SFM_CUDA(ctx, cudaSetDevice(ctx->device));
    SFM_CUDA(ctx, cudaMemcpyAsync(h_rows, ctx->d_flags.p + 1, sizeof(int32_t), cudaMemcpyDeviceToHost,
                                  static_cast<cudaStream_t>(stream)));
    SFM_CUDA(ctx, cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
    return 0;
}

}  // extern "C"
