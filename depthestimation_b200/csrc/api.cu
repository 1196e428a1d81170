// api.cu — the C ABI of libb200stereo.so (include/b200stereo.h): handle, scratch management and
// the stage graph  remap -> prefilter -> cost -> paths -> WTA -> median -> speckle -> post -> depth.
// Host-side counterpart of StereoCore (depthlib/stereo_core.py:44-75, 125-200, 212-293).
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <math.h>

#include <mutex>
#include <new>
#include <vector>

#include "common.cuh"

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <typename T> T* as() const { return (T*)p; }
};

struct B200sHandle {
    int device = 0;
    B200sParams params;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    char err[512];
    long long launches = 0;
    bool timing = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float last_ms = 0.f;
    // per-stage device timing (only when timing is on): marks are events recorded at stage ends
    enum { NSTAGE = B200S_NUM_STAGES, MAXMARK = 32 };
    cudaEvent_t mark_ev[MAXMARK] = {};
    int mark_stage[MAXMARK] = {};
    int mark_launches[MAXMARK] = {};
    int nmark = 0;
    long long mark_launch_base = 0;
    double stage_ms[NSTAGE] = {};
    long long stage_launches[NSTAGE] = {};
    // maps
    int mapH = 0, mapW = 0;
    DevBuf maps;  // 4 x H x W float
    // scratch
    DevBuf imgL, imgR, rectL, rectR, recL, recR, C, S, Calt, raw, med, labels, sizes, f0, f1, s16c, hostio, ck, keys;
    // input preparation of the frame pipeline (INTER_AREA / INTER_LINEAR results per side, gray images) and the
    // per-axis tap tables of the two resizes, cached by (source size, target size)
    DevBuf prepL[2], prepR[2], grayL, grayR, tabArea, tabLin;
    int tabAreaKey[4] = {0, 0, 0, 0}, tabLinKey[4] = {0, 0, 0, 0};
    size_t tabAreaOff[4] = {0, 0, 0, 0};
    int last_flow = 0;        // B200S_FLOW_* bits of the last matcher call
    // asynchronous host-batch pipeline: four slots, copies on their own streams so that the H2D of chunk k+1 and
    // the D2H of chunk k-1 overlap the kernels of chunk k
    // (more than two slots let chunks k+1, k+2 be uploaded while chunk k-1 is still being downloaded: with two, a download
    // slower than the kernels - eight GPUs sharing the host's PCIe / memory throughput - stalled the next upload and the GPU)
    enum { NSLOT = 4 };
    struct Slot {
        DevBuf inL, inR, out;
        cudaEvent_t in_done = nullptr, compute_done = nullptr, out_done = nullptr;
        bool busy = false;
        int rc = 0;
    } slot[NSLOT];
    cudaStream_t s_in = nullptr, s_out = nullptr;
    cudaStream_t s_aux = nullptr;                       // second kernel queue of small calls (row kernel pass 1 beside the vertical kernels)
    cudaEvent_t ev_cost = nullptr, ev_p1 = nullptr;
    unsigned next_ticket = 0;
    // submit and wait may come from two host threads (a producer that uploads and enqueues, a consumer that waits and
    // drains): the slot bookkeeping is guarded, the blocking event wait itself runs outside the lock
    std::mutex slot_mu;
    bool debug_flow = false;  // materialise the full S volume with the generic per-direction kernels (stage taps)
    // geometry of the last sgbm call (for debug taps)
    Eff last_eff;
    int last_B = 0;
    bool have_last = false;

    void set_error(int code, const char* fmt, ...) {
        (void)code;
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, sizeof(err), fmt, ap);
        va_end(ap);
    }
    void release_all() {
        DevBuf* all[] = {&maps, &imgL, &imgR, &rectL, &rectR, &recL, &recR, &C, &S, &Calt, &raw, &med,
                         &labels, &sizes, &f0, &f1, &s16c, &hostio, &ck, &keys, &prepL[0], &prepL[1], &prepR[0], &prepR[1],
                         &grayL, &grayR, &tabArea, &tabLin};
        for (DevBuf* b : all) b->release();
        for (Slot& sl : slot) { sl.inL.release(); sl.inR.release(); sl.out.release(); }
    }
};

static int validate(B200sHandle* h, const B200sParams* p) {
    if (!p) { if (h) h->set_error(B200S_E_INVALID, "params is NULL"); return B200S_E_INVALID; }
    auto fail = [&](const char* m) { if (h) h->set_error(B200S_E_INVALID, "%s", m); return B200S_E_INVALID; };
    if (p->numDisparities < 1 || p->numDisparities > 256) return fail("numDisparities must be in [1, 256]");
    int bs = p->blockSize > 0 ? p->blockSize : 5;
    if (bs % 2 == 0 || bs > 11) return fail("blockSize must be odd and <= 11");
    if (p->mode < 0 || p->mode > 3) return fail("mode must be one of sgbm(0), hh(1), sgbm_3way(2), hh4(3)");
    if (p->minDisparity < -4096 || p->minDisparity > 4096) return fail("minDisparity out of range");
    Eff e = make_eff(*p, 1, 1);
    if (e.P2e > 16384) return fail("P2 too large for the int16 aggregation (max 16384)");
    if ((long long)e.P2e + (long long)bs * bs * (2 * e.ftzero + 63) > 32767)
        return fail("blockSize/preFilterCap/P2 combination can overflow the int16 cost volume");
    if (p->mode == B200S_MODE_SGBM_3WAY && e.ue >= 100) return fail("uniquenessRatio must be < 100 in sgbm_3way mode");
    if (e.ue > 100) return fail("uniquenessRatio must be <= 100");
    return B200S_OK;
}

static StripePlan make_stripes(const Eff& e) {
    StripePlan sp;
    memset(&sp, 0, sizeof(sp));
    if (e.mode != B200S_MODE_SGBM_3WAY) {
        sp.n = 1; sp.a[0] = 0; sp.r0[0] = 0; sp.r1[0] = e.H; sp.alt_rows = 0;
        return sp;
    }
    const int nst = 4;
    const int sz = (e.H + nst - 1) / nst;
    const int ov = (e.bs / 2 + 1) + (sz + 9) / 10;  // ceil(0.1 * sz)
    sp.n = nst;
    sp.alt_rows = e.SH2;
    for (int n = 0; n < nst; n++) {
        int r0 = n * sz, r1 = h_imin((n + 1) * sz, e.H);
        if (r0 >= r1) { sp.a[n] = sp.r0[n] = sp.r1[n] = 0; continue; }
        sp.a[n] = h_imax(h_imin(r0 - ov, e.H), 0);
        sp.r0[n] = r0;
        sp.r1[n] = r1;
    }
    return sp;
}

extern "C" {

int b200s_abi_version(void) { return B200S_ABI_VERSION; }

int b200s_create(int device, const B200sParams* p, B200sHandle** out) {
    if (!out) return B200S_E_INVALID;
    *out = nullptr;
    int rc = validate(nullptr, p);
    if (rc) return rc;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return B200S_E_CUDA;
    if (cudaSetDevice(device) != cudaSuccess) return B200S_E_CUDA;
    B200sHandle* h = new (std::nothrow) B200sHandle();
    if (!h) return B200S_E_CUDA;
    h->err[0] = 0;
    h->device = device;
    h->params = *p;
    if (cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking) != cudaSuccess) { delete h; return B200S_E_CUDA; }
    h->stream = h->own_stream;
    cudaEventCreate(&h->ev0);
    cudaEventCreate(&h->ev1);
    cudaStreamCreateWithFlags(&h->s_in, cudaStreamNonBlocking);
    cudaStreamCreateWithFlags(&h->s_out, cudaStreamNonBlocking);
    cudaStreamCreateWithFlags(&h->s_aux, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&h->ev_cost, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&h->ev_p1, cudaEventDisableTiming);
    for (B200sHandle::Slot& sl : h->slot) {
        cudaEventCreateWithFlags(&sl.in_done, cudaEventDisableTiming);
        cudaEventCreateWithFlags(&sl.compute_done, cudaEventDisableTiming);
        cudaEventCreateWithFlags(&sl.out_done, cudaEventDisableTiming);
    }
    *out = h;
    return B200S_OK;
}

int b200s_configure(B200sHandle* h, const B200sParams* p) {
    if (!h) return B200S_E_INVALID;
    int rc = validate(h, p);
    if (rc) return rc;
    h->params = *p;
    h->have_last = false;
    return B200S_OK;
}

int b200s_destroy(B200sHandle* h) {
    if (!h) return B200S_OK;
    cudaSetDevice(h->device);
    cudaStreamSynchronize(h->stream);
    if (h->s_in) cudaStreamSynchronize(h->s_in);
    if (h->s_out) cudaStreamSynchronize(h->s_out);
    h->release_all();
    for (B200sHandle::Slot& sl : h->slot) {
        if (sl.in_done) cudaEventDestroy(sl.in_done);
        if (sl.compute_done) cudaEventDestroy(sl.compute_done);
        if (sl.out_done) cudaEventDestroy(sl.out_done);
    }
    if (h->s_aux) { cudaStreamSynchronize(h->s_aux); cudaStreamDestroy(h->s_aux); }
    if (h->ev_cost) cudaEventDestroy(h->ev_cost);
    if (h->ev_p1) cudaEventDestroy(h->ev_p1);
    if (h->s_in) cudaStreamDestroy(h->s_in);
    if (h->s_out) cudaStreamDestroy(h->s_out);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    delete h;
    return B200S_OK;
}

const char* b200s_last_error(const B200sHandle* h) { return h ? h->err : "null handle"; }

int b200s_set_stream(B200sHandle* h, void* s) {
    if (!h) return B200S_E_INVALID;
    h->stream = s ? (cudaStream_t)s : h->own_stream;
    return B200S_OK;
}

int b200s_sync(B200sHandle* h) {
    if (!h) return B200S_E_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->device));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return B200S_OK;
}

int b200s_preferred_batch(B200sHandle* h, int H, int W, int* batch) {
    if (!h || !batch || H <= 0 || W <= 0) return B200S_E_INVALID;
    CUDA_TRY(h, cudaSetDevice(h->device));
    Eff e = make_eff(h->params, H, W);
    *batch = 16;
    if (e.W1 <= 0) return B200S_OK;
    int granule = 0;
    if (e.mode == B200S_MODE_SGBM || e.mode == B200S_MODE_HH)
        if (launch_vert_family(e, 0, nullptr, nullptr, +1, 0, h->stream, &granule) != 0) granule = 0;
    // measured on B200: the row kernel wants whole waves (a partly filled last wave costs up to 10 % per frame) and
    // more than one of them (the second wave de-synchronises the phases of the CTAs: ~25 % faster per frame than a
    // single wave); three waves -> six still gains 1.5 % (kernel tails amortised), six -> nine nothing
    int cap = 6 * rows_capacity_lines(e) / H;
    // never more than ~70 % of the free HBM: C and S are 2 B/cell each, the O(N) buffers about 64 B/px
    size_t freeb = 0, totalb = 0;
    if (cudaMemGetInfo(&freeb, &totalb) == cudaSuccess) {
        const double per_frame = 4.0 * H * (double)e.W1 * e.Dp + 64.0 * H * W;
        const double fit = 0.7 * (double)freeb / per_frame;
        if (fit < cap) cap = fit < 1.0 ? 1 : (int)fit;
    }
    if (cap < 1) cap = 1;
    if (granule > cap) granule = 0;
    if (granule <= 0) { *batch = cap; return B200S_OK; }
    int b = granule;
    while (b + granule <= cap) b += granule;   // largest multiple of the cluster granule within six row-kernel waves
    *batch = b;
    return B200S_OK;
}
int b200s_set_debug_flow(B200sHandle* h, int on) { if (!h) return B200S_E_INVALID; h->debug_flow = on != 0; return B200S_OK; }
int b200s_enable_timing(B200sHandle* h, int on) { if (!h) return B200S_E_INVALID; h->timing = on != 0; return B200S_OK; }
float b200s_last_ms(const B200sHandle* h) { return h ? h->last_ms : 0.f; }
long long b200s_launch_count(const B200sHandle* h) { return h ? h->launches : 0; }

int b200s_geometry(const B200sHandle* h, int W, int* W1, int* minX1, int* Dp, int* INV) {
    if (!h) return B200S_E_INVALID;
    Eff e = make_eff(h->params, 1, W);
    if (W1) *W1 = e.W1;
    if (minX1) *minX1 = e.minX1;
    if (Dp) *Dp = e.Dp;
    if (INV) *INV = e.INV;
    return B200S_OK;
}

int b200s_set_maps(B200sHandle* h, const float* m1L, const float* m2L, const float* m1R, const float* m2R, int H, int W) {
    if (!h) return B200S_E_INVALID;
    if (!m1L || !m2L || !m1R || !m2R || H <= 0 || W <= 0) { h->set_error(B200S_E_INVALID, "set_maps: bad arguments"); return B200S_E_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    size_t n = (size_t)H * W;
    CUDA_TRY(h, h->maps.ensure(4 * n * sizeof(float)));
    const float* src[4] = {m1L, m2L, m1R, m2R};
    for (int i = 0; i < 4; i++)
        CUDA_TRY(h, cudaMemcpyAsync(h->maps.as<float>() + i * n, src[i], n * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    h->mapH = H; h->mapW = W;
    return B200S_OK;
}

}  // extern "C"

// ---- internal stage graph (device pointers) -----------------------------------------------------------

// end-of-stage mark: device time since the previous mark is attributed to `stage`
static void mark(B200sHandle* h, int stage) {
    if (!h->timing || h->nmark >= B200sHandle::MAXMARK) return;
    int i = h->nmark++;
    if (!h->mark_ev[i]) cudaEventCreate(&h->mark_ev[i]);
    cudaEventRecord(h->mark_ev[i], h->stream);
    h->mark_stage[i] = stage;
    h->mark_launches[i] = (int)(h->launches - h->mark_launch_base);
    h->mark_launch_base = h->launches;
}

static int check_launch(B200sHandle* h, const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { h->set_error(B200S_E_CUDA, "%s: %s", what, cudaGetErrorString(e)); return B200S_E_CUDA; }
    return B200S_OK;
}

static int dev_remap(B200sHandle* h, const uint8_t* dL, const uint8_t* dR, int B, int sH, int sW, int cn, uint8_t* oL, uint8_t* oR) {
    if (!h->maps.p) { h->set_error(B200S_E_STATE, "remap: rectification maps not set (call b200s_set_maps)"); return B200S_E_STATE; }
    size_t n = (size_t)h->mapH * h->mapW;
    const float* m = h->maps.as<float>();
    launch_remap(dL, B, sH, sW, cn, m, m + n, h->mapH, h->mapW, oL, h->stream);
    launch_remap(dR, B, sH, sW, cn, m + 2 * n, m + 3 * n, h->mapH, h->mapW, oR, h->stream);
    h->launches += 2;
    mark(h, B200S_STAGE_REMAP);
    return check_launch(h, "remap");
}

// cv2.StereoSGBM.compute on a dense device batch
static int dev_sgbm(B200sHandle* h, const uint8_t* dL, const uint8_t* dR, int B, int H, int W, int16_t* dDisp) {
    if (B <= 0 || H <= 0 || W <= 0) { h->set_error(B200S_E_INVALID, "sgbm: empty input"); return B200S_E_INVALID; }
    Eff e = make_eff(h->params, H, W);
    const size_t npx = (size_t)B * H * W;
    h->last_eff = e; h->last_B = B; h->have_last = true;
    h->last_flow = 0;
    if (e.W1 <= 0) {  // no column has a full disparity range: everything invalid (A.0)
        CUDA_TRY(h, h->raw.ensure(npx * 2));
        // rare path: host-side fill
        int16_t* tmp = (int16_t*)malloc(npx * 2);
        if (!tmp) { h->set_error(B200S_E_CUDA, "out of host memory"); return B200S_E_CUDA; }
        for (size_t i = 0; i < npx; i++) tmp[i] = (int16_t)e.INV;
        cudaError_t ce = cudaMemcpyAsync(dDisp, tmp, npx * 2, cudaMemcpyHostToDevice, h->stream);
        if (ce == cudaSuccess) ce = cudaMemcpyAsync(h->raw.p, tmp, npx * 2, cudaMemcpyHostToDevice, h->stream);
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(h->stream);
        free(tmp);
        if (ce != cudaSuccess) { h->set_error(B200S_E_CUDA, "fill: %s", cudaGetErrorString(ce)); return B200S_E_CUDA; }
        return B200S_OK;
    }
    const StripePlan sp = make_stripes(e);
    const size_t vol = (size_t)B * H * e.W1 * e.Dp * sizeof(uint16_t);
    CUDA_TRY(h, h->recL.ensure(npx * sizeof(uint2)));
    CUDA_TRY(h, h->recR.ensure(npx * sizeof(uint2)));
    CUDA_TRY(h, h->C.ensure(vol));
    CUDA_TRY(h, h->S.ensure(vol));
    if (sp.alt_rows > 0 && sp.n > 1)
        CUDA_TRY(h, h->Calt.ensure((size_t)B * B200S_MAX_STRIPES * sp.alt_rows * e.W1 * e.Dp * sizeof(uint16_t)));
    CUDA_TRY(h, h->raw.ensure(npx * 2));
    launch_prefilter(e, dL, dR, B, h->recL.as<uint2>(), h->recR.as<uint2>(), h->stream);
    h->launches++;
    mark(h, B200S_STAGE_PREFILTER);
    int rc = launch_cost(e, sp, h->recL.as<uint2>(), h->recR.as<uint2>(), B, h->C.as<uint16_t>(), h->Calt.as<uint16_t>(), h->stream);
    if (rc) { h->set_error(B200S_E_INVALID, "cost kernel: unsupported configuration (%d)", rc); return B200S_E_INVALID; }
    h->launches++;
    mark(h, B200S_STAGE_COST);
    if (h->debug_flow) {
        rc = launch_paths(e, sp, B, h->C.as<uint16_t>(), h->Calt.as<uint16_t>(), h->S.as<uint16_t>(), h->stream, &h->launches, 0, &h->last_flow);
        if (rc) { h->set_error(B200S_E_INVALID, "path kernel: unsupported configuration (%d)", rc); return B200S_E_INVALID; }
        mark(h, B200S_STAGE_PATHS);
        launch_wta(e, B, h->S.as<uint16_t>(), h->raw.as<int16_t>(), h->stream);
        h->launches++;
    } else {
        CUDA_TRY(h, h->ck.ensure(rows_ck_bytes(e, B)));
        CUDA_TRY(h, h->keys.ensure(npx * sizeof(unsigned)));
        // Few lines per call: the row kernel's first pass (C -> checkpoints of the right-to-left scan) needs nothing from
        // the vertical kernels, so it runs beside them on a second stream; its second pass follows both.
        // (sgbm / hh only: their vertical cluster kernel is long enough to hide it. Measured, one KITTI pair: sgbm 1.81 ->
        // 1.67 ms; sgbm_3way, whose vertical pass is 0.1 ms, 1.43 -> 1.48 ms with the split, so it keeps one launch)
        const bool split = rows_small_call(e, B) && h->s_aux && h->ev_cost && h->ev_p1 &&
                           (e.mode == B200S_MODE_SGBM || e.mode == B200S_MODE_HH);
        if (split) {
            CUDA_TRY(h, cudaEventRecord(h->ev_cost, h->stream));
            CUDA_TRY(h, cudaStreamWaitEvent(h->s_aux, h->ev_cost, 0));
            rc = launch_rows(e, B, h->C.as<uint16_t>(), h->S.as<uint16_t>(), h->ck.as<uint16_t>(), h->keys.as<unsigned>(),
                             h->raw.as<int16_t>(), h->s_aux, 1);
            if (rc) { h->set_error(B200S_E_INVALID, "row kernel: unsupported configuration (%d)", rc); return B200S_E_INVALID; }
            CUDA_TRY(h, cudaEventRecord(h->ev_p1, h->s_aux));
            h->launches++;
        }
        rc = launch_paths(e, sp, B, h->C.as<uint16_t>(), h->Calt.as<uint16_t>(), h->S.as<uint16_t>(), h->stream, &h->launches, 1, &h->last_flow);
        if (rc == -4) { h->set_error(B200S_E_CUDA, "vertical cluster kernel: planned but the launch failed on device %d (see stderr)", h->device); return B200S_E_CUDA; }
        if (rc) { h->set_error(B200S_E_INVALID, "path kernel: unsupported configuration (%d)", rc); return B200S_E_INVALID; }
        mark(h, B200S_STAGE_PATHS);
        if (split) CUDA_TRY(h, cudaStreamWaitEvent(h->stream, h->ev_p1, 0));
        rc = launch_rows(e, B, h->C.as<uint16_t>(), h->S.as<uint16_t>(), h->ck.as<uint16_t>(), h->keys.as<unsigned>(),
                         h->raw.as<int16_t>(), h->stream, split ? 2 : 0);
        if (rc) { h->set_error(B200S_E_INVALID, "row kernel: unsupported configuration (%d)", rc); return B200S_E_INVALID; }
        h->launches++;
        h->last_flow |= B200S_FLOW_ROWS_FUSED;
    }
    mark(h, B200S_STAGE_WTA);
    launch_median3_s16(h->raw.as<int16_t>(), dDisp, B, H, W, h->stream);
    h->launches++;
    mark(h, B200S_STAGE_MEDIAN);
    if (e.speckleWindowSize > 0) {
        CUDA_TRY(h, h->labels.ensure(npx * sizeof(int)));
        CUDA_TRY(h, h->sizes.ensure(npx * sizeof(int)));
        launch_speckle(dDisp, B, H, W, e.INV, e.speckleWindowSize, 16 * e.speckleRange, h->labels.as<int>(), h->sizes.as<int>(),
                       h->stream, &h->launches);
        mark(h, B200S_STAGE_SPECKLE);
    }
    return check_launch(h, "sgbm");
}

// postprocess_disparity (hole filling off) / fast-mode median on a dense device batch, result in `out`
// postprocess_disparity (depthlib/postprocess.py:86-124). The input is either float disparity (dIn) or, inside the
// fused pipeline, the matcher's own int16 x16 disparity (dIn16: (d*16).astype(int16) of d = v/16 is v itself, so
// the float round trip is skipped). When dp/depthOut are given and the last stage is the 3x3 median, depth is
// produced by that same kernel. Returns 1 in *depth_done when it was.
static int dev_post_impl(B200sHandle* h, const float* dIn, const int16_t* dIn16, int B, int H, int W,
                         const B200sPostParams& pp, float* out, const B200sDepthParams* dp, float* depthOut,
                         bool* depth_done) {
    const size_t n = (size_t)B * H * W;
    const int mode = pp.fastMode;   // B200S_POST_*
    if (depth_done) *depth_done = false;
    if (mode < 0 || mode > B200S_POST_NO_MEDIAN) { h->set_error(B200S_E_INVALID, "postprocess: bad fastMode/stage selector %d", mode); return B200S_E_INVALID; }
    const bool fuse_depth = dp && dp->enabled && depthOut;
    CUDA_TRY(h, h->f1.ensure(n * sizeof(float)));
    float* t = h->f1.as<float>();
    if (mode == B200S_POST_MEDIAN_ONLY || mode == B200S_POST_OUTLIERS_ONLY) {
        const float* src = dIn;
        if (!src) {   // int16 input: make the float view first
            launch_s16_to_f32_div16(dIn16, n, t, h->stream);
            h->launches++;
            src = t;
        }
        if (mode == B200S_POST_OUTLIERS_ONLY) launch_outliers(src, out, B, H, W, pp.outlierThreshold, h->stream);
        else if (fuse_depth) { launch_median3_f32_depth(src, out, B, H, W, *dp, depthOut, h->stream); if (depth_done) *depth_done = true; }
        else launch_median3_f32(src, out, B, H, W, h->stream);
        h->launches++;
        return check_launch(h, "post(single stage)");
    }
    CUDA_TRY(h, h->labels.ensure(n * sizeof(int)));
    CUDA_TRY(h, h->sizes.ensure(n * sizeof(int)));
    int16_t* s16;
    if (dIn16) {
        s16 = const_cast<int16_t*>(dIn16);          // filtered in place: the caller hands over a scratch copy
    } else {
        CUDA_TRY(h, h->s16c.ensure(n * 2));
        s16 = h->s16c.as<int16_t>();
        launch_f32_to_s16x16(dIn, n, s16, h->stream);
        h->launches++;
    }
    launch_speckle(s16, B, H, W, 0, pp.maxSpeckleSize, pp.maxDiff16, h->labels.as<int>(), h->sizes.as<int>(), h->stream, &h->launches);
    const bool outl = pp.applyOutlierRemoval && mode != B200S_POST_SPECKLE_ONLY;
    const bool med = (mode == B200S_POST_FULL);
    float* pre = med ? t : out;                     // what the median (if any) reads
    if (outl) launch_outliers_s16(s16, pre, B, H, W, pp.outlierThreshold, h->stream);
    else launch_s16_to_f32_div16(s16, n, pre, h->stream);
    h->launches++;
    if (med) {
        if (fuse_depth) { launch_median3_f32_depth(t, out, B, H, W, *dp, depthOut, h->stream); if (depth_done) *depth_done = true; }
        else launch_median3_f32(t, out, B, H, W, h->stream);
        h->launches++;
    }
    return check_launch(h, "post");
}

static int dev_post(B200sHandle* h, const float* dIn, int B, int H, int W, const B200sPostParams& pp, float* out) {
    return dev_post_impl(h, dIn, nullptr, B, H, W, pp, out, nullptr, nullptr, nullptr);
}

// dL / dR: dense batches of H x W frames with cn interleaved channels (1 = gray, 3 = as stored, converted with the
// BGR weights like cv2.cvtColor(COLOR_BGR2GRAY) at depthlib/rectify.py:116 / stereo_core.py:157-159)
static int dev_pipeline(B200sHandle* h, const uint8_t* dL, const uint8_t* dR, int cn, int B, int H, int W, int use_maps,
                        const B200sPostParams& pp, const B200sDepthParams& dp, float* dDisp, float* dDepth,
                        const uint8_t** rectLout, const uint8_t** rectRout) {
    const size_t npx = (size_t)B * H * W;
    const uint8_t* mL = dL;
    const uint8_t* mR = dR;
    if (cn == 3 && !use_maps) {      // no calibration: gray only (stereo_core.py:156-159)
        CUDA_TRY(h, h->grayL.ensure(npx));
        CUDA_TRY(h, h->grayR.ensure(npx));
        launch_bgr2gray(dL, npx, h->grayL.as<uint8_t>(), h->stream);
        launch_bgr2gray(dR, npx, h->grayR.as<uint8_t>(), h->stream);
        h->launches += 2;
        mark(h, B200S_STAGE_REMAP);
        mL = h->grayL.as<uint8_t>();
        mR = h->grayR.as<uint8_t>();
    }
    if (use_maps) {
        if (h->mapH != H || h->mapW != W) {
            h->set_error(B200S_E_INVALID, "pipeline: image size %dx%d does not match the rectification maps %dx%d", W, H, h->mapW, h->mapH);
            return B200S_E_INVALID;
        }
        CUDA_TRY(h, h->rectL.ensure(npx));
        CUDA_TRY(h, h->rectR.ensure(npx));
        int rc = dev_remap(h, dL, dR, B, H, W, cn, h->rectL.as<uint8_t>(), h->rectR.as<uint8_t>());
        if (rc) return rc;
        mL = h->rectL.as<uint8_t>();
        mR = h->rectR.as<uint8_t>();
    }
    if (rectLout) *rectLout = mL;
    if (rectRout) *rectRout = mR;
    const int x0 = h->params.numDisparities;  // crop [:, num_disp:] (stereo_core.py:168)
    if (x0 >= W) { h->set_error(B200S_E_INVALID, "pipeline: numDisparities >= image width"); return B200S_E_INVALID; }
    const int Wc = W - x0;
    CUDA_TRY(h, h->med.ensure(npx * 2));
    int rc = dev_sgbm(h, mL, mR, B, H, W, h->med.as<int16_t>());
    if (rc) return rc;
    const size_t nc = (size_t)B * H * Wc;
    // crop [:, num_disp:] in int16; everything after it works on the cropped frame (stereo_core.py:168-196)
    CUDA_TRY(h, h->s16c.ensure(nc * 2));
    launch_crop_s16(h->med.as<int16_t>(), B, H, W, x0, h->s16c.as<int16_t>(), h->stream);
    h->launches++;
    bool depth_done = false;
    rc = dev_post_impl(h, nullptr, h->s16c.as<int16_t>(), B, H, Wc, pp, dDisp, &dp, dDepth, &depth_done);
    if (rc) return rc;
    mark(h, B200S_STAGE_POST);
    if (dp.enabled && dDepth && !depth_done) {
        launch_depth(dDisp, nc, dp, dDepth, h->stream);
        h->launches++;
        mark(h, B200S_STAGE_DEPTH);
    }
    return check_launch(h, "pipeline");
}

struct Timed {
    B200sHandle* h;
    explicit Timed(B200sHandle* h_) : h(h_) {
        if (h->timing) {
            h->nmark = 0;
            h->mark_launch_base = h->launches;
            cudaEventRecord(h->ev0, h->stream);
        }
    }
    void stop() {
        if (h->timing) {
            cudaEventRecord(h->ev1, h->stream);
            cudaEventSynchronize(h->ev1);
            cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1);
            cudaEvent_t prev = h->ev0;
            for (int i = 0; i < h->nmark; i++) {
                float ms = 0.f;
                cudaEventElapsedTime(&ms, prev, h->mark_ev[i]);
                h->stage_ms[h->mark_stage[i]] += ms;
                h->stage_launches[h->mark_stage[i]] += h->mark_launches[i];
                prev = h->mark_ev[i];
            }
        }
    }
};

static int upload_u8(B200sHandle* h, DevBuf& buf, const uint8_t* src, int H, int W, size_t stride) {
    CUDA_TRY(h, buf.ensure((size_t)H * W));
    CUDA_TRY(h, cudaMemcpy2DAsync(buf.p, W, src, stride, W, H, cudaMemcpyHostToDevice, h->stream));
    return B200S_OK;
}

extern "C" {

int b200s_bgr2gray(B200sHandle* h, const uint8_t* img, int H, int W, size_t stride, uint8_t* gray) {
    if (!h) return B200S_E_INVALID;
    if (!img || !gray || H <= 0 || W <= 0 || stride < (size_t)3 * W) { h->set_error(B200S_E_INVALID, "bgr2gray: bad arguments"); return B200S_E_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    size_t n = (size_t)H * W;
    CUDA_TRY(h, h->hostio.ensure(3 * n));
    CUDA_TRY(h, h->imgL.ensure(n));
    CUDA_TRY(h, cudaMemcpy2DAsync(h->hostio.p, 3 * (size_t)W, img, stride, 3 * (size_t)W, H, cudaMemcpyHostToDevice, h->stream));
    launch_bgr2gray(h->hostio.as<uint8_t>(), n, h->imgL.as<uint8_t>(), h->stream);
    h->launches++;
    CUDA_TRY(h, cudaMemcpyAsync(gray, h->imgL.p, n, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return check_launch(h, "bgr2gray");
}

// DecimateAlpha table of one axis, exactly as the reference's dependency builds it (double precision on the host)
static void area_table(int ssize, int dsize, std::vector<int>& ofs, std::vector<int2>& taps) {
    const double scale = (double)ssize / dsize;
    ofs.clear(); taps.clear();
    auto push = [&](int si, double w) { float f = (float)w; int2 t; t.x = si; memcpy(&t.y, &f, 4); taps.push_back(t); };
    for (int d = 0; d < dsize; d++) {
        const double f1 = d * scale, f2 = f1 + scale;
        const double cell = scale < ssize - f1 ? scale : ssize - f1;
        int s1 = (int)ceil(f1), s2 = (int)floor(f2);
        if (s2 > ssize - 1) s2 = ssize - 1;
        if (s1 > s2) s1 = s2;
        ofs.push_back((int)taps.size());
        if (s1 - f1 > 1e-3) push(s1 - 1, (s1 - f1) / cell);
        for (int sx = s1; sx < s2; sx++) push(sx, 1.0 / cell);
        if (f2 - s2 > 1e-3) { double t = f2 - s2; if (t > 1.0) t = 1.0; if (t > cell) t = cell; push(s2, t / cell); }
    }
    ofs.push_back((int)taps.size());
}

int b200s_resize_area_u8(B200sHandle* h, const uint8_t* src, int sH, int sW, int cn, size_t stride, uint8_t* dst, int dH, int dW) {
    if (!h) return B200S_E_INVALID;
    if (!src || !dst || sH <= 0 || sW <= 0 || dH <= 0 || dW <= 0 || (cn != 1 && cn != 3) || stride < (size_t)cn * sW) { h->set_error(B200S_E_INVALID, "resize_area: bad arguments"); return B200S_E_INVALID; }
    if (dH > sH || dW > sW) { h->set_error(B200S_E_INVALID, "resize_area: only shrinking is supported (%dx%d -> %dx%d)", sW, sH, dW, dH); return B200S_E_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    const size_t ns = (size_t)sH * sW * cn, nd = (size_t)dH * dW * cn;
    CUDA_TRY(h, h->hostio.ensure(ns));
    CUDA_TRY(h, h->imgL.ensure(nd));
    CUDA_TRY(h, cudaMemcpy2DAsync(h->hostio.p, (size_t)cn * sW, src, stride, (size_t)cn * sW, sH, cudaMemcpyHostToDevice, h->stream));
    const int ix = sW / dW, iy = sH / dH;
    if (ix * dW == sW && iy * dH == sH) {
        launch_resize_area_int(h->hostio.as<uint8_t>(), 1, sH, sW, cn, ix, iy, h->imgL.as<uint8_t>(), dH, dW, h->stream);
    } else {
        std::vector<int> xo, yo; std::vector<int2> xt, yt;
        area_table(sW, dW, xo, xt);
        area_table(sH, dH, yo, yt);
        const size_t bx = xo.size() * 4, by = yo.size() * 4, tx = xt.size() * 8, ty = yt.size() * 8;
        auto al = [](size_t v) { return (v + 15) & ~(size_t)15; };
        CUDA_TRY(h, h->imgR.ensure(al(bx) + al(by) + al(tx) + al(ty)));
        char* base = h->imgR.as<char>();
        int* dxo = (int*)base; int* dyo = (int*)(base + al(bx));
        int2* dxt = (int2*)(base + al(bx) + al(by)); int2* dyt = (int2*)(base + al(bx) + al(by) + al(tx));
        CUDA_TRY(h, cudaMemcpyAsync(dxo, xo.data(), bx, cudaMemcpyHostToDevice, h->stream));
        CUDA_TRY(h, cudaMemcpyAsync(dyo, yo.data(), by, cudaMemcpyHostToDevice, h->stream));
        CUDA_TRY(h, cudaMemcpyAsync(dxt, xt.data(), tx, cudaMemcpyHostToDevice, h->stream));
        CUDA_TRY(h, cudaMemcpyAsync(dyt, yt.data(), ty, cudaMemcpyHostToDevice, h->stream));
        CUDA_TRY(h, cudaStreamSynchronize(h->stream));     // the host vectors go out of scope below
        launch_resize_area_tab(h->hostio.as<uint8_t>(), 1, sH, sW, cn, dxo, dxt, dyo, dyt, h->imgL.as<uint8_t>(), dH, dW, h->stream);
    }
    h->launches++;
    CUDA_TRY(h, cudaMemcpyAsync(dst, h->imgL.p, nd, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return check_launch(h, "resize_area");
}

// per-axis taps of cv2's INTER_LINEAR for uint8 (see oracle/imgops_oracle.c: orc_linear_tab)
static void linear_table(int ssize, int dsize, bool snap, std::vector<int4>& tab) {
    const double scale = 1.0 / ((double)dsize / ssize);
    tab.resize(dsize);
    for (int d = 0; d < dsize; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= (float)s;
        if (snap) {
            if (s < 0) { f = 0.f; s = 0; }
            if (s >= ssize - 1) { f = 0.f; s = ssize - 1; }
        }
        int c0 = (int)lrintf((1.f - f) * 2048.f), c1 = (int)lrintf(f * 2048.f);
        if (c0 > 32767) c0 = 32767;
        if (c1 > 32767) c1 = 32767;
        tab[d] = make_int4(s, c0, c1, 0);
    }
}

int b200s_resize_linear_u8(B200sHandle* h, const uint8_t* src, int sH, int sW, int cn, size_t stride, uint8_t* dst, int dH, int dW) {
    if (!h) return B200S_E_INVALID;
    if (!src || !dst || sH <= 0 || sW <= 0 || dH <= 0 || dW <= 0 || (cn != 1 && cn != 3) || stride < (size_t)cn * sW) { h->set_error(B200S_E_INVALID, "resize_linear: bad arguments"); return B200S_E_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    const size_t ns = (size_t)sH * sW * cn, nd = (size_t)dH * dW * cn;
    CUDA_TRY(h, h->hostio.ensure(ns));
    CUDA_TRY(h, h->imgL.ensure(nd));
    CUDA_TRY(h, cudaMemcpy2DAsync(h->hostio.p, (size_t)cn * sW, src, stride, (size_t)cn * sW, sH, cudaMemcpyHostToDevice, h->stream));
    std::vector<int4> tx, ty;
    linear_table(sW, dW, true, tx);
    linear_table(sH, dH, false, ty);
    CUDA_TRY(h, h->imgR.ensure((tx.size() + ty.size()) * sizeof(int4)));
    int4* dtx = h->imgR.as<int4>();
    int4* dty = dtx + tx.size();
    CUDA_TRY(h, cudaMemcpyAsync(dtx, tx.data(), tx.size() * sizeof(int4), cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(dty, ty.data(), ty.size() * sizeof(int4), cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));     // the host vectors go out of scope below
    launch_resize_linear(h->hostio.as<uint8_t>(), 1, sH, sW, cn, dtx, dty, h->imgL.as<uint8_t>(), dH, dW, h->stream);
    h->launches++;
    CUDA_TRY(h, cudaMemcpyAsync(dst, h->imgL.p, nd, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return check_launch(h, "resize_linear");
}

int b200s_remap_u8(B200sHandle* h, const uint8_t* L, const uint8_t* R, int srcH, int srcW, size_t stride, uint8_t* outL, uint8_t* outR) {
    if (!h) return B200S_E_INVALID;
    if (!L || !R || !outL || !outR || srcH <= 0 || srcW <= 0 || stride < (size_t)srcW) { h->set_error(B200S_E_INVALID, "remap: bad arguments"); return B200S_E_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    if (!h->maps.p) { h->set_error(B200S_E_STATE, "remap: rectification maps not set (call b200s_set_maps)"); return B200S_E_STATE; }
    int rc = upload_u8(h, h->imgL, L, srcH, srcW, stride); if (rc) return rc;
    rc = upload_u8(h, h->imgR, R, srcH, srcW, stride); if (rc) return rc;
    size_t n = (size_t)h->mapH * h->mapW;
    CUDA_TRY(h, h->rectL.ensure(n));
    CUDA_TRY(h, h->rectR.ensure(n));
    rc = dev_remap(h, h->imgL.as<uint8_t>(), h->imgR.as<uint8_t>(), 1, srcH, srcW, 1, h->rectL.as<uint8_t>(), h->rectR.as<uint8_t>());
    if (rc) return rc;
    CUDA_TRY(h, cudaMemcpyAsync(outL, h->rectL.p, n, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(outR, h->rectR.p, n, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return B200S_OK;
}

int b200s_sgbm_compute(B200sHandle* h, const uint8_t* L, const uint8_t* R, int H, int W, size_t stride, int16_t* disp16) {
    if (!h) return B200S_E_INVALID;
    if (!L || !R || !disp16 || H <= 0 || W <= 0 || stride < (size_t)W) { h->set_error(B200S_E_INVALID, "sgbm_compute: bad arguments"); return B200S_E_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    int rc = upload_u8(h, h->imgL, L, H, W, stride); if (rc) return rc;
    rc = upload_u8(h, h->imgR, R, H, W, stride); if (rc) return rc;
    size_t n = (size_t)H * W;
    CUDA_TRY(h, h->med.ensure(n * 2));
    Timed t(h);
    rc = dev_sgbm(h, h->imgL.as<uint8_t>(), h->imgR.as<uint8_t>(), 1, H, W, h->med.as<int16_t>());
    if (rc) return rc;
    t.stop();
    CUDA_TRY(h, cudaMemcpyAsync(disp16, h->med.p, n * 2, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return check_launch(h, "sgbm_compute");
}

int b200s_postprocess(B200sHandle* h, float* disp, int H, int W, const B200sPostParams* pp) {
    if (!h) return B200S_E_INVALID;
    if (!disp || !pp || H <= 0 || W <= 0) { h->set_error(B200S_E_INVALID, "postprocess: bad arguments"); return B200S_E_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    size_t n = (size_t)H * W;
    CUDA_TRY(h, h->f0.ensure(n * sizeof(float)));
    CUDA_TRY(h, h->hostio.ensure(n * sizeof(float)));
    CUDA_TRY(h, cudaMemcpyAsync(h->f0.p, disp, n * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    int rc = dev_post(h, h->f0.as<float>(), 1, H, W, *pp, h->hostio.as<float>());
    if (rc) return rc;
    CUDA_TRY(h, cudaMemcpyAsync(disp, h->hostio.p, n * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return check_launch(h, "postprocess");
}

int b200s_fill_nearest(B200sHandle* h, float* disp, const uint8_t* mask, int H, int W, int kernel_size) {
    if (!h) return B200S_E_INVALID;
    if (!disp || H <= 0 || W <= 0 || kernel_size < 1 || kernel_size > 31) { h->set_error(B200S_E_INVALID, "fill_nearest: bad arguments (1 <= kernel_size <= 31)"); return B200S_E_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    const size_t n = (size_t)H * W;
    CUDA_TRY(h, h->f0.ensure(n * sizeof(float)));
    CUDA_TRY(h, h->f1.ensure(n * sizeof(float)));
    CUDA_TRY(h, h->imgL.ensure(n));
    CUDA_TRY(h, cudaMemcpyAsync(h->f0.p, disp, n * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    std::vector<uint8_t> m;
    const uint8_t* hm = mask;
    if (!hm) {                                   // the reference's default mask: disparity <= 0
        m.resize(n);
        for (size_t i = 0; i < n; i++) m[i] = disp[i] <= 0.0f;
        hm = m.data();
    }
    CUDA_TRY(h, cudaMemcpyAsync(h->imgL.p, hm, n, cudaMemcpyHostToDevice, h->stream));
    const int nl = launch_fill_nearest(h->f0.as<float>(), h->f1.as<float>(), h->imgL.as<uint8_t>(), H, W, kernel_size, h->stream);
    if (nl < 0) { h->set_error(B200S_E_INVALID, "fill_nearest: unsupported kernel size"); return B200S_E_INVALID; }
    h->launches += nl;
    const float* res = (kernel_size & 1) ? h->f1.as<float>() : h->f0.as<float>();
    CUDA_TRY(h, cudaMemcpyAsync(disp, res, n * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));   // also keeps `m` alive until the upload has been consumed
    return check_launch(h, "fill_nearest");
}

int b200s_disp_to_depth(B200sHandle* h, const float* disp, size_t n, const B200sDepthParams* dp, float* depth) {
    if (!h) return B200S_E_INVALID;
    if (!disp || !depth || !dp || n == 0) { h->set_error(B200S_E_INVALID, "disp_to_depth: bad arguments"); return B200S_E_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    CUDA_TRY(h, h->f0.ensure(n * sizeof(float)));
    CUDA_TRY(h, h->f1.ensure(n * sizeof(float)));
    CUDA_TRY(h, cudaMemcpyAsync(h->f0.p, disp, n * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    launch_depth(h->f0.as<float>(), n, *dp, h->f1.as<float>(), h->stream);
    h->launches++;
    CUDA_TRY(h, cudaMemcpyAsync(depth, h->f1.p, n * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return check_launch(h, "disp_to_depth");
}

int b200s_pipeline(B200sHandle* h, const uint8_t* L, const uint8_t* R, int H, int W, size_t stride, int use_maps,
                   const B200sPostParams* pp, const B200sDepthParams* dp, float* disp, float* depth, uint8_t* rectL, uint8_t* rectR) {
    if (!h) return B200S_E_INVALID;
    if (!L || !R || !pp || !dp || H <= 0 || W <= 0 || stride < (size_t)W) { h->set_error(B200S_E_INVALID, "pipeline: bad arguments"); return B200S_E_INVALID; }
    if (!disp && !(dp->enabled && depth)) { h->set_error(B200S_E_INVALID, "pipeline: no output requested (disp is NULL and depth is NULL or disabled)"); return B200S_E_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    int rc = upload_u8(h, h->imgL, L, H, W, stride); if (rc) return rc;
    rc = upload_u8(h, h->imgR, R, H, W, stride); if (rc) return rc;
    const int Wc = W - h->params.numDisparities;
    if (Wc <= 0) { h->set_error(B200S_E_INVALID, "pipeline: numDisparities >= image width"); return B200S_E_INVALID; }
    size_t nc = (size_t)H * Wc;
    CUDA_TRY(h, h->hostio.ensure(2 * nc * sizeof(float)));
    float* dDisp = h->hostio.as<float>();
    float* dDepth = dDisp + nc;
    const uint8_t *rl = nullptr, *rr = nullptr;
    Timed t(h);
    rc = dev_pipeline(h, h->imgL.as<uint8_t>(), h->imgR.as<uint8_t>(), 1, 1, H, W, use_maps, *pp, *dp, dDisp, dDepth, &rl, &rr);
    if (rc) return rc;
    t.stop();
    if (disp) CUDA_TRY(h, cudaMemcpyAsync(disp, dDisp, nc * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    if (dp->enabled && depth) CUDA_TRY(h, cudaMemcpyAsync(depth, dDepth, nc * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    if (rectL) CUDA_TRY(h, cudaMemcpyAsync(rectL, rl, (size_t)H * W, cudaMemcpyDeviceToHost, h->stream));
    if (rectR) CUDA_TRY(h, cudaMemcpyAsync(rectR, rr, (size_t)H * W, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return check_launch(h, "pipeline");
}

int b200s_sgbm_compute_batch_dev(B200sHandle* h, const uint8_t* dL, const uint8_t* dR, int B, int H, int W, int16_t* dDisp16) {
    if (!h) return B200S_E_INVALID;
    if (!dL || !dR || !dDisp16) { h->set_error(B200S_E_INVALID, "sgbm_compute_batch_dev: bad arguments"); return B200S_E_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    Timed t(h);
    int rc = dev_sgbm(h, dL, dR, B, H, W, dDisp16);
    if (rc) return rc;
    t.stop();
    return B200S_OK;
}

int b200s_pipeline_batch_dev(B200sHandle* h, const uint8_t* dL, const uint8_t* dR, int B, int H, int W, int use_maps,
                             const B200sPostParams* pp, const B200sDepthParams* dp, float* dDisp, float* dDepth) {
    if (!h) return B200S_E_INVALID;
    if (!dL || !dR || !pp || !dp || !dDisp) { h->set_error(B200S_E_INVALID, "pipeline_batch_dev: bad arguments"); return B200S_E_INVALID; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    Timed t(h);
    int rc = dev_pipeline(h, dL, dR, 1, B, H, W, use_maps, *pp, *dp, dDisp, dDepth, nullptr, nullptr);
    if (rc) return rc;
    t.stop();
    return B200S_OK;
}

int b200s_debug_stage(B200sHandle* h, int which, int b, int16_t* out) {
    if (!h) return B200S_E_INVALID;
    if (!h->have_last || !out || b < 0 || b >= h->last_B) { h->set_error(B200S_E_STATE, "debug_stage: no previous sgbm call / bad frame"); return B200S_E_STATE; }
    CUDA_TRY(h, cudaSetDevice(h->device));
    const Eff& e = h->last_eff;
    if (which == 2) {
        size_t n = (size_t)e.H * e.W;
        CUDA_TRY(h, cudaMemcpyAsync(out, h->raw.as<int16_t>() + (size_t)b * n, n * 2, cudaMemcpyDeviceToHost, h->stream));
        CUDA_TRY(h, cudaStreamSynchronize(h->stream));
        return B200S_OK;
    }
    if (which != 0 && which != 1) { h->set_error(B200S_E_INVALID, "debug_stage: which must be 0, 1 or 2"); return B200S_E_INVALID; }
    if (e.W1 <= 0) { h->set_error(B200S_E_STATE, "debug_stage: empty cost domain"); return B200S_E_STATE; }
    size_t cells = (size_t)e.H * e.W1 * e.D;
    CUDA_TRY(h, h->hostio.ensure(cells * 2));
    const uint16_t* V = (which == 0 ? h->C.as<uint16_t>() : h->S.as<uint16_t>()) + (size_t)b * e.H * e.W1 * e.Dp;
    launch_debug_volume(V, e.H, e.W1, e.Dp, e.D, which == 0 ? e.P2e : 0, h->hostio.as<int16_t>(), h->stream);
    CUDA_TRY(h, cudaMemcpyAsync(out, h->hostio.p, cells * 2, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return check_launch(h, "debug_stage");
}

int b200s_stage_times(const B200sHandle* h, double* ms, long long* launches, int n) {
    if (!h) return B200S_E_INVALID;
    for (int i = 0; i < n && i < B200S_NUM_STAGES; i++) {
        if (ms) ms[i] = h->stage_ms[i];
        if (launches) launches[i] = h->stage_launches[i];
    }
    return B200S_NUM_STAGES;
}

int b200s_reset_stage_times(B200sHandle* h) {
    if (!h) return B200S_E_INVALID;
    for (int i = 0; i < B200S_NUM_STAGES; i++) { h->stage_ms[i] = 0.0; h->stage_launches[i] = 0; }
    return B200S_OK;
}

// ---- input preparation on the device ---------------------------------------------------------------
}  // extern "C"

// per-axis tables of the two resizes, built on the host once per (source, target) size and kept on the device
static int prep_area_tables(B200sHandle* h, int sH, int sW, int dH, int dW) {
    const int key[4] = {sH, sW, dH, dW};
    if (!memcmp(key, h->tabAreaKey, sizeof(key))) return B200S_OK;
    std::vector<int> xo, yo; std::vector<int2> xt, yt;
    area_table(sW, dW, xo, xt);
    area_table(sH, dH, yo, yt);
    auto al = [](size_t v) { return (v + 15) & ~(size_t)15; };
    const size_t bx = xo.size() * 4, by = yo.size() * 4, tx = xt.size() * 8, ty = yt.size() * 8;
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));              // an earlier call may still read the old tables
    CUDA_TRY(h, h->tabArea.ensure(al(bx) + al(by) + al(tx) + al(ty)));
    h->tabAreaOff[0] = 0; h->tabAreaOff[1] = al(bx); h->tabAreaOff[2] = al(bx) + al(by); h->tabAreaOff[3] = al(bx) + al(by) + al(tx);
    char* base = h->tabArea.as<char>();
    CUDA_TRY(h, cudaMemcpyAsync(base + h->tabAreaOff[0], xo.data(), bx, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(base + h->tabAreaOff[1], yo.data(), by, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(base + h->tabAreaOff[2], xt.data(), tx, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(base + h->tabAreaOff[3], yt.data(), ty, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));              // the host vectors go out of scope
    memcpy(h->tabAreaKey, key, sizeof(key));
    return B200S_OK;
}
static int prep_linear_tables(B200sHandle* h, int sH, int sW, int dH, int dW) {
    const int key[4] = {sH, sW, dH, dW};
    if (!memcmp(key, h->tabLinKey, sizeof(key))) return B200S_OK;
    std::vector<int4> tx, ty;
    linear_table(sW, dW, true, tx);
    linear_table(sH, dH, false, ty);
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    CUDA_TRY(h, h->tabLin.ensure((tx.size() + ty.size()) * sizeof(int4)));
    CUDA_TRY(h, cudaMemcpyAsync(h->tabLin.as<int4>(), tx.data(), tx.size() * sizeof(int4), cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(h->tabLin.as<int4>() + tx.size(), ty.data(), ty.size() * sizeof(int4), cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    memcpy(h->tabLinKey, key, sizeof(key));
    return B200S_OK;
}

// One side of a batch: src (B x sH x sW x cn) -> [INTER_AREA to aH x aW] -> [INTER_LINEAR to H x W], channels kept
// (gray conversion happens in the remap taps or in dev_pipeline). *out = the prepared frames (src itself if nothing ran).
static int dev_prep_side(B200sHandle* h, const uint8_t* src, int B, int sH, int sW, int cn, int aH, int aW, int H, int W,
                         DevBuf (&tmp)[2], const uint8_t** out) {
    const uint8_t* cur = src;
    int cH = sH, cW = sW;
    if (aH != cH || aW != cW) {            // cv2.resize(img, (aW, aH), INTER_AREA): depthlib/input.py:39-43, threaded_stereo.py:41-46
        CUDA_TRY(h, tmp[0].ensure((size_t)B * aH * aW * cn));
        const int ix = cW / aW, iy = cH / aH;
        if (ix * aW == cW && iy * aH == cH) {
            launch_resize_area_int(cur, B, cH, cW, cn, ix, iy, tmp[0].as<uint8_t>(), aH, aW, h->stream);
        } else {
            int rc = prep_area_tables(h, cH, cW, aH, aW);
            if (rc) return rc;
            char* base = h->tabArea.as<char>();
            launch_resize_area_tab(cur, B, cH, cW, cn, (const int*)(base + h->tabAreaOff[0]), (const int2*)(base + h->tabAreaOff[2]),
                                   (const int*)(base + h->tabAreaOff[1]), (const int2*)(base + h->tabAreaOff[3]),
                                   tmp[0].as<uint8_t>(), aH, aW, h->stream);
        }
        h->launches++;
        cur = tmp[0].as<uint8_t>(); cH = aH; cW = aW;
    }
    if (H != cH || W != cW) {              // _ensure_image_size: cv2.resize(image, size, INTER_LINEAR), depthlib/rectify.py:92-105
        int rc = prep_linear_tables(h, cH, cW, H, W);
        if (rc) return rc;
        CUDA_TRY(h, tmp[1].ensure((size_t)B * H * W * cn));
        launch_resize_linear(cur, B, cH, cW, cn, h->tabLin.as<int4>(), h->tabLin.as<int4>() + W, tmp[1].as<uint8_t>(), H, W, h->stream);
        h->launches++;
        cur = tmp[1].as<uint8_t>();
    }
    *out = cur;
    return B200S_OK;
}

static int frames_validate(B200sHandle* h, const B200sFrameIO* io, const B200sPostParams* pp, const B200sDepthParams* dp, int* aH, int* aW) {
    if (!io || !pp || !dp || !io->L || !io->R || io->B <= 0 || io->srcH <= 0 || io->srcW <= 0 || io->H <= 0 || io->W <= 0 ||
        (io->channels != 1 && io->channels != 3)) { h->set_error(B200S_E_INVALID, "pipeline_frames: bad arguments"); return B200S_E_INVALID; }
    *aH = io->areaH > 0 ? io->areaH : io->srcH;
    *aW = io->areaW > 0 ? io->areaW : io->srcW;
    if (*aH > io->srcH || *aW > io->srcW) { h->set_error(B200S_E_INVALID, "pipeline_frames: INTER_AREA only shrinks (%dx%d -> %dx%d)", io->srcW, io->srcH, *aW, *aH); return B200S_E_INVALID; }
    if (!io->disp && !io->disp16 && !(dp->enabled && io->depth)) { h->set_error(B200S_E_INVALID, "pipeline_frames: no output requested"); return B200S_E_INVALID; }
    if (io->W - h->params.numDisparities <= 0) { h->set_error(B200S_E_INVALID, "pipeline_frames: numDisparities >= image width"); return B200S_E_INVALID; }
    return B200S_OK;
}

extern "C" {

int b200s_pipeline_frames_submit(B200sHandle* h, const B200sFrameIO* io, const B200sPostParams* pp, const B200sDepthParams* dp,
                                 int* ticket) {
    if (!h) return B200S_E_INVALID;
    if (!ticket) { h->set_error(B200S_E_INVALID, "pipeline_frames: ticket is NULL"); return B200S_E_INVALID; }
    int aH = 0, aW = 0;
    int rc = frames_validate(h, io, pp, dp, &aH, &aW);
    if (rc) return rc;
    CUDA_TRY(h, cudaSetDevice(h->device));
    std::lock_guard<std::mutex> lk(h->slot_mu);
    const int B = io->B, H = io->H, W = io->W, cn = io->channels;
    const int Wc = W - h->params.numDisparities;
    const size_t nin = (size_t)B * io->srcH * io->srcW * cn, nc = (size_t)B * H * Wc;
    const unsigned tk = h->next_ticket;
    B200sHandle::Slot& sl = h->slot[tk % B200sHandle::NSLOT];
    if (sl.busy) { h->set_error(B200S_E_STATE, "pipeline submit: ticket %u of this slot has not been waited for", tk - B200sHandle::NSLOT); return B200S_E_STATE; }
    // slot output: float disparity, float depth, int16 disparity
    const size_t out_bytes = 2 * nc * sizeof(float) + nc * sizeof(int16_t);
    // growing a buffer frees device memory: nothing of this handle may still be running on it
    if (nin > sl.inL.cap || nin > sl.inR.cap || out_bytes > sl.out.cap) {
        CUDA_TRY(h, cudaStreamSynchronize(h->stream));
        CUDA_TRY(h, cudaStreamSynchronize(h->s_out));
        CUDA_TRY(h, sl.inL.ensure(nin));
        CUDA_TRY(h, sl.inR.ensure(nin));
        CUDA_TRY(h, sl.out.ensure(out_bytes));
    }
    // H2D on the copy-in stream (the slot's previous user has been waited for, so its kernels are done)
    CUDA_TRY(h, cudaMemcpyAsync(sl.inL.p, io->L, nin, cudaMemcpyHostToDevice, h->s_in));
    CUDA_TRY(h, cudaMemcpyAsync(sl.inR.p, io->R, nin, cudaMemcpyHostToDevice, h->s_in));
    CUDA_TRY(h, cudaEventRecord(sl.in_done, h->s_in));
    CUDA_TRY(h, cudaStreamWaitEvent(h->stream, sl.in_done, 0));
    float* dDisp = sl.out.as<float>();
    float* dDepth = dDisp + nc;
    int16_t* dDisp16 = (int16_t*)(dDepth + nc);
    {
        Timed t(h);
        const uint8_t *pL = nullptr, *pR = nullptr;
        rc = dev_prep_side(h, sl.inL.as<uint8_t>(), B, io->srcH, io->srcW, cn, aH, aW, H, W, h->prepL, &pL);
        if (rc) return rc;
        rc = dev_prep_side(h, sl.inR.as<uint8_t>(), B, io->srcH, io->srcW, cn, aH, aW, H, W, h->prepR, &pR);
        if (rc) return rc;
        rc = dev_pipeline(h, pL, pR, cn, B, H, W, io->use_maps, *pp, *dp, dDisp, dDepth, nullptr, nullptr);
        if (rc) return rc;
        if (io->disp16) {   // every post stage selects among the matcher's x16 values (or 0): the float map is exactly s16 / 16
            launch_f32_to_s16x16(dDisp, nc, dDisp16, h->stream);
            h->launches++;
            rc = check_launch(h, "disp16");
            if (rc) return rc;
        }
        t.stop();
    }
    CUDA_TRY(h, cudaEventRecord(sl.compute_done, h->stream));
    CUDA_TRY(h, cudaStreamWaitEvent(h->s_out, sl.compute_done, 0));
    // one plain copy per requested plane; planes nobody asked for stay on the device
    if (io->disp) CUDA_TRY(h, cudaMemcpyAsync(io->disp, dDisp, nc * sizeof(float), cudaMemcpyDeviceToHost, h->s_out));
    if (io->disp16) CUDA_TRY(h, cudaMemcpyAsync(io->disp16, dDisp16, nc * sizeof(int16_t), cudaMemcpyDeviceToHost, h->s_out));
    if (dp->enabled && io->depth) CUDA_TRY(h, cudaMemcpyAsync(io->depth, dDepth, nc * sizeof(float), cudaMemcpyDeviceToHost, h->s_out));
    CUDA_TRY(h, cudaEventRecord(sl.out_done, h->s_out));
    sl.busy = true;
    *ticket = (int)tk;
    h->next_ticket = tk + 1;
    return B200S_OK;
}

int b200s_pipeline_batch_submit(B200sHandle* h, const uint8_t* L, const uint8_t* R, int B, int H, int W, int use_maps,
                                const B200sPostParams* pp, const B200sDepthParams* dp, float* disp, float* depth,
                                int* ticket) {
    if (!h) return B200S_E_INVALID;
    B200sFrameIO io;
    memset(&io, 0, sizeof(io));
    io.L = L; io.R = R; io.B = B; io.srcH = H; io.srcW = W; io.channels = 1; io.H = H; io.W = W; io.use_maps = use_maps;
    io.disp = disp; io.depth = depth;
    return b200s_pipeline_frames_submit(h, &io, pp, dp, ticket);
}

int b200s_pipeline_batch_wait(B200sHandle* h, int ticket) {
    if (!h) return B200S_E_INVALID;
    B200sHandle::Slot& sl = h->slot[(unsigned)ticket % B200sHandle::NSLOT];
    const unsigned tk = (unsigned)ticket;
    {
        std::lock_guard<std::mutex> lk(h->slot_mu);
        if (!sl.busy || tk >= h->next_ticket || h->next_ticket - tk > B200sHandle::NSLOT) { h->set_error(B200S_E_STATE, "pipeline_batch_wait: ticket %d is not in flight", ticket); return B200S_E_STATE; }
    }
    CUDA_TRY(h, cudaSetDevice(h->device));
    CUDA_TRY(h, cudaEventSynchronize(sl.out_done));   // on failure the slot stays busy: its copies may still be pending
    std::lock_guard<std::mutex> lk(h->slot_mu);
    sl.busy = false;
    return check_launch(h, "pipeline_batch");
}

int b200s_pipeline_batch(B200sHandle* h, const uint8_t* L, const uint8_t* R, int B, int H, int W, int use_maps,
                         const B200sPostParams* pp, const B200sDepthParams* dp, float* disp, float* depth) {
    int ticket = 0;
    int rc = b200s_pipeline_batch_submit(h, L, R, B, H, W, use_maps, pp, dp, disp, depth, &ticket);
    if (rc) return rc;
    return b200s_pipeline_batch_wait(h, ticket);
}

int b200s_pipeline_frames(B200sHandle* h, const B200sFrameIO* io, const B200sPostParams* pp, const B200sDepthParams* dp) {
    int ticket = 0;
    int rc = b200s_pipeline_frames_submit(h, io, pp, dp, &ticket);
    if (rc) return rc;
    return b200s_pipeline_batch_wait(h, ticket);
}

int b200s_last_flow(const B200sHandle* h) { return h ? h->last_flow : 0; }

// 3x3 inverse the way the reference's dependency inverts small matrices (cofactors over the determinant, double)
static bool inv3(const double* a, double* o) {
    const double d = a[0] * (a[4] * a[8] - a[5] * a[7]) - a[1] * (a[3] * a[8] - a[5] * a[6]) + a[2] * (a[3] * a[7] - a[4] * a[6]);
    if (d == 0.0) return false;
    const double r = 1.0 / d;
    o[0] = (a[4] * a[8] - a[5] * a[7]) * r; o[1] = (a[2] * a[7] - a[1] * a[8]) * r; o[2] = (a[1] * a[5] - a[2] * a[4]) * r;
    o[3] = (a[5] * a[6] - a[3] * a[8]) * r; o[4] = (a[0] * a[8] - a[2] * a[6]) * r; o[5] = (a[2] * a[3] - a[0] * a[5]) * r;
    o[6] = (a[3] * a[7] - a[4] * a[6]) * r; o[7] = (a[1] * a[6] - a[0] * a[7]) * r; o[8] = (a[0] * a[4] - a[1] * a[3]) * r;
    return true;
}

int b200s_build_maps(B200sHandle* h, const double* K, const double* dist, int ndist, const double* R, const double* P,
                     int H, int W, float* map1, float* map2) {
    if (!h) return B200S_E_INVALID;
    if (!K || !map1 || !map2 || H <= 0 || W <= 0 || ndist < 0 || (ndist > 0 && !dist) ||
        !(ndist == 0 || ndist == 4 || ndist == 5 || ndist == 8 || ndist == 12 || ndist == 14)) {
        h->set_error(B200S_E_INVALID, "build_maps: bad arguments (distortion vectors of 0, 4, 5, 8, 12 or 14 coefficients)");
        return B200S_E_INVALID;
    }
    CUDA_TRY(h, cudaSetDevice(h->device));
    double q[34];
    const double I3[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    const double* Rm = R ? R : I3;
    double A[9];   // new camera matrix (first three columns of P, or K itself) times R
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) {
            double acc = 0.0;
            for (int k = 0; k < 3; k++) acc += (P ? P[r * 4 + k] : K[r * 3 + k]) * Rm[k * 3 + c];
            A[r * 3 + c] = acc;
        }
    if (!inv3(A, q)) { h->set_error(B200S_E_INVALID, "build_maps: P[:, :3] * R is singular"); return B200S_E_INVALID; }
    q[9] = K[0]; q[10] = K[4]; q[11] = K[2]; q[12] = K[5];
    double d[14] = {0};
    for (int i = 0; i < ndist; i++) d[i] = dist[i];
    q[13] = d[0]; q[14] = d[1]; q[15] = d[2]; q[16] = d[3]; q[17] = d[4]; q[18] = d[5]; q[19] = d[6]; q[20] = d[7];
    q[21] = d[8]; q[22] = d[9]; q[23] = d[10]; q[24] = d[11];
    // tilt (tauX, tauY): projection of the tilted sensor, identity when both are zero
    double* T = q + 25;
    for (int i = 0; i < 9; i++) T[i] = I3[i];
    if (d[12] != 0.0 || d[13] != 0.0) {
        const double cx = cos(d[12]), sx = sin(d[12]), cy = cos(d[13]), sy = sin(d[13]);
        const double RX[9] = {1, 0, 0, 0, cx, sx, 0, -sx, cx}, RY[9] = {cy, 0, -sy, 0, 1, 0, sy, 0, cy};
        double RXY[9];
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) { double acc = 0.0; for (int k = 0; k < 3; k++) acc += RY[r * 3 + k] * RX[k * 3 + c]; RXY[r * 3 + c] = acc; }
        const double PZ[9] = {RXY[8], 0, -RXY[2], 0, RXY[8], -RXY[5], 0, 0, 1};
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) { double acc = 0.0; for (int k = 0; k < 3; k++) acc += PZ[r * 3 + k] * RXY[k * 3 + c]; T[r * 3 + c] = acc; }
    }
    const size_t n = (size_t)H * W;
    CUDA_TRY(h, h->f0.ensure(n * sizeof(float)));
    CUDA_TRY(h, h->f1.ensure(n * sizeof(float)));
    launch_init_undistort_rectify_map(q, H, W, h->f0.as<float>(), h->f1.as<float>(), h->stream);
    h->launches++;
    CUDA_TRY(h, cudaMemcpyAsync(map1, h->f0.p, n * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaMemcpyAsync(map2, h->f1.p, n * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(h, cudaStreamSynchronize(h->stream));
    return check_launch(h, "build_maps");
}

int b200s_host_alloc(size_t bytes, void** out) {
    if (!out || bytes == 0) return B200S_E_INVALID;
    *out = nullptr;
    if (cudaHostAlloc(out, bytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); *out = nullptr; return B200S_E_CUDA; }
    return B200S_OK;
}
int b200s_host_free(void* p) {
    if (!p) return B200S_OK;
    if (cudaFreeHost(p) != cudaSuccess) { cudaGetLastError(); return B200S_E_CUDA; }
    return B200S_OK;
}

}  // extern "C"
