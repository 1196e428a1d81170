/*
 * b200stereo.h — C ABI of the B200-native stereo engine (libb200stereo.so).
 *
 * The reference (mspaintenjoyer/DepthEstimation, `depthlib`) has no FFI layer of its own: its
 * boundary is the Python class StereoCore and the cv2.* calls it makes. Each entry point below
 * replaces one of those calls on the per-frame path; the reference interface it stands in for is
 * cited as depthlib/<file>:<line>. All signatures are plain C (pointers + sizes, no torch types).
 *
 * Conventions
 *   - every function returns 0 on success, a negative B200S_E_* code otherwise;
 *     b200s_last_error(h) returns a human-readable message for the last failure on that handle
 *   - "host" entry points take caller-owned host buffers (row stride given in BYTES for u8
 *     images; other buffers are dense) and are synchronous: H2D copy -> kernels -> D2H copy
 *   - "_dev" entry points take DEVICE pointers to dense batches and only enqueue work on the
 *     handle's stream (call b200s_sync or synchronise the stream yourself)
 *   - a handle owns one CUDA stream + all scratch (cost volume C, aggregate S, labels ...);
 *     use one handle per host thread / per GPU, like one cv2.StereoSGBM object per thread
 *   - there is NO CPU fallback: without a CUDA device every call fails with B200S_E_CUDA
 */
#ifndef B200STEREO_H
#define B200STEREO_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200S_ABI_VERSION 1

enum {
    B200S_OK = 0,
    B200S_E_INVALID = -1,   /* bad argument / unsupported parameter combination */
    B200S_E_CUDA = -2,      /* CUDA runtime error (no device, OOM, launch failure) */
    B200S_E_STATE = -3      /* call order problem, e.g. remap before set_maps */
};

enum {
    B200S_STAGE_REMAP = 0, B200S_STAGE_PREFILTER, B200S_STAGE_COST, B200S_STAGE_PATHS, B200S_STAGE_WTA,
    B200S_STAGE_MEDIAN, B200S_STAGE_SPECKLE, B200S_STAGE_POST, B200S_STAGE_DEPTH, B200S_NUM_STAGES
};

/* cv2.STEREO_SGBM_MODE_* values, as mapped at depthlib/stereo_core.py:55-61 */
enum { B200S_MODE_SGBM = 0, B200S_MODE_HH = 1, B200S_MODE_SGBM_3WAY = 2, B200S_MODE_HH4 = 3 };

/* Arguments of cv2.StereoSGBM_create as issued at depthlib/stereo_core.py:63-75. */
typedef struct B200sParams {
    int32_t minDisparity;       /* sgbm_params['min_disp'] */
    int32_t numDisparities;     /* sgbm_params['num_disp'] (any value >= 1, need not be %16) */
    int32_t blockSize;          /* sgbm_params['block_size'], odd, 1..11 */
    int32_t P1;                 /* 8 * block_size^2  (stereo_core.py:51) */
    int32_t P2;                 /* 32 * block_size^2 (stereo_core.py:52) */
    int32_t disp12MaxDiff;
    int32_t preFilterCap;
    int32_t uniquenessRatio;
    int32_t speckleWindowSize;
    int32_t speckleRange;
    int32_t mode;               /* B200S_MODE_* */
} B200sParams;

/* Which stages of postprocess_disparity run (B200sPostParams.fastMode). 0 and 1 are the two paths
 * StereoCore._process_pair takes; 2-4 back the stand-alone functions of depthlib/postprocess.py. */
enum {
    B200S_POST_FULL = 0,           /* speckles -> outlier mask -> median (postprocess.py:120-171) */
    B200S_POST_MEDIAN_ONLY = 1,    /* fast_mode (stereo_core.py:171-173) */
    B200S_POST_SPECKLE_ONLY = 2,   /* filter_speckles (postprocess.py:6-35) */
    B200S_POST_OUTLIERS_ONLY = 3,  /* detect_outliers + masking (postprocess.py:37-70, 151-158) */
    B200S_POST_NO_MEDIAN = 4       /* speckles -> outlier mask; the caller fills holes, then runs 1 */
};

/* Keyword arguments of postprocess_disparity as passed at depthlib/stereo_core.py:175-184
 * (hole filling is not on the GPU path). */
typedef struct B200sPostParams {
    int32_t fastMode;             /* B200S_POST_* stage selector; 1 = StereoCore fast_mode: only
                                     medianBlur(3) (stereo_core.py:171-173) */
    int32_t maxSpeckleSize;       /* int(100 * downscale_factor) */
    int32_t maxDiff16;            /* int(max_diff * 16) (postprocess.py:30) */
    float   outlierThreshold;     /* 2.5 */
    int32_t applyOutlierRemoval;  /* 1 */
} B200sPostParams;

/* Arguments of StereoCore.disparity_to_depth (depthlib/stereo_core.py:234-272). */
typedef struct B200sDepthParams {
    int32_t enabled;      /* 0: focal_length/baseline unset -> no depth (stereo_core.py:191) */
    float   fB;           /* float32(f_pixels * baseline_m), product taken in double */
    float   doffs;
    float   eps;          /* StereoCore passes min_disp here (stereo_core.py:189,195) */
    float   maxDepth;
    int32_t hasMaxDepth;
} B200sDepthParams;

typedef struct B200sHandle B200sHandle;

int b200s_abi_version(void);

/* Lifetime. b200s_create replaces StereoCore._build_sgbm (depthlib/stereo_core.py:44-75): it
 * validates the matcher parameters and binds the handle to `device`. */
int b200s_create(int device, const B200sParams* p, B200sHandle** out);
int b200s_configure(B200sHandle* h, const B200sParams* p);
int b200s_destroy(B200sHandle* h);
const char* b200s_last_error(const B200sHandle* h);
/* Use an existing cudaStream_t (e.g. torch's current stream) instead of the handle's own. */
int b200s_set_stream(B200sHandle* h, void* cuda_stream);
int b200s_sync(B200sHandle* h);

/* Rectification maps: the four CV_32FC1 maps RectificationCache.get_maps returns
 * (depthlib/rectify.py:68-80). Uploaded once per calibration; host pointers, dense HxW. */
int b200s_set_maps(B200sHandle* h, const float* map1_L, const float* map2_L,
                   const float* map1_R, const float* map2_R, int H, int W);

/* cv2.initUndistortRectifyMap(K, dist, R, P, (W, H), cv2.CV_32FC1) at depthlib/rectify.py:68-73 (cached) and :229-234 (rebuilt
 * on every call when rectify_images gets no cache): the two float32 maps of ONE camera, computed per pixel in double
 * precision on the GPU, bit-exact. K: 3x3 camera matrix, dist: 0 / 4 / 5 / 8 / 12 / 14 distortion coefficients (k1 k2 p1 p2
 * k3 k4 k5 k6 s1 s2 s3 s4 tauX tauY), R: 3x3 rectification rotation (NULL = identity), P: 3x4 new projection matrix
 * (NULL = K), all row-major doubles; map1 / map2: dense H x W float32 (host). cv2.stereoRectify (3x3 algebra, once per
 * calibration, :63-66) stays on the host. */
int b200s_build_maps(B200sHandle* h, const double* K, const double* dist, int ndist, const double* R, const double* P,
                     int H, int W, float* map1, float* map2);

/* ---- host-buffer entry points (one pair) ------------------------------------------------ */

/* cv2.cvtColor(img, COLOR_BGR2GRAY) at depthlib/rectify.py:116, stereo_core.py:157-159. */
int b200s_bgr2gray(B200sHandle* h, const uint8_t* img, int H, int W, size_t stride, uint8_t* gray);

/* cv2.resize(img, (dW, dH), interpolation=cv2.INTER_AREA) at depthlib/input.py:39-43 and
 * depthlib/threaded_stereo.py:41-46 (the downscale_factor input prep): uint8, cn = 1 or 3 interleaved channels,
 * shrinking only (dH <= sH, dW <= sW; anything else is B200S_E_INVALID). dst is dense dH x dW x cn. */
int b200s_resize_area_u8(B200sHandle* h, const uint8_t* src, int sH, int sW, int cn, size_t stride,
                         uint8_t* dst, int dH, int dW);

/* cv2.resize(image, size, interpolation=cv2.INTER_LINEAR) at depthlib/rectify.py:92-105 (_ensure_image_size: the
 * input does not match the calibration size): uint8, cn = 1 or 3 interleaved channels, any scale. OpenCV's 11-bit
 * fixed-point bilinear, bit-exact. dst is dense dH x dW x cn. */
int b200s_resize_linear_u8(B200sHandle* h, const uint8_t* src, int sH, int sW, int cn, size_t stride,
                           uint8_t* dst, int dH, int dW);

/* 2x cv2.remap(gray, map1, map2, INTER_LINEAR) at depthlib/rectify.py:185-186. Source images
 * are srcH x srcW; outputs have the size of the maps (dense). */
int b200s_remap_u8(B200sHandle* h, const uint8_t* L, const uint8_t* R, int srcH, int srcW,
                   size_t stride, uint8_t* outL, uint8_t* outR);

/* self.sgbm.compute(L, R) at depthlib/stereo_core.py:231: int16 HxW, 16x fixed point. */
int b200s_sgbm_compute(B200sHandle* h, const uint8_t* L, const uint8_t* R, int H, int W,
                       size_t stride, int16_t* disp16);

/* postprocess_disparity(...) (depthlib/postprocess.py:120-171, hole filling off) or, with
 * fastMode, cv2.medianBlur(d, 3) (stereo_core.py:173). In place on a dense float32 H x W. */
int b200s_postprocess(B200sHandle* h, float* disp, int H, int W, const B200sPostParams* pp);

/* fill_holes(disparity, mask, method='nearest', kernel_size) (depthlib/postprocess.py:72-118, the dilation branch
 * :106-116): kernel_size times, every hole pixel takes the maximum over the kernel_size x kernel_size ellipse
 * (cv2.getStructuringElement(MORPH_ELLIPSE) / cv2.dilate). mask: H x W bytes, non-zero = hole, or NULL for the
 * reference's default (disparity <= 0). In place on a dense float32 H x W; 1 <= kernel_size <= 31.
 * (method='inpaint', Telea fast marching, stays host cv2: see DESIGN.md.) */
int b200s_fill_nearest(B200sHandle* h, float* disp, const uint8_t* mask, int H, int W, int kernel_size);

/* StereoCore.disparity_to_depth (depthlib/stereo_core.py:234-272) on n dense float32 values. */
int b200s_disp_to_depth(B200sHandle* h, const float* disp, size_t n, const B200sDepthParams* dp,
                        float* depth);

/* StereoCore.estimate_depth for an already-gray pair (depthlib/stereo_core.py:274-293):
 * [remap if maps are set and use_maps] -> sgbm -> crop [:, numDisparities:] -> post -> depth.
 * disp/depth are dense H x (W - numDisparities) float32; either may be NULL (not copied back).
 * rectL/rectR (dense HxW u8) may be NULL. */
int b200s_pipeline(B200sHandle* h, const uint8_t* L, const uint8_t* R, int H, int W, size_t stride,
                   int use_maps, const B200sPostParams* pp, const B200sDepthParams* dp,
                   float* disp, float* depth, uint8_t* rectL, uint8_t* rectR);

/* ---- device-resident batch entry points (frames dense, B x H x W) ------------------------- */

int b200s_sgbm_compute_batch_dev(B200sHandle* h, const uint8_t* dL, const uint8_t* dR, int B,
                                 int H, int W, int16_t* dDisp16);
int b200s_pipeline_batch_dev(B200sHandle* h, const uint8_t* dL, const uint8_t* dR, int B, int H,
                             int W, int use_maps, const B200sPostParams* pp,
                             const B200sDepthParams* dp, float* dDisp, float* dDepth);

/* Host-buffer batch: B dense frames in, B dense results out, one H2D + one D2H per buffer.
 * What a video/batch driver (depthlib/StereoDepthEstimatorVideo.py:84-148) calls per chunk. */
int b200s_pipeline_batch(B200sHandle* h, const uint8_t* L, const uint8_t* R, int B, int H, int W,
                         int use_maps, const B200sPostParams* pp, const B200sDepthParams* dp,
                         float* disp, float* depth);

/* The same call split in two, for a frame loop that keeps the GPU busy (the generator of
 * depthlib/StereoDepthEstimatorVideo.py:84-148 reads chunk k+1 while chunk k is computed): submit enqueues
 * H2D -> kernels -> D2H and returns a ticket at once; wait blocks until that chunk's results are in disp/depth.
 * At most four chunks may be in flight (submit k+4 requires wait k); the copies run on their own streams, so
 * the H2D of chunk k+1 and the D2H of chunk k-1 overlap the kernels of chunk k. L/R/disp/depth must stay
 * valid (and should be page-locked for the copies to be asynchronous) until the ticket has been waited for. */
int b200s_pipeline_batch_submit(B200sHandle* h, const uint8_t* L, const uint8_t* R, int B, int H, int W,
                                int use_maps, const B200sPostParams* pp, const B200sDepthParams* dp,
                                float* disp, float* depth, int* ticket);
int b200s_pipeline_batch_wait(B200sHandle* h, int ticket);
/* In b200s_pipeline, b200s_pipeline_batch and b200s_pipeline_batch_submit either output may be NULL: a plane nobody
 * asked for is not copied back (disp == NULL with depth enabled = depth only, what the reference's video generator
 * yields, depthlib/StereoDepthEstimatorVideo.py:103,130). Requesting nothing is B200S_E_INVALID. */

/* The general form of the batch call: frames as they come from the capture / file (depthlib/input.py:26-43,
 * depthlib/threaded_stereo.py:36-52), input preparation on the device, any subset of the outputs.
 *   input prep, in the reference's order: [cv2.resize(INTER_AREA) to areaW x areaH  (input.py:39-43, threaded_stereo.py:41-46)]
 *   -> [cv2.resize(INTER_LINEAR) to W x H when the size still differs from the calibration size  (_ensure_image_size,
 *   rectify.py:92-105)] -> gray with the BGR weights on the channels as stored (rectify.py:108-119, stereo_core.py:156-159;
 *   fused into the remap taps when use_maps) -> [remap] -> match -> crop -> post -> depth as b200s_pipeline.
 *   L, R: B dense frames of srcH x srcW x channels uint8 (channels 1 or 3). areaH/areaW 0 = no downscale.
 *   Outputs (host, each dense B x H x (W - numDisparities), each may be NULL): disp float32; disp16 int16 = disparity x 16
 *   (exact: every post stage selects among the matcher's fixed-point values, 2 B/px instead of 4); depth float32.
 * Asynchronous like b200s_pipeline_batch_submit (same slots, same tickets, wait with b200s_pipeline_batch_wait); the
 * submit calls and b200s_pipeline_batch_wait may be issued from two different host threads of one handle (a producer
 * and a consumer; every other entry point stays one-thread-per-handle);
 * b200s_pipeline_frames = submit + wait. */
typedef struct B200sFrameIO {
    const uint8_t* L;
    const uint8_t* R;
    int32_t B, srcH, srcW, channels;
    int32_t areaH, areaW;
    int32_t H, W;
    int32_t use_maps;
    float* disp;
    int16_t* disp16;
    float* depth;
} B200sFrameIO;
int b200s_pipeline_frames_submit(B200sHandle* h, const B200sFrameIO* io, const B200sPostParams* pp,
                                 const B200sDepthParams* dp, int* ticket);
int b200s_pipeline_frames(B200sHandle* h, const B200sFrameIO* io, const B200sPostParams* pp, const B200sDepthParams* dp);

/* Page-locked host memory for the frame loop's chunk buffers ("pinned double buffers"): the copies of the
 * asynchronous calls above only overlap the kernels when their host side is page-locked. Portable across devices. */
int b200s_host_alloc(size_t bytes, void** out);
int b200s_host_free(void* p);

/* ---- allocation-free variant ---------------------------------------------------------------- */

/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * cv2.filterSpeckles(img16, new_val, max_size, max_diff) on ONE dense int16 H x W image in DEVICE memory, the same
 * kernels as the speckle stage of b200s_sgbm_compute / b200s_postprocess, but the labels + sizes scratch lives in a
 * file-scope static __device__ array (16 MiB, held by every GPU that loads the library until the process exits)
 * instead of the handle's allocations. The call allocates nothing (no cudaMalloc / cudaFree / allocating library
 * routine) and needs no handle: it enqueues on `cuda_stream` (a cudaStream_t, NULL = default stream) of the calling
 * thread's CURRENT device, which must own d_src and d_dst; d_dst may equal d_src (in place). It does not synchronise.
 * Limit: H * W <= B200S_NOALLOC_MAX_PIXELS (2 Mi pixels, e.g. 1920 x 1080); larger images are rejected with
 * B200S_E_INVALID. All calls on a device share the scratch: calls whose kernels overlap on the GPU (different streams
 * without ordering, or several threads) race on it; serialise them. The first call per device resolves the array's
 * device address (cudaGetSymbolAddress); make that call outside a graph capture.
 * On failure a message is written to err (caller buffer of err_cap bytes, may be NULL). */
#define B200S_NOALLOC_MAX_PIXELS (1 << 21)
int b200s_filter_speckles_noalloc(const int16_t* d_src, int16_t* d_dst, int H, int W, int new_val, int max_size,
                                  int max_diff, void* cuda_stream, char* err, size_t err_cap);

/* ---- introspection used by tests / bench -------------------------------------------------- */

/* Cost-domain geometry for width W: W1 = columns with a full disparity range, minX1 = first of
 * them, Dp = padded disparity pitch of the internal volumes, INV = invalid marker. */
int b200s_geometry(const B200sHandle* h, int W, int* W1, int* minX1, int* Dp, int* INV);
/* Copy an internal stage of the LAST b200s_sgbm_compute* call (frame `b`) to the host:
 * which = 0: block cost C (int16 [H][W1][D], includes +P2 like the reference's C),
 *         1: aggregate S (int16 [H][W1][D]), 2: raw disparity before median/speckle (int16 HxW). */
int b200s_debug_stage(B200sHandle* h, int which, int b, int16_t* out);
/* Batch size (frames per *_batch* call) that fills the GPU without a partial last wave for HxW images
 * under the current parameters: a multiple of the number of frames the cluster kernel keeps
 * co-resident, within six waves of the row kernel and 70 % of the free device memory. A throughput hint, any B works. */
int b200s_preferred_batch(B200sHandle* h, int H, int W, int* batch);
/* on != 0: run the matcher with the generic one-direction-per-launch kernels that materialise the
 * full aggregate S (needed for the S tap above; slower, same results). Default off: horizontal
 * paths + winner selection are fused (rows.cu) and S holds only the vertical-family aggregate. */
int b200s_set_debug_flow(B200sHandle* h, int on);
/* Which kernels the last matcher call of this handle ran (bit set of B200S_FLOW_*): tests use it to prove that the
 * fused cluster kernel, not the generic per-direction fallback, ran on every device of a multi-GPU process. */
enum {
    B200S_FLOW_VERT_DOWN = 1,      /* k_vert, top-down family */
    B200S_FLOW_VERT_UP = 2,        /* k_vert, bottom-up family (hh) */
    B200S_FLOW_PATH_VERTICAL = 4,  /* generic k_path used for a vertical / diagonal direction */
    B200S_FLOW_ROWS_FUSED = 8      /* k_rows (horizontal paths + selection) */
};
int b200s_last_flow(const B200sHandle* h);
/* Number of kernels this handle has launched since creation (bench.py's gpu_launches). */
long long b200s_launch_count(const B200sHandle* h);
/* Device time in ms of the last *_dev / host call, measured with CUDA events on the handle's
 * stream (0 if timing is disabled). */
int b200s_enable_timing(B200sHandle* h, int on);
float b200s_last_ms(const B200sHandle* h);
/* Accumulated device ms and kernel launches per stage (B200S_STAGE_* order) while timing is on.
 * Returns B200S_NUM_STAGES. */
int b200s_stage_times(const B200sHandle* h, double* ms, long long* launches, int n);
int b200s_reset_stage_times(B200sHandle* h);

#ifdef __cplusplus
}
#endif
#endif /* B200STEREO_H */
