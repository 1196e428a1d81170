// common.cuh — shared definitions for the sm_100a stereo kernels (libb200stereo.so).
//
// Arithmetic follows SURVEY.md Appendix A/B (the verified restatement of what
// cv2.StereoSGBM.compute / cv2.remap / ... do when called from depthlib/stereo_core.py:63-75,231,
// depthlib/rectify.py:185-186 and depthlib/postprocess.py:6-171).
//
// Data layout in HBM (per handle, batch-major):
//   images        u8   [B][H][W]
//   prefilter rec u8x8 [2][B][H][W]   {g, g_lo, g_hi, r, r_lo, r_hi, 0, 0}  (A.1 + A.2 intervals)
//   C, S volumes  u16  [B][H][W1][Dp] d fastest, Dp = round_up(D, 8), or round_up(D, 16) when D > 64; C holds the block cost
//                                      WITHOUT the reference's +P2 (added back only for debug taps)
//   disparity     s16  [B][H][W]
// All volume values are in [0, 32767], so packed u16x2 arithmetic (VIMNMX.U16x2, VIMNMX3.U16x2,
// VIADDMNMX.U16x2, plain 32-bit adds where no half can carry) is exact.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/b200stereo.h"

#define B200S_MAX_STRIPES 4

struct Eff {
    int minD, D, Dp, maxD, bs, SW2, SH2, P1e, P2e, ue, d12e, ftzero, minX1, maxX1, W1, INV, mode;
    int H, W;
    int speckleWindowSize, speckleRange;
};

struct Eff;
static inline bool sum_cannot_saturate(const Eff& e);
static inline int h_imax(int a, int b) { return a > b ? a : b; }
static inline int h_imin(int a, int b) { return a < b ? a : b; }

// A.0 effective parameters
static inline Eff make_eff(const B200sParams& p, int H, int W) {
    Eff e;
    e.minD = p.minDisparity;
    e.D = p.numDisparities;
    // pitch of the d axis: whole 8-d lanes up to 64 disparities, whole 16-d lanes beyond (a lane never straddles the pitch)
    e.Dp = (e.D > 64) ? ((e.D + 15) & ~15) : ((e.D + 7) & ~7);
    e.maxD = e.minD + e.D;
    e.bs = p.blockSize > 0 ? p.blockSize : 5;
    e.SW2 = e.SH2 = e.bs / 2;
    e.P1e = p.P1 > 0 ? p.P1 : 2;
    e.P2e = h_imax(p.P2 > 0 ? p.P2 : 5, e.P1e + 1);
    e.ue = p.uniquenessRatio >= 0 ? p.uniquenessRatio : 10;
    e.d12e = p.disp12MaxDiff > 0 ? p.disp12MaxDiff : 1;
    e.ftzero = h_imax(p.preFilterCap, 15) | 1;
    e.minX1 = h_imax(e.maxD, 0);
    e.maxX1 = W + h_imin(e.minD, 0);
    e.W1 = e.maxX1 - e.minX1;
    e.INV = (e.minD - 1) * 16;
    e.mode = p.mode;
    e.H = H;
    e.W = W;
    e.speckleWindowSize = p.speckleWindowSize;
    e.speckleRange = p.speckleRange;
    return e;
}

// L(p,d) <= blockcost + P2 and blockcost <= bs^2 (2 ftzero + 63) (A.2-A.4), so the sum over all paths of a mode is
// bounded; when the bound is below 32768 the saturating adds of S never saturate and plain adds are exact.
static inline bool sum_cannot_saturate(const Eff& e) {
    const int npaths = (e.mode == B200S_MODE_HH) ? 8 : (e.mode == B200S_MODE_SGBM) ? 5 : (e.mode == B200S_MODE_HH4) ? 4 : 3;
    const long long lmax = (long long)e.bs * e.bs * (2 * e.ftzero + 63) + e.P2e;
    return npaths * lmax <= 32767;
}

// ---- packed 16x2 helpers (all single SASS instructions on sm_100a) --------------------------
__device__ __forceinline__ unsigned pk2(unsigned v) { return (v & 0xffffu) | (v << 16); }
__device__ __forceinline__ unsigned umin2(unsigned a, unsigned b) { return __vminu2(a, b); }
__device__ __forceinline__ unsigned umax2(unsigned a, unsigned b) { return __vmaxu2(a, b); }
__device__ __forceinline__ unsigned umin3_2(unsigned a, unsigned b, unsigned c) { return __vimin3_u16x2(a, b, c); }
// min(a + b, c) per u16 half
__device__ __forceinline__ unsigned uaddmin2(unsigned a, unsigned b, unsigned c) { return __viaddmin_u16x2(a, b, c); }
// (hi16 of a, lo16 of b) -> (lo, hi): the pair shifted down by one element
__device__ __forceinline__ unsigned shift_pair(unsigned a, unsigned b) { return __byte_perm(a, b, 0x5432); }

#define CUDA_TRY(h, expr)                                                                  \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            (h)->set_error(B200S_E_CUDA, "%s failed: %s", #expr, cudaGetErrorString(_e)); \
            return B200S_E_CUDA;                                                           \
        }                                                                                  \
    } while (0)

// ---- kernel launchers (defined in the .cu files) ---------------------------------------------
struct StripePlan {          // A.7: sgbm_3way row stripes; n == 1 and a[0] == 0 for the other modes
    int n;                   // number of stripes
    int a[B200S_MAX_STRIPES];   // first row computed by the stripe (overlap start)
    int r0[B200S_MAX_STRIPES];  // first row owned
    int r1[B200S_MAX_STRIPES];  // one past last row owned
    int alt_rows;            // rows of stripe-local cost per stripe (= SH2), 0 if none
};

void launch_prefilter(const Eff& e, const uint8_t* L, const uint8_t* R, int B, uint2* recL, uint2* recR,
                      cudaStream_t st);
// C volume for all rows (+ stripe-local alternative rows for sgbm_3way into Calt [stripe][alt_rows][W1][Dp] per frame)
int launch_cost(const Eff& e, const StripePlan& sp, const uint2* recL, const uint2* recR, int B, uint16_t* C,
                uint16_t* Calt, cudaStream_t st);
// all path directions of the mode, accumulated into S
int launch_paths(const Eff& e, const StripePlan& sp, int B, const uint16_t* C, const uint16_t* Calt, uint16_t* S,
                 cudaStream_t st, long long* launches, int vertical_only, int* flow);
// fused vertical family (vert.cu): 0 = launched, 1 = configuration does not fit (use launch_paths), -1 = launch error
int launch_vert_family(const Eff& e, int B, const uint16_t* C, uint16_t* S, int dy, int accumulate, cudaStream_t st,
                       int* clusters_out);
// fused horizontal paths + winner selection (rows.cu): C', Sv -> raw disparity
size_t rows_ck_bytes(const Eff& e, int B);
int rows_capacity_lines(const Eff& e);
// phase 0: both passes in one launch; 1: pass 1 only (C -> checkpoints); 2: pass 2 only
int launch_rows(const Eff& e, int B, const uint16_t* C, const uint16_t* Sv, uint16_t* ck, unsigned* keys, int16_t* raw,
                cudaStream_t st, int phase);
// few lines per call (a single pair): the call is bound by one line's serial chain, not by throughput
int rows_small_call(const Eff& e, int B);
void launch_wta(const Eff& e, int B, const uint16_t* S, int16_t* raw, cudaStream_t st);
void launch_median3_s16(const int16_t* src, int16_t* dst, int B, int H, int W, cudaStream_t st);
// speckle filter (B.3) in place; labels/sizes are int32 scratch of B*H*W each
void launch_speckle(int16_t* img, int B, int H, int W, int newVal, int maxSize, int maxDiff, int* labels, int* sizes,
                    cudaStream_t st, long long* launches);
// cn = 3: interleaved 3-channel source, gray conversion (B.2) fused into the taps
void launch_remap(const uint8_t* src, int B, int sH, int sW, int cn, const float* m1, const float* m2, int H, int W,
                  uint8_t* dst, cudaStream_t st);
void launch_bgr2gray(const uint8_t* img, size_t npix, uint8_t* dst, cudaStream_t st);
void launch_disp16_to_f32_crop(const int16_t* d16, int B, int H, int W, int x0, float* out, cudaStream_t st);
void launch_f32_to_s16x16(const float* d, size_t n, int16_t* out, cudaStream_t st);
void launch_s16_to_f32_div16(const int16_t* d, size_t n, float* out, cudaStream_t st);
void launch_outliers(const float* d, float* out, int B, int H, int W, float thr, cudaStream_t st);
void launch_outliers_s16(const int16_t* d16, float* out, int B, int H, int W, float thr, cudaStream_t st);
void launch_crop_s16(const int16_t* d16, int B, int H, int W, int x0, int16_t* out, cudaStream_t st);
void launch_median3_f32_depth(const float* src, float* dst, int B, int H, int W, const B200sDepthParams& dp, float* Z, cudaStream_t st);
void launch_median3_f32(const float* src, float* dst, int B, int H, int W, cudaStream_t st);
void launch_resize_linear(const uint8_t* src, int B, int sH, int sW, int cn, const int4* tx, const int4* ty, uint8_t* dst, int dH,
                          int dW, cudaStream_t st);
// k masked dilations (ping-pong a <-> b, result in b when k is odd, else a); returns the launch count or -1
int launch_fill_nearest(float* a, float* b, const uint8_t* mask, int H, int W, int k, cudaStream_t st);
void launch_depth(const float* disp, size_t n, const B200sDepthParams& dp, float* Z, cudaStream_t st);
// dense batches of B frames (blockIdx.y = frame)
void launch_resize_area_int(const uint8_t* src, int B, int sH, int sW, int cn, int ix, int iy, uint8_t* dst, int dH, int dW,
                            cudaStream_t st);
void launch_resize_area_tab(const uint8_t* src, int B, int sH, int sW, int cn, const int* xo, const int2* xt, const int* yo,
                            const int2* yt, uint8_t* dst, int dH, int dW, cudaStream_t st);
// cv2.initUndistortRectifyMap: params = MapParams of imgops.cu as 34 doubles (iR, fx, fy, u0, v0, k1 k2 p1 p2 k3 k4 k5 k6 s1..s4, tilt)
void launch_init_undistort_rectify_map(const double* params34, int H, int W, float* m1, float* m2, cudaStream_t st);
void launch_debug_volume(const uint16_t* V, int H, int W1, int Dp, int D, int add, int16_t* out, cudaStream_t st);
