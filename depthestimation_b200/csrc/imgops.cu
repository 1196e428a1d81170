// imgops.cu — the O(N) image ops on depthlib's stereo path:
//   remap         cv2.remap(gray, map1, map2, INTER_LINEAR)          depthlib/rectify.py:185-186
//   bgr2gray      cv2.cvtColor(img, COLOR_BGR2GRAY)                  depthlib/rectify.py:116, stereo_core.py:157-159
//   median3 s16   the medianBlur(3) inside StereoSGBM.compute (A.8)  depthlib/stereo_core.py:231
//   speckle       cv2.filterSpeckles (A.8 and postprocess.py:30)     depthlib/postprocess.py:6-35
//   outliers      detect_outliers + masking                          depthlib/postprocess.py:37-70,151-158
//   median3 f32   cv2.medianBlur(float32, 3)                         depthlib/postprocess.py:169, stereo_core.py:173
//   depth         StereoCore.disparity_to_depth                      depthlib/stereo_core.py:234-272
//   resize lin    cv2.resize(img, size, INTER_LINEAR) (size fix-up)  depthlib/rectify.py:92-105
//   fill nearest  fill_holes(method='nearest')                       depthlib/postprocess.py:106-116
// The float chain uses explicit round-to-nearest intrinsics (no FMA contraction, no fast math).
#include <string.h>

#include "common.cuh"

// ---- B.1 remap: fixed-point bilinear, 5 fractional bits, constant-0 border ---------------------
// Every O(N) kernel below gives a thread several pixels (or a CTA several rows): with one pixel per thread a batch is
// ~140 000 tiny CTAs per launch and the launch rate of CTAs (~0.7 per clock), not memory, sets the time (~95 us).
#define REMAP_PPT 4
// CN = 3: the source is the interleaved 3-channel frame and every tap is converted to gray on the fly (B.2), which
// is what remap of the cvtColor result reads at that tap (depthlib/rectify.py:182-186) without the gray image ever
// going to memory.
template <int CN>
__device__ __forceinline__ int remap_tap(const uint8_t* r, int x) {
    if (CN == 1) return r[x];
    const uint8_t* p = r + 3 * (size_t)x;
    return (3735 * p[0] + 19235 * p[1] + 9798 * p[2] + 16384) >> 15;
}
// A thread owns REMAP_PPT consecutive pixels and walks REMAP_FRAMES frames of the batch with them: the maps are the same
// for every frame, so the two float loads, the rounding to 5 fractional bits and the border tests are done once per
// pixel and used REMAP_FRAMES times (one launch over 150 KITTI frames moved 8 B of maps per 1-byte output through L2).
#define REMAP_FRAMES 4
template <int CN>
__global__ void __launch_bounds__(256) k_remap(const uint8_t* __restrict__ src, int sH, int sW,
                                               const float* __restrict__ m1, const float* __restrict__ m2, int H, int W,
                                               uint8_t* __restrict__ dst, int B) {
    const size_t n = (size_t)H * W;
    const size_t i0 = ((size_t)blockIdx.x * 256 + threadIdx.x) * REMAP_PPT;
    if (i0 >= n) return;
    const int np = (int)min((size_t)REMAP_PPT, n - i0);
    // 16-byte map loads, 4-byte stores: frames (and the second map) then start on aligned boundaries
    const bool vec = (np == REMAP_PPT) && ((n & 3) == 0) && ((((size_t)m1 | (size_t)m2) & 15) == 0) && (((size_t)dst & 3) == 0);
    float a1[REMAP_PPT], a2[REMAP_PPT];
    if (vec) {
        const float4 u = *(const float4*)(m1 + i0), v = *(const float4*)(m2 + i0);
        a1[0] = u.x; a1[1] = u.y; a1[2] = u.z; a1[3] = u.w;
        a2[0] = v.x; a2[1] = v.y; a2[2] = v.z; a2[3] = v.w;
    } else {
#pragma unroll
        for (int k = 0; k < REMAP_PPT; k++) { a1[k] = k < np ? m1[i0 + k] : -4.0f; a2[k] = k < np ? m2[i0 + k] : -4.0f; }
    }
    // per pixel: byte offsets of the two source rows and two columns (-1: outside, the tap reads as 0) and the weights
    int o0[REMAP_PPT], o1[REMAP_PPT], c0[REMAP_PPT], c1[REMAP_PPT], wgt[REMAP_PPT];
#pragma unroll
    for (int k = 0; k < REMAP_PPT; k++) {
        const int sx = __float2int_rn(__fmul_rn(a1[k], 32.0f));
        const int sy = __float2int_rn(__fmul_rn(a2[k], 32.0f));
        const int ix = sx >> 5, iy = sy >> 5;
        wgt[k] = (sx & 31) | ((sy & 31) << 8);
        c0[k] = (ix >= 0 && ix < sW) ? ix : -1;
        c1[k] = (ix + 1 >= 0 && ix + 1 < sW) ? ix + 1 : -1;
        o0[k] = (iy >= 0 && iy < sH) ? iy * sW * CN : -1;
        o1[k] = (iy + 1 >= 0 && iy + 1 < sH) ? (iy + 1) * sW * CN : -1;
    }
    const int f0 = blockIdx.y * REMAP_FRAMES, f1 = min(f0 + REMAP_FRAMES, B);
    for (int f = f0; f < f1; f++) {
        const uint8_t* s = src + (size_t)f * sH * sW * CN;
        unsigned out = 0u;
#pragma unroll
        for (int k = 0; k < REMAP_PPT; k++) {
            const int fx = wgt[k] & 31, fy = wgt[k] >> 8;
            int p00 = 0, p01 = 0, p10 = 0, p11 = 0;
            if (o0[k] >= 0) {
                const uint8_t* r = s + o0[k];
                if (c0[k] >= 0) p00 = remap_tap<CN>(r, c0[k]);
                if (c1[k] >= 0) p01 = remap_tap<CN>(r, c1[k]);
            }
            if (o1[k] >= 0) {
                const uint8_t* r = s + o1[k];
                if (c0[k] >= 0) p10 = remap_tap<CN>(r, c0[k]);
                if (c1[k] >= 0) p11 = remap_tap<CN>(r, c1[k]);
            }
            const int acc = (32 - fy) * ((32 - fx) * p00 + fx * p01) + fy * ((32 - fx) * p10 + fx * p11);
            out |= (unsigned)((acc * 32 + 16384) >> 15) << (8 * k);
        }
        uint8_t* d = dst + (size_t)f * n + i0;
        if (vec) *(unsigned*)d = out;
        else
            for (int k = 0; k < np; k++) d[k] = (uint8_t)(out >> (8 * k));
    }
}

void launch_remap(const uint8_t* src, int B, int sH, int sW, int cn, const float* m1, const float* m2, int H, int W,
                  uint8_t* dst, cudaStream_t st) {
    dim3 grid((unsigned)(((size_t)H * W + 256 * REMAP_PPT - 1) / (256 * REMAP_PPT)), (B + REMAP_FRAMES - 1) / REMAP_FRAMES);
    if (cn == 3) k_remap<3><<<grid, 256, 0, st>>>(src, sH, sW, m1, m2, H, W, dst, B);
    else k_remap<1><<<grid, 256, 0, st>>>(src, sH, sW, m1, m2, H, W, dst, B);
}

// ---- B.2 gray ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_bgr2gray(const uint8_t* __restrict__ img, size_t npix, uint8_t* __restrict__ dst) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    int c0 = img[3 * i], c1 = img[3 * i + 1], c2 = img[3 * i + 2];
    dst[i] = (uint8_t)((3735 * c0 + 19235 * c1 + 9798 * c2 + 16384) >> 15);
}
void launch_bgr2gray(const uint8_t* img, size_t npix, uint8_t* dst, cudaStream_t st) {
    k_bgr2gray<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(img, npix, dst);
}

// ---- 3x3 median, replicate border ------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void cswap(T& a, T& b) {
    T lo = a < b ? a : b, hi = a < b ? b : a;
    a = lo; b = hi;
}
template <typename T>
__device__ __forceinline__ T median9(T v0, T v1, T v2, T v3, T v4, T v5, T v6, T v7, T v8) {
    cswap(v1, v2); cswap(v4, v5); cswap(v7, v8);
    cswap(v0, v1); cswap(v3, v4); cswap(v6, v7);
    cswap(v1, v2); cswap(v4, v5); cswap(v7, v8);
    cswap(v0, v3); cswap(v5, v8); cswap(v4, v7);
    cswap(v3, v6); cswap(v1, v4); cswap(v2, v5);
    cswap(v4, v7); cswap(v4, v2); cswap(v6, v4);
    cswap(v4, v2);
    return v4;
}
// A CTA walks ROWS_PER_CTA rows of its 256 columns; the three taps of a source row are loaded once and roll through
// registers (3 loads per output instead of 9).
#define ROWS_PER_CTA 8
template <typename T, typename F>
__device__ __forceinline__ void median3_rows(const T* __restrict__ s, int H, int W, int x, int yb, F&& emit) {
    const int xm = max(x - 1, 0), xp = min(x + 1, W - 1);
    const T* r = s + (size_t)max(yb - 1, 0) * W;
    T a0 = r[xm], a1 = r[x], a2 = r[xp];
    r = s + (size_t)yb * W;
    T b0 = r[xm], b1 = r[x], b2 = r[xp];
    const int ye = min(yb + ROWS_PER_CTA, H);
    for (int y = yb; y < ye; y++) {
        r = s + (size_t)min(y + 1, H - 1) * W;
        const T c0 = r[xm], c1 = r[x], c2 = r[xp];
        emit(y, median9<T>(a0, a1, a2, b0, b1, b2, c0, c1, c2));
        a0 = b0; a1 = b1; a2 = b2; b0 = c0; b1 = c1; b2 = c2;
    }
}
template <typename T>
__global__ void __launch_bounds__(256) k_median3(const T* __restrict__ src, T* __restrict__ dst, int H, int W) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= W) return;
    const size_t frame = (size_t)blockIdx.z * H * W;
    median3_rows<T>(src + frame, H, W, x, blockIdx.y * ROWS_PER_CTA, [&](int y, T m) { dst[frame + (size_t)y * W + x] = m; });
}
void launch_median3_s16(const int16_t* src, int16_t* dst, int B, int H, int W, cudaStream_t st) {
    dim3 grid((W + 255) / 256, (H + ROWS_PER_CTA - 1) / ROWS_PER_CTA, B);
    k_median3<int16_t><<<grid, 256, 0, st>>>(src, dst, H, W);
}
void launch_median3_f32(const float* src, float* dst, int B, int H, int W, cudaStream_t st) {
    dim3 grid((W + 255) / 256, (H + ROWS_PER_CTA - 1) / ROWS_PER_CTA, B);
    k_median3<float><<<grid, 256, 0, st>>>(src, dst, H, W);
}

// ---- conversions ----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_d16_to_f32_crop(const int16_t* __restrict__ d16, int H, int W, int x0, float* __restrict__ out) {
    int Wc = W - x0;
    int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= Wc) return;
    size_t row = (size_t)blockIdx.z * H + blockIdx.y;
    out[row * Wc + x] = __fmul_rn((float)d16[row * W + x0 + x], 0.0625f);   // == / 16.0f exactly (power of two)
}
void launch_disp16_to_f32_crop(const int16_t* d16, int B, int H, int W, int x0, float* out, cudaStream_t st) {
    dim3 grid((W - x0 + 255) / 256, H, B);
    k_d16_to_f32_crop<<<grid, 256, 0, st>>>(d16, H, W, x0, out);
}
__global__ void __launch_bounds__(256) k_crop_s16(const int16_t* __restrict__ d16, int H, int W, int x0, int16_t* __restrict__ out) {
    const int Wc = W - x0;
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= Wc) return;
    const int yb = blockIdx.y * ROWS_PER_CTA, ye = min(yb + ROWS_PER_CTA, H);
    for (int y = yb; y < ye; y++) {
        const size_t row = (size_t)blockIdx.z * H + y;
        out[row * Wc + x] = d16[row * W + x0 + x];
    }
}
void launch_crop_s16(const int16_t* d16, int B, int H, int W, int x0, int16_t* out, cudaStream_t st) {
    dim3 grid((W - x0 + 255) / 256, (H + ROWS_PER_CTA - 1) / ROWS_PER_CTA, B);
    k_crop_s16<<<grid, 256, 0, st>>>(d16, H, W, x0, out);
}
// (d * 16).astype(int16): C truncation toward zero, then wrap to 16 bits (numpy behaviour for in-range values)
__global__ void __launch_bounds__(256) k_f32_to_s16x16(const float* __restrict__ d, size_t n, int16_t* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = (int16_t)__float2int_rz(__fmul_rn(d[i], 16.0f));
}
void launch_f32_to_s16x16(const float* d, size_t n, int16_t* out, cudaStream_t st) {
    k_f32_to_s16x16<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d, n, out);
}
__global__ void __launch_bounds__(256) k_s16_to_f32_div16(const int16_t* __restrict__ d, size_t n, float* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = __fmul_rn((float)d[i], 0.0625f);   // == / 16.0f exactly (a power of two: no rounding, no underflow for int16 inputs)
}
void launch_s16_to_f32_div16(const int16_t* d, size_t n, float* out, cudaStream_t st) {
    k_s16_to_f32_div16<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d, n, out);
}

// ---- B.3' outlier mask: 5x5 box mean of d and d^2 (REFLECT_101, double accumulation) --------------------
__device__ __forceinline__ int reflect101(int p, int n) {
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = (p < 0) ? -p : 2 * (n - 1) - p;
    return p;
}
// Separable form with the reference's exact summation order: per row the five taps are added left to right in
// double (row sum of d and of fl(d*d)), then the five row sums top to bottom; a CTA computes each row sum of its
// (64+0) x (16+4) halo tile once instead of five times. TIn = float, or int16 holding 16 * d (d = v / 16 exactly).
#define OT_W 64
#define OT_H 16
template <typename TIn> __device__ __forceinline__ float outl_val(TIn v);
template <> __device__ __forceinline__ float outl_val<float>(float v) { return v; }
template <> __device__ __forceinline__ float outl_val<int16_t>(int16_t v) { return __fmul_rn((float)v, 0.0625f); }   // == v / 16.0f exactly

template <typename TIn>
__global__ void __launch_bounds__(256) k_outliers(const TIn* __restrict__ d, float* __restrict__ out, int H, int W, float thr) {
    __shared__ float tile[OT_H + 4][OT_W + 4];
    __shared__ double rs[OT_H + 4][OT_W], rq[OT_H + 4][OT_W];
    const int x0 = blockIdx.x * OT_W, y0 = blockIdx.y * OT_H;
    const TIn* s = d + (size_t)blockIdx.z * H * W;
    for (int p = threadIdx.x; p < (OT_H + 4) * (OT_W + 4); p += 256) {
        const int ly = p / (OT_W + 4), lx = p % (OT_W + 4);
        const int y = reflect101(min(y0 + ly - 2, H + 1), H), x = reflect101(min(x0 + lx - 2, W + 1), W);
        tile[ly][lx] = outl_val<TIn>(s[(size_t)y * W + x]);
    }
    __syncthreads();
    for (int p = threadIdx.x; p < (OT_H + 4) * OT_W; p += 256) {
        const int ly = p / OT_W, lx = p % OT_W;
        double a = 0.0, q = 0.0;
#pragma unroll
        for (int dx = 0; dx < 5; dx++) {
            const float v = tile[ly][lx + dx];
            a = __dadd_rn(a, (double)v);
            q = __dadd_rn(q, (double)__fmul_rn(v, v));
        }
        rs[ly][lx] = a; rq[ly][lx] = q;
    }
    __syncthreads();
    for (int p = threadIdx.x; p < OT_H * OT_W; p += 256) {
        const int ly = p / OT_W, lx = p % OT_W, x = x0 + lx, y = y0 + ly;
        if (x >= W || y >= H) continue;
        double sum = 0.0, sumsq = 0.0;
#pragma unroll
        for (int dy = 0; dy < 5; dy++) { sum = __dadd_rn(sum, rs[ly + dy][lx]); sumsq = __dadd_rn(sumsq, rq[ly + dy][lx]); }
        const float mean = (float)__dmul_rn(sum, 1.0 / 25.0);
        const float msq = (float)__dmul_rn(sumsq, 1.0 / 25.0);
        const float var = __fsub_rn(msq, __fmul_rn(mean, mean));
        const float sd = __fsqrt_rn(fmaxf(var, 0.0f));
        const float v = tile[ly + 2][lx + 2];
        const float diff = fabsf(__fsub_rn(v, mean));
        const bool outl = (diff > __fmul_rn(thr, sd)) && (v > 0.0f);
        out[(size_t)blockIdx.z * H * W + (size_t)y * W + x] = outl ? 0.0f : v;
    }
}
void launch_outliers(const float* d, float* out, int B, int H, int W, float thr, cudaStream_t st) {
    dim3 grid((W + OT_W - 1) / OT_W, (H + OT_H - 1) / OT_H, B);
    k_outliers<float><<<grid, 256, 0, st>>>(d, out, H, W, thr);
}
void launch_outliers_s16(const int16_t* d16, float* out, int B, int H, int W, float thr, cudaStream_t st) {
    dim3 grid((W + OT_W - 1) / OT_W, (H + OT_H - 1) / OT_H, B);
    k_outliers<int16_t><<<grid, 256, 0, st>>>(d16, out, H, W, thr);
}

// ---- B.4 depth ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_depth(const float* __restrict__ disp, size_t n, float fB, float doffs, float eps,
                                               float maxDepth, int hasMax, float* __restrict__ Z) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float adj = __fadd_rn(disp[i], doffs);
    float z = __int_as_float(0x7f800000);
    if (adj > eps) z = __fdiv_rn(fB, adj);
    if (hasMax && z > maxDepth) z = maxDepth;
    Z[i] = z;
}
// 3x3 median of the float disparity and, from its result, the depth of the same pixel in one pass (pipeline tail)
__global__ void __launch_bounds__(256) k_median3_f32_depth(const float* __restrict__ src, float* __restrict__ dst, int H, int W,
                                                           float fB, float doffs, float eps, float maxDepth, int hasMax,
                                                           float* __restrict__ Z) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= W) return;
    const size_t frame = (size_t)blockIdx.z * H * W;
    median3_rows<float>(src + frame, H, W, x, blockIdx.y * ROWS_PER_CTA, [&](int y, float m) {
        const size_t i = frame + (size_t)y * W + x;
        dst[i] = m;
        float adj = __fadd_rn(m, doffs);
        float z = __int_as_float(0x7f800000);
        if (adj > eps) z = __fdiv_rn(fB, adj);
        if (hasMax && z > maxDepth) z = maxDepth;
        Z[i] = z;
    });
}
void launch_median3_f32_depth(const float* src, float* dst, int B, int H, int W, const B200sDepthParams& dp, float* Z, cudaStream_t st) {
    dim3 grid((W + 255) / 256, (H + ROWS_PER_CTA - 1) / ROWS_PER_CTA, B);
    k_median3_f32_depth<<<grid, 256, 0, st>>>(src, dst, H, W, dp.fB, dp.doffs, dp.eps, dp.maxDepth, dp.hasMaxDepth, Z);
}
void launch_depth(const float* disp, size_t n, const B200sDepthParams& dp, float* Z, cudaStream_t st) {
    k_depth<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(disp, n, dp.fB, dp.doffs, dp.eps, dp.maxDepth, dp.hasMaxDepth, Z);
}

// ---- debug tap: strip the Dp padding (and add P2 back for C) ------------------------------------------------------
__global__ void __launch_bounds__(256) k_debug_volume(const uint16_t* __restrict__ V, size_t npx, int Dp, int D, int add,
                                                      int16_t* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npx * D) return;
    size_t p = i / D;
    int d = (int)(i % D);
    out[i] = (int16_t)((int)V[p * Dp + d] + add);
}
void launch_debug_volume(const uint16_t* V, int H, int W1, int Dp, int D, int add, int16_t* out, cudaStream_t st) {
    size_t npx = (size_t)H * W1;
    k_debug_volume<<<(unsigned)((npx * D + 255) / 256), 256, 0, st>>>(V, npx, Dp, D, add, out);
}

// ---- input preparation: cv2.resize(..., INTER_AREA), shrinking (depthlib/input.py:39-43, threaded_stereo.py:41-46) --
// Integer factors ("area fast"): 2 x 2 -> (sum + 2) >> 2, else rint(sum * (float)(1 / area)). Otherwise the weighted
// tables built on the host (double precision, as the reference's dependency does): row value = sum_x S * alpha in
// float32 in table order, pixel = sum_y beta * row value, rint, clamp. No FMA anywhere (--fmad=false + explicit _rn).
__global__ void __launch_bounds__(256) k_resize_area_int(const uint8_t* __restrict__ src, int sW, int cn, int ix, int iy,
                                                         float scale, uint8_t* __restrict__ dst, int dH, int dW, size_t sframe) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)dH * dW * cn) return;
    src += (size_t)blockIdx.y * sframe;                      // blockIdx.y = frame of a dense batch
    dst += (size_t)blockIdx.y * dH * dW * cn;
    const int c = (int)(i % cn), x = (int)((i / cn) % dW), y = (int)(i / ((size_t)cn * dW));
    int sum = 0;
    for (int j = 0; j < iy; j++) {
        const uint8_t* r = src + ((size_t)(y * iy + j) * sW + (size_t)x * ix) * cn + c;
        for (int k = 0; k < ix; k++) sum += r[(size_t)k * cn];
    }
    int v = (ix == 2 && iy == 2) ? (sum + 2) >> 2 : __float2int_rn(__fmul_rn((float)sum, scale));
    dst[i] = (uint8_t)min(max(v, 0), 255);
}
__global__ void __launch_bounds__(256) k_resize_area_tab(const uint8_t* __restrict__ src, int sW, int cn,
                                                         const int* __restrict__ xo, const int2* __restrict__ xt,
                                                         const int* __restrict__ yo, const int2* __restrict__ yt,
                                                         uint8_t* __restrict__ dst, int dH, int dW, size_t sframe) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)dH * dW * cn) return;
    src += (size_t)blockIdx.y * sframe;
    dst += (size_t)blockIdx.y * dH * dW * cn;
    const int c = (int)(i % cn), x = (int)((i / cn) % dW), y = (int)(i / ((size_t)cn * dW));
    const int x0 = xo[x], x1 = xo[x + 1], y0 = yo[y], y1 = yo[y + 1];
    float sum = 0.f;
    for (int j = y0; j < y1; j++) {
        const int2 ty = yt[j];                               // {source row, weight bits}
        const uint8_t* r = src + (size_t)ty.x * sW * cn + c;
        float buf = 0.f;
        for (int k = x0; k < x1; k++) {
            const int2 tx = xt[k];
            buf = __fadd_rn(buf, __fmul_rn((float)r[(size_t)tx.x * cn], __int_as_float(tx.y)));
        }
        const float t = __fmul_rn(__int_as_float(ty.y), buf);
        sum = (j == y0) ? t : __fadd_rn(sum, t);
    }
    dst[i] = (uint8_t)min(max(__float2int_rn(sum), 0), 255);
}
void launch_resize_area_int(const uint8_t* src, int B, int sH, int sW, int cn, int ix, int iy, uint8_t* dst, int dH, int dW,
                            cudaStream_t st) {
    const size_t n = (size_t)dH * dW * cn;
    dim3 grid((unsigned)((n + 255) / 256), B);
    k_resize_area_int<<<grid, 256, 0, st>>>(src, sW, cn, ix, iy, (float)(1.0 / (ix * iy)), dst, dH, dW, (size_t)sH * sW * cn);
}
void launch_resize_area_tab(const uint8_t* src, int B, int sH, int sW, int cn, const int* xo, const int2* xt, const int* yo,
                            const int2* yt, uint8_t* dst, int dH, int dW, cudaStream_t st) {
    const size_t n = (size_t)dH * dW * cn;
    dim3 grid((unsigned)((n + 255) / 256), B);
    k_resize_area_tab<<<grid, 256, 0, st>>>(src, sW, cn, xo, xt, yo, yt, dst, dH, dW, (size_t)sH * sW * cn);
}

// ---- fill_holes(method='nearest'): one dilation by the k x k ellipse, applied under the hole mask -----------------------
// out = mask ? max over the element (anchored at its centre, pixels outside the image ignored) : in. The element is
// given per row as its half-width (EllipseRows::ext, -1 = empty row); even sizes are clipped on the right (k - 1 - r).
struct EllipseRows { int k; int ext[31]; };
__global__ void __launch_bounds__(256) k_dilate_masked(const float* __restrict__ in, const uint8_t* __restrict__ mask, int H, int W,
                                                       EllipseRows el, float* __restrict__ out) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= W) return;
    const int r = el.k / 2;
    const int yb = blockIdx.y * ROWS_PER_CTA, ye = min(yb + ROWS_PER_CTA, H);
    for (int y = yb; y < ye; y++) {
        const size_t i0 = (size_t)y * W + x;
        float v = in[i0];
        if (mask[i0]) {
            bool have = false;
            float m = 0.0f;
            for (int i = 0; i < el.k; i++) {
                const int yy = y + i - r, e = el.ext[i];
                if (e < 0 || yy < 0 || yy >= H) continue;
                const int x1 = min(x + min(e, el.k - 1 - r), W - 1);
                for (int xx = max(x - e, 0); xx <= x1; xx++) {
                    const float u = in[(size_t)yy * W + xx];
                    if (!have || u > m) { m = u; have = true; }
                }
            }
            if (have) v = m;
        }
        out[i0] = v;
    }
}
int launch_fill_nearest(float* a, float* b, const uint8_t* mask, int H, int W, int k, cudaStream_t st) {
    if (k < 1 || k > 31) return -1;
    EllipseRows el;
    el.k = k;
    const int r = k / 2;
    const double inv_r2 = r ? 1.0 / ((double)r * r) : 0.0;
    for (int i = 0; i < 31; i++) {
        const int dy = i - r;
        el.ext[i] = (i < k && abs(dy) <= r) ? (int)lrint(r * sqrt((r * r - dy * dy) * inv_r2)) : -1;
    }
    dim3 grid((W + 255) / 256, (H + ROWS_PER_CTA - 1) / ROWS_PER_CTA, 1);
    for (int it = 0; it < k; it++) {       // ping-pong a -> b -> a ...; the result of k passes is in (k odd ? b : a)
        k_dilate_masked<<<grid, 256, 0, st>>>((it & 1) ? b : a, mask, H, W, el, (it & 1) ? a : b);
    }
    return k;
}

// ---- cv2.resize(INTER_LINEAR) for uint8: 11-bit fixed-point bilinear, per-axis tap tables built on the host ----------
// tap = {source index, coefficient of it, coefficient of the next one}; x indices are already snapped to the border,
// y indices are clipped here with their weights kept (that is how cv2 treats the two axes, see the oracle).
__global__ void __launch_bounds__(256) k_resize_linear(const uint8_t* __restrict__ src, int sH, int sW, int cn,
                                                       const int4* __restrict__ tx, const int4* __restrict__ ty,
                                                       uint8_t* __restrict__ dst, int dH, int dW) {
    const size_t n = (size_t)dH * dW * cn;
    src += (size_t)blockIdx.y * sH * sW * cn;                // blockIdx.y = frame of a dense batch
    dst += (size_t)blockIdx.y * n;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const size_t i = ((size_t)blockIdx.x * 4 + k) * 256 + threadIdx.x;
        if (i >= n) return;
        const int c = (int)(i % cn), x = (int)((i / cn) % dW), y = (int)(i / ((size_t)cn * dW));
        const int4 ax = tx[x], ay = ty[y];
        const int x0 = ax.x, x1 = min(x0 + 1, sW - 1);
        const int y0 = min(max(ay.x, 0), sH - 1), y1 = min(max(ay.x + 1, 0), sH - 1);
        const uint8_t* r0 = src + (size_t)y0 * sW * cn + c;
        const uint8_t* r1 = src + (size_t)y1 * sW * cn + c;
        const int h0 = r0[(size_t)x0 * cn] * ax.y + r0[(size_t)x1 * cn] * ax.z;
        const int h1 = r1[(size_t)x0 * cn] * ax.y + r1[(size_t)x1 * cn] * ax.z;
        const int v = (((ay.y * (h0 >> 4)) >> 16) + ((ay.z * (h1 >> 4)) >> 16) + 2) >> 2;
        dst[i] = (uint8_t)min(max(v, 0), 255);
    }
}
void launch_resize_linear(const uint8_t* src, int B, int sH, int sW, int cn, const int4* tx, const int4* ty, uint8_t* dst, int dH,
                          int dW, cudaStream_t st) {
    const size_t n = (size_t)dH * dW * cn;
    dim3 grid((unsigned)((n + 1023) / 1024), B);
    k_resize_linear<<<grid, 256, 0, st>>>(src, sH, sW, cn, tx, ty, dst, dH, dW);
}

// ---- cv2.initUndistortRectifyMap(K, dist, R, P, size, CV_32FC1) (depthlib/rectify.py:68-73, 229-234) --------------------
// Per pixel, in double precision with the operation order of the reference's dependency (no FMA contraction): the ray
// of rectified pixel (j, i) through iR = (P[:, :3] R)^-1, the distortion model (k1..k6, p1, p2, s1..s4, tilt), the
// camera matrix; float32 maps out. p = {iR[9], fx, fy, u0, v0, k1, k2, p1, p2, k3, k4, k5, k6, s1, s2, s3, s4, tilt[9]}.
struct MapParams { double ir[9], fx, fy, u0, v0, k1, k2, p1, p2, k3, k4, k5, k6, s1, s2, s3, s4, tilt[9]; };
#define DM(a, b) __dmul_rn(a, b)
#define DA(a, b) __dadd_rn(a, b)
__global__ void __launch_bounds__(256) k_init_undistort_rectify_map(const MapParams p, int H, int W, float* __restrict__ m1,
                                                                    float* __restrict__ m2) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)H * W) return;
    const double i = (double)(int)(idx / W), j = (double)(int)(idx % W);
    const double _x = DA(DA(DM(i, p.ir[1]), p.ir[2]), DM(j, p.ir[0]));
    const double _y = DA(DA(DM(i, p.ir[4]), p.ir[5]), DM(j, p.ir[3]));
    const double _w = DA(DA(DM(i, p.ir[7]), p.ir[8]), DM(j, p.ir[6]));
    const double w = __ddiv_rn(1.0, _w), x = DM(_x, w), y = DM(_y, w);
    const double x2 = DM(x, x), y2 = DM(y, y), r2 = DA(x2, y2), _2xy = DM(DM(2.0, x), y);
    const double num = DA(1.0, DM(DA(DM(DA(DM(p.k3, r2), p.k2), r2), p.k1), r2));
    const double den = DA(1.0, DM(DA(DM(DA(DM(p.k6, r2), p.k5), r2), p.k4), r2));
    const double kr = __ddiv_rn(num, den);
    const double xd = DA(DA(DA(DA(DM(x, kr), DM(p.p1, _2xy)), DM(p.p2, DA(r2, DM(2.0, x2)))), DM(p.s1, r2)), DM(DM(p.s2, r2), r2));
    const double yd = DA(DA(DA(DA(DM(y, kr), DM(p.p1, DA(r2, DM(2.0, y2)))), DM(p.p2, _2xy)), DM(p.s3, r2)), DM(DM(p.s4, r2), r2));
    const double t0 = DA(DA(DM(p.tilt[0], xd), DM(p.tilt[1], yd)), p.tilt[2]);
    const double t1 = DA(DA(DM(p.tilt[3], xd), DM(p.tilt[4], yd)), p.tilt[5]);
    const double t2 = DA(DA(DM(p.tilt[6], xd), DM(p.tilt[7], yd)), p.tilt[8]);
    const double inv = (t2 != 0.0) ? __ddiv_rn(1.0, t2) : 1.0;
    m1[idx] = (float)DA(DM(DM(p.fx, inv), t0), p.u0);
    m2[idx] = (float)DA(DM(DM(p.fy, inv), t1), p.v0);
}
void launch_init_undistort_rectify_map(const double* params34, int H, int W, float* m1, float* m2, cudaStream_t st) {
    MapParams p;
    static_assert(sizeof(MapParams) == 34 * sizeof(double), "parameter block");
    memcpy(&p, params34, sizeof(p));
    const size_t n = (size_t)H * W;
    k_init_undistort_rectify_map<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(p, H, W, m1, m2);
}
