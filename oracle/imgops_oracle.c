/*
 * oracle/imgops_oracle.c — TEST INFRASTRUCTURE ONLY (never on the product path).
 *
 * Scalar CPU restatement of the O(N) image ops on depthlib's stereo hot path:
 *   orc_remap_u8       cv2.remap(gray, map1, map2, INTER_LINEAR)      depthlib/rectify.py:185-186
 *   orc_bgr2gray       cv2.cvtColor(img, COLOR_BGR2GRAY)              depthlib/rectify.py:116, stereo_core.py:157-159
 *   orc_median3_f32    cv2.medianBlur(float32, 3)                     depthlib/postprocess.py:169, stereo_core.py:173
 *   orc_box5_mean      cv2.boxFilter(f32, -1, (5,5))                  depthlib/postprocess.py:59,63
 *   orc_postprocess    postprocess_disparity (hole filling off)       depthlib/postprocess.py:120-171
 *   orc_disp_to_depth  StereoCore.disparity_to_depth                  depthlib/stereo_core.py:234-272
 *   orc_resize_linear  cv2.resize(img, size, INTER_LINEAR), uint8 (the size fix-up)   depthlib/rectify.py:92-105
 *   orc_fill_nearest   fill_holes(method='nearest'): k x dilate(ellipse k x k) under the hole mask   depthlib/postprocess.py:106-116
 * The arithmetic of the cv2 calls lives in the opencv-python wheel (not under /root/reference); it is
 * restated from SURVEY.md Appendix B and pinned against the live cv2 binary by tests/test_oracle_vs_cv2.py.
 * Compile with -ffp-contract=off: the float32 chain must not be fused into FMAs.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

void orc_filter_speckles(int16_t *img, int H, int W, int newVal, int maxSize, int maxDiff);

/* B.1: fixed-point bilinear gather, 5 fractional bits, constant-0 border */
void orc_remap_u8(const uint8_t *src, int sH, int sW, const float *map1, const float *map2,
                  int H, int W, uint8_t *dst)
{
    for (size_t i = 0; i < (size_t)H * W; i++) {
        int sx = (int)lrintf(map1[i] * 32.0f); /* round-half-even of the float32 product */
        int sy = (int)lrintf(map2[i] * 32.0f);
        int ix = sx >> 5, iy = sy >> 5, fx = sx & 31, fy = sy & 31;
        int p00 = 0, p01 = 0, p10 = 0, p11 = 0;
        if (iy >= 0 && iy < sH) {
            if (ix >= 0 && ix < sW) p00 = src[(size_t)iy * sW + ix];
            if (ix + 1 >= 0 && ix + 1 < sW) p01 = src[(size_t)iy * sW + ix + 1];
        }
        if (iy + 1 >= 0 && iy + 1 < sH) {
            if (ix >= 0 && ix < sW) p10 = src[(size_t)(iy + 1) * sW + ix];
            if (ix + 1 >= 0 && ix + 1 < sW) p11 = src[(size_t)(iy + 1) * sW + ix + 1];
        }
        int acc = (32 - fy) * (32 - fx) * p00 + (32 - fy) * fx * p01 + fy * (32 - fx) * p10 + fy * fx * p11;
        dst[i] = (uint8_t)((acc * 32 + 16384) >> 15);
    }
}

/* B.2: weights applied to the channels as stored */
void orc_bgr2gray(const uint8_t *img, size_t npix, uint8_t *dst)
{
    for (size_t i = 0; i < npix; i++)
        dst[i] = (uint8_t)((3735 * img[3 * i] + 19235 * img[3 * i + 1] + 9798 * img[3 * i + 2] + 16384) >> 15);
}

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

static inline void fswap(float *a, float *b) { if (*a > *b) { float t = *a; *a = *b; *b = t; } }

void orc_median3_f32(const float *src, int H, int W, float *dst)
{
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            float v[9];
            int k = 0;
            for (int dy = -1; dy <= 1; dy++)
                for (int dx = -1; dx <= 1; dx++)
                    v[k++] = src[(size_t)clampi(y + dy, 0, H - 1) * W + clampi(x + dx, 0, W - 1)];
            for (int i = 1; i < 9; i++) { /* insertion sort: 9 elements */
                float t = v[i];
                int j = i - 1;
                while (j >= 0 && v[j] > t) { v[j + 1] = v[j]; j--; }
                v[j + 1] = t;
            }
            dst[(size_t)y * W + x] = v[4];
        }
}

static inline int reflect101(int p, int n)
{
    if (n == 1) return 0;
    while (p < 0 || p >= n) {
        if (p < 0) p = -p;
        else p = 2 * (n - 1) - p;
    }
    return p;
}

/* 5x5 normalised box mean, BORDER_REFLECT_101, double accumulation, one rounding to float32 */
void orc_box5_mean(const float *src, int H, int W, float *dst)
{
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            double s = 0.0;
            for (int dy = -2; dy <= 2; dy++)
                for (int dx = -2; dx <= 2; dx++)
                    s += (double)src[(size_t)reflect101(y + dy, H) * W + reflect101(x + dx, W)];
            dst[(size_t)y * W + x] = (float)(s * (1.0 / 25.0));
        }
}

/* postprocess.postprocess_disparity with apply_hole_filling=False */
void orc_postprocess(const float *disp, int H, int W, int maxSpeckle, int maxDiff16, float thr,
                     int outliers, float *out)
{
    size_t N = (size_t)H * W;
    int16_t *d16 = malloc(N * sizeof(int16_t));
    float *d = malloc(N * sizeof(float)), *sq = malloc(N * sizeof(float));
    float *mean = malloc(N * sizeof(float)), *msq = malloc(N * sizeof(float));
    for (size_t i = 0; i < N; i++) d16[i] = (int16_t)(int32_t)(disp[i] * 16.0f); /* astype(int16) truncates */
    orc_filter_speckles(d16, H, W, 0, maxSpeckle, maxDiff16);
    for (size_t i = 0; i < N; i++) d[i] = (float)d16[i] / 16.0f;
    if (outliers) {
        for (size_t i = 0; i < N; i++) sq[i] = d[i] * d[i];
        orc_box5_mean(d, H, W, mean);
        orc_box5_mean(sq, H, W, msq);
        for (size_t i = 0; i < N; i++) {
            float m2 = mean[i] * mean[i];
            float var = msq[i] - m2;
            float sd = sqrtf(var > 0.0f ? var : 0.0f);
            float diff = fabsf(d[i] - mean[i]);
            float lim = thr * sd;
            if (diff > lim && d[i] > 0.0f) sq[i] = 0.0f; else sq[i] = d[i];
        }
        memcpy(d, sq, N * sizeof(float));
    }
    orc_median3_f32(d, H, W, out);
    free(d16); free(d); free(sq); free(mean); free(msq);
}

/* B.4 */
void orc_disp_to_depth(const float *disp, size_t n, float fB, float doffs, float eps,
                       float max_depth, int has_max, float *Z)
{
    for (size_t i = 0; i < n; i++) {
        float adj = disp[i] + doffs;
        float z = INFINITY;
        if (adj > eps) z = fB / adj;
        if (has_max && z > max_depth) z = max_depth;
        Z[i] = z;
    }
}

/* cv2.resize(src, (dW, dH), interpolation=INTER_AREA), uint8, cn interleaved channels, shrinking only
 * (depthlib/input.py:39-43, threaded_stereo.py:41-46). Restated from the published OpenCV algorithm (imgproc resize:
 * "area fast" for integer factors, DecimateAlpha tables otherwise) and pinned against the live binary by
 * tests/test_oracle_vs_cv2.py:
 *   both factors integer, 2 x 2:   (a + b + c + d + 2) >> 2
 *   both factors integer, other:   rint( sum * (float)(1.0 / area) )               (float32 product, ties to even)
 *   otherwise: per axis a table of (source index, float weight): with scale = ssize / dsize (double),
 *     f1 = d * scale, f2 = f1 + scale, cell = min(scale, ssize - f1), s1 = ceil(f1), s2 = min(floor(f2), ssize - 1),
 *     s1 = min(s1, s2); [s1 - 1: (s1 - f1) / cell if s1 - f1 > 1e-3], [s1 .. s2 - 1: 1 / cell],
 *     [s2: min(min(f2 - s2, 1), cell) / cell if f2 - s2 > 1e-3]; row value = sum_x S * alpha accumulated in float32 in
 *     table order, pixel = sum_y beta * row value (first term assigned, the rest added), then rint and clamp. */
typedef struct { int si; float w; } OrcAreaTap;

static int orc_area_tab(int ssize, int dsize, OrcAreaTap *tab, int *ofs)
{
    double scale = (double)ssize / dsize;
    int k = 0;
    for (int d = 0; d < dsize; d++) {
        double f1 = d * scale, f2 = f1 + scale;
        double cell = scale < ssize - f1 ? scale : ssize - f1;
        int s1 = (int)ceil(f1), s2 = (int)floor(f2);
        if (s2 > ssize - 1) s2 = ssize - 1;
        if (s1 > s2) s1 = s2;
        ofs[d] = k;
        if (s1 - f1 > 1e-3) { tab[k].si = s1 - 1; tab[k].w = (float)((s1 - f1) / cell); k++; }
        for (int s = s1; s < s2; s++) { tab[k].si = s; tab[k].w = (float)(1.0 / cell); k++; }
        if (f2 - s2 > 1e-3) {
            double t = f2 - s2; if (t > 1.0) t = 1.0; if (t > cell) t = cell;
            tab[k].si = s2; tab[k].w = (float)(t / cell); k++;
        }
    }
    ofs[dsize] = k;
    return k;
}

int orc_resize_area_u8(const uint8_t *src, int sH, int sW, int cn, uint8_t *dst, int dH, int dW)
{
    if (dH > sH || dW > sW || dH <= 0 || dW <= 0 || cn <= 0) return -1;
    const int ix = sW / dW, iy = sH / dH;
    if (ix * dW == sW && iy * dH == sH) {             /* integer factors: "area fast" */
        const int area = ix * iy;
        const float scale = (float)(1.0 / area);
        for (int y = 0; y < dH; y++)
            for (int x = 0; x < dW; x++)
                for (int c = 0; c < cn; c++) {
                    int sum = 0;
                    for (int j = 0; j < iy; j++)
                        for (int i = 0; i < ix; i++) sum += src[((size_t)(y * iy + j) * sW + x * ix + i) * cn + c];
                    int v;
                    if (ix == 2 && iy == 2) v = (sum + 2) >> 2;
                    else { float f = (float)sum * scale; v = (int)lrintf(f); }
                    dst[((size_t)y * dW + x) * cn + c] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
                }
        return 0;
    }
    OrcAreaTap *xt = (OrcAreaTap *)malloc(sizeof(OrcAreaTap) * (size_t)(sW + 2 * dW + 2));
    OrcAreaTap *yt = (OrcAreaTap *)malloc(sizeof(OrcAreaTap) * (size_t)(sH + 2 * dH + 2));
    int *xo = (int *)malloc(sizeof(int) * (size_t)(dW + 1)), *yo = (int *)malloc(sizeof(int) * (size_t)(dH + 1));
    orc_area_tab(sW, dW, xt, xo);
    orc_area_tab(sH, dH, yt, yo);
    for (int y = 0; y < dH; y++)
        for (int x = 0; x < dW; x++)
            for (int c = 0; c < cn; c++) {
                float sum = 0.f;
                for (int j = yo[y]; j < yo[y + 1]; j++) {
                    float buf = 0.f;
                    for (int i = xo[x]; i < xo[x + 1]; i++)
                        buf = buf + (float)src[((size_t)yt[j].si * sW + xt[i].si) * cn + c] * xt[i].w;
                    float t = yt[j].w * buf;
                    sum = (j == yo[y]) ? t : sum + t;
                }
                long v = lrintf(sum);
                dst[((size_t)y * dW + x) * cn + c] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
            }
    free(xt); free(yt); free(xo); free(yo);
    return 0;
}

/* ---- fill_holes(method='nearest') (depthlib/postprocess.py:106-116) -------------------------------------------
 * kernel = cv2.getStructuringElement(MORPH_ELLIPSE, (k, k)); k times: filled = where(mask, cv2.dilate(filled, kernel), filled).
 * The ellipse rows follow OpenCV's getStructuringElement: with r = c = k / 2, row i (dy = i - r) holds ones in
 * [c - dx, min(c + dx, k - 1)] with dx = round(c * sqrt((r*r - dy*dy) / (r*r))) (k = 1: the single pixel). cv2.dilate = maximum over
 * the element anchored at its centre; pixels outside the image do not take part. (The distance transform the reference
 * also computes there is never used.) ext[i] = dx of row i, or -1 for an empty row. */
int orc_ellipse_rows(int k, int *ext)
{
    if (k < 1) return -1;
    int r = k / 2, c = k / 2;
    double inv_r2 = r ? 1.0 / ((double)r * r) : 0.0;
    for (int i = 0; i < k; i++) {
        int dy = i - r;
        ext[i] = -1;
        if (abs(dy) <= r) ext[i] = (int)lrint(c * sqrt((r * r - dy * dy) * inv_r2));
    }
    return 0;
}

int orc_fill_nearest_f32(float *img, const uint8_t *mask, int H, int W, int k)
{
    if (k < 1 || k > 63) return -1;
    int ext[64];
    orc_ellipse_rows(k, ext);
    size_t n = (size_t)H * W;
    float *tmp = (float *)malloc(n * sizeof(float));
    if (!tmp) return -2;
    int r = k / 2;
    for (int it = 0; it < k; it++) {
        for (int y = 0; y < H; y++)
            for (int x = 0; x < W; x++) {
                float v = img[(size_t)y * W + x];
                if (mask[(size_t)y * W + x]) {
                    int have = 0;
                    float m = 0.0f;
                    for (int i = 0; i < k; i++) {
                        int yy = y + i - r;
                        if (ext[i] < 0 || yy < 0 || yy >= H) continue;
                        const int right = ext[i] < k - 1 - r ? ext[i] : k - 1 - r;   /* even k: the element is clipped on the right */
                        for (int xx = x - ext[i]; xx <= x + right; xx++) {
                            if (xx < 0 || xx >= W) continue;
                            float u = img[(size_t)yy * W + xx];
                            if (!have || u > m) { m = u; have = 1; }
                        }
                    }
                    if (have) v = m;
                }
                tmp[(size_t)y * W + x] = v;
            }
        memcpy(img, tmp, n * sizeof(float));
    }
    free(tmp);
    return 0;
}

/* ---- cv2.resize(img, (dW, dH), interpolation=cv2.INTER_LINEAR) for uint8 (depthlib/rectify.py:92-105) ---------------
 * OpenCV's fixed-point bilinear: per axis f = (float)((d + 0.5) * scale - 0.5), s = floor(f), f -= s, 11-bit
 * coefficients cvRound((1 - f) * 2048), cvRound(f * 2048). Along x an index outside [0, sW-1) snaps to the border
 * with f = 0; along y only the row INDEX is clipped (the weights stay), which rounds differently at the border.
 * Horizontal pass in int (no shift), vertical pass ((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2. */
typedef struct { int ofs; int c0, c1; } OrcLinTap;
static int orc_round_half_even(float v) { return (int)lrintf(v); }
void orc_linear_tab(int ssize, int dsize, int snap, OrcLinTap *tab)
{
    const double scale = 1.0 / ((double)dsize / ssize);
    for (int d = 0; d < dsize; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= (float)s;
        if (snap) {
            if (s < 0) { f = 0.f; s = 0; }
            if (s >= ssize - 1) { f = 0.f; s = ssize - 1; }
        }
        int c0 = orc_round_half_even((1.f - f) * 2048.f), c1 = orc_round_half_even(f * 2048.f);
        if (c0 > 32767) c0 = 32767;
        if (c1 > 32767) c1 = 32767;
        tab[d].ofs = s; tab[d].c0 = c0; tab[d].c1 = c1;
    }
}
int orc_resize_linear_u8(const uint8_t *src, int sH, int sW, int cn, uint8_t *dst, int dH, int dW)
{
    if (sH <= 0 || sW <= 0 || dH <= 0 || dW <= 0 || cn < 1) return -1;
    OrcLinTap *tx = (OrcLinTap *)malloc(sizeof(OrcLinTap) * (size_t)dW), *ty = (OrcLinTap *)malloc(sizeof(OrcLinTap) * (size_t)dH);
    if (!tx || !ty) { free(tx); free(ty); return -2; }
    orc_linear_tab(sW, dW, 1, tx);
    orc_linear_tab(sH, dH, 0, ty);
    for (int y = 0; y < dH; y++) {
        const int y0 = clampi(ty[y].ofs, 0, sH - 1), y1 = clampi(ty[y].ofs + 1, 0, sH - 1);
        for (int x = 0; x < dW; x++) {
            const int x0 = tx[x].ofs, x1 = x0 + 1 < sW ? x0 + 1 : sW - 1;
            for (int c = 0; c < cn; c++) {
                const int h0 = src[((size_t)y0 * sW + x0) * cn + c] * tx[x].c0 + src[((size_t)y0 * sW + x1) * cn + c] * tx[x].c1;
                const int h1 = src[((size_t)y1 * sW + x0) * cn + c] * tx[x].c0 + src[((size_t)y1 * sW + x1) * cn + c] * tx[x].c1;
                int v = (((ty[y].c0 * (h0 >> 4)) >> 16) + ((ty[y].c1 * (h1 >> 4)) >> 16) + 2) >> 2;
                dst[((size_t)y * dW + x) * cn + c] = (uint8_t)clampi(v, 0, 255);
            }
        }
    }
    free(tx); free(ty);
    return 0;
}
