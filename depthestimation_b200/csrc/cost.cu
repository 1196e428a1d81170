// cost.cu — prefilter (A.1), Birchfield–Tomasi pixel cost (A.2) and bs x bs block sum (A.3).
//
// Replaces the cost-volume part of cv2.StereoSGBM.compute as called at
// depthlib/stereo_core.py:231. Output: C[b][y][xc][d] (u16, d fastest, pitch Dp), WITHOUT the
// reference's +P2 term (the path recursion is rewritten so it is never needed, see paths.cu).
//
// Kernel k_cost: one CTA = a strip of XT cost columns x all disparities, marching down a row
// segment. Thread = one disparity pair (packed 16x2). Per source row the thread slides along x
// with a register ring of the last BS pixel costs (horizontal box sum), and keeps XT running
// vertical sums in registers, fed by a shared-memory ring of the last BS horizontal-sum rows.
// Everything is packed-16x2 integer SIMD (VIADD.16x2 / VIADDMNMX.S16x2.RELU / VIMNMX.S16x2).
#include <stdlib.h>

#include "common.cuh"

// ------------------------------------------------------------------------------------------------
// prefilter (A.1): Sobel-x clipped to [-ftzero, ftzero] + ftzero, raw sample, and their Birchfield-Tomasi half-sample
// intervals -> 8-byte record {g, g_lo, g_hi, r, r_lo, r_hi, 0, 0}; border columns hold ftzero in both channels
// ------------------------------------------------------------------------------------------------
// Four pixels per thread: the Sobel response and the raw sample are needed at x-1 .. x+4 (six columns, eight source
// columns x-2 .. x+5 per row), each evaluated once instead of three times.
#if COST_PART <= 0
#define PF_ROWS 8   // rows per CTA (one row per CTA made a batch ~170 000 tiny CTAs: CTA launch rate bound)
__global__ void __launch_bounds__(128) k_prefilter(const uint8_t* __restrict__ L, const uint8_t* __restrict__ R,
                                                   int H, int W, int ftzero, uint2* __restrict__ recL,
                                                   uint2* __restrict__ recR) {
    const int x0 = 4 * (blockIdx.x * blockDim.x + threadIdx.x);
    const int z = blockIdx.z;  // 2*b + image
    if (x0 >= W) return;
    const int b = z >> 1;
    const uint8_t* img = ((z & 1) ? R : L) + (size_t)b * H * W;
    uint2* rec = ((z & 1) ? recR : recL) + (size_t)b * H * W;
    int xx[8];                 // source columns x0-2 .. x0+5, clamped reads (clamped values are never used: borders give ftzero)
#pragma unroll
    for (int k = 0; k < 8; k++) xx[k] = min(max(x0 - 2 + k, 0), W - 1);
    const bool vec = (W & 1) == 0;   // rows (and frames) of 8-byte records then start on 16-byte boundaries
    const int yb = blockIdx.y * PF_ROWS, ye = min(yb + PF_ROWS, H);
    // the three source rows roll through registers: one new row of eight bytes per output row
    int nn[8], cc[8], ss[8];
    {
        const uint8_t* n = img + (size_t)max(yb - 1, 0) * W;
        const uint8_t* c = img + (size_t)yb * W;
#pragma unroll
        for (int k = 0; k < 8; k++) { nn[k] = n[xx[k]]; cc[k] = c[xx[k]]; }
    }
    for (int y = yb; y < ye; y++) {
        const uint8_t* sp = img + (size_t)min(y + 1, H - 1) * W;
#pragma unroll
        for (int k = 0; k < 8; k++) ss[k] = sp[xx[k]];
        int g[6], r[6];            // columns x0-1 .. x0+4
#pragma unroll
        for (int k = 0; k < 6; k++) {
            const int x = x0 - 1 + k;
            const bool border = (x <= 0 || x >= W - 1);
            const int v = 2 * (cc[k + 2] - cc[k]) + (nn[k + 2] - nn[k]) + (ss[k + 2] - ss[k]);
            g[k] = border ? ftzero : min(max(v, -ftzero), ftzero) + ftzero;
            r[k] = border ? ftzero : cc[k + 1];
        }
        uint2 o[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int x = x0 + k;
            const int g0 = g[k + 1], r0 = r[k + 1];
            int glo = g0, ghi = g0, rlo = r0, rhi = r0;
            if (x > 0) {
                const int ga = (g0 + g[k]) >> 1, ra = (r0 + r[k]) >> 1;
                glo = min(glo, ga); ghi = max(ghi, ga); rlo = min(rlo, ra); rhi = max(rhi, ra);
            }
            if (x < W - 1) {
                const int gb = (g0 + g[k + 2]) >> 1, rb = (r0 + r[k + 2]) >> 1;
                glo = min(glo, gb); ghi = max(ghi, gb); rlo = min(rlo, rb); rhi = max(rhi, rb);
            }
            o[k].x = (unsigned)g0 | ((unsigned)glo << 8) | ((unsigned)ghi << 16) | ((unsigned)r0 << 24);
            o[k].y = (unsigned)rlo | ((unsigned)rhi << 8);
        }
        uint2* dst = rec + (size_t)y * W + x0;
        if (vec && x0 + 3 < W) {   // 32 contiguous bytes per thread as two 16-byte stores (whole sectors)
            ((uint4*)dst)[0] = make_uint4(o[0].x, o[0].y, o[1].x, o[1].y);
            ((uint4*)dst)[1] = make_uint4(o[2].x, o[2].y, o[3].x, o[3].y);
        } else {
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (x0 + k < W) dst[k] = o[k];
        }
#pragma unroll
        for (int k = 0; k < 8; k++) { nn[k] = cc[k]; cc[k] = ss[k]; }
    }
}

void launch_prefilter(const Eff& e, const uint8_t* L, const uint8_t* R, int B, uint2* recL, uint2* recR,
                      cudaStream_t st) {
    dim3 grid((e.W + 511) / 512, (e.H + PF_ROWS - 1) / PF_ROWS, 2 * B);
    k_prefilter<<<grid, 128, 0, st>>>(L, R, e.H, e.W, e.ftzero, recL, recR);
}

#endif  // COST_PART <= 0 (prefilter)

// ------------------------------------------------------------------------------------------------
// block cost
// ------------------------------------------------------------------------------------------------
struct CostArgs {
    const uint2* recL;
    const uint2* recR;
    uint16_t* C;
    uint16_t* Calt;
    int H, W, W1, D, Dp, minD, minX1, SH2, mode;
    int RS, nreg;           // rows per regular segment, number of regular segments
    StripePlan sp;          // alt segments: one per stripe >= 1 when sp.alt_rows > 0
};

// right-image arrays in shared memory, stored reversed in x so that increasing d walks forward
// (the negated samples -v are not stored: lo1 - v is formed as (lo1 + 1) + ~v, one LOP3 instead of an LDS + PRMT)
enum { RA_VG = 0, RA_LOG, RA_NHIG, RA_VR, RA_LOR, RA_NHIR, RA_COUNT };

#ifndef COST_MINB
#define COST_MINB 4
#endif
#ifndef COST_BIAS
#define COST_BIAS 0   // 1: negated samples are stored as (NB - v) so that the packed sums need no per-half carry
#endif                //    handling and are plain 32-bit adds (ptxas places them on the FMA pipe); 0: VIADD.16x2. Measured: 0 is 5 % faster
#define NB 0x1000u
#if COST_BIAS
#define NEG16(v) (NB - (v))
#define LO1 0u
#else
#define NEG16(v) ((0u - (v)) & 0xffffu)
#define LO1 1u   // the left interval's lower end is stored + 1 (pairs with ~v of the right sample)
#endif
#ifndef COST_XT
#define COST_XT 20          // cost columns per strip; measured on B200: 20 beats 16 (by 3 %), 24 and 32
#endif

// Compile-time geometry of one CTA. NTS = stride of the per-thread shared-memory arrays (a multiple of 16 >= Dp / 2);
// every shared-memory and (when Dp / 2 == NTS) global offset inside the slide loop is then an immediate.
template <int BS, int NTS>
struct CostCfg {
    static constexpr int XT = COST_XT;
    static constexpr int NXI = XT + BS - 1;                              // pixel-cost columns per row (with halo)
    static constexpr int RW = (((NXI + 2 * NTS + 3) / 2 + 1) + 1) & ~1;  // words per right-image array
    static constexpr int RING = BS * XT * NTS;                           // [BS][XT][NTS] horizontal sums
    static constexpr size_t SMEM = (size_t)(RING + 2 * (NXI * 8 + RA_COUNT * RW)) * 4;   // ring + two staging buffers
};

// One source row of the strip: pixel cost -> horizontal box sum -> vertical running sum V[].
// INTERIOR: no column of the strip (halo included) is clamped, so the reversed right index of column i is the
// constant NXI - 1 - i.
template <int BS, int NTS, bool INTERIOR>
__device__ __forceinline__ void cost_row(const unsigned* __restrict__ sL, const unsigned* __restrict__ sRd,
                                         unsigned* __restrict__ ringd, unsigned (&V)[COST_XT], const int xa,
                                         const int W1, const int minX1, const int x_max) {
    using G = CostCfg<BS, NTS>;
    unsigned pr[BS];
#pragma unroll
    for (int j = 0; j < BS; j++) pr[j] = 0u;
    unsigned hs = 0u;
    // right-image words: word k of an array = (half[2k], half[2k+1]); the thread needs the pair starting at half
    // e + 2 dp. Interior strips walk e = NXI-1 .. 0, so consecutive columns share a word: one LDS per array every
    // second column plus a byte permute on the odd ones (instead of two alignment copies and one LDS per column).
    unsigned wlo[RA_COUNT], whi[RA_COUNT];
#pragma unroll
    for (int i = 0; i < G::NXI; i++) {
        unsigned rv[RA_COUNT];
        if (INTERIOR) {
            const int e = G::NXI - 1 - i;                                  // compile-time after unrolling
            if (e & 1) {
                if (i == 0) {
#pragma unroll
                    for (int arr = 0; arr < RA_COUNT; arr++) whi[arr] = sRd[arr * G::RW + (e + 1) / 2];
                }
#pragma unroll
                for (int arr = 0; arr < RA_COUNT; arr++) {
                    wlo[arr] = sRd[arr * G::RW + (e - 1) / 2];
                    rv[arr] = __byte_perm(wlo[arr], whi[arr], 0x5432);     // (half[e], half[e+1])
                }
            } else {
                if (i == 0) {
#pragma unroll
                    for (int arr = 0; arr < RA_COUNT; arr++) wlo[arr] = sRd[arr * G::RW + e / 2];
                }
#pragma unroll
                for (int arr = 0; arr < RA_COUNT; arr++) { rv[arr] = wlo[arr]; whi[arr] = wlo[arr]; }
            }
        } else {
            const int xc = min(max(xa - BS / 2 + i, 0), W1 - 1);
            const int e = x_max - (xc + minX1);                            // uniform, run time
            const unsigned* base = sRd + (e >> 1);
            const unsigned sh = (unsigned)(e & 1) * 16u;
#pragma unroll
            for (int arr = 0; arr < RA_COUNT; arr++)
                rv[arr] = __funnelshift_r(base[arr * G::RW], base[arr * G::RW + 1], sh);
        }
        const uint4 l0 = *(const uint4*)(sL + i * 8);                 // {u, -u, lo1, -hi1} gradient channel
        const uint4 l1 = *(const uint4*)(sL + i * 8 + 4);             // raw channel
        const unsigned vg = rv[RA_VG], log2_ = rv[RA_LOG], nhig = rv[RA_NHIG];
        const unsigned vr = rv[RA_VR], lor = rv[RA_LOR], nhir = rv[RA_NHIR];
#if COST_BIAS
        const unsigned nvg = pk2(NB) - vg, nvr = pk2(NB) - vr;
#else
        const unsigned nvg = ~vg, nvr = ~vr;                          // -v - 1 per half; the left record holds lo1 + 1
#endif
        // c0 = max(0, u - hi2, lo2 - u); c1 = max(0, v - hi1, lo1 - v); cost = min(c0, c1)
#if COST_BIAS
        // every term carries the bias NB once: NB + (u - hi2), NB + (lo2 - u), ...; it is removed (with the clamp at 0)
        // after the min. pk2(pixel) < 256 and NB = 4096, so no half ever carries into its neighbour.
        unsigned c0 = __viaddmax_s16x2(nhig, l0.x, log2_ + l0.y);
        unsigned c1 = __viaddmax_s16x2(vg, l0.w, nvg + l0.z);
        const unsigned cg = __viaddmax_s16x2_relu(__vmins2(c0, c1), pk2(0x10000u - NB), 0u);
        c0 = __viaddmax_s16x2(nhir, l1.x, lor + l1.y);
        c1 = __viaddmax_s16x2(vr, l1.w, nvr + l1.z);
        const unsigned cr = (__viaddmax_s16x2_relu(__vmins2(c0, c1), pk2(0x10000u - NB), 0u) >> 2) & 0x3fff3fffu;
#else
        unsigned c0 = __viaddmax_s16x2_relu(nhig, l0.x, __vadd2(log2_, l0.y));
        unsigned c1 = __viaddmax_s16x2_relu(vg, l0.w, __vadd2(nvg, l0.z));
        const unsigned cg = __vmins2(c0, c1);
        c0 = __viaddmax_s16x2_relu(nhir, l1.x, __vadd2(lor, l1.y));
        c1 = __viaddmax_s16x2_relu(vr, l1.w, __vadd2(nvr, l1.z));
        const unsigned cr = (__vmins2(c0, c1) >> 2) & 0x3fff3fffu;
#endif
        const unsigned pix = cg + cr;
        hs = hs + pix - pr[i % BS];
        pr[i % BS] = pix;
        if (i >= BS - 1) {
            const int xo = i - (BS - 1);
            unsigned* rp = ringd + xo * NTS;
            const unsigned old = *rp;
            *rp = hs;
            V[xo] = V[xo] + hs - old;
        }
    }
}

template <int BS, int NTS, bool FULL>
__global__ void __launch_bounds__(NTS, COST_MINB * 4 / ((NTS + 31) / 32)) k_cost(const CostArgs a) {   // COST_MINB * 4 warps per SM
    using G = CostCfg<BS, NTS>;
    constexpr int XT = G::XT, NXI = G::NXI, RW = G::RW;
    extern __shared__ __align__(16) unsigned smem[];
    const int NT = FULL ? NTS : (int)blockDim.x;   // Dp / 2 threads, one disparity pair each
    const int dp = threadIdx.x;
    const int xa = blockIdx.x * XT;                // first cost column of the strip
    const int b = blockIdx.z;

    // ---- which rows does this CTA produce? -------------------------------------------------
    int y0, y1, ylo;
    uint16_t* out;
    {
        int seg = blockIdx.y;
        if (seg < a.nreg) {
            y0 = seg * a.RS; y1 = min(y0 + a.RS, a.H); ylo = 0;
            out = a.C + (size_t)b * a.H * a.W1 * a.Dp;           // indexed by absolute row
        } else {
            int n = seg - a.nreg + 1;                              // stripe-local rows (A.7)
            y0 = a.sp.a[n]; y1 = min(y0 + a.sp.alt_rows, a.sp.r1[n]); ylo = y0;
            out = a.Calt + ((size_t)(b * B200S_MAX_STRIPES + n) * a.sp.alt_rows) * a.W1 * a.Dp
                  - (size_t)y0 * a.W1 * a.Dp;                      // so that out[row y] lands at local row y - y0
        }
    }
    if (y0 >= y1) return;

    // ---- shared memory carve-up ---------------------------------------------------------------
    const int xc_min = min(max(xa - BS / 2, 0), a.W1 - 1);
    const int xc_max = min(max(xa + XT - 1 + BS / 2, 0), a.W1 - 1);
    const int x_min = xc_min + a.minX1, x_max = xc_max + a.minX1;      // left-image columns touched
    const int NR = (x_max - x_min) + 2 * NT + 2;                      // right-image samples (reversed)
    const bool interior = (xa - BS / 2 >= 0) && (xa + XT - 1 + BS / 2 <= a.W1 - 1);
    unsigned* ring = smem;                                             // [BS][XT][NTS]
    unsigned* sL = ring + G::RING;                                     // 2 x { [NXI][8] left, [RA_COUNT][RW] right }

    for (int i = threadIdx.x; i < G::RING; i += NT) ring[i] = 0u;

    unsigned V[XT];
#pragma unroll
    for (int i = 0; i < XT; i++) V[i] = 0u;

    const size_t frame = (size_t)b * a.H * a.W;
    const int xr_hi = x_max - a.minD;  // right x for d index 0 at the right-most left pixel
    const int s_begin = y0 - a.SH2, s_end = y1 - 1 + a.SH2;
    const int ostride = FULL ? NTS : NT;                               // u32 words between cost columns
    int slot = 0;

    // Staging is double-buffered and software-pipelined: the records of row s+1 are fetched from global memory
    // into registers before the slide over row s starts, and unpacked into the other sL/sR buffer after it, so
    // one barrier per row suffices and the global latency hides behind the slide.
    constexpr int STG = NXI * 8 + RA_COUNT * RW;                       // words per staging buffer (sL + sR)
    uint2 pfL = make_uint2(0u, 0u), pfR[2][2];
    auto fetch = [&](const int s) {                                    // first NT left columns, first 2 NT right words
        const int yc = min(max(s, ylo), a.H - 1);
        const size_t row = frame + (size_t)yc * a.W;
        if ((int)threadIdx.x < NXI) {
            const int xc = min(max(xa - BS / 2 + (int)threadIdx.x, 0), a.W1 - 1);
            pfL = __ldg(a.recL + row + xc + a.minX1);
        }
#pragma unroll
        for (int it = 0; it < 2; it++) {
            const int k = threadIdx.x + it * NT;
#pragma unroll
            for (int q = 0; q < 2; q++) {
                const int xr = xr_hi - (2 * k + q);
                pfR[it][q] = make_uint2(0u, 0u);
                if (2 * k < NR + 2 && xr >= 0 && xr < a.W) pfR[it][q] = __ldg(a.recR + row + xr);
            }
        }
    };
    auto put_left = [&](unsigned* bL, const int i, const uint2 r) {    // replicated pairs {u, -u, lo1, -hi1} for g and r
        const unsigned g = r.x & 255u, glo = (r.x >> 8) & 255u, ghi = (r.x >> 16) & 255u;
        const unsigned rr = r.x >> 24, rlo = r.y & 255u, rhi = (r.y >> 8) & 255u;
        uint4 w0, w1;
        w0.x = pk2(g); w0.y = pk2(NEG16(g)); w0.z = pk2(glo + LO1); w0.w = pk2(NEG16(ghi));
        w1.x = pk2(rr); w1.y = pk2(NEG16(rr)); w1.z = pk2(rlo + LO1); w1.w = pk2(NEG16(rhi));
        *(uint4*)(bL + i * 8) = w0;
        *(uint4*)(bL + i * 8 + 4) = w1;
    };
    auto put_right = [&](unsigned* bR, const int k, const uint2 r0, const uint2 r1) {   // word k = halves 2k, 2k+1
        unsigned v0[RA_COUNT], v1[RA_COUNT];
#pragma unroll
        for (int q = 0; q < 2; q++) {
            const uint2 r = q ? r1 : r0;                                // an absent pixel is the all-zero record
            const unsigned g = r.x & 255u, glo = (r.x >> 8) & 255u, ghi = (r.x >> 16) & 255u;
            const unsigned rr = r.x >> 24, rlo = r.y & 255u, rhi = (r.y >> 8) & 255u;
            unsigned* v = q ? v1 : v0;
            v[RA_VG] = g; v[RA_LOG] = glo; v[RA_NHIG] = NEG16(ghi);
            v[RA_VR] = rr; v[RA_LOR] = rlo; v[RA_NHIR] = NEG16(rhi);
        }
#pragma unroll
        for (int arr = 0; arr < RA_COUNT; arr++) bR[arr * RW + k] = v0[arr] | (v1[arr] << 16);
    };
    auto stage = [&](const int s, unsigned* bL) {                      // registers (and, for tiny D, global) -> shared
        unsigned* bR = bL + NXI * 8;
        if ((int)threadIdx.x < NXI) put_left(bL, threadIdx.x, pfL);
#pragma unroll
        for (int it = 0; it < 2; it++) {
            const int k = threadIdx.x + it * NT;
            if (2 * k < NR + 2) put_right(bR, k, pfR[it][0], pfR[it][1]);
        }
        // strips wider than the thread count (very small D): the rest comes straight from global memory
        if (NT < NXI || 4 * NT < NR + 2) {
            const int yc = min(max(s, ylo), a.H - 1);
            const size_t row = frame + (size_t)yc * a.W;
            for (int i = threadIdx.x + NT; i < NXI; i += NT) {
                const int xc = min(max(xa - BS / 2 + i, 0), a.W1 - 1);
                put_left(bL, i, a.recL[row + xc + a.minX1]);
            }
            for (int k = threadIdx.x + 2 * NT; 2 * k < NR + 2; k += NT) {
                uint2 r[2];
#pragma unroll
                for (int q = 0; q < 2; q++) {
                    const int xr = xr_hi - (2 * k + q);
                    r[q] = (xr >= 0 && xr < a.W) ? __ldg(a.recR + row + xr) : make_uint2(0u, 0u);
                }
                put_right(bR, k, r[0], r[1]);
            }
        }
    };

    fetch(s_begin);
    stage(s_begin, sL);
    __syncthreads();
    for (int s = s_begin; s <= s_end; s++) {
        const int par = (s - s_begin) & 1;
        const unsigned* cL = sL + par * STG;
        if (s < s_end) fetch(s + 1);

        // -- slide along the strip
        unsigned* ringd = ring + slot * (XT * NTS) + dp;
        if (interior) cost_row<BS, NTS, true>(cL, cL + NXI * 8 + dp, ringd, V, xa, a.W1, a.minX1, x_max);
        else cost_row<BS, NTS, false>(cL, cL + NXI * 8 + dp, ringd, V, xa, a.W1, a.minX1, x_max);
        slot = (slot + 1 == BS) ? 0 : slot + 1;

        // -- emit the output row whose window just completed
        if (s - s_begin >= BS - 1) {
            const int y = s - a.SH2;
            const bool flat = (a.mode == B200S_MODE_HH4) && (y > 0) && (y + a.SH2 >= a.H);
            unsigned* orow = (unsigned*)(out + ((size_t)y * a.W1 + xa) * a.Dp) + dp;
            if (interior && !flat) {
#pragma unroll
                for (int xo = 0; xo < XT; xo++) orow[xo * ostride] = V[xo];
            } else {
#pragma unroll
                for (int xo = 0; xo < XT; xo++)
                    if (xa + xo < a.W1) orow[xo * ostride] = flat ? 0u : V[xo];
            }
        }
        if (s < s_end) stage(s + 1, sL + (par ^ 1) * STG);
        __syncthreads();   // row s+1 is staged, and everybody is done reading row s
    }
}

template <int BS, int NTS>
static int launch_cost_t(const CostArgs& a, int B, int nseg, cudaStream_t st) {
    using G = CostCfg<BS, NTS>;
    const int NT = a.Dp / 2;
    if (G::SMEM > 220 * 1024) return -1;
    dim3 grid((a.W1 + G::XT - 1) / G::XT, nseg, B);
    if (NT == NTS) {
        cudaFuncSetAttribute(k_cost<BS, NTS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G::SMEM);
        k_cost<BS, NTS, true><<<grid, NT, G::SMEM, st>>>(a);
    } else {
        cudaFuncSetAttribute(k_cost<BS, NTS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G::SMEM);
        k_cost<BS, NTS, false><<<grid, NT, G::SMEM, st>>>(a);
    }
    return 0;
}

// Build partitioning (Makefile): the kernel instantiations of one block size each compile in their own translation
// unit (-DCOST_PART=<block size>), the host side in part 0; without COST_PART everything is in this one file.
#ifndef COST_PART
#define COST_PART -1
#endif
template <int BS>
int launch_cost_bs(const CostArgs& a, int B, int nseg, cudaStream_t st) {
    const int NT = a.Dp / 2;
    if (NT <= 16) return launch_cost_t<BS, 16>(a, B, nseg, st);
    if (NT <= 32) return launch_cost_t<BS, 32>(a, B, nseg, st);
    if (NT <= 48) return launch_cost_t<BS, 48>(a, B, nseg, st);
    if (NT <= 64) return launch_cost_t<BS, 64>(a, B, nseg, st);
    // strides that follow D instead of jumping to 128: the ring of horizontal sums is BS * XT * NTS words, so D = 140
    // (example_stereo.py) at NTS = 128 left room for four CTAs per SM instead of six
    if (NT <= 72) return launch_cost_t<BS, 72>(a, B, nseg, st);   // D = 129 .. 144 exactly: compile-time global offsets too
    if (NT <= 80) return launch_cost_t<BS, 80>(a, B, nseg, st);
    if (NT <= 96) return launch_cost_t<BS, 96>(a, B, nseg, st);
    if (NT <= 112) return launch_cost_t<BS, 112>(a, B, nseg, st);
    return launch_cost_t<BS, 128>(a, B, nseg, st);
}

#if COST_PART > 0
template int launch_cost_bs<COST_PART>(const CostArgs&, int, int, cudaStream_t);
#else
#if COST_PART == 0
extern template int launch_cost_bs<1>(const CostArgs&, int, int, cudaStream_t);
extern template int launch_cost_bs<3>(const CostArgs&, int, int, cudaStream_t);
extern template int launch_cost_bs<5>(const CostArgs&, int, int, cudaStream_t);
extern template int launch_cost_bs<7>(const CostArgs&, int, int, cudaStream_t);
extern template int launch_cost_bs<9>(const CostArgs&, int, int, cudaStream_t);
extern template int launch_cost_bs<11>(const CostArgs&, int, int, cudaStream_t);
#endif
int launch_cost(const Eff& e, const StripePlan& sp, const uint2* recL, const uint2* recR, int B, uint16_t* C,
                uint16_t* Calt, cudaStream_t st) {
    CostArgs a;
    a.recL = recL; a.recR = recR; a.C = C; a.Calt = Calt;
    a.H = e.H; a.W = e.W; a.W1 = e.W1; a.D = e.D; a.Dp = e.Dp; a.minD = e.minD; a.minX1 = e.minX1;
    a.SH2 = e.SH2; a.mode = e.mode;
    static int rs_env = -1;
    if (rs_env < 0) { const char* v = getenv("B200S_COST_RS"); rs_env = v ? atoi(v) : 0; }
    // rows per segment: every segment re-computes SH2 rows above and below it, so long segments are cheaper per row
    // (measured, KITTI batch: 32 -> 54.5, 64 -> 51.7, 128 -> 50.3 us per frame) as long as the grid still fills the GPU
    a.RS = 64;
    const long long strips = (long long)((e.W1 + COST_XT - 1) / COST_XT) * B;
    if (strips * ((e.H + 127) / 128) >= 32LL * 148) a.RS = 128;
    // a single pair (or a few small frames): shorter segments, so that there are several CTAs per SM to overlap
    // (one KITTI pair: 336 CTAs of 2 warps on 148 SMs at 64 rows per segment)
    else if (strips * ((e.H + 63) / 64) < 8LL * 148) a.RS = (strips * ((e.H + 31) / 32) < 8LL * 148) ? 16 : 32;
    if (rs_env > 0) a.RS = rs_env;
    a.nreg = (e.H + a.RS - 1) / a.RS;
    a.sp = sp;
    int nseg = a.nreg + ((sp.alt_rows > 0 && sp.n > 1) ? sp.n - 1 : 0);
    if (e.Dp / 2 > 128) return -2;  // D > 256 not supported by this kernel
    switch (e.bs) {
        case 1: return launch_cost_bs<1>(a, B, nseg, st);
        case 3: return launch_cost_bs<3>(a, B, nseg, st);
        case 5: return launch_cost_bs<5>(a, B, nseg, st);
        case 7: return launch_cost_bs<7>(a, B, nseg, st);
        case 9: return launch_cost_bs<9>(a, B, nseg, st);
        case 11: return launch_cost_bs<11>(a, B, nseg, st);
        default: return -3;
    }
}
#endif  // COST_PART <= 0
