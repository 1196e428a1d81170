// wta.cu — per-row winner selection (A.6): argmin over d, uniqueness (scalar rule and the
// sgbm_3way SIMD rule), right-view disparity disp2, parabola sub-pixel, left-right check.
//
// Replaces the disparity-selection loop of cv2.StereoSGBM.compute (depthlib/stereo_core.py:231).
// One CTA per (row, frame). A group of LW lanes holds S(y, xc, 0..Dp) in registers (8 d / lane).
// disp2 is built with shared-memory atomicMin on the key (minS << 16 | 0xFFFF - best): the
// reference scans xc from right to left and only replaces on a strictly smaller cost, i.e. the
// winner for a right-view pixel is the smallest minS, ties to the largest xc (= largest best).
#include "common.cuh"

struct WtaArgs {
    const uint16_t* S;
    int16_t* raw;
    int H, W, W1, D, Dp, minD, minX1, maxX1, INV, ue, d12e, mode;
};

__device__ __forceinline__ int elem8(const unsigned (&v)[4], int k) {
    unsigned r = v[0];
    r = (k >> 1) == 1 ? v[1] : r;
    r = (k >> 1) == 2 ? v[2] : r;
    r = (k >> 1) == 3 ? v[3] : r;
    return (k & 1) ? (int)(r >> 16) : (int)(r & 0xffffu);
}

template <int LW>
__global__ void __launch_bounds__(256) k_wta(const WtaArgs a) {
    extern __shared__ unsigned wsm[];
    unsigned* keys = wsm;                      // [W]
    short* outrow = (short*)(wsm + a.W);       // [W]
    const int y = blockIdx.x, b = blockIdx.y;
    for (int x = threadIdx.x; x < a.W; x += blockDim.x) {
        keys[x] = 0xffffffffu;
        outrow[x] = (short)a.INV;
    }
    __syncthreads();

    const int gl = threadIdx.x % LW, g = threadIdx.x / LW, ngroups = blockDim.x / LW;
    const int gbase = (threadIdx.x % 32) / LW * LW;
    const unsigned gmask = (LW == 32) ? 0xffffffffu : (((1u << LW) - 1u) << gbase);
    const bool active = gl * 8 < a.Dp;
    unsigned pm[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        int d = gl * 8 + 2 * j;
        pm[j] = (d >= a.D ? 0x7fffu : 0u) | (d + 1 >= a.D ? 0x7fff0000u : 0u);
    }
    const int d0 = 8 * (a.D / 8);
    const bool simd3 = (a.mode == B200S_MODE_SGBM_3WAY) && d0 > 0;
    const int qlane = a.D / 8;  // lane holding the scalar tail d0..D-1 (if D % 8 != 0)
    const uint16_t* Srow = a.S + ((size_t)b * a.H + y) * a.W1 * a.Dp;

    for (int xc = g; xc < a.W1; xc += ngroups) {
        unsigned v[4] = {0x7fff7fffu, 0x7fff7fffu, 0x7fff7fffu, 0x7fff7fffu};
        if (active) {
            uint4 q = *(const uint4*)(Srow + (size_t)xc * a.Dp + gl * 8);
            v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
        }
        int ev[8];
#pragma unroll
        for (int j = 0; j < 4; j++) { ev[2 * j] = (int)(v[j] & 0xffffu); ev[2 * j + 1] = (int)(v[j] >> 16); }
        // padded disparities never win: compare against 0x7fff-forced copies
        unsigned w[4];
#pragma unroll
        for (int j = 0; j < 4; j++) w[j] = umax2(v[j], pm[j]);

        int minS, best;
        if (!simd3) {
            unsigned mn = umin2(umin3_2(w[0], w[1], w[2]), w[3]);
#pragma unroll
            for (int o = LW / 2; o > 0; o >>= 1) mn = umin2(mn, __shfl_xor_sync(gmask, mn, o, LW));
            minS = (int)min(mn & 0xffffu, mn >> 16);
            int cand = 0x7fffffff;
#pragma unroll
            for (int k = 7; k >= 0; k--)
                if (gl * 8 + k < a.D && ev[k] == minS) cand = gl * 8 + k;
#pragma unroll
            for (int o = LW / 2; o > 0; o >>= 1) cand = min(cand, __shfl_xor_sync(gmask, cand, o, LW));
            best = cand;  // lowest d reaching the minimum
        } else {
            // sgbm_3way: 8-lane SIMD minimum over d < d0 (within SIMD lane j = d % 8 the LAST d
            // reaching the lane minimum is kept; across lanes the lowest such d), then a scalar
            // tail d >= d0 that only wins when strictly smaller.
            const bool full = (gl < qlane);
            unsigned lm[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                lm[j] = full ? v[j] : 0x7fff7fffu;
#pragma unroll
                for (int o = LW / 2; o > 0; o >>= 1) lm[j] = umin2(lm[j], __shfl_xor_sync(gmask, lm[j], o, LW));
            }
            int lmin[8];
#pragma unroll
            for (int j = 0; j < 4; j++) { lmin[2 * j] = (int)(lm[j] & 0xffffu); lmin[2 * j + 1] = (int)(lm[j] >> 16); }
            minS = 32767;
#pragma unroll
            for (int k = 0; k < 8; k++) minS = min(minS, lmin[k]);
            best = -1;
#pragma unroll
            for (int k = 0; k < 8; k++) {
                unsigned bits = (__ballot_sync(gmask, full && ev[k] == lmin[k]) >> gbase) & ((LW == 32) ? 0xffffffffu : ((1u << LW) - 1u));
                if (lmin[k] == minS) {
                    // lanes can tie at 32767 with no full lane matching only if there are no full lanes (d0 == 0), excluded
                    int hl = 31 - __clz(bits);
                    int dk = hl * 8 + k;
                    if (best < 0 || dk < best) best = dk;
                }
            }
            if (a.D > d0) {  // scalar tail, evaluated by the lane that holds it, then broadcast
                int ms = minS, bs_ = best;
#pragma unroll
                for (int k = 0; k < 8; k++)
                    if (d0 + k < a.D && ev[k] < ms) { ms = ev[k]; bs_ = d0 + k; }
                minS = __shfl_sync(gmask, ms, qlane, LW);
                best = __shfl_sync(gmask, bs_, qlane, LW);
            }
        }

        // ---- uniqueness -------------------------------------------------------------------
        bool viol = false;
        if (a.mode == B200S_MODE_SGBM_3WAY) {
            if (a.ue > 0) {
                const int thr = (100 * minS) / (100 - a.ue);
                const int t16 = (int)(short)(thr + 1);
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    int d = gl * 8 + k;
                    if (d < a.D && (d < best - 1 || d > best + 1)) {
                        if (d < d0) viol |= (ev[k] < t16);
                        else viol |= (ev[k] * (100 - a.ue) < minS * 100);
                    }
                }
            }
        } else {
#pragma unroll
            for (int k = 0; k < 8; k++) {
                int d = gl * 8 + k;
                if (d < a.D && (d < best - 1 || d > best + 1)) viol |= (ev[k] * (100 - a.ue) < minS * 100);
            }
        }
        const bool reject = __ballot_sync(gmask, viol) != 0u;
        // ---- neighbours of the winner for the parabola fit (shuffles outside the branch) -------
        const int dm = max(best - 1, 0), dq = min(best + 1, a.Dp - 1);
        const int Sm = __shfl_sync(gmask, elem8(v, dm & 7), dm >> 3, LW);
        const int Sq = __shfl_sync(gmask, elem8(v, dq & 7), dq >> 3, LW);
        if (reject) continue;
        if (gl == 0) {
            const int x2 = xc + a.minX1 - best - a.minD;
            if (minS < 32767) atomicMin(&keys[x2], ((unsigned)minS << 16) | (0xffffu - (unsigned)best));
            int d16 = best * 16;
            if (best > 0 && best < a.D - 1) {
                int den2 = max(Sm + Sq - 2 * minS, 1);
                d16 += ((Sm - Sq) * 16 + den2) / (2 * den2);  // C truncating division
            }
            outrow[xc + a.minX1] = (short)(d16 + a.minD * 16);
        }
    }
    __syncthreads();
    // ---- left-right check + store ------------------------------------------------------------
    int16_t* orow = a.raw + ((size_t)b * a.H + y) * a.W;
    for (int x = threadIdx.x; x < a.W; x += blockDim.x) {
        int d1 = outrow[x];
        if (x >= a.minX1 && x < a.maxX1 && d1 != a.INV) {
            const int _d = d1 >> 4, d_ = (d1 + 15) >> 4;
            const int _x = x - _d, x_ = x - d_;
            bool c1 = false, c2 = false;
            if (_x >= 0 && _x < a.W) {
                unsigned k = keys[_x];
                int v2 = (k == 0xffffffffu) ? (int)(short)a.INV : (int)(short)((0xffffu - (k & 0xffffu)) + a.minD);
                c1 = v2 >= a.minD && abs(v2 - _d) > a.d12e;
            }
            if (x_ >= 0 && x_ < a.W) {
                unsigned k = keys[x_];
                int v2 = (k == 0xffffffffu) ? (int)(short)a.INV : (int)(short)((0xffffu - (k & 0xffffu)) + a.minD);
                c2 = v2 >= a.minD && abs(v2 - d_) > a.d12e;
            }
            if (c1 && c2) d1 = a.INV;
        }
        orow[x] = (int16_t)d1;
    }
}

void launch_wta(const Eff& e, int B, const uint16_t* S, int16_t* raw, cudaStream_t st) {
    WtaArgs a;
    a.S = S; a.raw = raw;
    a.H = e.H; a.W = e.W; a.W1 = e.W1; a.D = e.D; a.Dp = e.Dp; a.minD = e.minD; a.minX1 = e.minX1;
    a.maxX1 = e.maxX1; a.INV = e.INV; a.ue = e.ue; a.d12e = e.d12e; a.mode = e.mode;
    int LW = 2;
    while (LW * 8 < e.Dp) LW *= 2;
    dim3 grid(e.H, B);
    size_t smem = (size_t)e.W * 4 + (size_t)e.W * 2 + 16;
    switch (LW) {
        case 2: k_wta<2><<<grid, 256, smem, st>>>(a); break;
        case 4: k_wta<4><<<grid, 256, smem, st>>>(a); break;
        case 8: k_wta<8><<<grid, 256, smem, st>>>(a); break;
        case 16: k_wta<16><<<grid, 256, smem, st>>>(a); break;
        default: k_wta<32><<<grid, 256, smem, st>>>(a); break;
    }
}
