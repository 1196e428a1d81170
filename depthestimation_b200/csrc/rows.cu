// rows.cu — fused horizontal aggregation + winner selection, one scan line (image row) per lane group.
//
// Replaces, for every mode, the two horizontal paths (x-1,y) / (x+1,y) of A.5 and the per-row
// selection of A.6 inside cv2.StereoSGBM.compute (depthlib/stereo_core.py:231). Inputs: block cost
// C' and the vertical-family aggregate Sv; output: the raw int16 disparity row (before the 3x3
// median). The final S = Sv + L(->) + L(<-) never goes to HBM.
//
// A row's two scans run in opposite directions, so L(<-) of a column is needed long after it was
// produced. Instead of spilling a 2V scratch volume the kernel checkpoints:
//   pass 1: scan <- over the whole row, keep only L(<-) after every K columns (checkpoint, 2*Dp bytes);
//   pass 2: for each K-column chunk, left to right: re-run <- inside the chunk from its checkpoint into the
//           chunk's per-pixel records in shared memory (C of the chunk is kept there too), scan -> over the chunk,
//           S = Sv + L(->) + L(<-) replaces L(<-) in the records, then select_chunk picks every pixel's winner,
//           one lane per pixel (A.6, or A.7's SIMD rule for sgbm_3way).
// HBM traffic: C twice + Sv once + the checkpoints = 6.7 B/cell instead of the 14 B/cell of two read-modify-write
// passes plus a selection pass.
#include <stdlib.h>

#include "sgm_step.cuh"

#ifndef ROW_K
#define ROW_K 6  // columns per chunk (a multiple of 3, the depth of the load ring)
#endif

struct RowArgs {
    const uint16_t* C;
    const uint16_t* Sv;
    uint16_t* ck;      // checkpoints [B*H][nck][Dp]
    unsigned* keys;    // right-view candidates [B*H][W]
    int16_t* raw;      // [B*H][W]
    int H, W, W1, D, Dp, P1, P2, minD, minX1, maxX1, INV, ue, d12e, mode, nck, nlines, nosat;
    int phase;         // 0: both passes; 1: pass 1 only (C -> checkpoints, needs no Sv); 2: pass 2 only (checkpoints are there)
};

#ifndef ROWS_PF_L1
#define ROWS_PF_L1 0
#endif
#if ROWS_PF_L1
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
#else
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
#endif
#ifndef ROW_PD
#define ROW_PD 12  // pass-1 L2 prefetch distance in columns (measured: 3 -> 72.8, 6 -> 67.9, 12 -> 66.2, 18 -> 66.6, 24 -> 67.1 us per frame)
#endif
#ifndef ROWS_PFC
#define ROWS_PFC 1   // pass-2 L2 prefetch runs that many chunks ahead
#endif
#ifndef ROWS_PF2
#define ROWS_PF2 3   // pass-2 L2 prefetch of the next chunk: bit 0 = C, bit 1 = Sv
#endif
// L2 residency hints: thousands of lines stream their C and Sv through L2 once per pass; everything that is used for
// the last time is read evict_first so that the prefetched columns ahead of the scans survive (measured: 72.6 -> 67.9 us
// per frame).
#ifndef ROWS_HINTS
#define ROWS_HINTS 1
#endif
__device__ __forceinline__ uint64_t l2_policy_first() {
    uint64_t p; asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p;
}
template <int NR>
__device__ __forceinline__ void load_regs_pol(const uint16_t* p, unsigned (&r)[NR], const uint64_t pol) {
#if ROWS_HINTS
#pragma unroll
    for (int q = 0; q < NR / 4; q++)
        asm volatile("ld.global.nc.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;"
                     : "=r"(r[4 * q]), "=r"(r[4 * q + 1]), "=r"(r[4 * q + 2]), "=r"(r[4 * q + 3])
                     : "l"(p + 8 * q), "l"(pol));
#else
    load_regs_nc<NR>(p, r);
#endif
}

#ifndef ROWS_THREADS
#define ROWS_THREADS 128
#endif
template <int LW, int NR>
struct RowGeom {
    static constexpr int NCH = LW * NR / 4;                                   // 16-byte units of one pixel's disparities
    static constexpr int NCHP = NCH >= 8 ? (NCH + 7) & ~7 : NCH;              // a record is a whole number of 8-unit swizzle blocks
};
#ifndef ROWS_MINB
#define ROWS_MINB 4   // 4 CTAs (16 warps) per SM: 57 KB of shared memory each, <= 128 registers
#endif

// The winner of every pixel of a chunk is selected AFTER the chunk's -> scan, one lane per pixel (select_chunk below).
// W3 = true: sgbm_3way's selection rule (the binary's 8-wide SIMD minimum + scalar tail, A.7); false: A.6.
// NS: the sum of all paths cannot reach 32768, plain adds instead of saturating ones.
template <int LW, int NR, bool PAD, bool W3, bool NS>
__global__ void __launch_bounds__(ROWS_THREADS, (NR == 4 && LW >= 16) ? 2 : ROWS_MINB) k_rows(const RowArgs a) {
    constexpr int DPL = 2 * NR;
    // The small-call instantiations (8 disparities per lane on 16 / 32 lanes: few lines in flight, each line's serial
    // chain IS the run time) load a whole phase ahead: while a chunk's <- re-scan runs, its six Sv columns are in
    // flight, while its -> scan runs, the six C columns of the next chunk are (two sets of ROW_K x NR registers; ncu of
    // one pair per call had 39 % of the stall samples on the first use of a column that was requested two steps before).
    constexpr bool DEEP = (NR == 4 && LW >= 16);
    constexpr bool WR = (LW == 16);   // every lane of the warp runs the same loops (below): two 16-lane groups reduce with two full-warp CREDUX
    using LG = LaneGroup<LW>;
    constexpr int RPW = LG::RPW;                    // lines per warp
    constexpr int NCH = LW * NR / 4;                // 16-byte units per pixel record (all disparities of one pixel)
    constexpr int NCHP = RowGeom<LW, NR>::NCHP;     // units a record occupies (padded to whole swizzle blocks)
    constexpr int NPW = ROW_K * RPW;                // pixels per warp and chunk
    constexpr int SWZ = NCH >= 8 ? 7 : NCH - 1;     // unit swizzle: logical unit c of record p sits at (c & ~SWZ) + ((c ^ p) & SWZ)
    const int NWARP = (int)blockDim.x / 32;
    const int LINES = NWARP * RPW;                  // lines per CTA (128 threads for throughput, one warp per CTA when few lines are in flight)
    extern __shared__ __align__(16) uint4 rsm[];    // [warps][NPW][NCHP] records: L(<-), then S, of the current chunk; then C of the chunk per line
    const int lane = threadIdx.x & 31;
    const bool idle = !LG::POW2 && lane >= LG::USED;            // lanes beyond the last whole group of the warp
    const int gl = LG::POW2 ? (int)threadIdx.x % LW : (idle ? lane - LG::USED : lane % LW);
    const int lic = LG::POW2 ? (int)threadIdx.x / LW : ((int)threadIdx.x >> 5) * RPW + (idle ? RPW - 1 : lane / LW);   // line in CTA
    const int line_raw = blockIdx.x * LINES + lic;
    const bool valid = line_raw < a.nlines && !idle;
    const int line = valid ? line_raw : a.nlines - 1;   // keep the warp converged: surplus groups shadow the last line
    // every lane of the warp runs the same loops, so shuffles use the full mask with width LW (no
    // divergence check per shuffle)
    const unsigned gmask = 0xffffffffu;
    const int Dp = PAD ? a.Dp : LW * DPL;           // no padded disparities: the pitch is a compile-time constant
    const bool active = gl * DPL < Dp && !idle;
    constexpr bool nosat = NS;
    // one RECORD per pixel = all its disparities (NCH units of 8), records of a warp ordered column-major
    // (p = column-in-chunk * RPW + line-in-warp). Lane gl's q-th unit (disparities gl*DPL + 8q ..) is logical unit
    // gl + LW*q, stored at unit ^ (p & SWZ): conflict-free both for the lane groups that write a record during the
    // scans (8 lanes, 8 different units) and for the lanes that each read a whole record during selection.
    uint4* Lw = rsm + (size_t)(lic / RPW) * NPW * NCHP;
    const int rw = lic % RPW;
    uint4* Cb = rsm + (size_t)NWARP * NPW * NCHP + (size_t)lic * ROW_K * (NR / 4) * LW;    // C of the current chunk
    // where this lane's q-th unit (logical unit gl + LW q) sits inside a record of swizzle phase sw
    auto unit_at = [&](const int q, const int sw) {
        if (LG::POW2 && NCHP == NCH) return (gl ^ sw) + q * LW;      // LW is a multiple of the swizzle block (or smaller than it)
        const int c = gl + LW * q;
        return (c & ~SWZ) + ((c ^ sw) & SWZ);
    };

    unsigned pm[NR];
    make_padmask<NR>(pm, gl, a.D);
    const unsigned P1P1 = pk2((unsigned)a.P1), P2mP1 = (unsigned)(a.P2 - a.P1);
    const int d0 = 8 * (a.D / 8);
    const bool tail3 = (a.mode == B200S_MODE_SGBM_3WAY) && (a.D % 8 != 0) && d0 > 0;
    const int quirkL = tail3 ? 2 : 0;               // the (x+1, y) path ignores the d-1 neighbour at d0 (A.7)

    const uint16_t* Crow = a.C + (size_t)line * a.W1 * Dp + gl * DPL;
    const uint16_t* Srow = a.Sv + (size_t)line * a.W1 * Dp + gl * DPL;
    uint16_t* ck = a.ck + (size_t)line * a.nck * Dp + gl * DPL;
    unsigned* keys = a.keys + (size_t)line * a.W;
    int16_t* out = a.raw + (size_t)line * a.W;

    if (valid && a.phase != 1)
        for (int x = gl; x < a.W; x += LW) { keys[x] = 0xffffffffu; out[x] = (int16_t)a.INV; }
    __syncwarp();

    unsigned L[NR];
    unsigned m;
    // Chunks: chunk 0 = [0, r0) with r0 in [1, K] (the ragged one, handled by a short prologue), then full chunks of
    // K columns. Checkpoint j (j >= 1) is L(<-) after column r0 + (j-1) K, i.e. the state a re-scan of chunk j-1 starts from.
    static_assert(ROW_K % 3 == 0, "ROW_K must be a multiple of the prefetch ring (3 buffers)");
    const int r0 = a.W1 - ROW_K * ((a.W1 - 1) / ROW_K);
    // A step takes as long as the memory system needs to deliver its operands (~1600 cycles measured with one column
    // of look-ahead), so every scan runs over a ring of THREE register buffers: the loads of columns x+1 and x+2 are
    // in flight while column x is processed, also across the phase and chunk boundaries of pass 2.
    unsigned X0[NR], X1[NR], X2[NR];
#pragma unroll
    for (int j = 0; j < NR; j++) { X0[j] = 0u; X1[j] = 0u; X2[j] = 0u; }
    const uint64_t pol_first = l2_policy_first();
    auto ldC = [&](const int x, unsigned (&dst)[NR]) { if (active) load_regs_pol<NR>(Crow + (size_t)x * Dp, dst, pol_first); };
    auto ldS = [&](const int x, unsigned (&dst)[NR]) { if (active) load_regs_pol<NR>(Srow + (size_t)x * Dp, dst, pol_first); };
    // ---- pass 1: <- over the row down to column r0, checkpoint L at every chunk start -------------------
    // (a.phase, uniform: with few lines per call the two passes are two launches, and pass 1, which needs only C,
    // runs beside the vertical kernels on a second stream)
    sgm_reset<NR>(L, m);
    if (a.phase != 2) {
        // W1 - r0 is a multiple of ROW_K: the pass runs whole groups of ROW_K columns, each ending on a checkpoint
        // column, so neither the ring position nor the checkpoint test depends on x
        auto step1 = [&](const unsigned (&Cc)[NR]) { sgm_step<LW, NR, PAD, WR>(L, m, Cc, pm, gmask, gl, P1P1, P2mP1, quirkL, d0); };
        int x = a.W1 - 1;
        const uint16_t* pC = Crow + (size_t)x * Dp;                                   // C of column x
        uint16_t* pck = ck + (size_t)((x - (ROW_K - 1) - r0) / ROW_K + 1) * Dp;       // checkpoint of the group that starts at x
        if (!DEEP && x >= r0 && active) {
            load_regs_pol<NR>(pC, X0, pol_first);
            load_regs_pol<NR>(pC - Dp, X1, pol_first);
        }
        if constexpr (DEEP) {
            unsigned G0[ROW_K][NR], G1[ROW_K][NR];                                    // this group's columns / the next group's
#pragma unroll
            for (int i = 0; i < ROW_K; i++)
#pragma unroll
                for (int j = 0; j < NR; j++) { G0[i][j] = 0u; G1[i][j] = 0u; }
            if (x >= r0 && active) {
#pragma unroll
                for (int i = 0; i < ROW_K; i++) load_regs_pol<NR>(pC - (size_t)i * Dp, G0[i], pol_first);
            }
            for (; x >= r0; x -= ROW_K, pC -= (size_t)ROW_K * Dp, pck -= Dp) {
                if (active && x - ROW_K >= r0) {
#pragma unroll
                    for (int i = 0; i < ROW_K; i++) load_regs_pol<NR>(pC - (size_t)(ROW_K + i) * Dp, G1[i], pol_first);
                }
#pragma unroll
                for (int i = 0; i < ROW_K; i++) step1(G0[i]);
                if (active && valid) store_regs<NR>(pck, L);
#pragma unroll
                for (int i = 0; i < ROW_K; i++)
#pragma unroll
                    for (int j = 0; j < NR; j++) G0[i][j] = G1[i][j];
            }
        } else
        for (; x >= r0; x -= ROW_K, pC -= (size_t)ROW_K * Dp, pck -= Dp) {
            const bool more = x - ROW_K >= r0;                                        // another group follows
#pragma unroll
            for (int i = 0; i < ROW_K; i += 3) {
                if (active) {
                    load_regs_pol<NR>(pC - (size_t)(i + 2) * Dp, X2, pol_first);
                    if (x - i - ROW_PD - 2 >= r0) {
                        prefetch_l2(pC - (size_t)(i + ROW_PD) * Dp);
                        prefetch_l2(pC - (size_t)(i + ROW_PD + 1) * Dp);
                        prefetch_l2(pC - (size_t)(i + ROW_PD + 2) * Dp);
                    }
                }
                step1(X0);
                if (active && (i + 3 < ROW_K || more)) load_regs_pol<NR>(pC - (size_t)(i + 3) * Dp, X0, pol_first);
                step1(X1);
                if (active && (i + 3 < ROW_K || more)) load_regs_pol<NR>(pC - (size_t)(i + 4) * Dp, X1, pol_first);
                step1(X2);
            }
            if (active && valid) store_regs<NR>(pck, L);
        }
    }
    if (a.phase == 1) return;
    __syncwarp();
    // ---- pass 2: chunk by chunk, left to right ------------------------------------------------------
    unsigned Lr[NR];   // L(->)
    unsigned mr;
    sgm_reset<NR>(Lr, mr);
    auto stepb = [&](const int x, const int xs, const unsigned (&Cc)[NR]) {     // one <- step; L and C kept for the -> scan
        sgm_step<LW, NR, PAD, WR>(L, m, Cc, pm, gmask, gl, P1P1, P2mP1, quirkL, d0);
        const int pix = (x - xs) * RPW + rw;
        uint4* rec = Lw + pix * NCHP;
#pragma unroll
        for (int q = 0; q < NR / 4; q++) {
            if (LG::POW2 || !idle) {   // idle lanes shadow the warp's last group: they must not touch its records
                rec[unit_at(q, pix & SWZ)] = make_uint4(L[4 * q], L[4 * q + 1], L[4 * q + 2], L[4 * q + 3]);
                Cb[((x - xs) * (NR / 4) + q) * LW + gl] = make_uint4(Cc[4 * q], Cc[4 * q + 1], Cc[4 * q + 2], Cc[4 * q + 3]);
            }
        }
    };
    auto stepf = [&](const int x, const int xs, unsigned (&Sc)[NR]) {   // one -> step + winner; Sc holds Sv on entry
        unsigned Cc[NR];
#pragma unroll
        for (int q = 0; q < NR / 4; q++) {
            const uint4 t = Cb[((x - xs) * (NR / 4) + q) * LW + gl];
            Cc[4 * q] = t.x; Cc[4 * q + 1] = t.y; Cc[4 * q + 2] = t.z; Cc[4 * q + 3] = t.w;
        }
        sgm_step<LW, NR, PAD, WR>(Lr, mr, Cc, pm, gmask, gl, P1P1, P2mP1, 0, d0);
        const int pix = (x - xs) * RPW + rw;
        uint4* rec = Lw + pix * NCHP;
#pragma unroll
        for (int q = 0; q < NR / 4; q++) {
            const uint4 t = rec[unit_at(q, pix & SWZ)];
            if (nosat) {   // the sum of all paths provably stays below 32768: plain adds on the packed halves
#if SGM_FMA_ADDS
                Sc[4 * q] = __vadd2(__vadd2(Sc[4 * q], Lr[4 * q]), t.x); Sc[4 * q + 1] = __vadd2(__vadd2(Sc[4 * q + 1], Lr[4 * q + 1]), t.y);
                Sc[4 * q + 2] = __vadd2(__vadd2(Sc[4 * q + 2], Lr[4 * q + 2]), t.z); Sc[4 * q + 3] = __vadd2(__vadd2(Sc[4 * q + 3], Lr[4 * q + 3]), t.w);
#else
                Sc[4 * q] += Lr[4 * q] + t.x; Sc[4 * q + 1] += Lr[4 * q + 1] + t.y;
                Sc[4 * q + 2] += Lr[4 * q + 2] + t.z; Sc[4 * q + 3] += Lr[4 * q + 3] + t.w;
#endif
            } else {
                Sc[4 * q] = uaddmin2(uaddmin2(Sc[4 * q], Lr[4 * q], 0x7fff7fffu), t.x, 0x7fff7fffu);
                Sc[4 * q + 1] = uaddmin2(uaddmin2(Sc[4 * q + 1], Lr[4 * q + 1], 0x7fff7fffu), t.y, 0x7fff7fffu);
                Sc[4 * q + 2] = uaddmin2(uaddmin2(Sc[4 * q + 2], Lr[4 * q + 2], 0x7fff7fffu), t.z, 0x7fff7fffu);
                Sc[4 * q + 3] = uaddmin2(uaddmin2(Sc[4 * q + 3], Lr[4 * q + 3], 0x7fff7fffu), t.w, 0x7fff7fffu);
            }
        }
        // S(p, .) replaces L(<-) in the pixel's record (padded disparities forced to 0x7fff); the winner is
        // selected once the chunk's -> scan is done (select_chunk)
#pragma unroll
        for (int q = 0; q < NR / 4; q++) {
            uint4 w;
            w.x = PAD ? umax2(Sc[4 * q], pm[4 * q]) : Sc[4 * q];
            w.y = PAD ? umax2(Sc[4 * q + 1], pm[4 * q + 1]) : Sc[4 * q + 1];
            w.z = PAD ? umax2(Sc[4 * q + 2], pm[4 * q + 2]) : Sc[4 * q + 2];
            w.w = PAD ? umax2(Sc[4 * q + 3], pm[4 * q + 3]) : Sc[4 * q + 3];
            if (LG::POW2 || !idle) rec[unit_at(q, pix & SWZ)] = w;
        }
    };
    // ---- A.6 steps 1-4 for the pixels of chunk [xs, xs + ncols), one lane per pixel ---------------------------------
    // The scans leave S(p, .) of every pixel of the chunk in shared memory. Selection per pixel is scalar work
    // (the winning d, its two neighbours, the smallest S outside [best-1, best+1], the uniqueness test, the
    // right-view key, the sub-pixel parabola): done by the lane group it cost ~250 warp instructions per COLUMN
    // with one lane in eight busy; done here it costs about as much per CHUNK, NPW pixels at a time.
    auto select_chunk = [&](const int xs, const int ncols) {
        __syncwarp();                                       // every record of the chunk is complete
        const int wline0 = blockIdx.x * LINES + (lic - rw);   // first line of this warp
#pragma unroll 1
        for (int pix = threadIdx.x & 31; pix < NPW; pix += 32) {
            const int k = pix / RPW, r = pix % RPW;
            const int pline = wline0 + r;
            if (k >= ncols || pline >= a.nlines) continue;
            uint4* pb = Lw + pix * NCHP;
            const int sw = pix & SWZ;
            auto unit_of = [&](const int j) {                 // record unit that holds disparities 8j .. 8j+7
                const int c = (NR == 8) ? (j >> 1) + LW * (j & 1) : j;
                return (c & ~SWZ) + ((c ^ sw) & SWZ);
            };
            auto unit_ct = [&](const int j) -> uint4 {        // the same for a compile-time j: immediate offsets
                const int c = (NR == 8) ? (j >> 1) + LW * (j & 1) : j;
                return pb[(c & ~SWZ) + ((c & SWZ) ^ sw)];
            };
            int minS, best;
            if constexpr (!W3) {
                // A.6: minimum and the first block of 8 disparities that reaches it: min over (value << 8 | block)
                unsigned kmin = 0xffffffffu;
#pragma unroll
                for (int j = 0; j < NCH; j++) {
                    const uint4 v = unit_ct(j);
                    const unsigned m2 = umin2(umin3_2(v.x, v.y, v.z), v.w);
                    kmin = min(kmin, (min(m2 & 0xffffu, m2 >> 16) << 8) | (unsigned)j);
                }
                minS = (int)(kmin >> 8);
                const int jb = (int)(kmin & 0xffu);
                const uint4 v = pb[unit_of(jb)];
                const unsigned pk = pk2((unsigned)minS);
                const unsigned e0 = v.x ^ pk, e1 = v.y ^ pk, e2 = v.z ^ pk, e3 = v.w ^ pk;
                int hb = 7;                                   // some half equals minS; the lowest one wins
                if ((e3 & 0xffffu) == 0u) hb = 6;
                if ((e2 >> 16) == 0u) hb = 5;
                if ((e2 & 0xffffu) == 0u) hb = 4;
                if ((e1 >> 16) == 0u) hb = 3;
                if ((e1 & 0xffffu) == 0u) hb = 2;
                if ((e0 >> 16) == 0u) hb = 1;
                if ((e0 & 0xffffu) == 0u) hb = 0;
                best = 8 * jb + hb;
            } else {
                // A.7 (sgbm_3way): the minimum over the NB whole blocks is taken 8 SIMD lanes wide (SIMD lane = d % 8 =
                // the half's position in a unit); inside a SIMD lane the LAST d reaching the lane minimum is kept,
                // across SIMD lanes the lowest such d; the scalar tail d >= 8 NB only wins when strictly smaller.
                const int NB = a.D >> 3;
                unsigned l0 = 0x7fff7fffu, l1 = l0, l2 = l0, l3 = l0;
#pragma unroll
                for (int j = 0; j < NCH; j++) {
                    if (j < NB) {
                        const uint4 v = unit_ct(j);
                        l0 = umin2(l0, v.x); l1 = umin2(l1, v.y); l2 = umin2(l2, v.z); l3 = umin2(l3, v.w);
                    }
                }
                const unsigned mm = umin2(umin2(l0, l1), umin2(l2, l3));
                minS = (int)min(mm & 0xffffu, mm >> 16);
                // Which SIMD lanes reach minS (almost always one), and in each of them the LAST block that does: one
                // 16-bit load and a compare per block of that lane only, instead of five packed operations on each of the
                // four registers of every block.
                unsigned cand = 0u;
                {
                    const unsigned lq[4] = {l0, l1, l2, l3};
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        cand |= (unsigned)((int)(lq[q] & 0xffffu) == minS) << (2 * q);
                        cand |= (unsigned)((int)(lq[q] >> 16) == minS) << (2 * q + 1);
                    }
                }
                if (NB == 0) cand = 0u;                       // no whole block: only the scalar tail below
                best = 0x7fffffff;
                const unsigned short* phs = (const unsigned short*)pb;
                while (cand) {
                    const int l = __ffs((int)cand) - 1;
                    cand &= cand - 1u;
                    const unsigned short* pl = phs + l;
                    int jl = 0;
#pragma unroll
                    for (int j = 0; j < NCH; j++) {
                        if (j < NB) {
                            const int c = (NR == 8) ? (j >> 1) + LW * (j & 1) : j;
                            if ((int)pl[8 * ((c & ~SWZ) + ((c & SWZ) ^ sw))] == minS) jl = j;
                        }
                    }
                    best = min(best, 8 * jl + l);
                }
                if (a.D > 8 * NB) {                           // scalar tail: at most 7 disparities, all in unit NB
                    const uint4 v = pb[unit_of(NB)];
                    const unsigned t[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int kk = 0; kk < 8; kk++) {
                        const int val = (kk & 1) ? (int)(t[kk >> 1] >> 16) : (int)(t[kk >> 1] & 0xffffu);
                        if (8 * NB + kk < a.D && val < minS) { minS = val; best = 8 * NB + kk; }
                    }
                }
            }
            // the winner's neighbours, then knock the three out for the uniqueness test
            unsigned short* ph = (unsigned short*)pb;
            const int dm = max(best - 1, 0), dq = min(best + 1, LW * DPL - 1);
            unsigned short* pm_ = ph + 8 * unit_of(dm >> 3) + (dm & 7);
            unsigned short* pq_ = ph + 8 * unit_of(dq >> 3) + (dq & 7);
            const int Sm = *pm_, Sq = *pq_;
            if (best > 0) *pm_ = 0xffffu;
            ph[8 * unit_of(best >> 3) + (best & 7)] = 0xffffu;
            if (best + 1 < LW * DPL) *pq_ = 0xffffu;
            bool reject;
            if constexpr (!W3) {
                unsigned acc = 0xffffffffu;
#pragma unroll
                for (int j = 0; j < NCH; j++) {
                    const uint4 v = unit_ct(j);
                    acc = umin3_2(acc, v.x, v.y);
                    acc = umin3_2(acc, v.z, v.w);
                }
                const int min2 = (int)min(acc & 0xffffu, acc >> 16);
                const int nexcl = 1 + (best > 0) + (best < a.D - 1);
                reject = (a.D > nexcl) && (min2 * (100 - a.ue) < minS * 100);
            } else {
                reject = false;
                if (a.ue > 0) {
                    // SIMD part d < 8 NB: S[d] < (int16)(thr + 1); scalar tail: S[d] * (100 - ue) < minS * 100
                    const int NB = a.D >> 3;
                    unsigned la = 0xffffffffu, lb = 0xffffffffu;
#pragma unroll
                    for (int j = 0; j < NCH; j++) {
                        const uint4 v = unit_ct(j);
                        const unsigned m2 = umin2(umin3_2(v.x, v.y, v.z), v.w);
                        if (j < NB) la = umin2(la, m2); else lb = umin2(lb, m2);
                    }
                    const int m2a = (int)min(la & 0xffffu, la >> 16), m2b = (int)min(lb & 0xffffu, lb >> 16);
                    const int thr = (100 * minS) / (100 - a.ue);
                    const int t16 = (int)(short)(thr + 1);
                    if (m2a != 0xffff && m2a < t16) reject = true;
                    if (m2b != 0xffff && m2b < 0x7fff + 1 && (m2b * (100 - a.ue) < minS * 100)) {
                        // padded disparities (forced to 0x7fff) are not candidates: make sure a real tail element exists
                        const int lo = 8 * NB, hi = a.D - 1;
                        int ntail = hi - lo + 1;
                        for (int d = best - 1; d <= best + 1; d++) ntail -= (d >= lo && d <= hi);
                        if (ntail > 0) reject = true;
                    }
                }
            }
            if (reject) continue;
            const int x = xs + k;
            const size_t prow = (size_t)pline * a.W;
            const int x2 = x + a.minX1 - best - a.minD;
            if (minS < 32767) atomicMin(a.keys + prow + x2, ((unsigned)minS << 16) | (0xffffu - (unsigned)best));   // RED.MIN
            int d16 = best * 16;
            if (best > 0 && best < a.D - 1) {
                const int den2 = max(Sm + Sq - 2 * minS, 1);
                // C truncating division. |Sm - Sq| <= den2, so |quotient| <= 8.5 and a non-integer quotient is at least
                // 2^-17 away from an integer: the correctly rounded float quotient truncates to the exact result.
                d16 += (int)__fdiv_rn((float)((Sm - Sq) * 16 + den2), (float)(2 * den2));
            }
            a.raw[prow + x + a.minX1] = (int16_t)(d16 + a.minD * 16);
        }
        __syncwarp();                                       // the records are free for the next chunk
    };
    // checkpoint of the chunk that ends at xe -> L (in flight while other work runs); nothing to load at the row end
    auto fetch_ck = [&](const int xe) {
#pragma unroll
        for (int j = 0; j < NR; j++) L[j] = 0x7fff7fffu;
        if (active && xe < a.W1) load_regs<NR>(ck + (size_t)((xe - r0) / ROW_K + 1) * Dp, L);
    };
    auto start_from_ck = [&](const int xe) {
        if (xe < a.W1) {
            m = group_min<LW, WR>(local_min2<NR>(L), gmask, gl);
        } else {
            sgm_reset<NR>(L, m);
        }
    };
    auto prefetch_chunk = [&](const int xs) {    // bring C and Sv of chunk [xs, xs+K) towards L2
#if ROWS_PF2
#pragma unroll
        for (int c0 = 0; c0 < ROW_K; c0 += LW) {
            const int c = c0 + gl;
            if (active && xs + c < a.W1 && c < ROW_K) {
                if (ROWS_PF2 & 1) prefetch_l2(Crow + (size_t)(xs + c) * Dp - gl * DPL);
                if (ROWS_PF2 & 2) prefetch_l2(Srow + (size_t)(xs + c) * Dp - gl * DPL);
                if (Dp > 64) {
                    if (ROWS_PF2 & 1) prefetch_l2(Crow + (size_t)(xs + c) * Dp - gl * DPL + 64);
                    if (ROWS_PF2 & 2) prefetch_l2(Srow + (size_t)(xs + c) * Dp - gl * DPL + 64);
                }
                if (Dp > 128) {
                    if (ROWS_PF2 & 1) prefetch_l2(Crow + (size_t)(xs + c) * Dp - gl * DPL + 128);
                    if (ROWS_PF2 & 2) prefetch_l2(Srow + (size_t)(xs + c) * Dp - gl * DPL + 128);
                }
            }
        }
#endif
    };
    {   // chunk 0 = [0, r0): short and ragged, simple loops (once per row)
        fetch_ck(r0);
#pragma unroll
        for (int c = 0; c < ROWS_PFC; c++) prefetch_chunk(r0 + c * ROW_K);
        start_from_ck(r0);
        for (int x = r0 - 1; x >= 0; x--) {
            ldC(x, X2);
            stepb(x, 0, X2);
        }
        __syncwarp();
        if (r0 < a.W1) {   // set up chunk 1: its checkpoint and its first two <- columns
            fetch_ck(r0 + ROW_K);
            if (!DEEP) {
                ldC(r0 + ROW_K - 1, X0);
                ldC(r0 + ROW_K - 2, X1);
            }
        }
        for (int x = 0; x < r0; x++) {
            ldS(x, X2);
            stepf(x, 0, X2);
        }
        select_chunk(0, r0);
        __syncwarp();
    }
    if constexpr (DEEP) {
        unsigned XC[ROW_K][NR], XS[ROW_K][NR];
#pragma unroll
        for (int i = 0; i < ROW_K; i++)
#pragma unroll
            for (int j = 0; j < NR; j++) { XC[i][j] = 0u; XS[i][j] = 0u; }
        if (r0 < a.W1) {
#pragma unroll
            for (int i = 0; i < ROW_K; i++) ldC(r0 + ROW_K - 1 - i, XC[i]);       // <- order of the first full chunk
        }
        for (int xs = r0; xs < a.W1; xs += ROW_K) {
            const int xe = xs + ROW_K;
            start_from_ck(xe);
            prefetch_chunk(xe + (ROWS_PFC - 1) * ROW_K);
#pragma unroll
            for (int i = 0; i < ROW_K; i++) ldS(xs + i, XS[i]);                   // this chunk's Sv, needed after the re-scan
#pragma unroll
            for (int i = 0; i < ROW_K; i++) stepb(xe - 1 - i, xs, XC[i]);
            __syncwarp();
            if (xe < a.W1) {
                fetch_ck(xe + ROW_K);
#pragma unroll
                for (int i = 0; i < ROW_K; i++) ldC(xe + ROW_K - 1 - i, XC[i]);   // the next chunk's C, needed after the -> scan
            }
#pragma unroll
            for (int i = 0; i < ROW_K; i++) stepf(xs + i, xs, XS[i]);
            select_chunk(xs, ROW_K);
            __syncwarp();
        }
    } else
    for (int xs = r0; xs < a.W1; xs += ROW_K) {
        // full chunk [xs, xe): X0 = C(xe-1), X1 = C(xe-2) and L = checkpoint(xe) were requested a phase ago
        const int xe = xs + ROW_K;
        start_from_ck(xe);
        prefetch_chunk(xe + (ROWS_PFC - 1) * ROW_K);
#pragma unroll 1   // keep the code small: the body inlines three full steps already
        for (int i = 0; i < ROW_K; i += 3) {
            const int x = xe - 1 - i;
            ldC(x - 2, X2);
            stepb(x, xs, X0);
            if (i + 3 < ROW_K) ldC(x - 3, X0); else ldS(xs, X0);          // ... or Sv of the first -> column
            stepb(x - 1, xs, X1);
            if (i + 3 < ROW_K) ldC(x - 4, X1); else ldS(xs + 1, X1);
            stepb(x - 2, xs, X2);
        }
        __syncwarp();
        if (xe < a.W1) fetch_ck(xe + ROW_K);                             // L is dead until the next chunk
#pragma unroll 1
        for (int i = 0; i < ROW_K; i += 3) {
            const int x = xs + i;
            ldS(x + 2, X2);
            stepf(x, xs, X0);
            if (i + 3 < ROW_K) ldS(x + 3, X0); else if (xe < a.W1) ldC(xe + ROW_K - 1, X0);   // ... or the next chunk's first <- columns
            stepf(x + 1, xs, X1);
            if (i + 3 < ROW_K) ldS(x + 4, X1); else if (xe < a.W1) ldC(xe + ROW_K - 2, X1);
            stepf(x + 2, xs, X2);
        }
        select_chunk(xs, ROW_K);
        __syncwarp();
    }
    // ---- A.6 step 5: left-right check on the finished row, four pixels per lane in flight ------------------
    __syncwarp();
    if (valid) {
        constexpr int UB = 4;
        for (int xb = a.minX1 + gl; xb < a.maxX1; xb += UB * LW) {
            int d1[UB];
            unsigned ka[UB], kb[UB];
#pragma unroll
            for (int u = 0; u < UB; u++) {
                const int x = xb + u * LW;
                d1[u] = (x < a.maxX1) ? (int)__ldcg(&out[x]) : a.INV;
            }
#pragma unroll
            for (int u = 0; u < UB; u++) {
                const int x = xb + u * LW;
                const int _x = x - (d1[u] >> 4), x_ = x - ((d1[u] + 15) >> 4);
                ka[u] = kb[u] = 0xffffffffu;
                if (d1[u] != a.INV) {
                    if (_x >= 0 && _x < a.W) ka[u] = __ldcg(&keys[_x]);   // keys are updated by RED at L2: read them there
                    if (x_ >= 0 && x_ < a.W) kb[u] = __ldcg(&keys[x_]);
                }
            }
#pragma unroll
            for (int u = 0; u < UB; u++) {
                const int x = xb + u * LW;
                if (d1[u] == a.INV) continue;
                const int _d = d1[u] >> 4, d_ = (d1[u] + 15) >> 4;
                const int _x = x - _d, x_ = x - d_;
                bool c1 = false, c2 = false;
                if (_x >= 0 && _x < a.W) {
                    const int v2 = (ka[u] == 0xffffffffu) ? (int)(short)a.INV : (int)(short)((0xffffu - (ka[u] & 0xffffu)) + a.minD);
                    c1 = v2 >= a.minD && abs(v2 - _d) > a.d12e;
                }
                if (x_ >= 0 && x_ < a.W) {
                    const int v2 = (kb[u] == 0xffffffffu) ? (int)(short)a.INV : (int)(short)((0xffffu - (kb[u] & 0xffffu)) + a.minD);
                    c2 = v2 >= a.minD && abs(v2 - d_) > a.d12e;
                }
                if (c1 && c2) out[x] = (int16_t)a.INV;
            }
        }
    }
}

template <int LW, int NR>
static size_t rows_smem() {   // L / S records + C of the chunk, for ROWS_THREADS threads
    constexpr int NW = ROWS_THREADS / 32, RPW = LaneGroup<LW>::RPW;
    return ((size_t)NW * ROW_K * RPW * RowGeom<LW, NR>::NCHP + (size_t)NW * RPW * ROW_K * (NR / 4) * LW) * sizeof(uint4);
}

template <int LW, int NR, bool PAD, bool W3, bool NS>
static void launch_rows_k(const RowArgs& a, int grid, size_t smem, cudaStream_t st) {
    int sms = 0, dev = 0;   // per call: the current device is the handle's, and a process may drive several
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaFuncSetAttribute(k_rows<LW, NR, PAD, W3, NS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    // a single frame (or a few small ones) leaves most SMs without a CTA at 128 threads: spread the lines one warp per CTA
    if (grid < 2 * sms) {
        constexpr int L1 = LaneGroup<LW>::RPW;
        k_rows<LW, NR, PAD, W3, NS><<<(a.nlines + L1 - 1) / L1, 32, smem / (ROWS_THREADS / 32), st>>>(a);
    } else {
        k_rows<LW, NR, PAD, W3, NS><<<grid, ROWS_THREADS, smem, st>>>(a);
    }
}

template <int LW, int NR>
void launch_rows_t(const RowArgs& a, cudaStream_t st) {
    constexpr int LINES = ROWS_THREADS / 32 * LaneGroup<LW>::RPW;
    const bool w3 = (a.mode == B200S_MODE_SGBM_3WAY) && (8 * (a.D / 8) > 0);
    const size_t smem = rows_smem<LW, NR>();
    const int grid = (a.nlines + LINES - 1) / LINES;
    if (a.D == LW * 2 * NR) {   // no padded disparities anywhere: drop the pad clamps
        if (w3) { if (a.nosat) launch_rows_k<LW, NR, false, true, true>(a, grid, smem, st); else launch_rows_k<LW, NR, false, true, false>(a, grid, smem, st); }
        else if (a.nosat) launch_rows_k<LW, NR, false, false, true>(a, grid, smem, st);
        else launch_rows_k<LW, NR, false, false, false>(a, grid, smem, st);
    } else {
        RowArgs b = a;
        b.nosat = 0;            // padded halves hold 0x7fff: their sums must keep saturating
        if (w3) launch_rows_k<LW, NR, true, true, false>(b, grid, smem, st);
        else launch_rows_k<LW, NR, true, false, false>(b, grid, smem, st);
    }
}

template <int LW, int NR>
int rows_capacity_t() {
    constexpr int LINES = ROWS_THREADS / 32 * LaneGroup<LW>::RPW;
    const size_t smem = rows_smem<LW, NR>();
    cudaFuncSetAttribute(k_rows<LW, NR, true, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int nb = 0, dev = 0, sms = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_rows<LW, NR, true, false, false>, ROWS_THREADS, smem) != cudaSuccess) { cudaGetLastError(); return 0; }
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return nb * sms * LINES;
}

// Build partitioning (Makefile): the lane-group instantiations are spread over six translation units
// (-DROWS_PART=1..6), the host side is part 0; without ROWS_PART everything is in this one file.
#ifndef ROWS_PART
#define ROWS_PART -1
#endif
#if ROWS_PART == 1
template void launch_rows_t<2, 4>(const RowArgs&, cudaStream_t);
template int rows_capacity_t<2, 4>();
#elif ROWS_PART == 0
extern template void launch_rows_t<2, 4>(const RowArgs&, cudaStream_t);
extern template int rows_capacity_t<2, 4>();
#endif
#if ROWS_PART == 2
template void launch_rows_t<4, 4>(const RowArgs&, cudaStream_t);
template int rows_capacity_t<4, 4>();
#elif ROWS_PART == 0
extern template void launch_rows_t<4, 4>(const RowArgs&, cudaStream_t);
extern template int rows_capacity_t<4, 4>();
#endif
#if ROWS_PART == 3
template void launch_rows_t<8, 4>(const RowArgs&, cudaStream_t);
template int rows_capacity_t<8, 4>();
#elif ROWS_PART == 0
extern template void launch_rows_t<8, 4>(const RowArgs&, cudaStream_t);
extern template int rows_capacity_t<8, 4>();
#endif
#if ROWS_PART == 4
template void launch_rows_t<16, 4>(const RowArgs&, cudaStream_t);
template int rows_capacity_t<16, 4>();
#elif ROWS_PART == 0
extern template void launch_rows_t<16, 4>(const RowArgs&, cudaStream_t);
extern template int rows_capacity_t<16, 4>();
#endif
#if ROWS_PART == 5
template void launch_rows_t<32, 4>(const RowArgs&, cudaStream_t);
template int rows_capacity_t<32, 4>();
#elif ROWS_PART == 0
extern template void launch_rows_t<32, 4>(const RowArgs&, cudaStream_t);
extern template int rows_capacity_t<32, 4>();
#endif
#if ROWS_PART == 6
template void launch_rows_t<5, 8>(const RowArgs&, cudaStream_t);
template int rows_capacity_t<5, 8>();
#elif ROWS_PART == 0
extern template void launch_rows_t<5, 8>(const RowArgs&, cudaStream_t);
extern template int rows_capacity_t<5, 8>();
#endif
#if ROWS_PART == 1
template void launch_rows_t<6, 8>(const RowArgs&, cudaStream_t);
template int rows_capacity_t<6, 8>();
#elif ROWS_PART == 0
extern template void launch_rows_t<6, 8>(const RowArgs&, cudaStream_t);
extern template int rows_capacity_t<6, 8>();
#endif
#if ROWS_PART == 2
template void launch_rows_t<7, 8>(const RowArgs&, cudaStream_t);
template int rows_capacity_t<7, 8>();
#elif ROWS_PART == 0
extern template void launch_rows_t<7, 8>(const RowArgs&, cudaStream_t);
extern template int rows_capacity_t<7, 8>();
#endif
#if ROWS_PART == 3
template void launch_rows_t<8, 8>(const RowArgs&, cudaStream_t);
template int rows_capacity_t<8, 8>();
#elif ROWS_PART == 0
extern template void launch_rows_t<8, 8>(const RowArgs&, cudaStream_t);
extern template int rows_capacity_t<8, 8>();
#endif
#if ROWS_PART == 4
template void launch_rows_t<9, 8>(const RowArgs&, cudaStream_t);
template int rows_capacity_t<9, 8>();
#elif ROWS_PART == 0
extern template void launch_rows_t<9, 8>(const RowArgs&, cudaStream_t);
extern template int rows_capacity_t<9, 8>();
#endif
#if ROWS_PART == 5
template void launch_rows_t<10, 8>(const RowArgs&, cudaStream_t);
template int rows_capacity_t<10, 8>();
#elif ROWS_PART == 0
extern template void launch_rows_t<10, 8>(const RowArgs&, cudaStream_t);
extern template int rows_capacity_t<10, 8>();
#endif
#if ROWS_PART == 6
template void launch_rows_t<16, 8>(const RowArgs&, cudaStream_t);
template int rows_capacity_t<16, 8>();
#elif ROWS_PART == 0
extern template void launch_rows_t<16, 8>(const RowArgs&, cudaStream_t);
extern template int rows_capacity_t<16, 8>();
#endif
#if ROWS_PART == 1
template void launch_rows_t<32, 8>(const RowArgs&, cudaStream_t);
template int rows_capacity_t<32, 8>();
#elif ROWS_PART == 0
extern template void launch_rows_t<32, 8>(const RowArgs&, cudaStream_t);
extern template int rows_capacity_t<32, 8>();
#endif
#if ROWS_PART <= 0
// image rows the row kernel keeps in flight at once (one wave); a batch of B frames is B*H rows
int rows_capacity_lines(const Eff& e) {
    if (e.Dp > 64) {
        switch ((e.Dp + 15) / 16) {
            case 5: return rows_capacity_t<5, 8>();
            case 6: return rows_capacity_t<6, 8>();
            case 7: return rows_capacity_t<7, 8>();
            case 8: return rows_capacity_t<8, 8>();
            case 9: return rows_capacity_t<9, 8>();
            case 10: return rows_capacity_t<10, 8>();
            case 11: case 12: case 13: case 14: case 15: case 16: return rows_capacity_t<16, 8>();
            default: return rows_capacity_t<32, 8>();
        }
    }
    int LW = 2;
    while (LW * 8 < e.Dp) LW *= 2;
    return LW == 2 ? rows_capacity_t<2, 4>() : (LW == 4 ? rows_capacity_t<4, 4>() : rows_capacity_t<8, 4>());
}

size_t rows_ck_bytes(const Eff& e, int B) {
    int nck = (e.W1 + ROW_K - 1) / ROW_K + 1;
    return (size_t)B * e.H * nck * e.Dp * sizeof(uint16_t);
}

int rows_small_call(const Eff& e, int B) {
    static int small_env = -1;
    if (small_env < 0) { const char* v = getenv("B200S_ROWS_SMALL"); small_env = v ? atoi(v) : 2; }   // 0 never, 1 always, 2 auto
    if (!small_env) return 0;
    if (small_env == 1) return 1;
    int sms = 0, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long long nlines = (long long)B * e.H;
    const long long warps = (e.Dp <= 128) ? (nlines + 1) / 2 : nlines;
    // measured (KITTI pairs per call, small / regular variant, pairs/s): sgbm 4: 2330 / 1841, 8: 3690 / 3078, 12: 4630 / 4104,
    // 16: 3904 / 3880; sgbm_3way 4: 2428 / 2076, 8: 3577 / 3234, 12: 3446 / 4064
    const bool vertical_cluster = (e.mode == B200S_MODE_SGBM || e.mode == B200S_MODE_HH);   // pass 1 then hides beside it
    // with the two-reduction group minimum (sgm_step.cuh) the small variant of sgbm keeps up with the regular one at every
    // call size and wins up to ~50 KITTI pairs (pairs/s, small / regular: 20: 4650 / 4600, 32: 4760 / 4220, 48: 4990 /
    // 4780, 75: 5167 / 5171, 150: 5220 / 5226): taken up to 64 warps per SM at D <= 128
    return warps <= (vertical_cluster ? (e.Dp <= 128 ? 64LL : 16LL) : 11LL) * sms;
}

int launch_rows(const Eff& e, int B, const uint16_t* C, const uint16_t* Sv, uint16_t* ck, unsigned* keys, int16_t* raw,
                cudaStream_t st, int phase) {
    RowArgs a;
    a.phase = phase;
    a.C = C; a.Sv = Sv; a.ck = ck; a.keys = keys; a.raw = raw;
    a.H = e.H; a.W = e.W; a.W1 = e.W1; a.D = e.D; a.Dp = e.Dp; a.P1 = e.P1e; a.P2 = e.P2e; a.minD = e.minD;
    a.minX1 = e.minX1; a.maxX1 = e.maxX1; a.INV = e.INV; a.ue = e.ue; a.d12e = e.d12e; a.mode = e.mode;
    a.nck = (e.W1 + ROW_K - 1) / ROW_K + 1;
    a.nosat = sum_cannot_saturate(e) ? 1 : 0;
    a.nlines = B * e.H;
    if (e.Dp > 512) return -2;
    // Few lines in the call (a single pair, a handful of small frames): the GPU is far from full and the time is the
    // length of one line's serial scan, which a lane group issues at well under one instruction per clock. Twice the
    // lanes per line with half the registers each halves the instructions on that critical path (and costs ~1.5x the
    // issue slots per line, which is why full batches keep 16 disparities per lane).
    if (e.Dp > 64 && e.Dp <= 256 && rows_small_call(e, B)) {
        if (e.Dp <= 128) launch_rows_t<16, 4>(a, st); else launch_rows_t<32, 4>(a, st);
        return 0;
    }
    if (e.Dp > 64) {             // 16 disparities per lane; the lane count follows D (a warp carries 32 / LW lines)
        switch ((e.Dp + 15) / 16) {
            case 5: launch_rows_t<5, 8>(a, st); break;     // D  65 ..  80: 6 lines per warp
            case 6: launch_rows_t<6, 8>(a, st); break;     // D  81 ..  96: 5
            case 7: launch_rows_t<7, 8>(a, st); break;     // D  97 .. 112: 4
            case 8: launch_rows_t<8, 8>(a, st); break;     // D 113 .. 128: 4
            case 9: launch_rows_t<9, 8>(a, st); break;     // D 129 .. 144: 3 (example_stereo.py: D = 140)
            case 10: launch_rows_t<10, 8>(a, st); break;   // D 145 .. 160: 3
            case 11: case 12: case 13: case 14: case 15: case 16: launch_rows_t<16, 8>(a, st); break;   // 2 lines per warp either way
            default: launch_rows_t<32, 8>(a, st); break;
        }
    } else {                     // 8 disparities per lane
        int LW = 2;
        while (LW * 8 < e.Dp) LW *= 2;
        switch (LW) {
            case 2: launch_rows_t<2, 4>(a, st); break;
            case 4: launch_rows_t<4, 4>(a, st); break;
            default: launch_rows_t<8, 4>(a, st); break;
        }
    }
    return 0;
}
#endif  // ROWS_PART <= 0
