// sgm_step.cuh — one step of the SGM path recursion (A.4) for a group of LW lanes holding
// 2*NR disparities per lane (NR packed u16x2 registers, d contiguous within the lane).
//
//   L(p,d) = C'(p,d) + min(L(q,d), L(q,d-1)+P1, L(q,d+1)+P1, m+P2) - m,   m = min_d L(q,d)
//
// with C' the block cost without the reference's +P2 (same L values as A.4). Per register:
//   2 PRMT (d-1 / d+1 neighbours) + VIMNMX3.U16x2 + VIADDMNMX.U16x2 + 2 integer add/sub + VIMNMX.U16x2
// Values stay in [0, 32767] (pads are forced to 0x7fff), so no u16 half ever carries or borrows.
#pragma once
#include "common.cuh"

#ifndef SGM_FMA_ADDS
#define SGM_FMA_ADDS 1
#endif
#ifndef SGM_WARP_REDUX
#define SGM_WARP_REDUX 1   // 0: shuffle tree also for whole-warp groups (for comparison)
#endif

// ---- 16-byte global loads / stores of a lane's NR packed registers --------------------------------
template <int NR>
__device__ __forceinline__ void load_regs(const uint16_t* p, unsigned (&r)[NR]) {
#pragma unroll
    for (int q = 0; q < NR / 4; q++) {
        uint4 t = *(const uint4*)(p + 8 * q);
        r[4 * q] = t.x; r[4 * q + 1] = t.y; r[4 * q + 2] = t.z; r[4 * q + 3] = t.w;
    }
}
template <int NR>
__device__ __forceinline__ void load_regs_nc(const uint16_t* p, unsigned (&r)[NR]) {
#pragma unroll
    for (int q = 0; q < NR / 4; q++) {
        uint4 t = __ldg((const uint4*)(p + 8 * q));
        r[4 * q] = t.x; r[4 * q + 1] = t.y; r[4 * q + 2] = t.z; r[4 * q + 3] = t.w;
    }
}
template <int NR>
__device__ __forceinline__ void store_regs(uint16_t* p, const unsigned (&r)[NR]) {
#pragma unroll
    for (int q = 0; q < NR / 4; q++) *(uint4*)(p + 8 * q) = make_uint4(r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3]);
}

// ---- lane groups ----------------------------------------------------------------------------------------------------------
// LW lanes hold the disparities of one pixel. LW need not be a power of two: a warp then carries RPW = 32 / LW whole groups
// in its first RPW * LW lanes (the rest idle), e.g. 3 groups of 9 lanes for D = 140 instead of 2 groups of 16 with 7 lanes
// idle each. Neighbour exchange uses absolute lane indices (the group-edge lanes are patched by the caller anyway); the
// group minimum of a non-power-of-two group is a shuffle-down tree with out-of-group sources ignored, then a broadcast.
template <int LW> struct LaneGroup {
    static constexpr bool POW2 = (LW & (LW - 1)) == 0;
    static constexpr int RPW = 32 / LW;                       // groups per warp
    static constexpr int USED = RPW * LW;                     // lanes of a warp that belong to a group
    static constexpr int P2 = LW <= 1 ? 1 : LW <= 2 ? 2 : LW <= 4 ? 4 : LW <= 8 ? 8 : LW <= 16 ? 16 : 32;   // next power of two
};
template <int LW>
__device__ __forceinline__ unsigned group_mask() {
    if (LW == 32) return 0xffffffffu;
    return ((1u << LW) - 1u) << ((threadIdx.x & 31) / LW * LW);
}
// value of the previous / next lane (garbage at the group edges: callers replace it there)
template <int LW>
__device__ __forceinline__ unsigned grp_prev(unsigned v, unsigned gmask) {
    if (LaneGroup<LW>::POW2) return __shfl_up_sync(gmask, v, 1, LW);
    return __shfl_sync(0xffffffffu, v, (int)(threadIdx.x & 31) - 1);
}
template <int LW>
__device__ __forceinline__ unsigned grp_next(unsigned v, unsigned gmask) {
    if (LaneGroup<LW>::POW2) return __shfl_down_sync(gmask, v, 1, LW);
    return __shfl_sync(0xffffffffu, v, (int)(threadIdx.x & 31) + 1);
}

template <int LW>
__device__ __forceinline__ unsigned group_min2(unsigned v, unsigned gmask, const int gl = 0) {
    if constexpr (LaneGroup<LW>::POW2) {
#pragma unroll
        for (int o = LW / 2; o > 0; o >>= 1) v = umin2(v, __shfl_xor_sync(gmask, v, o, LW));
        return v;
    } else {
#pragma unroll
        for (int o = LaneGroup<LW>::P2 / 2; o > 0; o >>= 1) {
            const unsigned t = __shfl_down_sync(0xffffffffu, v, o);
            if (gl + o < LW) v = umin2(v, t);
        }
        return __shfl_sync(0xffffffffu, v, (int)(threadIdx.x & 31) - gl);   // group lane 0 holds the minimum
    }
}

// minimum over both halves of v over the lane group, as a scalar. A group that is the whole warp uses the hardware warp
// reduction (CREDUX.MIN: ~40 cycles on the dependent path against ~170 for five shuffle + min levels, tools/ubench/redux.cu;
// with a sub-warp member mask the same intrinsic serialises over the groups and is slower than the shuffle tree).
// WR (the caller guarantees that all 32 lanes of the warp are converged here): two groups of 16 take two FULL-warp
// reductions, each group masking the other's lanes out with 0xffffffff, and every lane keeps its own group's result:
// two pipelined CREDUX against four shuffle + min levels on the critical path of a single line's scan.
template <int LW, bool WR = false>
__device__ __forceinline__ unsigned group_min(unsigned v, unsigned gmask, const int gl = 0) {
    if constexpr (LW == 32 && SGM_WARP_REDUX) {
        return __reduce_min_sync(0xffffffffu, min(v & 0xffffu, v >> 16));
    } else if constexpr (LW == 16 && WR && SGM_WARP_REDUX) {
        const unsigned s = min(v & 0xffffu, v >> 16);
        const bool up = (threadIdx.x & 16u) != 0u;
        const unsigned lo = __reduce_min_sync(0xffffffffu, up ? 0xffffffffu : s);
        const unsigned hi = __reduce_min_sync(0xffffffffu, up ? s : 0xffffffffu);
        return up ? hi : lo;
    } else {
        const unsigned mn = group_min2<LW>(v, gmask, gl);
        return min(mn & 0xffffu, mn >> 16);
    }
}

template <int NR>
__device__ __forceinline__ unsigned local_min2(const unsigned (&v)[NR]) {
    unsigned m = v[0];
#pragma unroll
    for (int j = 1; j + 1 < NR; j += 2) m = umin3_2(m, v[j], v[j + 1]);
    if ((NR & 1) == 0) m = umin2(m, v[NR - 1]);
    return m;
}

// pad mask: 0x7fff in every half whose disparity index is >= D
template <int NR>
__device__ __forceinline__ void make_padmask(unsigned (&pm)[NR], int gl, int D) {
#pragma unroll
    for (int j = 0; j < NR; j++) {
        int d = gl * 2 * NR + 2 * j;
        pm[j] = (d >= D ? 0x7fffu : 0u) | (d + 1 >= D ? 0x7fff0000u : 0u);
    }
}

// quirk: sgbm_3way D%8 tail artefact (A.7): 0 none, 1 "d-1 neighbour of d0 comes from the
// already-updated current column", 2 "d-1 neighbour of d0 is ignored". d0 = 8*floor(D/8).
// PAD = false: every lane of the group holds only real disparities (D == LW * 2 * NR), the pad clamp is dropped.
template <int LW, int NR, bool PAD = true>
__device__ __forceinline__ unsigned sgm_step_core(unsigned (&L)[NR], const unsigned m, const unsigned (&Cq)[NR],
                                         const unsigned (&pm)[NR], const unsigned gmask, const int gl,
                                         const unsigned P1P1, const unsigned P2mP1, const int quirk, const int d0) {
    unsigned up = grp_prev<LW>(L[NR - 1], gmask);
    unsigned dn = grp_next<LW>(L[0], gmask);
    if (gl == 0) up = 0x7fff7fffu;
    if (gl == LW - 1) dn = 0x7fff7fffu;
    const unsigned mm = pk2(m);
    const unsigned dmp = pk2(m + P2mP1);
    unsigned Ln[NR];
#if SGM_FMA_ADDS
    const unsigned nmm = pk2(0x10000u - m);                                // -m per half (wraps inside the half)
#endif
#pragma unroll
    for (int j = 0; j < NR; j++) {
        unsigned bq = shift_pair(j == 0 ? up : L[j - 1], L[j]);            // (L[d-1], L[d])
        unsigned cn = shift_pair(L[j], j == NR - 1 ? dn : L[j + 1]);       // (L[d+1], L[d+2])
        unsigned tt = umin3_2(bq, cn, dmp);
        tt = uaddmin2(tt, P1P1, L[j]);
#if SGM_FMA_ADDS
        // C + (tt - m) as two VIADD.16x2: they issue on the FMA pipe, which is idle, while the one IADD3 this replaces
        // competes with the min / max / permute instructions for the half-rate ALU pipe (tools/ubench/pipes2.cu)
        const unsigned ln = __vadd2(__vadd2(tt, nmm), Cq[j]);
#else
        const unsigned ln = Cq[j] + tt - mm;
#endif
        Ln[j] = PAD ? umax2(ln, pm[j]) : ln;
    }
    if (PAD && quirk) {
        // d0 is a multiple of 8, so it sits in register 0 or (NR == 8) register 4 of lane d0 / (2 NR)
        const unsigned upn = grp_prev<LW>(Ln[NR - 1], gmask);
        const int ql = d0 / (2 * NR), jq = (d0 % (2 * NR)) / 2;
        if (gl == ql) {
            if (jq == 0) {
                unsigned bq = shift_pair(quirk == 1 ? upn : 0x7fff7fffu, L[0]);
                unsigned cn = shift_pair(L[0], NR == 1 ? dn : L[NR > 1 ? 1 : 0]);
                unsigned tt = uaddmin2(umin3_2(bq, cn, dmp), P1P1, L[0]);
                Ln[0] = umax2(Cq[0] + tt - mm, pm[0]);
            } else if (NR >= 8) {
                constexpr int J = NR / 2;
                unsigned bq = shift_pair(quirk == 1 ? Ln[J - 1] : 0x7fff7fffu, L[J]);
                unsigned cn = shift_pair(L[J], L[J + 1 < NR ? J + 1 : J]);
                unsigned tt = uaddmin2(umin3_2(bq, cn, dmp), P1P1, L[J]);
                Ln[J] = umax2(Cq[J] + tt - mm, pm[J]);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < NR; j++) L[j] = Ln[j];
    return local_min2<NR>(Ln);   // packed (min over even d, min over odd d) of this lane; the caller reduces it
}


// full step: the minimum over d of the new column is reduced across the lane group here
template <int LW, int NR, bool PAD = true, bool WR = false>
__device__ __forceinline__ void sgm_step(unsigned (&L)[NR], unsigned& m, const unsigned (&Cq)[NR],
                                         const unsigned (&pm)[NR], const unsigned gmask, const int gl,
                                         const unsigned P1P1, const unsigned P2mP1, const int quirk, const int d0) {
    m = group_min<LW, WR>(sgm_step_core<LW, NR, PAD>(L, m, Cq, pm, gmask, gl, P1P1, P2mP1, quirk, d0), gmask, gl);
}

template <int NR>
__device__ __forceinline__ void sgm_reset(unsigned (&L)[NR], unsigned& m) {
#pragma unroll
    for (int j = 0; j < NR; j++) L[j] = 0u;
    m = 0u;
}

// dynamic element fetch from a register array (k in [0, 2*NR)), kept in registers (mask-and-or, no local memory)
template <int NR>
__device__ __forceinline__ int elem_of(const unsigned (&v)[NR], int k) {
    unsigned r = 0u;
    const int kk = k >> 1;
#pragma unroll
    for (int j = 0; j < NR; j++) r |= v[j] & (0u - (unsigned)(kk == j));
    return (k & 1) ? (int)(r >> 16) : (int)(r & 0xffffu);
}
