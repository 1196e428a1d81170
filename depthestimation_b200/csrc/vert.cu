// vert.cu — the three paths of a vertical family ((x-1,y-+1), (x,y-+1), (x+1,y-+1), A.5) in ONE pass:
// C' is read once and the family sum Sv is written once (4 B/cell instead of 16 B/cell for three
// read-modify-write launches of the generic line-scan kernel).
//
// The pass is row-synchronous: row y needs, per column x, L(x-1), L(x), L(x+1) of the previous row.
// One thread-block CLUSTER owns one frame; CTA r of the cluster owns a strip of NC columns and keeps
// the previous row's L of all three directions in shared memory (ping-pong for the diagonals, in
// place for the vertical path). The only data that crosses strips is one column of L per diagonal
// per row: the group that computes a strip-edge column writes it straight from registers into the
// neighbour CTA's shared memory with st.async (DSMEM), which also completes transaction bytes on an mbarrier in
// the neighbour: the group that consumes the halo one row later waits on that mbarrier only. Inside a CTA one
// __syncthreads per row orders the rows. There is NO cluster-wide barrier per row: barrier.cluster.arrive.release
// would wait for every outstanding global store of S (measured: 40 % of all stall samples were membar waits).
// Two halo slots per direction suffice: a CTA can only run one row ahead of its neighbour, because it needs the
// neighbour's halo of row t to start row t + 1, and the neighbour sends that only after consuming ours of row t - 1.
// No global-memory halo, co-scheduling guaranteed by the cluster launch.
//
// Replaces part of the aggregation inside cv2.StereoSGBM.compute (depthlib/stereo_core.py:231) for
// MODE_SGBM (top-down family) and MODE_HH (top-down, then bottom-up with accumulate).
#include <cooperative_groups.h>
#include <stdlib.h>

#include <mutex>
#include <type_traits>
#include <vector>

#include "sgm_step.cuh"

namespace cg = cooperative_groups;

struct VertArgs {
    const uint16_t* C;
    uint16_t* S;
    int H, W1, D, Dp, P1, P2;
    int NC;          // columns per CTA (last CTA of the cluster may own fewer)
    int KIT, nrec;   // column slots per thread group, shared-memory records (NC + 1: own columns + the right halo)
    int CS;          // cluster size = strips per frame
    int dy;          // +1: top-down (predecessor row y-1), -1: bottom-up (predecessor row y+1)
    int accumulate;  // 0: S = sum, 1: S = sat(S + sum)
    int nosat;       // the sum of all paths cannot reach 32768: plain adds
};

// threads per CTA: 512 (one CTA per SM, wide strips) or 288 (two CTAs per SM, narrow strips: two row pipelines per SM
// that drift against each other and fill each other's barrier / latency gaps); picked by the planner
#define VERT_THREADS_WIDE 512
#define VERT_THREADS_NARROW 288
#define VERT_THREADS_XWIDE 576   // latency plan only: 72 column groups per CTA, so a 16-CTA cluster covers 1152 columns in ONE slot
#define VERT_KMAX 4
#ifndef VERT_PAIRSYNC
#define VERT_PAIRSYNC 1   // 1: rows are ordered warp to neighbouring warp through progress counters; 0: one CTA barrier per row
#endif
#define VERT_TAIL_BYTES (4 * sizeof(unsigned long long) + 32 * sizeof(unsigned))   // 4 halo mbarriers + per-warp row counters
#define VERT_MAX_DEVICES 64
#ifndef VERT_UREG
#define VERT_UREG 1   // 1: keep the vertical path's L in registers instead of shared memory
#endif

// ---- mbarrier / st.async helpers (sm_90+ PTX) ---------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned map_cluster(unsigned addr, unsigned rank) {
    unsigned r; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank)); return r;
}
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned bar, unsigned parity) {
    unsigned ok;
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void st_async_v4(unsigned raddr, const uint4 v, unsigned rbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [%0], {%1, %2, %3, %4}, [%5];"
                 ::"r"(raddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "r"(rbar) : "memory");
}
__device__ __forceinline__ void st_async_b32(unsigned raddr, const unsigned v, unsigned rbar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];"
                 ::"r"(raddr), "r"(v), "r"(rbar) : "memory");
}

// Shared memory: one RECORD per column index 0..nrec-1, all offsets compile-time:
//   A[2]  L of the (x-1) diagonal, ping-pong by row parity; record i holds own column i-1 (record 0 = halo from the left CTA)
//   B[2]  L of the (x+1) diagonal, ping-pong; record i holds own column i (record NC = halo from the right CTA)
//   U     L of the vertical path, in place; record i holds own column i
//   m[5]  the minima of A[0], A[1], B[0], B[1], U
// A field is [NR/4][LW] uint4 (16-byte chunk c of lane gl at c*LW + gl): conflict-free for LDS.128 / STS.128.
// Everything starts zeroed, which IS the reset state (L = 0, m = 0) of the first row and of the predecessors
// outside the cost domain (x-1 < 0, x+1 >= W1): those slots are never written, so no special cases remain.
template <int LW, int NR, int VT>
struct VertLayout {
    static constexpr int COLQ = (NR / 4) * LW;       // uint4 per field
    static constexpr int FLD = COLQ * 16;            // bytes per field
    // (with VERT_UREG the vertical path's L lives in registers: no U field, records are a fifth smaller)
    static constexpr int NFLD = VERT_UREG ? 4 : 5;
    static constexpr int OFF_A = 0, OFF_B = 2 * FLD, OFF_U = 4 * FLD, OFF_M = NFLD * FLD;
    static constexpr int REC = NFLD * FLD + 32;      // bytes per record
    static constexpr int G = VT / 32 * LaneGroup<LW>::RPW;   // column groups per CTA (whole lane groups per warp, sgm_step.cuh)
};

// CL = true: the strip state is sized NC + 1 records instead of KIT * G + 1 and the surplus groups of the last column
// slot are redirected to the strip's last column (a run-time record offset instead of an immediate: measurably slower,
// so it is only used for shapes whose state would not fit otherwise, e.g. 1920 wide with 256 disparities)
template <int LW, int NR, bool PAD, int VT, bool CL>
__global__ void __launch_bounds__(VT, (VT <= 320 ? 2 : 1)) k_vert(const VertArgs a) {
    using VL = VertLayout<LW, NR, VT>;
    constexpr int DPL = 2 * NR, G = VL::G, FLD = VL::FLD, REC = VL::REC;
    constexpr int OFF_A = VL::OFF_A, OFF_B = VL::OFF_B, OFF_U = VL::OFF_U, OFF_M = VL::OFF_M;
    cg::cluster_group cluster = cg::this_cluster();
    const int r = (int)cluster.block_rank();
    const int b = blockIdx.x / a.CS;
    using LG = LaneGroup<LW>;
    const int lane = threadIdx.x & 31;
    const bool idle = !LG::POW2 && lane >= LG::USED;            // lanes beyond the warp's last whole group: no loads, no stores
    const int gl = LG::POW2 ? (int)threadIdx.x % LW : (idle ? lane - LG::USED : lane % LW);
    const int g = LG::POW2 ? (int)threadIdx.x / LW : ((int)threadIdx.x >> 5) * LG::RPW + (idle ? LG::RPW - 1 : lane / LW);
    const unsigned gmask = 0xffffffffu;
    const int x0 = r * a.NC;
    const int NCr = min(a.NC, a.W1 - x0);     // >= 1 by construction of CS / NC
    const int KIT = a.KIT;

    extern __shared__ __align__(16) unsigned char vsm[];
    unsigned long long* bars = (unsigned long long*)(vsm + (size_t)a.nrec * REC);   // [4]: halo A slot 0/1, halo B slot 0/1
    // rows finished by every warp of the CTA. A column only exchanges L with the columns beside it, which belong to
    // the same warp or to the warp before / after it (cyclically: a thread's next column slot continues at group 0),
    // so a warp may start row t + 1 as soon as it and its two neighbours have finished row t: reads of row t - 1 data
    // are over (the slot it overwrites) and row t data is there. Warps drift by a row against each other instead of
    // all meeting at a CTA barrier per row (measured: 13 % of the stall samples were barrier waits, and the warps then
    // hit the ALU pipe all at once).
    unsigned* const done = (unsigned*)(bars + 4);
    const int wid = (int)threadIdx.x >> 5, NWARP = VT / 32;
    const unsigned done_l = smem_u32(done + (wid + NWARP - 1) % NWARP), done_r = smem_u32(done + (wid + 1) % NWARP), done_me = smem_u32(done + wid);
    if (threadIdx.x < 32) done[threadIdx.x] = 0u;
    for (int i = threadIdx.x; i < a.nrec * (REC / 16); i += VT) ((uint4*)vsm)[i] = make_uint4(0u, 0u, 0u, 0u);
    unsigned char* const tb = vsm + g * REC + gl * 16;   // this thread's chunk 0 of field 0 of record g
    unsigned char* const tm = vsm + g * REC + OFF_M;     // minima of record g

    const int lastrec = NCr - 1 - g;                     // record offset (relative to g) of the strip's last column
    const bool has_right = r + 1 < a.CS, has_left = r > 0;
    constexpr unsigned HALO_BYTES = (unsigned)FLD + 4u;                 // one column of L plus its minimum
    const unsigned barA = smem_u32(bars), barB = smem_u32(bars + 2);    // + 8 * slot
    // shared::cluster addresses of the neighbours' halo records and barriers
    unsigned rec_right = 0, barA_right = 0, rec_left = 0, barB_left = 0;
    if (has_right) { rec_right = map_cluster(smem_u32(vsm), r + 1); barA_right = map_cluster(barA, r + 1); }
    if (has_left) { rec_left = map_cluster(smem_u32(vsm) + (unsigned)a.NC * REC, r - 1); barB_left = map_cluster(barB, r - 1); }
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < 4; i++) mbar_init(smem_u32(bars + i), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        // arm both slots of each direction that will receive data (rows 0 and 1 of the neighbour)
        if (has_left && a.H > 1) { mbar_expect_tx(barA, HALO_BYTES); if (a.H > 2) mbar_expect_tx(barA + 8, HALO_BYTES); }
        if (has_right && a.H > 1) { mbar_expect_tx(barB, HALO_BYTES); if (a.H > 2) mbar_expect_tx(barB + 8, HALO_BYTES); }
    }

    const bool active = (PAD ? (gl * DPL < a.Dp) : true) && !idle;
    unsigned pm[NR];
    make_padmask<NR>(pm, gl, a.D);
    const unsigned P1P1 = pk2((unsigned)a.P1), P2mP1 = (unsigned)(a.P2 - a.P1);
    const uint16_t* const Cf = a.C + (size_t)b * a.H * a.W1 * a.Dp;
    uint16_t* const Sf = a.S + (size_t)b * a.H * a.W1 * a.Dp;
    const unsigned rowstep = (unsigned)(a.dy * a.W1 * a.Dp);           // wraps for dy < 0; offsets are modulo 2^32

    // element offset (within the frame) of this thread's 16-byte chunk 0, per column slot; C of the next row is
    // held in one register set per column slot (software pipeline across the row boundary)
    unsigned off[VERT_KMAX];
    unsigned Cb[VERT_KMAX][NR];
#if VERT_UREG
    unsigned Ur[VERT_KMAX][NR], mUr[VERT_KMAX];   // L and minimum of the vertical path never leave the thread
#pragma unroll
    for (int k = 0; k < VERT_KMAX; k++) {
        mUr[k] = 0u;
#pragma unroll
        for (int j = 0; j < NR; j++) Ur[k][j] = 0u;
    }
#endif
    {
        const int y = (a.dy > 0) ? 0 : a.H - 1;
#pragma unroll
        for (int k = 0; k < VERT_KMAX; k++) {
#pragma unroll
            for (int j = 0; j < NR; j++) Cb[k][j] = 0u;
            const int ic = min(k * G + g, NCr - 1);                      // surplus groups shadow the last column
            off[k] = (unsigned)((y * a.W1 + x0 + ic) * a.Dp + gl * DPL);
            if (k < KIT && active) load_regs_nc<NR>(Cf + off[k], Cb[k]);
        }
    }
    __syncthreads();
    cluster.sync();   // every CTA of the cluster is running, zeroed and armed before anyone writes remote shared memory

    // one row; Q = parity of t (written slot), P = 1 - Q (slot of the previous row)
    auto row = [&](auto Qc, const int t) {
        constexpr int Q = decltype(Qc)::value, P = Q ^ 1;
#pragma unroll
        for (int k = 0; k < VERT_KMAX; k++) {
            if (k < KIT) {
                const int i = k * G + g;
                const bool valid = i < NCr && !idle;
                const int ro = CL ? min(k * G, lastrec) * REC : k * G * REC;   // CL: surplus groups shadow record NCr-1
                unsigned char* const rb = tb + ro;
                unsigned char* const rm = tm + ro;
                const unsigned (&Cq)[NR] = Cb[k];
                unsigned LA[NR], LB[NR], LU[NR];
                // halo of row t-1 from the neighbours: only the warps that own a strip-edge column wait
                if (t > 0) {
                    const unsigned ph = (unsigned)((t - 1) >> 1) & 1u;   // slot P is in its ((t-1)/2)-th use
                    const bool eL = (i == 0) && has_left, eR = (i == NCr - 1) && has_right;
                    if (__any_sync(0xffffffffu, eL)) {
                        while (!mbar_try_wait(barA + 8 * P, ph)) {}
                        // re-arm the slot for row t+1 of the neighbour (row H-1 is never sent)
                        if (eL && gl == 0 && !idle && t + 1 < a.H - 1) mbar_expect_tx(barA + 8 * P, HALO_BYTES);
                    }
                    if (__any_sync(0xffffffffu, eR)) {
                        while (!mbar_try_wait(barB + 8 * P, ph)) {}
                        if (eR && gl == 0 && !idle && t + 1 < a.H - 1) mbar_expect_tx(barB + 8 * P, HALO_BYTES);
                    }
                }
#pragma unroll
                for (int c = 0; c < NR / 4; c++) {
                    const uint4 va = *(const uint4*)(rb + OFF_A + P * FLD + c * LW * 16);
                    const uint4 vb = *(const uint4*)(rb + REC + OFF_B + P * FLD + c * LW * 16);
#if !VERT_UREG
                    const uint4 vu = *(const uint4*)(rb + OFF_U + c * LW * 16);
#endif
                    LA[4 * c] = va.x; LA[4 * c + 1] = va.y; LA[4 * c + 2] = va.z; LA[4 * c + 3] = va.w;
                    LB[4 * c] = vb.x; LB[4 * c + 1] = vb.y; LB[4 * c + 2] = vb.z; LB[4 * c + 3] = vb.w;
#if !VERT_UREG
                    LU[4 * c] = vu.x; LU[4 * c + 1] = vu.y; LU[4 * c + 2] = vu.z; LU[4 * c + 3] = vu.w;
#endif
                }
                unsigned ma = *(const unsigned*)(rm + P * 4);
                unsigned mb = *(const unsigned*)(rm + REC + 8 + P * 4);
#if VERT_UREG
                unsigned mu = mUr[k];
#pragma unroll
                for (int j = 0; j < NR; j++) LU[j] = Ur[k][j];
#else
                unsigned mu = *(const unsigned*)(rm + 16);
#endif
                sgm_step<LW, NR, PAD>(LA, ma, Cq, pm, gmask, gl, P1P1, P2mP1, 0, 0);
                sgm_step<LW, NR, PAD>(LB, mb, Cq, pm, gmask, gl, P1P1, P2mP1, 0, 0);
                sgm_step<LW, NR, PAD>(LU, mu, Cq, pm, gmask, gl, P1P1, P2mP1, 0, 0);
                const unsigned o = off[k];
                off[k] = o + rowstep;
                // C of the next row for this column slot: a whole row of work hides the latency
                if (active && t + 1 < a.H) load_regs_nc<NR>(Cf + off[k], Cb[k]);
#if VERT_UREG
                mUr[k] = mu;
#pragma unroll
                for (int j = 0; j < NR; j++) Ur[k][j] = LU[j];
#endif
                if (valid) {
#pragma unroll
                    for (int c = 0; c < NR / 4; c++) {
                        *(uint4*)(rb + REC + OFF_A + Q * FLD + c * LW * 16) = make_uint4(LA[4 * c], LA[4 * c + 1], LA[4 * c + 2], LA[4 * c + 3]);
                        *(uint4*)(rb + OFF_B + Q * FLD + c * LW * 16) = make_uint4(LB[4 * c], LB[4 * c + 1], LB[4 * c + 2], LB[4 * c + 3]);
#if !VERT_UREG
                        *(uint4*)(rb + OFF_U + c * LW * 16) = make_uint4(LU[4 * c], LU[4 * c + 1], LU[4 * c + 2], LU[4 * c + 3]);
#endif
                    }
                    if (gl == 0) {
                        *(unsigned*)(rm + REC + Q * 4) = ma;
                        *(unsigned*)(rm + 8 + Q * 4) = mb;
#if !VERT_UREG
                        *(unsigned*)(rm + 16) = mu;
#endif
                    }
                    // strip-edge columns go straight to the neighbour CTA's halo records (DSMEM, st.async + tx bytes)
                    if (t + 1 < a.H) {
                        if (i == NCr - 1 && has_right) {     // -> record 0, field A[Q] of the right neighbour
                            const unsigned dst = rec_right + (unsigned)(OFF_A + Q * FLD + gl * 16);
#pragma unroll
                            for (int c = 0; c < NR / 4; c++)
                                st_async_v4(dst + (unsigned)(c * LW * 16),
                                            make_uint4(LA[4 * c], LA[4 * c + 1], LA[4 * c + 2], LA[4 * c + 3]), barA_right + 8 * Q);
                            if (gl == 0) st_async_b32(rec_right + (unsigned)(OFF_M + Q * 4), ma, barA_right + 8 * Q);
                        }
                        if (i == 0 && has_left) {            // -> record NC, field B[Q] of the left neighbour
                            const unsigned dst = rec_left + (unsigned)(OFF_B + Q * FLD + gl * 16);
#pragma unroll
                            for (int c = 0; c < NR / 4; c++)
                                st_async_v4(dst + (unsigned)(c * LW * 16),
                                            make_uint4(LB[4 * c], LB[4 * c + 1], LB[4 * c + 2], LB[4 * c + 3]), barB_left + 8 * Q);
                            if (gl == 0) st_async_b32(rec_left + (unsigned)(OFF_M + 8 + Q * 4), mb, barB_left + 8 * Q);
                        }
                    }
                    if (active) {
                        uint16_t* sp = Sf + o;
                        unsigned So[NR];
#pragma unroll
                        for (int j = 0; j < NR; j++)
#if SGM_FMA_ADDS
                            So[j] = a.nosat ? __vadd2(__vadd2(LA[j], LU[j]), LB[j]) : uaddmin2(uaddmin2(LA[j], LU[j], 0x7fff7fffu), LB[j], 0x7fff7fffu);
#else
                            So[j] = a.nosat ? LA[j] + LU[j] + LB[j] : uaddmin2(uaddmin2(LA[j], LU[j], 0x7fff7fffu), LB[j], 0x7fff7fffu);
#endif
                        if (a.accumulate) {
                            unsigned Sold[NR];
                            load_regs<NR>(sp, Sold);
#pragma unroll
                            for (int j = 0; j < NR; j++) So[j] = a.nosat ? So[j] + Sold[j] : uaddmin2(So[j], Sold[j], 0x7fff7fffu);
                        }
                        store_regs<NR>(sp, So);
                    }
                }
            }
        }
#if VERT_PAIRSYNC == 2
        // progress counters in shared memory polled by lane 0 (measured: 4x SLOWER, 7.8 -> 31.5 ms per 150 frames: the
        // polling warps take the issue slots and the shared-memory port of the working ones)
        __syncwarp();
        if (lane == 0) {
            const unsigned row1 = (unsigned)t + 1u;
            unsigned v;
            asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(done_me), "r"(row1) : "memory");
            do { asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"(done_l) : "memory"); } while (v < row1);
            do { asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"(done_r) : "memory"); } while (v < row1);
        }
        __syncwarp();
#elif VERT_PAIRSYNC == 1
        // row t complete in this warp: meet the two neighbouring warps on hardware barriers (one per pair of adjacent
        // warps, ids 0..15: only for an even number of at most 16 warps, where the ring of pairs can be walked in two
        // rounds without a cycle); otherwise the whole CTA
        if constexpr (VT / 32 <= 16 && (VT / 32) % 2 == 0) {
            constexpr int NW_ = VT / 32;
            const int lo = (wid + NW_ - 1) % NW_;                       // pair (wid-1, wid) has id wid-1, pair (wid, wid+1) id wid
            const int first = (wid & 1) ? lo : wid, second = (wid & 1) ? wid : lo;
            asm volatile("bar.sync %0, 64;" ::"r"(first) : "memory");
            asm volatile("bar.sync %0, 64;" ::"r"(second) : "memory");
        } else {
            __syncthreads();
        }
#else
        __syncthreads();   // row t complete in this CTA: ping-pong fields may flip (halos are ordered by the mbarriers)
#endif
    };
    for (int t = 0; t < a.H; t += 2) {
        row(std::integral_constant<int, 0>{}, t);
        if (t + 1 < a.H) row(std::integral_constant<int, 1>{}, t + 1);
    }
    cluster.sync();        // nobody leaves while a neighbour could still address its shared memory
}

// ---- launcher ------------------------------------------------------------------------------------
struct VertPlan { int CS, NC, KIT, clusters, threads, clamp; size_t smem; double score; };

static cudaLaunchConfig_t vert_cfg(int threads, int CS, int B, size_t smem, cudaStream_t st, cudaLaunchAttribute* attr) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(CS * B), 1, 1);
    cfg.blockDim = dim3((unsigned)threads, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)CS; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    return cfg;
}

// Pick the cluster size (strips per frame) for one CTA width: frames in flight (co-resident clusters, asked from the
// driver) per column iteration of a thread, among the sizes whose strip state fits; ties go to the fuller column slots.
// latency = true: the plan for a few frames per call instead (every cluster resident at once anyway): the fewest
// column slots per CTA and row, i.e. the frame spread over as many SMs as a cluster allows. A row of a strip costs its
// CTA ~24 ALU-pipe cycles per column and direction, so one KITTI frame on 9 SMs is bound by those 9 SMs' issue rate
// (measured: 831 us per frame at one frame per call against 53 us per frame in a full batch).
template <int LW, int NR, int VT>
static bool plan_vert(const Eff& e, VertPlan& best, bool latency = false, std::vector<VertPlan>* all = nullptr) {
    using VL = VertLayout<LW, NR, VT>;
    constexpr int G = VL::G;
    // function attributes are per DEVICE: opt every device in once (a process may drive several GPUs from several threads)
    static bool configured[VERT_MAX_DEVICES] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= VERT_MAX_DEVICES) { cudaGetLastError(); return false; }
    if (!configured[dev]) {
        auto cfgk = [](auto kern) {
            if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess) { cudaGetLastError(); return false; }
            if (cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) { cudaGetLastError(); return false; }
            return true;
        };
        if (!cfgk(k_vert<LW, NR, true, VT, false>) || !cfgk(k_vert<LW, NR, false, VT, false>) ||
            !cfgk(k_vert<LW, NR, true, VT, true>) || !cfgk(k_vert<LW, NR, false, VT, true>)) return false;
        configured[dev] = true;
    }
    static int cs_env = -1;
    if (cs_env < 0) { const char* v = getenv("B200S_VERT_CS"); cs_env = v ? atoi(v) : 0; }
    best.score = 0.0;
    for (int CS = 1; CS <= 16; CS++) {
        if (cs_env > 0 && CS != cs_env) continue;
        int NC = (e.W1 + CS - 1) / CS;
        if ((CS - 1) * NC >= e.W1) continue;                       // last strip would be empty
        int KIT = (NC + G - 1) / G;
        if (KIT > VERT_KMAX) continue;
        size_t smem = (size_t)(KIT * G + 1) * VL::REC + VERT_TAIL_BYTES;   // a record per group and slot + the right halo
        int clamp = 0;
        if (smem > 225 * 1024) {                                   // own columns + the right halo only (k_vert<.., CL = true>)
            smem = (size_t)(NC + 1) * VL::REC + VERT_TAIL_BYTES;
            clamp = 1;
        }
        if (smem > 225 * 1024) continue;
        cudaLaunchAttribute attr[1];
        cudaLaunchConfig_t cfg = vert_cfg(VT, CS, 1, smem, 0, attr);
        int ncl = 0;
        if (cudaOccupancyMaxActiveClusters(&ncl, k_vert<LW, NR, true, VT, false>, &cfg) != cudaSuccess) { cudaGetLastError(); continue; }
        if (ncl <= 0) continue;
        double score = (double)ncl / KIT * (clamp ? 0.8 : 1.0) + 1e-3 * (double)e.W1 / ((double)CS * KIT * G);
        if (latency) {
            score = 1.0 / ((double)KIT * (clamp ? 1.25 : 1.0)) + 1e-3 / ((double)KIT * G) + 1e-6 * CS;   // serial slots per thread first
            // one CTA per SM: two narrow CTAs of a cluster sharing an SM would share its issue slots again
            if (smem < 116 * 1024) smem = 116 * 1024;
        }
        if (all) { VertPlan p; p.score = score; p.CS = CS; p.NC = NC; p.KIT = KIT; p.smem = smem; p.clusters = ncl; p.threads = VT; p.clamp = clamp; all->push_back(p); }
        if (score > best.score) { best.score = score; best.CS = CS; best.NC = NC; best.KIT = KIT; best.smem = smem; best.clusters = ncl; best.threads = VT; best.clamp = clamp; }
    }
    return best.score > 0.0;
}

template <int LW, int NR, int VT>
static cudaError_t launch_vert_vt(const VertPlan& pl, const VertArgs& a, int B, bool pad, cudaStream_t st) {
    cudaLaunchAttribute attr[1];
    cudaLaunchConfig_t cfg = vert_cfg(VT, pl.CS, B, pl.smem, st, attr);
    if (pl.clamp) return pad ? cudaLaunchKernelEx(&cfg, k_vert<LW, NR, true, VT, true>, a) : cudaLaunchKernelEx(&cfg, k_vert<LW, NR, false, VT, true>, a);
    return pad ? cudaLaunchKernelEx(&cfg, k_vert<LW, NR, true, VT, false>, a) : cudaLaunchKernelEx(&cfg, k_vert<LW, NR, false, VT, false>, a);
}

template <int LW, int NR>
int launch_vert_t(const Eff& e, int B, const uint16_t* C, uint16_t* S, int dy, int accumulate, cudaStream_t st,
                  int* clusters_out) {
    // the plan depends on (device, W1, Dp) only; the last one per device is cached. Several host threads (one handle per
    // thread, possibly on different GPUs) come through here: the cache is guarded and used through a local copy.
    struct Cache { VertPlan plan, lat; std::vector<VertPlan> all; int W1 = -1, Dp = -1, sms = 0; bool ok = false, lat_ok = false; };
    static Cache cache[VERT_MAX_DEVICES];
    static std::mutex mu;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= VERT_MAX_DEVICES) { cudaGetLastError(); return 1; }
    VertPlan cached;
    bool cok;
    int sms = 0;
    {
        std::lock_guard<std::mutex> lk(mu);
        Cache& c = cache[dev];
        if (c.W1 != e.W1 || c.Dp != e.Dp) {
            static int vt_env = -1;
            if (vt_env < 0) { const char* v = getenv("B200S_VERT_VT"); vt_env = v ? atoi(v) : 0; }
            VertPlan wide, narrow;
            c.all.clear();
            const bool okw = (vt_env == 0 || vt_env == VERT_THREADS_WIDE) && plan_vert<LW, NR, VERT_THREADS_WIDE>(e, wide, false, &c.all);
            const bool okn = (vt_env == 0 || vt_env == VERT_THREADS_NARROW) && plan_vert<LW, NR, VERT_THREADS_NARROW>(e, narrow, false, &c.all);
            c.ok = okw || okn;
            if (c.ok) c.plan = (okn && (!okw || narrow.score > wide.score)) ? narrow : wide;
            VertPlan xwide;
            const bool lw = (vt_env == 0 || vt_env == VERT_THREADS_WIDE) && plan_vert<LW, NR, VERT_THREADS_WIDE>(e, wide, true);
            const bool ln = (vt_env == 0 || vt_env == VERT_THREADS_NARROW) && plan_vert<LW, NR, VERT_THREADS_NARROW>(e, narrow, true);
            const bool lx = (vt_env == 0 || vt_env == VERT_THREADS_XWIDE) && plan_vert<LW, NR, VERT_THREADS_XWIDE>(e, xwide, true);
            c.lat_ok = lw || ln || lx;
            if (c.lat_ok) {
                c.lat.score = -1.0;
                if (lw && wide.score > c.lat.score) c.lat = wide;
                if (ln && narrow.score > c.lat.score) c.lat = narrow;
                if (lx && xwide.score > c.lat.score) c.lat = xwide;
            }
            cudaDeviceGetAttribute(&c.sms, cudaDevAttrMultiProcessorCount, dev);
            c.W1 = e.W1; c.Dp = e.Dp;
            if (getenv("B200S_VERBOSE") && c.ok)
                fprintf(stderr, "[b200s] dev %d k_vert<%d,%d>: W1=%d Dp=%d -> %d threads, cluster %d x %d cols (%d per thread group), %zu B smem%s, %d clusters co-resident\n",
                        dev, LW, NR, e.W1, e.Dp, c.plan.threads, c.plan.CS, c.plan.NC, c.plan.KIT, c.plan.smem, c.plan.clamp ? " (clamped records)" : "", c.plan.clusters);
        }
        cached = c.plan;
        cok = c.ok;
        // The throughput plan maximises frames in flight per column slot, which is right when the call fills whole waves of
        // clusters. A call whose frame count is not a multiple of the co-resident cluster count pays a whole extra wave
        // for the remainder (16 KITTI pairs on 15 co-resident clusters: two waves): take the plan with the fewest
        // (waves x column slots per thread) for THIS frame count, the throughput score breaking ties.
        if (cok && B > 0 && B % c.plan.clusters != 0) {
            auto cost = [&](const VertPlan& p) { return (double)((B + p.clusters - 1) / p.clusters) * p.KIT * (p.clamp ? 1.25 : 1.0); };
            for (const VertPlan& p : c.all)
                if (cost(p) < cost(cached) - 1e-9 || (cost(p) < cost(cached) + 1e-9 && p.score > cached.score)) cached = p;
        }
        static int lat_env = -1;
        if (lat_env < 0) { const char* v = getenv("B200S_VERT_LATENCY"); lat_env = v ? atoi(v) : 1; }
        // few frames in the call (at most half the SMs busy even with the widest clusters) and a latency plan that
        // holds every column of its strip in ONE slot per thread: take it. Measured, KITTI pairs per call: 1 -> 1.97 /
        // 1.79 ms, 4 -> 2.37 / 2.18 ms per call (throughput plan / 16 CTAs x 576 threads); from 8 pairs on the throughput
        // plan wins, and a latency plan that still needs several slots loses (1920x1080 D=256: 10.5 -> 12.3 ms)
        if (cok && c.lat_ok && lat_env && B > 0 && 2LL * B * c.lat.CS <= c.sms && c.lat.KIT == 1 && c.plan.KIT > 1)
            cached = c.lat;
        sms = c.sms;
    }
    if (!cok) return 1;                                            // does not fit: caller falls back to k_path
    // The clamped-record plan (wide frames at D > 128: every column of a strip in shared memory, several per thread) on a
    // call that keeps under a third of the SMs busy: the generic line kernel, one launch per direction with the lines
    // spread over every SM, is faster there. Measured per pair, 1920x1080 D = 256 hh: 1 pair 9.14 -> 5.66 ms, 2 pairs
    // 4.63 -> 3.69, 4 pairs 2.34 -> 2.42 (cluster kernel kept). Unclamped plans stay: 1920x1080 D = 128 sgbm, one pair,
    // 2.07 (cluster) against 2.57 ms; KITTI 0.66 against 0.87.
    // (B200S_VERT_SMALL_GENERIC=0, read on every call so that a test can switch it, keeps the cluster kernel.)
    if (B > 0 && cached.clamp && 3LL * B * cached.CS <= sms) {
        const char* v = getenv("B200S_VERT_SMALL_GENERIC");
        if (!v || v[0] != '0') return 1;
    }
    if (clusters_out) *clusters_out = cached.clusters;
    if (B <= 0) return 0;                                          // plan query only
    VertArgs a;
    a.C = C; a.S = S; a.H = e.H; a.W1 = e.W1; a.D = e.D; a.Dp = e.Dp; a.P1 = e.P1e; a.P2 = e.P2e;
    a.NC = cached.NC; a.CS = cached.CS; a.dy = dy; a.accumulate = accumulate; a.nosat = (sum_cannot_saturate(e) && e.D == LW * 2 * NR) ? 1 : 0;   // padded halves (0x7fff) must keep saturating
    a.KIT = cached.KIT; a.nrec = cached.clamp ? cached.NC + 1 : cached.KIT * (cached.threads / 32 * LaneGroup<LW>::RPW) + 1;   // G column groups per CTA
    const bool pad = e.D != LW * 2 * NR;
    cudaError_t le = (cached.threads == VERT_THREADS_WIDE) ? launch_vert_vt<LW, NR, VERT_THREADS_WIDE>(cached, a, B, pad, st)
                   : (cached.threads == VERT_THREADS_XWIDE) ? launch_vert_vt<LW, NR, VERT_THREADS_XWIDE>(cached, a, B, pad, st)
                                                            : launch_vert_vt<LW, NR, VERT_THREADS_NARROW>(cached, a, B, pad, st);
    if (le != cudaSuccess) {   // the planner said it fits: a failed launch is an error, not a reason to fall back silently
        fprintf(stderr, "[b200s] k_vert launch failed on device %d: %s\n", dev, cudaGetErrorString(le));
        cudaGetLastError();
        return -1;
    }
    return 0;
}

// Build partitioning (Makefile): the lane-group instantiations are spread over six translation units
// (-DVERT_PART=1..6), the host side is part 0; without VERT_PART everything is in this one file.
#ifndef VERT_PART
#define VERT_PART -1
#endif
#define VERT_SIG const Eff&, int, const uint16_t*, uint16_t*, int, int, cudaStream_t, int*
#if VERT_PART == 1
template int launch_vert_t<2, 4>(VERT_SIG);
#elif VERT_PART == 0
extern template int launch_vert_t<2, 4>(VERT_SIG);
#endif
#if VERT_PART == 2
template int launch_vert_t<4, 4>(VERT_SIG);
#elif VERT_PART == 0
extern template int launch_vert_t<4, 4>(VERT_SIG);
#endif
#if VERT_PART == 3
template int launch_vert_t<8, 4>(VERT_SIG);
#elif VERT_PART == 0
extern template int launch_vert_t<8, 4>(VERT_SIG);
#endif
#if VERT_PART == 4
template int launch_vert_t<5, 8>(VERT_SIG);
#elif VERT_PART == 0
extern template int launch_vert_t<5, 8>(VERT_SIG);
#endif
#if VERT_PART == 5
template int launch_vert_t<6, 8>(VERT_SIG);
#elif VERT_PART == 0
extern template int launch_vert_t<6, 8>(VERT_SIG);
#endif
#if VERT_PART == 6
template int launch_vert_t<7, 8>(VERT_SIG);
#elif VERT_PART == 0
extern template int launch_vert_t<7, 8>(VERT_SIG);
#endif
#if VERT_PART == 1
template int launch_vert_t<8, 8>(VERT_SIG);
#elif VERT_PART == 0
extern template int launch_vert_t<8, 8>(VERT_SIG);
#endif
#if VERT_PART == 2
template int launch_vert_t<9, 8>(VERT_SIG);
#elif VERT_PART == 0
extern template int launch_vert_t<9, 8>(VERT_SIG);
#endif
#if VERT_PART == 3
template int launch_vert_t<10, 8>(VERT_SIG);
#elif VERT_PART == 0
extern template int launch_vert_t<10, 8>(VERT_SIG);
#endif
#if VERT_PART == 4
template int launch_vert_t<16, 8>(VERT_SIG);
#elif VERT_PART == 0
extern template int launch_vert_t<16, 8>(VERT_SIG);
#endif
#if VERT_PART == 5
template int launch_vert_t<32, 8>(VERT_SIG);
#elif VERT_PART == 0
extern template int launch_vert_t<32, 8>(VERT_SIG);
#endif
#if VERT_PART <= 0
// returns 0 when launched, 1 when the configuration does not fit this kernel (use the generic kernel), -1 when the
// launch itself failed (a CUDA error the caller must report).
// B == 0 only plans; *clusters_out = frames this kernel processes concurrently (preferred batch granule).
int launch_vert_family(const Eff& e, int B, const uint16_t* C, uint16_t* S, int dy, int accumulate, cudaStream_t st,
                       int* clusters_out) {
    static int disabled = -1;
    if (disabled < 0) { const char* v = getenv("B200S_NO_VERT"); disabled = (v && v[0] == '1') ? 1 : 0; }
    if (disabled) return 1;
    if (e.Dp % 8 != 0 || e.Dp > 512) return 1;
    if ((size_t)e.H * e.W1 * e.Dp >= (1ull << 31)) return 1;   // 32-bit element offsets inside a frame
    if (e.Dp > 64) {               // 16 disparities per lane, the lane count follows D (sgm_step.cuh, LaneGroup)
        switch ((e.Dp + 15) / 16) {
            case 5: return launch_vert_t<5, 8>(e, B, C, S, dy, accumulate, st, clusters_out);
            case 6: return launch_vert_t<6, 8>(e, B, C, S, dy, accumulate, st, clusters_out);
            case 7: return launch_vert_t<7, 8>(e, B, C, S, dy, accumulate, st, clusters_out);
            case 8: return launch_vert_t<8, 8>(e, B, C, S, dy, accumulate, st, clusters_out);
            case 9: return launch_vert_t<9, 8>(e, B, C, S, dy, accumulate, st, clusters_out);
            case 10: return launch_vert_t<10, 8>(e, B, C, S, dy, accumulate, st, clusters_out);
            case 11: case 12: case 13: case 14: case 15: case 16: return launch_vert_t<16, 8>(e, B, C, S, dy, accumulate, st, clusters_out);
            default: return launch_vert_t<32, 8>(e, B, C, S, dy, accumulate, st, clusters_out);
        }
    }
    int LW = 2;
    while (LW * 8 < e.Dp) LW *= 2;
    switch (LW) {
        case 2: return launch_vert_t<2, 4>(e, B, C, S, dy, accumulate, st, clusters_out);
        case 4: return launch_vert_t<4, 4>(e, B, C, S, dy, accumulate, st, clusters_out);
        default: return launch_vert_t<8, 4>(e, B, C, S, dy, accumulate, st, clusters_out);
    }
}
#endif  // VERT_PART <= 0
