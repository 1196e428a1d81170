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

#include "sgm_step.cuh"

namespace cg = cooperative_groups;

struct VertArgs {
    const uint16_t* C;
    uint16_t* S;
    int H, W1, D, Dp, P1, P2;
    int NC;          // columns per CTA (last CTA of the cluster may own fewer)
    int CS;          // cluster size = strips per frame
    int dy;          // +1: top-down (predecessor row y-1), -1: bottom-up (predecessor row y+1)
    int accumulate;  // 0: S = sum, 1: S = sat(S + sum)
};

#define VERT_THREADS 512
#define VERT_KMAX 4

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

template <int NR>
__device__ __forceinline__ void lds_regs(const uint4* p, int LW, unsigned (&r)[NR]) {
#pragma unroll
    for (int q = 0; q < NR / 4; q++) {
        const uint4 t = p[q * LW];
        r[4 * q] = t.x; r[4 * q + 1] = t.y; r[4 * q + 2] = t.z; r[4 * q + 3] = t.w;
    }
}
template <int NR>
__device__ __forceinline__ void sts_regs(uint4* p, int LW, const unsigned (&r)[NR]) {
#pragma unroll
    for (int q = 0; q < NR / 4; q++) p[q * LW] = make_uint4(r[4 * q], r[4 * q + 1], r[4 * q + 2], r[4 * q + 3]);
}

// shared-memory column layout: [NR/4][LW] uint4 (16-byte chunk q of lane gl at q*LW + gl): conflict-free
template <int LW, int NR, bool PAD>
__global__ void __launch_bounds__(VERT_THREADS, 1) k_vert(const VertArgs a) {
    constexpr int DPL = 2 * NR;
    constexpr int G = VERT_THREADS / LW;      // column groups per CTA
    constexpr int COLQ = (NR / 4) * LW;       // uint4 per column
    cg::cluster_group cluster = cg::this_cluster();
    const int r = (int)cluster.block_rank();
    const int b = blockIdx.x / a.CS;
    const int gl = threadIdx.x % LW, g = threadIdx.x / LW;
    const unsigned gmask = 0xffffffffu;
    const int x0 = r * a.NC;
    const int NCr = min(a.NC, a.W1 - x0);     // >= 1 by construction of CS / NC
    const int KIT = (a.NC + G - 1) / G;

    extern __shared__ __align__(16) uint4 vsm[];
    // A[par][0..NC]  : L of the (x-1) diagonal; index c+1 = own column c, index 0 = halo from the left CTA
    // Bd[par][0..NC] : L of the (x+1) diagonal; index c = own column c, index NC = halo from the right CTA
    // U[0..NC-1]     : L of the vertical path, in place
    uint4* A = vsm;
    uint4* Bd = A + (size_t)2 * (a.NC + 1) * COLQ;
    uint4* U = Bd + (size_t)2 * (a.NC + 1) * COLQ;
    unsigned* mA = (unsigned*)(U + (size_t)a.NC * COLQ);
    unsigned* mB = mA + 2 * (a.NC + 1);
    unsigned* mU = mB + 2 * (a.NC + 1);
    unsigned long long* bars = (unsigned long long*)(mU + a.NC + (a.NC & 1));   // [4]: halo A slot 0/1, halo B slot 0/1

    const bool has_right = r + 1 < a.CS, has_left = r > 0;
    constexpr unsigned HALO_BYTES = (unsigned)COLQ * 16u + 4u;          // one column of L plus its minimum
    const unsigned barA = smem_u32(bars), barB = smem_u32(bars + 2);    // + 8 * slot
    // shared::cluster addresses of the neighbours' halo slots and barriers
    unsigned A_right = 0, mA_right = 0, barA_right = 0, B_left = 0, mB_left = 0, barB_left = 0;
    if (has_right) {
        A_right = map_cluster(smem_u32(A), r + 1); mA_right = map_cluster(smem_u32(mA), r + 1);
        barA_right = map_cluster(barA, r + 1);
    }
    if (has_left) {
        B_left = map_cluster(smem_u32(Bd), r - 1); mB_left = map_cluster(smem_u32(mB), r - 1);
        barB_left = map_cluster(barB, r - 1);
    }
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < 4; i++) mbar_init(smem_u32(bars + i), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        // arm both slots of each direction that will receive data (rows 0 and 1 of the neighbour)
        if (has_left && a.H > 1) { mbar_expect_tx(barA, HALO_BYTES); if (a.H > 2) mbar_expect_tx(barA + 8, HALO_BYTES); }
        if (has_right && a.H > 1) { mbar_expect_tx(barB, HALO_BYTES); if (a.H > 2) mbar_expect_tx(barB + 8, HALO_BYTES); }
    }

    const bool active = gl * DPL < a.Dp;
    unsigned pm[NR];
    make_padmask<NR>(pm, gl, a.D);
    const unsigned P1P1 = pk2((unsigned)a.P1), P2mP1 = (unsigned)(a.P2 - a.P1);
    const size_t frame = (size_t)b * a.H * a.W1 * a.Dp;

    // C of the next row, one register set per column iteration (software pipeline across the row barrier)
    unsigned Cb[VERT_KMAX][NR];
    {
        const int y = (a.dy > 0) ? 0 : a.H - 1;
#pragma unroll
        for (int k = 0; k < VERT_KMAX; k++) {
#pragma unroll
            for (int j = 0; j < NR; j++) Cb[k][j] = 0u;
            const int ic = min(k * G + g, NCr - 1);
            if (k < KIT && active) load_regs_nc<NR>(a.C + frame + ((size_t)y * a.W1 + x0 + ic) * a.Dp + gl * DPL, Cb[k]);
        }
    }
    cluster.sync();   // every CTA of the cluster is running before anyone writes remote shared memory

    for (int t = 0; t < a.H; t++) {
        const int y = (a.dy > 0) ? t : a.H - 1 - t;
        const int yn = (a.dy > 0) ? y + 1 : y - 1;          // next row (for the prefetch)
        const int q = t & 1, p = q ^ 1;
        const uint4* Ap = A + (size_t)p * (a.NC + 1) * COLQ;
        const uint4* Bp = Bd + (size_t)p * (a.NC + 1) * COLQ;
        uint4* Aq = A + (size_t)q * (a.NC + 1) * COLQ;
        uint4* Bq = Bd + (size_t)q * (a.NC + 1) * COLQ;
        const unsigned* mAp = mA + p * (a.NC + 1);
        const unsigned* mBp = mB + p * (a.NC + 1);
        unsigned* mAq = mA + q * (a.NC + 1);
        unsigned* mBq = mB + q * (a.NC + 1);
#pragma unroll
        for (int k = 0; k < VERT_KMAX; k++) {
            if (k < KIT) {
                const int i = k * G + g;
                const bool valid = i < NCr;
                const int ic = valid ? i : NCr - 1;           // surplus groups shadow the last column (uniform flow)
                const int x = x0 + ic;
                const unsigned (&Cq)[NR] = Cb[k];
                unsigned LA[NR], LB[NR], LU[NR], ma, mb, mu;
                // halo of row t-1 from the neighbours: only the warps that own a strip-edge column wait
                if (t > 0) {
                    const unsigned ph = (unsigned)((t - 1) >> 1) & 1u;   // slot p is in its ((t-1)/2)-th use
                    if (__any_sync(0xffffffffu, ic == 0 && has_left)) {
                        while (!mbar_try_wait(barA + 8 * p, ph)) {}
                        // re-arm the slot for row t+1 of the neighbour (rows H-1 are never sent)
                        if (ic == 0 && gl == 0 && valid && t + 1 < a.H - 1) mbar_expect_tx(barA + 8 * p, HALO_BYTES);
                    }
                    if (__any_sync(0xffffffffu, ic == NCr - 1 && has_right)) {
                        while (!mbar_try_wait(barB + 8 * p, ph)) {}
                        if (ic == NCr - 1 && gl == 0 && valid && t + 1 < a.H - 1) mbar_expect_tx(barB + 8 * p, HALO_BYTES);
                    }
                }
                // predecessor outside the cost domain (first row, or x-1 / x+1 off the edge): L = 0, m = 0
                if (t == 0 || x == 0) sgm_reset<NR>(LA, ma);
                else { lds_regs<NR>(Ap + (size_t)ic * COLQ + gl, LW, LA); ma = mAp[ic]; }
                if (t == 0 || x == a.W1 - 1) sgm_reset<NR>(LB, mb);
                else { lds_regs<NR>(Bp + (size_t)(ic + 1) * COLQ + gl, LW, LB); mb = mBp[ic + 1]; }
                if (t == 0) sgm_reset<NR>(LU, mu);
                else { lds_regs<NR>(U + (size_t)ic * COLQ + gl, LW, LU); mu = mU[ic]; }
                sgm_step<LW, NR, PAD>(LA, ma, Cq, pm, gmask, gl, P1P1, P2mP1, 0, 0);
                sgm_step<LW, NR, PAD>(LB, mb, Cq, pm, gmask, gl, P1P1, P2mP1, 0, 0);
                sgm_step<LW, NR, PAD>(LU, mu, Cq, pm, gmask, gl, P1P1, P2mP1, 0, 0);
                // C of the next row for this column slot: a whole row of work hides the latency
                if (active && t + 1 < a.H)
                    load_regs_nc<NR>(a.C + frame + ((size_t)yn * a.W1 + x) * a.Dp + gl * DPL, Cb[k]);
                if (valid) {
                    sts_regs<NR>(Aq + (size_t)(ic + 1) * COLQ + gl, LW, LA);
                    sts_regs<NR>(Bq + (size_t)ic * COLQ + gl, LW, LB);
                    sts_regs<NR>(U + (size_t)ic * COLQ + gl, LW, LU);
                    if (gl == 0) { mAq[ic + 1] = ma; mBq[ic] = mb; mU[ic] = mu; }
                    // strip-edge columns go straight to the neighbour CTA's halo slots (DSMEM, st.async + tx bytes)
                    if (t + 1 < a.H) {
                        if (ic == NCr - 1 && has_right) {
                            const unsigned dst = A_right + (unsigned)((q * (a.NC + 1)) * COLQ + gl) * 16u;
#pragma unroll
                            for (int c = 0; c < NR / 4; c++)
                                st_async_v4(dst + (unsigned)(c * LW) * 16u,
                                            make_uint4(LA[4 * c], LA[4 * c + 1], LA[4 * c + 2], LA[4 * c + 3]), barA_right + 8 * q);
                            if (gl == 0) st_async_b32(mA_right + (unsigned)(q * (a.NC + 1)) * 4u, ma, barA_right + 8 * q);
                        }
                        if (ic == 0 && has_left) {
                            const unsigned dst = B_left + (unsigned)((q * (a.NC + 1) + a.NC) * COLQ + gl) * 16u;
#pragma unroll
                            for (int c = 0; c < NR / 4; c++)
                                st_async_v4(dst + (unsigned)(c * LW) * 16u,
                                            make_uint4(LB[4 * c], LB[4 * c + 1], LB[4 * c + 2], LB[4 * c + 3]), barB_left + 8 * q);
                            if (gl == 0) st_async_b32(mB_left + (unsigned)(q * (a.NC + 1) + a.NC) * 4u, mb, barB_left + 8 * q);
                        }
                    }
                    if (active) {
                        uint16_t* sp = a.S + frame + ((size_t)y * a.W1 + x) * a.Dp + gl * DPL;
                        unsigned So[NR];
#pragma unroll
                        for (int j = 0; j < NR; j++) So[j] = uaddmin2(uaddmin2(LA[j], LU[j], 0x7fff7fffu), LB[j], 0x7fff7fffu);
                        if (a.accumulate) {
                            unsigned Sold[NR];
                            load_regs<NR>(sp, Sold);
#pragma unroll
                            for (int j = 0; j < NR; j++) So[j] = uaddmin2(So[j], Sold[j], 0x7fff7fffu);
                        }
                        store_regs<NR>(sp, So);
                    }
                }
            }
        }
        __syncthreads();   // row t complete in this CTA: ping-pong buffers may flip (halos are ordered by the mbarriers)
    }
    cluster.sync();        // nobody leaves while a neighbour could still address its shared memory
}

// ---- launcher ------------------------------------------------------------------------------------
struct VertPlan { int CS, NC, clusters; size_t smem; };

template <int LW, int NR>
static cudaLaunchConfig_t vert_cfg(int CS, int B, size_t smem, cudaStream_t st, cudaLaunchAttribute* attr) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(CS * B), 1, 1);
    cfg.blockDim = dim3(VERT_THREADS, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)CS; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    return cfg;
}

// Pick the cluster size (strips per frame) that keeps most of the chip busy: co-resident clusters (asked
// from the driver) x cluster size x fill of the column slots, among the sizes whose strip state fits.
template <int LW, int NR>
static bool plan_vert(const Eff& e, VertPlan& best) {
    constexpr int G = VERT_THREADS / LW;
    constexpr int COLQ = (NR / 4) * LW;
    double best_score = 0.0;
    for (int CS = 1; CS <= 16; CS++) {
        int NC = (e.W1 + CS - 1) / CS;
        if ((CS - 1) * NC >= e.W1) continue;                       // last strip would be empty
        size_t smem = ((size_t)4 * (NC + 1) + NC) * COLQ * sizeof(uint4) + ((size_t)4 * (NC + 1) + NC + 1) * sizeof(unsigned) + 4 * sizeof(unsigned long long) + 64;
        int KIT = (NC + G - 1) / G;
        if (smem > 225 * 1024 || KIT > VERT_KMAX) continue;
        cudaLaunchAttribute attr[1];
        cudaLaunchConfig_t cfg = vert_cfg<LW, NR>(CS, 1, smem, 0, attr);
        int ncl = 0;
        if (cudaOccupancyMaxActiveClusters(&ncl, k_vert<LW, NR, true>, &cfg) != cudaSuccess) { cudaGetLastError(); continue; }
        if (ncl <= 0) continue;
        double score = (double)ncl * e.W1 / ((double)KIT * G);     // useful column slots in flight
        if (score > best_score) { best_score = score; best.CS = CS; best.NC = NC; best.smem = smem; best.clusters = ncl; }
    }
    return best_score > 0.0;
}

template <int LW, int NR>
static int launch_vert_t(const Eff& e, int B, const uint16_t* C, uint16_t* S, int dy, int accumulate, cudaStream_t st,
                         int* clusters_out) {
    static bool configured = false;
    if (!configured) {
        if (cudaFuncSetAttribute(k_vert<LW, NR, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess) { cudaGetLastError(); return 1; }
        if (cudaFuncSetAttribute(k_vert<LW, NR, true>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) { cudaGetLastError(); return 1; }
        if (cudaFuncSetAttribute(k_vert<LW, NR, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024) != cudaSuccess) { cudaGetLastError(); return 1; }
        if (cudaFuncSetAttribute(k_vert<LW, NR, false>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) { cudaGetLastError(); return 1; }
        configured = true;
    }
    // the plan depends on (W1, Dp) only; cache the last one
    static VertPlan cached; static int cW1 = -1, cDp = -1; static bool cok = false;
    if (cW1 != e.W1 || cDp != e.Dp) {
        cok = plan_vert<LW, NR>(e, cached); cW1 = e.W1; cDp = e.Dp;
        if (getenv("B200S_VERBOSE") && cok)
            fprintf(stderr, "[b200s] k_vert<%d,%d>: W1=%d Dp=%d -> cluster %d x %d cols, %zu B smem, %d clusters co-resident\n",
                    LW, NR, e.W1, e.Dp, cached.CS, cached.NC, cached.smem, cached.clusters);
    }
    if (!cok) return 1;                                            // does not fit: caller falls back to k_path
    if (clusters_out) *clusters_out = cached.clusters;
    if (B <= 0) return 0;                                          // plan query only
    VertArgs a;
    a.C = C; a.S = S; a.H = e.H; a.W1 = e.W1; a.D = e.D; a.Dp = e.Dp; a.P1 = e.P1e; a.P2 = e.P2e;
    a.NC = cached.NC; a.CS = cached.CS; a.dy = dy; a.accumulate = accumulate;
    cudaLaunchAttribute attr[1];
    cudaLaunchConfig_t cfg = vert_cfg<LW, NR>(cached.CS, B, cached.smem, st, attr);
    cudaError_t le = (e.D == LW * 2 * NR) ? cudaLaunchKernelEx(&cfg, k_vert<LW, NR, false>, a)
                                          : cudaLaunchKernelEx(&cfg, k_vert<LW, NR, true>, a);
    if (le != cudaSuccess) { cudaGetLastError(); return 1; }
    return 0;
}

// returns 0 when launched, 1 when the configuration does not fit this kernel (use the generic kernel).
// B == 0 only plans; *clusters_out = frames this kernel processes concurrently (preferred batch granule).
int launch_vert_family(const Eff& e, int B, const uint16_t* C, uint16_t* S, int dy, int accumulate, cudaStream_t st,
                       int* clusters_out) {
    static int disabled = -1;
    if (disabled < 0) { const char* v = getenv("B200S_NO_VERT"); disabled = (v && v[0] == '1') ? 1 : 0; }
    if (disabled) return 1;
    if (e.Dp % 8 != 0 || e.Dp > 512) return 1;
    if (e.Dp > 64) {               // 16 disparities per lane
        int LW = 8;
        while (LW * 16 < e.Dp) LW *= 2;
        switch (LW) {
            case 8: return launch_vert_t<8, 8>(e, B, C, S, dy, accumulate, st, clusters_out);
            case 16: return launch_vert_t<16, 8>(e, B, C, S, dy, accumulate, st, clusters_out);
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
