// vert.cu — the three paths of a vertical family ((x-1,y-+1), (x,y-+1), (x+1,y-+1), A.5) in ONE pass:
// C' is read once and the family sum Sv is written once (4 B/cell instead of 16 B/cell for three
// read-modify-write launches of the generic line-scan kernel).
//
// The pass is row-synchronous: row y needs, per column x, L(x-1), L(x), L(x+1) of the previous row.
// One thread-block CLUSTER owns one frame; CTA r of the cluster owns a strip of NC columns and keeps
// the previous row's L of all three directions in shared memory (ping-pong for the diagonals, in
// place for the vertical path). The only data that crosses strips is one column of L per diagonal
// per row: the group that computes a strip-edge column writes it straight from registers into the
// neighbour CTA's shared memory (DSMEM, st.shared::cluster), then one cluster barrier per row.
// No global-memory halo, no inter-CTA spin, co-scheduling guaranteed by the cluster launch.
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

    uint4* A_right = nullptr; unsigned* mA_right = nullptr;   // neighbour CTA r+1
    uint4* B_left = nullptr;  unsigned* mB_left = nullptr;    // neighbour CTA r-1
    if (r + 1 < a.CS) { A_right = cluster.map_shared_rank(A, r + 1); mA_right = cluster.map_shared_rank(mA, r + 1); }
    if (r > 0) { B_left = cluster.map_shared_rank(Bd, r - 1); mB_left = cluster.map_shared_rank(mB, r - 1); }

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
                    // strip-edge columns go straight to the neighbour CTA's halo slots (DSMEM)
                    if (ic == NCr - 1 && A_right != nullptr) {
                        sts_regs<NR>(A_right + (size_t)q * (a.NC + 1) * COLQ + gl, LW, LA);
                        if (gl == 0) mA_right[q * (a.NC + 1)] = ma;
                    }
                    if (ic == 0 && B_left != nullptr) {
                        sts_regs<NR>(B_left + ((size_t)q * (a.NC + 1) + a.NC) * COLQ + gl, LW, LB);
                        if (gl == 0) mB_left[q * (a.NC + 1) + a.NC] = mb;
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
        cluster.sync();   // row t complete everywhere: halos delivered, ping-pong buffers may flip
    }
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
        size_t smem = ((size_t)4 * (NC + 1) + NC) * COLQ * sizeof(uint4) + ((size_t)4 * (NC + 1) + NC) * sizeof(unsigned) + 64;
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
