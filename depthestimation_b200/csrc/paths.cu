// paths.cu — semi-global path aggregation (A.4, A.5, A.7), generic line-scan kernel.
//
// Replaces the aggregation loops of cv2.StereoSGBM.compute (depthlib/stereo_core.py:231).
//
// Recursion, rewritten so that the cost volume need not carry the reference's +P2:
//   L_r(p,d) = C'(p,d) + min(L_r(q,d), L_r(q,d-1)+P1, L_r(q,d+1)+P1, m+P2) - m,  m = min_d L_r(q,d)
// with C' = C - P2 (identical values to A.4). All terms are in [0, 65535], so the packed u16x2
// DPX instructions are exact:  t = VIMNMX3.U16x2(b, c, m+P2-P1); t = VIADDMNMX.U16x2(t, P1, a);
// L = C' + t - m (plain 32-bit add/sub: no half can carry or borrow).
//
// One "line" = one scan line of one direction; LW lanes cooperate on it, 8 disparities per lane
// (one 16-byte load of C per lane and step). Every line is independent of every other line:
//   horizontal dirs : line = image row
//   vertical dirs   : line = cost column (per sgbm_3way stripe)
//   diagonal dirs   : line g visits x = (g +- t) mod W1 at step t and restarts when it wraps —
//                     the predecessor is then outside the cost domain, which is exactly the
//                     reference's border condition (L = 0, m = 0)
// S is accumulated with a saturating add at 32767 (VIADDMNMX.U16x2).
#include "common.cuh"

struct PathArgs {
    const uint16_t* C;
    const uint16_t* Calt;
    uint16_t* S;
    int H, W1, D, Dp, P1, P2;
    int px, py;   // predecessor offset (A.5)
    int first;    // 1: S = L (S not read)
    int quirk;    // sgbm_3way D%8 tail artefact: 0 none, 1 vertical path, 2 px=+1 path (A.7)
    StripePlan sp;
};

template <int LW>
__global__ void __launch_bounds__(128) k_path(const PathArgs a) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    const int gl = threadIdx.x % LW;            // lane within the line group
    const int line = tid / LW;
    const int b = blockIdx.z;
    const bool horizontal = (a.py == 0);
    const int nlines = horizontal ? a.H : a.W1 * a.sp.n;
    if (line >= nlines) return;
    const unsigned gmask = (LW == 32) ? 0xffffffffu : (((1u << LW) - 1u) << ((threadIdx.x % 32) / LW * LW));

    // ---- line geometry --------------------------------------------------------------------
    int x, y, nsteps, ya = 0, racc = 0, stripe = 0;
    if (horizontal) {
        y = line; nsteps = a.W1; x = (a.px < 0) ? 0 : a.W1 - 1;
    } else {
        stripe = line / a.W1;
        x = line - stripe * a.W1;
        ya = a.sp.a[stripe];
        racc = a.sp.r0[stripe];
        const int yb = a.sp.r1[stripe];
        nsteps = yb - ya;
        y = (a.py < 0) ? ya : yb - 1;
    }
    const int sx = -a.px, sy = -a.py;

    const bool active = (gl * 8 < a.Dp);
    unsigned pm[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        int d = gl * 8 + 2 * j;
        pm[j] = (d >= a.D ? 0x7fffu : 0u) | (d + 1 >= a.D ? 0x7fff0000u : 0u);
    }
    const unsigned P1P1 = pk2((unsigned)a.P1);
    const int qlane = a.D / 8;  // lane holding the partial tail d0..D-1 (only used when quirk)

    unsigned L[4] = {0x7fff7fffu, 0x7fff7fffu, 0x7fff7fffu, 0x7fff7fffu};
    unsigned m = 0;
    const size_t frameC = (size_t)b * a.H * a.W1 * a.Dp;

    for (int t = 0; t < nsteps; t++) {
        bool reset = (t == 0);
        if (!horizontal) reset = reset || (a.px < 0 && x == 0) || (a.px > 0 && x == a.W1 - 1);
        // ---- loads -----------------------------------------------------------------------
        uint4 cq = make_uint4(0, 0, 0, 0), sq = make_uint4(0, 0, 0, 0);
        const bool acc = horizontal || (y >= racc);
        size_t off = frameC + ((size_t)y * a.W1 + x) * a.Dp + gl * 8;
        if (active) {
            const uint16_t* cp = a.C + off;
            if (a.Calt != nullptr && stripe > 0 && (y - ya) < a.sp.alt_rows)
                cp = a.Calt + (((size_t)(b * B200S_MAX_STRIPES + stripe) * a.sp.alt_rows + (y - ya)) * a.W1 + x) * a.Dp +
                     gl * 8;
            cq = __ldg((const uint4*)cp);
            if (acc && !a.first) sq = *(const uint4*)(a.S + off);
        }
        if (reset) {
            L[0] = L[1] = L[2] = L[3] = 0u;
            m = 0;
        }
        // ---- neighbours d-1 / d+1 across lanes ----------------------------------------------
        unsigned up = __shfl_up_sync(gmask, L[3], 1, LW);
        unsigned dn = __shfl_down_sync(gmask, L[0], 1, LW);
        if (gl == 0) up = 0x7fff7fffu;
        if (gl == LW - 1) dn = 0x7fff7fffu;
        const unsigned mm = pk2(m);
        const unsigned dmp = pk2(m + (unsigned)(a.P2 - a.P1));
        const unsigned Cq[4] = {cq.x, cq.y, cq.z, cq.w};
        unsigned Ln[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            unsigned bq = shift_pair(j == 0 ? up : L[j - 1], L[j]);       // (L[d-1], L[d])
            unsigned cn = shift_pair(L[j], j == 3 ? dn : L[j + 1]);       // (L[d+1], L[d+2])
            unsigned tt = umin3_2(bq, cn, dmp);
            tt = uaddmin2(tt, P1P1, L[j]);
            Ln[j] = umax2(Cq[j] + tt - mm, pm[j]);
        }
        if (a.quirk) {
            unsigned upn = __shfl_up_sync(gmask, Ln[3], 1, LW);
            if (gl == qlane) {
                unsigned bq = shift_pair(a.quirk == 1 ? upn : 0x7fff7fffu, L[0]);
                unsigned cn = shift_pair(L[0], L[1]);
                unsigned tt = umin3_2(bq, cn, dmp);
                tt = uaddmin2(tt, P1P1, L[0]);
                Ln[0] = umax2(Cq[0] + tt - mm, pm[0]);
            }
        }
        if (!active) Ln[0] = Ln[1] = Ln[2] = Ln[3] = 0x7fff7fffu;
        // ---- m = min over d ------------------------------------------------------------------
        unsigned mn = umin2(umin3_2(Ln[0], Ln[1], Ln[2]), Ln[3]);
#pragma unroll
        for (int o = LW / 2; o > 0; o >>= 1) mn = umin2(mn, __shfl_xor_sync(gmask, mn, o, LW));
        m = min(mn & 0xffffu, mn >> 16);
#pragma unroll
        for (int j = 0; j < 4; j++) L[j] = Ln[j];
        // ---- S ---------------------------------------------------------------------------------
        if (active && acc) {
            uint4 so;
            if (a.first) {
                so = make_uint4(Ln[0], Ln[1], Ln[2], Ln[3]);
            } else {
                so.x = uaddmin2(sq.x, Ln[0], 0x7fff7fffu);
                so.y = uaddmin2(sq.y, Ln[1], 0x7fff7fffu);
                so.z = uaddmin2(sq.z, Ln[2], 0x7fff7fffu);
                so.w = uaddmin2(sq.w, Ln[3], 0x7fff7fffu);
            }
            *(uint4*)(a.S + off) = so;
        }
        // ---- advance ---------------------------------------------------------------------------
        if (horizontal) {
            x += sx;
        } else {
            y += sy;
            x += sx;
            if (x >= a.W1) x -= a.W1;
            if (x < 0) x += a.W1;
        }
    }
}

static void launch_one(const PathArgs& a, int B, cudaStream_t st) {
    int LW = 2;
    while (LW * 8 < a.Dp) LW *= 2;
    const int nlines = (a.py == 0) ? a.H : a.W1 * a.sp.n;
    const int threads = 128;
    dim3 grid((unsigned)(((size_t)nlines * LW + threads - 1) / threads), 1, B);
    switch (LW) {
        case 2: k_path<2><<<grid, threads, 0, st>>>(a); break;
        case 4: k_path<4><<<grid, threads, 0, st>>>(a); break;
        case 8: k_path<8><<<grid, threads, 0, st>>>(a); break;
        case 16: k_path<16><<<grid, threads, 0, st>>>(a); break;
        default: k_path<32><<<grid, threads, 0, st>>>(a); break;
    }
}

int launch_paths(const Eff& e, const StripePlan& sp, int B, const uint16_t* C, const uint16_t* Calt, uint16_t* S,
                 cudaStream_t st, long long* launches, int vertical_only, int* flow) {
    if (e.Dp > 256) return -2;
    PathArgs a;
    a.C = C; a.Calt = nullptr; a.S = S;
    a.H = e.H; a.W1 = e.W1; a.D = e.D; a.Dp = e.Dp; a.P1 = e.P1e; a.P2 = e.P2e;
    a.quirk = 0;
    StripePlan whole;
    whole.n = 1; whole.a[0] = 0; whole.r0[0] = 0; whole.r1[0] = e.H; whole.alt_rows = 0;
    for (int i = 1; i < B200S_MAX_STRIPES; i++) whole.a[i] = whole.r0[i] = whole.r1[i] = 0;
    a.sp = whole;
    const bool tail = (e.mode == B200S_MODE_SGBM_3WAY) && (e.D % 8 != 0) && (e.D >= 8);
    int dirs[8][2];
    int nd = 0;
    auto add = [&](int px, int py) { dirs[nd][0] = px; dirs[nd][1] = py; nd++; };
    switch (e.mode) {
        case B200S_MODE_SGBM: add(-1, 0); add(1, 0); add(-1, -1); add(0, -1); add(1, -1); break;
        case B200S_MODE_HH:
            add(-1, 0); add(1, 0); add(-1, -1); add(0, -1); add(1, -1); add(-1, 1); add(0, 1); add(1, 1); break;
        case B200S_MODE_HH4: add(-1, 0); add(1, 0); add(0, -1); add(0, 1); break;
        default: add(-1, 0); add(1, 0); add(0, -1); break;  // sgbm_3way
    }
    bool first = true;
    // sgbm / hh: each vertical family ((x-1|x|x+1, y-1) and, for hh, (.., y+1)) runs as ONE cluster kernel
    // (vert.cu) when the strip state fits in shared memory; otherwise one generic launch per direction
    bool fused_down = false, fused_up = false;
    if (vertical_only && (e.mode == B200S_MODE_SGBM || e.mode == B200S_MODE_HH)) {
        int rv = launch_vert_family(e, B, C, S, +1, 0, st, nullptr);
        if (rv < 0) return -4;                         // the cluster kernel was planned but its launch failed
        if (rv == 0) {
            fused_down = true; first = false;
            if (launches) (*launches)++;
            if (flow) *flow |= B200S_FLOW_VERT_DOWN;
            if (e.mode == B200S_MODE_HH) {
                rv = launch_vert_family(e, B, C, S, -1, 1, st, nullptr);
                if (rv < 0) return -4;
                if (rv == 0) {
                    fused_up = true;
                    if (launches) (*launches)++;
                    if (flow) *flow |= B200S_FLOW_VERT_UP;
                }
            }
        }
    }
    for (int k = 0; k < nd; k++) {
        a.px = dirs[k][0]; a.py = dirs[k][1];
        if (vertical_only && a.py == 0) continue;  // the two horizontal paths live in rows.cu
        if ((fused_down && a.py == -1) || (fused_up && a.py == 1)) continue;
        a.first = first;
        first = false;
        a.quirk = 0;
        a.sp = whole;
        a.Calt = nullptr;
        if (e.mode == B200S_MODE_SGBM_3WAY) {
            if (tail && a.py == 0 && a.px == 1) a.quirk = 2;
            if (a.py != 0) {
                a.sp = sp;
                a.Calt = (sp.alt_rows > 0 && sp.n > 1) ? Calt : nullptr;
                if (tail) a.quirk = 1;
            }
        }
        launch_one(a, B, st);
        if (launches) (*launches)++;
        if (flow && a.py != 0) *flow |= B200S_FLOW_PATH_VERTICAL;
    }
    return 0;
}
