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
#include "sgm_step.cuh"

struct PathArgs {
    const uint16_t* C;
    const uint16_t* Calt;
    uint16_t* S;
    int H, W1, D, Dp, P1, P2;
    int px, py;   // predecessor offset (A.5)
    int first;    // 1: S = L (S not read)
    int quirk;    // sgbm_3way D%8 tail artefact: 0 none, 1 vertical path, 2 px=+1 path (A.7)
    int plain;    // product flow (S is consumed by the row kernel): the unpadded instantiation may be used when D == Dp
    StripePlan sp;
};

// LW lanes x 2 NR disparities per lane; a warp carries 32 / LW lines (LW need not be a power of two, sgm_step.cuh).
// PAD = false: no padded disparities at all (D == Dp == LW * 2 * NR) and S goes to the row kernel: no clamp per step
template <int LW, int NR, bool PAD>
__global__ void __launch_bounds__(128) k_path(const PathArgs a) {
    using LG = LaneGroup<LW>;
    constexpr int DPL = 2 * NR;
    const int lane = threadIdx.x & 31;
    const bool idle = !LG::POW2 && lane >= LG::USED;
    const int gl = LG::POW2 ? (int)threadIdx.x % LW : (idle ? lane - LG::USED : lane % LW);   // lane within the line group
    const int wline = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * LG::RPW;        // first line of this warp
    const int b = blockIdx.z;
    const bool horizontal = (a.py == 0);
    const int nlines = horizontal ? a.H : a.W1 * a.sp.n;
    if (wline >= nlines) return;                                                              // whole warps only: shuffles below
    const int line_raw = wline + (LG::POW2 ? lane / LW : (idle ? LG::RPW - 1 : lane / LW));
    const bool live = line_raw < nlines && !idle;
    const int line = live ? line_raw : nlines - 1;                                            // surplus groups shadow the last line
    const unsigned gmask = 0xffffffffu;

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
    // every lane of the warp runs the longest line of the warp (stripes differ in length): shuffles stay converged
    int wsteps = nsteps;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wsteps = max(wsteps, __shfl_xor_sync(0xffffffffu, wsteps, o));
    const int sx = -a.px, sy = -a.py;

    const bool active = (gl * DPL < a.Dp) && live;
    unsigned pm[NR];
    make_padmask<NR>(pm, gl, a.D);
    const unsigned P1P1 = pk2((unsigned)a.P1), P2mP1 = (unsigned)(a.P2 - a.P1);
    const int d0 = 8 * (a.D / 8);

    unsigned L[NR];
#pragma unroll
    for (int j = 0; j < NR; j++) L[j] = 0x7fff7fffu;
    unsigned m = 0;
    const size_t frameC = (size_t)b * a.H * a.W1 * a.Dp;
    auto c_ptr = [&](const int xx, const int yy) -> const uint16_t* {
        if (a.Calt != nullptr && stripe > 0 && (yy - ya) < a.sp.alt_rows)
            return a.Calt + (((size_t)(b * B200S_MAX_STRIPES + stripe) * a.sp.alt_rows + (yy - ya)) * a.W1 + xx) * a.Dp + gl * DPL;
        return a.C + frameC + ((size_t)yy * a.W1 + xx) * a.Dp + gl * DPL;
    };
    // C of the next step is requested before the current step is computed (one step of look-ahead in registers)
    unsigned Cq[NR], Cn[NR];
#pragma unroll
    for (int j = 0; j < NR; j++) { Cq[j] = 0u; Cn[j] = 0u; }
    if (active && nsteps > 0) load_regs_nc<NR>(c_ptr(x, y), Cn);

    for (int t = 0; t < wsteps; t++) {
        const bool in_line = t < nsteps;
        bool reset = (t == 0);
        if (!horizontal) reset = reset || (a.px < 0 && x == 0) || (a.px > 0 && x == a.W1 - 1);
        // ---- loads -----------------------------------------------------------------------
        unsigned Sq[NR];
#pragma unroll
        for (int j = 0; j < NR; j++) { Cq[j] = Cn[j]; Sq[j] = 0u; }
        const bool acc = horizontal || (y >= racc);
        const size_t off = frameC + ((size_t)y * a.W1 + x) * a.Dp + gl * DPL;
        int xn = x, yn = y;
        if (horizontal) {
            xn += sx;
        } else {
            yn += sy;
            xn += sx;
            if (xn >= a.W1) xn -= a.W1;
            if (xn < 0) xn += a.W1;
        }
        if (active && in_line) {
            if (t + 1 < nsteps) load_regs_nc<NR>(c_ptr(xn, yn), Cn);
            if (acc && !a.first) load_regs<NR>(a.S + off, Sq);
        }
        if (reset) sgm_reset<NR>(L, m);
        sgm_step<LW, NR, PAD>(L, m, Cq, pm, gmask, gl, P1P1, P2mP1, a.quirk, d0);
        // ---- S ---------------------------------------------------------------------------------
        if (active && in_line && acc) {
            unsigned So[NR];
#pragma unroll
            for (int j = 0; j < NR; j++) So[j] = a.first ? L[j] : uaddmin2(Sq[j], L[j], 0x7fff7fffu);
            store_regs<NR>(a.S + off, So);
        }
        x = xn; y = yn;
    }
}

template <int LW, int NR>
static void launch_path_t(const PathArgs& a, int B, int nlines, cudaStream_t st) {
    const int threads = 128;
    const int per_cta = (threads / 32) * LaneGroup<LW>::RPW;
    dim3 grid((unsigned)((nlines + per_cta - 1) / per_cta), 1, B);
    if (a.plain && a.D == LW * 2 * NR) k_path<LW, NR, false><<<grid, threads, 0, st>>>(a);
    else k_path<LW, NR, true><<<grid, threads, 0, st>>>(a);
}

static void launch_one(const PathArgs& a, int B, cudaStream_t st) {
    const int nlines = (a.py == 0) ? a.H : a.W1 * a.sp.n;
    if (a.Dp > 64) {                       // 16 disparities per lane, the lane count follows D
        switch ((a.Dp + 15) / 16) {
            case 5: launch_path_t<5, 8>(a, B, nlines, st); break;
            case 6: launch_path_t<6, 8>(a, B, nlines, st); break;
            case 7: launch_path_t<7, 8>(a, B, nlines, st); break;
            case 8: launch_path_t<8, 8>(a, B, nlines, st); break;
            case 9: launch_path_t<9, 8>(a, B, nlines, st); break;
            case 10: launch_path_t<10, 8>(a, B, nlines, st); break;
            default: launch_path_t<16, 8>(a, B, nlines, st); break;
        }
        return;
    }
    switch ((a.Dp + 7) / 8) {              // 8 disparities per lane
        case 1: case 2: launch_path_t<2, 4>(a, B, nlines, st); break;
        case 3: launch_path_t<3, 4>(a, B, nlines, st); break;
        case 4: launch_path_t<4, 4>(a, B, nlines, st); break;
        case 5: launch_path_t<5, 4>(a, B, nlines, st); break;
        case 6: launch_path_t<6, 4>(a, B, nlines, st); break;
        case 7: launch_path_t<7, 4>(a, B, nlines, st); break;
        default: launch_path_t<8, 4>(a, B, nlines, st); break;
    }
}

int launch_paths(const Eff& e, const StripePlan& sp, int B, const uint16_t* C, const uint16_t* Calt, uint16_t* S,
                 cudaStream_t st, long long* launches, int vertical_only, int* flow) {
    if (e.Dp > 256) return -2;
    PathArgs a;
    a.C = C; a.Calt = nullptr; a.S = S;
    a.H = e.H; a.W1 = e.W1; a.D = e.D; a.Dp = e.Dp; a.P1 = e.P1e; a.P2 = e.P2e;
    a.quirk = 0;
    a.plain = vertical_only ? 1 : 0;   // the debug flow's selection kernel reads S as is: it keeps the clamped steps
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
