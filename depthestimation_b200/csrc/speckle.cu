// speckle.cu — cv2.filterSpeckles: 4-connected components over pixels != newVal,
// neighbours joined iff |p - q| <= maxDiff; components with size <= maxSize are set to newVal.
// Used twice per frame: inside StereoSGBM.compute (A.8, depthlib/stereo_core.py:231) and in
// postprocess.filter_speckles (depthlib/postprocess.py:6-35).
//
// Tiled union-find: (A) each CTA labels a 64x16 tile entirely in shared memory (horizontal runs by
// warp ballot, vertical links by shared-memory atomicMin, redundant links skipped), and writes
// per pixel the global index of its tile-local root plus, at roots, the local component size;
// (B) only tile-border pixels link across tiles in global memory; (C) tile-local roots add their
// size to the global root; (D) every pixel looks its root's size up. The result does not depend on
// traversal order, so it is bit-identical to the sequential flood fill.
#include "common.cuh"

#define TW 64
#define TH 16

__device__ __forceinline__ int uf_find(const int* lab, int i) {
    int p = ((volatile const int*)lab)[i];
    while (p != i) { i = p; p = ((volatile const int*)lab)[i]; }
    return i;
}
__device__ __forceinline__ void uf_union(int* lab, int a, int b) {
    while (true) {
        a = uf_find(lab, a);
        b = uf_find(lab, b);
        if (a == b) return;
        if (a < b) { int t = a; a = b; b = t; }   // hook the larger root under the smaller
        int old = atomicMin(&lab[a], b);
        if (old == a) return;
        a = old;
    }
}
__device__ __forceinline__ bool joined(int p, int q, int newVal, int maxDiff) {
    return p != newVal && q != newVal && abs(p - q) <= maxDiff;
}

__global__ void __launch_bounds__(256) k_ccl_tile(const int16_t* __restrict__ img, int H, int W, int newVal, int maxDiff,
                                                  int* __restrict__ lab, int* __restrict__ sizes) {
    __shared__ short val[TH][TW + 1];
    __shared__ int par[TH * TW];
    __shared__ int cnt[TH * TW];
    const int x0 = blockIdx.x * TW, y0 = blockIdx.y * TH;
    const size_t frame = (size_t)blockIdx.z * H * W;
    const int16_t* src = img + frame;
    for (int p = threadIdx.x; p < TH * TW; p += 256) {
        int lx = p % TW, ly = p / TW, x = x0 + lx, y = y0 + ly;
        val[ly][lx] = (x < W && y < H) ? src[(size_t)y * W + x] : (short)newVal;
        cnt[p] = 0;
    }
    __syncthreads();
    // horizontal runs: a warp covers 32 consecutive pixels of a row, so the start of a pixel's run inside that
    // segment comes from one ballot ("joined to the left neighbour") and a count-leading-zeros; a run that reaches
    // the segment start continues through that pixel's own link to the previous segment (chain of length 2)
    static_assert(TW % 32 == 0 && (TH * TW) % 256 == 0, "tile layout");
    int piece[TH * TW / 256];   // > 0: this pixel starts a run piece of that many valid pixels inside its 32-pixel segment
#pragma unroll
    for (int k = 0; k < TH * TW / 256; k++) {
        const int p = threadIdx.x + 256 * k;
        const int lx = p % TW, ly = p / TW, lane = threadIdx.x & 31;
        const bool jl = lx > 0 && joined(val[ly][lx], val[ly][lx - 1], newVal, maxDiff);
        const unsigned bits = __ballot_sync(0xffffffffu, jl);
        const unsigned zeros = ~bits & (0xffffffffu >> (31 - lane));      // run starts at or below this lane
        // no start in the segment (lane 0 is joined to the previous segment): hang on lane 0, which links to p - 1
        const int start = zeros ? 31 - __clz(zeros) : 0;
        par[p] = (lane == 0) ? (jl ? p - 1 : p) : p - lane + start;
        const unsigned above = (lane == 31) ? 0u : (~bits & (0xffffffffu << (lane + 1)));   // run starts above this lane
        const int len = (above ? __ffs(above) - 1 : 32) - lane;
        piece[k] = ((lane == 0 || !jl) && val[ly][lx] != newVal) ? len : 0;
    }
    // Fast paths for the two common tiles of a disparity map: every pixel equal to newVal (nothing to label), and
    // every pixel joined to its left and upper neighbour inside the tile (ONE component: a smooth surface). Both skip
    // the union-find phases; the result is what they would produce (root = the tile's first pixel).
    {
        bool all_new = true, all_joined = (x0 + TW <= W && y0 + TH <= H);
#pragma unroll
        for (int k = 0; k < TH * TW / 256; k++) {
            const int p = threadIdx.x + 256 * k;
            const int lx = p % TW, ly = p / TW;
            const int v = val[ly][lx];
            all_new = all_new && (v == newVal);
            all_joined = all_joined && (lx == 0 || joined(v, val[ly][lx - 1], newVal, maxDiff)) &&
                         (ly == 0 || joined(v, val[ly - 1][lx], newVal, maxDiff)) && v != newVal;
        }
        // (the barrier reductions return a truth value, not a bitwise result: one per question)
        const bool one = __syncthreads_and(all_joined ? 1 : 0) != 0;
        const bool none = __syncthreads_and(all_new ? 1 : 0) != 0;
        if (one || none) {
#pragma unroll
            for (int k = 0; k < TH * TW / 256; k++) {
                const int p = threadIdx.x + 256 * k;
                const int lx = p % TW, ly = p / TW, x = x0 + lx, y = y0 + ly;
                if (x < W && y < H) {
                    const size_t gi = frame + (size_t)y * W + x;
                    lab[gi] = one ? (int)(frame + (size_t)y0 * W + x0) : -1;
                    sizes[gi] = (one && p == 0) ? TH * TW : 0;
                }
            }
            return;
        }
    }
    // vertical links; skipped when the left neighbour pair already carries the same link
    for (int p = threadIdx.x; p < TH * TW; p += 256) {
        int lx = p % TW, ly = p / TW;
        if (ly == 0) continue;
        int v = val[ly][lx], u = val[ly - 1][lx];
        if (!joined(v, u, newVal, maxDiff)) continue;
        if (lx > 0) {
            int vl = val[ly][lx - 1], ul = val[ly - 1][lx - 1];
            if (joined(v, vl, newVal, maxDiff) && joined(u, ul, newVal, maxDiff) && joined(vl, ul, newVal, maxDiff)) continue;
        }
        uf_union(par, p, p - TW);
    }
    __syncthreads();
    int root[TH * TW / 256];
#pragma unroll
    for (int k = 0; k < TH * TW / 256; k++) {
        int p = threadIdx.x + 256 * k;
        int lx = p % TW, ly = p / TW;
        root[k] = (val[ly][lx] != newVal) ? uf_find(par, p) : -1;
    }
    // local component sizes: one shared-memory atomic per run piece
#pragma unroll
    for (int k = 0; k < TH * TW / 256; k++)
        if (piece[k] > 0) atomicAdd(&cnt[root[k]], piece[k]);
    __syncthreads();
#pragma unroll
    for (int k = 0; k < TH * TW / 256; k++) {
        int p = threadIdx.x + 256 * k;
        int lx = p % TW, ly = p / TW, x = x0 + lx, y = y0 + ly;
        if (x < W && y < H) {
            size_t gi = frame + (size_t)y * W + x;
            int r = root[k];
            lab[gi] = (r >= 0) ? (int)(frame + (size_t)(y0 + r / TW) * W + (x0 + r % TW)) : -1;
            sizes[gi] = (r == p) ? cnt[p] : 0;
        }
    }
}

// links across tile borders: item < nh: horizontal border rows (y = TH*k), else vertical border columns (x = TW*k)
__global__ void __launch_bounds__(256) k_ccl_border(const int16_t* __restrict__ img, int H, int W, int newVal, int maxDiff,
                                                    int* lab) {
    const int nby = (H - 1) / TH, nbx = (W - 1) / TW;  // number of interior borders
    const int nh = nby * W, nv = nbx * H;
    int item = blockIdx.x * blockDim.x + threadIdx.x;
    if (item >= nh + nv) return;
    const size_t frame = (size_t)blockIdx.z * H * W;
    const int16_t* s = img + frame;
    if (item < nh) {
        int y = (item / W + 1) * TH, x = item % W;
        int v = s[(size_t)y * W + x], u = s[(size_t)(y - 1) * W + x];
        if (!joined(v, u, newVal, maxDiff)) return;
        if (x % TW != 0) {
            int vl = s[(size_t)y * W + x - 1], ul = s[(size_t)(y - 1) * W + x - 1];
            if (joined(v, vl, newVal, maxDiff) && joined(u, ul, newVal, maxDiff) && joined(vl, ul, newVal, maxDiff)) return;
        }
        uf_union(lab, (int)(frame + (size_t)y * W + x), (int)(frame + (size_t)(y - 1) * W + x));
    } else {
        item -= nh;
        int x = (item / H + 1) * TW, y = item % H;
        int v = s[(size_t)y * W + x], l = s[(size_t)y * W + x - 1];
        if (!joined(v, l, newVal, maxDiff)) return;
        if (y % TH != 0) {
            int vu = s[(size_t)(y - 1) * W + x], lu = s[(size_t)(y - 1) * W + x - 1];
            if (joined(v, vu, newVal, maxDiff) && joined(l, lu, newVal, maxDiff) && joined(vu, lu, newVal, maxDiff)) return;
        }
        uf_union(lab, (int)(frame + (size_t)y * W + x), (int)(frame + (size_t)y * W + x - 1));
    }
}

// tile-local roots hand their size to the global root (CCL_IPT block-strided items per thread: a batch would otherwise
// be ~140 000 one-item CTAs, and CTA launch rate, not memory, would set the time)
#define CCL_IPT 8
__global__ void __launch_bounds__(256) k_ccl_count(size_t n, int* lab, int* sizes) {
#pragma unroll
    for (int k = 0; k < CCL_IPT; k++) {
        const size_t i = ((size_t)blockIdx.x * CCL_IPT + k) * 256 + threadIdx.x;
        if (i >= n) return;
        int c = sizes[i];
        if (c == 0) continue;             // not a tile-local root
        int r = uf_find(lab, (int)i);
        if (r != (int)i) {
            atomicAdd(&sizes[r], c);
            lab[i] = r;                   // roots do not change in this kernel
        }
    }
}

__global__ void __launch_bounds__(256) k_ccl_apply(int16_t* img, size_t n, int newVal, int maxSize, const int* lab,
                                                   const int* __restrict__ sizes) {
#pragma unroll
    for (int k = 0; k < CCL_IPT; k++) {
        const size_t i = ((size_t)blockIdx.x * CCL_IPT + k) * 256 + threadIdx.x;
        if (i >= n) return;
        int l = lab[i];
        if (l < 0) continue;
        if (sizes[l] > maxSize) continue; // the tile-local part alone is big enough: the component stays (most pixels end here)
        int r = uf_find(lab, l);
        if (sizes[r] <= maxSize) img[i] = (int16_t)newVal;
    }
}

void launch_speckle(int16_t* img, int B, int H, int W, int newVal, int maxSize, int maxDiff, int* labels, int* sizes,
                    cudaStream_t st, long long* launches) {
    size_t n = (size_t)B * H * W;
    unsigned blocks = (unsigned)((n + 256 * CCL_IPT - 1) / (256 * CCL_IPT));
    dim3 tg((W + TW - 1) / TW, (H + TH - 1) / TH, B);
    k_ccl_tile<<<tg, 256, 0, st>>>(img, H, W, newVal, maxDiff, labels, sizes);
    int items = ((H - 1) / TH) * W + ((W - 1) / TW) * H;
    int nl = 3;
    if (items > 0) {
        dim3 bg((items + 255) / 256, 1, B);
        k_ccl_border<<<bg, 256, 0, st>>>(img, H, W, newVal, maxDiff, labels);
        nl++;
    }
    k_ccl_count<<<blocks, 256, 0, st>>>(n, labels, sizes);
    k_ccl_apply<<<blocks, 256, 0, st>>>(img, n, newVal, maxSize, labels, sizes);
    if (launches) *launches += nl;
}

// ---- allocation-free entry point -----------------------------------------------------------------
// b200s_filter_speckles_noalloc (include/b200stereo.h): the same four kernels, with labels + sizes in a file-scope
// device array instead of the handle's cudaMalloc'd scratch, for callers that may not allocate (CUDA graph capture,
// serving loops). No counterpart in the reference.
//
// g_noalloc_scratch is a HOST-side stand-in in host code, not a device address: the device address is taken once per
// device with cudaGetSymbolAddress (below) and that pointer is what the kernels receive. Every GPU that loads the
// library holds the array (2 x 4 B x B200S_NOALLOC_MAX_PIXELS = 16 MiB) until the process exits. All calls on a
// device share it: k_ccl_tile writes every label and size the later kernels read (nothing has to start at zero, so
// the previous call's contents are harmless), but two calls whose kernels overlap on the GPU race on it.
#include <mutex>
#include <stdarg.h>

static __device__ __align__(256) int g_noalloc_scratch[2 * (size_t)B200S_NOALLOC_MAX_PIXELS];

#define NOALLOC_MAX_DEVICES 64
static std::mutex g_noalloc_mu;
static int* g_noalloc_ptr[NOALLOC_MAX_DEVICES];   // device address of g_noalloc_scratch, per device ordinal

static int noalloc_fail(char* err, size_t cap, int code, const char* fmt, ...) {
    if (err && cap) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, cap, fmt, ap);
        va_end(ap);
    }
    return code;
}

// once per device, with that device current
static cudaError_t noalloc_scratch(int dev, int** out) {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        int* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return cudaSuccess;
}

extern "C" int b200s_filter_speckles_noalloc(const int16_t* d_src, int16_t* d_dst, int H, int W, int new_val, int max_size,
                                             int max_diff, void* cuda_stream, char* err, size_t err_cap) {
    if (err && err_cap) err[0] = 0;
    if (!d_src || !d_dst) return noalloc_fail(err, err_cap, B200S_E_INVALID, "filter_speckles_noalloc: null device pointer");
    if (H <= 0 || W <= 0) return noalloc_fail(err, err_cap, B200S_E_INVALID, "filter_speckles_noalloc: bad size %dx%d", W, H);
    if ((long long)H * W > (long long)B200S_NOALLOC_MAX_PIXELS)
        return noalloc_fail(err, err_cap, B200S_E_INVALID,
                            "filter_speckles_noalloc: %dx%d = %lld pixels exceeds the static scratch (B200S_NOALLOC_MAX_PIXELS = %d); "
                            "use b200s_postprocess / b200s_sgbm_compute, which allocate",
                            W, H, (long long)H * W, (int)B200S_NOALLOC_MAX_PIXELS);
    if (new_val < -32768 || new_val > 32767 || max_diff < 0)
        return noalloc_fail(err, err_cap, B200S_E_INVALID, "filter_speckles_noalloc: new_val %d / max_diff %d out of range", new_val, max_diff);
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return noalloc_fail(err, err_cap, B200S_E_CUDA, "cudaGetDevice failed: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return noalloc_fail(err, err_cap, B200S_E_INVALID, "filter_speckles_noalloc: device ordinal %d not supported", dev);
    int* scratch = nullptr;
    e = noalloc_scratch(dev, &scratch);
    if (e != cudaSuccess) return noalloc_fail(err, err_cap, B200S_E_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch) failed on device %d: %s", dev, cudaGetErrorString(e));
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const size_t n = (size_t)H * W;
    if (d_dst != d_src) {
        e = cudaMemcpyAsync(d_dst, d_src, n * sizeof(int16_t), cudaMemcpyDeviceToDevice, st);
        if (e != cudaSuccess) return noalloc_fail(err, err_cap, B200S_E_CUDA, "device-to-device copy failed: %s", cudaGetErrorString(e));
    }
    launch_speckle(d_dst, 1, H, W, new_val, max_size, max_diff, scratch, scratch + B200S_NOALLOC_MAX_PIXELS, st, nullptr);
    e = cudaGetLastError();
    if (e != cudaSuccess) return noalloc_fail(err, err_cap, B200S_E_CUDA, "speckle kernels failed to launch: %s", cudaGetErrorString(e));
    return B200S_OK;
}
