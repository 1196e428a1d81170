// redux.cu — group minimum of a packed u16x2 value over LW lanes: shuffle tree (what group_min2 does) against
// redux.sync.min.u32 with the group's member mask. Latency (one warp, dependent chain) and throughput (32 warps per SM on
// every SM), cycles per reduction.   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o redux redux.cu && ./redux
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned umin2(unsigned a, unsigned b) { return __vminu2(a, b); }

template <int LW> __device__ __forceinline__ unsigned tree(unsigned v, unsigned gmask) {
#pragma unroll
    for (int o = LW / 2; o > 0; o >>= 1) v = umin2(v, __shfl_xor_sync(gmask, v, o, LW));
    return min(v & 0xffffu, v >> 16);
}
template <int LW> __device__ __forceinline__ unsigned redux(unsigned v, unsigned gmask) {
    return __reduce_min_sync(gmask, min(v & 0xffffu, v >> 16));
}

template <int LW, bool REDUX>
__global__ void k(unsigned* out, long long* cyc, int iters) {
    const unsigned lane = threadIdx.x & 31;
    const unsigned gmask = LW == 32 ? 0xffffffffu : (((1u << LW) - 1u) << (lane / LW * LW));
    unsigned v = (threadIdx.x * 2654435761u) | 0x00010001u, acc = 0;
    const long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        const unsigned m = REDUX ? redux<LW>(v, gmask) : tree<LW>(v, gmask);
        v = (v + m * 0x00010001u + 0x00030005u) & 0x7fff7fffu;      // the next value depends on the minimum (as in the scan)
        acc += m;
    }
    const long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc + v;
}

// correctness: both give the same minimum per group
template <int LW> __global__ void check(int* bad) {
    const unsigned lane = threadIdx.x & 31;
    const unsigned gmask = LW == 32 ? 0xffffffffu : (((1u << LW) - 1u) << (lane / LW * LW));
    for (unsigned s = 0; s < 1000; s++) {
        const unsigned v = ((threadIdx.x + 1) * 2654435761u * (s + 1)) & 0x7fff7fffu;
        if (tree<LW>(v, gmask) != redux<LW>(v, gmask)) atomicAdd(bad, 1);
    }
}

template <int LW, bool REDUX> void run(const char* name) {
    unsigned* out; long long* cyc; cudaMalloc(&out, 148 * 1024 * 4 * 4); cudaMalloc(&cyc, 8);
    const int iters = 20000;
    long long h;
    k<LW, REDUX><<<1, 32>>>(out, cyc, iters); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double lat = (double)h / iters;
    k<LW, REDUX><<<148 * 4, 256>>>(out, cyc, iters); cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    // 32 warps per SM = 8 per scheduler: cycles per reduction per scheduler
    printf("%-22s LW=%2d  latency %6.1f cycles per dependent step   throughput %6.2f cycles per warp-step per scheduler\n", name, LW, lat,
           (double)h / iters / 8.0);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    int* bad; cudaMalloc(&bad, 4); cudaMemset(bad, 0, 4);
    check<8><<<2, 128>>>(bad); check<16><<<2, 128>>>(bad); check<32><<<2, 128>>>(bad); check<4><<<2, 128>>>(bad);
    int hb; cudaMemcpy(&hb, bad, 4, cudaMemcpyDeviceToHost);
    printf("mismatches tree vs redux: %d (%s)\n", hb, cudaGetErrorString(cudaGetLastError()));
    run<8, false>("shuffle tree"); run<8, true>("redux.sync.min");
    run<16, false>("shuffle tree"); run<16, true>("redux.sync.min");
    run<32, false>("shuffle tree"); run<32, true>("redux.sync.min");
    return 0;
}
