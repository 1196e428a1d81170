// Issue-rate micro-benchmark for the packed 16x2 integer instructions the matcher is built from (sm_100a).
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pipes pipes.cu && ./pipes
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 4096
template <int OP>
__global__ void k(unsigned* out, unsigned seed) {
    unsigned a[8];
#pragma unroll
    for (int j = 0; j < 8; j++) a[j] = seed + threadIdx.x * 7 + j;
    unsigned b = seed * 3 + 1, c = seed ^ 0x12341234u;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < ITER; i++) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (OP == 0) a[j] = __vminu2(a[j], b + j);                         // VIMNMX.U16x2
            if (OP == 1) a[j] = __vimin3_u16x2(a[j], b + j, c);                // VIMNMX3.U16x2
            if (OP == 2) a[j] = __viaddmin_u16x2(a[j], b, c + j);              // VIADDMNMX.U16x2
            if (OP == 3) a[j] = __byte_perm(a[j], b + j, 0x5432);              // PRMT
            if (OP == 4) a[j] = a[j] + b + (c ^ j);                            // IADD3 / LOP3
            if (OP == 5) a[j] = __vadd2(a[j], b + j);                          // VIADD.16x2
            if (OP == 6) a[j] = a[j] * 3 + b;                                  // IMAD
            if (OP == 7) a[j] = min(a[j], b + j);                              // VIMNMX.U32
            if (OP == 8) a[j] = __viaddmax_s16x2_relu(a[j], b, c + j);         // VIADDMNMX.S16x2.RELU
        }
    }
    long long t1 = clock64();
    unsigned r = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) r ^= a[j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (unsigned)(t1 - t0);
}
template <int OP>
void run(const char* name, unsigned* d) {
    for (int warps = 4; warps <= 16; warps *= 2) {
        k<OP><<<148, warps * 32>>>(d, 5);
        cudaDeviceSynchronize();
        unsigned cyc; cudaMemcpy(&cyc, d, 4, cudaMemcpyDeviceToHost);
        double per = (double)cyc / (ITER * 8.0) / (warps / 4.0);   // cycles per warp-instruction per SMSP
        printf("%-28s warps/SM=%2d  cycles per warp-instr per SMSP: %.2f\n", name, warps, per);
    }
}
int main() {
    unsigned* d; cudaMalloc(&d, 148 * 1024 * 4);
    run<0>("VIMNMX.U16x2", d); run<1>("VIMNMX3.U16x2", d); run<2>("VIADDMNMX.U16x2", d); run<3>("PRMT", d);
    run<4>("IADD3+LOP3 (2 instr)", d); run<5>("VIADD.16x2", d); run<6>("IMAD", d); run<7>("VIMNMX.U32", d);
    run<8>("VIADDMNMX.S16x2.RELU", d);
    return 0;
}
