// Which pipe does each candidate instruction issue on, and do two of them co-issue? (sm_100a)
// For every pair (A, B): 4 independent chains of A and 4 of B per thread; cycles per warp-instruction per SMSP.
// A pair that shares a pipe costs the sum of the two rates; a pair on different pipes costs the max.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pipes2 pipes2.cu && ./pipes2
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#define ITER 2048
enum { VMIN2, IMAD, HMNMX2, HADD2, IADD3, PRMT, LOP3, IMADADD, VIADDMNMX, HFMA2, IMADHI, SHF, FMNMX, VMIN3, FADD, IMADSHL, VIADD2, NOPS };
static const char* NAMES[NOPS] = {"VIMNMX.U16x2", "IMAD", "HMNMX2", "HADD2", "IADD3", "PRMT", "LOP3", "IMAD(x*one+y)", "VIADDMNMX.U16x2",
                                  "HFMA2", "IMAD.HI", "SHF", "FMNMX", "VIMNMX3.U16x2", "FADD", "IMAD(x*65536+y)", "VIADD.16x2"};
template <int OP>
__device__ __forceinline__ unsigned op(unsigned a, unsigned b, unsigned c, unsigned one) {
    unsigned r;
    if (OP == VMIN2) return __vminu2(a, b);
    if (OP == IMAD) return a * 3 + b;
    if (OP == HMNMX2) { asm volatile("min.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
    if (OP == HADD2) { asm volatile("add.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
    if (OP == IADD3) return a + b + c;
    if (OP == PRMT) return __byte_perm(a, b, 0x5432);
    if (OP == LOP3) return (a & b) ^ c;
    if (OP == IMADADD) return a * one + b;
    if (OP == VIADDMNMX) return __viaddmin_u16x2(a, b, c);
    if (OP == HFMA2) { asm volatile("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
    if (OP == IMADHI) return __umulhi(a, b) + c;
    if (OP == SHF) return __funnelshift_r(a, b, 16);
    if (OP == FMNMX) return __float_as_uint(fminf(__uint_as_float(a), __uint_as_float(b)));
    if (OP == VMIN3) return __vimin3_u16x2(a, b, c);
    if (OP == FADD) return __float_as_uint(__uint_as_float(a) + __uint_as_float(b));
    if (OP == IMADSHL) return a * 65536u + b;
    if (OP == VIADD2) return __vadd2(a, b);
    return a;
}
template <int A, int B>
__global__ void k(unsigned* out, unsigned seed, unsigned one) {
    unsigned a[8];
#pragma unroll
    for (int j = 0; j < 8; j++) a[j] = seed + threadIdx.x * 7 + j;
    unsigned b = seed * 3 + 1, c = seed ^ 0x12341234u;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < ITER; i++) {
#pragma unroll
        for (int j = 0; j < 4; j++) {
            a[j] = op<A>(a[j], b + j, c, one);
            a[j + 4] = op<B>(a[j + 4], b + j, c, one);
        }
    }
    long long t1 = clock64();
    unsigned r = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) r ^= a[j];
    out[blockIdx.x * blockDim.x + threadIdx.x + 1] = r;
    if (threadIdx.x == 0 && blockIdx.x == 0) out[0] = (unsigned)(t1 - t0);
}
template <int A, int B>
void run(unsigned* d) {
    const int warps = 16;
    k<A, B><<<148, warps * 32>>>(d, 5, 1);
    cudaDeviceSynchronize();
    unsigned cyc; cudaMemcpy(&cyc, d, 4, cudaMemcpyDeviceToHost);
    double per = (double)cyc / (ITER * 8.0) / (warps / 4.0);
    printf("%-18s + %-18s : %.2f cycles per warp-instr per SMSP\n", NAMES[A], NAMES[B], per);
}
template <int A> void row(unsigned* d) {
    run<A, A>(d); run<VMIN2, A>(d); run<IMAD, A>(d);
}
int main() {
    unsigned* d; cudaMalloc(&d, (148 * 1024 + 1) * 4);
    row<VMIN2>(d); row<IMAD>(d); row<HMNMX2>(d); row<HADD2>(d); row<IADD3>(d); row<PRMT>(d); row<LOP3>(d); row<IMADADD>(d);
    row<VIADDMNMX>(d); row<HFMA2>(d); row<IMADHI>(d); row<SHF>(d); row<FMNMX>(d); row<VMIN3>(d); row<FADD>(d); row<IMADSHL>(d); row<VIADD2>(d);
    return 0;
}
