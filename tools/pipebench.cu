// pipebench.cu -- integer pipe throughput of one B200 SM (thread-ops per clock per SM), to settle the
// denominators of the integer-issue roofline ("INT32 lanes per SM ... measure in the same run").
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pipebench pipebench.cu && ./pipebench
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 4096
#define UNROLL 8

template <int OP>
__global__ void __launch_bounds__(1024) k(uint32_t *out, long long *cyc, uint32_t seed) {
    __shared__ uint32_t tab[8192];
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) tab[i] = i * 2654435761u;
    __syncthreads();
    uint32_t a[UNROLL];
#pragma unroll
    for (int i = 0; i < UNROLL; i++) a[i] = seed + threadIdx.x * 31 + i;
    uint32_t b = seed ^ 0x9E3779B9u, c = seed * 7 + 1;
    long long t0 = clock64();
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < UNROLL; i++) {
            if (OP == 0) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c));
            if (OP == 1) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
            if (OP == 2) asm volatile("shf.l.wrap.b32 %0, %0, %1, 7;" : "+r"(a[i]) : "r"(b));
            if (OP == 3) asm volatile("popc.b32 %0, %0;" : "+r"(a[i]));
            if (OP == 4) asm volatile("bfind.u32 %0, %0;" : "+r"(a[i]));
            if (OP == 5) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
            if (OP == 6) asm volatile("prmt.b32 %0, %0, %1, 0x3210;" : "+r"(a[i]) : "r"(b));
            if (OP == 7) { a[i] = tab[a[i] & 8191]; }                                  // random LDS (dependent)
            if (OP == 8) { asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c));
                           asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[i]) : "r"(b), "r"(c)); }   // alu+fma mix
            if (OP == 9) asm volatile("shl.b32 %0, %0, 1;" : "+r"(a[i]));
            if (OP == 11) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
            if (OP == 12) asm volatile("min.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b));
            if (OP == 13) asm volatile("add.u32 %0, %0, %1; add.u32 %0, %0, %2;" : "+r"(a[i]) : "r"(b), "r"(c));
            if (OP == 14) asm volatile("shr.u32 %0, %0, 1;" : "+r"(a[i]));
            if (OP == 15) { asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c));
                            asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(a[i]) : "r"(b)); }
            if (OP == 16) { asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[i]) : "r"(b), "r"(c));
                            asm volatile("popc.b32 %0, %0;" : "+r"(a[i])); }
            if (OP == 17) asm volatile("mad.lo.u32 %0, %0, 1, %1;" : "+r"(a[i]) : "r"(b));
            if (OP == 18) asm volatile("{.reg .pred p; setp.ne.u32 p, %1, 0; @p mov.u32 %0, %2;}" : "+r"(a[i]) : "r"(b), "r"(c));
            if (OP == 10) { asm volatile("{.reg .pred p; setp.gt.u32 p, %0, %1; selp.u32 %0, %0, %2, p;}" : "+r"(a[i]) : "r"(b), "r"(c)); }
        }
    }
    long long t1 = clock64();
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < UNROLL; i++) s ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char *name, int ops_per_inner, int sms) {
    const int threads = 1024, blocks = sms * 2;
    uint32_t *out; long long *cyc;
    cudaMalloc(&out, sizeof(uint32_t) * threads * blocks);
    cudaMalloc(&cyc, sizeof(long long) * blocks);
    k<OP><<<blocks, threads>>>(out, cyc, 12345);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<OP><<<blocks, threads>>>(out, cyc, 777);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h[1024]; cudaMemcpy(h, cyc, sizeof(long long) * blocks, cudaMemcpyDeviceToHost);
    long long mx = 0; for (int i = 0; i < blocks; i++) if (h[i] > mx) mx = h[i];
    // two 1024-thread blocks are resident per SM and run concurrently
    double ops_sm = 2.0 * threads * (double)ITERS * UNROLL * ops_per_inner;
    printf("%-28s %8.2f thread-ops/clk/SM   (%.3f ms, %.1f Tops/s chip)\n", name, ops_sm / (double)mx, ms,
           ops_sm * sms / (ms * 1e-3) / 1e12);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    printf("%s, %d SMs, clockRate %d kHz\n", p.name, p.multiProcessorCount, p.clockRate);
    int sms = p.multiProcessorCount;
    run<0>("LOP3", 1, sms); run<1>("IADD", 1, sms); run<2>("SHF", 1, sms); run<9>("SHL imm", 1, sms);
    run<6>("PRMT", 1, sms); run<3>("POPC", 1, sms); run<4>("FLO (bfind)", 1, sms); run<5>("IMAD", 1, sms);
    run<10>("ISETP+SEL", 2, sms); run<11>("IMAD.HI (mul.hi)", 1, sms); run<12>("VIMNMX (min)", 1, sms);
    run<13>("IADD x2 (3-input?)", 2, sms); run<14>("SHR imm", 1, sms); run<15>("LOP3+IMAD.HI mix", 2, sms);
    run<16>("LOP3+POPC mix", 2, sms); run<17>("mad.lo x*1+b", 1, sms); run<18>("ISETP+@p MOV", 2, sms); run<8>("LOP3+IMAD mix", 2, sms); run<7>("LDS random (8192 words)", 1, sms);
    return 0;
}
