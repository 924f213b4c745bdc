// pipe_probe.cu -- issue-rate probe for the instruction classes of the BVH8 node step on sm_100a.
// Each test runs 8 independent dependency chains of one instruction class (or a fixed mix) per thread, 16 warps per SM
// (4 per scheduler), and reports warp instructions per cycle per scheduler (SMSP) from clock64().  The question it
// answers: which pipe does PRMT / FMNMX / FFMA / FFMA2 / HADD2.F32 issue to, and at what rate, so that the node step's
// instruction mix can be balanced.   build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdint>
#include <cstdio>
#include <cuda_fp16.h>

constexpr int kIters = 4096;

#define CHAINS8(M) M(0) M(1) M(2) M(3) M(4) M(5) M(6) M(7)

template <int OP>
__global__ void __launch_bounds__(512, 1) probe(float* out, long long* cycles, float fb, float fc, uint32_t ub)
{
    float    f[8], g[8];
    uint32_t u[8];
    unsigned long long d[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        f[i] = (float)(threadIdx.x + i) * 1e-3f;
        g[i] = f[i] + 1.0f;
        u[i] = threadIdx.x * 2654435761u + i;
        d[i] = ((unsigned long long)u[i] << 32) | u[i];
    }
    unsigned long long b2 = ((unsigned long long)__float_as_uint(fb) << 32) | __float_as_uint(fb);
    unsigned long long c2 = ((unsigned long long)__float_as_uint(fc) << 32) | __float_as_uint(fc);
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < kIters; ++it) {
        if (OP == 0) {
#define M(i) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(fb), "f"(fc));
            CHAINS8(M)
#undef M
        } else if (OP == 1) {
#define M(i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(d[i]) : "l"(b2), "l"(c2));
            CHAINS8(M)
#undef M
        } else if (OP == 2) {
#define M(i) asm volatile("prmt.b32 %0, %0, %1, 0x7604;" : "+r"(u[i]) : "r"(ub));
            CHAINS8(M)
#undef M
        } else if (OP == 3) {
#define M(i) asm volatile("max.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(fb));
            CHAINS8(M)
#undef M
        } else if (OP == 4) {
#define M(i) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(fb), "f"(fc));
            CHAINS8(M)
#undef M
        } else if (OP == 5) {
#define M(i) asm volatile("{.reg .f16 h; .reg .f32 x; mov.b32 {h, _}, %0; cvt.f32.f16 x, h; mov.b32 %0, x;}" : "+r"(u[i]));
            CHAINS8(M)
#undef M
        } else if (OP == 6) {
#define M(i) asm volatile("mad.lo.u32 %0, %0, %1, %1;" : "+r"(u[i]) : "r"(ub));
            CHAINS8(M)
#undef M
        } else if (OP == 7) {
#define M(i) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[i]) : "r"(ub), "r"(ub >> 3));
            CHAINS8(M)
#undef M
        } else if (OP == 8) { // FFMA + PRMT 1:1
#define M(i) asm volatile("fma.rn.f32 %0, %0, %2, %3; prmt.b32 %1, %1, %4, 0x7604;" : "+f"(f[i]), "+r"(u[i]) : "f"(fb), "f"(fc), "r"(ub));
            CHAINS8(M)
#undef M
        } else if (OP == 9) { // FFMA + 2 PRMT
#define M(i) asm volatile("fma.rn.f32 %0, %0, %2, %3; prmt.b32 %1, %1, %4, 0x7604; prmt.b32 %1, %1, %4, 0x7614;" : "+f"(f[i]), "+r"(u[i]) : "f"(fb), "f"(fc), "r"(ub));
            CHAINS8(M)
#undef M
        } else if (OP == 10) { // FFMA2 + PRMT 1:1
#define M(i) asm volatile("fma.rn.f32x2 %0, %0, %2, %3; prmt.b32 %1, %1, %4, 0x7604;" : "+l"(d[i]), "+r"(u[i]) : "l"(b2), "l"(c2), "r"(ub));
            CHAINS8(M)
#undef M
        } else if (OP == 11) { // the node step's per-child mix today: 6 PRMT, 6 FFMA, 2 FMNMX, 2 FMNMX3, 1 FMUL
#define M(i)                                                                                                         \
    asm volatile("{.reg .b32 a,b,c; .reg .f32 x,y,z;\n"                                                              \
                 "mov.b32 %1, %0;\n"                                                                                  \
                 "prmt.b32 a, %1, %4, 0x7604; prmt.b32 b, %1, %4, 0x7614; prmt.b32 c, %1, %4, 0x7624;\n"             \
                 "mov.b32 x, a; mov.b32 y, b; mov.b32 z, c;\n"                                                       \
                 "fma.rn.f32 x, x, %2, %3; fma.rn.f32 y, y, %2, %3; fma.rn.f32 z, z, %2, %3;\n"                      \
                 "max.f32 z, z, %3; max.f32 %0, x, y, z;\n"                                                          \
                 "prmt.b32 a, %1, %4, 0x7634; prmt.b32 b, %1, %4, 0x7605; prmt.b32 c, %1, %4, 0x7615;\n"             \
                 "mov.b32 x, a; mov.b32 y, b; mov.b32 z, c;\n"                                                       \
                 "fma.rn.f32 x, x, %2, %3; fma.rn.f32 y, y, %2, %3; fma.rn.f32 z, z, %2, %3;\n"                      \
                 "min.f32 x, x, y, z; mul.f32 x, x, %2; min.f32 x, x, %3; add.f32 %0, %0, x;}"                       \
                 : "+f"(f[i]), "+r"(u[i]) : "f"(fb), "f"(fc), "r"(ub));
            CHAINS8(M)
#undef M
        } else if (OP == 12) { // proposed mix: 3 PRMT (half2 pairs), 6 HADD2.F32, 3 FFMA2, 2 FMNMX, 2 FMNMX3
#define M(i)                                                                                                         \
    asm volatile("{.reg .b32 a,b,c; .reg .f16 h0,h1; .reg .f32 x0,x1,y0,y1,z0,z1; .reg .b64 p,q,r;\n"                \
                 "mov.b32 %1, %0;\n"                                                                                  \
                 "prmt.b32 a, %1, %4, 0x7170; prmt.b32 b, %1, %4, 0x7372; prmt.b32 c, %1, %4, 0x7071;\n"             \
                 "mov.b32 {h0,h1}, a; cvt.f32.f16 x0, h0; cvt.f32.f16 x1, h1;\n"                                     \
                 "mov.b32 {h0,h1}, b; cvt.f32.f16 y0, h0; cvt.f32.f16 y1, h1;\n"                                     \
                 "mov.b32 {h0,h1}, c; cvt.f32.f16 z0, h0; cvt.f32.f16 z1, h1;\n"                                     \
                 "mov.b64 p, {x0,x1}; mov.b64 q, {y0,y1}; mov.b64 r, {z0,z1};\n"                                     \
                 "fma.rn.f32x2 p, p, %5, %6; fma.rn.f32x2 q, q, %5, %6; fma.rn.f32x2 r, r, %5, %6;\n"                \
                 "mov.b64 {x0,x1}, p; mov.b64 {y0,y1}, q; mov.b64 {z0,z1}, r;\n"                                     \
                 "max.f32 z0, z0, %3; max.f32 x0, x0, y0, z0; min.f32 x1, x1, y1, z1; min.f32 x1, x1, %3;\n"         \
                 "add.f32 x0, x0, x1; add.f32 %0, %0, x0;}"                                                          \
                 : "+f"(f[i]), "+r"(u[i]) : "f"(fb), "f"(fc), "r"(ub), "l"(b2), "l"(c2));
            CHAINS8(M)
#undef M
        } else if (OP == 13) {
#define M(i) asm volatile("add.f32 %0, %0, %1;" : "+f"(f[i]) : "f"(fb));
            CHAINS8(M)
#undef M
        } else if (OP == 14) { // FSETP + predicated LOP3 (mask assembly)
#define M(i) asm volatile("{.reg .pred p; setp.le.f32 p, %2, %3; @p or.b32 %1, %1, %4; add.f32 %0, %0, %2;}" : "+f"(f[i]), "+r"(u[i]) : "f"(f[(i + 1) & 7]), "f"(fc), "r"(ub));
            CHAINS8(M)
#undef M
        } else if (OP == 15) { // FMNMX + FFMA 1:1
#define M(i) asm volatile("max.f32 %0, %0, %2; fma.rn.f32 %1, %1, %2, %3;" : "+f"(f[i]), "+f"(g[i]) : "f"(fb), "f"(fc));
            CHAINS8(M)
#undef M
        }
    }
    const long long t1 = clock64();
    float acc = 0.0f;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc += f[i] + g[i] + __uint_as_float(u[i]) + __uint_as_float((uint32_t)d[i]) + __uint_as_float((uint32_t)(d[i] >> 32));
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, double insts_per_iter, float* d_out, long long* d_cyc, int blocks)
{
    probe<OP><<<blocks, 512>>>(d_out, d_cyc, 1.0000001f, 1e-9f, 0x47000000u);
    probe<OP><<<blocks, 512>>>(d_out, d_cyc, 1.0000001f, 1e-9f, 0x47000000u);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    long long* h = new long long[blocks];
    cudaMemcpy(h, d_cyc, sizeof(long long) * blocks, cudaMemcpyDeviceToHost);
    double mx = 0;
    for (int i = 0; i < blocks; ++i) mx = h[i] > mx ? (double)h[i] : mx;
    delete[] h;
    // 16 warps per SM = 4 per SMSP; warp instructions per SMSP = 4 * kIters * insts_per_iter
    const double ipc = 4.0 * kIters * insts_per_iter / mx;
    printf("%-44s %7.3f warp-inst/cycle/SMSP  (%.0f cycles, %.0f insts/iter)\n", name, ipc, mx, insts_per_iter);
}

int main()
{
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float*     d_out;
    long long* d_cyc;
    cudaMalloc(&d_out, sizeof(float) * sms * 512);
    cudaMalloc(&d_cyc, sizeof(long long) * sms);
    run<0>("FFMA (3 registers)", 8, d_out, d_cyc, sms);
    run<1>("FFMA2 (f32x2)", 8, d_out, d_cyc, sms);
    run<2>("PRMT", 8, d_out, d_cyc, sms);
    run<3>("FMNMX", 8, d_out, d_cyc, sms);
    run<4>("FMNMX3", 8, d_out, d_cyc, sms);
    run<5>("HADD2.F32 (cvt.f32.f16)", 8, d_out, d_cyc, sms);
    run<6>("IMAD", 8, d_out, d_cyc, sms);
    run<7>("LOP3", 8, d_out, d_cyc, sms);
    run<13>("FADD", 8, d_out, d_cyc, sms);
    run<8>("FFMA + PRMT (1:1)", 16, d_out, d_cyc, sms);
    run<9>("FFMA + 2 PRMT", 24, d_out, d_cyc, sms);
    run<10>("FFMA2 + PRMT (1:1)", 16, d_out, d_cyc, sms);
    run<15>("FMNMX + FFMA (1:1)", 16, d_out, d_cyc, sms);
    run<14>("FSETP + @P LOP3 + FADD", 24, d_out, d_cyc, sms);
    run<11>("node-step child mix today (19 insts)", 8 * 19, d_out, d_cyc, sms);
    run<12>("node-step child mix proposed (19 insts)", 8 * 19, d_out, d_cyc, sms);
    return 0;
}
