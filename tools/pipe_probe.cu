// pipe_probe.cu -- design tool (not product code): issue rate of the SASS ops a compressed-wide-node box test is made of,
// in warp-instructions per clock per SM (4.0 = one per clock per SM sub-partition).  Each thread runs 8 independent
// dependency chains of one op; 1024 threads per SM hide the latency.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o tools/pipe_probe tools/pipe_probe.cu && tools/pipe_probe
#include <cstdio>
#include <cuda_runtime.h>

enum { OP_FFMA, OP_FMNMX, OP_FMNMX3, OP_PRMT, OP_I2F_U8, OP_I2FP, OP_F16, OP_LOP3, OP_SHF, OP_IMAD, OP_SEL, OP_FSETP_SEL, OP_POPC, OP_FLO,
       OP_MIX_DEC_I2F, OP_MIX_DEC_PRMT, OP_MIX_DEC_BF16, OP_N };
static const char* NAMES[] = { "FFMA", "FMNMX", "FMNMX3", "PRMT", "I2F.U8", "I2FP.F32.U32", "HADD2.F32 (f16->f32)", "LOP3", "SHF", "IMAD", "SEL",
                               "FSETP+SEL", "POPC", "FLO (bfind)", "mix: I2F.U8 + FFMA", "mix: PRMT + FFMA", "mix: (SHF|LOP3) + FFMA" };

template <int OP>
__device__ __forceinline__ void step(unsigned& x, float& f, unsigned c, float g, float h)
{
    if (OP == OP_FFMA) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(g), "f"(h));
    if (OP == OP_FMNMX) asm volatile("max.f32 %0, %0, %1;" : "+f"(f) : "f"(g));
    if (OP == OP_FMNMX3) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(g), "f"(h));
    if (OP == OP_PRMT) asm volatile("prmt.b32 %0, %0, %1, 0x7440;" : "+r"(x) : "r"(c));
    if (OP == OP_I2F_U8) { asm volatile("cvt.rn.f32.u8 %0, %1;" : "=f"(f) : "r"(x)); x = __float_as_uint(f); }
    if (OP == OP_I2FP) { asm volatile("cvt.rn.f32.u32 %0, %1;" : "=f"(f) : "r"(x)); x = __float_as_uint(f); }
    if (OP == OP_F16) { asm volatile("{ .reg .b16 lo, hi; mov.b32 {lo, hi}, %1; cvt.f32.f16 %0, lo; }" : "=f"(f) : "r"(x)); x = __float_as_uint(f); }
    if (OP == OP_LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x) : "r"(c), "r"(__float_as_uint(g)));
    if (OP == OP_SHF) asm volatile("shf.l.wrap.b32 %0, %0, %1, %2;" : "+r"(x) : "r"(c), "r"(5u));
    if (OP == OP_IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x) : "r"(c), "r"(__float_as_uint(g)));
    if (OP == OP_SEL) asm volatile("{ .reg .pred p; setp.ne.u32 p, %1, 0; selp.b32 %0, %0, %2, p; }" : "+r"(x) : "r"(c), "r"(__float_as_uint(g)));
    if (OP == OP_FSETP_SEL) asm volatile("{ .reg .pred p; setp.le.f32 p, %0, %1; selp.f32 %0, %0, %2, p; }" : "+f"(f) : "f"(g), "f"(h));
    if (OP == OP_POPC) asm volatile("popc.b32 %0, %0;" : "+r"(x));
    if (OP == OP_FLO) asm volatile("bfind.u32 %0, %0;" : "+r"(x));
    if (OP == OP_MIX_DEC_I2F) { float t; asm volatile("cvt.rn.f32.u8 %0, %1;" : "=f"(t) : "r"(x)); asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(f) : "f"(t), "f"(g)); }
    if (OP == OP_MIX_DEC_PRMT) { unsigned t; asm volatile("prmt.b32 %0, %1, %2, 0x7440;" : "=r"(t) : "r"(x), "r"(c)); asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(f) : "f"(__uint_as_float(t)), "f"(g)); }
    if (OP == OP_MIX_DEC_BF16) { unsigned t, u; asm volatile("shl.b32 %0, %1, 16;" : "=r"(t) : "r"(x)); asm volatile("and.b32 %0, %1, 0xffff0000;" : "=r"(u) : "r"(x));
                                 asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(f) : "f"(__uint_as_float(t)), "f"(g));
                                 asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(f) : "f"(__uint_as_float(u)), "f"(g)); }
}

template <int OP>
__global__ void __launch_bounds__(256) k_probe(unsigned* out, unsigned seed, int iters, float g, float h)
{
    unsigned x[8]; float f[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { x[i] = seed * (threadIdx.x + 1) + i; f[i] = (float)(threadIdx.x + i); }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 8; r++)
#pragma unroll
            for (int i = 0; i < 8; i++) step<OP>(x[i], f[i], seed, g, h);
    }
    unsigned acc = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) acc += x[i] + __float_as_uint(f[i]);
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int OP>
static void run(unsigned* out, int sms, double ghz)
{
    const int iters = 2000, grid = sms * 4;
    k_probe<OP><<<grid, 256>>>(out, 12345u, 10, 1.0001f, 0.5f);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    k_probe<OP><<<grid, 256>>>(out, 12345u, iters, 1.0001f, 0.5f);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms = 0; cudaEventElapsedTime(&ms, a, b);
    int per_step = OP == OP_MIX_DEC_BF16 ? 4 : (OP >= OP_MIX_DEC_I2F || OP == OP_SEL || OP == OP_FSETP_SEL ? 2 : 1);
    const double winst = (double)grid * 8 /*warps*/ * iters * 64.0 * per_step;
    printf("%-26s %8.3f ms  %6.2f warp-inst/clk/SM (SASS ops per step: %d)\n", NAMES[OP], ms, winst / (ms * 1e-3 * ghz * 1e9 * sms), per_step);
}

int main()
{
    int sms = 0, khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double ghz = khz * 1e-6;
    printf("SMs %d, clock %.3f GHz (nominal max; rates assume the clock holds)\n", sms, ghz);
    unsigned* out; cudaMalloc(&out, sizeof(unsigned) * sms * 4 * 256);
    run<OP_FFMA>(out, sms, ghz); run<OP_FMNMX>(out, sms, ghz); run<OP_FMNMX3>(out, sms, ghz); run<OP_PRMT>(out, sms, ghz);
    run<OP_I2F_U8>(out, sms, ghz); run<OP_I2FP>(out, sms, ghz); run<OP_F16>(out, sms, ghz); run<OP_LOP3>(out, sms, ghz);
    run<OP_SHF>(out, sms, ghz); run<OP_IMAD>(out, sms, ghz); run<OP_SEL>(out, sms, ghz); run<OP_FSETP_SEL>(out, sms, ghz);
    run<OP_POPC>(out, sms, ghz); run<OP_FLO>(out, sms, ghz);
    run<OP_MIX_DEC_I2F>(out, sms, ghz); run<OP_MIX_DEC_PRMT>(out, sms, ghz); run<OP_MIX_DEC_BF16>(out, sms, ghz);
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
