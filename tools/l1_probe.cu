// L1 data-pipe probe: how many cycles does a warp-wide LDG.128 take when every lane reads a DIFFERENT 128-byte line
// (one BVH node per lane), as a function of the 16-byte offset inside the line that the lanes use?
//   mode 0: all lanes use the same offset k (what an SoA node layout does: "load field k of my node")
//   mode 1: lane L uses offset (k + L) & 7 (fields rotated per lane)
//   mode 2: coalesced: lanes read consecutive 16-byte words (4 lines per request)
//   mode 3: same offset, LDG.64 x2 instead of LDG.128
//   mode 4: lane L uses offset (k + (L >> 2)) & 7 (rotation per group of four lanes)
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/l1_probe tools/l1_probe.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(128, 8) probe(const float4* __restrict__ nodes, int n_nodes, int iters, float* out, long long* cyc)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // each warp walks its own window of 32 nodes (4 KB), hopping pseudo-randomly inside a 256-node (32 KB) CTA region: L1 resident
    const int base = 0;    // every CTA uses the same 32 KB: L1 resident on every SM
    unsigned s = lane * 7 + warp * 61;
    float acc = 0.f;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        s = s * 1664525u + 1013904223u;
        const int node = base + ((s >> 8) & 255);
        const float4* p = nodes + (size_t)node * 8;
#pragma unroll
        for (int k = 0; k < 7; k++) {
            float4 v;
            if (MODE == 0) v = __ldg(p + k);
            else if (MODE == 1) v = __ldg(p + ((k + lane) & 7));
            else if (MODE == 2) v = __ldg(nodes + (size_t)base * 8 + ((((s >> 8) & 63) * 32 + lane + k * 32) & 2047));
            else if (MODE == 3) { const float2* q = (const float2*)(p + k); float2 a = __ldg(q), b = __ldg(q + 1); v = make_float4(a.x, a.y, b.x, b.y); }
            else v = __ldg(p + ((k + (lane >> 2)) & 7));
            acc += v.x + v.y + v.z + v.w;
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main()
{
    const int n_nodes = 1 << 16, iters = 2000, grid = 148 * 8;
    float4* nodes; float* out; long long* cyc;
    cudaMalloc(&nodes, sizeof(float4) * 8 * n_nodes);
    cudaMemset(nodes, 0, sizeof(float4) * 8 * n_nodes);
    cudaMalloc(&out, sizeof(float) * grid * 128);
    cudaMalloc(&cyc, sizeof(long long) * grid);
    for (int mode = 0; mode < 5; mode++) {
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            switch (mode) {
            case 0: probe<0><<<grid, 128>>>(nodes, n_nodes, iters, out, cyc); break;
            case 1: probe<1><<<grid, 128>>>(nodes, n_nodes, iters, out, cyc); break;
            case 2: probe<2><<<grid, 128>>>(nodes, n_nodes, iters, out, cyc); break;
            case 3: probe<3><<<grid, 128>>>(nodes, n_nodes, iters, out, cyc); break;
            default: probe<4><<<grid, 128>>>(nodes, n_nodes, iters, out, cyc); break;
            }
            cudaEventRecord(e1); cudaEventSynchronize(e1);
        }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        // 32 warps per SM, each issuing iters * 7 requests
        const double req_per_sm = 32.0 * iters * 7, cycles = ms * 1e-3 * 1.965e9;
        printf("mode %d: %.3f ms, %.2f SM cycles per warp-wide request (%.1f bytes/cycle/SM)\n", mode, ms, cycles / req_per_sm, 512.0 * req_per_sm / cycles);
    }
    printf("err=%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
