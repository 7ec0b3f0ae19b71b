
// mix_probe3.cu — does op-major ordering (operand reuse) lift the sweep body's issue rate? (diagnostic)
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

#ifdef PIN
__device__ __forceinline__ float fsub(float a, float b) { float d; asm volatile("sub.rn.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b)); return d; }
__device__ __forceinline__ float fmul(float a, float b) { float d; asm volatile("mul.rn.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b)); return d; }
__device__ __forceinline__ float ffma(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
#else
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float ffma(float a, float b, float c) { return __fmaf_rn(a, b, c); }
#endif

// G = spheres processed op-major together (G = 1 is the sphere-major baseline)
template <int G>
__global__ void __launch_bounds__(256, 4) k_mix(const float* in, unsigned* out, int iters) {
    __shared__ float4 sm[512];
    for (int i = threadIdx.x; i < 512; i += 256) sm[i] = make_float4(in[i & 63], in[(i + 1) & 63], in[(i + 2) & 63], in[(i + 3) & 63]);
    __syncthreads();
    float ox = in[threadIdx.x & 63], oy = in[(threadIdx.x + 7) & 63], oz = in[(threadIdx.x + 9) & 63];
    float dx = in[(threadIdx.x + 11) & 63], dy = in[(threadIdx.x + 13) & 63], dz = in[(threadIdx.x + 17) & 63], a = in[(threadIdx.x + 19) & 63];
    unsigned mask = 0;
    for (int it = 0; it < iters; it++) {
        const float4* base = sm + (it & 15) * 32;
#pragma unroll
        for (int j = 0; j < 32; j += G) {
            float4 s[G];
            float ocx[G], ocy[G], ocz[G], hb[G], cc[G], ac[G], disc[G];
#pragma unroll
            for (int g = 0; g < G; g++) s[g] = base[j + g];
#pragma unroll
            for (int g = 0; g < G; g++) ocx[g] = fsub(ox, s[g].x);
#pragma unroll
            for (int g = 0; g < G; g++) ocy[g] = fsub(oy, s[g].y);
#pragma unroll
            for (int g = 0; g < G; g++) ocz[g] = fsub(oz, s[g].z);
#pragma unroll
            for (int g = 0; g < G; g++) hb[g] = fmul(ocx[g], dx);
#pragma unroll
            for (int g = 0; g < G; g++) cc[g] = fmul(ocx[g], ocx[g]);
#pragma unroll
            for (int g = 0; g < G; g++) hb[g] = ffma(ocy[g], dy, hb[g]);
#pragma unroll
            for (int g = 0; g < G; g++) cc[g] = ffma(ocy[g], ocy[g], cc[g]);
#pragma unroll
            for (int g = 0; g < G; g++) hb[g] = ffma(ocz[g], dz, hb[g]);
#pragma unroll
            for (int g = 0; g < G; g++) cc[g] = ffma(ocz[g], ocz[g], cc[g]);
#pragma unroll
            for (int g = 0; g < G; g++) cc[g] = fsub(cc[g], s[g].w);
#pragma unroll
            for (int g = 0; g < G; g++) ac[g] = fmul(a, cc[g]);
#pragma unroll
            for (int g = 0; g < G; g++) disc[g] = ffma(hb[g], hb[g], -ac[g]);
#pragma unroll
            for (int g = 0; g < G; g++) mask = __funnelshift_l(__float_as_uint(disc[g]), mask, 1);
        }
        ox = __fadd_rn(ox, 0.001f);
    }
    out[blockIdx.x * 256 + threadIdx.x] = mask;
}

template <int G>
void run(const char* name, int nsm, const float* d_in, unsigned* d_out) {
    int iters = 4000;
    int grid = nsm * 4;
    cudaFuncAttributes fa; CK(cudaFuncGetAttributes(&fa, k_mix<G>));
    k_mix<G><<<grid, 256>>>(d_in, d_out, 10);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int r = 0; r < 3; r++) {
        CK(cudaEventRecord(e0));
        k_mix<G><<<grid, 256>>>(d_in, d_out, iters);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    double tests = (double)iters * 32 * 8.0;
    double cycles = best * 1e-3 * 1.965e9;
    printf("%-28s regs=%3d ms=%8.3f  cycles/warp-test=%6.3f  IPC=%.3f  frac_of_FP32_peak=%.3f\n", name, fa.numRegs, best, cycles / tests,
           14 * tests / cycles, 8.5 * tests / cycles);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int nsm = p.multiProcessorCount;
    float h[64];
    for (int i = 0; i < 64; i++) h[i] = 0.5f + 0.01f * i;
    float* d_in; unsigned* d_out;
    CK(cudaMalloc(&d_in, sizeof(h))); CK(cudaMemcpy(d_in, h, sizeof(h), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_out, nsm * 4 * 256 * 4));
    run<1>("G=1 sphere-major", nsm, d_in, d_out);
    run<2>("G=2 op-major", nsm, d_in, d_out);
    run<4>("G=4 op-major", nsm, d_in, d_out);
    run<8>("G=8 op-major", nsm, d_in, d_out);
    return 0;
}
