
// mix_probe2.cu — issue-slot cost of each ingredient of the sweep body (diagnostic).
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ float ffma(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ float fadd_as_fma(float a, float b) { float d; asm volatile("fma.rn.f32 %0, %1, 0f3F800000, %2;" : "=f"(d) : "f"(a), "f"(b)); return d; }
__device__ __forceinline__ float fmul_as_fma(float a, float b) { float d; asm volatile("fma.rn.f32 %0, %1, %2, 0f80000000;" : "=f"(d) : "f"(a), "f"(b)); return d; }

// LDS: sphere from shared memory (else from 8 register-resident spheres); ALLFMA: adds/muls issued as FFMA;
// SINK: 0 = FMNMX per test, 1 = SHF per test, 2 = FMNMX3 per two tests
template <bool LDS, bool ALLFMA, int SINK>
__global__ void __launch_bounds__(256, 4) k_mix(const float* in, unsigned* out, int iters) {
    __shared__ float4 sm[512];
    for (int i = threadIdx.x; i < 512; i += 256) sm[i] = make_float4(in[i & 63], in[(i + 1) & 63], in[(i + 2) & 63], in[(i + 3) & 63]);
    __syncthreads();
    float ox = in[threadIdx.x & 63], oy = in[(threadIdx.x + 7) & 63], oz = in[(threadIdx.x + 9) & 63];
    float dx = in[(threadIdx.x + 11) & 63], dy = in[(threadIdx.x + 13) & 63], dz = in[(threadIdx.x + 17) & 63], a = in[(threadIdx.x + 19) & 63];
    float one = in[63] * 0.0f + 1.0f, nzero = -(in[62] * 0.0f);
    float4 rs[8];
#pragma unroll
    for (int i = 0; i < 8; i++) rs[i] = sm[(threadIdx.x + i) & 511];
    unsigned mask = 0;
    float facc = -1e30f;
    for (int it = 0; it < iters; it++) {
        float prev = 0.f;
#pragma unroll
        for (int j = 0; j < 32; j++) {
            float4 s = LDS ? sm[(it & 15) * 32 + j] : rs[j & 7];
            float ocx, ocy, ocz, hb, cc, disc;
            if (ALLFMA) {
                ocx = fadd_as_fma(ox, -s.x); ocy = fadd_as_fma(oy, -s.y); ocz = fadd_as_fma(oz, -s.z);
                hb = ffma(ocz, dz, ffma(ocy, dy, fmul_as_fma(ocx, dx)));
                cc = fadd_as_fma(ffma(ocz, ocz, ffma(ocy, ocy, fmul_as_fma(ocx, ocx))), -s.w);
                disc = ffma(hb, hb, -fmul_as_fma(a, cc));
            } else {
                ocx = __fsub_rn(ox, s.x); ocy = __fsub_rn(oy, s.y); ocz = __fsub_rn(oz, s.z);
                hb = __fmaf_rn(ocz, dz, __fmaf_rn(ocy, dy, __fmul_rn(ocx, dx)));
                cc = __fsub_rn(__fmaf_rn(ocz, ocz, __fmaf_rn(ocy, ocy, __fmul_rn(ocx, ocx))), s.w);
                disc = __fmaf_rn(hb, hb, -__fmul_rn(a, cc));
            }
            if (SINK == 0) facc = fmaxf(facc, disc);
            else if (SINK == 1) mask = __funnelshift_l(__float_as_uint(disc), mask, 1);
            else { if (j & 1) facc = fmaxf(fmaxf(facc, disc), prev); else prev = disc; }
        }
        ox = __fadd_rn(ox, 0.001f);
    }
    out[blockIdx.x * 256 + threadIdx.x] = mask + __float_as_uint(facc);
}

template <bool LDS, bool ALLFMA, int SINK>
void run(const char* name, int nsm, const float* d_in, unsigned* d_out, double other_per_test) {
    int iters = 4000;
    int grid = nsm * 4;
    k_mix<LDS, ALLFMA, SINK><<<grid, 256>>>(d_in, d_out, 10);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int r = 0; r < 3; r++) {
        CK(cudaEventRecord(e0));
        k_mix<LDS, ALLFMA, SINK><<<grid, 256>>>(d_in, d_out, iters);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    double tests = (double)iters * 32 * 8.0;              // warp-tests per SMSP
    double cycles = best * 1e-3 * 1.965e9;
    printf("%-44s ms=%8.3f  cycles/warp-test=%6.3f  instr/test=%5.2f  IPC=%.3f  frac_of_FP32_peak=%.3f\n", name, best, cycles / tests,
           12 + other_per_test, (12 + other_per_test) * tests / cycles, 8.5 * tests / cycles);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int nsm = p.multiProcessorCount;
    float h[64];
    for (int i = 0; i < 64; i++) h[i] = 0.5f + 0.01f * i;
    float* d_in; unsigned* d_out;
    CK(cudaMalloc(&d_in, sizeof(h))); CK(cudaMemcpy(d_in, h, sizeof(h), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_out, nsm * 4 * 256 * 4));
    run<false, false, 0>("regs, mixed body, FMNMX", nsm, d_in, d_out, 1);
    run<false, true, 0>("regs, all-FFMA body, FMNMX", nsm, d_in, d_out, 1);
    run<true, false, 0>("LDS, mixed body, FMNMX", nsm, d_in, d_out, 2);
    run<true, true, 0>("LDS, all-FFMA body, FMNMX", nsm, d_in, d_out, 2);
    run<true, false, 1>("LDS, mixed body, SHF", nsm, d_in, d_out, 2);
    run<true, true, 1>("LDS, all-FFMA body, SHF", nsm, d_in, d_out, 2);
    run<true, false, 2>("LDS, mixed body, FMNMX3 per 2", nsm, d_in, d_out, 1.5);
    run<true, true, 2>("LDS, all-FFMA body, FMNMX3 per 2", nsm, d_in, d_out, 1.5);
    run<false, false, 2>("regs, mixed body, FMNMX3 per 2", nsm, d_in, d_out, 0.5);
    return 0;
}
