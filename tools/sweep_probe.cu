// sweep_probe.cu — design-space probe for the closest-hit sweep inner loop (diagnostic, not product code).
// Measures cycles per ray-sphere test for sweep variants on synthetic rays/spheres:
//   RAYS  : rays per thread (1 or 2) — one LDS.128 per sphere is shared by RAYS tests
//   UNR   : spheres per unrolled body
//   W     : warps per block configuration is fixed at 256 threads; blocks/SM chosen by registers
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -o sweep_probe sweep_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

template <int RAYS, int UNR, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_sweep(const float4* __restrict__ g_sph, int padded, const float* __restrict__ rays,
                                                         int reps, unsigned* __restrict__ out, long long* cyc) {
    extern __shared__ float4 s_sph[];
    for (int i = threadIdx.x; i < padded; i += THREADS) s_sph[i] = g_sph[i];
    __syncthreads();
    float ox[RAYS], oy[RAYS], oz[RAYS], dx[RAYS], dy[RAYS], dz[RAYS], a[RAYS];
    long long tid = (long long)blockIdx.x * THREADS + threadIdx.x;
#pragma unroll
    for (int r = 0; r < RAYS; r++) {
        const float* p = rays + (tid * RAYS + r) * 6;
        ox[r] = p[0]; oy[r] = p[1]; oz[r] = p[2]; dx[r] = p[3]; dy[r] = p[4]; dz[r] = p[5];
        a[r] = __fmaf_rn(dz[r], dz[r], __fmaf_rn(dy[r], dy[r], __fmul_rn(dx[r], dx[r])));
    }
    unsigned acc = 0;
    long long t0 = clock64();
    for (int rep = 0; rep < reps; rep++) {
        for (int blk = 0; blk < padded; blk += 32) {
            unsigned mask[RAYS];
#pragma unroll
            for (int r = 0; r < RAYS; r++) mask[r] = 0;
#pragma unroll 1
            for (int sub = 0; sub < 32; sub += UNR) {
#pragma unroll
                for (int j = 0; j < UNR; j++) {
                    float4 s = s_sph[blk + sub + j];
#pragma unroll
                    for (int r = 0; r < RAYS; r++) {
                        float ocx = __fsub_rn(ox[r], s.x), ocy = __fsub_rn(oy[r], s.y), ocz = __fsub_rn(oz[r], s.z);
                        float hb = __fmaf_rn(ocz, dz[r], __fmaf_rn(ocy, dy[r], __fmul_rn(ocx, dx[r])));
                        float cc = __fsub_rn(__fmaf_rn(ocz, ocz, __fmaf_rn(ocy, ocy, __fmul_rn(ocx, ocx))), s.w);
                        float disc = __fmaf_rn(hb, hb, -__fmul_rn(a[r], cc));
                        mask[r] = __funnelshift_l(__float_as_uint(disc), mask[r], 1);
                    }
                }
            }
#pragma unroll
            for (int r = 0; r < RAYS; r++) acc += __popc(~mask[r]);
        }
#pragma unroll
        for (int r = 0; r < RAYS; r++) ox[r] = __fadd_rn(ox[r], 0.001f);   // new ray each repetition
    }
    long long t1 = clock64();
    out[tid] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int RAYS, int UNR, int THREADS, int MINB>
void run(const char* name, int nsm, const float4* d_sph, int padded, const float* d_rays, unsigned* d_out, long long* d_cyc) {
    int reps = 40;
    size_t smem = (size_t)padded * sizeof(float4);
    CK(cudaFuncSetAttribute(k_sweep<RAYS, UNR, THREADS, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_sweep<RAYS, UNR, THREADS, MINB>, THREADS, smem));
    int grid = nsm * occ;
    cudaFuncAttributes fa; CK(cudaFuncGetAttributes(&fa, k_sweep<RAYS, UNR, THREADS, MINB>));
    k_sweep<RAYS, UNR, THREADS, MINB><<<grid, THREADS, smem>>>(d_sph, padded, d_rays, 2, d_out, d_cyc);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int it = 0; it < 3; it++) {
        CK(cudaEventRecord(e0));
        k_sweep<RAYS, UNR, THREADS, MINB><<<grid, THREADS, smem>>>(d_sph, padded, d_rays, reps, d_out, d_cyc);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    double tests = (double)grid * THREADS * RAYS * (double)padded * reps;
    double tps = tests / (best * 1e-3);
    // cycles per warp-test per SMSP: warps per SMSP = occ*THREADS/32/4
    long long c0; CK(cudaMemcpy(&c0, d_cyc, 8, cudaMemcpyDeviceToHost));
    double warp_tests_per_smsp = (double)occ * THREADS / 32.0 / 4.0 * RAYS * (double)padded * reps;
    printf("%-34s regs=%3d occ=%d blocks/SM  ms=%7.3f  tests/s=%.3e  TFLOP/s(17)=%6.2f  cyc/warp-test/SMSP=%6.2f\n", name, fa.numRegs, occ,
           best, tps, tps * 17e-12, (double)c0 / warp_tests_per_smsp);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int nsm = p.multiProcessorCount;
    const int padded = 512;
    std::vector<float4> sph(padded);
    srand(1);
    auto rf = []() { return (float)rand() / RAND_MAX; };
    for (int i = 0; i < padded; i++) sph[i] = make_float4(rf() * 22 - 11, 0.2f, rf() * 22 - 11, 0.04f);
    size_t nthreads = (size_t)nsm * 2048;
    std::vector<float> rays(nthreads * 2 * 6);
    for (size_t i = 0; i < rays.size() / 6; i++) {
        rays[6 * i] = rf() * 20 - 10; rays[6 * i + 1] = rf() * 3; rays[6 * i + 2] = rf() * 20 - 10;
        rays[6 * i + 3] = rf() * 2 - 1; rays[6 * i + 4] = rf() * 2 - 1; rays[6 * i + 5] = rf() * 2 - 1;
    }
    float4* d_sph; float* d_rays; unsigned* d_out; long long* d_cyc;
    CK(cudaMalloc(&d_sph, padded * sizeof(float4))); CK(cudaMemcpy(d_sph, sph.data(), padded * sizeof(float4), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_rays, rays.size() * 4)); CK(cudaMemcpy(d_rays, rays.data(), rays.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_out, nthreads * 4)); CK(cudaMalloc(&d_cyc, 8 * nsm * 16));
    printf("device %s SMs=%d\n", p.name, nsm);
    run<1, 32, 256, 4>("1 ray  unroll32 256x4", nsm, d_sph, padded, d_rays, d_out, d_cyc);
    run<1, 16, 256, 4>("1 ray  unroll16 256x4", nsm, d_sph, padded, d_rays, d_out, d_cyc);
    run<1, 8, 256, 4>("1 ray  unroll8  256x4", nsm, d_sph, padded, d_rays, d_out, d_cyc);
    run<1, 4, 256, 4>("1 ray  unroll4  256x4", nsm, d_sph, padded, d_rays, d_out, d_cyc);
    run<1, 32, 256, 8>("1 ray  unroll32 256x8 (32 regs)", nsm, d_sph, padded, d_rays, d_out, d_cyc);
    run<1, 8, 256, 8>("1 ray  unroll8  256x8 (32 regs)", nsm, d_sph, padded, d_rays, d_out, d_cyc);
    run<1, 32, 256, 2>("1 ray  unroll32 256x2", nsm, d_sph, padded, d_rays, d_out, d_cyc);
    run<1, 8, 256, 2>("1 ray  unroll8  256x2", nsm, d_sph, padded, d_rays, d_out, d_cyc);
    run<2, 32, 256, 4>("2 rays unroll32 256x4", nsm, d_sph, padded, d_rays, d_out, d_cyc);
    run<2, 16, 256, 4>("2 rays unroll16 256x4", nsm, d_sph, padded, d_rays, d_out, d_cyc);
    run<2, 8, 256, 4>("2 rays unroll8  256x4", nsm, d_sph, padded, d_rays, d_out, d_cyc);
    run<2, 4, 256, 4>("2 rays unroll4  256x4", nsm, d_sph, padded, d_rays, d_out, d_cyc);
    run<2, 16, 256, 2>("2 rays unroll16 256x2", nsm, d_sph, padded, d_rays, d_out, d_cyc);
    run<2, 8, 256, 2>("2 rays unroll8  256x2", nsm, d_sph, padded, d_rays, d_out, d_cyc);
    run<2, 4, 256, 2>("2 rays unroll4  256x2", nsm, d_sph, padded, d_rays, d_out, d_cyc);
    return 0;
}
