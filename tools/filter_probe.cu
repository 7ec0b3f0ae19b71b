// filter_probe.cu — design-space probe for the candidate-filter sweep loop (diagnostic, not product code).
// Where should the warp-uniform sphere operand of the 6 chain FFMAs come from?
//   SRC 0: shared memory (LDS.128 -> 4 registers; every chain FFMA then reads 3 registers)
//   SRC 1: __constant__ array (LDCU.128 -> 4 uniform registers; every chain FFMA reads 2 registers + 1 uniform)
//   SRC 2: kernel parameter space (__grid_constant__ struct)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -o filter_probe filter_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

constexpr int NS = 512;
__constant__ float4 c_sph[NS];
struct ParamScene { float4 s[NS]; };

template <int SRC, int RAYS, int UNR, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_sweep(const float4* __restrict__ g_sph, const float* __restrict__ rays, int reps,
                                                         unsigned* __restrict__ out, const __grid_constant__ ParamScene ps) {
    extern __shared__ float4 s_sph[];
    if (SRC == 0) {
        for (int i = threadIdx.x; i < NS; i += THREADS) s_sph[i] = g_sph[i];
        __syncthreads();
    }
    float hx[RAYS], hy[RAYS], hz[RAYS], nod[RAYS], gx[RAYS], gy[RAYS], gz[RAYS], oo[RAYS];
    long long tid = (long long)blockIdx.x * THREADS + threadIdx.x;
#pragma unroll
    for (int r = 0; r < RAYS; r++) {
        const float* p = rays + (tid * RAYS + r) * 6;
        float il = rsqrtf(p[3] * p[3] + p[4] * p[4] + p[5] * p[5]);
        hx[r] = p[3] * il; hy[r] = p[4] * il; hz[r] = p[5] * il;
        nod[r] = -(p[0] * hx[r] + p[1] * hy[r] + p[2] * hz[r]);
        gx[r] = -2 * p[0]; gy[r] = -2 * p[1]; gz[r] = -2 * p[2];
        oo[r] = p[0] * p[0] + p[1] * p[1] + p[2] * p[2];
    }
    unsigned acc = 0;
    for (int rep = 0; rep < reps; rep++) {
        for (int blk = 0; blk < NS; blk += 32) {
            unsigned mask[RAYS];
#pragma unroll
            for (int r = 0; r < RAYS; r++) mask[r] = 0;
#pragma unroll 1
            for (int sub = 0; sub < 32; sub += UNR) {
#pragma unroll
                for (int j = 0; j < UNR; j++) {
                    float4 s = SRC == 0 ? s_sph[blk + sub + j] : SRC == 1 ? c_sph[blk + sub + j] : ps.s[blk + sub + j];
#pragma unroll
                    for (int r = 0; r < RAYS; r++) {
                        float h = __fmaf_rn(s.x, hx[r], __fmaf_rn(s.y, hy[r], __fmaf_rn(s.z, hz[r], nod[r])));
                        float g = __fmaf_rn(s.x, gx[r], __fmaf_rn(s.y, gy[r], __fmaf_rn(s.z, gz[r], oo[r])));
                        float w = __fadd_rn(g, s.w);
                        float q = __fmaf_rn(h, h, -w);
                        mask[r] = __funnelshift_l(__float_as_uint(q), mask[r], 1);
                    }
                }
            }
#pragma unroll
            for (int r = 0; r < RAYS; r++) acc += __popc(~mask[r]);
        }
#pragma unroll
        for (int r = 0; r < RAYS; r++) nod[r] = __fadd_rn(nod[r], 0.001f);   // new ray each repetition
    }
    out[tid] = acc;
}

template <int SRC, int RAYS, int UNR, int THREADS, int MINB>
void run(const char* name, int nsm, const float4* d_sph, const float* d_rays, unsigned* d_out, const ParamScene& ps) {
    int reps = 40;
    size_t smem = SRC == 0 ? (size_t)NS * sizeof(float4) : 0;
    auto k = k_sweep<SRC, RAYS, UNR, THREADS, MINB>;
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, THREADS, smem));
    int grid = nsm * occ;
    cudaFuncAttributes fa; CK(cudaFuncGetAttributes(&fa, k));
    k<<<grid, THREADS, smem>>>(d_sph, d_rays, 2, d_out, ps);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int it = 0; it < 3; it++) {
        CK(cudaEventRecord(e0));
        k<<<grid, THREADS, smem>>>(d_sph, d_rays, reps, d_out, ps);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    double tests = (double)grid * THREADS * RAYS * (double)NS * reps;
    double tps = tests / (best * 1e-3);
    double cyc = 1.965e9 * (best * 1e-3) / ((double)occ * THREADS / 32.0 / 4.0 * RAYS * (double)NS * reps);
    printf("%-40s regs=%3d occ=%d  ms=%7.3f  tests/s=%.3e  TFLOP/s(17)=%6.2f  cyc/warp-test/SMSP@1965=%5.2f\n", name, fa.numRegs, occ, best, tps,
           tps * 17e-12, cyc);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int nsm = p.multiProcessorCount;
    static ParamScene ps;
    srand(1);
    auto rf = []() { return (float)rand() / RAND_MAX; };
    for (int i = 0; i < NS; i++) ps.s[i] = make_float4(rf() * 22 - 11, 0.2f, rf() * 22 - 11, rf() * 200);
    size_t nthreads = (size_t)nsm * 2048;
    std::vector<float> rays(nthreads * 2 * 6);
    for (size_t i = 0; i < rays.size() / 6; i++) {
        rays[6 * i] = rf() * 20 - 10; rays[6 * i + 1] = rf() * 3; rays[6 * i + 2] = rf() * 20 - 10;
        rays[6 * i + 3] = rf() * 2 - 1; rays[6 * i + 4] = rf() * 2 - 1; rays[6 * i + 5] = rf() * 2 - 1;
    }
    float4* d_sph; float* d_rays; unsigned* d_out;
    CK(cudaMalloc(&d_sph, NS * sizeof(float4))); CK(cudaMemcpy(d_sph, ps.s, NS * sizeof(float4), cudaMemcpyHostToDevice));
    CK(cudaMemcpyToSymbol(c_sph, ps.s, NS * sizeof(float4)));
    CK(cudaMalloc(&d_rays, rays.size() * 4)); CK(cudaMemcpy(d_rays, rays.data(), rays.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_out, nthreads * 4));
    printf("device %s SMs=%d\n", p.name, nsm);
    run<0, 1, 32, 256, 4>("smem  1 ray  unroll32 256x4", nsm, d_sph, d_rays, d_out, ps);
    run<0, 2, 16, 256, 4>("smem  2 rays unroll16 256x4", nsm, d_sph, d_rays, d_out, ps);
    run<0, 2, 16, 256, 3>("smem  2 rays unroll16 256x3", nsm, d_sph, d_rays, d_out, ps);
    run<1, 1, 32, 256, 4>("const 1 ray  unroll32 256x4", nsm, d_sph, d_rays, d_out, ps);
    run<1, 1, 16, 256, 4>("const 1 ray  unroll16 256x4", nsm, d_sph, d_rays, d_out, ps);
    run<1, 1, 8, 256, 4>("const 1 ray  unroll8  256x4", nsm, d_sph, d_rays, d_out, ps);
    run<1, 2, 16, 256, 4>("const 2 rays unroll16 256x4", nsm, d_sph, d_rays, d_out, ps);
    run<1, 2, 8, 256, 4>("const 2 rays unroll8  256x4", nsm, d_sph, d_rays, d_out, ps);
    run<1, 1, 32, 256, 8>("const 1 ray  unroll32 256x8", nsm, d_sph, d_rays, d_out, ps);
    run<2, 1, 32, 256, 4>("param 1 ray  unroll32 256x4", nsm, d_sph, d_rays, d_out, ps);
    run<2, 1, 8, 256, 4>("param 1 ray  unroll8  256x4", nsm, d_sph, d_rays, d_out, ps);
    run<2, 2, 16, 256, 4>("param 2 rays unroll16 256x4", nsm, d_sph, d_rays, d_out, ps);
    return 0;
}
