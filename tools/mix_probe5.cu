
// mix_probe5.cu — is the sweep bound by shared-memory -> register return bandwidth? Same instruction count, different LDS width.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

// WIDTH: 16 = LDS.128, 8 = LDS.64, 4 = LDS.32 per test (missing components come from registers); 0 = LDC (constant bank) 128-bit
__constant__ float4 c_sph[512];

template <int WIDTH, int RAYS>
__global__ void __launch_bounds__(256, 4) k_mix(const float* in, unsigned* out, int iters) {
    __shared__ float4 sm[512];
    for (int i = threadIdx.x; i < 512; i += 256) sm[i] = make_float4(in[i & 63], in[(i + 1) & 63], in[(i + 2) & 63], in[(i + 3) & 63]);
    __syncthreads();
    float ox[RAYS], oy[RAYS], oz[RAYS], dx[RAYS], dy[RAYS], dz[RAYS], a[RAYS];
    unsigned mask[RAYS];
#pragma unroll
    for (int r = 0; r < RAYS; r++) {
        ox[r] = in[(threadIdx.x + r) & 63]; oy[r] = in[(threadIdx.x + 7 + r) & 63]; oz[r] = in[(threadIdx.x + 9 + r) & 63];
        dx[r] = in[(threadIdx.x + 11 + r) & 63]; dy[r] = in[(threadIdx.x + 13 + r) & 63]; dz[r] = in[(threadIdx.x + 17 + r) & 63];
        a[r] = in[(threadIdx.x + 19 + r) & 63]; mask[r] = 0;
    }
    float rz = in[5], rw = in[6], ry = in[7];
    for (int it = 0; it < iters; it++) {
        const float4* base = sm + (it & 15) * 32;
#pragma unroll
        for (int j = 0; j < 32; j++) {
            float4 s;
            if (WIDTH == 16) s = base[j];
            else if (WIDTH == 8) { float2 t = reinterpret_cast<const float2*>(base + j)[0]; s = make_float4(t.x, t.y, rz, rw); }
            else if (WIDTH == 4) { float t = reinterpret_cast<const float*>(base + j)[0]; s = make_float4(t, ry, rz, rw); }
            else s = c_sph[(it & 15) * 32 + j];
#pragma unroll
            for (int r = 0; r < RAYS; r++) {
                float ocx = __fsub_rn(ox[r], s.x), ocy = __fsub_rn(oy[r], s.y), ocz = __fsub_rn(oz[r], s.z);
                float hb = __fmaf_rn(ocz, dz[r], __fmaf_rn(ocy, dy[r], __fmul_rn(ocx, dx[r])));
                float cc = __fsub_rn(__fmaf_rn(ocz, ocz, __fmaf_rn(ocy, ocy, __fmul_rn(ocx, ocx))), s.w);
                float disc = __fmaf_rn(hb, hb, -__fmul_rn(a[r], cc));
                mask[r] = __funnelshift_l(__float_as_uint(disc), mask[r], 1);
            }
        }
#pragma unroll
        for (int r = 0; r < RAYS; r++) ox[r] = __fadd_rn(ox[r], 0.001f);
    }
    unsigned m = 0;
#pragma unroll
    for (int r = 0; r < RAYS; r++) m += mask[r];
    out[blockIdx.x * 256 + threadIdx.x] = m;
}

template <int WIDTH, int RAYS>
void run(const char* name, int nsm, const float* d_in, unsigned* d_out) {
    int iters = 4000 / RAYS;
    cudaFuncAttributes fa; CK(cudaFuncGetAttributes(&fa, k_mix<WIDTH, RAYS>));
    int occ = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_mix<WIDTH, RAYS>, 256, 0));
    int grid = nsm * occ;
    k_mix<WIDTH, RAYS><<<grid, 256>>>(d_in, d_out, 10);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int r = 0; r < 3; r++) {
        CK(cudaEventRecord(e0));
        k_mix<WIDTH, RAYS><<<grid, 256>>>(d_in, d_out, iters);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    double tests = (double)iters * 32 * RAYS * (occ * 2.0);   // warp-tests per SMSP (occ blocks * 8 warps / 4 SMSP)
    double cycles = best * 1e-3 * 1.965e9;
    printf("%-34s regs=%3d occ=%d ms=%8.3f  cycles/warp-test=%6.3f  frac_of_FP32_peak=%.3f\n", name, fa.numRegs, occ, best, cycles / tests,
           8.5 * tests / cycles);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int nsm = p.multiProcessorCount;
    float h[64];
    for (int i = 0; i < 64; i++) h[i] = 0.5f + 0.01f * i;
    float* d_in; unsigned* d_out;
    CK(cudaMalloc(&d_in, sizeof(h))); CK(cudaMemcpy(d_in, h, sizeof(h), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_out, nsm * 8 * 256 * 4));
    float4 hc[512]; for (int i = 0; i < 512; i++) hc[i] = make_float4(h[i & 63], h[(i + 1) & 63], h[(i + 2) & 63], h[(i + 3) & 63]);
    CK(cudaMemcpyToSymbol(c_sph, hc, sizeof(hc)));
    run<16, 1>("1 ray  LDS.128", nsm, d_in, d_out);
    run<8, 1>("1 ray  LDS.64", nsm, d_in, d_out);
    run<4, 1>("1 ray  LDS.32", nsm, d_in, d_out);
    run<0, 1>("1 ray  constant bank", nsm, d_in, d_out);
    run<16, 2>("2 rays LDS.128", nsm, d_in, d_out);
    run<8, 2>("2 rays LDS.64", nsm, d_in, d_out);
    run<0, 2>("2 rays constant bank", nsm, d_in, d_out);
    run<16, 3>("3 rays LDS.128", nsm, d_in, d_out);
    run<16, 4>("4 rays LDS.128", nsm, d_in, d_out);
    run<0, 4>("4 rays constant bank", nsm, d_in, d_out);
    return 0;
}
