// mix_probe.cu — which ingredient of the sweep body costs issue slots? (diagnostic)
// Runs the 12-op discriminant body on register operands with optional LDS.128 / SHF, and pure FADD / FMUL streams.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

// MODE bit0: LDS.128 per body, bit1: SHF per body.  MODE 8: pure FADD, 9: pure FMUL, 10: FADD/FMUL/FFMA round robin
template <int MODE>
__global__ void __launch_bounds__(256, 4) k_mix(const float* in, unsigned* out, int iters) {
    __shared__ float4 sm[512];
    for (int i = threadIdx.x; i < 512; i += 256) sm[i] = make_float4(in[i & 63], in[(i + 1) & 63], in[(i + 2) & 63], in[(i + 3) & 63]);
    __syncthreads();
    float ox = in[threadIdx.x & 63], oy = in[(threadIdx.x + 7) & 63], oz = in[(threadIdx.x + 9) & 63];
    float dx = in[(threadIdx.x + 11) & 63], dy = in[(threadIdx.x + 13) & 63], dz = in[(threadIdx.x + 17) & 63], a = in[(threadIdx.x + 19) & 63];
    unsigned mask = 0;
    float facc = 0.f;
    if (MODE >= 8) {
        float acc[8];
#pragma unroll
        for (int i = 0; i < 8; i++) acc[i] = ox + i;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int u = 0; u < 48; u++) {
                int i = u & 7;
                if (MODE == 8) acc[i] = __fadd_rn(acc[i], dx);
                else if (MODE == 9) acc[i] = __fmul_rn(acc[i], dx);
                else {
                    if (u % 3 == 0) acc[i] = __fadd_rn(acc[i], dx);
                    else if (u % 3 == 1) acc[i] = __fmul_rn(acc[i], dy);
                    else acc[i] = __fmaf_rn(acc[i], dy, dz);
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 8; i++) facc += acc[i];
        out[blockIdx.x * 256 + threadIdx.x] = __float_as_uint(facc);
        return;
    }
    float4 s = sm[threadIdx.x & 1];
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int j = 0; j < 32; j++) {
            if (MODE & 1) s = sm[(it & 15) * 32 + j];
            else { s.x = __int_as_float(__float_as_int(s.x) ^ 0); }
            float ocx = __fsub_rn(ox, s.x), ocy = __fsub_rn(oy, s.y), ocz = __fsub_rn(oz, s.z);
            float hb = __fmaf_rn(ocz, dz, __fmaf_rn(ocy, dy, __fmul_rn(ocx, dx)));
            float cc = __fsub_rn(__fmaf_rn(ocz, ocz, __fmaf_rn(ocy, ocy, __fmul_rn(ocx, ocx))), s.w);
            float disc = __fmaf_rn(hb, hb, -__fmul_rn(a, cc));
            if (MODE & 2) mask = __funnelshift_l(__float_as_uint(disc), mask, 1);
            else if (MODE & 4) facc = fmaxf(facc, disc);
            else { ox = disc; }   // keep the body alive with a dependency and no extra instruction
        }
    }
    out[blockIdx.x * 256 + threadIdx.x] = mask + __float_as_uint(facc) + __float_as_uint(ox);
}

template <int MODE>
void run(const char* name, int nsm, const float* d_in, unsigned* d_out, double fp_instr_per_body, double bodies_per_iter) {
    int iters = 4000;
    int grid = nsm * 4;
    k_mix<MODE><<<grid, 256>>>(d_in, d_out, 10);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int r = 0; r < 3; r++) {
        CK(cudaEventRecord(e0));
        k_mix<MODE><<<grid, 256>>>(d_in, d_out, iters);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    // FP32-pipe warp-instructions per SMSP per cycle at 1965 MHz
    double warps_per_smsp = 8.0;
    double instr = (double)iters * bodies_per_iter * fp_instr_per_body * warps_per_smsp;
    double cycles = best * 1e-3 * 1.965e9;
    printf("%-40s ms=%8.3f  FP-pipe instr/clk/SMSP=%.3f\n", name, best, instr / cycles);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int nsm = p.multiProcessorCount;
    float h[64];
    for (int i = 0; i < 64; i++) h[i] = 0.5f + 0.01f * i;
    float* d_in; unsigned* d_out;
    CK(cudaMalloc(&d_in, sizeof(h))); CK(cudaMemcpy(d_in, h, sizeof(h), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_out, nsm * 4 * 256 * 4));
    run<0>("12-op body, registers only", nsm, d_in, d_out, 12, 32);
    run<1>("12-op body + LDS.128", nsm, d_in, d_out, 12, 32);
    run<2>("12-op body + SHF", nsm, d_in, d_out, 12, 32);
    run<3>("12-op body + LDS.128 + SHF", nsm, d_in, d_out, 12, 32);
    run<4>("12-op body + FMNMX", nsm, d_in, d_out, 12, 32);
    run<5>("12-op body + LDS.128 + FMNMX", nsm, d_in, d_out, 12, 32);
    run<8>("pure FADD", nsm, d_in, d_out, 1, 48);
    run<9>("pure FMUL", nsm, d_in, d_out, 1, 48);
    run<10>("FADD/FMUL/FFMA round robin", nsm, d_in, d_out, 1, 48);
    return 0;
}
