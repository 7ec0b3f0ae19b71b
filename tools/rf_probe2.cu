
// rf_probe2.cu — packed FFMA2/FADD2/FMUL2 throughput vs fresh 64-bit operand reads.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

template <int MODE>
__global__ void __launch_bounds__(256, 4) k_rf(const float* in, float* out, int iters) {
    float2 acc[8], x[8], y[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        acc[i] = make_float2(in[(threadIdx.x + i) & 63], in[(threadIdx.x + i + 1) & 63]);
        x[i] = make_float2(in[(threadIdx.x + 8 + i) & 63], in[(threadIdx.x + 9 + i) & 63]);
        y[i] = make_float2(in[(threadIdx.x + 16 + i) & 63], in[(threadIdx.x + 17 + i) & 63]);
    }
    float2 m = make_float2(in[3], in[5]), b = make_float2(in[4], in[6]);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (MODE == 0) acc[i] = __ffma2_rn(acc[i], m, b);
                else if (MODE == 1) acc[i] = __ffma2_rn(acc[i], x[i], b);
                else if (MODE == 2) acc[i] = __ffma2_rn(x[i], y[i], acc[i]);
                else if (MODE == 3) acc[i] = __fadd2_rn(acc[i], x[i]);
                else if (MODE == 4) acc[i] = __fmul2_rn(acc[i], x[i]);
                else if (MODE == 5) {   // alternate packed 3-read FFMA2 with two scalar 2-read FADDs
                    acc[i] = __ffma2_rn(x[i], y[i], acc[i]);
                    x[i].x = __fadd_rn(x[i].x, m.x);
                    y[i].y = __fadd_rn(y[i].y, m.y);
                }
            }
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += acc[i].x + acc[i].y + x[i].x + y[i].y;
    out[blockIdx.x * 256 + threadIdx.x] = s;
}

template <int MODE>
void run(const char* name, int nsm, const float* d_in, float* d_out, double fp_lane_ops_per_inner) {
    int iters = 10000;
    cudaFuncAttributes fa; CK(cudaFuncGetAttributes(&fa, k_rf<MODE>));
    int occ = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_rf<MODE>, 256, 0));
    int grid = nsm * occ;
    k_rf<MODE><<<grid, 256>>>(d_in, d_out, 10);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int r = 0; r < 3; r++) {
        CK(cudaEventRecord(e0));
        k_rf<MODE><<<grid, 256>>>(d_in, d_out, iters);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    double inner = (double)iters * 64 * (occ * 2.0);   // inner bodies per SMSP
    double cycles = best * 1e-3 * 1.965e9;
    printf("%-44s regs=%3d occ=%d ms=%8.3f  cycles/inner=%.3f  scalar-op-equivalents/cycle=%.3f\n", name, fa.numRegs, occ, best, cycles / inner,
           fp_lane_ops_per_inner * inner / cycles);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int nsm = p.multiProcessorCount;
    float h[64];
    for (int i = 0; i < 64; i++) h[i] = 0.5f + 0.01f * i;
    float* d_in; float* d_out;
    CK(cudaMalloc(&d_in, sizeof(h))); CK(cudaMemcpy(d_in, h, sizeof(h), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_out, nsm * 8 * 256 * 4));
    run<0>("FFMA2 acc*m+b (1 fresh pair)", nsm, d_in, d_out, 2);
    run<1>("FFMA2 acc*x[i]+b (2 fresh pairs)", nsm, d_in, d_out, 2);
    run<2>("FFMA2 x[i]*y[i]+acc (3 fresh pairs)", nsm, d_in, d_out, 2);
    run<3>("FADD2 acc+x[i] (2 fresh pairs)", nsm, d_in, d_out, 2);
    run<4>("FMUL2 acc*x[i] (2 fresh pairs)", nsm, d_in, d_out, 2);
    run<5>("FFMA2(3 fresh) + 2 scalar FADD", nsm, d_in, d_out, 4);
    return 0;
}
