
// rf_probe.cu — register-file operand bandwidth: FFMA/FADD throughput vs distinct register operands per instruction.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

// MODE 0: acc = fma(acc, m, b)        (m, b shared by all: reuse-cache friendly)   1 fresh read
// MODE 1: acc = fma(acc, x[i], b)     2 fresh reads
// MODE 2: acc = fma(x[i], y[i], acc)  3 fresh reads
// MODE 3: acc = acc + x[i]            2 fresh reads (FADD)
// MODE 4: acc = acc + b               1 fresh read (FADD)
// MODE 5: acc = fma(x[i], x[i], acc)  2 distinct (square)
template <int MODE>
__global__ void __launch_bounds__(256, 4) k_rf(const float* in, float* out, int iters) {
    float acc[8], x[8], y[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = in[(threadIdx.x + i) & 63]; x[i] = in[(threadIdx.x + 8 + i) & 63]; y[i] = in[(threadIdx.x + 16 + i) & 63]; }
    float m = in[3], b = in[4];
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 8; u++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (MODE == 0) acc[i] = __fmaf_rn(acc[i], m, b);
                else if (MODE == 1) acc[i] = __fmaf_rn(acc[i], x[i], b);
                else if (MODE == 2) acc[i] = __fmaf_rn(x[i], y[i], acc[i]);
                else if (MODE == 3) acc[i] = __fadd_rn(acc[i], x[i]);
                else if (MODE == 4) acc[i] = __fadd_rn(acc[i], b);
                else acc[i] = __fmaf_rn(x[i], x[i], acc[i]);
            }
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += acc[i];
    out[blockIdx.x * 256 + threadIdx.x] = s;
}

template <int MODE>
void run(const char* name, int nsm, const float* d_in, float* d_out) {
    int iters = 20000;
    int grid = nsm * 4;
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
    double instr = (double)iters * 64 * 8.0;   // per SMSP
    double cycles = best * 1e-3 * 1.965e9;
    printf("%-40s ms=%8.3f  IPC/SMSP=%.3f\n", name, best, instr / cycles);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int nsm = p.multiProcessorCount;
    float h[64];
    for (int i = 0; i < 64; i++) h[i] = 0.5f + 0.01f * i;
    float* d_in; float* d_out;
    CK(cudaMalloc(&d_in, sizeof(h))); CK(cudaMemcpy(d_in, h, sizeof(h), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_out, nsm * 4 * 256 * 4));
    run<0>("FFMA acc*m+b (1 fresh read)", nsm, d_in, d_out);
    run<1>("FFMA acc*x[i]+b (2 fresh reads)", nsm, d_in, d_out);
    run<2>("FFMA x[i]*y[i]+acc (3 fresh reads)", nsm, d_in, d_out);
    run<5>("FFMA x[i]*x[i]+acc (2 distinct)", nsm, d_in, d_out);
    run<3>("FADD acc+x[i] (2 fresh reads)", nsm, d_in, d_out);
    run<4>("FADD acc+b (1 fresh read)", nsm, d_in, d_out);
    return 0;
}
