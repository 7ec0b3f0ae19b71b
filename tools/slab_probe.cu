// slab_probe.cu — design-space probe for stage 1 of the two-stage sweep (diagnostic, not product code).
//   MODE 0: FP32 slab test, 3 FFMA + SHF per pair, sphere triples {c_a, c_b, R'} in parameter space (the product's form)
//   MODE 1: packed half2 slab test, 3 HFMA2 + SHF + LOP3 per TWO pairs, triples {c_a, c_b, R'} as half2 of two spheres
// Question: does HFMA2 with a uniform-register operand issue at one per cycle, i.e. is MODE 1 worth its looser margin?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -o slab_probe slab_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

constexpr int NS = 512;
struct ParamScene {
    float2 f[NS * 3 / 2];        // MODE 0: three floats per sphere
    uint32_t h[NS * 3 / 2];      // MODE 1: per sphere pair (k, k+16 of a block): half2 c_a, half2 c_b, half2 R'
};

template <int MODE, int RAYS, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_slab(const float* __restrict__ rays, int reps, unsigned* __restrict__ out,
                                                        const __grid_constant__ ParamScene ps) {
    float ua[RAYS], ub[RAYS], nou[RAYS];
    __half2 hua[RAYS], hub[RAYS], hnou[RAYS];
    long long tid = (long long)blockIdx.x * THREADS + threadIdx.x;
#pragma unroll
    for (int r = 0; r < RAYS; r++) {
        const float* p = rays + (tid * RAYS + r) * 6;
        float il = rsqrtf(p[3] * p[3] + p[5] * p[5]);
        ua[r] = p[5] * il; ub[r] = -p[3] * il;
        nou[r] = -(p[0] * ua[r] + p[2] * ub[r]);
        hua[r] = __float2half2_rn(ua[r]); hub[r] = __float2half2_rn(ub[r]); hnou[r] = __float2half2_rn(nou[r]);
    }
    unsigned acc = 0;
    for (int rep = 0; rep < reps; rep++) {
#pragma unroll 1
        for (int blk = 0; blk < NS / 32; blk++) {
            unsigned mask[RAYS];
#pragma unroll
            for (int r = 0; r < RAYS; r++) mask[r] = 0;
            if (MODE == 0) {
#pragma unroll
                for (int j = 0; j < 16; j++) {
                    const float2 A = ps.f[3 * (blk * 16 + j)], B = ps.f[3 * (blk * 16 + j) + 1], C = ps.f[3 * (blk * 16 + j) + 2];
#pragma unroll
                    for (int r = 0; r < RAYS; r++) {
                        float p = __fmaf_rn(A.x, ua[r], __fmaf_rn(A.y, ub[r], nou[r]));
                        mask[r] = __funnelshift_l(__float_as_uint(__fmaf_rn(-p, p, B.x)), mask[r], 1);
                    }
#pragma unroll
                    for (int r = 0; r < RAYS; r++) {
                        float p = __fmaf_rn(B.y, ua[r], __fmaf_rn(C.x, ub[r], nou[r]));
                        mask[r] = __funnelshift_l(__float_as_uint(__fmaf_rn(-p, p, C.y)), mask[r], 1);
                    }
                }
            } else {
#pragma unroll
                for (int j = 0; j < 16; j++) {
                    const uint32_t a = ps.h[3 * (blk * 16 + j)], b = ps.h[3 * (blk * 16 + j) + 1], c = ps.h[3 * (blk * 16 + j) + 2];
                    const __half2 ca = *reinterpret_cast<const __half2*>(&a), cb = *reinterpret_cast<const __half2*>(&b),
                                  rp = *reinterpret_cast<const __half2*>(&c);
#pragma unroll
                    for (int r = 0; r < RAYS; r++) {
                        __half2 p = __hfma2(ca, hua[r], __hfma2(cb, hub[r], hnou[r]));
                        __half2 q = __hfma2(__hneg2(p), p, rp);
                        uint32_t qb = *reinterpret_cast<uint32_t*>(&q);
                        mask[r] = ((mask[r] >> 1) & 0x7fff7fffu) | (qb & 0x80008000u);
                    }
                }
            }
#pragma unroll
            for (int r = 0; r < RAYS; r++) acc += __popc(~mask[r]);
        }
#pragma unroll
        for (int r = 0; r < RAYS; r++) { nou[r] = __fadd_rn(nou[r], 0.001f); hnou[r] = __float2half2_rn(nou[r]); }
    }
    out[tid] = acc;
}

template <int MODE, int RAYS, int THREADS, int MINB>
void run(const char* name, int nsm, const float* d_rays, unsigned* d_out, const ParamScene& ps) {
    int reps = 200;
    auto k = k_slab<MODE, RAYS, THREADS, MINB>;
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, THREADS, 0));
    int grid = nsm * occ;
    cudaFuncAttributes fa; CK(cudaFuncGetAttributes(&fa, k));
    k<<<grid, THREADS>>>(d_rays, 2, d_out, ps);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int it = 0; it < 3; it++) {
        CK(cudaEventRecord(e0));
        k<<<grid, THREADS>>>(d_rays, reps, d_out, ps);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    std::vector<unsigned> h(16);
    CK(cudaMemcpy(h.data(), d_out, 64, cudaMemcpyDeviceToHost));
    double tests = (double)grid * THREADS * RAYS * (double)NS * reps;
    double tps = tests / (best * 1e-3);
    double cyc = 1.965e9 * (best * 1e-3) / ((double)occ * THREADS / 32.0 / 4.0 * RAYS * (double)NS * reps);
    printf("%-34s regs=%3d occ=%d ms=%7.3f tests/s=%.3e cyc/warp-test/SMSP@1965=%5.2f  (acc0=%u)\n", name, fa.numRegs, occ, best, tps, cyc, h[0]);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int nsm = p.multiProcessorCount;
    static ParamScene ps;
    srand(1);
    auto rf = []() { return (float)rand() / RAND_MAX; };
    float* f = reinterpret_cast<float*>(ps.f);
    for (int i = 0; i < NS; i++) { f[3 * i] = rf() * 22 - 11; f[3 * i + 1] = rf() * 22 - 11; f[3 * i + 2] = 0.04f; }
    for (int blk = 0; blk < NS / 32; blk++)
        for (int j = 0; j < 16; j++) {
            int lo = blk * 32 + j, hi = lo + 16;
            __half2 a = __floats2half2_rn(f[3 * lo], f[3 * hi]), b = __floats2half2_rn(f[3 * lo + 1], f[3 * hi + 1]),
                    c = __floats2half2_rn(f[3 * lo + 2], f[3 * hi + 2]);
            ps.h[3 * (blk * 16 + j)] = *reinterpret_cast<uint32_t*>(&a);
            ps.h[3 * (blk * 16 + j) + 1] = *reinterpret_cast<uint32_t*>(&b);
            ps.h[3 * (blk * 16 + j) + 2] = *reinterpret_cast<uint32_t*>(&c);
        }
    size_t nthreads = (size_t)nsm * 2048;
    std::vector<float> rays(nthreads * 4 * 6);
    for (size_t i = 0; i < rays.size() / 6; i++) {
        rays[6 * i] = rf() * 20 - 10; rays[6 * i + 1] = rf() * 3; rays[6 * i + 2] = rf() * 20 - 10;
        rays[6 * i + 3] = rf() * 2 - 1; rays[6 * i + 4] = rf() * 2 - 1; rays[6 * i + 5] = rf() * 2 - 1;
    }
    float* d_rays; unsigned* d_out;
    CK(cudaMalloc(&d_rays, rays.size() * 4)); CK(cudaMemcpy(d_rays, rays.data(), rays.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_out, nthreads * 4));
    printf("device %s SMs=%d\n", p.name, nsm);
    run<0, 2, 256, 4>("fp32  2 rays 256x4", nsm, d_rays, d_out, ps);
    run<0, 4, 256, 4>("fp32  4 rays 256x4", nsm, d_rays, d_out, ps);
    run<1, 2, 256, 4>("half2 2 rays 256x4", nsm, d_rays, d_out, ps);
    run<1, 4, 256, 4>("half2 4 rays 256x4", nsm, d_rays, d_out, ps);
    run<1, 2, 256, 8>("half2 2 rays 256x8", nsm, d_rays, d_out, ps);
    return 0;
}
