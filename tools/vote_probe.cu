// vote_probe.cu — design-space probes for stage 1 of the two-stage sweep (diagnostic, not product code).
//   MODE 0: the product's LDS-fed form: rays per lane (RAYS = 2), sphere triples {c_a, c_b, R'} broadcast from shared memory
//           (three LDS.128 per four spheres), 3 FFMA + SHF per pair.
//   MODE 1: transposed: a lane holds K = 8 spheres {c_a, c_b, rho} of a 256-sphere chunk in registers (coalesced LDG.128 from
//           global), the warp's 64 rays' constants {ua, ub, nou} are broadcast one ray at a time from shared memory (one
//           LDS.128 per ray and chunk), per (ray, sphere slot): 2 FFMA + FSETP(|p| <= rho) + VOTE.ballot; the ray's owner lane
//           stores the 8 ballots (= the mask words of the chunk's 8 blocks) with two STS.128.
// Question: does FSETP + VOTE issue at the rate of FFMA + SHF (is VOTE a half-rate pipe shared with the ALU?), and what is the
// cost per 32 pairs of either form at NS = 512 and NS = 10240 spheres?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -o vote_probe vote_probe.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

// MODE 0 ---------------------------------------------------------------------------------------------------------------
template <int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_lds(const float4* __restrict__ g_s1, int n_spheres, const float* __restrict__ rays, int reps,
                                                       unsigned* __restrict__ out) {
    extern __shared__ float4 sm[];
    for (int i = threadIdx.x; i < n_spheres * 3 / 4; i += THREADS) sm[i] = g_s1[i];
    __syncthreads();
    constexpr int RAYS = 2;
    float ua[RAYS], ub[RAYS], nou[RAYS];
    long long tid = (long long)blockIdx.x * THREADS + threadIdx.x;
#pragma unroll
    for (int r = 0; r < RAYS; r++) {
        const float* p = rays + (tid * RAYS + r) * 6;
        float il = rsqrtf(p[3] * p[3] + p[5] * p[5]);
        ua[r] = p[5] * il; ub[r] = -p[3] * il;
        nou[r] = -(p[0] * ua[r] + p[2] * ub[r]);
    }
    unsigned acc = 0;
    for (int rep = 0; rep < reps; rep++) {
#pragma unroll 1
        for (int blk = 0; blk < n_spheres / 32; blk++) {
            unsigned mask[RAYS];
#pragma unroll
            for (int r = 0; r < RAYS; r++) mask[r] = 0;
            const float4* q = sm + blk * 24;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const float4 A = q[3 * j], B = q[3 * j + 1], C = q[3 * j + 2];
#define SLAB(ca, cb, rp)                                                                                              \
    _Pragma("unroll") for (int r = 0; r < RAYS; r++) {                                                                \
        float p = __fmaf_rn(ca, ua[r], __fmaf_rn(cb, ub[r], nou[r]));                                                  \
        mask[r] = __funnelshift_l(__float_as_uint(__fmaf_rn(-p, p, rp)), mask[r], 1);                                  \
    }
                SLAB(A.x, A.y, A.z) SLAB(A.w, B.x, B.y) SLAB(B.z, B.w, C.x) SLAB(C.y, C.z, C.w)
            }
#pragma unroll
            for (int r = 0; r < RAYS; r++) acc += __popc(~mask[r]);
        }
#pragma unroll
        for (int r = 0; r < RAYS; r++) nou[r] = __fadd_rn(nou[r], 0.001f);
    }
    out[tid] = acc;
}

// MODE 2 ---------------------------------------------------------------------------------------------------------------
// LDS-fed, but ONE RAY AT A TIME over each 32-sphere block: the block's triples are read twice (volatile LDS.128, so the
// two passes are not merged), and within a pass the FFMAs of consecutive spheres share the ray's constants in the same
// operand slots (operand-reuse cache): T_j = cb_j*ub + nou (1 fresh register), P_j = ca_j*ua + T_j (2 fresh), Q_j (2 fresh).
__device__ __forceinline__ float4 lds_v4_volatile(const float4* p) {
    float4 v;
    const unsigned a = (unsigned)__cvta_generic_to_shared(p);
    asm volatile("ld.volatile.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
template <int THREADS, int MINB, int GROUP>
__global__ void __launch_bounds__(THREADS, MINB) k_lds_seq(const float4* __restrict__ g_s1, int n_spheres, const float* __restrict__ rays, int reps,
                                                           unsigned* __restrict__ out) {
    extern __shared__ float4 sm[];
    for (int i = threadIdx.x; i < n_spheres * 3 / 4; i += THREADS) sm[i] = g_s1[i];
    __syncthreads();
    constexpr int RAYS = 2;
    float ua[RAYS], ub[RAYS], nou[RAYS];
    long long tid = (long long)blockIdx.x * THREADS + threadIdx.x;
#pragma unroll
    for (int r = 0; r < RAYS; r++) {
        const float* p = rays + (tid * RAYS + r) * 6;
        float il = rsqrtf(p[3] * p[3] + p[5] * p[5]);
        ua[r] = p[5] * il; ub[r] = -p[3] * il;
        nou[r] = -(p[0] * ua[r] + p[2] * ub[r]);
    }
    unsigned acc = 0;
    for (int rep = 0; rep < reps; rep++) {
#pragma unroll 1
        for (int blk = 0; blk < n_spheres / 32; blk++) {
            unsigned mask[RAYS];
            const float4* q = sm + blk * 24;
#pragma unroll
            for (int r = 0; r < RAYS; r++) {
                unsigned m = 0;
#pragma unroll
                for (int j0 = 0; j0 < 8; j0 += GROUP) {
                    float ca[4 * GROUP], cb[4 * GROUP], rp[4 * GROUP], t[4 * GROUP];
#pragma unroll
                    for (int j = 0; j < GROUP; j++) {
                        const float4 A = lds_v4_volatile(q + 3 * (j0 + j)), B = lds_v4_volatile(q + 3 * (j0 + j) + 1), C = lds_v4_volatile(q + 3 * (j0 + j) + 2);
                        ca[4 * j] = A.x; cb[4 * j] = A.y; rp[4 * j] = A.z;
                        ca[4 * j + 1] = A.w; cb[4 * j + 1] = B.x; rp[4 * j + 1] = B.y;
                        ca[4 * j + 2] = B.z; cb[4 * j + 2] = B.w; rp[4 * j + 2] = C.x;
                        ca[4 * j + 3] = C.y; cb[4 * j + 3] = C.z; rp[4 * j + 3] = C.w;
                    }
#pragma unroll
                    for (int k = 0; k < 4 * GROUP; k++) t[k] = __fmaf_rn(cb[k], ub[r], nou[r]);
#pragma unroll
                    for (int k = 0; k < 4 * GROUP; k++) t[k] = __fmaf_rn(ca[k], ua[r], t[k]);
#pragma unroll
                    for (int k = 0; k < 4 * GROUP; k++) m = __funnelshift_l(__float_as_uint(__fmaf_rn(-t[k], t[k], rp[k])), m, 1);
                }
                mask[r] = m;
            }
#pragma unroll
            for (int r = 0; r < RAYS; r++) acc += __popc(~mask[r]);
        }
#pragma unroll
        for (int r = 0; r < RAYS; r++) nou[r] = __fadd_rn(nou[r], 0.001f);
    }
    out[tid] = acc;
}

// MODE 3 ---------------------------------------------------------------------------------------------------------------
// LDS-fed with PAIRED LANES: a lane sweeps 4 rays (its own two and its partner's, lane ^ 1) over HALF of the blocks (even
// lanes the even blocks, odd lanes the odd ones): one LDS.128 group serves four rays, so only one FFMA in four meets a
// sphere operand that is not in the operand-reuse cache.  Same pair-tests per lane as MODE 0.
template <int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB) k_lds_pair(const float4* __restrict__ g_s1, int n_spheres, const float* __restrict__ rays, int reps,
                                                            unsigned* __restrict__ out) {
    extern __shared__ float4 sm[];
    for (int i = threadIdx.x; i < n_spheres * 3 / 4; i += THREADS) sm[i] = g_s1[i];
    __syncthreads();
    constexpr int RAYS = 4;
    float ua[RAYS], ub[RAYS], nou[RAYS];
    long long tid = (long long)blockIdx.x * THREADS + threadIdx.x;
#pragma unroll
    for (int r = 0; r < 2; r++) {
        const float* p = rays + (tid * 2 + r) * 6;
        float il = rsqrtf(p[3] * p[3] + p[5] * p[5]);
        ua[r] = p[5] * il; ub[r] = -p[3] * il;
        nou[r] = -(p[0] * ua[r] + p[2] * ub[r]);
    }
#pragma unroll
    for (int r = 0; r < 2; r++) {
        ua[2 + r] = __shfl_xor_sync(0xffffffffu, ua[r], 1); ub[2 + r] = __shfl_xor_sync(0xffffffffu, ub[r], 1);
        nou[2 + r] = __shfl_xor_sync(0xffffffffu, nou[r], 1);
    }
    unsigned acc = 0;
    const int odd = threadIdx.x & 1;
    for (int rep = 0; rep < reps; rep++) {
#pragma unroll 1
        for (int blk = 0; blk < n_spheres / 32; blk += 2) {
            unsigned mask[RAYS];
#pragma unroll
            for (int r = 0; r < RAYS; r++) mask[r] = 0;
            const float4* q = sm + (blk + odd) * 24;
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const float4 A = q[3 * j], B = q[3 * j + 1], C = q[3 * j + 2];
                SLAB(A.x, A.y, A.z) SLAB(A.w, B.x, B.y) SLAB(B.z, B.w, C.x) SLAB(C.y, C.z, C.w)
            }
#pragma unroll
            for (int r = 0; r < RAYS; r++) acc += __popc(~mask[r]);
        }
#pragma unroll
        for (int r = 0; r < RAYS; r++) nou[r] = __fadd_rn(nou[r], 0.001f);
    }
    out[tid] = acc;
}

// MODE 1 ---------------------------------------------------------------------------------------------------------------
// g_t: per chunk of 256 spheres, per lane, 8 spheres x {c_a, c_b, rho} = 24 floats = 6 float4 (lane-major: coalesced LDG.128)
template <int THREADS, int MINB, int UNROLL>
__global__ void __launch_bounds__(THREADS, MINB) k_vote(const float4* __restrict__ g_t, int n_spheres, const float* __restrict__ rays, int reps,
                                                        unsigned* __restrict__ out) {
    extern __shared__ float4 sm[];
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    float4* ray_tab = sm + warp * 64;                                   // 64 rays x {ua, ub, nou, -}
    uint4* mask_tab = reinterpret_cast<uint4*>(sm + (THREADS / 32) * 64) + warp * 128;   // 64 rays x 8 words
    long long tid = (long long)blockIdx.x * THREADS + threadIdx.x;
    float nou_own[2];
#pragma unroll
    for (int r = 0; r < 2; r++) {
        const float* p = rays + (tid * 2 + r) * 6;
        float il = rsqrtf(p[3] * p[3] + p[5] * p[5]);
        float ua = p[5] * il, ub = -p[3] * il;
        nou_own[r] = -(p[0] * ua + p[2] * ub);
        ray_tab[r * 32 + lane] = make_float4(ua, ub, nou_own[r], 0.f);
    }
    __syncwarp();
    unsigned acc = 0;
    const int n_chunks = n_spheres / 256;
    for (int rep = 0; rep < reps; rep++) {
#pragma unroll 1
        for (int c = 0; c < n_chunks; c++) {
            float ca[8], cb[8], rho[8];
            {
                const float4* src = g_t + ((size_t)c * 32 + lane) * 6;
                float4 v[6];
#pragma unroll
                for (int i = 0; i < 6; i++) v[i] = __ldg(src + i);
                const float* f = reinterpret_cast<const float*>(v);
#pragma unroll
                for (int k = 0; k < 8; k++) { ca[k] = f[3 * k]; cb[k] = f[3 * k + 1]; rho[k] = f[3 * k + 2]; }
            }
#pragma unroll UNROLL
            for (int ray = 0; ray < 64; ray++) {
                const float4 rc = ray_tab[ray];                         // LDS.128 broadcast
                unsigned m[8];
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const float p = __fmaf_rn(ca[k], rc.x, __fmaf_rn(cb[k], rc.y, rc.z));
                    m[k] = __ballot_sync(0xffffffffu, fabsf(p) <= rho[k]);
                }
                if (lane == (unsigned)(ray & 31)) {
                    mask_tab[ray * 2] = make_uint4(m[0], m[1], m[2], m[3]);
                    mask_tab[ray * 2 + 1] = make_uint4(m[4], m[5], m[6], m[7]);
                }
            }
            __syncwarp();
            // owner side: every lane reads the masks of its two rays (what the refine walk would start from)
#pragma unroll
            for (int r = 0; r < 2; r++) {
                const uint4 a = mask_tab[(r * 32 + lane) * 2], b = mask_tab[(r * 32 + lane) * 2 + 1];
                acc += __popc(a.x) + __popc(a.y) + __popc(a.z) + __popc(a.w) + __popc(b.x) + __popc(b.y) + __popc(b.z) + __popc(b.w);
            }
            __syncwarp();
        }
#pragma unroll
        for (int r = 0; r < 2; r++) { nou_own[r] = __fadd_rn(nou_own[r], 0.001f); ray_tab[r * 32 + lane].z = nou_own[r]; }
        __syncwarp();
    }
    out[tid] = acc;
}

static double report(const char* name, int regs, int occ, int threads, float ms, double pairs_per_thread, int grid, unsigned acc0) {
    double tests = (double)grid * threads * pairs_per_thread;
    double cyc = 1.965e9 * (ms * 1e-3) / ((double)occ * threads / 32.0 / 4.0 * pairs_per_thread);
    printf("%-44s regs=%3d occ=%d ms=%8.3f tests/s=%.3e cycles per 32 pairs per SMSP @1965 = %5.2f  (acc0=%u)\n", name, regs, occ, ms, tests / (ms * 1e-3),
           cyc, acc0);
    return cyc;
}

template <class K>
void run(const char* name, K k, int threads, size_t smem, int nsm, const float4* d_s, int ns, const float* d_rays, unsigned* d_out, int reps) {
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, threads, smem));
    if (occ == 0) { printf("%-44s does not fit\n", name); return; }
    int grid = nsm * occ;
    cudaFuncAttributes fa; CK(cudaFuncGetAttributes(&fa, k));
    k<<<grid, threads, smem>>>(d_s, ns, d_rays, 1, d_out);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int it = 0; it < 3; it++) {
        CK(cudaEventRecord(e0));
        k<<<grid, threads, smem>>>(d_s, ns, d_rays, reps, d_out);
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    unsigned h0 = 0;
    CK(cudaMemcpy(&h0, d_out, 4, cudaMemcpyDeviceToHost));
    report(name, fa.numRegs, occ, threads, best, 2.0 * ns * reps, grid, h0);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int nsm = p.multiProcessorCount;
    srand(1);
    auto rf = []() { return (float)rand() / RAND_MAX; };
    const int NMAX = 10240;
    std::vector<float> s1(NMAX * 3), st(NMAX * 3);
    for (int i = 0; i < NMAX; i++) { s1[3 * i] = rf() * 100 - 50; s1[3 * i + 1] = rf() * 100 - 50; s1[3 * i + 2] = 0.04f; }
    for (int c = 0; c < NMAX / 256; c++)                     // transposed layout: lane j, slot k = sphere c*256 + k*32 + j
        for (int j = 0; j < 32; j++)
            for (int k = 0; k < 8; k++) {
                const int s = c * 256 + k * 32 + j;
                float* d = &st[((size_t)(c * 32 + j) * 8 + k) * 3];
                d[0] = s1[3 * s]; d[1] = s1[3 * s + 1]; d[2] = 0.2f;
            }
    size_t nthreads = (size_t)nsm * 2048;
    std::vector<float> rays(nthreads * 2 * 6);
    for (size_t i = 0; i < rays.size() / 6; i++) {
        rays[6 * i] = rf() * 20 - 10; rays[6 * i + 1] = rf() * 3; rays[6 * i + 2] = rf() * 20 - 10;
        rays[6 * i + 3] = rf() * 2 - 1; rays[6 * i + 4] = rf() * 2 - 1; rays[6 * i + 5] = rf() * 2 - 1;
    }
    float4 *d_s1, *d_st; float* d_rays; unsigned* d_out;
    CK(cudaMalloc(&d_s1, s1.size() * 4)); CK(cudaMemcpy(d_s1, s1.data(), s1.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_st, st.size() * 4)); CK(cudaMemcpy(d_st, st.data(), st.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_rays, rays.size() * 4)); CK(cudaMemcpy(d_rays, rays.data(), rays.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&d_out, nthreads * 4));
    printf("device %s SMs=%d\n", p.name, nsm);
    auto vsm = [](int threads) { return (size_t)(threads / 32) * (64 * 16 + 128 * 16); };
    for (int ns : {512, 10240}) {
        const int reps = ns == 512 ? 400 : 20;
        printf("-- %d spheres\n", ns);
        run("LDS-fed 2 rays/lane 640x1", k_lds<640, 1>, 640, (size_t)ns * 12, nsm, d_s1, ns, d_rays, d_out, reps);
        if (ns == 512) run("LDS-fed 2 rays/lane 256x4", k_lds<256, 4>, 256, (size_t)ns * 12, nsm, d_s1, ns, d_rays, d_out, reps);
        run("LDS-fed paired lanes (4 rays, half blocks) 640x1", k_lds_pair<640, 1>, 640, (size_t)ns * 12, nsm, d_s1, ns, d_rays, d_out, reps);
        run("LDS-fed paired lanes (4 rays, half blocks) 576x1", k_lds_pair<576, 1>, 576, (size_t)ns * 12, nsm, d_s1, ns, d_rays, d_out, reps);
        run("LDS-fed ray-sequential group 1 640x1", k_lds_seq<640, 1, 1>, 640, (size_t)ns * 12, nsm, d_s1, ns, d_rays, d_out, reps);
        run("LDS-fed ray-sequential group 2 640x1", k_lds_seq<640, 1, 2>, 640, (size_t)ns * 12, nsm, d_s1, ns, d_rays, d_out, reps);
        run("LDS-fed ray-sequential group 4 640x1", k_lds_seq<640, 1, 4>, 640, (size_t)ns * 12, nsm, d_s1, ns, d_rays, d_out, reps);
        run("transposed vote 640x1 unroll 2", k_vote<640, 1, 2>, 640, vsm(640), nsm, d_st, ns, d_rays, d_out, reps);
        run("transposed vote 640x1 unroll 4", k_vote<640, 1, 4>, 640, vsm(640), nsm, d_st, ns, d_rays, d_out, reps);
        run("transposed vote 256x4 unroll 2", k_vote<256, 4, 2>, 256, vsm(256), nsm, d_st, ns, d_rays, d_out, reps);
        run("transposed vote 256x4 unroll 4", k_vote<256, 4, 4>, 256, vsm(256), nsm, d_st, ns, d_rays, d_out, reps);
        run("transposed vote 512x2 unroll 4", k_vote<512, 2, 4>, 512, vsm(512), nsm, d_st, ns, d_rays, d_out, reps);
    }
    return 0;
}
