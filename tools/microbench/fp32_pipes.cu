// Microbenchmark: FP32-pipe, packed f32x2 and MUFU.RSQ issue rates on sm_100a.
// Answers the design question in DESIGN.md ("scalar FFMA vs FFMA2 inner loop").
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o fp32_pipes fp32_pipes.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
    fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int ITERS = 4096;

template <int ILP>
__global__ void __launch_bounds__(256) k_ffma(float* out, float b, float c) {
    float a[ILP];
#pragma unroll
    for (int k = 0; k < ILP; ++k) a[k] = threadIdx.x * 1e-3f + k;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int k = 0; k < ILP; ++k) a[k] = fmaf(a[k], b, c);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < ILP; ++k) s += a[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// 3 distinct, varying register operands per FFMA (a = a*x + y with x,y per-k registers)
template <int ILP>
__global__ void __launch_bounds__(256) k_ffma3(float* out, float b, float c) {
    float a[ILP], x[ILP], y[ILP];
#pragma unroll
    for (int k = 0; k < ILP; ++k) { a[k] = threadIdx.x * 1e-3f + k; x[k] = b + k * 1e-7f; y[k] = c + k; }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int k = 0; k < ILP; ++k) a[k] = fmaf(a[k], x[k], y[(k + 1) % ILP]);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < ILP; ++k) s += a[k] + x[k] + y[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void __launch_bounds__(256) k_ffma2(float* out, float b, float c) {
    float2 a[ILP];
    float2 bb = make_float2(b, b * 1.0001f), cc = make_float2(c, c * 0.999f);
#pragma unroll
    for (int k = 0; k < ILP; ++k) a[k] = make_float2(threadIdx.x * 1e-3f + k, k * 0.5f);
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int k = 0; k < ILP; ++k) a[k] = __ffma2_rn(a[k], bb, cc);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < ILP; ++k) s += a[k].x + a[k].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void __launch_bounds__(256) k_fadd2(float* out, float b, float c) {
    float2 a[ILP];
    float2 bb = make_float2(b, b * 1.0001f);
#pragma unroll
    for (int k = 0; k < ILP; ++k) a[k] = make_float2(threadIdx.x * 1e-3f + k, k * 0.5f);
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int k = 0; k < ILP; ++k) a[k] = __fadd2_rn(a[k], bb);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < ILP; ++k) s += a[k].x + a[k].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int ILP>
__global__ void __launch_bounds__(256) k_rsq(float* out, float b, float c) {
    float a[ILP];
#pragma unroll
    for (int k = 0; k < ILP; ++k) a[k] = threadIdx.x * 1e-3f + k + 1.0f;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int k = 0; k < ILP; ++k) a[k] = rsqrtf(a[k]);   // MUFU.RSQ (+ denormal fix-up unless ftz)
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < ILP; ++k) s += a[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// MUFU.RSQ + NF dependent FFMAs per rsq, to see whether MUFU issue co-schedules with the FMA pipe.
template <int ILP, int NF>
__global__ void __launch_bounds__(256) k_mix(float* out, float b, float c) {
    float a[ILP];
#pragma unroll
    for (int k = 0; k < ILP; ++k) a[k] = threadIdx.x * 1e-3f + k + 1.0f;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int k = 0; k < ILP; ++k) {
            float r;
            asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a[k]));
#pragma unroll
            for (int f = 0; f < NF; ++f) r = fmaf(r, b, c);
            a[k] = r;
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < ILP; ++k) s += a[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// FSETP+FSEL (select) per FFMA: is compare/select on a different pipe from FFMA?
template <int ILP>
__global__ void __launch_bounds__(256) k_ffma_sel(float* out, float b, float c) {
    float a[ILP];
#pragma unroll
    for (int k = 0; k < ILP; ++k) a[k] = threadIdx.x * 1e-3f + k;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int k = 0; k < ILP; ++k) { float t = fmaf(a[k], b, c); a[k] = (t >= c) ? t : 0.0f; }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < ILP; ++k) s += a[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}


// K independent FFMA2 + NM independent MUFU.RSQ per iteration: does a MUFU issue cost FMA-pipe cycles?
template <int K, int NM>
__global__ void __launch_bounds__(256) k_ffma2_mufu(float* out, float b, float c) {
    float2 a[K];
    float m[NM > 0 ? NM : 1];
    float2 bb = make_float2(b, b * 1.0001f), cc = make_float2(c, c * 0.999f);
#pragma unroll
    for (int k = 0; k < K; ++k) a[k] = make_float2(threadIdx.x * 1e-3f + k, k * 0.5f);
#pragma unroll
    for (int k = 0; k < NM; ++k) m[k] = threadIdx.x * 1e-3f + k + 1.0f;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int k = 0; k < K; ++k) {
            a[k] = __ffma2_rn(a[k], bb, cc);
            if (k < NM) asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(m[k]) : "f"(m[k]));
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) s += a[k].x + a[k].y;
#pragma unroll
    for (int k = 0; k < NM; ++k) s += m[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// The pair-symmetric inner step itself, registers only (no shared memory, no shuffles, no loop-carried addressing):
// per (i-body, j-pair) 3 FADD2 + 3 FFMA2 + 2 MUFU.RSQ + 2 FMUL2 + 6 FFMA2 with the kernel's dependency chains
// (DESIGN.md 4.1b).  R i-bodies x NP j-pairs independent chains per iteration.  Lane-ops counted for the 14 packed
// instructions only: 128/clk/SM = FMA pipe full.  This is the ceiling of the instruction stream, before any data movement.
template <int R, int NP>
__global__ void __launch_bounds__(256) k_symstep(float* out, float b, float c) {
    float2 nx[R], ny[R], nz[R], ax[R], ay[R], az[R];
    float2 xj[NP], yj[NP], zj[NP], jx[NP], jy[NP], jz[NP];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        nx[r] = make_float2(-(threadIdx.x * 1e-3f + r), -(threadIdx.x * 1e-3f + r));
        ny[r] = make_float2(-(r * 0.37f + b), -(r * 0.37f + b));
        nz[r] = make_float2(-(r * 0.11f + c), -(r * 0.11f + c));
        ax[r] = ay[r] = az[r] = make_float2(0.f, 0.f);
    }
#pragma unroll
    for (int h = 0; h < NP; ++h) {
        xj[h] = make_float2(100.f + h + threadIdx.x, 101.f + h); yj[h] = make_float2(50.f + h, 51.f - h);
        zj[h] = make_float2(25.f + h, 26.f + h); jx[h] = jy[h] = jz[h] = make_float2(0.f, 0.f);
    }
    const float2 eps2 = make_float2(c, c), drift = make_float2(b * 1e-3f, b * 1e-3f);
    for (int it = 0; it < ITERS / 8; ++it) {
#pragma unroll
        for (int h = 0; h < NP; ++h) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float2 dx = __fadd2_rn(xj[h], nx[r]), dy = __fadd2_rn(yj[h], ny[r]), dz = __fadd2_rn(zj[h], nz[r]);
                float2 r2 = __ffma2_rn(dx, dx, eps2);
                r2 = __ffma2_rn(dy, dy, r2);
                r2 = __ffma2_rn(dz, dz, r2);
                float2 ri;
                asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(ri.x) : "f"(r2.x));
                asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(ri.y) : "f"(r2.y));
                const float2 t = __fmul2_rn(__fmul2_rn(ri, ri), ri);
                ax[r] = __ffma2_rn(t, dx, ax[r]); ay[r] = __ffma2_rn(t, dy, ay[r]); az[r] = __ffma2_rn(t, dz, az[r]);
                jx[h] = __ffma2_rn(t, dx, jx[h]); jy[h] = __ffma2_rn(t, dy, jy[h]); jz[h] = __ffma2_rn(t, dz, jz[h]);
            }
            xj[h] = __fadd2_rn(xj[h], drift);      // "the next pair": keeps the loop body from being hoisted (1 extra FADD2 per R)
        }
    }
    float s = 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) s += ax[r].x + ax[r].y + ay[r].x + ay[r].y + az[r].x + az[r].y;
#pragma unroll
    for (int h = 0; h < NP; ++h) s += jx[h].x + jx[h].y + jy[h].x + jy[h].y + jz[h].x + jz[h].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename K>
static void run(const char* name, K kern, int blocks_per_sm, double ops_per_thread_iter, int sms, float* d_out) {
    int grid = sms * blocks_per_sm;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int w = 0; w < 3; ++w) kern<<<grid, 256>>>(d_out, 0.9999f, 1e-3f);
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        CK(cudaEventRecord(e0));
        kern<<<grid, 256>>>(d_out, 0.9999f, 1e-3f);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    double ops = (double)grid * 256 * ITERS * ops_per_thread_iter;  // lane-ops
    double per_s = ops / (best * 1e-3);
    // lane-ops per clock per SM at the nominal 1965 MHz
    printf("%-28s blocks/SM=%d  %.3f ms  %.3e lane-ops/s  %.1f lane-ops/clk/SM@1965MHz\n",
           name, blocks_per_sm, best, per_s, per_s / sms / 1.965e9);
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int sms = p.multiProcessorCount;
    printf("device %s, %d SMs, clockRate %d kHz\n", p.name, sms, p.clockRate);
    float* d_out; CK(cudaMalloc(&d_out, (size_t)sms * 8 * 256 * sizeof(float)));
    // the symmetric inner step: 8 warps per SM as in the shipped kernel (its registers allow one 256-thread block)
    for (int bps : {1}) {
        run("symstep R=8 NP=4 (packed lanes)", k_symstep<8, 4>, bps, 8 * 4 * 28 / 8.0, sms, d_out);
        run("symstep R=8 NP=2 (packed lanes)", k_symstep<8, 2>, bps, 8 * 2 * 28 / 8.0, sms, d_out);
        run("symstep R=4 NP=4 (packed lanes)", k_symstep<4, 4>, bps, 4 * 4 * 28 / 8.0, sms, d_out);
    }
    for (int bps : {2, 4}) {
        run("ffma  (reg,const,const) ILP8", k_ffma<8>, bps, 8, sms, d_out);
        run("ffma3 (3 distinct regs) ILP8", k_ffma3<8>, bps, 8, sms, d_out);
        run("ffma2 packed ILP8 (x2 lanes)", k_ffma2<8>, bps, 16, sms, d_out);
        run("ffma2 packed ILP4 (x2 lanes)", k_ffma2<4>, bps, 8, sms, d_out);
        run("fadd2 packed ILP8 (x2 lanes)", k_fadd2<8>, bps, 16, sms, d_out);
        run("rsqrtf ILP8", k_rsq<8>, bps, 8, sms, d_out);
        run("mix rsq + 6 ffma ILP8", k_mix<8, 6>, bps, 8 * 7, sms, d_out);
        run("mix rsq + 12 ffma ILP8", k_mix<8, 12>, bps, 8 * 13, sms, d_out);
        run("ffma + setp/sel ILP8 (ffma only)", k_ffma_sel<8>, bps, 8, sms, d_out);
        // lane-ops counted for the FFMA2 only (2 per lane): 128 = FMA pipe full
        run("12 ffma2 + 0 mufu (ffma2 lanes)", k_ffma2_mufu<12, 0>, bps, 24, sms, d_out);
        run("12 ffma2 + 1 mufu (ffma2 lanes)", k_ffma2_mufu<12, 1>, bps, 24, sms, d_out);
        run("12 ffma2 + 2 mufu (ffma2 lanes)", k_ffma2_mufu<12, 2>, bps, 24, sms, d_out);
        run("12 ffma2 + 3 mufu (ffma2 lanes)", k_ffma2_mufu<12, 3>, bps, 24, sms, d_out);
        run("12 ffma2 + 4 mufu (ffma2 lanes)", k_ffma2_mufu<12, 4>, bps, 24, sms, d_out);
        run("6 ffma2 + 1 mufu (ffma2 lanes)", k_ffma2_mufu<6, 1>, bps, 12, sms, d_out);
    }
    CK(cudaFree(d_out));
    return 0;
}
