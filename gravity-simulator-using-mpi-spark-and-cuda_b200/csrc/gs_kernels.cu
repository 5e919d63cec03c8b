// gs_kernels.cu -- hand-written sm_100a kernels of the direct all-pairs gravity engine.
//
// Replaces the reference's calculate_forces_kernel + calculate_force_between (cuda.cu:32-60),
// its serial host update (cuda.cu:63-78) and mpi.c's force row + integrator (mpi.c:196-216).
//
// Force kernel (force_step_kernel):
//   * persistent CTAs, one static "stream-K" slice of the (i-tile x j-tile) unit space each
//     (gs_layout.h), so all 148 SMs finish together whatever N is;
//   * j-tiles (512 bodies, 8 KB) streamed into a 4-stage shared-memory ring by TMA bulk copies
//     (cp.async.bulk -> UBLKCP) that complete on mbarriers; one elected thread refills a stage two
//     tiles after the CTA's warps released it (empty mbarrier), nobody waits on __syncthreads;
//   * R i-bodies register-blocked per thread; every 128-bit broadcast LDS delivers one j-PAIR as
//     (even,odd) register pairs, consumed by packed FADD2/FFMA2/FMUL2 (two interactions per
//     instruction) and two MUFU.RSQ;
//   * two-level fp32 accumulation: per-tile packed partials flushed into per-body outer sums
//     (keeps the per-body relative error ~1e-6);
//   * the Euler / leapfrog velocity+position update is fused into the epilogue of every i-tile
//     that one CTA saw completely; the few split tiles leave partial sums in scratch and are
//     finished by fixup_kernel.
// There are no tensor-core instructions: the interaction is not a dense contraction.
#include "gs_kernels.h"
#include "gs_device.cuh"

namespace gs {

constexpr size_t kTileBytes = (size_t)kJTile * sizeof(float4);
constexpr size_t kSmemBytes = kStages * kTileBytes + 2 * kStages * sizeof(uint64_t);

// KIND: 0 scalar, 1 packed f32x2, 2 packed f32x2 with the uniform-mass hoist.
template <int R, int THREADS, int MINB, int CUT, int KIND, int UNROLL>
__global__ void __launch_bounds__(THREADS, MINB) force_step_kernel(const StepParams p) {
    static_assert(R % 2 == 0, "threads own whole i-pairs");
    constexpr int kThreads = THREADS;     // shadows gs::kThreads: this variant's CTA size
    constexpr int ITILE = kThreads * R;
    constexpr int NP = R / 2;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4* tiles = reinterpret_cast<float4*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + kStages * kTileBytes);
    uint64_t* empty = full + kStages;

    const int tid = threadIdx.x;
    const int64_t U = (int64_t)p.ni_tiles * p.nj_tiles;
    const int64_t u0 = cta_unit_begin(U, gridDim.x, blockIdx.x);
    const int64_t u1 = cta_unit_begin(U, gridDim.x, blockIdx.x + 1);
    const int nunits = (int)(u1 - u0);
    if (nunits <= 0) return;
    const int64_t first_tile = u0 / p.nj_tiles;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], kThreads / 32);
        }
        fence_mbar_init();
    }
    __syncthreads();

    // physical source of the k-th unit's j-tile
    auto tile_src = [&](int k) -> const float4* {
        int v = (int)((u0 + k) % p.nj_tiles);
        int phys = v + p.j_rot;
        if (phys >= p.nj_total) phys -= p.nj_total;
        return p.pos_old + (size_t)phys * kJTile;
    };
    if (tid == 0) {
        for (int k = 0; k < kStages && k < nunits; ++k) {
            mbar_arrive_expect_tx(&full[k], (uint32_t)kTileBytes);
            tma_bulk_g2s(tiles + (size_t)k * kJTile, tile_src(k), (uint32_t)kTileBytes, &full[k]);
        }
    }

    const float2 eps2 = make_float2(p.eps2, p.eps2);
    int k = 0;
    while (k < nunits) {
        const int64_t u = u0 + k;
        const int64_t it = u / p.nj_tiles;
        const int jv = (int)(u - it * p.nj_tiles);
        const int seg = min(p.nj_tiles - jv, nunits - k);
        const bool whole = (jv == 0) && (seg == p.nj_tiles);

        // this thread's R i-bodies (NP pairs), negated and duplicated for the packed subtract
        float2 nxi[R], nyi[R], nzi[R];
        Acc<R> acc;
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            const int pair_local = (int)it * (ITILE / 2) + q * kThreads + tid;
            const int pair_global = (p.i_begin >> 1) + pair_local;
            const float4 A = p.pos_old[2 * pair_global], B = p.pos_old[2 * pair_global + 1];
            nxi[2 * q] = make_float2(-A.x, -A.x); nxi[2 * q + 1] = make_float2(-A.y, -A.y);
            nyi[2 * q] = make_float2(-A.z, -A.z); nyi[2 * q + 1] = make_float2(-A.w, -A.w);
            nzi[2 * q] = make_float2(-B.x, -B.x); nzi[2 * q + 1] = make_float2(-B.y, -B.y);
        }
#pragma unroll
        for (int b = 0; b < R; ++b) acc.x[b] = acc.y[b] = acc.z[b] = 0.f;

        for (int t = 0; t < seg; ++t, ++k) {
            const int stage = k % kStages;
            if (tid == 0 && k >= kRefillLag) {   // refill the stage used kRefillLag tiles ago
                const int kr = k - kRefillLag, kn = kr + kStages;
                if (kn < nunits) {
                    const int sr = kr % kStages;
                    mbar_wait(&empty[sr], (uint32_t)((kr / kStages) & 1));
                    mbar_arrive_expect_tx(&full[sr], (uint32_t)kTileBytes);
                    tma_bulk_g2s(tiles + (size_t)sr * kJTile, tile_src(kn), (uint32_t)kTileBytes, &full[sr]);
                }
            }
            mbar_wait(&full[stage], (uint32_t)((k / kStages) & 1));
            const float4* T = tiles + (size_t)stage * kJTile;
            if (KIND == 0) tile_scalar<R, CUT>(T, nxi, nyi, nzi, eps2, p.cut2, acc);
            else tile_packed<R, CUT, KIND == 2, UNROLL>(T, nxi, nyi, nzi, eps2, p.cut2, acc);
            __syncwarp();
            if ((tid & 31) == 0) mbar_arrive(&empty[stage]);
        }

        // epilogue of this (i-tile, j-range) segment
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            const int in_tile = 2 * (q * kThreads + tid);            // first body of the pair, in the tile
            const float w = (KIND == 2) ? p.mu_uniform : 1.f;        // hoisted mass factor
            const float4 a0 = make_float4(w * acc.x[2 * q], w * acc.y[2 * q], w * acc.z[2 * q], 0.f);
            const float4 a1 = make_float4(w * acc.x[2 * q + 1], w * acc.y[2 * q + 1], w * acc.z[2 * q + 1], 0.f);
            if (whole) {
                finish_pair<false>(p, (int)it * (ITILE / 2) + q * kThreads + tid, a0, a1);
            } else {
                const int slot = (it == first_tile) ? 0 : 1;
                float4* dst = p.partial + ((size_t)blockIdx.x * 2 + slot) * ITILE + in_tile;
                dst[0] = a0; dst[1] = a1;
            }
        }
    }
}

// Finishes i-tiles after the force kernel:
//   split tiles   : sums the per-CTA partials in fixed CTA order, then finish_pair;
//   whole tiles   : untouched (already finished in the force kernel), unless `whole_acc` is given --
//                   then the force kernel only stored their accelerations there and this kernel finishes
//                   them too (multi-GPU: integration + peer stores happen here, for every body).
__global__ void __launch_bounds__(256) fixup_kernel(const StepParams p, int grid_main, int ITILE,
                                                    const float4* __restrict__ whole_acc, const FinishExtras ex) {
    const int pair_local = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair_local >= p.ni_tiles * (ITILE / 2)) return;
    const int64_t it = pair_local / (ITILE / 2);
    const int q = pair_local - (int)it * (ITILE / 2);
    const TileOwners o = tile_owners(p.ni_tiles, p.nj_tiles, grid_main, it);
    if (!o.split) {
        if (whole_acc == nullptr) return;   // finished inside the force kernel
        StepParams pw = p;
        pw.acc_in = nullptr;                // the force kernel already added the upstream accelerations
        finish_pair<true>(pw, pair_local, whole_acc[2 * pair_local], whole_acc[2 * pair_local + 1], &ex);
        return;
    }
    const int64_t U = (int64_t)p.ni_tiles * p.nj_tiles;
    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0;
    for (int c = o.first; c <= o.last; ++c) {   // fixed order: bit-reproducible
        if (cta_unit_begin(U, grid_main, c) >= cta_unit_begin(U, grid_main, c + 1)) continue;   // idle CTA
        const int slot = partial_slot(p.ni_tiles, p.nj_tiles, grid_main, c, it);
        const float4* src = p.partial + ((size_t)c * 2 + slot) * ITILE + 2 * q;
        const float4 s0 = src[0], s1 = src[1];
        a0.x += s0.x; a0.y += s0.y; a0.z += s0.z;
        a1.x += s1.x; a1.y += s1.y; a1.z += s1.z;
    }
    finish_pair<true>(p, pair_local, a0, a1, &ex);
}

// ------------------------------------------------------------------------------------------------
// dispatch
// ------------------------------------------------------------------------------------------------
//            id  R  threads minB packed unroll
#define GS_VARIANTS(X) \
    X(1, 4, 256, 2, 1, 2)  \
    X(2, 4, 256, 2, 0, 2)  \
    X(3, 8, 256, 1, 1, 2)  \
    X(4, 8, 256, 1, 0, 2)  \
    X(5, 2, 256, 2, 1, 2)  \
    X(6, 6, 256, 1, 1, 2)  \
    X(7, 4, 512, 1, 1, 2)  \
    X(8, 4, 256, 2, 1, 4)  \
    X(9, 4, 256, 2, 1, 1)  \
    X(10, 4, 128, 4, 1, 2) \
    X(11, 6, 128, 3, 1, 2) \
    X(12, 8, 128, 2, 1, 2) \
    X(13, 2, 512, 2, 1, 2) \
    X(14, 6, 256, 1, 1, 1) \
    X(15, 8, 128, 2, 1, 1) \
    X(16, 2, 128, 5, 1, 2) \
    X(17, 8, 384, 1, 1, 1)

bool variant_info(int variant, VariantInfo* out) {
    switch (variant) {
#define GS_X(ID, R_, T_, MB_, PK_, UN_) \
    case ID: if (out) { out->R = R_; out->threads = T_; out->unroll = UN_; out->packed = PK_; out->i_tile = R_ * T_; } return true;
        GS_VARIANTS(GS_X)
#undef GS_X
        default: return false;
    }
}

size_t force_smem_bytes() { return kSmemBytes; }

typedef void (*ForceKernel)(const StepParams);
template <int R, int T, int MB, int KIND, int UN>
static ForceKernel pick_cut(int cut) {
    switch (cut) {
        case kSoftened: return (ForceKernel)force_step_kernel<R, T, MB, kSoftened, KIND, UN>;
        case kCutoff: return (ForceKernel)force_step_kernel<R, T, MB, kCutoff, KIND, UN>;
        case kCutoffTracked: return (ForceKernel)force_step_kernel<R, T, MB, kCutoffTracked, KIND, UN>;
        default: return nullptr;
    }
}

static ForceKernel pick_force(int variant, int cut, bool uniform) {
    switch (variant) {
#define GS_X(ID, R, T, MB, PK, UN) \
    case ID: return (PK && uniform) ? pick_cut<R, T, MB, (PK ? 2 : 0), UN>(cut) : pick_cut<R, T, MB, PK, UN>(cut);
        GS_VARIANTS(GS_X)
#undef GS_X
        default: return nullptr;
    }
}

cudaError_t force_kernel_occupancy(int variant, int cut_mode, bool uniform, int* ctas_per_sm, int* regs) {
    ForceKernel k = pick_force(variant, cut_mode, uniform);
    VariantInfo vi;
    if (!k || !variant_info(variant, &vi)) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes);
    if (e != cudaSuccess) return e;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(ctas_per_sm, (const void*)k, vi.threads, kSmemBytes);
    if (e != cudaSuccess) return e;
    cudaFuncAttributes fa;
    e = cudaFuncGetAttributes(&fa, (const void*)k);
    if (e != cudaSuccess) return e;
    *regs = fa.numRegs;
    return cudaSuccess;
}

cudaError_t launch_force(int variant, int cut_mode, const StepParams& p, int grid, cudaStream_t s) {
    const bool uniform = p.mu_uniform != 0.f;
    ForceKernel k = pick_force(variant, cut_mode, uniform);
    VariantInfo vi;
    if (!k || !variant_info(variant, &vi)) return cudaErrorInvalidValue;
    k<<<grid, vi.threads, kSmemBytes, s>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_fixup(int variant, const StepParams& p, int grid_main, const float4* whole_acc,
                         const FinishExtras* extras, cudaStream_t s) {
    VariantInfo vi;
    if (!variant_info(variant, &vi)) return cudaErrorInvalidValue;
    const int pairs = p.ni_tiles * (vi.i_tile / 2);
    const int blocks = (pairs + 255) / 256;
    FinishExtras none;
    none.peers.n = 0; none.pos_lo = nullptr; none.vel_lo = nullptr; none.acc_lo = nullptr;
    fixup_kernel<<<blocks, 256, 0, s>>>(p, grid_main, vi.i_tile, whole_acc, extras ? *extras : none);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// fp64 force pass (accuracy mode) and the plain finishing kernel
// ------------------------------------------------------------------------------------------------
constexpr int kF64Threads = 128;

__global__ void __launch_bounds__(kF64Threads) force_f64_kernel(const float4* __restrict__ pos,
                                                                const float4* __restrict__ pos_lo_all,
                                                                const double* __restrict__ mass, double G, int n_pad,
                                                                int i_begin, double eps2, double cut2,
                                                                float4* __restrict__ acc_hi, float4* __restrict__ acc_lo) {
    __shared__ double tx[kJTile], ty[kJTile], tz[kJTile], tmu[kJTile];
    auto body = [&](int i, double& x, double& y, double& z, double& mu) {
        const float4 A = pos[2 * (i >> 1)], B = pos[2 * (i >> 1) + 1];
        const bool odd = (i & 1) != 0;
        x = odd ? A.y : A.x; y = odd ? A.w : A.z; z = odd ? B.y : B.x;
        mu = G * mass[i];                        // G*m in fp64 (mpi.c:66), not the fp32 mu of the packed record
        if (pos_lo_all != nullptr) {
            const float4 LA = pos_lo_all[2 * (i >> 1)], LB = pos_lo_all[2 * (i >> 1) + 1];
            x += odd ? LA.y : LA.x; y += odd ? LA.w : LA.z; z += odd ? LB.y : LB.x;
        }
    };
    const int il = blockIdx.x * kF64Threads + threadIdx.x;
    double xi, yi, zi, mui;
    body(i_begin + il, xi, yi, zi, mui);
    double ax = 0.0, ay = 0.0, az = 0.0;
    for (int j0 = 0; j0 < n_pad; j0 += kJTile) {
        __syncthreads();
        for (int t = threadIdx.x; t < kJTile; t += kF64Threads) body(j0 + t, tx[t], ty[t], tz[t], tmu[t]);
        __syncthreads();
#pragma unroll 4
        for (int t = 0; t < kJTile; ++t) {
            const double dx = tx[t] - xi, dy = ty[t] - yi, dz = tz[t] - zi;
            const double d2 = dx * dx + dy * dy + dz * dz;
            const double inv = rsqrt(d2 + eps2);
            // eps > 0: d = 0 contributes exactly 0; eps == 0: "r < 1e-10 contributes nothing" (mpi.c:64), incl. the body itself
            const double w = (d2 >= cut2 && tmu[t] != 0.0) ? tmu[t] * inv * inv * inv : 0.0;
            ax = fma(w, dx, ax); ay = fma(w, dy, ay); az = fma(w, dz, az);
        }
    }
    const float hx = (float)ax, hy = (float)ay, hz = (float)az;
    acc_hi[il] = make_float4(hx, hy, hz, 0.f);
    acc_lo[il] = make_float4((float)(ax - hx), (float)(ay - hy), (float)(az - hz), 0.f);
}

cudaError_t launch_force_f64(const float4* pos, const float4* pos_lo_all, const double* mass, double G, int n_pad,
                             int i_begin, int n_local_pad, double eps2, double cut2, float4* acc_hi, float4* acc_lo,
                             cudaStream_t s) {
    force_f64_kernel<<<n_local_pad / kF64Threads, kF64Threads, 0, s>>>(pos, pos_lo_all, mass, G, n_pad, i_begin, eps2, cut2,
                                                                      acc_hi, acc_lo);
    return cudaGetLastError();
}

__global__ void __launch_bounds__(256) finish_acc_kernel(const StepParams p, int pairs, const float4* __restrict__ acc,
                                                        const FinishExtras ex) {
    const int pair_local = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair_local >= pairs) return;
    finish_pair<true>(p, pair_local, acc[2 * pair_local], acc[2 * pair_local + 1], &ex);
}

cudaError_t launch_finish_acc(const StepParams& p, int n_local_pad, const float4* acc, const FinishExtras* extras,
                              cudaStream_t s) {
    const int pairs = n_local_pad / 2;
    FinishExtras none;
    none.peers.n = 0; none.pos_lo = nullptr; none.vel_lo = nullptr; none.acc_lo = nullptr;
    finish_acc_kernel<<<(pairs + 255) / 256, 256, 0, s>>>(p, pairs, acc, extras ? *extras : none);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// peer-memory exchange: flags
// ------------------------------------------------------------------------------------------------
__global__ void signal_peers_kernel(PeerFlags f, int world, int my_rank, uint32_t value) {
    const int q = threadIdx.x;
    if (q < world && q != my_rank) {
        __threadfence_system();                                  // earlier kernels' peer stores are ordered before the flag
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f.slot[q]), "r"(value) : "memory");
    }
}

__global__ void wait_flags_kernel(const uint32_t* my_flags, int world, int my_rank, uint32_t value,
                                  unsigned long long timeout_ns, uint32_t* err) {
    const int q = threadIdx.x;
    if (q >= world || q == my_rank) return;
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(my_flags + q) : "memory");
        if ((int32_t)(v - value) >= 0) break;                    // counters only grow (wrap-safe compare)
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t - t0 > timeout_ns) { *err = 1u; break; }           // never hang the GPU on a lost peer
        __nanosleep(200);
    }
}

// Both halves of a flag round in one launch: lane q raises my counter on peer q, then waits for peer q's counter here.
__global__ void signal_wait_kernel(PeerFlags f, const uint32_t* my_flags, int world, int my_rank, uint32_t value,
                                   unsigned long long timeout_ns, uint32_t* err) {
    const int q = threadIdx.x;
    if (q >= world || q == my_rank) return;
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f.slot[q]), "r"(value) : "memory");
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(my_flags + q) : "memory");
        if ((int32_t)(v - value) >= 0) break;
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t - t0 > timeout_ns) { *err = 1u; break; }
        __nanosleep(100);
    }
}
cudaError_t launch_signal_wait(const PeerFlags& f, const uint32_t* my_flags, int world, int my_rank, uint32_t value,
                               unsigned long long timeout_ns, uint32_t* err, cudaStream_t s) {
    signal_wait_kernel<<<1, 32, 0, s>>>(f, my_flags, world, my_rank, value, timeout_ns, err);
    return cudaGetLastError();
}

cudaError_t launch_signal_peers(const PeerFlags& f, int world, int my_rank, uint32_t value, cudaStream_t s) {
    signal_peers_kernel<<<1, 32, 0, s>>>(f, world, my_rank, value);
    return cudaGetLastError();
}
cudaError_t launch_wait_flags(const uint32_t* my_flags, int world, int my_rank, uint32_t value,
                              unsigned long long timeout_ns, uint32_t* err, cudaStream_t s) {
    wait_flags_kernel<<<1, 32, 0, s>>>(my_flags, world, my_rank, value, timeout_ns, err);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// AoS <-> packed pairs
// ------------------------------------------------------------------------------------------------
// Bodies [body_first, body_first + body_count) (even bounds): the record of body i is aos[(i - rec_first) * 7];
// bodies >= n are padding.  Packed positions go to every buffer in `dst` (full upload: this GPU's two buffers;
// slice upload: both buffers of every GPU, over NVLink), velocities / masses / low-order words stay local.
// mu_minmax[0] = min, mu_minmax[1] = max over the real bodies of the bit pattern of mu: equal <=> all masses equal.
template <typename T>
__global__ void aos_to_packed_kernel(const T* __restrict__ aos, int rec_first, int n, int body_first, int body_count,
                                     double G, UploadTargets dst, float4* vel_local, double* mass, int i_begin,
                                     int n_local_pad, float4* pos_lo_local, float4* vel_lo_local, uint32_t* mu_minmax) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t lo_bits = 0xffffffffu, hi_bits = 0u;
    if (k < body_count / 2) {
        const int pair = (body_first >> 1) + k;
        float x[2], y[2], z[2], vx[2], vy[2], vz[2], mu[2];
        float lo[2][6];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int i = 2 * pair + h;
            if (i < n) {
                const T* r = aos + (size_t)(i - rec_first) * 7;
                x[h] = (float)r[0]; y[h] = (float)r[1]; z[h] = (float)r[2];
                vx[h] = (float)r[3]; vy[h] = (float)r[4]; vz[h] = (float)r[5];
                const float hi[6] = {x[h], y[h], z[h], vx[h], vy[h], vz[h]};
#pragma unroll
                for (int c = 0; c < 6; ++c) lo[h][c] = (float)((double)r[c] - (double)hi[c]);   // 0 for f32 records
                const double m = (double)r[6];
                mu[h] = (float)(G * m);          // never form m_i*m_j: G*m <= 1.33e20 fits fp32
                mass[i] = m;
                const uint32_t bits = __float_as_uint(mu[h]);
                lo_bits = min(lo_bits, bits); hi_bits = max(hi_bits, bits);
            } else {                             // padding: massless, at rest, so far away that rinv^3 underflows to 0
                x[h] = y[h] = z[h] = kPadPosition;
                vx[h] = vy[h] = vz[h] = mu[h] = 0.f;
#pragma unroll
                for (int c = 0; c < 6; ++c) lo[h][c] = 0.f;
                mass[i] = 0.0;
            }
        }
        const float4 PA = make_float4(x[0], x[1], y[0], y[1]), PB = make_float4(z[0], z[1], mu[0], mu[1]);
        for (int q = 0; q < dst.n; ++q) { dst.pos[q][2 * pair] = PA; dst.pos[q][2 * pair + 1] = PB; }
        const int first = 2 * pair;
        if (first >= i_begin && first < i_begin + n_local_pad) {
            const int pl = (first - i_begin) >> 1;
            vel_local[2 * pl] = make_float4(vx[0], vx[1], vy[0], vy[1]);
            vel_local[2 * pl + 1] = make_float4(vz[0], vz[1], 0.f, 0.f);
            if (pos_lo_local != nullptr) {
                pos_lo_local[2 * pl] = make_float4(lo[0][0], lo[1][0], lo[0][1], lo[1][1]);
                pos_lo_local[2 * pl + 1] = make_float4(lo[0][2], lo[1][2], 0.f, 0.f);
                vel_lo_local[2 * pl] = make_float4(lo[0][3], lo[1][3], lo[0][4], lo[1][4]);
                vel_lo_local[2 * pl + 1] = make_float4(lo[0][5], lo[1][5], 0.f, 0.f);
            }
        }
    }
    lo_bits = __reduce_min_sync(0xffffffffu, lo_bits);
    hi_bits = __reduce_max_sync(0xffffffffu, hi_bits);
    if ((threadIdx.x & 31) == 0 && lo_bits <= hi_bits) { atomicMin(mu_minmax, lo_bits); atomicMax(mu_minmax + 1, hi_bits); }
}

// Bodies [body_first, body_first + count): pos is indexed globally, vel / lo words with (body - vel_first),
// records are written to aos[(body - body_first) * 7].
template <typename T>
__global__ void packed_to_aos_kernel(const float4* __restrict__ pos, const float4* __restrict__ vel, int vel_first,
                                     const double* __restrict__ mass, int body_first, int count, T* aos,
                                     const float4* __restrict__ pos_lo, const float4* __restrict__ vel_lo) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (2 * k >= count) return;
    const int pair = (body_first >> 1) + k, vp = pair - (vel_first >> 1);
    const float4 PA = pos[2 * pair], PB = pos[2 * pair + 1];
    const float4 VA = vel[2 * vp], VB = vel[2 * vp + 1];
    float4 LA = make_float4(0.f, 0.f, 0.f, 0.f), LB = LA, WA = LA, WB = LA;
    if (pos_lo != nullptr) { LA = pos_lo[2 * vp]; LB = pos_lo[2 * vp + 1]; WA = vel_lo[2 * vp]; WB = vel_lo[2 * vp + 1]; }
    {
        T* r = aos + (size_t)(2 * k) * 7;
        r[0] = (T)((double)PA.x + LA.x); r[1] = (T)((double)PA.z + LA.z); r[2] = (T)((double)PB.x + LB.x);
        r[3] = (T)((double)VA.x + WA.x); r[4] = (T)((double)VA.z + WA.z); r[5] = (T)((double)VB.x + WB.x);
        r[6] = (T)mass[2 * pair];
    }
    if (2 * k + 1 < count) {
        T* r = aos + (size_t)(2 * k + 1) * 7;
        r[0] = (T)((double)PA.y + LA.y); r[1] = (T)((double)PA.w + LA.w); r[2] = (T)((double)PB.y + LB.y);
        r[3] = (T)((double)VA.y + WA.y); r[4] = (T)((double)VA.w + WA.w); r[5] = (T)((double)VB.y + WB.y);
        r[6] = (T)mass[2 * pair + 1];
    }
}

__global__ void scale_positions_kernel(const float4* __restrict__ src, float4* __restrict__ dst, int npairs, float scale) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npairs) return;
    const float4 A = src[2 * p], B = src[2 * p + 1];
    dst[2 * p] = make_float4(A.x * scale, A.y * scale, A.z * scale, A.w * scale);
    dst[2 * p + 1] = make_float4(B.x * scale, B.y * scale, B.z, B.w);            // {z0, z1, mu0, mu1}
}
cudaError_t launch_scale_positions(const float4* src, float4* dst, int n_pad, float scale, cudaStream_t s) {
    scale_positions_kernel<<<(n_pad / 2 + 255) / 256, 256, 0, s>>>(src, dst, n_pad / 2, scale);
    return cudaGetLastError();
}

// {min, max} -> {min, ~max} behind them, so that ONE min-reduction over ranks gives both
__global__ void pack_minmax_kernel(uint32_t* mm) { mm[2] = mm[0]; mm[3] = ~mm[1]; }
cudaError_t launch_pack_minmax(uint32_t* mm4, cudaStream_t s) {
    pack_minmax_kernel<<<1, 1, 0, s>>>(mm4);
    return cudaGetLastError();
}

cudaError_t launch_aos_to_packed(const void* aos, int record_bytes, int rec_first, int n, int body_first,
                                 int body_count, double G, const UploadTargets& dst, float4* vel_local, double* mass,
                                 int i_begin, int n_local_pad, float4* pos_lo_local, float4* vel_lo_local,
                                 uint32_t* mu_minmax, cudaStream_t s) {
    const int blocks = (body_count / 2 + 255) / 256;
    if (blocks <= 0) return cudaErrorInvalidValue;
    if (record_bytes == 28)
        aos_to_packed_kernel<float><<<blocks, 256, 0, s>>>((const float*)aos, rec_first, n, body_first, body_count, G, dst,
                                                           vel_local, mass, i_begin, n_local_pad, pos_lo_local,
                                                           vel_lo_local, mu_minmax);
    else if (record_bytes == 56)
        aos_to_packed_kernel<double><<<blocks, 256, 0, s>>>((const double*)aos, rec_first, n, body_first, body_count, G,
                                                            dst, vel_local, mass, i_begin, n_local_pad, pos_lo_local,
                                                            vel_lo_local, mu_minmax);
    else
        return cudaErrorInvalidValue;
    return cudaGetLastError();
}

cudaError_t launch_packed_to_aos(const float4* pos, const float4* vel, int vel_first, const double* mass,
                                 int body_first, int count, void* aos, int record_bytes, const float4* pos_lo,
                                 const float4* vel_lo, cudaStream_t s) {
    const int blocks = ((count + 1) / 2 + 255) / 256;
    if (blocks <= 0) return cudaSuccess;
    if (record_bytes == 28)
        packed_to_aos_kernel<float><<<blocks, 256, 0, s>>>(pos, vel, vel_first, mass, body_first, count, (float*)aos, pos_lo, vel_lo);
    else if (record_bytes == 56)
        packed_to_aos_kernel<double><<<blocks, 256, 0, s>>>(pos, vel, vel_first, mass, body_first, count, (double*)aos, pos_lo, vel_lo);
    else
        return cudaErrorInvalidValue;
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// energy diagnostics (not in the reference)
// ------------------------------------------------------------------------------------------------
constexpr int kEThreads = 256;

__device__ __forceinline__ double block_sum(double v, double* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[w] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x == 0)
        for (int i = 0; i < kEThreads / 32; ++i) t += red[i];
    return t;   // valid in thread 0
}

// One i-PAIR per thread (256 bodies per CTA) against every j-tile, packed f32x2 like the force kernels: per j-pair
// and i-body 3 FADD2 + 3 FFMA2 + 2 MUFU.RSQ + 1 FFMA2; fp32 pair terms, one fp32 partial per j-tile, fp64 above.
// SOFT (eps > 0): no select at all -- the body's own term mu_i / eps is subtracted afterwards.
// !SOFT (eps == 0): pairs with r^2 < cut2 (the body itself, coincident bodies) contribute nothing (cuda.cu:39, mpi.c:64).
constexpr int kPotThreads = 128;
template <bool SOFT>
__global__ void __launch_bounds__(kPotThreads) potential_kernel(const float4* __restrict__ pos, int n_pad, int i_begin,
                                                                float eps2, float cut2, double* out_partials) {
    __shared__ __align__(16) float4 tile[kJTile];
    __shared__ double red[kPotThreads / 32];
    const int pair = ((i_begin + blockIdx.x * 2 * kPotThreads) >> 1) + threadIdx.x;
    const float4 IA = pos[2 * pair], IB = pos[2 * pair + 1];
    const float2 nx[2] = {make_float2(-IA.x, -IA.x), make_float2(-IA.y, -IA.y)};
    const float2 ny[2] = {make_float2(-IA.z, -IA.z), make_float2(-IA.w, -IA.w)};
    const float2 nz[2] = {make_float2(-IB.x, -IB.x), make_float2(-IB.y, -IB.y)};
    const float2 e2 = make_float2(eps2, eps2);
    double pot[2] = {0.0, 0.0};
    for (int j0 = 0; j0 < n_pad; j0 += kJTile) {
        __syncthreads();
        for (int t = threadIdx.x; t < kJTile; t += kPotThreads) tile[t] = pos[j0 + t];
        __syncthreads();
        float2 part[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
#pragma unroll 4
        for (int jp = 0; jp < kJTile / 2; ++jp) {
            const float4 A = tile[2 * jp], B = tile[2 * jp + 1];
            const float2 xj = make_float2(A.x, A.y), yj = make_float2(A.z, A.w);
            const float2 zj = make_float2(B.x, B.y), mj = make_float2(B.z, B.w);
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                const float2 dx = __fadd2_rn(xj, nx[b]), dy = __fadd2_rn(yj, ny[b]), dz = __fadd2_rn(zj, nz[b]);
                float2 r2 = __ffma2_rn(dx, dx, e2);
                r2 = __ffma2_rn(dy, dy, r2);
                r2 = __ffma2_rn(dz, dz, r2);
                float2 ri = make_float2(rsqrt_approx(r2.x), rsqrt_approx(r2.y));
                if (!SOFT) {
                    ri.x = (r2.x >= cut2) ? ri.x : 0.f;
                    ri.y = (r2.y >= cut2) ? ri.y : 0.f;
                }
                part[b] = __ffma2_rn(mj, ri, part[b]);
            }
        }
        pot[0] += (double)part[0].x + (double)part[0].y;
        pot[1] += (double)part[1].x + (double)part[1].y;
    }
    const double mu0 = IB.z, mu1 = IB.w;
    if (SOFT) {                  // the term of the body with itself: d = 0, so r2 was exactly eps2
        const double self = (double)rsqrt_approx(eps2);
        pot[0] -= mu0 * self;
        pot[1] -= mu1 * self;
    }
    double v = mu0 * pot[0] + mu1 * pot[1];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int i = 0; i < kPotThreads / 32; ++i) t += red[i];
        out_partials[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(kEThreads) kinetic_kernel(const float4* __restrict__ vel_local,
                                                            const double* __restrict__ mass, int i_begin,
                                                            double* out4) {
    __shared__ double red[kEThreads / 32];
    const int il = blockIdx.x * kEThreads + threadIdx.x;   // local body
    const float4 VA = vel_local[2 * (il >> 1)], VB = vel_local[2 * (il >> 1) + 1];
    const bool odd = (il & 1) != 0;
    const double vx = odd ? VA.y : VA.x, vy = odd ? VA.w : VA.z, vz = odd ? VB.y : VB.x;
    const double m = mass[i_begin + il];
    const double e = block_sum(m * (vx * vx + vy * vy + vz * vz), red);
    const double px = block_sum(m * vx, red);
    const double py = block_sum(m * vy, red);
    const double pz = block_sum(m * vz, red);
    if (threadIdx.x == 0) {
        out4[4 * blockIdx.x + 0] = e; out4[4 * blockIdx.x + 1] = px;
        out4[4 * blockIdx.x + 2] = py; out4[4 * blockIdx.x + 3] = pz;
    }
}

// Device-side end of the energy pass: sums the per-CTA partials in a fixed order (bit-reproducible) and writes one
// sample {step, time, KE, PE, px, py, pz, 0} -- this rank's share; nothing is reduced on the host.
__global__ void __launch_bounds__(kEThreads) energy_sample_kernel(const double* __restrict__ pot_partials,
                                                                  const double* __restrict__ kin4, int egrid, double G,
                                                                  double step, double time, double* __restrict__ out8) {
    __shared__ double red[kEThreads / 32];
    double v[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    for (int b = threadIdx.x; b < egrid; b += kEThreads) {
        v[0] += pot_partials[b];
#pragma unroll
        for (int q = 0; q < 4; ++q) v[1 + q] += kin4[4 * b + q];
    }
    double t[5];
#pragma unroll
    for (int q = 0; q < 5; ++q) t[q] = block_sum(v[q], red);
    if (threadIdx.x == 0) {
        out8[0] = step; out8[1] = time;
        out8[2] = 0.5 * t[1];                  // KE = 1/2 sum m v^2
        out8[3] = -0.5 * t[0] / G;             // U = -(1/(2G)) sum_i mu_i sum_{j != i} mu_j / d
        out8[4] = t[2]; out8[5] = t[3]; out8[6] = t[4]; out8[7] = 0.0;
    }
}
cudaError_t launch_energy_sample(const double* pot_partials, const double* kin4, int egrid, double G, double step,
                                 double time, double* out8, cudaStream_t s) {
    energy_sample_kernel<<<1, kEThreads, 0, s>>>(pot_partials, kin4, egrid, G, step, time, out8);
    return cudaGetLastError();
}

int energy_grid(int n_local_pad) { return n_local_pad / kEThreads; }

cudaError_t launch_potential(const float4* pos, int n_pad, int i_begin, int n_local_pad, float eps2, float cut2,
                             double* out_partials, cudaStream_t s) {
    static_assert(2 * kPotThreads == kEThreads, "one partial per 256 bodies (energy_grid, gs_noalloc.cu)");
    if (eps2 > 0.f) potential_kernel<true><<<energy_grid(n_local_pad), kPotThreads, 0, s>>>(pos, n_pad, i_begin, eps2, cut2, out_partials);
    else potential_kernel<false><<<energy_grid(n_local_pad), kPotThreads, 0, s>>>(pos, n_pad, i_begin, eps2, cut2, out_partials);
    return cudaGetLastError();
}
cudaError_t launch_kinetic(const float4* vel_local, const double* mass, int i_begin, int n_local_pad, double* out4,
                           cudaStream_t s) {
    kinetic_kernel<<<energy_grid(n_local_pad), kEThreads, 0, s>>>(vel_local, mass, i_begin, out4);
    return cudaGetLastError();
}

}  // namespace gs
