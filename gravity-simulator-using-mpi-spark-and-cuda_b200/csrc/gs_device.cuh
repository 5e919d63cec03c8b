// gs_device.cuh -- device-side building blocks shared by the force kernels (gs_kernels.cu, gs_symmetric.cu):
// mbarrier / TMA PTX helpers, the per-pair finishing step (integration, peer stores), and the j-tile inner loops.
#pragma once
#include "gs_kernels.h"

namespace gs {

// ------------------------------------------------------------------------------------------------
// PTX helpers: mbarrier + TMA bulk copy
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "GS_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra GS_DONE;\n"
        "bra GS_WAIT;\n"
        "GS_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// 1-D TMA bulk copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void tma_bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ float rsqrt_approx(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));   // MUFU.RSQ, nothing else
    return r;
}

// ------------------------------------------------------------------------------------------------
// Order-independent sums of floats: wide fixed point.
// The j-side sums of the pair-symmetric pass are added up by many CTAs (and GPUs) in an order that depends on
// timing.  Floating-point atomics would make the last bits of every run different; integer addition is
// associative, so each float is added EXACTLY into a 160-bit fixed-point number (kFxLimbs 64-bit words holding
// base-2^32 digits, least significant bit 2^kFxLsb) with two integer atomics.  The result does not depend on the
// order of the additions: the default pass is bit-reproducible.
// Window: magnitudes in [2^-80, 2^71) (8e-25 .. 2.4e21 in whatever units the accelerations have); smaller
// contributions are truncated towards -inf bit by bit, larger ones saturate -- both deterministically.
// ------------------------------------------------------------------------------------------------
constexpr int kFxLimbs = 5;
constexpr int kFxLsb = -80;
constexpr int kFxWordsPerBody = 3 * kFxLimbs;     // x, y, z

// limbs[k * stride] is limb k
template <bool SYS>
__device__ __forceinline__ void fx_add(long long* limbs, int stride, float v) {
    const int bits = __float_as_int(v);
    const int e = (bits >> 23) & 0xff;
    if (e == 0) return;                                     // zero (denormals count as zero)
    long long m = (long long)((bits & 0x7fffff) | 0x800000);
    int s = e - 150 - kFxLsb;                               // v = +-m * 2^(e - 150) = +-m * 2^s * 2^kFxLsb
    if (e == 0xff || s > 32 * (kFxLimbs - 2) + 31) { s = 32 * (kFxLimbs - 2) + 31; m = 0xffffff; }   // inf/NaN/huge: saturate
    if (bits < 0) m = -m;
    if (s < 0) {
        if (s < -24) return;
        m >>= -s;                                           // arithmetic shift: floor
        s = 0;
    }
    const int k = s >> 5;
    const long long X = m << (s & 31);                      // |X| < 2^55
    const unsigned long long lo = (unsigned long long)(X & 0xffffffffll);
    const long long hi = X >> 32;                           // X = hi * 2^32 + lo exactly
    unsigned long long* w = reinterpret_cast<unsigned long long*>(limbs) + (size_t)k * stride;
    if (SYS) {
        if (lo) atomicAdd_system(w, lo);
        if (hi) atomicAdd_system(w + stride, (unsigned long long)hi);
    } else {
        if (lo) atomicAdd(w, lo);
        if (hi) atomicAdd(w + stride, (unsigned long long)hi);
    }
}
// The accumulated value, rounded once; reads the limbs in a fixed order (deterministic).
__device__ __forceinline__ float fx_value(const long long* limbs) {
    double v = 0.0;
#pragma unroll
    for (int k = kFxLimbs - 1; k >= 0; --k) v += ldexp((double)limbs[k], 32 * k + kFxLsb);
    return (float)v;
}

// ------------------------------------------------------------------------------------------------
// finishing one i-pair: add upstream accelerations, then write them out or integrate
//   v += a * kick_dt ; x += v_new * drift_dt          (cuda.cu:73-76, mpi.c:207-215)
// ------------------------------------------------------------------------------------------------
// PEERS: also store the new positions into every peer GPU's buffer (the exchange, fused with the
// integration).  Only the finishing kernel is built with PEERS = true: compiled into the force kernel the
// extra loop perturbs ptxas' schedule of the hot inner loop (measured -2.5 %), so on several GPUs the
// force kernel hands its accelerations to finish_kernel instead (16 B per body, negligible).
// Error-free pieces of the compensated update (GS_FLAG_COMPENSATED): the state is kept as hi + lo fp32 pairs,
// so x += v*dt no longer rounds to the 16 km ulp of a 3e11 m coordinate every step.
__device__ __forceinline__ void two_sum(float a, float b, float& s, float& e) {
    s = __fadd_rn(a, b);
    const float bb = __fadd_rn(s, -a);
    e = __fadd_rn(__fadd_rn(a, -__fadd_rn(s, -bb)), __fadd_rn(b, -bb));
}
// (hi, lo) += inc_hi + inc_lo
__device__ __forceinline__ void comp_add(float& hi, float& lo, float inc_hi, float inc_lo) {
    float s, e;
    two_sum(hi, inc_hi, s, e);
    e = __fadd_rn(e, __fadd_rn(lo, inc_lo));
    hi = __fadd_rn(s, e);
    lo = __fadd_rn(e, -__fadd_rn(hi, -s));
}
// v += (a + alo)*kick ; x += v*drift, both compensated (alo: low-order word of the acceleration, 0 unless the forces
// were summed in fp64)
__device__ __forceinline__ void comp_step(float a, float alo, float kick, float drift, bool do_drift, float& v, float& vlo,
                                          float& x, float& xlo) {
    const float dv = __fmul_rn(a, kick);
    comp_add(v, vlo, dv, __fmaf_rn(alo, kick, __fmaf_rn(a, kick, -dv)));
    if (!do_drift) return;
    const float dx = __fmul_rn(v, drift);
    comp_add(x, xlo, dx, __fadd_rn(__fmaf_rn(v, drift, -dx), __fmul_rn(vlo, drift)));
}

template <bool PEERS>
__device__ __forceinline__ void finish_pair(const StepParams& p, int pair_local, float4 a0, float4 a1,
                                            const FinishExtras* ex = nullptr) {
    if (p.acc_in != nullptr) {
        const float4 b0 = p.acc_in[2 * pair_local], b1 = p.acc_in[2 * pair_local + 1];
        a0.x += b0.x; a0.y += b0.y; a0.z += b0.z;
        a1.x += b1.x; a1.y += b1.y; a1.z += b1.z;
    }
    const int body0 = p.i_begin + 2 * pair_local;   // padding bodies stay put (massless, at rest)
    if (body0 >= p.n) a0 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (body0 + 1 >= p.n) a1 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p.final_mode == kFinalAccOut) {
        a0.w = 0.f; a1.w = 0.f;
        p.acc_out[2 * pair_local] = a0;
        p.acc_out[2 * pair_local + 1] = a1;
        return;
    }
    const int pair_global = (p.i_begin >> 1) + pair_local;
    float4 VA = p.vel[2 * pair_local], VB = p.vel[2 * pair_local + 1];
    float4 PA = p.pos_old[2 * pair_global], PB = p.pos_old[2 * pair_global + 1];
    if (PEERS && ex->pos_lo != nullptr) {   // compensated integration (finish kernel only)
        float4 LA = ex->pos_lo[2 * pair_local], LB = ex->pos_lo[2 * pair_local + 1];
        float4 WA = ex->vel_lo[2 * pair_local], WB = ex->vel_lo[2 * pair_local + 1];
        const bool drift = (p.final_mode != kFinalKickOnly);
        float4 l0 = make_float4(0.f, 0.f, 0.f, 0.f), l1 = l0;
        if (ex->acc_lo != nullptr && body0 < p.n) l0 = ex->acc_lo[2 * pair_local];
        if (ex->acc_lo != nullptr && body0 + 1 < p.n) l1 = ex->acc_lo[2 * pair_local + 1];
        comp_step(a0.x, l0.x, p.kick_dt, p.drift_dt, drift, VA.x, WA.x, PA.x, LA.x);
        comp_step(a1.x, l1.x, p.kick_dt, p.drift_dt, drift, VA.y, WA.y, PA.y, LA.y);
        comp_step(a0.y, l0.y, p.kick_dt, p.drift_dt, drift, VA.z, WA.z, PA.z, LA.z);
        comp_step(a1.y, l1.y, p.kick_dt, p.drift_dt, drift, VA.w, WA.w, PA.w, LA.w);
        comp_step(a0.z, l0.z, p.kick_dt, p.drift_dt, drift, VB.x, WB.x, PB.x, LB.x);
        comp_step(a1.z, l1.z, p.kick_dt, p.drift_dt, drift, VB.y, WB.y, PB.y, LB.y);
        p.vel[2 * pair_local] = VA; p.vel[2 * pair_local + 1] = VB;
        ex->vel_lo[2 * pair_local] = WA; ex->vel_lo[2 * pair_local + 1] = WB;
        if (!drift) return;
        ex->pos_lo[2 * pair_local] = LA; ex->pos_lo[2 * pair_local + 1] = LB;
    } else {
        VA.x = fmaf(a0.x, p.kick_dt, VA.x); VA.y = fmaf(a1.x, p.kick_dt, VA.y);
        VA.z = fmaf(a0.y, p.kick_dt, VA.z); VA.w = fmaf(a1.y, p.kick_dt, VA.w);
        VB.x = fmaf(a0.z, p.kick_dt, VB.x); VB.y = fmaf(a1.z, p.kick_dt, VB.y);
        p.vel[2 * pair_local] = VA; p.vel[2 * pair_local + 1] = VB;
        if (p.final_mode == kFinalKickOnly) return;
        PA.x = fmaf(VA.x, p.drift_dt, PA.x); PA.y = fmaf(VA.y, p.drift_dt, PA.y);
        PA.z = fmaf(VA.z, p.drift_dt, PA.z); PA.w = fmaf(VA.w, p.drift_dt, PA.w);
        PB.x = fmaf(VB.x, p.drift_dt, PB.x); PB.y = fmaf(VB.y, p.drift_dt, PB.y);
    }
    p.pos_new[2 * pair_global] = PA; p.pos_new[2 * pair_global + 1] = PB;   // PB.zw = mu, unchanged
    if (PEERS) {
        // same 32 B to every peer over NVLink (replaces MPI_Allgatherv, mpi.c:227-231); made visible to the
        // peers by the flag protocol of signal_peers_kernel / wait_flags_kernel
        for (int q = 0; q < ex->peers.n; ++q) {
            float4* dst = ex->peers.pos_new[q] + 2 * (size_t)pair_global;
            dst[0] = PA; dst[1] = PB;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// the force kernel
// ------------------------------------------------------------------------------------------------
template <int R>
struct Acc {
    float x[R], y[R], z[R];
};

// One j-tile against this thread's R i-bodies, packed f32x2: per j-PAIR and i-body
//   3 FADD2 + 3 FFMA2 + 2 MUFU.RSQ + 3 FMUL2 + 3 FFMA2
//   SELECT: + 2 FSETP + 2 FSEL, the reference's "r < 1e-10 contributes nothing" (cuda.cu:39, mpi.c:64), exact;
//   TRACK : + 1 FMNMX3 instead: only remember the smallest r^2 seen per i-body (the caller redoes the tile
//           with SELECT when that is below the cutoff -- i.e. for the tile that holds the body itself).
// UNI (all bodies have the same mass): the mu_j factor is hoisted out of the sum, 2 FMUL2 instead of 3;
// the caller multiplies the finished sums by mu.  Padding bodies sit at 1e18, where rinv^3 underflows to 0.
template <int R, bool SELECT, bool TRACK, bool UNI, int UNROLL>
__device__ __forceinline__ void tile_sums(const float4* __restrict__ T, const float2 (&nxi)[R],
                                          const float2 (&nyi)[R], const float2 (&nzi)[R], float2 eps2, float cut2,
                                          float2 (&ax)[R], float2 (&ay)[R], float2 (&az)[R], float (&rmin)[R]) {
#pragma unroll
    for (int b = 0; b < R; ++b) ax[b] = ay[b] = az[b] = make_float2(0.f, 0.f);
#pragma unroll UNROLL
    for (int jp = 0; jp < kJTile / 2; ++jp) {
        const float4 A = T[2 * jp], B = T[2 * jp + 1];
        const float2 xj = make_float2(A.x, A.y), yj = make_float2(A.z, A.w);
        const float2 zj = make_float2(B.x, B.y), mj = make_float2(B.z, B.w);
#pragma unroll
        for (int b = 0; b < R; ++b) {
            const float2 dx = __fadd2_rn(xj, nxi[b]);
            const float2 dy = __fadd2_rn(yj, nyi[b]);
            const float2 dz = __fadd2_rn(zj, nzi[b]);
            float2 r2 = __ffma2_rn(dx, dx, eps2);
            r2 = __ffma2_rn(dy, dy, r2);
            r2 = __ffma2_rn(dz, dz, r2);
            float2 ri = make_float2(rsqrt_approx(r2.x), rsqrt_approx(r2.y));
            if (SELECT) {
                ri.x = (r2.x >= cut2) ? ri.x : 0.f;
                ri.y = (r2.y >= cut2) ? ri.y : 0.f;
            }
            if (TRACK) asm("min.f32 %0, %0, %1, %2;" : "+f"(rmin[b]) : "f"(r2.x), "f"(r2.y));   // FMNMX3
            const float2 ri2 = __fmul2_rn(ri, ri);
            float2 s;
            if (UNI) {
                s = __fmul2_rn(ri2, ri);
            } else {
                s = __fmul2_rn(mj, ri);          // (mu * rinv) * rinv^2: no overflow in SI units
                s = __fmul2_rn(s, ri2);
            }
            ax[b] = __ffma2_rn(s, dx, ax[b]);
            ay[b] = __ffma2_rn(s, dy, ay[b]);
            az[b] = __ffma2_rn(s, dz, az[b]);
        }
    }
}

template <int R, int CUT, bool UNI, int UNROLL>
__device__ __forceinline__ void tile_packed(const float4* __restrict__ T, const float2 (&nxi)[R],
                                            const float2 (&nyi)[R], const float2 (&nzi)[R], float2 eps2,
                                            float cut2, Acc<R>& out) {
    float2 ax[R], ay[R], az[R];
    float rmin[R];
    if (CUT == kCutoffTracked) {
        // fast pass without the select; its sums are only kept if no pair of this thread came closer than the
        // cutoff.  Otherwise (the j-tile containing the body itself, coincident bodies) they may hold inf/NaN
        // and the tile is recomputed with the exact rule.  Same result as kCutoff, ~10 % faster.
#pragma unroll
        for (int b = 0; b < R; ++b) rmin[b] = 3.0e38f;
        tile_sums<R, false, true, UNI, UNROLL>(T, nxi, nyi, nzi, eps2, cut2, ax, ay, az, rmin);
        bool close = false;
#pragma unroll
        for (int b = 0; b < R; ++b) close |= !(rmin[b] >= cut2);
        if (close) tile_sums<R, true, false, UNI, UNROLL>(T, nxi, nyi, nzi, eps2, cut2, ax, ay, az, rmin);
    } else {
        tile_sums<R, CUT == kCutoff, false, UNI, UNROLL>(T, nxi, nyi, nzi, eps2, cut2, ax, ay, az, rmin);
    }
#pragma unroll
    for (int b = 0; b < R; ++b) {   // level-2 accumulation (flush the tile partials)
        out.x[b] += ax[b].x + ax[b].y;
        out.y[b] += ay[b].x + ay[b].y;
        out.z[b] += az[b].x + az[b].y;
    }
}

// Scalar FP32 variant of the same tile (12 FP32-pipe + 1 MUFU issue slots per interaction), kept
// as the measured comparison for the packed path.
template <int R, int CUT>
__device__ __forceinline__ void tile_scalar(const float4* __restrict__ T, const float2 (&nxi)[R],
                                            const float2 (&nyi)[R], const float2 (&nzi)[R], float2 eps2,
                                            float cut2, Acc<R>& out) {
    float ax[R], ay[R], az[R], bx[R], by[R], bz[R];
#pragma unroll
    for (int b = 0; b < R; ++b) ax[b] = ay[b] = az[b] = bx[b] = by[b] = bz[b] = 0.f;
#pragma unroll 2
    for (int jp = 0; jp < kJTile / 2; ++jp) {
        const float4 A = T[2 * jp], B = T[2 * jp + 1];
#pragma unroll
        for (int b = 0; b < R; ++b) {
            {
                const float dx = A.x + nxi[b].x, dy = A.z + nyi[b].x, dz = B.x + nzi[b].x;
                float r2 = fmaf(dx, dx, eps2.x); r2 = fmaf(dy, dy, r2); r2 = fmaf(dz, dz, r2);
                float ri = rsqrt_approx(r2);
                if (CUT != kSoftened) ri = (r2 >= cut2) ? ri : 0.f;
                const float s = (B.z * ri) * (ri * ri);
                ax[b] = fmaf(s, dx, ax[b]); ay[b] = fmaf(s, dy, ay[b]); az[b] = fmaf(s, dz, az[b]);
            }
            {
                const float dx = A.y + nxi[b].x, dy = A.w + nyi[b].x, dz = B.y + nzi[b].x;
                float r2 = fmaf(dx, dx, eps2.x); r2 = fmaf(dy, dy, r2); r2 = fmaf(dz, dz, r2);
                float ri = rsqrt_approx(r2);
                if (CUT != kSoftened) ri = (r2 >= cut2) ? ri : 0.f;
                const float s = (B.w * ri) * (ri * ri);
                bx[b] = fmaf(s, dx, bx[b]); by[b] = fmaf(s, dy, by[b]); bz[b] = fmaf(s, dz, bz[b]);
            }
        }
    }
#pragma unroll
    for (int b = 0; b < R; ++b) {
        out.x[b] += ax[b] + bx[b];
        out.y[b] += ay[b] + by[b];
        out.z[b] += az[b] + bz[b];
    }
}


}  // namespace gs
