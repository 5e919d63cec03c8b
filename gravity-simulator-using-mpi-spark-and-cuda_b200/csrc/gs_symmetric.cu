// gs_symmetric.cu -- pair-symmetric force pass (the default; GS_FLAG_DETERMINISTIC turns it off): every unordered
// pair of bodies in different tiles is evaluated ONCE and applied to both bodies (Newton's third law), which the
// reference's cuda.cu also does (cuda.cu:43-49) -- there with a data race, here without.
//
// Why it pays: the kernel is bound by the FP32 FMA pipe.  A (i-body, j-pair) step costs 12 packed
// instructions + 2 MUFU for two directed interactions in the row-gather kernel; sharing d, r^2, rsqrt and rinv^3
// between the two directions costs 14 + 2 for four.  The usual objection -- the j-side sums need a reduction across
// the 32 lanes that all look at the same j -- is avoided by ROTATING the j-pairs over the lanes: in iteration `it`
// lane l works on j-pair (it + l) mod 32 of the current 128-pair chunk (and on that pair + 32, + 64, + 96), so no two
// lanes of a warp touch the same j at the same time.  A pair's running sums follow it from lane to lane by warp
// shuffles and are stored once per chunk into a per-warp shared-memory array.
// After a j-tile the warps' arrays are added up and added into the owner's j-side sums in global memory -- local
// memory, or NVLink peer memory when the j-bodies belong to another GPU -- as exact wide fixed point with 64-bit
// integer atomics (gs_device.cuh, fx_add): the result does not depend on the order of arrival, so the pass is
// bit-reproducible.
//
// Work: rows of (i-tile x j-tile) units; row I starts with its `diag` = i_tile/j_tile own j-tiles, done as
// plain (one-directional, all ordered pairs) units, followed by symmetric units for every later j-tile of the
// rank's slice (earlier ones are covered by earlier rows) and, on P GPUs, for the next floor((P-1)/2) slices and
// half of the opposite one (SymParams, sym_row_units/sym_unit_tile in gs_kernels.h).  Units are handed to the persistent CTAs as contiguous
// ranges of equal estimated cost (host-side plan, build_sym_plan); i-side sums of a row split between CTAs go
// through the same partial-sum scratch as the row-gather kernel; sym_finish_kernel adds i-side and j-side sums
// and finishes (integration or acceleration output).
#include <cstdlib>
#include <vector>

#include "gs_device.cuh"

namespace gs {

static_assert(kSymAccjBytesPerBody == kFxWordsPerBody * (int)sizeof(long long), "accj sizing");
constexpr int kSymDefaultNPair = 4;
constexpr size_t kSymTileBytes = (size_t)kJTile * sizeof(float4);
constexpr size_t kSymAccBytesPerWarp = (size_t)(kJTile / 2) * (sizeof(float4) + sizeof(float2));
constexpr size_t sym_smem_for(int threads) {
    return kStages * kSymTileBytes + 2 * kStages * sizeof(uint64_t) + 64 + (size_t)(threads / 32) * kSymAccBytesPerWarp;
}

// How the reference's "r < 1e-10 contributes nothing" (cuda.cu:39, mpi.c:64) is applied inside a symmetric unit
// when eps == 0:
//   kSymSelect: rinv = (r^2 >= 1e-20) ? rsqrt(r^2) : 0 on every interaction (2 FSETP + 2 FSEL per j-pair and i-body);
//   kSymTrack : no select and no extra instruction.  The kernel reads positions scaled by 2^-10 (SymParams::pos is
//               then a scaled copy, kSymOutScale = 2^-20 undoes it: powers of two, exact), so rinv is 2^10 times larger
//               and rinv^3 OVERFLOWS to +inf for every pair with r < 1.46e-10 -- a superset of the pairs the rule
//               removes (r < 1e-10).  inf (or NaN = inf * 0) then poisons the i-side sums of that lane; the warp
//               checks its tile sums once per tile and redoes the unit with kSymSelect if any is not finite.
//               In a symmetric unit the i-tile and the j-tile are different tiles, so a body never meets itself
//               there: a redo needs two distinct bodies within 1.46e-10 of each other (in fp32: coincident bodies,
//               or tiny separations next to the origin).
enum SymCut : int { kSymSoft = 0, kSymSelect = 1, kSymTrack = 2 };

// One j-tile, both directions.  NPAIR j-pairs per lane and iteration (pairs q, q+32, ... of a 32*NPAIR-pair chunk,
// q = (it + lane) mod 32): NPAIR * R independent chains per iteration.
// How a j-pair's running sums follow the pair from lane to lane (it is worked on by lane l in iteration `it` and by
// lane l-1 in iteration it+1):
//   SHFL = true : they stay in registers and move with six 32-bit shuffles per pair and iteration; after the 32
//                 iterations of a chunk every pair has been round the warp once and its sums are stored, once, into
//                 this warp's accA/accB ({ax0,ax1,ay0,ay1}, {az0,az1} per j-pair; plain stores, nothing to clear);
//   SHFL = false: read-modify-write of accA/accB every iteration with a __syncwarp() in between (the first version
//                 of this pass; kept for A/B measurements: the round trip through shared memory costs ~9 %).
// Returns true (warp-uniform, kSymTrack only) when the unit has to be redone with the select; `out` is then untouched
// and the warp's accA/accB hold garbage.
template <int R, int CUT, bool UNI, int NPAIR, bool SHFL>
__device__ __forceinline__ bool sym_tile(const float4* __restrict__ T, float4* __restrict__ accA,
                                         float2* __restrict__ accB, int lane, const float2 (&nxi)[R],
                                         const float2 (&nyi)[R], const float2 (&nzi)[R], const float2 (&nmui)[R],
                                         float2 eps2, float cut2, Acc<R>& out) {
    float2 ax[R], ay[R], az[R];
#pragma unroll
    for (int b = 0; b < R; ++b) ax[b] = ay[b] = az[b] = make_float2(0.f, 0.f);
    const int from = (lane + 1) & 31;       // the lane that worked on "my next pair" in this iteration
    // one j-pair (records A, B) against the R i-bodies, both directions
    auto pair_step = [&](const float4& A, const float4& B, float2& jx, float2& jy, float2& jz) {
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
            if (CUT == kSymSelect) {
                ri.x = (r2.x >= cut2) ? ri.x : 0.f;
                ri.y = (r2.y >= cut2) ? ri.y : 0.f;
            }
            const float2 ri2 = __fmul2_rn(ri, ri);
            const float2 t = __fmul2_rn(ri2, ri);                       // rinv^3, shared by both directions
            const float2 s = UNI ? t : __fmul2_rn(mj, t);               // on i:  +mu_j t d
            const float2 u = UNI ? t : __fmul2_rn(nmui[b], t);          // on j:  -mu_i t d  (UNI: sign and mu at the flush)
            ax[b] = __ffma2_rn(s, dx, ax[b]);
            ay[b] = __ffma2_rn(s, dy, ay[b]);
            az[b] = __ffma2_rn(s, dz, az[b]);
            jx = __ffma2_rn(u, dx, jx);
            jy = __ffma2_rn(u, dy, jy);
            jz = __ffma2_rn(u, dz, jz);
        }
    };
    constexpr int NCHUNK = kJTile / (64 * NPAIR);
    if (SHFL) {
        for (int chunk = 0; chunk < NCHUNK; ++chunk) {
            float2 jx[NPAIR], jy[NPAIR], jz[NPAIR];
#pragma unroll
            for (int h = 0; h < NPAIR; ++h) jx[h] = jy[h] = jz[h] = make_float2(0.f, 0.f);
#pragma unroll 1
            for (int it = 0; it < 32; ++it) {
                const int p0 = chunk * (32 * NPAIR) + ((it + lane) & 31);
                float4 A[NPAIR], B[NPAIR];
#pragma unroll
                for (int h = 0; h < NPAIR; ++h) { A[h] = T[2 * (p0 + 32 * h)]; B[h] = T[2 * (p0 + 32 * h) + 1]; }
#pragma unroll
                for (int h = 0; h < NPAIR; ++h) pair_step(A[h], B[h], jx[h], jy[h], jz[h]);
                // (measured: issuing each group's shuffles right after the group, with the next iteration's first
                // loads prefetched, is 2.5 % slower -- ptxas sinks them to the end anyway and the loop gets longer)
#pragma unroll
                for (int h = 0; h < NPAIR; ++h) {
                    jx[h].x = __shfl_sync(0xffffffffu, jx[h].x, from); jx[h].y = __shfl_sync(0xffffffffu, jx[h].y, from);
                    jy[h].x = __shfl_sync(0xffffffffu, jy[h].x, from); jy[h].y = __shfl_sync(0xffffffffu, jy[h].y, from);
                    jz[h].x = __shfl_sync(0xffffffffu, jz[h].x, from); jz[h].y = __shfl_sync(0xffffffffu, jz[h].y, from);
                }
            }
            // after 32 shuffles lane l holds the finished sums of pair l (+ 32h) of this chunk
#pragma unroll
            for (int h = 0; h < NPAIR; ++h) {
                accA[chunk * (32 * NPAIR) + lane + 32 * h] = make_float4(jx[h].x, jx[h].y, jy[h].x, jy[h].y);
                accB[chunk * (32 * NPAIR) + lane + 32 * h] = jz[h];
            }
        }
    } else {
        for (int chunk = 0; chunk < NCHUNK; ++chunk) {
#pragma unroll 1
            for (int it = 0; it < 32; ++it) {
                const int p0 = chunk * (32 * NPAIR) + ((it + lane) & 31);
                float4 A[NPAIR], B[NPAIR];
                float2 jx[NPAIR], jy[NPAIR], jz[NPAIR];
#pragma unroll
                for (int h = 0; h < NPAIR; ++h) {
                    A[h] = T[2 * (p0 + 32 * h)]; B[h] = T[2 * (p0 + 32 * h) + 1];
                    const float4 JA = accA[p0 + 32 * h];
                    jx[h] = make_float2(JA.x, JA.y); jy[h] = make_float2(JA.z, JA.w); jz[h] = accB[p0 + 32 * h];
                }
#pragma unroll
                for (int h = 0; h < NPAIR; ++h) pair_step(A[h], B[h], jx[h], jy[h], jz[h]);
#pragma unroll
                for (int h = 0; h < NPAIR; ++h) {
                    accA[p0 + 32 * h] = make_float4(jx[h].x, jx[h].y, jy[h].x, jy[h].y);
                    accB[p0 + 32 * h] = jz[h];
                }
                __syncwarp();     // next iteration another lane picks these pairs up
            }
        }
    }
    if (CUT == kSymTrack) {
        float2 c = make_float2(0.f, 0.f);         // inf and NaN survive a sum
#pragma unroll
        for (int b = 0; b < R; ++b) c = __fadd2_rn(__fadd2_rn(c, ax[b]), __fadd2_rn(ay[b], az[b]));
        const float cs = c.x + c.y;
        if (__any_sync(0xffffffffu, !(fabsf(cs) <= 3.0e38f))) return true;
    }
#pragma unroll
    for (int b = 0; b < R; ++b) {
        out.x[b] += ax[b].x + ax[b].y;
        out.y[b] += ay[b].x + ay[b].y;
        out.z[b] += az[b].x + az[b].y;
    }
    return false;
}

template <int R, int THREADS, int CUT, bool UNI, int NPAIR, bool SHFL>
__global__ void __launch_bounds__(THREADS, 1) sym_force_kernel(const SymParams p) {
    constexpr int ITILE = THREADS * R;
    constexpr int NP = R / 2;
    constexpr int WARPS = THREADS / 32;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4* tiles = reinterpret_cast<float4*>(smem_raw);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + kStages * kSymTileBytes);
    uint64_t* empty = full + kStages;
    float4* accA_all = reinterpret_cast<float4*>(smem_raw + kStages * kSymTileBytes + 2 * kStages * sizeof(uint64_t) + 64);
    float2* accB_all = reinterpret_cast<float2*>(accA_all + (size_t)WARPS * (kJTile / 2));

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const SymCta me = p.ctas[blockIdx.x];
    const int nunits = me.count;
    if (nunits <= 0) return;
    float4* accA = accA_all + warp * (kJTile / 2);
    float2* accB = accB_all + warp * (kJTile / 2);

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], WARPS);
        }
        fence_mbar_init();
    }
    if (!SHFL)
        for (int e = lane; e < kJTile / 2; e += 32) {
            accA[e] = make_float4(0.f, 0.f, 0.f, 0.f);
            accB[e] = make_float2(0.f, 0.f);
        }
    __syncthreads();

    // producer cursor: the next unit whose j-tile has to be loaded (only thread 0 uses it)
    int prow = me.row0, pk = me.k0, loaded = 0;
    auto load_next = [&](int stage) {
        const int jtile = sym_unit_tile(p, prow, pk);
        mbar_arrive_expect_tx(&full[stage], (uint32_t)kSymTileBytes);
        tma_bulk_g2s(tiles + (size_t)stage * kJTile, p.pos + (size_t)jtile * kJTile, (uint32_t)kSymTileBytes, &full[stage]);
        ++loaded;
        if (++pk >= sym_row_units(p, prow)) { ++prow; pk = 0; }
    };
    if (tid == 0)
        for (int s = 0; s < kStages && s < nunits; ++s) load_next(s);

    const float2 eps2 = make_float2(p.eps2, p.eps2);
    int row = me.row0, kk = me.k0, k = 0;
    while (k < nunits) {
        const int row_units = sym_row_units(p, row);
        const int seg = min(row_units - kk, nunits - k);
        const bool whole = (kk == 0) && (seg == row_units);

        float2 nxi[R], nyi[R], nzi[R], nmui[R];
        Acc<R> acc;
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            const int pair = (p.i_begin >> 1) + row * (ITILE / 2) + q * THREADS + tid;
            const float4 A = p.pos[2 * pair], B = p.pos[2 * pair + 1];
            nxi[2 * q] = make_float2(-A.x, -A.x); nxi[2 * q + 1] = make_float2(-A.y, -A.y);
            nyi[2 * q] = make_float2(-A.z, -A.z); nyi[2 * q + 1] = make_float2(-A.w, -A.w);
            nzi[2 * q] = make_float2(-B.x, -B.x); nzi[2 * q + 1] = make_float2(-B.y, -B.y);
            nmui[2 * q] = make_float2(-B.z, -B.z); nmui[2 * q + 1] = make_float2(-B.w, -B.w);
        }
#pragma unroll
        for (int b = 0; b < R; ++b) acc.x[b] = acc.y[b] = acc.z[b] = 0.f;

        for (int t = 0; t < seg; ++t, ++k, ++kk) {
            const int stage = k % kStages;
            if (tid == 0 && k >= kRefillLag && loaded < nunits) {   // refill the stage used kRefillLag units ago
                const int kr = k - kRefillLag, sr = kr % kStages;
                mbar_wait(&empty[sr], (uint32_t)((kr / kStages) & 1));
                load_next(sr);
            }
            mbar_wait(&full[stage], (uint32_t)((k / kStages) & 1));
            const float4* T = tiles + (size_t)stage * kJTile;
            const bool symmetric = (kk >= p.diag);
            if (!symmetric) {
                // the row's own j-tiles hold the i-bodies themselves: one direction, select on every interaction
                tile_packed<R, (CUT == kSymSoft ? kSoftened : kCutoff), UNI, 2>(T, nxi, nyi, nzi, eps2, p.cut2, acc);
            } else if (sym_tile<R, CUT, UNI, NPAIR, SHFL>(T, accA, accB, lane, nxi, nyi, nzi, nmui, eps2, p.cut2, acc)) {
                // kSymTrack only: something of this warp came within the cutoff (coincident bodies) -- exact redo
                if (!SHFL) {
                    for (int e = lane; e < kJTile / 2; e += 32) {
                        accA[e] = make_float4(0.f, 0.f, 0.f, 0.f);
                        accB[e] = make_float2(0.f, 0.f);
                    }
                    __syncwarp();
                }
                sym_tile<R, kSymSelect, UNI, (CUT == kSymTrack ? 2 : NPAIR), SHFL>(T, accA, accB, lane, nxi, nyi, nzi, nmui,
                                                                                 eps2, p.cut2, acc);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[stage]);
            if (symmetric) {
                // j-side sums of this tile: add the warps' arrays, send them off, clear them
                __syncthreads();
                const int jtile = sym_unit_tile(p, row, kk);
                const int owner = jtile / p.slice_tiles;                   // whose bodies these are: local or NVLink peer memory
                // accj layout: per j-tile, word (component * kFxLimbs + limb) * kJTile + body: consecutive lanes =
                // consecutive bodies = consecutive words, so each of the integer atomics below is one coalesced request
                long long* const jbase = p.accj[owner] + (size_t)kFxWordsPerBody * (size_t)(jtile - owner * p.slice_tiles) * kJTile;
                const float w = (UNI ? -p.mu_uniform : 1.f) * (CUT == kSymTrack ? kSymOutScale : 1.f);
                for (int body = tid; body < kJTile; body += THREADS) {
                    const int e = body >> 1;
                    const bool odd = (body & 1) != 0;
                    float sx = 0.f, sy = 0.f, sz = 0.f;
#pragma unroll
                    for (int wv = 0; wv < WARPS; ++wv) {
                        const float4 a = accA_all[wv * (kJTile / 2) + e];
                        const float2 b = accB_all[wv * (kJTile / 2) + e];
                        sx += odd ? a.y : a.x; sy += odd ? a.w : a.z; sz += odd ? b.y : b.x;
                    }
                    // Exact fixed-point adds (gs_device.cuh): the owner's j-side sums do not depend on the order in
                    // which CTAs and GPUs arrive.  System scope when other GPUs add into the same words over NVLink
                    // (only system-scope atomics are atomic across GPUs; SASS: RED.E.ADD.64...SYS); on one GPU
                    // device scope is enough.
                    long long* dst = jbase + body;
                    if (p.world > 1) {
                        fx_add<true>(dst, kJTile, w * sx); fx_add<true>(dst + kFxLimbs * kJTile, kJTile, w * sy);
                        fx_add<true>(dst + 2 * kFxLimbs * kJTile, kJTile, w * sz);
                    } else {
                        fx_add<false>(dst, kJTile, w * sx); fx_add<false>(dst + kFxLimbs * kJTile, kJTile, w * sy);
                        fx_add<false>(dst + 2 * kFxLimbs * kJTile, kJTile, w * sz);
                    }
                }
                if (!SHFL) {
                    __syncthreads();
                    for (int e = tid; e < WARPS * (kJTile / 2); e += THREADS) {
                        accA_all[e] = make_float4(0.f, 0.f, 0.f, 0.f);
                        accB_all[e] = make_float2(0.f, 0.f);
                    }
                }
                __syncthreads();
            }
        }

        // i-side sums of this row segment
#pragma unroll
        for (int q = 0; q < NP; ++q) {
            const int in_tile = 2 * (q * THREADS + tid);
            const float w = (UNI ? p.mu_uniform : 1.f) * (CUT == kSymTrack ? kSymOutScale : 1.f);
            const float4 a0 = make_float4(w * acc.x[2 * q], w * acc.y[2 * q], w * acc.z[2 * q], 0.f);
            const float4 a1 = make_float4(w * acc.x[2 * q + 1], w * acc.y[2 * q + 1], w * acc.z[2 * q + 1], 0.f);
            float4* dst = whole ? p.acc_whole + (size_t)row * ITILE + in_tile
                                : p.partial + ((size_t)blockIdx.x * 2 + (row == me.row0 ? 0 : 1)) * ITILE + in_tile;
            dst[0] = a0; dst[1] = a1;
        }
        ++row; kk = 0;
    }
}

// a_i = (i-side sums: whole row or per-CTA partials in fixed order) + (j-side sums), then integrate / output.
__global__ void __launch_bounds__(256) sym_finish_kernel(const StepParams p, const SymParams sp, const int2* rows,
                                                        int ITILE, const FinishExtras ex) {
    const int pair_local = blockIdx.x * blockDim.x + threadIdx.x;
    if (pair_local >= sp.ni_tiles * (ITILE / 2)) return;
    const int row = pair_local / (ITILE / 2);
    const int q = pair_local - row * (ITILE / 2);
    const int2 owners = rows[row];
    float4 a0, a1;
    if (owners.x == owners.y) {
        a0 = sp.acc_whole[2 * (size_t)pair_local];
        a1 = sp.acc_whole[2 * (size_t)pair_local + 1];
    } else {
        a0 = make_float4(0.f, 0.f, 0.f, 0.f); a1 = a0;
        for (int c = owners.x; c <= owners.y; ++c) {
            const SymCta cta = sp.ctas[c];
            if (cta.count <= 0) continue;
            const float4* src = sp.partial + ((size_t)c * 2 + (row == cta.row0 ? 0 : 1)) * ITILE + 2 * q;
            const float4 s0 = src[0], s1 = src[1];
            a0.x += s0.x; a0.y += s0.y; a0.z += s0.z;
            a1.x += s1.x; a1.y += s1.y; a1.z += s1.z;
        }
    }
    const int body0 = 2 * pair_local;                                    // local bodies body0, body0 + 1
    long long* mine = sp.accj[sp.rank] + (size_t)kFxWordsPerBody * (size_t)(body0 / kJTile) * kJTile + (body0 % kJTile);
    float j[6];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        long long l0[kFxLimbs], l1[kFxLimbs];
#pragma unroll
        for (int k = 0; k < kFxLimbs; ++k) {
            longlong2* w = reinterpret_cast<longlong2*>(mine + (size_t)(c * kFxLimbs + k) * kJTile);
            const longlong2 v = *w;
            l0[k] = v.x; l1[k] = v.y;
            *w = make_longlong2(0, 0);                                     // ready for the next pass
        }
        j[c] = fx_value(l0); j[3 + c] = fx_value(l1);
    }
    a0.x += j[0]; a0.y += j[1]; a0.z += j[2];
    a1.x += j[3]; a1.y += j[4]; a1.z += j[5];
    finish_pair<true>(p, pair_local, a0, a1, &ex);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
// Contiguous unit ranges of equal estimated cost (plain unit = cost_plain, symmetric unit = cost_sym); only the
// geometry fields of `sp` are read.
void build_sym_plan(const SymParams& sp, int grid, int cost_plain, int cost_sym, std::vector<SymCta>& ctas,
                    std::vector<int2>& rows) {
    const int ni = sp.ni_tiles, diag = sp.diag;
    auto row_units = [&](int I) { return (int64_t)sym_row_units(sp, I); };
    auto row_cost = [&](int I) { return (int64_t)diag * cost_plain + (row_units(I) - diag) * cost_sym; };
    std::vector<int64_t> prefix(ni + 1, 0), uprefix(ni + 1, 0);
    for (int I = 0; I < ni; ++I) { prefix[I + 1] = prefix[I] + row_cost(I); uprefix[I + 1] = uprefix[I] + row_units(I); }
    const int64_t W = prefix[ni];
    // first unit (row, k) whose start cost is >= c*W/grid
    std::vector<int> brow(grid + 1), bk(grid + 1);
    int I = 0;
    for (int c = 0; c <= grid; ++c) {
        const int64_t B = (c == grid) ? W : (int64_t)(((__int128)W * c) / grid);
        while (I < ni && prefix[I + 1] <= B) ++I;
        if (I >= ni) { brow[c] = ni; bk[c] = 0; continue; }
        const int64_t off = B - prefix[I];
        int64_t k;
        if (off <= (int64_t)diag * cost_plain) k = (off + cost_plain - 1) / cost_plain;
        else k = diag + (off - (int64_t)diag * cost_plain + cost_sym - 1) / cost_sym;
        if (k >= row_units(I)) { brow[c] = I + 1; bk[c] = 0; } else { brow[c] = I; bk[c] = (int)k; }
    }
    ctas.assign(grid, SymCta{0, 0, 0, 0});
    rows.assign(ni, int2{grid, -1});
    for (int c = 0; c < grid; ++c) {
        const int64_t g0 = uprefix[brow[c]] + bk[c], g1 = uprefix[brow[c + 1]] + bk[c + 1];
        ctas[c].row0 = brow[c]; ctas[c].k0 = bk[c]; ctas[c].count = (int)(g1 - g0);
        if (g1 <= g0) continue;
        const int last_row = (bk[c + 1] == 0) ? brow[c + 1] - 1 : brow[c + 1];
        for (int r = brow[c]; r <= last_row && r < ni; ++r) {
            if (c < rows[r].x) rows[r].x = c;
            if (c > rows[r].y) rows[r].y = c;
        }
    }
}

typedef void (*SymKernel)(const SymParams);
static int env_int(const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; }
// measurement knobs: j-pairs per lane and iteration; 0 = j-side sums through shared memory (old scheme)
// (read at every launch, not cached: the probes switch them inside one process)
static int sym_npair() { const int v = env_int("GRAVSIM_SYM_NPAIR", 0); return (v == 2 || v == 4) ? v : kSymDefaultNPair; }
static bool sym_shfl() { return env_int("GRAVSIM_SYM_SHFL", 1) != 0; }
static int sym_cut_of(int cut_mode) {
    return cut_mode == kSoftened ? kSymSoft : cut_mode == kCutoff ? kSymSelect : kSymTrack;
}
template <int RR, int TH, int NP, bool SH>
static SymKernel pick_sym_rn(int cut, bool uni) {
    switch (sym_cut_of(cut)) {
        case kSymSoft: return uni ? (SymKernel)sym_force_kernel<RR, TH, kSymSoft, true, NP, SH> : (SymKernel)sym_force_kernel<RR, TH, kSymSoft, false, NP, SH>;
        case kSymSelect: return uni ? (SymKernel)sym_force_kernel<RR, TH, kSymSelect, true, NP, SH> : (SymKernel)sym_force_kernel<RR, TH, kSymSelect, false, NP, SH>;
        default: return uni ? (SymKernel)sym_force_kernel<RR, TH, kSymTrack, true, NP, SH> : (SymKernel)sym_force_kernel<RR, TH, kSymTrack, false, NP, SH>;
    }
}
template <int RR, int TH>
static SymKernel pick_sym_t(int cut, bool uni) {
    const int np = sym_npair();
    if (sym_shfl()) return np == 4 ? pick_sym_rn<RR, TH, 4, true>(cut, uni) : pick_sym_rn<RR, TH, 2, true>(cut, uni);
    return np == 4 ? pick_sym_rn<RR, TH, 4, false>(cut, uni) : pick_sym_rn<RR, TH, 2, false>(cut, uni);
}
// (i-bodies per thread, threads per CTA) pairs the pass is built for
static SymKernel pick_sym(int R, int threads, int cut, bool uni) {
    if (threads == 256 && R == 4) return pick_sym_t<4, 256>(cut, uni);
    if (threads == 256 && R == 6) return pick_sym_t<6, 256>(cut, uni);
    if (threads == 256 && R == 8) return pick_sym_t<8, 256>(cut, uni);
    if (threads == 384 && R == 8) return sym_shfl() ? pick_sym_rn<8, 384, 2, true>(cut, uni) : pick_sym_rn<8, 384, 2, false>(cut, uni);
    if (threads == 512 && R == 4) return pick_sym_t<4, 512>(cut, uni);
    return nullptr;
}

bool sym_supported(int R, int threads) { return pick_sym(R, threads, kSoftened, false) != nullptr; }
size_t sym_smem_bytes(int threads) { return sym_smem_for(threads); }

cudaError_t sym_kernel_setup(int R, int threads, int cut_mode, int* ctas_per_sm, int* regs) {
    int best = 1 << 30, r = 0;
    const size_t smem = sym_smem_for(threads);
    for (int uni = 0; uni < 2; ++uni) {
        SymKernel k = pick_sym(R, threads, cut_mode, uni != 0);
        if (!k) return cudaErrorInvalidValue;
        cudaError_t e = cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        int n = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, (const void*)k, threads, smem);
        if (e != cudaSuccess) return e;
        cudaFuncAttributes fa;
        e = cudaFuncGetAttributes(&fa, (const void*)k);
        if (e != cudaSuccess) return e;
        if (n < best) best = n;
        if (fa.numRegs > r) r = fa.numRegs;
    }
    *ctas_per_sm = best;
    *regs = r;
    return cudaSuccess;
}

cudaError_t launch_sym_force(int R, int threads, int cut_mode, const SymParams& p, int grid, cudaStream_t s) {
    SymKernel k = pick_sym(R, threads, cut_mode, p.mu_uniform != 0.f);
    if (!k) return cudaErrorInvalidValue;
    k<<<grid, threads, sym_smem_for(threads), s>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_sym_finish(const StepParams& p, const SymParams& sp, const int2* rows, int i_tile,
                              const FinishExtras* extras, cudaStream_t s) {
    const int pairs = sp.ni_tiles * (i_tile / 2);
    FinishExtras none;
    none.peers.n = 0; none.pos_lo = nullptr; none.vel_lo = nullptr; none.acc_lo = nullptr;
    sym_finish_kernel<<<(pairs + 255) / 256, 256, 0, s>>>(p, sp, rows, i_tile, extras ? *extras : none);
    return cudaGetLastError();
}

}  // namespace gs
