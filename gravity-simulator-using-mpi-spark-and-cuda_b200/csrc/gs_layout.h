// gs_layout.h -- device data layout and the static stream-K work partition.
// Shared by host and device code (and unit-tested on the CPU through gs_plan_*).
//
// HBM layout ("packed pairs"): bodies are stored two at a time as two float4,
//     rec[2p]   = { x[2p], x[2p+1], y[2p], y[2p+1] }
//     rec[2p+1] = { z[2p], z[2p+1], mu[2p], mu[2p+1] }      mu = G * mass
// i.e. 16 B per body, a structure-of-arrays at pair granularity.  A 128-bit shared-memory load
// yields two (even,odd) register pairs that feed the sm_100 packed-FP32 instructions
// (FADD2/FMUL2/FFMA2) directly, one j-pair per instruction.  Velocities use the same shape
// ({vx,vx',vy,vy'}, {vz,vz',0,0}).  Body range [b0,b1) with even b0 is float4 range [b0,b1).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define GS_HD __host__ __device__ __forceinline__
#else
#define GS_HD static inline
#endif

namespace gs {

constexpr int kThreads = 256;     // threads per CTA of the force kernel
constexpr int kJTile   = 512;     // bodies per j-tile = one TMA bulk copy of 8 KB
constexpr int kStages  = 4;       // smem ring depth
constexpr int kRefillLag = 2;     // a stage is refilled this many tiles after its use

// Work unit u in [0, ni_tiles*nj_tiles): i-tile u / nj_tiles, (virtual) j-tile u % nj_tiles.
// Persistent CTA c of `grid` owns the contiguous unit range [c*U/grid, (c+1)*U/grid).
GS_HD int64_t cta_unit_begin(int64_t total_units, int grid, int cta) {
    return (total_units * (int64_t)cta) / grid;
}

// CTA that owns unit u (largest c with cta_unit_begin(c) <= u).
GS_HD int unit_owner(int64_t total_units, int grid, int64_t u) {
    int c = (int)((u * (int64_t)grid) / total_units);
    if (c >= grid) c = grid - 1;
    while (c + 1 < grid && cta_unit_begin(total_units, grid, c + 1) <= u) ++c;
    while (c > 0 && cta_unit_begin(total_units, grid, c) > u) --c;
    return c;
}

// An i-tile is "split" when its nj_tiles units are not all owned by one CTA.  Split tiles are
// finished by the fix-up kernel from per-CTA partial sums; whole tiles are finished (integrated)
// inside the force kernel.  A CTA touches at most two split tiles: its first and its last.
struct TileOwners { int first, last; bool split; };
GS_HD TileOwners tile_owners(int64_t ni_tiles, int64_t nj_tiles, int grid, int64_t it) {
    int64_t U = ni_tiles * nj_tiles;
    TileOwners o;
    o.first = unit_owner(U, grid, it * nj_tiles);
    o.last  = unit_owner(U, grid, (it + 1) * nj_tiles - 1);
    o.split = (o.first != o.last);
    return o;
}

// Scratch slot of CTA c's partial sum for i-tile `it`: 0 if `it` is the tile its range starts
// in, else 1 (then it is necessarily the tile its range ends in).
GS_HD int partial_slot(int64_t ni_tiles, int64_t nj_tiles, int grid, int cta, int64_t it) {
    int64_t U = ni_tiles * nj_tiles;
    int64_t first_tile = cta_unit_begin(U, grid, cta) / nj_tiles;
    return (it == first_tile) ? 0 : 1;
}

}  // namespace gs
