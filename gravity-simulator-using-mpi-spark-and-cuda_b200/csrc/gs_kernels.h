// gs_kernels.h -- launch wrappers of the sm_100a kernels (gs_kernels.cu) used by the C-ABI (gs_api.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "gs_layout.h"

namespace gs {

enum FinalMode : int { kFinalIntegrate = 0,   // v += a*kick_dt; x += v*drift_dt -> pos_new
                       kFinalAccOut = 1,      // write accelerations, state untouched
                       kFinalKickOnly = 2 };  // v += a*kick_dt only (leapfrog velocity synchronisation)
enum CutMode : int { kSoftened = 0,        // eps > 0
                     kCutoff = 1,          // eps == 0: reference r<1e-10 rule by a select on every interaction
                     kCutoffTracked = 2 }; // eps == 0: same result; select-free fast pass + exact redo of the rare tiles

// Kernel variants (gs_config.variant / env GRAVSIM_VARIANT): i-bodies per thread R, CTA size, inner
// unroll, packed f32x2 or scalar arithmetic.  The table lives in gs_kernels.cu (GS_VARIANTS).
constexpr int kNumVariants = 17;
constexpr int kDefaultVariant = 1;
constexpr float kPadPosition = 1.0e18f;   // padding bodies: (1e18,1e18,1e18), mu = 0
struct VariantInfo { int R, threads, unroll, packed, i_tile; };
bool variant_info(int variant, VariantInfo* out);     // false if the variant does not exist

constexpr int kSymAccjBytesPerBody = 3 * 5 * 8;   // = kFxWordsPerBody * sizeof(long long), checked in gs_symmetric.cu
constexpr int kMaxPeers = 7;   // direct peer-memory exchange supports up to 8 GPUs (one NVSwitch domain)

struct StepParams {
    const float4* pos_old;   // packed pairs, n_pad bodies (all ranks' bodies)
    float4* pos_new;         // packed pairs, n_pad bodies; only this rank's slice is written
    float4* vel;             // packed pairs, this rank's slice only (local index)
    const float4* acc_in;    // optional: accelerations to add before finishing (local index)
    float4* acc_out;         // kFinalAccOut: accelerations (x,y,z,0) per local body
    float4* partial;         // scratch [grid][2][i_tile]: partial sums of split i-tiles
    int n;                   // real bodies; bodies >= n are padding and are never moved
    int i_begin;             // first body of this rank's slice (global index, multiple of the i-tile)
    int ni_tiles;            // i-tiles in this rank's slice
    int nj_tiles;            // j-tiles covered by this launch (virtual index v in [0,nj_tiles))
    int nj_total;            // n_pad / kJTile
    int j_rot;               // physical j-tile = (v + j_rot) mod nj_total
    float eps2;              // softening^2 (kSoftened) or 0
    float cut2;              // kCutoff: pairs with r^2 < cut2 contribute nothing (1e-20)
    float kick_dt;           // v += a * kick_dt
    float drift_dt;          // x += v_new * drift_dt
    float mu_uniform;        // != 0: every real body has this G*m; use the hoisted-mass kernel
    int final_mode;          // FinalMode
};

// Peers' pos_new buffers (NVLink peer memory, same indexing as StepParams::pos_new).  Kept out of StepParams:
// growing the force kernel's parameter block changes ptxas' schedule of its inner loop (measured -2.5 %).
struct PeerTargets { int n; float4* pos_new[kMaxPeers]; };
// What only the finish kernel needs: peers, and the low-order words of the compensated integration
// (GS_FLAG_COMPENSATED: x = pos + pos_lo, v = vel + vel_lo; null otherwise).  Local (slice) indexing.
// acc_lo (nullable): low-order words of the accelerations (GS_FLAG_FP64_FORCES), used by the compensated update.
struct FinishExtras { PeerTargets peers; float4* pos_lo; float4* vel_lo; const float4* acc_lo; };

size_t force_smem_bytes();
// Max resident CTAs per SM of the force kernel for this variant / cut mode, and its registers.
cudaError_t force_kernel_occupancy(int variant, int cut_mode, bool uniform, int* ctas_per_sm, int* regs);
cudaError_t launch_force(int variant, int cut_mode, const StepParams& p, int grid, cudaStream_t s);
// Finishes the i-tiles whose units were split across CTAs (sums the partials, then integrates
// or writes accelerations).  `grid_main` is the force kernel's grid.
// whole_acc != nullptr: also finish the whole tiles, whose accelerations the force kernel left there.
cudaError_t launch_fixup(int variant, const StepParams& p, int grid_main, const float4* whole_acc,
                         const FinishExtras* extras, cudaStream_t s);

// GS_FLAG_FP64_FORCES: accelerations of this rank's bodies summed and evaluated in fp64 (the precision of mpi.c:59-73),
// one i-body per thread, every ordered pair; positions are pos (+ pos_lo_all when the low-order words of ALL bodies
// are on this GPU, else nullptr), masses are the original doubles of ALL bodies (mu = G*m in fp64).
// Output as hi + lo float4 pairs per local body.
cudaError_t launch_force_f64(const float4* pos, const float4* pos_lo_all, const double* mass, double G, int n_pad,
                             int i_begin, int n_local_pad, double eps2, double cut2, float4* acc_hi, float4* acc_lo,
                             cudaStream_t s);
// Finishes every local pair from accelerations left in `acc` (integration / output / kick, peers, compensated update).
cudaError_t launch_finish_acc(const StepParams& p, int n_local_pad, const float4* acc, const FinishExtras* extras,
                              cudaStream_t s);

// Peer-memory exchange (replaces the per-step collective when every GPU can map its peers' buffers):
// the integrating epilogue stores new positions straight into every peer's position buffer over
// NVLink; afterwards each rank raises a step counter in every peer's flag array, and before a pass
// that reads remote bodies it waits until all peers' counters have reached the step it needs.
struct PeerFlags { uint32_t* slot[kMaxPeers + 1]; };   // slot[q] = address of "my" counter inside rank q's flag array
cudaError_t launch_signal_peers(const PeerFlags& f, int world, int my_rank, uint32_t value, cudaStream_t s);
// Spins (bounded by timeout_ns) until flags[q] >= value for all q != my_rank; on timeout sets *err = 1.
cudaError_t launch_wait_flags(const uint32_t* my_flags, int world, int my_rank, uint32_t value,
                              unsigned long long timeout_ns, uint32_t* err, cudaStream_t s);

// signal + wait of one flag round in a single launch
cudaError_t launch_signal_wait(const PeerFlags& f, const uint32_t* my_flags, int world, int my_rank, uint32_t value,
                               unsigned long long timeout_ns, uint32_t* err, cudaStream_t s);

// ---- pair-symmetric force pass (gs_symmetric.cu; the default) --------------------------------------------------
struct SymCta { int row0, k0, count, pad; };   // a persistent CTA's first unit (row, index in the row) and unit count
constexpr float kSymPosScale = 1.0f / 1024.0f;              // tracked cutoff: positions scaled by 2^-10 ...
constexpr float kSymOutScale = 1.0f / 1048576.0f;           // ... make every sum 2^20 too large (mu t d -> mu (2^30 t)(2^-10 d))
struct SymParams {
    const float4* pos;       // packed pairs, n_pad bodies (every rank's); kCutoffTracked: the copy scaled by kSymPosScale
    float4* acc_whole;       // i-side sums of rows handled by a single CTA (float4 per local body)
    float4* partial;         // [grid][2][i_tile]: i-side sums of rows split between CTAs
    long long* accj[kMaxPeers + 1];   // per rank: j-side sums of ITS bodies, 3 x kFxLimbs words of wide fixed point per
                                      // local body (gs_device.cuh, fx_add), stored per j-tile as [component][limb][body];
                                      // mine is local memory, the others NVLink peer memory; exact integer atomics,
                                      // all zero before the pass
    const SymCta* ctas;      // [grid]
    int n, i_begin;          // real bodies; first body of this rank's slice
    int ni_tiles;            // rows = i-tiles of this rank's slice
    int slice_tiles;         // j-tiles per rank slice (= ni_tiles * diag)
    int diag;                // i_tile / j_tile: that many plain units open every row
    int rank, world;
    int full_slices;         // floor((world-1)/2): slices rank+1 .. rank+full_slices (mod world) are taken whole
    int half_mode, half_rows;   // world even: the opposite slice is shared with its owner --
                                //   1: my rows < half_rows take all of its tiles; 2: all my rows take its tiles from
                                //   half_rows*diag on; 0: world odd or 1
    float eps2, cut2, mu_uniform;
};
// Units of row I of `rank`: own tiles from the row's own on, then the whole slices, then the shared opposite slice.
GS_HD int sym_row_units(const SymParams& p, int row) {
    int n = p.slice_tiles - p.diag * row + p.full_slices * p.slice_tiles;
    if (p.half_mode == 1) n += (row < p.half_rows) ? p.slice_tiles : 0;
    else if (p.half_mode == 2) n += p.slice_tiles - p.half_rows * p.diag;
    return n;
}
// Global j-tile of unit k of row `row`.
GS_HD int sym_unit_tile(const SymParams& p, int row, int k) {
    const int own = p.slice_tiles - p.diag * row;
    if (k < own) return p.rank * p.slice_tiles + p.diag * row + k;
    k -= own;
    if (k < p.full_slices * p.slice_tiles) {
        const int d = k / p.slice_tiles + 1;
        return ((p.rank + d) % p.world) * p.slice_tiles + (k - (d - 1) * p.slice_tiles);
    }
    k -= p.full_slices * p.slice_tiles;
    const int opposite = (p.rank + p.world / 2) % p.world;
    return opposite * p.slice_tiles + (p.half_mode == 2 ? p.half_rows * p.diag : 0) + k;
}
bool sym_supported(int R, int threads);     // is the pass built for this (i-bodies per thread, CTA size)?
size_t sym_smem_bytes(int threads);
// cut_mode: CutMode (kCutoffTracked = select-free pass + exact redo of the units where two bodies coincide)
cudaError_t sym_kernel_setup(int R, int threads, int cut_mode, int* ctas_per_sm, int* regs);
cudaError_t launch_sym_force(int R, int threads, int cut_mode, const SymParams& p, int grid, cudaStream_t s);
cudaError_t launch_sym_finish(const StepParams& p, const SymParams& sp, const int2* rows, int i_tile,
                              const FinishExtras* extras, cudaStream_t s);

// AoS <-> packed-pair conversion on the device. `aos` is the raw byte image of the reference's
// 28-byte (f32) or 56-byte (f64) records.  Bodies >= n are padding (mu = 0, at kPadPosition).
// Covers bodies [body_first, body_first + body_count) (even bounds); the record of body i is record (i - rec_first).
// Packed positions are stored into every buffer of `dst`: this GPU's two position buffers for a full upload, both
// buffers of EVERY GPU (NVLink peer memory) when a rank uploads only its slice.
// pos_lo_local / vel_lo_local (nullable): low-order words, i.e. what rounding the record to fp32 dropped.
// mu_minmax: two words, preset to {0xffffffff, 0}: min and max bit pattern of mu over the real bodies.
struct UploadTargets { int n; float4* pos[2 * (kMaxPeers + 1)]; };
cudaError_t launch_aos_to_packed(const void* aos, int record_bytes, int rec_first, int n, int body_first,
                                 int body_count, double G, const UploadTargets& dst, float4* vel_local, double* mass,
                                 int i_begin, int n_local_pad, float4* pos_lo_local, float4* vel_lo_local,
                                 uint32_t* mu_minmax, cudaStream_t s);
cudaError_t launch_pack_minmax(uint32_t* mm4, cudaStream_t s);
// dst = src with x, y, z multiplied by `scale` (a power of two: exact), mu unchanged; n_pad bodies.
cudaError_t launch_scale_positions(const float4* src, float4* dst, int n_pad, float scale, cudaStream_t s);      // mm4[2] = mm4[0], mm4[3] = ~mm4[1]
// Records of bodies [body_first, body_first + count) into aos[0..count): pos and mass are indexed globally, vel and
// the (nullable) low-order words with (body - vel_first).
cudaError_t launch_packed_to_aos(const float4* pos, const float4* vel, int vel_first, const double* mass,
                                 int body_first, int count, void* aos, int record_bytes, const float4* pos_lo,
                                 const float4* vel_lo, cudaStream_t s);

// Energy diagnostics (fp64 accumulation).  potential: out_partials[gridDim] doubles, one per
// CTA, of sum_i mu_i * sum_{j != i} mu_j / sqrt(r^2 + eps^2) over this rank's i-bodies (the caller
// sums them and multiplies by -1/(2G)).  kinetic: out4[grid*4] = {sum m v^2, sum m vx, m vy, m vz}.
int energy_grid(int n_local_pad);
cudaError_t launch_potential(const float4* pos, int n_pad, int i_begin, int n_local_pad,
                             float eps2, float cut2, double* out_partials, cudaStream_t s);
cudaError_t launch_kinetic(const float4* vel_local, const double* mass, int i_begin, int n_local_pad,
                           double* out4, cudaStream_t s);

// One sample {step, time, KE, PE, px, py, pz, 0} (this rank's share) from the partials, reduced on the device.
cudaError_t launch_energy_sample(const double* pot_partials, const double* kin4, int egrid, double G, double step,
                                 double time, double* out8, cudaStream_t s);

}  // namespace gs
