/*
 * gravsim.h -- C-ABI of libgravsim.so, the B200-native direct all-pairs gravity engine.
 *
 * This is the drop-in boundary for the hot path of
 * pdpatel13/Gravity-Simulator-using-MPI-Spark-and-CUDA: force evaluation + integration + the
 * per-step position exchange.  The reference has no plugin/FFI layer (force and update are
 * free functions called from main() in the same translation unit), so each entry point cites
 * the reference statements it replaces.  Plain pointers and sizes only; no C++ or torch types.
 *
 * Conventions
 *   - every call returns GS_OK (0) or a negative GS_E* code, never aborts or throws;
 *     gs_strerror() names the code, gs_last_cuda_error() gives the cudaError_t behind GS_ECUDA.
 *   - the library owns all device memory, streams and NCCL communicators behind gs_ctx;
 *     host buffers stay caller-owned and are only touched during the call.
 *   - a gs_ctx is used from one host thread at a time; one ctx drives one GPU.  With world_size > 1 every rank
 *     needs its own host thread or process, as the ranks of mpirun have: uploads, downloads and gs_energy
 *     synchronise while they wait for the peers, so one thread driving several ranks in turn would wait for itself
 *     (and get GS_ENCCL after the peer timeout).
 *   - there is NO CPU fallback: without a usable CUDA device gs_create fails with GS_ENOGPU.
 *
 * Body records accepted/produced (the reference's own structs, byte for byte):
 *   f32: 7 floats  {pos[3], vel[3], mass} = 28 B   (cuda.cu:14-29 `struct Particle`)
 *   f64: 7 doubles {pos, vel, mass}       = 56 B   (mpi.c:11-21 `Vector3`/`Particle`)
 */
#ifndef GRAVSIM_H
#define GRAVSIM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GS_OK        0
#define GS_EINVAL   -1   /* bad argument / config                                  */
#define GS_ECUDA    -2   /* a CUDA runtime call failed; see gs_last_cuda_error()   */
#define GS_ENOMEM   -3   /* host or device allocation failed                       */
#define GS_ENCCL    -4   /* NCCL missing, a collective failed, or a peer timed out */
#define GS_ESTATE   -5   /* call out of order (e.g. step before upload)            */
#define GS_ENOGPU   -6   /* no CUDA device: this library has no CPU path           */
#define GS_EIO      -7   /* state-file I/O failed                                  */
#define GS_ETOOBIG  -8   /* problem larger than a fixed-size buffer allows         */

#define GS_G_SI 6.67430e-11          /* cuda.cu:11, mpi.c:9, pyspark.py:46 */
#define GS_NCCL_ID_BYTES 128

typedef struct gs_ctx gs_ctx;

enum { GS_INTEGRATOR_EULER = 0,      /* v += a dt; x += v dt (new v): cuda.cu:73-76, mpi.c:206-215 */
       GS_INTEGRATOR_LEAPFROG = 1 }; /* kick-drift-kick, 1 force pass per step (extension)         */

/* Replaces the compile-time constants of cuda.cu:11,121-123,155 and mpi.c:9,146-148. */
typedef struct gs_config {
    uint32_t struct_size;   /* = sizeof(gs_config); set by gs_config_init                       */
    int32_t  n;             /* number of bodies (cuda.cu:121; mpi.c:107)                         */
    double   dt;            /* time step (cuda.cu:123, mpi.c:148); default 3600                  */
    double   G;             /* gravitational constant; default GS_G_SI                           */
    double   eps;           /* Plummer softening length. 0 = the reference law: bare Newton,
                               pairs with r < 1e-10 contribute nothing (cuda.cu:39, mpi.c:64)    */
    int32_t  integrator;    /* GS_INTEGRATOR_*                                                   */
    int32_t  device;        /* CUDA device ordinal for this ctx; -1 = current device             */
    int32_t  rank;          /* this ctx's slice of the i-bodies, as mpi.c:184-187                */
    int32_t  world_size;    /* number of slices (GPUs); 1 = no exchange                          */
    const void *nccl_id;    /* world_size > 1: GS_NCCL_ID_BYTES from gs_nccl_unique_id(), same on
                               every rank (the caller broadcasts it; replaces MPI_Init mpi.c:142) */
    void    *stream;        /* cudaStream_t to run on; NULL = library-owned stream               */
    uint32_t flags;         /* GS_FLAG_*                                                         */
    int32_t  variant;       /* kernel variant; 0 = default (see gs_info.variant)                 */
} gs_config;

#define GS_FLAG_TWO_PASS   2u   /* one GPU: run the multi-GPU local-then-remote two-pass schedule  */
#define GS_FLAG_NO_UNIFORM_MASS 4u /* keep the general kernel even when all masses are equal        */
#define GS_FLAG_OVERLAP    8u   /* multi-GPU: own j-slice first, remote slices after the exchange  */
#define GS_FLAG_NCCL_EXCHANGE 16u /* multi-GPU: per-step ncclAllGather instead of peer-memory stores */
#define GS_FLAG_DETERMINISTIC 128u /* use the row-gather kernel (every ORDERED pair like mpi.c:196-204, no atomics of
                                   any kind, per-body sums in a fixed order) instead of the default pair-symmetric
                                   pass.  0.7x the speed.  The default pass is bit-reproducible as well (its
                                   cross-CTA / cross-GPU sums are exact integer atomics), but only for a fixed
                                   configuration (N, GPU count); its rounding differs from the row-gather kernel's */
#define GS_FLAG_COMPENSATED 32u /* positions and velocities as hi+lo fp32 pairs: the integration no longer
                                   rounds to the state's fp32 ulp each step (forces stay fp32); f64 records
                                   keep their low-order bits through upload/download                    */

#define GS_FLAG_FP64_FORCES 256u /* accuracy mode: every ordered pair evaluated and summed in fp64, the precision of
                                   mpi.c:59-73 (positions are the fp32 state, or its hi+lo pairs with
                                   GS_FLAG_COMPENSATED on one GPU); accelerations reach the integrator as hi+lo fp32
                                   pairs.  ~8x slower than the default pass; with GS_FLAG_COMPENSATED trajectories
                                   follow mpi.c's fp64 ones to ~1e-12                                            */

typedef struct gs_info {
    int32_t n, n_pad;             /* bodies, padded bodies (multiple of the tile sizes)          */
    int32_t i_begin, i_end;       /* this rank's i-bodies [i_begin, i_end) clipped to n          */
    int32_t sm_count, grid;       /* SMs and persistent CTAs of the force kernel                 */
    int32_t i_tile, j_tile;       /* bodies per i-tile (CTA) and per j-tile (smem stage)         */
    int32_t stages, variant;      /* TMA pipeline depth; kernel variant actually used            */
    int32_t regs_per_thread;      /* of the force kernel (cudaFuncGetAttributes)                 */
    int32_t threads;              /* threads per CTA                                             */
    int64_t scratch_bytes;        /* device scratch for split-tile partial sums                  */
    int32_t peer_exchange;        /* 1: positions travel by peer-memory stores; 0: ncclAllGather  */
    int32_t uniform_mass;         /* 1: hoisted-mass kernel in use (all masses equal)            */
    int32_t symmetric;            /* 1: pair-symmetric force pass in use                         */
    int32_t reserved;
} gs_info;

/* ---- life cycle -------------------------------------------------------------------------- */

void gs_config_init(gs_config *cfg);   /* fills defaults: dt 3600, G SI, eps 0, Euler, 1 GPU */

/* Replaces cudaMalloc x2 (cuda.cu:145-149) and MPI init + partition (mpi.c:142-144,184-187).
 * Allocates the packed float4 position (x2, double-buffered) and velocity arrays on the GPU,
 * the split-tile scratch, streams/events and, when world_size > 1, the NCCL communicator. */
int gs_create(gs_ctx **out, const gs_config *cfg);
void gs_destroy(gs_ctx *ctx);          /* reference frees nothing on the GPU (cuda.cu:175-177) */

/* rank 0 calls this and ships the bytes to the other ranks (replaces mpirun's wire-up). */
int gs_nccl_unique_id(void *out_id /* GS_NCCL_ID_BYTES */);

/* ---- data in / out ----------------------------------------------------------------------- */

/* Replace the H2D copy cuda.cu:151 and the MPI_Bcast of the particle array mpi.c:182.  Every
 * rank passes the FULL array of n records; AoS -> device layout is done on the GPU. */
int gs_upload_aos_f32(gs_ctx *ctx, const void *particles28, int n);
int gs_upload_aos_f64(gs_ctx *ctx, const void *particles56, int n);

/* Replace the final-state read-out (cuda.cu:173 via host vector; mpi.c:250-257 after the last
 * MPI_Allgatherv).  Every rank receives the FULL array (collective when world_size > 1).
 * With GS_INTEGRATOR_LEAPFROG velocities are synchronised to the positions' time first. */
int gs_download_aos_f32(gs_ctx *ctx, void *particles28, int n);
int gs_download_aos_f64(gs_ctx *ctx, void *particles56, int n);

/* The same for ONE rank's own records: rank r passes / receives exactly the records of its i-bodies,
 * first = gs_info.i_begin, count = gs_info.i_end - gs_info.i_begin (else GS_EINVAL) -- the slice mpi.c:184-187
 * assigns to it and mpi.c:196-216 updates.  Collective like the full-array calls when world_size > 1, but each GPU
 * moves n/world_size records over PCIe; the packed positions reach the other GPUs by NVLink stores (this replaces the
 * MPI_Bcast of mpi.c:182 by "every rank contributes its part", as MPI_Allgatherv does at mpi.c:227-231).  A download
 * of a slice touches no other GPU. */
int gs_upload_slice_f32(gs_ctx *ctx, const void *my_particles28, int first, int count);
int gs_upload_slice_f64(gs_ctx *ctx, const void *my_particles56, int first, int count);
int gs_download_slice_f32(gs_ctx *ctx, void *my_particles28, int first, int count);
int gs_download_slice_f64(gs_ctx *ctx, void *my_particles56, int first, int count);

/* ---- the hot path ------------------------------------------------------------------------ */

/* Replaces the step loops cuda.cu:155-163 (kernel, sync, D2H, host update) and mpi.c:191-237
 * (force rows, in-place update, MPI_Allgatherv, MPI_Barrier).  By default every unordered pair is
 * evaluated once and applied to both bodies (cuda.cu:43-49, without its race, and bit-reproducible;
 * GS_FLAG_DETERMINISTIC evaluates ordered pairs like mpi.c:196-204).  All bodies advance from the
 * same snapshot (synchronous step; see DESIGN.md on mpi.c's in-place order).  Asynchronous:
 * returns after enqueueing; no per-step host synchronisation or D2H. */
int gs_step(gs_ctx *ctx, int nsteps);

/* gs_step that also records the positions of all n bodies every `every` steps (after steps every, 2*every,
 * ...): xyz_host receives nsteps/every frames of n x 3 floats (at most capacity_snapshots, else GS_EINVAL).
 * Takes the place of the per-step `trajectories.append(positions)` of pyspark.py:104-121 (cuda.cu and mpi.c
 * keep final positions only).  Frames leave through a copy stream and pinned staging while the following
 * steps compute; the call returns when every frame is in xyz_host. */
int gs_step_trajectory(gs_ctx *ctx, int nsteps, int every, float *xyz_host, int capacity_snapshots,
                       int *n_snapshots);

/* Accelerations of THIS rank's bodies [i_begin, i_end) on the current frozen snapshot, 3 floats
 * per body; state unchanged.  The reference only exposes forces[] via the D2H at cuda.cu:160
 * and F/m at cuda.cu:70, mpi.c:206; this is the hook the <=1e-5 parity check consumes. */
int gs_accel(gs_ctx *ctx, float *acc3_host);
/* The same as doubles; beyond fp32 only with GS_FLAG_FP64_FORCES (hi + lo words), where it reproduces
 * sum_j calculate_force()/m_i (mpi.c:199-206) to fp64 rounding. */
int gs_accel_f64(gs_ctx *ctx, double *acc3_host);

/* Kinetic, potential energy and linear momentum (fp64 accumulation).  Not in the reference.
 * Collective when world_size > 1 (every rank gets the global values). */
int gs_energy(gs_ctx *ctx, double *kinetic, double *potential, double momentum[3]);

/* On-device energy / momentum monitor (not in the reference, which computes no energies).
 * After gs_monitor_start, gs_step() takes a sample after every `every`-th step: {step, time, KE, PE, px, py, pz, 0}
 * (8 doubles), computed, reduced (over the GPU and, with ncclAllReduce, over the ranks) and stored entirely on the
 * device -- no host synchronisation, no host arithmetic, the step loop stays asynchronous.  At most `capacity` samples
 * are kept (later ones are dropped); an upload clears them; every = 0 switches the monitor off.  gs_monitor_read
 * synchronises and copies the samples taken so far.  A sample costs one O(N^2) potential pass, and with the leapfrog a
 * velocity synchronisation (one extra force pass) -- choose `every` accordingly.  Collective when world_size > 1. */
int gs_monitor_start(gs_ctx *ctx, int every, int capacity);
int gs_monitor_read(gs_ctx *ctx, double *samples8, int capacity, int *count);

/* Potential energy WITHOUT any allocation and without a gs_ctx.  (No counterpart in the reference:
 * it computes no energies and allocates as it pleases, cuda.cu:145-149,159.)
 * For callers that may not call cudaMalloc -- a hook inside a loop whose device memory comes from a
 * pool or a caching allocator, a latency-critical loop that cannot afford cudaMalloc's device-wide
 * synchronisation, a process at its memory limit.  The kernel's per-CTA partial sums go to a fixed
 * 128 KiB __device__ array inside the library (present on every GPU that loads it, until exit)
 * instead of a cudaMalloc'ed buffer, hence:
 *   - n_pad <= GS_NOALLOC_MAX_BODIES, else GS_ETOOBIG and nothing is launched;
 *   - every call on a GPU shares that array: calls that can overlap on the GPU (two streams, two
 *     threads) race and must be serialised by the caller; stale contents are never read;
 *   - the first call on each device resolves the array's address; make it before any stream capture;
 *   - the array used is the CURRENT device's copy: pos_packed_dev, potential_dev and `stream` must belong to
 *     the current device (cudaSetDevice first); pointers of another device are refused with GS_EINVAL.
 * pos_packed_dev: n_pad bodies in the packed-pair layout of DESIGN.md section 3, caller-owned device
 * memory ({x0,x1,y0,y1},{z0,z1,mu0,mu1} per pair, mu = (float)(G*mass); bodies beyond the real ones
 * all-zero); n_pad a multiple of 512.  potential_dev: one double in device memory, written on
 * `stream` (a cudaStream_t, NULL = default stream); same value as gs_energy()'s potential. */
#define GS_NOALLOC_MAX_BODIES 4194304
int gs_potential_energy_noalloc(const void *pos_packed_dev, int n_pad, double G, double eps,
                                double *potential_dev, void *stream);

int gs_sync(gs_ctx *ctx);              /* cudaDeviceSynchronize at cuda.cu:157, per ctx */
int gs_set_dt(gs_ctx *ctx, double dt);

/* ---- introspection ----------------------------------------------------------------------- */

int gs_get_info(const gs_ctx *ctx, gs_info *info);
uint64_t gs_kernel_launches(const gs_ctx *ctx);    /* kernels this ctx has launched so far    */
int gs_last_cuda_error(const gs_ctx *ctx);         /* cudaError_t of the last GS_ECUDA        */
const char *gs_strerror(int code);
const char *gs_version(void);

/* Pinned host buffers for callers that want full-speed uploads (cudaHostAlloc). */
int gs_host_alloc(void **ptr, size_t bytes);
int gs_host_free(void *ptr);

/* ---- work partition (host logic, usable without a GPU; see DESIGN.md "stream-K") ---------- */

/* i-body slice of `rank` out of `world_size` over n_pad padded bodies: the block partition of
 * mpi.c:184-187 on tile granularity. */
int gs_plan_slice(int n, int world_size, int rank, int variant, int *n_pad, int *i_begin_pad, int *i_end_pad);
/* Unit range [u0,u1) of persistent CTA `cta` out of `grid` over ni_tiles x nj_tiles work units. */
int gs_plan_cta_range(int64_t ni_tiles, int64_t nj_tiles, int grid, int cta, int64_t *u0, int64_t *u1);
/* For i-tile `it`: first/last CTA touching it and whether it is split across CTAs. */
int gs_plan_tile_owners(int64_t ni_tiles, int64_t nj_tiles, int grid, int64_t it,
                        int *cta_first, int *cta_last, int *is_split);
int gs_plan_tiles(int variant, int *i_tile, int *j_tile);
/* Pair-symmetric pass: rows of `rank` (one per i-tile), units of a row (the first n_plain are
 * one-directional units on the row's own j-tiles), the global j-tile of a unit, and the unit range of a
 * persistent CTA.  Every unordered pair of bodies is covered exactly once over all ranks. */
int gs_plan_sym_row(int n, int world_size, int rank, int variant, int row, int *n_rows, int *n_units, int *n_plain);
int gs_plan_sym_unit(int n, int world_size, int rank, int variant, int row, int k, int *j_tile);
int gs_plan_sym_cta(int n, int world_size, int rank, int variant, int grid, int cta, int *row0, int *k0, int *count);

/* ---- host utilities: initial conditions, state files, reference-format dumps -------------- */
/* (plain C, no GPU needed) */

/* Sun/Earth/Mars (cuda.cu:82-93 == mpi.c:76-94) followed by n-3 bodies with uniform position in
 * [-3e11,3e11]^3 m, velocity in [-3e4,3e4]^3 m/s and mass in [1e23,1e25] kg (cuda.cu:129-131,
 * mpi.c:98-104), from a SEEDED counter-based generator (the reference seeds from entropy/clock:
 * cuda.cu:127, mpi.c:96).  Values are fp32-representable so f32 and f64 records agree exactly. */
int gs_ic_reference_uniform_f32(void *particles28, int n, uint64_t seed);
int gs_ic_reference_uniform_f64(void *particles56, int n, uint64_t seed);
/* Plummer sphere in N-body units (G = 1, total mass 1, scale radius 1, equal masses 1/n,
 * radius truncated at rmax scale radii); Aarseth-Henon-Wielen sampling.  Not in the reference. */
int gs_ic_plummer_f32(void *particles28, int n, uint64_t seed, double rmax);
/* Two Plummer systems of n/2 bodies each on an offset approach orbit: separation `sep`,
 * impact parameter `impact`, relative speed `vrel` (N-body units, total mass 1). */
int gs_ic_two_galaxy_f32(void *particles28, int n, uint64_t seed, double sep, double impact, double vrel);

/* Binary state file: 64-byte header (magic "GSSTATE1", n, record bytes 28|56, step, time, dt, G,
 * eps) followed by n reference-layout records.  Doubles as checkpoint/resume. */
typedef struct gs_state_header {
    char     magic[8];
    int32_t  n;
    int32_t  record_bytes;
    int64_t  step;
    double   time, dt, G, eps;
    uint8_t  reserved[8];
} gs_state_header;
int gs_state_write(const char *path, const gs_state_header *hdr, const void *records);
int gs_state_read_header(const char *path, gs_state_header *hdr);
int gs_state_read(const char *path, gs_state_header *hdr, void *records, size_t capacity_bytes);

/* Final-position dumps in the reference's two text formats. `stream` is a FILE*.
 *   cuda: "Particle i: (x, y, z)" with std::fixed << setprecision(14)   (cuda.cu:99-110)
 *   mpi : "Particle i: (%e, %e, %e)"                                     (mpi.c:249-257) */
int gs_dump_positions_cuda_format(void *stream, const void *particles28, int n, int print_particles);
int gs_dump_positions_mpi_format(void *stream, const void *particles56, int n);

#ifdef __cplusplus
}
#endif
#endif /* GRAVSIM_H */
