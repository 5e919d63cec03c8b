// gs_api.cu -- the C-ABI of libgravsim.so (include/gravsim.h): context, streams, step scheduling,
// multi-GPU exchange.  Every entry point returns an error code; nothing throws across the boundary.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>
#include <unistd.h>

#include "../../include/gravsim.h"
#include "gs_kernels.h"
#include "gs_nccl_dyn.h"

using namespace gs;

namespace gs {
void build_sym_plan(const SymParams& sp, int grid, int cost_plain, int cost_sym, std::vector<SymCta>& ctas,
                    std::vector<int2>& rows);
}

struct gs_ctx {
    gs_config cfg;
    int device = 0;
    int n = 0, n_pad = 0, slice = 0, i_begin = 0;   // slice = padded bodies per rank
    int variant = 0, R = 0, threads = 0, i_tile = 0, cut_mode = 0;
    float mu_uniform = 0.f;       // != 0: all real bodies share this G*m (hoisted-mass kernel)
    int sm_count = 0, ctas_per_sm = 0, grid = 0, regs = 0;
    float eps2 = 0.f, cut2 = 0.f;
    float4* pos[2] = {nullptr, nullptr};
    int cur = 0;
    float4* vel = nullptr;        // this rank's slice
    float4* vel_full = nullptr;   // world_size > 1: gathered velocities for download (lazy)
    float4* pos_lo = nullptr;     // GS_FLAG_COMPENSATED: low-order words of this rank's positions / velocities
    float4* vel_lo = nullptr;
    float4* lo_full[2] = {nullptr, nullptr};   // ... gathered for download when world_size > 1 (lazy)
    float4* acc_local = nullptr;  // world_size > 1: accelerations from the local j-slice
    float4* acc_stage = nullptr;  // peer-memory mode: accelerations handed from the force kernel to the finish kernel
    float4* acc_lo = nullptr;     // GS_FLAG_FP64_FORCES: low-order words of the accelerations
    float4* acc_out = nullptr;    // gs_accel output (lazy)
    float4* partial = nullptr;    // split-tile scratch
    size_t partial_bytes = 0;
    double* mass = nullptr;       // n_pad doubles: original masses (returned bit-exactly by downloads)
    bool mass_full = true;        // false after a slice upload: only this rank's part of mass[] is valid
    uint32_t* d_mu_mm = nullptr;  // {min, max} bit pattern of mu over the uploaded bodies (equal-mass detection on the device)
    uint32_t* h_mu_mm = nullptr;  // pinned mirror (2 words) + 2 words of presets
    void* d_aos = nullptr;        // staging for the reference-layout records
    size_t d_aos_bytes = 0;
    double* d_energy = nullptr;   // per-CTA partials: [egrid] potential + [egrid*4] kinetic/momentum + 8
    double* h_energy = nullptr;   // pinned mirror
    float4* h_acc = nullptr;      // pinned staging for gs_accel
    size_t h_acc_bytes = 0;
    cudaStream_t stream = nullptr, comm_stream = nullptr;
    bool own_stream = false;
    cudaEvent_t ev_integ = nullptr, ev_gather[2] = {nullptr, nullptr};
    ncclComm_t comm = nullptr;
    bool uploaded = false;
    bool vel_synced = true;       // leapfrog: velocities at the same time as positions
    uint64_t launches = 0;
    int last_cuda_error = 0;
    int64_t steps_done = 0;
    double sim_time = 0.0;        // sum of the dt of the steps since the last upload
    // on-device energy / momentum monitor (gs_monitor_start): samples stay in device memory until read
    double* d_log = nullptr;
    int mon_every = 0, mon_capacity = 0, mon_count = 0;
    bool split_full = false, split_local = false, split_remote = false;   // does a pass need the fix-up kernel
    // peer-memory exchange (world_size > 1, all peers mappable): see setup_peer_exchange()
    bool p2p = false;
    float4* peer_pos[2][kMaxPeers + 1] = {};   // [buffer][rank]: every rank's pos buffers as seen from this GPU
    uint32_t* flags = nullptr;                 // my flag array: flags[q] = exchanges rank q has completed into my buffers
    uint32_t* peer_flags[kMaxPeers + 1] = {};  // rank q's flag array as seen from this GPU
    void* ipc_opened[4 * (kMaxPeers + 1)] = {};
    int n_ipc_opened = 0;
    uint32_t* d_comm_err = nullptr;            // set by the wait kernel on timeout
    uint32_t xstep = 0;                        // exchanges issued so far (same on every rank)
    bool wait_pending = false;                 // pos[cur] still needs the peers' flags before remote bodies are read
    unsigned long long comm_timeout_ns = 20000000000ull;
    // trajectory recording (gs_step_trajectory): snapshots leave on a copy stream while the next steps run
    cudaStream_t copy_stream = nullptr;
    float4* h_snap[2] = {nullptr, nullptr};
    cudaEvent_t ev_snap_ready[2] = {nullptr, nullptr}, ev_snap_done[2] = {nullptr, nullptr};
    bool buf_copy_pending[2] = {false, false};   // pos[b] is still being read by a snapshot copy
    // pair-symmetric pass (the default)
    bool symmetric = false;
    int sym_grid = 0, sym_regs = 0, sym_cut_mode = 0;
    SymCta* d_sym_ctas = nullptr;
    int2* d_sym_rows = nullptr;
    float4* pos_scaled = nullptr;               // tracked cutoff (eps == 0): positions x 2^-10, rewritten every pass
    long long* d_accj = nullptr;                // j-side sums of MY bodies, wide fixed point (peers add into it over NVLink)
    long long* peer_accj[kMaxPeers + 1] = {};   // every rank's accj as seen from this GPU
    float4* sym_partial = nullptr;
    bool want_sym = false, p2p_atomics = false;
    uint32_t fstep = 0;                         // symmetric passes whose remote atomics were fenced so far
    int buf_copy_event[2] = {0, 0};
};

namespace {

#define GS_CUDA(ctx, call)                                  \
    do {                                                    \
        cudaError_t e__ = (call);                           \
        if (e__ != cudaSuccess) {                           \
            (ctx)->last_cuda_error = (int)e__;              \
            return GS_ECUDA;                                \
        }                                                   \
    } while (0)


bool any_split(int64_t ni, int64_t nj, int grid) {
    for (int64_t it = 0; it < ni; ++it)
        if (tile_owners(ni, nj, grid, it).split) return true;
    return false;
}

int ensure_aos(gs_ctx* c, size_t bytes) {
    if (c->d_aos_bytes >= bytes) return GS_OK;
    if (c->d_aos) cudaFree(c->d_aos);
    c->d_aos = nullptr; c->d_aos_bytes = 0;
    GS_CUDA(c, cudaMalloc(&c->d_aos, bytes));
    c->d_aos_bytes = bytes;
    return GS_OK;
}

// Geometry of the pair-symmetric pass for this rank (see SymParams): own triangle, the next
// floor((P-1)/2) slices whole, and for even P the opposite slice shared half/half with its owner.
void fill_sym_geometry(const gs_ctx* c, SymParams* sp) {
    memset(sp, 0, sizeof(*sp));
    const int P = c->cfg.world_size, r = c->cfg.rank;
    sp->n = c->n; sp->i_begin = c->i_begin;
    sp->ni_tiles = c->slice / c->i_tile;
    sp->slice_tiles = c->slice / kJTile;
    sp->diag = c->i_tile / kJTile;
    sp->rank = r; sp->world = P;
    sp->full_slices = (P - 1) / 2;
    sp->half_rows = sp->ni_tiles / 2;
    sp->half_mode = (P > 1 && P % 2 == 0) ? (r < P / 2 ? 1 : 2) : 0;
}

// Orders "every rank's bodies of pos[cur] are in this GPU's memory" before whatever is enqueued next on
// the engine's stream.
int wait_positions(gs_ctx* c) {
    if (c->cfg.world_size <= 1) return GS_OK;
    if (c->p2p) {
        if (!c->wait_pending) return GS_OK;
        GS_CUDA(c, launch_wait_flags(c->flags, c->cfg.world_size, c->cfg.rank, c->xstep, c->comm_timeout_ns,
                                     c->d_comm_err, c->stream));
        c->launches++;
        c->wait_pending = false;
        return GS_OK;
    }
    GS_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_gather[c->cur], 0));
    return GS_OK;
}

// One force pass over this rank's i-bodies on the snapshot pos[cur].
// world_size > 1 and overlap enabled: local j-slice first (needs nothing from the network), then
// the remote slices once the all-gather of pos[cur] has landed.
int force_pass(gs_ctx* c, int final_mode, float kick, float drift, float4* acc_out) {
    StepParams p;
    memset(&p, 0, sizeof(p));
    p.pos_old = c->pos[c->cur];
    p.pos_new = c->pos[c->cur ^ 1];
    p.vel = c->vel;
    p.partial = c->partial;
    p.i_begin = c->i_begin;
    p.n = c->n;
    p.ni_tiles = c->slice / c->i_tile;
    p.nj_total = c->n_pad / kJTile;
    p.eps2 = c->eps2;
    p.cut2 = c->cut2;
    p.kick_dt = kick;
    p.drift_dt = drift;
    p.mu_uniform = c->mu_uniform;
    const int W = c->cfg.world_size;
    FinishExtras extras;
    extras.peers.n = 0;
    extras.pos_lo = c->pos_lo;
    extras.vel_lo = c->vel_lo;
    extras.acc_lo = nullptr;
    if (c->p2p && final_mode == kFinalIntegrate) {   // fused exchange: the finish kernel stores into every peer's pos[nxt]
        for (int q = 0; q < W; ++q)
            if (q != c->cfg.rank) extras.peers.pos_new[extras.peers.n++] = c->peer_pos[c->cur ^ 1][q];
    }
    // two passes (own j-slice first, the rest once the peers' bodies are in): GS_FLAG_OVERLAP on several
    // GPUs, or GS_FLAG_TWO_PASS (same code path on one GPU, for tests)
    const bool two_pass = ((W > 1) && (c->cfg.flags & GS_FLAG_OVERLAP)) ||
                          ((W == 1) && (c->cfg.flags & GS_FLAG_TWO_PASS) && p.nj_total >= 2);
    int rc;
    // On several GPUs with the peer-memory exchange the integrating pass is finished by the finish kernel
    // for EVERY tile (integration + stores into all peers' buffers in one kernel): the force kernel then
    // only leaves accelerations in acc_stage.
    auto run = [&](StepParams q, bool need_fixup) -> int {
        const bool defer = (c->p2p && q.final_mode == kFinalIntegrate) ||
                           (c->pos_lo != nullptr && q.final_mode != kFinalAccOut);   // compensated update lives there too
        StepParams f = q;
        if (defer) { f.final_mode = kFinalAccOut; f.acc_out = c->acc_stage; }
        GS_CUDA(c, launch_force(c->variant, c->cut_mode, f, c->grid, c->stream)); c->launches++;
        if (need_fixup || defer) {
            GS_CUDA(c, launch_fixup(c->variant, q, c->grid, defer ? c->acc_stage : nullptr, defer ? &extras : nullptr,
                                    c->stream));
            c->launches++;
        }
        return GS_OK;
    };
    if (c->cfg.flags & GS_FLAG_FP64_FORCES) {
        // accuracy mode: every ordered pair in fp64 (mpi.c:59-73's precision), then the common finishing kernel
        if ((rc = wait_positions(c)) != GS_OK) return rc;
        p.nj_tiles = p.nj_total; p.j_rot = 0;
        p.acc_in = nullptr; p.acc_out = acc_out; p.final_mode = final_mode;
        const float4* lo_all = (W == 1) ? c->pos_lo : nullptr;     // low-order words of all bodies are only here on one GPU
        if (!c->mass_full) {     // after a slice upload every rank only holds its own masses
            if (nccl_api().AllGather(c->mass + c->i_begin, c->mass, (size_t)c->slice, ncclDouble, c->comm, c->stream) != ncclSuccess)
                return GS_ENCCL;
            c->mass_full = true;
        }
        GS_CUDA(c, launch_force_f64(c->pos[c->cur], lo_all, c->mass, c->cfg.G, c->n_pad, c->i_begin, c->slice, c->cfg.eps * c->cfg.eps,
                                    c->cfg.eps > 0.0 ? 0.0 : 1e-20, c->acc_stage, c->acc_lo, c->stream));
        extras.acc_lo = c->acc_lo;
        GS_CUDA(c, launch_finish_acc(p, c->slice, c->acc_stage, &extras, c->stream));
        c->launches += 2;
        return GS_OK;
    }
    if (!two_pass) {
        if ((rc = wait_positions(c)) != GS_OK) return rc;
        p.nj_tiles = p.nj_total; p.j_rot = 0;
        p.acc_in = nullptr; p.acc_out = acc_out; p.final_mode = final_mode;
        // every unordered pair once, both directions; finished by sym_finish_kernel.  On several GPUs only the
        // integrating pass (it is collective: peers add into my j-side sums and I into theirs).
        if (c->symmetric && (W == 1 || final_mode == kFinalIntegrate)) {
            SymParams sp;
            fill_sym_geometry(c, &sp);
            sp.pos = c->pos[c->cur]; sp.acc_whole = c->acc_stage; sp.partial = c->sym_partial;
            for (int q = 0; q < W; ++q) sp.accj[q] = c->peer_accj[q];
            sp.ctas = c->d_sym_ctas; sp.eps2 = c->eps2; sp.cut2 = c->cut2; sp.mu_uniform = c->mu_uniform;
            if (c->sym_cut_mode == kCutoffTracked) {
                // the tracked cutoff finds close pairs by overflow: the kernel reads a copy scaled by 2^-10
                // (all n_pad bodies, 32 B each: ~10 us at N = 1M); the kernel rescales its sums by kSymOutScale
                GS_CUDA(c, launch_scale_positions(c->pos[c->cur], c->pos_scaled, c->n_pad, kSymPosScale, c->stream));
                c->launches++;
                sp.pos = c->pos_scaled;
                sp.cut2 = c->cut2 * kSymPosScale * kSymPosScale;
            }
            GS_CUDA(c, launch_sym_force(c->R, c->threads, c->sym_cut_mode, sp, c->sym_grid, c->stream)); c->launches++;
            if (W > 1) {    // my j-side sums are complete once every peer's force kernel has finished
                PeerFlags f;
                for (int q = 0; q < W; ++q) f.slot[q] = c->peer_flags[q] + 16 + c->cfg.rank;
                c->fstep++;
                GS_CUDA(c, launch_signal_wait(f, c->flags + 16, W, c->cfg.rank, c->fstep, c->comm_timeout_ns, c->d_comm_err,
                                              c->stream));
                c->launches++;
            }
            GS_CUDA(c, launch_sym_finish(p, sp, c->d_sym_rows, c->i_tile, &extras, c->stream)); c->launches++;
            return GS_OK;
        }
        return run(p, c->split_full);
    }
    const int nj_local = (W > 1) ? c->slice / kJTile : p.nj_total / 2;
    const int rot_local = (W > 1) ? c->cfg.rank * nj_local : p.nj_total / 4;
    // pass L: j-tiles of this rank's own slice
    p.nj_tiles = nj_local; p.j_rot = rot_local;
    p.acc_in = nullptr; p.acc_out = c->acc_local; p.final_mode = kFinalAccOut;
    if ((rc = run(p, c->split_local)) != GS_OK) return rc;
    // pass R: everything else, after the exchange
    if ((rc = wait_positions(c)) != GS_OK) return rc;
    p.nj_tiles = p.nj_total - nj_local; p.j_rot = (rot_local + nj_local) % p.nj_total;
    p.acc_in = c->acc_local; p.acc_out = acc_out; p.final_mode = final_mode;
    return run(p, c->split_remote);
}

// Position exchange after an integrating pass wrote this rank's slice of pos[nxt]
// (replaces MPI_Allgatherv + MPI_Barrier, mpi.c:227-236; 16 B/body instead of 56).
//  peer-memory mode: the data already went out from the kernel's epilogue; only raise my counter on every peer.
//  NCCL mode (fallback): in-place ncclAllGather on the comm stream.
int exchange_positions(gs_ctx* c, int nxt) {
    if (c->cfg.world_size <= 1) return GS_OK;
    if (c->p2p) {
        PeerFlags f;
        for (int q = 0; q < c->cfg.world_size; ++q) f.slot[q] = c->peer_flags[q] + c->cfg.rank;
        c->xstep++;
        if (c->cfg.flags & GS_FLAG_OVERLAP) {      // own-slice work of the next pass may run before the wait
            GS_CUDA(c, launch_signal_peers(f, c->cfg.world_size, c->cfg.rank, c->xstep, c->stream));
            c->wait_pending = true;
        } else {                                   // one launch for both halves of the round
            GS_CUDA(c, launch_signal_wait(f, c->flags, c->cfg.world_size, c->cfg.rank, c->xstep, c->comm_timeout_ns,
                                          c->d_comm_err, c->stream));
            c->wait_pending = false;
        }
        c->launches++;
        return GS_OK;
    }
    NcclApi& nc = nccl_api();
    GS_CUDA(c, cudaEventRecord(c->ev_integ, c->stream));
    GS_CUDA(c, cudaStreamWaitEvent(c->comm_stream, c->ev_integ, 0));
    const float4* send = c->pos[nxt] + c->i_begin;
    ncclResult_t r = nc.AllGather(send, c->pos[nxt], (size_t)c->slice * 4, ncclFloat, c->comm, c->comm_stream);
    if (r != ncclSuccess) return GS_ENCCL;
    GS_CUDA(c, cudaEventRecord(c->ev_gather[nxt], c->comm_stream));
    return GS_OK;
}

// Leapfrog keeps velocities half a step behind/ahead of the positions between steps; bring them
// to the positions' time with a half kick from the current accelerations.
int sync_velocities(gs_ctx* c) {
    if (c->cfg.integrator != GS_INTEGRATOR_LEAPFROG || c->vel_synced) return GS_OK;
    int rc = force_pass(c, kFinalKickOnly, (float)(0.5 * c->cfg.dt), 0.f, nullptr);
    if (rc != GS_OK) return rc;
    c->vel_synced = true;
    return GS_OK;
}

// Measured on B200 at N = 1M (profiles/r1_variant_probe_n1048576.txt): softened -> R=6 x 256 threads (72.0 % general,
// 78.7 % equal masses), cutoff -> R=8 x 256 threads (65.2 %).  R=6 has a 1536-body i-tile, which pads some
// N / GPU-count combinations more than the 2048-body tile of R=8 does: charge each its padded work.
int default_variant(int n, int world, int cut_mode, uint32_t flags) {
    int pad6 = 0, pad3 = 0, pad1 = 0;
    if (gs_plan_slice(n, world, 0, 6, &pad6, nullptr, nullptr) != GS_OK ||
        gs_plan_slice(n, world, 0, 3, &pad3, nullptr, nullptr) != GS_OK ||
        gs_plan_slice(n, world, 0, 1, &pad1, nullptr, nullptr) != GS_OK)
        return kDefaultVariant;
    // small problems: too few (i-tile x j-tile) units to feed 148 SMs -> 512-body i-tiles, 4 CTAs per SM
    const int64_t units6 = ((int64_t)pad6 / world / 1536) * (pad6 / kJTile);
    if (units6 < 8 * 148) {
        // tiny problems: even 512-body i-tiles leave SMs idle -> 256-body i-tiles (R=2 x 128 threads)
        const int64_t units10 = ((int64_t)pad6 / world / 512) * (pad6 / kJTile);
        if (units10 < 4 * 148) return 16;
        // in between (N ~ 14K..30K per GPU-row band): the pair-symmetric pass on 1024-body i-tiles, 2 CTAs per SM
        // (profiles/r2_small_n_probe.txt: N = 16,384 58.6 % vs 54-56 % for the row-gather variants)
        const int64_t units1 = ((int64_t)pad1 / world / 1024) * (pad1 / kJTile);
        if (!(flags & GS_FLAG_DETERMINISTIC) && world <= kMaxPeers + 1 && units1 >= 400) return 1;
        return 10;
    }
    if (!(flags & GS_FLAG_DETERMINISTIC) && world <= kMaxPeers + 1) {
        // pair-symmetric pass (256 threads; measured at N = 1M, equal masses: R = 8 4.12e12, R = 6 3.88e12,
        // R = 4 3.66e12 interactions/s): the fastest after charging each its padded work
        const double cost3 = (double)pad3 * pad3, cost6 = (double)pad6 * pad6 / 0.94, cost1 = (double)pad1 * pad1 / 0.89;
        if (cost3 <= cost6 && cost3 <= cost1) return 3;
        return cost6 <= cost1 ? 6 : 1;
    }
    if (cut_mode != kSoftened) return 3;
    const double cost6 = (double)pad6 * pad6, cost3 = (double)pad3 * pad3 / 0.98;
    return cost6 <= cost3 ? 6 : 3;
}

// ---- peer-memory exchange set-up ----------------------------------------------------------------
// Every rank publishes IPC handles (other processes) and raw pointers (other threads of this process)
// of its two position buffers and its flag array; NCCL is only the courier here, once.  If any rank
// cannot map any peer, all ranks stay on the NCCL all-gather path together.
struct PeerBlob {
    long long pid;
    int device, ok;
    cudaIpcMemHandle_t handle[4];      // pos[0], pos[1], flags, j-side sums (pair-symmetric pass; may be absent)
    void* raw[4];
    char pad[512 - 16 - 4 * sizeof(cudaIpcMemHandle_t) - 4 * sizeof(void*)];
};
static_assert(sizeof(PeerBlob) == 512, "PeerBlob is exchanged as 512 raw bytes");

void close_peer_mappings(gs_ctx* c) {
    for (int k = 0; k < c->n_ipc_opened; ++k) cudaIpcCloseMemHandle(c->ipc_opened[k]);
    c->n_ipc_opened = 0;
}

int setup_peer_exchange(gs_ctx* c) {
    const int W = c->cfg.world_size, me = c->cfg.rank;
    if (W > kMaxPeers + 1 || (c->cfg.flags & GS_FLAG_NCCL_EXCHANGE)) return GS_OK;
    NcclApi& nc = nccl_api();
    if (const char* t = getenv("GRAVSIM_COMM_TIMEOUT_S")) c->comm_timeout_ns = (unsigned long long)(atof(t) * 1e9);
    GS_CUDA(c, cudaMalloc((void**)&c->flags, 64 * sizeof(uint32_t)));
    GS_CUDA(c, cudaMalloc((void**)&c->d_comm_err, sizeof(uint32_t)));
    GS_CUDA(c, cudaMemsetAsync(c->flags, 0, 64 * sizeof(uint32_t), c->stream));
    GS_CUDA(c, cudaMemsetAsync(c->d_comm_err, 0, sizeof(uint32_t), c->stream));
    PeerBlob mine;
    memset(&mine, 0, sizeof(mine));
    mine.pid = (long long)getpid();
    mine.device = c->device;
    mine.ok = 1;
    void* ptrs[4] = {c->pos[0], c->pos[1], c->flags, c->d_accj};
    const int nh = c->d_accj ? 4 : 3;
    for (int k = 0; k < nh; ++k) {
        mine.raw[k] = ptrs[k];
        if (cudaIpcGetMemHandle(&mine.handle[k], ptrs[k]) != cudaSuccess) { cudaGetLastError(); mine.ok = 0; }
    }
    PeerBlob* d_blobs = nullptr;
    PeerBlob* h_blobs = new (std::nothrow) PeerBlob[W];
    if (!h_blobs) return GS_ENOMEM;
    int* d_ok = nullptr;
    int rc = GS_OK, all_ok = 0;
    do {
        if (cudaMalloc((void**)&d_blobs, sizeof(PeerBlob) * W) != cudaSuccess ||
            cudaMalloc((void**)&d_ok, sizeof(int)) != cudaSuccess) { rc = GS_ENOMEM; break; }
        if (cudaMemcpyAsync(d_blobs + me, &mine, sizeof(mine), cudaMemcpyHostToDevice, c->stream) != cudaSuccess ||
            nc.AllGather(d_blobs + me, d_blobs, sizeof(PeerBlob), ncclChar, c->comm, c->stream) != ncclSuccess ||
            cudaMemcpyAsync(h_blobs, d_blobs, sizeof(PeerBlob) * W, cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
            cudaStreamSynchronize(c->stream) != cudaSuccess) { rc = GS_ENCCL; break; }
        int ok = 1, atomics = 1;
        for (int q = 0; q < W && ok; ++q) ok = h_blobs[q].ok;
        for (int q = 0; q < W && ok; ++q) {
            if (q == me) {
                c->peer_pos[0][q] = c->pos[0]; c->peer_pos[1][q] = c->pos[1]; c->peer_flags[q] = c->flags;
                continue;
            }
            void* mapped[4] = {nullptr, nullptr, nullptr, nullptr};
            int native = 0;      // remote atomics over NVLink, needed by the pair-symmetric pass
            if (cudaDeviceGetP2PAttribute(&native, cudaDevP2PAttrNativeAtomicSupported, c->device, h_blobs[q].device) !=
                    cudaSuccess || !native) { cudaGetLastError(); atomics = 0; }
            if (h_blobs[q].pid == mine.pid) {        // another thread of this process: plain peer access
                int can = 0;
                if (cudaDeviceCanAccessPeer(&can, c->device, h_blobs[q].device) != cudaSuccess || !can) { ok = 0; break; }
                cudaError_t e = cudaDeviceEnablePeerAccess(h_blobs[q].device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { ok = 0; break; }
                cudaGetLastError();
                for (int k = 0; k < nh; ++k) mapped[k] = h_blobs[q].raw[k];
            } else {                                 // another process: CUDA IPC
                for (int k = 0; k < nh && ok; ++k) {
                    if (cudaIpcOpenMemHandle(&mapped[k], h_blobs[q].handle[k], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                        cudaGetLastError(); ok = 0;
                    } else {
                        c->ipc_opened[c->n_ipc_opened++] = mapped[k];
                    }
                }
                if (!ok) break;
            }
            c->peer_pos[0][q] = (float4*)mapped[0]; c->peer_pos[1][q] = (float4*)mapped[1];
            c->peer_flags[q] = (uint32_t*)mapped[2];
            c->peer_accj[q] = (long long*)mapped[3];
        }
        if (ok) ok = atomics ? 2 : 1;      // min over ranks: 0 no peer access, 1 stores only, 2 stores + atomics
        // unanimous decision
        if (cudaMemcpyAsync(d_ok, &ok, sizeof(int), cudaMemcpyHostToDevice, c->stream) != cudaSuccess ||
            nc.AllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, c->comm, c->stream) != ncclSuccess ||
            cudaMemcpyAsync(&all_ok, d_ok, sizeof(int), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
            cudaStreamSynchronize(c->stream) != cudaSuccess) { rc = GS_ENCCL; break; }
    } while (0);
    cudaFree(d_blobs); cudaFree(d_ok);
    delete[] h_blobs;
    if (rc != GS_OK) return rc;
    if (all_ok) {
        if (!c->acc_stage) GS_CUDA(c, cudaMalloc((void**)&c->acc_stage, (size_t)c->slice * sizeof(float4)));
        c->p2p = true;
        c->p2p_atomics = (all_ok == 2);
    } else {
        close_peer_mappings(c);
    }
    return GS_OK;
}

// Collective farewell in peer-memory mode: nobody unmaps or frees while a peer may still store into it.
void peer_exchange_goodbye(gs_ctx* c) {
    if (!c->p2p) return;
    PeerFlags f;
    for (int q = 0; q < c->cfg.world_size; ++q) f.slot[q] = c->peer_flags[q] + c->cfg.rank;
    c->xstep++;
    launch_signal_peers(f, c->cfg.world_size, c->cfg.rank, c->xstep, c->stream);
    launch_wait_flags(c->flags, c->cfg.world_size, c->cfg.rank, c->xstep, 5000000000ull, c->d_comm_err, c->stream);
    cudaStreamSynchronize(c->stream);      // a peer that is gone already costs the 5 s above, nothing else: we are leaving too
}

// Second half of the pair-symmetric set-up, once it is known whether peers can be reached with atomics.
int setup_symmetric(gs_ctx* c) {
    if (!c->want_sym) return GS_OK;
    if (c->cfg.world_size > 1 && !(c->p2p && c->p2p_atomics)) return GS_OK;     // stay on the row-gather kernel
    // eps == 0: select-free pass + exact redo of the rare units where two bodies coincide; in a symmetric unit no
    // body meets itself, so (unlike the row-gather kernel's diagonal tiles) there is no systematic extra work
    const char* ecm = getenv("GRAVSIM_CUTMODE");
    c->sym_cut_mode = (c->cfg.eps > 0.0) ? kSoftened : (ecm && !strcmp(ecm, "select")) ? kCutoff : kCutoffTracked;
    int ctas = 0;
    cudaError_t e = sym_kernel_setup(c->R, c->threads, c->sym_cut_mode, &ctas, &c->sym_regs);
    if (e != cudaSuccess || ctas < 1) { c->last_cuda_error = (int)e; return GS_ECUDA; }
    c->sym_grid = c->sm_count * ctas;
    const char* eg = getenv("GRAVSIM_GRID");
    if (eg && atoi(eg) > 0 && atoi(eg) <= c->sym_grid) c->sym_grid = atoi(eg);
    SymParams geo;
    fill_sym_geometry(c, &geo);
    std::vector<SymCta> hc;
    std::vector<int2> hr;
    const int cost_plain = 100, cost_sym = 133;           // issue cycles per unit, relative (34 vs 26 per j-pair and i-body)
    build_sym_plan(geo, c->sym_grid, cost_plain, cost_sym, hc, hr);
    const size_t slice_bytes = (size_t)c->slice * sizeof(float4);
    GS_CUDA(c, cudaMalloc((void**)&c->d_sym_ctas, hc.size() * sizeof(SymCta)));
    GS_CUDA(c, cudaMalloc((void**)&c->d_sym_rows, hr.size() * sizeof(int2)));
    GS_CUDA(c, cudaMalloc((void**)&c->sym_partial, (size_t)c->sym_grid * 2 * c->i_tile * sizeof(float4)));
    if (c->sym_cut_mode == kCutoffTracked) GS_CUDA(c, cudaMalloc((void**)&c->pos_scaled, (size_t)c->n_pad * sizeof(float4)));
    if (!c->acc_stage) GS_CUDA(c, cudaMalloc((void**)&c->acc_stage, slice_bytes));
    GS_CUDA(c, cudaMemcpy(c->d_sym_ctas, hc.data(), hc.size() * sizeof(SymCta), cudaMemcpyHostToDevice));
    GS_CUDA(c, cudaMemcpy(c->d_sym_rows, hr.data(), hr.size() * sizeof(int2), cudaMemcpyHostToDevice));
    c->symmetric = true;
    return GS_OK;
}

int set_device(gs_ctx* c) {
    GS_CUDA(c, cudaSetDevice(c->device));
    return GS_OK;
}

// In peer-memory mode a wait kernel that gave up on a lost peer only leaves a mark: turn it into an error code
// after every host synchronisation, so that no entry point hands back data computed from stale peer positions.
int check_peers(gs_ctx* c) {
    if (!c->p2p) return GS_OK;
    uint32_t err = 0;
    GS_CUDA(c, cudaMemcpy(&err, c->d_comm_err, sizeof(err), cudaMemcpyDeviceToHost));
    return err ? GS_ENCCL : GS_OK;
}


}  // namespace

extern "C" {

const char* gs_version(void) { return "gravsim-b200 0.1 (sm_100a)"; }

const char* gs_strerror(int code) {
    switch (code) {
        case GS_OK: return "ok";
        case GS_EINVAL: return "invalid argument";
        case GS_ECUDA: return "CUDA runtime error (see gs_last_cuda_error)";
        case GS_ENOMEM: return "out of memory";
        case GS_ENCCL: return "NCCL unavailable, collective failed or peer exchange timed out";
        case GS_ESTATE: return "call out of order";
        case GS_ENOGPU: return "no CUDA device available (libgravsim has no CPU path)";
        case GS_EIO: return "state-file I/O error";
        case GS_ETOOBIG: return "problem exceeds a fixed-size buffer (gs_potential_energy_noalloc: at most 4194304 padded bodies)";
        default: return "unknown error";
    }
}

void gs_config_init(gs_config* cfg) {
    if (!cfg) return;
    memset(cfg, 0, sizeof(*cfg));
    cfg->struct_size = (uint32_t)sizeof(gs_config);
    cfg->dt = 3600.0;                 // cuda.cu:123, mpi.c:148
    cfg->G = GS_G_SI;
    cfg->eps = 0.0;
    cfg->integrator = GS_INTEGRATOR_EULER;
    cfg->device = -1;
    cfg->rank = 0;
    cfg->world_size = 1;
}

int gs_plan_tiles(int variant, int* i_tile, int* j_tile) {
    if (variant == 0) variant = kDefaultVariant;
    VariantInfo vi;
    if (!variant_info(variant, &vi)) return GS_EINVAL;
    if (i_tile) *i_tile = vi.i_tile;
    if (j_tile) *j_tile = kJTile;
    return GS_OK;
}

int gs_plan_slice(int n, int world_size, int rank, int variant, int* n_pad, int* i_begin_pad, int* i_end_pad) {
    if (n <= 0 || world_size < 1 || rank < 0 || rank >= world_size) return GS_EINVAL;
    if (variant == 0) variant = kDefaultVariant;
    VariantInfo vi;
    if (!variant_info(variant, &vi)) return GS_EINVAL;
    // every rank's slice is a whole number of i-tiles and of j-tiles
    int64_t tile = vi.i_tile;
    int64_t unit = tile;
    while (unit % kJTile != 0) unit += tile;
    unit *= world_size;
    const int64_t np = ((n + unit - 1) / unit) * unit;
    if (np > INT32_MAX) return GS_EINVAL;
    const int slice = (int)(np / world_size);
    if (n_pad) *n_pad = (int)np;
    if (i_begin_pad) *i_begin_pad = rank * slice;
    if (i_end_pad) *i_end_pad = (rank + 1) * slice;
    return GS_OK;
}

int gs_plan_cta_range(int64_t ni_tiles, int64_t nj_tiles, int grid, int cta, int64_t* u0, int64_t* u1) {
    if (ni_tiles <= 0 || nj_tiles <= 0 || grid <= 0 || cta < 0 || cta >= grid) return GS_EINVAL;
    const int64_t U = ni_tiles * nj_tiles;
    if (u0) *u0 = cta_unit_begin(U, grid, cta);
    if (u1) *u1 = cta_unit_begin(U, grid, cta + 1);
    return GS_OK;
}

int gs_plan_tile_owners(int64_t ni_tiles, int64_t nj_tiles, int grid, int64_t it, int* cta_first, int* cta_last,
                        int* is_split) {
    if (ni_tiles <= 0 || nj_tiles <= 0 || grid <= 0 || it < 0 || it >= ni_tiles) return GS_EINVAL;
    const TileOwners o = tile_owners(ni_tiles, nj_tiles, grid, it);
    if (cta_first) *cta_first = o.first;
    if (cta_last) *cta_last = o.last;
    if (is_split) *is_split = o.split ? 1 : 0;
    return GS_OK;
}

// ---- pair-symmetric plan, host side (CPU-testable) ----------------------------------------------------
static int sym_geometry_for(int n, int world, int rank, int variant, SymParams* sp) {
    VariantInfo vi;
    if (variant == 0) variant = 3;
    if (!variant_info(variant, &vi) || world < 1 || rank < 0 || rank >= world) return GS_EINVAL;
    int n_pad = 0, b = 0, e = 0;
    if (gs_plan_slice(n, world, rank, variant, &n_pad, &b, &e) != GS_OK) return GS_EINVAL;
    gs_ctx tmp;
    tmp.cfg.world_size = world; tmp.cfg.rank = rank;
    tmp.n = n; tmp.i_begin = b; tmp.slice = e - b; tmp.i_tile = vi.i_tile;
    fill_sym_geometry(&tmp, sp);
    return GS_OK;
}

int gs_plan_sym_row(int n, int world_size, int rank, int variant, int row, int* n_rows, int* n_units, int* n_plain) {
    SymParams sp;
    int rc = sym_geometry_for(n, world_size, rank, variant, &sp);
    if (rc != GS_OK) return rc;
    if (row < 0 || row >= sp.ni_tiles) return GS_EINVAL;
    if (n_rows) *n_rows = sp.ni_tiles;
    if (n_units) *n_units = sym_row_units(sp, row);
    if (n_plain) *n_plain = sp.diag;
    return GS_OK;
}

int gs_plan_sym_unit(int n, int world_size, int rank, int variant, int row, int k, int* j_tile) {
    SymParams sp;
    int rc = sym_geometry_for(n, world_size, rank, variant, &sp);
    if (rc != GS_OK) return rc;
    if (row < 0 || row >= sp.ni_tiles || k < 0 || k >= sym_row_units(sp, row) || !j_tile) return GS_EINVAL;
    *j_tile = sym_unit_tile(sp, row, k);
    return GS_OK;
}

int gs_plan_sym_cta(int n, int world_size, int rank, int variant, int grid, int cta, int* row0, int* k0, int* count) {
    SymParams sp;
    int rc = sym_geometry_for(n, world_size, rank, variant, &sp);
    if (rc != GS_OK) return rc;
    if (grid < 1 || cta < 0 || cta >= grid) return GS_EINVAL;
    std::vector<SymCta> hc;
    std::vector<int2> hr;
    build_sym_plan(sp, grid, 100, 133, hc, hr);
    if (row0) *row0 = hc[cta].row0;
    if (k0) *k0 = hc[cta].k0;
    if (count) *count = hc[cta].count;
    return GS_OK;
}

int gs_nccl_unique_id(void* out_id) {
    if (!out_id) return GS_EINVAL;
    NcclApi& nc = nccl_api();
    if (!nc.ok) return GS_ENCCL;
    ncclUniqueId id;
    if (nc.GetUniqueId(&id) != ncclSuccess) return GS_ENCCL;
    static_assert(sizeof(id) == GS_NCCL_ID_BYTES, "ncclUniqueId size");
    memcpy(out_id, &id, sizeof(id));
    return GS_OK;
}

int gs_host_alloc(void** ptr, size_t bytes) {
    if (!ptr) return GS_EINVAL;
    return cudaHostAlloc(ptr, bytes, cudaHostAllocDefault) == cudaSuccess ? GS_OK : GS_ENOMEM;
}
int gs_host_free(void* ptr) { return cudaFreeHost(ptr) == cudaSuccess ? GS_OK : GS_ECUDA; }

void gs_destroy(gs_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    peer_exchange_goodbye(c);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->comm_stream) cudaStreamSynchronize(c->comm_stream);
    close_peer_mappings(c);
    cudaFree(c->flags); cudaFree(c->d_comm_err);
    if (c->comm) nccl_api().CommDestroy(c->comm);
    cudaFree(c->pos[0]); cudaFree(c->pos[1]); cudaFree(c->vel); cudaFree(c->vel_full);
    cudaFree(c->d_sym_ctas); cudaFree(c->d_sym_rows); cudaFree(c->d_accj); cudaFree(c->sym_partial); cudaFree(c->pos_scaled);
    cudaFree(c->pos_lo); cudaFree(c->vel_lo); cudaFree(c->lo_full[0]); cudaFree(c->lo_full[1]);
    cudaFree(c->acc_local); cudaFree(c->acc_stage); cudaFree(c->acc_lo); cudaFree(c->acc_out); cudaFree(c->partial); cudaFree(c->mass);
    cudaFree(c->d_aos); cudaFree(c->d_energy); cudaFree(c->d_mu_mm); cudaFree(c->d_log);
    if (c->h_energy) cudaFreeHost(c->h_energy);
    if (c->h_mu_mm) cudaFreeHost(c->h_mu_mm);
    if (c->h_acc) cudaFreeHost(c->h_acc);
    if (c->ev_integ) cudaEventDestroy(c->ev_integ);
    for (int b = 0; b < 2; ++b)
        if (c->ev_gather[b]) cudaEventDestroy(c->ev_gather[b]);
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    if (c->comm_stream) cudaStreamDestroy(c->comm_stream);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    for (int b = 0; b < 2; ++b) {
        if (c->h_snap[b]) cudaFreeHost(c->h_snap[b]);
        if (c->ev_snap_ready[b]) cudaEventDestroy(c->ev_snap_ready[b]);
        if (c->ev_snap_done[b]) cudaEventDestroy(c->ev_snap_done[b]);
    }
    delete c;
}

int gs_create(gs_ctx** out, const gs_config* cfg) {
    if (!out || !cfg) return GS_EINVAL;
    *out = nullptr;
    if (cfg->struct_size != sizeof(gs_config)) return GS_EINVAL;
    if (cfg->n <= 0 || cfg->world_size < 1 || cfg->rank < 0 || cfg->rank >= cfg->world_size) return GS_EINVAL;
    if (!(cfg->G > 0.0) || !std::isfinite(cfg->dt) || !(cfg->eps >= 0.0)) return GS_EINVAL;
    if (cfg->integrator != GS_INTEGRATOR_EULER && cfg->integrator != GS_INTEGRATOR_LEAPFROG) return GS_EINVAL;
    if (cfg->world_size > 1 && !cfg->nccl_id) return GS_EINVAL;

    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) return GS_ENOGPU;

    gs_ctx* c = new (std::nothrow) gs_ctx();
    if (!c) return GS_ENOMEM;
    c->cfg = *cfg;
    c->cfg.nccl_id = nullptr;
    int rc = GS_OK;
    auto fail = [&](int code) { gs_destroy(c); return code; };

    if (cfg->device >= 0) c->device = cfg->device;
    else if (cudaGetDevice(&c->device) != cudaSuccess) return fail(GS_ENOGPU);
    if (c->device >= ndev) return fail(GS_EINVAL);
    if ((rc = set_device(c)) != GS_OK) return fail(rc);

    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, c->device) != cudaSuccess) return fail(GS_ECUDA);
    if (prop.major < 10) return fail(GS_ENOGPU);   // sm_100a code only
    c->sm_count = prop.multiProcessorCount;

    c->cut_mode = (cfg->eps > 0.0) ? kSoftened : kCutoffTracked;
    const char* ecm = getenv("GRAVSIM_CUTMODE");                        // "select" | "tracked": tests, A/B runs
    if (c->cut_mode != kSoftened && ecm && !strcmp(ecm, "select")) c->cut_mode = kCutoff;
    c->variant = cfg->variant;
    if (c->variant == 0) c->variant = default_variant(cfg->n, cfg->world_size, c->cut_mode, c->cfg.flags);
    VariantInfo vi;
    if (!variant_info(c->variant, &vi)) return fail(GS_EINVAL);
    c->R = vi.R;
    c->threads = vi.threads;
    c->i_tile = vi.i_tile;
    c->eps2 = (float)(cfg->eps * cfg->eps);
    c->cut2 = (c->cut_mode != kSoftened) ? 1e-20f : 0.f;   // (1e-10)^2: cuda.cu:39, mpi.c:64

    c->n = cfg->n;
    int ib = 0, ie = 0;
    if (gs_plan_slice(c->n, cfg->world_size, cfg->rank, c->variant, &c->n_pad, &ib, &ie) != GS_OK) return fail(GS_EINVAL);
    c->i_begin = ib;
    c->slice = ie - ib;

    // both flavours of the kernel (general / hoisted uniform mass) must fit the same persistent grid
    int ctas_u = 0, regs_u = 0;
    cudaError_t e = force_kernel_occupancy(c->variant, c->cut_mode, false, &c->ctas_per_sm, &c->regs);
    if (e == cudaSuccess) e = force_kernel_occupancy(c->variant, c->cut_mode, true, &ctas_u, &regs_u);
    if (e != cudaSuccess || c->ctas_per_sm < 1 || ctas_u < 1) { c->last_cuda_error = (int)e; return fail(GS_ECUDA); }
    if (ctas_u < c->ctas_per_sm) c->ctas_per_sm = ctas_u;
    c->grid = c->sm_count * c->ctas_per_sm;
    const char* eg = getenv("GRAVSIM_GRID");
    if (eg && atoi(eg) > 0 && atoi(eg) <= c->grid) c->grid = atoi(eg);

    if (c->cut_mode == kCutoffTracked && !(ecm && !strcmp(ecm, "tracked"))) {
        // The tracked mode recomputes the i_tile/j_tile "diagonal" units of every i-tile (those that contain the
        // bodies themselves) with the select.  The static partition does not know that: when a CTA only has a few
        // dozen units, the CTAs holding diagonal units become a tail (N = 65,536: +14 %).  Use the plain select
        // kernel (65 % of peak instead of 67 %) unless the extra work is below 1 % of a CTA's share.
        const double units_per_cta = (double)(c->slice / c->i_tile) * (c->n_pad / kJTile) / c->grid;
        if ((double)(c->i_tile / kJTile) > 0.01 * units_per_cta) {
            c->cut_mode = kCutoff;
            int ctas = 0, regs = 0, ctas2 = 0, regs2 = 0;
            e = force_kernel_occupancy(c->variant, c->cut_mode, false, &ctas, &regs);
            if (e == cudaSuccess) e = force_kernel_occupancy(c->variant, c->cut_mode, true, &ctas2, &regs2);
            if (e != cudaSuccess || ctas < 1 || ctas2 < 1) { c->last_cuda_error = (int)e; return fail(GS_ECUDA); }
            c->ctas_per_sm = ctas < ctas2 ? ctas : ctas2;
            c->regs = regs;
            c->grid = c->sm_count * c->ctas_per_sm;
            if (eg && atoi(eg) > 0 && atoi(eg) <= c->grid) c->grid = atoi(eg);
        }
    }
    {
        const int64_t ni = c->slice / c->i_tile, nj = c->n_pad / kJTile, njl = c->slice / kJTile;
        c->split_full = any_split(ni, nj, c->grid);
        if (cfg->world_size > 1) {
            c->split_local = any_split(ni, njl, c->grid);
            c->split_remote = any_split(ni, nj - njl, c->grid);
        } else if ((c->cfg.flags & GS_FLAG_TWO_PASS) && nj >= 2) {
            c->split_local = any_split(ni, nj / 2, c->grid);
            c->split_remote = any_split(ni, nj - nj / 2, c->grid);
        }
    }

    auto dmalloc = [&](void** p, size_t bytes) { return cudaMalloc(p, bytes) == cudaSuccess; };
    const size_t pos_bytes = (size_t)c->n_pad * sizeof(float4);
    const size_t vel_bytes = (size_t)c->slice * sizeof(float4);
    c->partial_bytes = (size_t)c->grid * 2 * c->i_tile * sizeof(float4);
    const int egrid = energy_grid(c->slice);
    if (!dmalloc((void**)&c->pos[0], pos_bytes) || !dmalloc((void**)&c->pos[1], pos_bytes) ||
        !dmalloc((void**)&c->vel, vel_bytes) || !dmalloc((void**)&c->partial, c->partial_bytes) ||
        !dmalloc((void**)&c->mass, (size_t)c->n_pad * sizeof(double)) ||
        !dmalloc((void**)&c->d_energy, ((size_t)egrid * 5 + 8) * sizeof(double)))
        return fail(GS_ENOMEM);
    if ((c->cfg.flags & GS_FLAG_COMPENSATED) &&
        (!dmalloc((void**)&c->pos_lo, vel_bytes) || !dmalloc((void**)&c->vel_lo, vel_bytes) ||
         (!c->acc_stage && !dmalloc((void**)&c->acc_stage, vel_bytes))))
        return fail(GS_ENOMEM);
    if ((c->cfg.flags & GS_FLAG_FP64_FORCES) &&
        ((!c->acc_stage && !dmalloc((void**)&c->acc_stage, vel_bytes)) || !dmalloc((void**)&c->acc_lo, vel_bytes)))
        return fail(GS_ENOMEM);
    // pair-symmetric pass: wanted unless the caller asks for bit-reproducible sums; needs a 256-thread R=4/6 variant
    c->want_sym = !(c->cfg.flags & (GS_FLAG_DETERMINISTIC | GS_FLAG_FP64_FORCES)) && sym_supported(c->R, c->threads) &&
                  cfg->world_size <= kMaxPeers + 1;
    if (c->want_sym) {      // allocated now so that setup_peer_exchange() can publish it to the peers
        const size_t accj_bytes = (size_t)c->slice * kSymAccjBytesPerBody;
        if (!dmalloc((void**)&c->d_accj, accj_bytes)) return fail(GS_ENOMEM);
        if (cudaMemset(c->d_accj, 0, accj_bytes) != cudaSuccess) return fail(GS_ECUDA);
        c->peer_accj[cfg->rank] = c->d_accj;
    }
    if ((cfg->world_size > 1 || (c->cfg.flags & GS_FLAG_TWO_PASS)) && !dmalloc((void**)&c->acc_local, vel_bytes))
        return fail(GS_ENOMEM);
    if (cudaHostAlloc((void**)&c->h_energy, ((size_t)egrid * 5 + 8) * sizeof(double), cudaHostAllocDefault) !=
        cudaSuccess)
        return fail(GS_ENOMEM);
    if (!dmalloc((void**)&c->d_mu_mm, 4 * sizeof(uint32_t)) ||
        cudaHostAlloc((void**)&c->h_mu_mm, 4 * sizeof(uint32_t), cudaHostAllocDefault) != cudaSuccess)
        return fail(GS_ENOMEM);
    c->h_mu_mm[2] = 0xffffffffu; c->h_mu_mm[3] = 0u;      // presets of {min, max}

    if (cfg->stream) { c->stream = (cudaStream_t)cfg->stream; c->own_stream = false; }
    else {
        if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) return fail(GS_ECUDA);
        c->own_stream = true;
    }
    if (cudaEventCreateWithFlags(&c->ev_integ, cudaEventDisableTiming) != cudaSuccess) return fail(GS_ECUDA);
    for (int b = 0; b < 2; ++b)
        if (cudaEventCreateWithFlags(&c->ev_gather[b], cudaEventDisableTiming) != cudaSuccess) return fail(GS_ECUDA);

    if (cfg->world_size > 1) {
        NcclApi& nc = nccl_api();
        if (!nc.ok) return fail(GS_ENCCL);
        if (cudaStreamCreateWithFlags(&c->comm_stream, cudaStreamNonBlocking) != cudaSuccess) return fail(GS_ECUDA);
        ncclUniqueId id;
        memcpy(&id, cfg->nccl_id, sizeof(id));
        if (nc.CommInitRank(&c->comm, cfg->world_size, id, cfg->rank) != ncclSuccess) return fail(GS_ENCCL);
        if ((rc = setup_peer_exchange(c)) != GS_OK) return fail(rc);
    }
    if ((rc = setup_symmetric(c)) != GS_OK) return fail(rc);
    *out = c;
    return GS_OK;
}

// Are all uploaded masses equal?  Decided from the {min, max} bit patterns of mu the upload kernel left in
// d_mu_mm (on several GPUs after a slice upload: min over ranks of {min, ~max}).  Needs the value on the host because
// it picks the kernel (mass factor hoisted out of the inner loop), hence one 8-byte D2H and a stream synchronisation.
static int decide_uniform_mass(gs_ctx* c, bool reduce_over_ranks) {
    c->mu_uniform = 0.f;
    if (c->cfg.flags & GS_FLAG_NO_UNIFORM_MASS) return GS_OK;
    if (reduce_over_ranks && c->cfg.world_size > 1) {
        if (nccl_api().AllReduce(c->d_mu_mm + 2, c->d_mu_mm + 2, 2, ncclUint32, ncclMin, c->comm, c->stream) != ncclSuccess)
            return GS_ENCCL;
    }
    GS_CUDA(c, cudaMemcpyAsync(c->h_mu_mm, c->d_mu_mm + 2, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    const uint32_t lo = c->h_mu_mm[0], hi = ~c->h_mu_mm[1];
    float mu;
    memcpy(&mu, &lo, 4);
    if (lo == hi && mu > 0.f && std::isfinite(mu)) c->mu_uniform = mu;
    return GS_OK;
}

static int reset_after_upload(gs_ctx* c) {
    if (c->copy_stream) GS_CUDA(c, cudaStreamSynchronize(c->copy_stream));
    c->buf_copy_pending[0] = c->buf_copy_pending[1] = false;
    c->uploaded = true;
    c->vel_synced = true;
    c->steps_done = 0;
    c->sim_time = 0.0;
    c->mon_count = 0;
    return GS_OK;
}

static int upload_common(gs_ctx* c, const void* host, int n, int record_bytes) {
    if (!c || !host) return GS_EINVAL;
    if (n != c->n) return GS_EINVAL;
    int rc = set_device(c);
    if (rc != GS_OK) return rc;
    const size_t bytes = (size_t)n * record_bytes;
    if ((rc = ensure_aos(c, bytes)) != GS_OK) return rc;
    GS_CUDA(c, cudaMemcpyAsync(c->d_aos, host, bytes, cudaMemcpyHostToDevice, c->stream));
    GS_CUDA(c, cudaMemcpyAsync(c->d_mu_mm, c->h_mu_mm + 2, 2 * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
    c->cur = 0;
    // both buffers must carry mu and the padding: remote slices are filled by the exchange, the own slice by the
    // integrator, but the padding bodies of a single-GPU run are only ever written here
    UploadTargets dst;
    dst.n = 2; dst.pos[0] = c->pos[0]; dst.pos[1] = c->pos[1];
    GS_CUDA(c, launch_aos_to_packed(c->d_aos, record_bytes, 0, n, 0, c->n_pad, c->cfg.G, dst, c->vel, c->mass, c->i_begin,
                                    c->slice, c->pos_lo, c->vel_lo, c->d_mu_mm, c->stream));
    GS_CUDA(c, launch_pack_minmax(c->d_mu_mm, c->stream));
    c->launches += 2;
    c->mass_full = true;
    if (c->cfg.world_size > 1) {
        // every rank uploaded the full array: pos[0] is already complete everywhere.
        if (c->p2p) {
            // still an exchange epoch: no peer may start storing the next step into my buffers before I am
            // done with the old contents (my stream has passed this point), and vice versa.
            if ((rc = exchange_positions(c, 0)) != GS_OK) return rc;
        } else {
            GS_CUDA(c, cudaEventRecord(c->ev_integ, c->stream));
            GS_CUDA(c, cudaStreamWaitEvent(c->comm_stream, c->ev_integ, 0));
            GS_CUDA(c, cudaEventRecord(c->ev_gather[0], c->comm_stream));
        }
    }
    if ((rc = decide_uniform_mass(c, false)) != GS_OK) return rc;     // every rank saw every mass
    return reset_after_upload(c);
}

int gs_upload_aos_f32(gs_ctx* ctx, const void* particles28, int n) { return upload_common(ctx, particles28, n, 28); }
int gs_upload_aos_f64(gs_ctx* ctx, const void* particles56, int n) { return upload_common(ctx, particles56, n, 56); }

// A rank uploads only the records of its own i-bodies (mpi.c:184-187: rank r owns [start_idx, start_idx + local_n));
// the packed positions reach the other GPUs over NVLink instead of every rank pushing the whole array through PCIe.
static int upload_slice_common(gs_ctx* c, const void* host, int first, int count, int record_bytes) {
    if (!c || (!host && count > 0)) return GS_EINVAL;
    const int lo = c->i_begin < c->n ? c->i_begin : c->n;
    const int hi = (c->i_begin + c->slice) < c->n ? (c->i_begin + c->slice) : c->n;
    if (first != lo || count != hi - lo) return GS_EINVAL;
    int rc = set_device(c);
    if (rc != GS_OK) return rc;
    const int W = c->cfg.world_size;
    const size_t bytes = (size_t)count * record_bytes;
    if (count > 0) {
        if ((rc = ensure_aos(c, bytes)) != GS_OK) return rc;
        GS_CUDA(c, cudaMemcpyAsync(c->d_aos, host, bytes, cudaMemcpyHostToDevice, c->stream));
    }
    GS_CUDA(c, cudaMemcpyAsync(c->d_mu_mm, c->h_mu_mm + 2, 2 * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
    c->cur = 0;
    UploadTargets dst;
    dst.n = 0;
    if (W > 1 && c->p2p) {
        // Epoch 1: nobody stores into a peer before that peer's stream is past everything that read its buffers
        // (a gs_accel / gs_energy after the last step); epoch 2 (below): the stores have landed everywhere.
        if ((rc = wait_positions(c)) != GS_OK) return rc;
        if ((rc = exchange_positions(c, 0)) != GS_OK) return rc;
        if ((rc = wait_positions(c)) != GS_OK) return rc;
        for (int q = 0; q < W; ++q) { dst.pos[dst.n++] = c->peer_pos[0][q]; dst.pos[dst.n++] = c->peer_pos[1][q]; }
    } else {
        dst.pos[dst.n++] = c->pos[0];
        dst.pos[dst.n++] = c->pos[1];
    }
    GS_CUDA(c, launch_aos_to_packed(c->d_aos, record_bytes, first, c->n, c->i_begin, c->slice, c->cfg.G, dst, c->vel,
                                    c->mass, c->i_begin, c->slice, c->pos_lo, c->vel_lo, c->d_mu_mm, c->stream));
    GS_CUDA(c, launch_pack_minmax(c->d_mu_mm, c->stream));
    c->launches += 2;
    c->mass_full = (W == 1);
    if (W > 1) {
        if (c->p2p) {
            if ((rc = exchange_positions(c, 0)) != GS_OK) return rc;
        } else {        // no peer access: in-place all-gather of both buffers
            NcclApi& nc = nccl_api();
            for (int b = 0; b < 2; ++b)
                if (nc.AllGather(c->pos[b] + c->i_begin, c->pos[b], (size_t)c->slice * 4, ncclFloat, c->comm, c->stream) != ncclSuccess)
                    return GS_ENCCL;
            GS_CUDA(c, cudaEventRecord(c->ev_integ, c->stream));
            GS_CUDA(c, cudaStreamWaitEvent(c->comm_stream, c->ev_integ, 0));
            GS_CUDA(c, cudaEventRecord(c->ev_gather[0], c->comm_stream));
        }
    }
    if ((rc = decide_uniform_mass(c, true)) != GS_OK) return rc;
    return reset_after_upload(c);
}

int gs_upload_slice_f32(gs_ctx* ctx, const void* my28, int first, int count) { return upload_slice_common(ctx, my28, first, count, 28); }
int gs_upload_slice_f64(gs_ctx* ctx, const void* my56, int first, int count) { return upload_slice_common(ctx, my56, first, count, 56); }

static int energy_enqueue(gs_ctx* c, double* out8);

static int one_step(gs_ctx* c) {
    int rc;
    const float dt = (float)c->cfg.dt;
    const int nxt = c->cur ^ 1;
    if (c->buf_copy_pending[nxt]) {      // a recorded snapshot of that buffer must have left before it is overwritten
        GS_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_snap_done[c->buf_copy_event[nxt]], 0));
        c->buf_copy_pending[nxt] = false;
    }
    float kick = dt;
    if (c->cfg.integrator == GS_INTEGRATOR_LEAPFROG && c->vel_synced) kick = 0.5f * dt;   // opening half kick
    if ((rc = force_pass(c, kFinalIntegrate, kick, dt, nullptr)) != GS_OK) return rc;
    if ((rc = exchange_positions(c, nxt)) != GS_OK) return rc;
    c->cur = nxt;
    if (c->cfg.integrator == GS_INTEGRATOR_LEAPFROG) c->vel_synced = false;
    c->steps_done++;
    c->sim_time += c->cfg.dt;
    if (c->mon_every > 0 && c->steps_done % c->mon_every == 0 && c->mon_count < c->mon_capacity) {
        // the sample is computed and stored on the device; with the leapfrog this closes the half kick first
        if ((rc = energy_enqueue(c, c->d_log + (size_t)8 * c->mon_count)) != GS_OK) return rc;
        c->mon_count++;
    }
    return GS_OK;
}

int gs_step(gs_ctx* c, int nsteps) {
    if (!c || nsteps < 0) return GS_EINVAL;
    if (!c->uploaded) return GS_ESTATE;
    int rc = set_device(c);
    if (rc != GS_OK) return rc;
    for (int s = 0; s < nsteps; ++s)
        if ((rc = one_step(c)) != GS_OK) return rc;
    // order the last exchange on the caller-visible stream too, so an event recorded there after
    // gs_step covers every byte of the step (positions are then valid on every rank)
    if (c->cfg.world_size > 1 && nsteps > 0 && (rc = wait_positions(c)) != GS_OK) return rc;
    return GS_OK;
}

static void unpack_snapshot(const float4* packed, int n, float* xyz) {
    for (int i = 0; i < n; ++i) {
        const float4 A = packed[2 * (i >> 1)], B = packed[2 * (i >> 1) + 1];
        const bool odd = (i & 1) != 0;
        xyz[3 * (size_t)i + 0] = odd ? A.y : A.x;
        xyz[3 * (size_t)i + 1] = odd ? A.w : A.z;
        xyz[3 * (size_t)i + 2] = odd ? B.y : B.x;
    }
}

int gs_step_trajectory(gs_ctx* c, int nsteps, int every, float* xyz_host, int capacity_snapshots, int* n_snapshots) {
    if (!c || nsteps < 0 || every < 1 || !xyz_host) return GS_EINVAL;
    if (!c->uploaded) return GS_ESTATE;
    const int count = nsteps / every;
    if (count > capacity_snapshots) return GS_EINVAL;
    int rc = set_device(c);
    if (rc != GS_OK) return rc;
    const size_t snap_bytes = (size_t)c->n_pad * sizeof(float4);
    if (!c->copy_stream) {
        GS_CUDA(c, cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
        for (int b = 0; b < 2; ++b) {
            GS_CUDA(c, cudaHostAlloc((void**)&c->h_snap[b], snap_bytes, cudaHostAllocDefault));
            GS_CUDA(c, cudaEventCreateWithFlags(&c->ev_snap_ready[b], cudaEventDisableTiming));
            GS_CUDA(c, cudaEventCreateWithFlags(&c->ev_snap_done[b], cudaEventDisableTiming));
        }
    }
    const size_t frame = (size_t)c->n * 3;
    int taken = 0, unpacked = 0;
    for (int s = 1; s <= nsteps; ++s) {
        if ((rc = one_step(c)) != GS_OK) return rc;
        if (s % every != 0) continue;
        const int b = taken & 1;
        if (taken >= 2) {                       // staging buffer b still holds snapshot taken-2: unpack it first
            GS_CUDA(c, cudaEventSynchronize(c->ev_snap_done[b]));
            unpack_snapshot(c->h_snap[b], c->n, xyz_host + frame * unpacked++);
        }
        if ((rc = wait_positions(c)) != GS_OK) return rc;            // every rank's bodies are in pos[cur]
        GS_CUDA(c, cudaEventRecord(c->ev_snap_ready[b], c->stream));
        GS_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_snap_ready[b], 0));
        GS_CUDA(c, cudaMemcpyAsync(c->h_snap[b], c->pos[c->cur], snap_bytes, cudaMemcpyDeviceToHost, c->copy_stream));
        GS_CUDA(c, cudaEventRecord(c->ev_snap_done[b], c->copy_stream));
        if (c->cfg.world_size > 1) {
            // peers store into this buffer two steps from now, gated only by my step counter: keep it simple
            // and let the engine's stream wait for the copy right away
            GS_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_snap_done[b], 0));
        } else {
            c->buf_copy_pending[c->cur] = true;   // one_step() waits only when this buffer is about to be overwritten
            c->buf_copy_event[c->cur] = b;
        }
        ++taken;
    }
    while (unpacked < taken) {
        const int b = unpacked & 1;
        GS_CUDA(c, cudaEventSynchronize(c->ev_snap_done[b]));
        unpack_snapshot(c->h_snap[b], c->n, xyz_host + frame * unpacked++);
    }
    if (c->cfg.world_size > 1 && nsteps > 0 && (rc = wait_positions(c)) != GS_OK) return rc;
    if (n_snapshots) *n_snapshots = taken;
    return taken > 0 ? check_peers(c) : GS_OK;        // the frames were synchronised above
}

static int accel_common(gs_ctx* c, float* acc3_f32, double* acc3_f64) {
    if (!c || (!acc3_f32 && !acc3_f64)) return GS_EINVAL;
    if (!c->uploaded) return GS_ESTATE;
    int rc = set_device(c);
    if (rc != GS_OK) return rc;
    const size_t bytes = (size_t)c->slice * sizeof(float4);
    if (!c->acc_out) GS_CUDA(c, cudaMalloc((void**)&c->acc_out, bytes));
    if (c->h_acc_bytes < 2 * bytes) {
        if (c->h_acc) cudaFreeHost(c->h_acc);
        c->h_acc = nullptr; c->h_acc_bytes = 0;
        GS_CUDA(c, cudaHostAlloc((void**)&c->h_acc, 2 * bytes, cudaHostAllocDefault));
        c->h_acc_bytes = 2 * bytes;
    }
    if ((rc = force_pass(c, kFinalAccOut, 0.f, 0.f, c->acc_out)) != GS_OK) return rc;
    int n_local = c->n - c->i_begin;
    if (n_local > c->slice) n_local = c->slice;
    if (n_local < 0) n_local = 0;
    const bool have_lo = (c->cfg.flags & GS_FLAG_FP64_FORCES) != 0;
    GS_CUDA(c, cudaMemcpyAsync(c->h_acc, c->acc_out, (size_t)n_local * sizeof(float4), cudaMemcpyDeviceToHost,
                               c->stream));
    if (have_lo)
        GS_CUDA(c, cudaMemcpyAsync(c->h_acc + c->slice, c->acc_lo, (size_t)n_local * sizeof(float4), cudaMemcpyDeviceToHost,
                                   c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    if ((rc = check_peers(c)) != GS_OK) return rc;
    for (int i = 0; i < n_local; ++i) {
        const float4 h = c->h_acc[i];
        const float4 l = have_lo ? c->h_acc[c->slice + i] : make_float4(0.f, 0.f, 0.f, 0.f);
        if (acc3_f32) { acc3_f32[3 * (size_t)i + 0] = h.x; acc3_f32[3 * (size_t)i + 1] = h.y; acc3_f32[3 * (size_t)i + 2] = h.z; }
        if (acc3_f64) {
            acc3_f64[3 * (size_t)i + 0] = (double)h.x + l.x; acc3_f64[3 * (size_t)i + 1] = (double)h.y + l.y;
            acc3_f64[3 * (size_t)i + 2] = (double)h.z + l.z;
        }
    }
    return GS_OK;
}

int gs_accel(gs_ctx* c, float* acc3_host) { return accel_common(c, acc3_host, nullptr); }
int gs_accel_f64(gs_ctx* c, double* acc3_host) { return accel_common(c, nullptr, acc3_host); }

// Enqueues one energy / momentum sample into out8 (device memory, 8 doubles: step, time, KE, PE, px, py, pz, 0):
// potential pass (O(N^2), fp32 pair terms, fp64 sums), kinetic pass, device-side reduction and, on several GPUs, an
// in-place ncclAllReduce of the five sums.  No host synchronisation, no host arithmetic.
static int energy_enqueue(gs_ctx* c, double* out8) {
    int rc;
    if ((rc = sync_velocities(c)) != GS_OK) return rc;
    if ((rc = wait_positions(c)) != GS_OK) return rc;
    const int egrid = energy_grid(c->slice);
    double* d_pot = c->d_energy;
    double* d_kin = c->d_energy + egrid;
    GS_CUDA(c, launch_potential(c->pos[c->cur], c->n_pad, c->i_begin, c->slice, c->eps2, c->cut2, d_pot, c->stream));
    GS_CUDA(c, launch_kinetic(c->vel, c->mass, c->i_begin, c->slice, d_kin, c->stream));
    GS_CUDA(c, launch_energy_sample(d_pot, d_kin, egrid, c->cfg.G, (double)c->steps_done, c->sim_time, out8, c->stream));
    c->launches += 3;
    if (c->cfg.world_size > 1 &&
        nccl_api().AllReduce(out8 + 2, out8 + 2, 5, ncclDouble, ncclSum, c->comm, c->stream) != ncclSuccess)
        return GS_ENCCL;
    return GS_OK;
}

int gs_energy(gs_ctx* c, double* kinetic, double* potential, double momentum[3]) {
    if (!c) return GS_EINVAL;
    if (!c->uploaded) return GS_ESTATE;
    int rc = set_device(c);
    if (rc != GS_OK) return rc;
    double* d_out = c->d_energy + (size_t)energy_grid(c->slice) * 5;
    if ((rc = energy_enqueue(c, d_out)) != GS_OK) return rc;
    GS_CUDA(c, cudaMemcpyAsync(c->h_energy, d_out, 8 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    if ((rc = check_peers(c)) != GS_OK) return rc;
    if (kinetic) *kinetic = c->h_energy[2];
    if (potential) *potential = c->h_energy[3];
    if (momentum) { momentum[0] = c->h_energy[4]; momentum[1] = c->h_energy[5]; momentum[2] = c->h_energy[6]; }
    return GS_OK;
}

int gs_monitor_start(gs_ctx* c, int every, int capacity) {
    if (!c || every < 0 || capacity < 0) return GS_EINVAL;
    int rc = set_device(c);
    if (rc != GS_OK) return rc;
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    cudaFree(c->d_log);
    c->d_log = nullptr;
    c->mon_every = c->mon_capacity = c->mon_count = 0;
    if (every == 0 || capacity == 0) return GS_OK;           // monitor off
    GS_CUDA(c, cudaMalloc((void**)&c->d_log, (size_t)capacity * 8 * sizeof(double)));
    c->mon_every = every; c->mon_capacity = capacity;
    return GS_OK;
}

int gs_monitor_read(gs_ctx* c, double* samples8, int capacity, int* count) {
    if (!c || !count || capacity < 0 || (capacity > 0 && !samples8)) return GS_EINVAL;
    int rc = set_device(c);
    if (rc != GS_OK) return rc;
    const int n = c->mon_count < capacity ? c->mon_count : capacity;
    if (n > 0) GS_CUDA(c, cudaMemcpyAsync(samples8, c->d_log, (size_t)n * 8 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    *count = n;
    return check_peers(c);
}

static int download_common(gs_ctx* c, void* host, int n, int record_bytes) {
    if (!c || !host) return GS_EINVAL;
    if (!c->uploaded) return GS_ESTATE;
    if (n != c->n) return GS_EINVAL;
    int rc = set_device(c);
    if (rc != GS_OK) return rc;
    if ((rc = sync_velocities(c)) != GS_OK) return rc;
    const float4* vel_full = c->vel;
    const float4 *pos_lo_full = c->pos_lo, *vel_lo_full = c->vel_lo;
    if (c->cfg.world_size > 1) {
        NcclApi& nc = nccl_api();
        if (!c->vel_full) GS_CUDA(c, cudaMalloc((void**)&c->vel_full, (size_t)c->n_pad * sizeof(float4)));
        if ((rc = wait_positions(c)) != GS_OK) return rc;
        if (nc.AllGather(c->vel, c->vel_full, (size_t)c->slice * 4, ncclFloat, c->comm, c->stream) != ncclSuccess)
            return GS_ENCCL;
        vel_full = c->vel_full;
        if (!c->mass_full) {     // after a slice upload every rank only holds its own masses
            if (nc.AllGather(c->mass + c->i_begin, c->mass, (size_t)c->slice, ncclDouble, c->comm, c->stream) != ncclSuccess)
                return GS_ENCCL;
            c->mass_full = true;
        }
        if (c->pos_lo) {
            const float4* src[2] = {c->pos_lo, c->vel_lo};
            for (int k = 0; k < 2; ++k) {
                if (!c->lo_full[k]) GS_CUDA(c, cudaMalloc((void**)&c->lo_full[k], (size_t)c->n_pad * sizeof(float4)));
                if (nc.AllGather(src[k], c->lo_full[k], (size_t)c->slice * 4, ncclFloat, c->comm, c->stream) != ncclSuccess)
                    return GS_ENCCL;
            }
            pos_lo_full = c->lo_full[0]; vel_lo_full = c->lo_full[1];
        }
    }
    const size_t bytes = (size_t)n * record_bytes;
    if ((rc = ensure_aos(c, bytes)) != GS_OK) return rc;
    GS_CUDA(c, launch_packed_to_aos(c->pos[c->cur], vel_full, 0, c->mass, 0, n, c->d_aos, record_bytes, pos_lo_full,
                                    vel_lo_full, c->stream));
    c->launches++;
    GS_CUDA(c, cudaMemcpyAsync(host, c->d_aos, bytes, cudaMemcpyDeviceToHost, c->stream));
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    return check_peers(c);
}

int gs_download_aos_f32(gs_ctx* ctx, void* particles28, int n) { return download_common(ctx, particles28, n, 28); }
int gs_download_aos_f64(gs_ctx* ctx, void* particles56, int n) { return download_common(ctx, particles56, n, 56); }

// This rank's own records only: nothing crosses between GPUs (positions, velocities and masses of the slice are local).
static int download_slice_common(gs_ctx* c, void* host, int first, int count, int record_bytes) {
    if (!c || (!host && count > 0)) return GS_EINVAL;
    if (!c->uploaded) return GS_ESTATE;
    const int lo = c->i_begin < c->n ? c->i_begin : c->n;
    const int hi = (c->i_begin + c->slice) < c->n ? (c->i_begin + c->slice) : c->n;
    if (first != lo || count != hi - lo) return GS_EINVAL;
    int rc = set_device(c);
    if (rc != GS_OK) return rc;
    if ((rc = sync_velocities(c)) != GS_OK) return rc;       // collective with GS_INTEGRATOR_LEAPFROG
    if (count > 0) {
        const size_t bytes = (size_t)count * record_bytes;
        if ((rc = ensure_aos(c, bytes)) != GS_OK) return rc;
        GS_CUDA(c, launch_packed_to_aos(c->pos[c->cur], c->vel, c->i_begin, c->mass, c->i_begin, count, c->d_aos,
                                        record_bytes, c->pos_lo, c->vel_lo, c->stream));
        c->launches++;
        GS_CUDA(c, cudaMemcpyAsync(host, c->d_aos, bytes, cudaMemcpyDeviceToHost, c->stream));
    }
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    return check_peers(c);
}

int gs_download_slice_f32(gs_ctx* ctx, void* my28, int first, int count) { return download_slice_common(ctx, my28, first, count, 28); }
int gs_download_slice_f64(gs_ctx* ctx, void* my56, int first, int count) { return download_slice_common(ctx, my56, first, count, 56); }

int gs_sync(gs_ctx* c) {
    if (!c) return GS_EINVAL;
    int rc = set_device(c);
    if (rc != GS_OK) return rc;
    GS_CUDA(c, cudaStreamSynchronize(c->stream));
    if (c->comm_stream) GS_CUDA(c, cudaStreamSynchronize(c->comm_stream));
    return check_peers(c);      // a peer never raised its flag: the wait kernel gave up instead of hanging the GPU
}

int gs_set_dt(gs_ctx* c, double dt) {
    if (!c || !std::isfinite(dt)) return GS_EINVAL;
    if (c->cfg.integrator == GS_INTEGRATOR_LEAPFROG && !c->vel_synced) {
        int rc = set_device(c);
        if (rc != GS_OK) return rc;
        if ((rc = sync_velocities(c)) != GS_OK) return rc;   // close the step with the old dt first
    }
    c->cfg.dt = dt;
    return GS_OK;
}

int gs_get_info(const gs_ctx* c, gs_info* info) {
    if (!c || !info) return GS_EINVAL;
    memset(info, 0, sizeof(*info));
    info->n = c->n; info->n_pad = c->n_pad;
    info->i_begin = c->i_begin < c->n ? c->i_begin : c->n;
    const int ie = c->i_begin + c->slice;
    info->i_end = ie < c->n ? ie : c->n;
    info->sm_count = c->sm_count; info->grid = c->grid;
    info->i_tile = c->i_tile; info->j_tile = kJTile;
    info->stages = kStages; info->variant = c->variant;
    info->regs_per_thread = c->regs; info->threads = c->threads;
    info->scratch_bytes = (int64_t)c->partial_bytes;
    info->peer_exchange = c->p2p ? 1 : 0;
    info->uniform_mass = c->mu_uniform != 0.f ? 1 : 0;
    info->symmetric = c->symmetric ? 1 : 0;
    return GS_OK;
}

uint64_t gs_kernel_launches(const gs_ctx* c) { return c ? c->launches : 0; }
int gs_last_cuda_error(const gs_ctx* c) { return c ? c->last_cuda_error : 0; }

}  // extern "C"
