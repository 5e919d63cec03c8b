// gs_noalloc.cu -- gs_potential_energy_noalloc(): the potential-energy pass for callers that are
// not allowed to allocate GPU memory.
//
// Who needs this: code that runs where cudaMalloc/cudaFree are off limits -- a monitoring hook
// inside somebody else's time loop whose device memory is handed out once by a pool (a framework's
// caching allocator, a job that pre-partitions the whole GPU), a latency-critical loop that cannot
// take cudaMalloc's implicit device-wide synchronisation, or a process running so close to the
// memory limit that an extra allocation may fail.  gs_energy() needs a gs_ctx, and gs_create()
// allocates; this entry point needs neither.
//
// How: the per-CTA partial sums that the potential kernel leaves in global memory live in ONE
// file-scope __device__ array of fixed size instead of a cudaMalloc'ed buffer.  Consequences the
// caller has to know:
//   * the array is part of the library image: every GPU that loads libgravsim.so carries its
//     128 KiB until the process exits, used or not;
//   * it bounds the problem: at most GS_NOALLOC_MAX_BODIES (4,194,304) padded bodies, beyond
//     that the call returns GS_ETOOBIG and does nothing;
//   * all calls on one GPU share it.  Two calls whose kernels may overlap on the GPU (different
//     streams, or different host threads) race on it and give wrong sums -- serialise them.
//     Nothing in it has to be zero beforehand: each call overwrites exactly the entries it reads
//     back, so leftovers from an earlier (larger) call are never seen;
//   * the first call on each device looks the array's device address up (cudaGetSymbolAddress)
//     and remembers it.  Make that first call during set-up, before starting any stream capture;
//   * "each device" means the CURRENT device of the calling thread: the scratch used is the copy
//     that lives on it, so pos_packed_dev, potential_dev and the stream must all belong to the
//     current device (cudaSetDevice first when the process drives several GPUs).  The call checks
//     this for the two pointers and refuses (GS_EINVAL) rather than touch another GPU's array.
// The call itself only launches two kernels on the given stream: no cudaMalloc, no cudaFree, no
// library routine that allocates, no host synchronisation.
#include <mutex>

#include "../../include/gravsim.h"
#include "gs_kernels.h"

namespace {

constexpr int kPartialBodies = 256;   // bodies per CTA of the potential kernel = one partial each
constexpr int kMaxPartials = GS_NOALLOC_MAX_BODIES / kPartialBodies;   // 16384 doubles = 128 KiB
constexpr int kMaxDevices = 64;

// Scratch shared by every call on this GPU.  In host code this NAME is only a host-side handle of
// the symbol, not a device address: the address is fetched per device below and passed to the
// kernels as an ordinary pointer argument.
static __device__ __align__(256) double g_noalloc_partials[kMaxPartials];

std::mutex g_addr_mutex;
double* g_partials_addr[kMaxDevices];   // per device ordinal; nullptr until that device's first call

// Sums partials[0..count) in a fixed order (bit-reproducible) and writes scale * sum.
__global__ void __launch_bounds__(256) noalloc_reduce_kernel(const double* __restrict__ partials, int count,
                                                             double scale, double* __restrict__ out) {
    __shared__ double sh[256];
    double v = 0.0;
    for (int k = threadIdx.x; k < count; k += 256) v += partials[k];
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if (threadIdx.x < w) sh[threadIdx.x] += sh[threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[0] = scale * sh[0];
}

// Device address of the scratch array on the CURRENT device; resolved once per device.
int scratch_address(double** out) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return GS_ENOGPU;
    if (dev < 0 || dev >= kMaxDevices) return GS_EINVAL;
    std::lock_guard<std::mutex> lock(g_addr_mutex);
    if (g_partials_addr[dev] == nullptr) {
        double* p = nullptr;
        if (cudaGetSymbolAddress((void**)&p, g_noalloc_partials) != cudaSuccess || p == nullptr) {
            cudaGetLastError();
            return GS_ECUDA;
        }
        g_partials_addr[dev] = p;
    }
    *out = g_partials_addr[dev];
    return GS_OK;
}

}  // namespace

extern "C" int gs_potential_energy_noalloc(const void* pos_packed_dev, int n_pad, double G, double eps,
                                           double* potential_dev, void* stream) {
    if (!pos_packed_dev || !potential_dev || n_pad <= 0 || !(G > 0.0) || !(eps >= 0.0)) return GS_EINVAL;
    if (n_pad % gs::kJTile != 0) return GS_EINVAL;
    if (n_pad > GS_NOALLOC_MAX_BODIES) return GS_ETOOBIG;      // would overrun the fixed scratch
    double* scratch = nullptr;
    const int rc = scratch_address(&scratch);
    if (rc != GS_OK) return rc;
    // the scratch just resolved lives on the current device: both buffers have to live there too
    int dev = -1;
    cudaGetDevice(&dev);
    const void* ptrs[2] = {pos_packed_dev, potential_dev};
    for (const void* q : ptrs) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, q) != cudaSuccess) { cudaGetLastError(); return GS_EINVAL; }
        if (at.type != cudaMemoryTypeDevice || at.device != dev) return GS_EINVAL;
    }
    cudaStream_t s = (cudaStream_t)stream;
    const float eps2 = (float)(eps * eps);
    const float cut2 = (eps > 0.0) ? 0.f : 1e-20f;             // same pair rule as gs_energy
    if (gs::launch_potential((const float4*)pos_packed_dev, n_pad, 0, n_pad, eps2, cut2, scratch, s) != cudaSuccess)
        return GS_ECUDA;
    noalloc_reduce_kernel<<<1, 256, 0, s>>>(scratch, n_pad / kPartialBodies, -0.5 / G, potential_dev);
    return cudaPeekAtLastError() == cudaSuccess ? GS_OK : GS_ECUDA;     // this launch only; earlier sticky errors are the caller's
}
