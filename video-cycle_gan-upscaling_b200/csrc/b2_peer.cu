// b2_peer.cu — one-shot sum-all-reduce of small fp32 vectors over NVLink peer memory, for the cross-replica
// BatchNormalization sums of the data-parallel training step (the reference trains on one GPU:
// BatchNormalization at model.py:20, 23, 284, 840-872 sees the whole batch, which the replicas reproduce by summing
// their per-channel sum / sum-of-squares, forward and backward, ~70 times per step).
//
// These reductions sit on the critical path (the next kernel needs the result) and carry at most a few KB, so their cost
// is latency, not bandwidth.  Every rank owns one region, allocated with cudaMalloc and opened by its peers through CUDA
// IPC:   stage[2][8][kPeerMaxElems] of {value, epoch} pairs (8 bytes each).
// One launch (a single CTA on the compute stream) of epoch e, parity p = e & 1, on rank r: thread i
//   1. stores {x[i], e} as ONE 8-byte word into stage[p][r][i] of every rank's region (remote stores are fire-and-forget
//      over NVLink; an aligned 8-byte store arrives whole, so the epoch doubles as the "valid" flag of its value - the
//      flagged-word scheme of low-latency collectives: no fence, no separate flag, one one-way trip),
//   2. spins on its own stage[p][q][i] for q = 0..world-1 until the epoch matches, summing the values in rank order - the
//      same order on every rank, so all replicas get identical bits.
// Two parities suffice: a rank reaches epoch e+2 only after every peer has delivered e+1, i.e. has finished reading e.
#include "b2_host.h"
#include "b2_peer.cuh"

#include <string.h>

namespace {

using b2::set_error;
using b2::kPeerMaxElems;
using b2::kPeerMaxWorld;
using b2::kPeerRegionBytes;

__global__ void __launch_bounds__(512) peer_allreduce_kernel(float* data, int n, b2::PeerBases pb) {
    b2::pdl_entry();
    b2::peer_resolve_epoch(pb);
    for (int i = threadIdx.x; i < n; i += blockDim.x) b2::peer_put(pb, i, data[i]);
    for (int i = threadIdx.x; i < n; i += blockDim.x) data[i] = b2::peer_get_sum(pb, i);
}

// `iters` dependent exchanges in one launch (tools/peer_probe.py): what one exchange costs once launch latency, the host
// and the skew between the ranks' streams are out of the picture.
__global__ void __launch_bounds__(512) peer_probe_kernel(float* data, int n, b2::PeerBases pb, int iters, float inv_world,
                                                         long long* cycles) {
    b2::pdl_entry();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        for (int i = threadIdx.x; i < n; i += blockDim.x) b2::peer_put(pb, i, data[i]);
        for (int i = threadIdx.x; i < n; i += blockDim.x) data[i] = b2::peer_get_sum(pb, i) * inv_world;
        pb.epoch = pb.epoch % 0x7FFFFFFEu + 1u;
    }
    __syncthreads();
    if (threadIdx.x == 0) cycles[0] = clock64() - t0;
}

}  // namespace

namespace b2 {
int make_peer_bases(void* const* regions, int rank, int world, unsigned epoch, PeerBases* out, const unsigned* epoch_base) {
    memset(out, 0, sizeof(*out));
    out->rank = 0; out->world = 1; out->epoch = epoch; out->epoch_base = epoch_base;
    if (!regions || world <= 1) {
        if (regions && world == 1 && regions[0]) out->base[0] = reinterpret_cast<unsigned char*>(regions[0]);
        return 0;
    }
    if (world > kPeerMaxWorld || rank < 0 || rank >= world)
        return set_error("peer exchange: rank %d of %d (at most %d peers)", rank, world, kPeerMaxWorld);
    for (int p = 0; p < world; ++p) {
        if (!regions[p]) return set_error("peer exchange: region of rank %d is null", p);
        out->base[p] = reinterpret_cast<unsigned char*>(regions[p]);
    }
    out->rank = rank; out->world = world;
    return 0;
}
}  // namespace b2

extern "C" {

int b2_peer_max_elems(void) { return kPeerMaxElems; }

int b2_peer_alloc(void** region, unsigned char* handle64) {
    if (!region || !handle64) return set_error("b2_peer_alloc: null pointer");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, kPeerRegionBytes);
    if (e != cudaSuccess) return set_error("b2_peer_alloc: cudaMalloc: %s", cudaGetErrorString(e));
    e = cudaMemset(p, 0, kPeerRegionBytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        return set_error("b2_peer_alloc: %s", cudaGetErrorString(e));
    }
    memcpy(handle64, &h, 64);
    *region = p;
    return 0;
}

int b2_peer_open(const unsigned char* handle64, void** region) {
    if (!region || !handle64) return set_error("b2_peer_open: null pointer");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return set_error("b2_peer_open: cudaIpcOpenMemHandle: %s", cudaGetErrorString(e));
    *region = p;
    return 0;
}

int b2_peer_close(void* region, int opened) {
    if (!region) return 0;
    cudaError_t e = opened ? cudaIpcCloseMemHandle(region) : cudaFree(region);
    if (e != cudaSuccess) return set_error("b2_peer_close: %s", cudaGetErrorString(e));
    return 0;
}

int b2_peer_exchange_probe(float* data, int n, void* const* regions, int rank, int world, unsigned epoch0, int iters,
                           long long* cycles, void* stream) {
    if (!data || !regions || !cycles) return set_error("b2_peer_exchange_probe: null pointer");
    if (n < 1 || n > kPeerMaxElems || iters < 1) return set_error("b2_peer_exchange_probe: bad sizes");
    b2::PeerBases pb;
    if (int rc = b2::make_peer_bases(regions, rank, world, epoch0, &pb)) return rc;
    const int threads = n >= 512 ? 512 : ((n + 31) / 32) * 32;
    b2::launch_k(peer_probe_kernel, 1, threads, 0, reinterpret_cast<cudaStream_t>(stream), data, n, pb, iters, 1.0f / float(pb.world), cycles);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("b2_peer_exchange_probe launch: %s", cudaGetErrorString(e));
    return 0;
}

int b2_peer_allreduce_f32(float* data, int n, void* const* regions, int rank, int world, unsigned epoch, const unsigned* epoch_base,
                          void* stream) {
    if (!data || !regions) return set_error("b2_peer_allreduce_f32: null pointer");
    if (n < 0 || n > kPeerMaxElems) return set_error("b2_peer_allreduce_f32: %d elements exceed %d", n, kPeerMaxElems);
    b2::PeerBases pb;
    if (world < 1) return set_error("b2_peer_allreduce_f32: rank %d of %d (at most %d peers)", rank, world, kPeerMaxWorld);
    if (int rc = b2::make_peer_bases(regions, rank, world, epoch, &pb, epoch_base)) return rc;
    pb.world = world;      // world == 1 still runs the exchange through the own region (identity; exercised by the tests)
    if (n == 0) return 0;
    const int threads = n >= 512 ? 512 : ((n + 31) / 32) * 32;
    b2::launch_k(peer_allreduce_kernel, 1, threads, 0, reinterpret_cast<cudaStream_t>(stream), data, n, pb);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("b2_peer_allreduce_f32 launch: %s", cudaGetErrorString(e));
    return 0;
}

}  // extern "C"
