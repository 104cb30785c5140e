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

#include <string.h>

namespace {

using b2::set_error;

constexpr int kPeerMaxWorld = 8;
constexpr int kPeerMaxElems = 4096;
constexpr size_t kPeerRegionBytes = size_t(2) * kPeerMaxWorld * kPeerMaxElems * sizeof(uint2);

struct PeerBases {
    unsigned char* base[kPeerMaxWorld];
};

__device__ __forceinline__ void st_pair(uint2* p, unsigned v, unsigned flag) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(v), "r"(flag) : "memory");
}
__device__ __forceinline__ uint2 ld_pair(const uint2* p) {
    uint2 v;
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(512) peer_allreduce_kernel(float* data, int n, PeerBases peers, int rank, int world,
                                                              unsigned epoch) {
    const size_t par = size_t(epoch & 1u) * kPeerMaxWorld * kPeerMaxElems;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const unsigned v = __float_as_uint(data[i]);
        for (int p = 1; p <= world; ++p) {                      // peers first, own region last
            const int q = (rank + p) % world;
            st_pair(reinterpret_cast<uint2*>(peers.base[q]) + par + size_t(rank) * kPeerMaxElems + i, v, epoch);
        }
    }
    const uint2* mine = reinterpret_cast<const uint2*>(peers.base[rank]) + par;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        float s = 0.f;
        for (int q = 0; q < world; ++q) {
            uint2 w;
            do {
                w = ld_pair(mine + size_t(q) * kPeerMaxElems + i);
            } while (w.y != epoch);
            s += __uint_as_float(w.x);
        }
        data[i] = s;
    }
}

}  // namespace

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

int b2_peer_allreduce_f32(float* data, int n, void* const* regions, int rank, int world, unsigned epoch, void* stream) {
    if (!data || !regions) return set_error("b2_peer_allreduce_f32: null pointer");
    if (world < 1 || world > kPeerMaxWorld || rank < 0 || rank >= world)
        return set_error("b2_peer_allreduce_f32: rank %d of %d (at most %d peers)", rank, world, kPeerMaxWorld);
    if (n < 0 || n > kPeerMaxElems) return set_error("b2_peer_allreduce_f32: %d elements exceed %d", n, kPeerMaxElems);
    if (n == 0) return 0;
    PeerBases pb;
    memset(&pb, 0, sizeof(pb));
    for (int p = 0; p < world; ++p) {
        if (!regions[p]) return set_error("b2_peer_allreduce_f32: region of rank %d is null", p);
        pb.base[p] = reinterpret_cast<unsigned char*>(regions[p]);
    }
    const int threads = n >= 512 ? 512 : ((n + 31) / 32) * 32;
    peer_allreduce_kernel<<<1, threads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(data, n, pb, rank, world, epoch);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return set_error("b2_peer_allreduce_f32 launch: %s", cudaGetErrorString(e));
    return 0;
}

}  // extern "C"
