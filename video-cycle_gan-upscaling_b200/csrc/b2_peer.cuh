// b2_peer.cuh — device side of the NVLink peer-memory exchange (see b2_peer.cu for the protocol): every rank owns a
// region stage[2][kPeerMaxWorld][kPeerMaxElems] of {value, epoch} words; a value is delivered to every peer with ONE
// aligned 8-byte store whose epoch half doubles as its valid flag.  Shared by the stand-alone all-reduce launch and by
// the BatchNorm finalize kernels (b2_bn_train.cu), which fold the exchange into the kernel that needs its result.
#pragma once
#include <stdint.h>

namespace b2 {

constexpr int kPeerMaxWorld = 8;
constexpr int kPeerMaxElems = 4096;
constexpr size_t kPeerRegionBytes = size_t(2) * kPeerMaxWorld * kPeerMaxElems * sizeof(uint2);

struct PeerBases {
    unsigned char* base[kPeerMaxWorld];
    int rank, world;      // world <= 1: no exchange
    unsigned epoch;
    // CUDA-graph replay: a captured launch bakes its arguments, so the epoch of a replayed exchange is epoch (the launch's
    // position inside the captured step, baked) + *epoch_base (a device word the host sets before every replay).  Null in
    // a plain launch.
    const unsigned* epoch_base;
};

// Host: validates and packs the caller's region table; world <= 1 or regions == null gives an inactive descriptor.
int make_peer_bases(void* const* regions, int rank, int world, unsigned epoch, PeerBases* out, const unsigned* epoch_base = nullptr);

#ifdef __CUDACC__
// First statement of every kernel that exchanges: resolve the epoch of a graph-replayed launch.
__device__ __forceinline__ void peer_resolve_epoch(PeerBases& pb) {
    if (pb.epoch_base) pb.epoch += *pb.epoch_base;
}
__device__ __forceinline__ void peer_st_pair(uint2* p, unsigned v, unsigned flag) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(v), "r"(flag) : "memory");
}
__device__ __forceinline__ uint2 peer_ld_pair(const uint2* p) {
    uint2 v;
    asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
    return v;
}
// Element i of this rank's vector goes to slot [parity][rank][i] of every region (peers first, own region last).
__device__ __forceinline__ void peer_put(const PeerBases& pb, int i, float x) {
    const size_t par = size_t(pb.epoch & 1u) * kPeerMaxWorld * kPeerMaxElems;
    const unsigned v = __float_as_uint(x);
    for (int p = 1; p <= pb.world; ++p) {
        const int q = (pb.rank + p) % pb.world;
        peer_st_pair(reinterpret_cast<uint2*>(pb.base[q]) + par + size_t(pb.rank) * kPeerMaxElems + i, v, pb.epoch);
    }
}
// Sum over the ranks of element i, in rank order (identical bits on every replica); spins until every peer delivered.
__device__ __forceinline__ float peer_get_sum(const PeerBases& pb, int i) {
    const uint2* mine = reinterpret_cast<const uint2*>(pb.base[pb.rank]) + size_t(pb.epoch & 1u) * kPeerMaxWorld * kPeerMaxElems;
    float s = 0.f;
    for (int q = 0; q < pb.world; ++q) {
        uint2 w;
        do {
            w = peer_ld_pair(mine + size_t(q) * kPeerMaxElems + i);
        } while (w.y != pb.epoch);
        s += __uint_as_float(w.x);
    }
    return s;
}
#endif

}  // namespace b2
