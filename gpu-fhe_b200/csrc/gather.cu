// The path's one collective -- the all-gather of result limbs to every GPU of the box -- done by the library itself with
// NVLink peer stores instead of a separate NCCL call: a rank's result block is pushed into the same position of every
// peer's full [L, B, N] array by one kernel (16-byte coalesced stores on peer-mapped pointers), which can be enqueued limb
// by limb behind the kernels that produce each limb, on the stream of the caller's choice.  No kernel ever waits for
// another GPU: completion is established by the caller with one stream-ordered collective on a flag after the pushes
// (gpu-fhe_b200/shard.py, PeerGather.finish), so nothing here can hang on a peer.
//
// Peer memory is plain cudaMalloc memory exported with CUDA IPC handles (one process per GPU, as torchrun launches them).
// The reference has no collective at all; the north-star asks for "the final all-gather of limbs
// back to the caller's device".
#include "../../include/gpufhe_b200.h"
#include <cstdlib>
#include <cstring>
#include "plan.h"

namespace gfhe {

struct PeerPtrs { void* p[GFHE_MAX_PEERS]; };

// Each thread moves four 16-byte words per trip: four loads in flight, then the same four words to every destination.
__global__ void __launch_bounds__(256)
k_push_gather(const uint4* __restrict__ src, PeerPtrs dst, u32 ndst, size_t n16, size_t off16) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n16; i += 4 * stride) {
        const uint4 v0 = __ldcs(src + i), v1 = __ldcs(src + i + stride), v2 = __ldcs(src + i + 2 * stride), v3 = __ldcs(src + i + 3 * stride);
        for (u32 d = 0; d < ndst; ++d) {
            uint4* o = reinterpret_cast<uint4*>(dst.p[d]) + off16;
            o[i] = v0; o[i + stride] = v1; o[i + 2 * stride] = v2; o[i + 3 * stride] = v3;
        }
    }
    for (; i < n16; i += stride) {
        const uint4 v = __ldcs(src + i);
        for (u32 d = 0; d < ndst; ++d) (reinterpret_cast<uint4*>(dst.p[d]) + off16)[i] = v;
    }
}

}  // namespace gfhe

using namespace gfhe;

extern "C" {

int gfhe_peer_alloc(int device, size_t bytes, void** d_ptr, unsigned char* handle64) {
    if (!d_ptr || !handle64 || bytes == 0) return GFHE_EARG;
    DeviceGuard g(device);
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) return (int)e;
    cudaIpcMemHandle_t h;
    if ((e = cudaIpcGetMemHandle(&h, p)) != cudaSuccess) { cudaFree(p); return (int)e; }
    static_assert(sizeof(h) == 64, "CUDA IPC handle size");
    memcpy(handle64, &h, 64);
    *d_ptr = p;
    return 0;
}

int gfhe_peer_open(int device, const unsigned char* handle64, void** d_ptr) {
    if (!d_ptr || !handle64) return GFHE_EARG;
    DeviceGuard g(device);
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return (int)e;
    *d_ptr = p;
    return 0;
}

int gfhe_peer_close(int device, void* d_ptr) {
    if (!d_ptr) return 0;
    DeviceGuard g(device);
    return (int)cudaIpcCloseMemHandle(d_ptr);
}

int gfhe_peer_free(int device, void* d_ptr) {
    if (!d_ptr) return 0;
    DeviceGuard g(device);
    return (int)cudaFree(d_ptr);
}

int gfhe_push_gather(int device, const void* d_src, size_t bytes, void* const* dst_bases, uint32_t ndst, size_t dst_offset_bytes, uint32_t ctas,
                     void* stream) {
    if (!d_src || !dst_bases || ndst == 0 || ndst > GFHE_MAX_PEERS) return GFHE_EARG;
    if (bytes == 0) return 0;
    if ((bytes & 15u) || (dst_offset_bytes & 15u) || (reinterpret_cast<uintptr_t>(d_src) & 15u)) return GFHE_EALIGN;
    PeerPtrs pp{};
    for (uint32_t d = 0; d < ndst; ++d) {
        if (!dst_bases[d] || (reinterpret_cast<uintptr_t>(dst_bases[d]) & 15u)) return GFHE_EALIGN;
        pp.p[d] = dst_bases[d];
    }
    DeviceGuard g(device);
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const size_t n16 = bytes / 16;
    // A copy kernel whose stores wait on NVLink.  Measured on 8 GPUs, 32 MiB per rank to 7 peers (profiles/r02_run15_push_probe_n8.txt):
    // alone, one CTA per SM is fastest (0.372 ms; 0.404 with 32 CTAs, 0.416 with 592); pushed limb by limb UNDER the next limb's
    // kernels, 32 CTAs are (0.474 ms per step against 0.550 with 148) -- every resident copy CTA takes an SM slot away from
    // the transforms.  `ctas` = 0: one per SM; the caller passes 32 for pushes it overlaps.  GFHE_PUSH_CTAS overrides both.
    static const int ctas_env = [] { const char* env = std::getenv("GFHE_PUSH_CTAS"); return env ? std::atoi(env) : 0; }();
    size_t cap = ctas_env > 0 ? (size_t)ctas_env : (ctas ? (size_t)ctas : (size_t)sms);
    size_t blocks = (n16 + 256 * 4 - 1) / (256 * 4);
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    k_push_gather<<<(unsigned)blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(static_cast<const uint4*>(d_src), pp, ndst, n16, dst_offset_bytes / 16);
    return (int)cudaGetLastError();
}

}  // extern "C"
