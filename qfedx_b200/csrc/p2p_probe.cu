// NVLink bandwidth probes run next to the fused swap (K6) so that its roofline is a number measured in the same job,
// on the same GPUs, with the same access pattern - not a quoted figure.  Plain 16-byte loads / stores on mapped peer
// pointers (symmetric heap), no multicast, no flags: the host orders the launches (barrier + synchronize around the
// timed loop) and the data is never interpreted, so concurrent probes on different GPUs need nothing from each other.
//   mode 0  ring pull      every rank reads the whole shard of rank+1          (one-directional peer copy per link pair)
//   mode 1  all-to-all pull  dst[(s << sb) + j] = src_s[(rank << sb) + j]      (the swap's access pattern, no gates)
//   mode 2  all-to-all push  dst_s[(rank << sb) + j] = src[(s << sb) + j]      (same exchange with stores)
// No reference counterpart: /root/reference has no communication layer.
#include "common.cuh"
#include "../../include/qfedx_b200.h"

namespace qfx {

struct ProbeArgs {
    const float4* const* src;   // [world] every rank's source shard
    float4* const* dst;         // [world] every rank's destination shard
    size_t n_vec;               // float4 elements per shard
    int rank, world, mode;
};

template <int UNROLL>
__global__ void __launch_bounds__(256) p2p_probe_kernel(const ProbeArgs a) {
    const size_t tid = (size_t)blockIdx.x * 256 + threadIdx.x, stride = (size_t)gridDim.x * 256;
    if (a.mode == 0) {
        const float4* __restrict__ src = a.src[(a.rank + 1) % a.world];
        float4* __restrict__ dst = a.dst[a.rank];
        for (size_t i = tid; i < a.n_vec; i += stride * UNROLL) {
            float4 v[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) if (i + u * stride < a.n_vec) v[u] = src[i + u * stride];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) if (i + u * stride < a.n_vec) dst[i + u * stride] = v[u];
        }
        return;
    }
    const size_t slice = a.n_vec / a.world;
    for (size_t j = tid; j < slice; j += stride) {
        // peers are visited starting from the next rank so that at any moment the W ranks aim at W different targets
        for (int t0 = 0; t0 < a.world; t0 += UNROLL) {
            float4 v[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const int s = (a.rank + 1 + t0 + u) % a.world;
                if (t0 + u < a.world)
                    v[u] = a.mode == 1 ? a.src[s][(size_t)a.rank * slice + j] : a.src[a.rank][(size_t)s * slice + j];
            }
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const int s = (a.rank + 1 + t0 + u) % a.world;
                if (t0 + u < a.world) {
                    if (a.mode == 1) a.dst[a.rank][(size_t)s * slice + j] = v[u];
                    else a.dst[s][(size_t)a.rank * slice + j] = v[u];
                }
            }
        }
    }
}

}  // namespace qfx

using namespace qfx;

extern "C" int qfx_p2p_probe(const float* const* src_table, float* const* dst_table, size_t shard_bytes, int rank,
                             int world, int mode, int unroll, int ctas_per_sm, cudaStream_t stream) {
    QFX_ARG_CHECK(src_table && dst_table && shard_bytes % (16 * (size_t)world) == 0 && world >= 2 && world <= 16 &&
                  rank >= 0 && rank < world && mode >= 0 && mode <= 2, QFX_ERR_BAD_ARG, "bad probe arguments");
    ProbeArgs a;
    a.src = reinterpret_cast<const float4* const*>(src_table);
    a.dst = reinterpret_cast<float4* const*>(dst_table);
    a.n_vec = shard_bytes / 16; a.rank = rank; a.world = world; a.mode = mode;
    const int blocks = device_sm_count() * (ctas_per_sm > 0 ? ctas_per_sm : 4);
    if (unroll >= 8) p2p_probe_kernel<8><<<blocks, 256, 0, stream>>>(a);
    else if (unroll >= 4) p2p_probe_kernel<4><<<blocks, 256, 0, stream>>>(a);
    else if (unroll >= 2) p2p_probe_kernel<2><<<blocks, 256, 0, stream>>>(a);
    else p2p_probe_kernel<1><<<blocks, 256, 0, stream>>>(a);
    return check_launch("p2p_probe_kernel");
}
