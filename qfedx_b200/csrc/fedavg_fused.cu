// K7: per-round federated aggregation in ONE cooperative launch.
//
//   delta_k = wrap(theta_k - theta_g)            (angle entries wrapped to [-pi, pi))
//   s_k     = min(1, C / ||delta_k||_2)           per-client L2 clip
//   v_k     = w_k/sum(w) * (s_k * delta_k + sigma*C * N(0, I))
//   y_k     = fixed_point(v_k) + sum_{j>k} m_kj - sum_{j<k} m_jk      (pairwise masks, mod 2^64)
//   theta_g <- theta_g + server_lr * decode(sum_k y_k)                 (masks cancel exactly)
//
// Restates /root/reference/src/CFed/Classical_FL.py:66-81 (FedAvg by sample count; "weights"
// mode reproduces it: theta_g <- sum_k w_k theta_k / sum w) and implements the planned DP
// clip + Gaussian noise and pairwise-mask secure aggregation of /root/reference/ROADMAP.md:45-62,
// 137-141.  Phase 1 reduces per-client squared norms into scratch, a grid barrier follows,
// phase 2 does everything else per element.  With world > 1 the clients of this launch are one
// rank's share: the rank's partial sum is PUSHED into every peer's receive slot with plain
// stores over NVLink (peer pointers from the symmetric heap), a system-scope flag is released
// per peer, and after acquiring all peers' flags every rank sums the slots and applies the
// identical server update - an all-reduce with no NCCL call.
#include <cooperative_groups.h>
#include "common.cuh"
#include "../../include/qfedx_b200.h"

namespace cg = cooperative_groups;

namespace qfx {

constexpr int AGG_THREADS = 256;
constexpr int AGG_MAX_WORLD = 16;
constexpr double FIXED_SCALE = 4294967296.0;  // 2^32
constexpr uint64_t TAG_NOISE = 0x4e4f495345ull;
constexpr uint64_t TAG_MASK = 0x4d41534b53ull;

struct AggArgs {
    const float* client_params;  // [K_local, D]
    const float* global_in;      // [D]
    float* global_out;           // [D]
    const float* weights;        // [K_global] sample counts
    const int* participate;      // [K_global] 0/1 or null (all participate)
    const int* client_gids;      // [K_local] global client id of each local row, null = identity
    int k_local, k_global, dim;
    int mode;                    // 0 weights (no clip/noise), 1 delta
    int n_wrap;                  // first n_wrap entries of delta are angles to wrap
    float clip, sigma, server_lr;
    unsigned long long seed;
    int round;
    int secagg;
    float* norms_out;            // [K_local] pre-clip norms, nullable
    float* scratch;              // [K_local] squared-norm accumulators (zeroed by the launcher)
    long long* masked_out;       // [K_local, D] masked fixed-point messages, nullable (inspection)
    const unsigned long long* pair_keys;   // [K_global, K_global] pair secrets from the key agreement (entry lo*K+hi), or
                                           // null: every pair's mask stream is derived from the job seed
    // multi-GPU
    int rank, world;
    long long* const* recv_slots;  // [world] device pointers: recv_slots[r] = base of rank r's slot array
                                   // [world][dim] int64 (secagg) or viewed as float (plain)
    unsigned int* const* flags;    // [world] device pointers: flags[r] = rank r's flag array [world]
    unsigned int epoch;            // value written into flags this round (monotonic)
    int two_shot;                  // 0: every rank pushes its partial to every peer and reduces W slots locally (one
                                   //    barrier; best for latency-bound sizes).  1: reduce-scatter by peer stores during the
                                   //    sweep over the client rows (slice o of the partial goes straight into rank o's
                                   //    buffer), local reduce + update of the own slice, then every rank gathers the other
                                   //    slices by peer loads (two barriers, 2 (W-1)/W D words over NVLink per rank
                                   //    instead of (W-1) D; buffer: (W + 1) * ceil(D / W) words per rank).  2: NVLS -
                                   //    the partial stays in this rank's own multicast-bound buffer, the owner of a slice
                                   //    gets the sum of all W partials from the switch (multimem.ld_reduce) and broadcasts
                                   //    the updated slice with one multimem.st: (W-1)/W D + D/W words per direction and
                                   //    rank, half of the peer-store exchange, and no reduction work on the SMs
    void* nvls_partial;            // this rank's partial row [dim] (float, or int64 with secagg): unicast address
    const void* nvls_partial_mc;   // multicast address of the same buffer (every rank's copy behind one address)
    float* nvls_out;               // [dim] result row in the multicast-bound buffer: unicast address
    float* nvls_out_mc;            // its multicast address
};

// Multicast (NVLS) accesses: the address names the same offset of every rank's buffer; the switch sums the W copies
// on a load and replicates a store to all of them.
__device__ __forceinline__ void mc_ld_reduce_f32x4(const float* mc, float (&x)[4]) {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(x[0]), "=f"(x[1]), "=f"(x[2]), "=f"(x[3]) : "l"(mc) : "memory");
}
__device__ __forceinline__ float mc_ld_reduce_f32(const float* mc) {
    float v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long mc_ld_reduce_u64(const long long* mc) {
    unsigned long long v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u64 %0, [%1];" : "=l"(v) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mc_st_f32x4(float* mc, const float (&x)[4]) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 ::"l"(mc), "f"(x[0]), "f"(x[1]), "f"(x[2]), "f"(x[3]) : "memory");
}
__device__ __forceinline__ void mc_st_f32(float* mc, float v) {
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(mc), "f"(v) : "memory");
}

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ float delta_of(const AggArgs& a, float x, int d, float g) {
    if (a.mode == 0) return x;
    x -= g;
    return d < a.n_wrap ? wrap_pi(x) : x;
}

constexpr int AGG_KCHUNK = 8;   // clients whose norms are accumulated in registers in one sweep

// Dynamic shared memory: [mask keys u64 k_local*k_global (optional)] [noise keys u64 k_local]
// [coef f32 k_local] [wnoise f32 k_local] [gid i32 k_local] [participates u8 k_global]
__host__ __device__ inline size_t agg_smem_bytes(int k_local, int k_global, bool mask_keys) {
    size_t b = (mask_keys ? (size_t)k_local * k_global * 8 : 0) + (size_t)k_local * (8 + 4 + 4 + 4) + (size_t)k_global;
    return (b + 15) & ~(size_t)15;
}

// VEC consecutive floats per access: one 16-byte transaction when the row pitch allows it, else scalars.
template <int VEC> struct VecIO;
template <> struct VecIO<1> {
    static __device__ __forceinline__ void ldg(const float* p, float (&x)[1]) { x[0] = __ldg(p); }
    static __device__ __forceinline__ void ld(const float* p, float (&x)[1]) { x[0] = *p; }
    static __device__ __forceinline__ void ldcv(const float* p, float (&x)[1]) { x[0] = __ldcv(p); }
    static __device__ __forceinline__ void ldcv(const long long* p, long long (&x)[1]) { x[0] = __ldcv(p); }
    static __device__ __forceinline__ void st(float* p, const float (&x)[1]) { *p = x[0]; }
    static __device__ __forceinline__ void st(long long* p, const unsigned long long (&x)[1]) { *p = (long long)x[0]; }
};
template <> struct VecIO<4> {
    static __device__ __forceinline__ void ldg(const float* p, float (&x)[4]) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(p));
        x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
    }
    static __device__ __forceinline__ void ld(const float* p, float (&x)[4]) {
        const float4 v = *reinterpret_cast<const float4*>(p);
        x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
    }
    static __device__ __forceinline__ void ldcv(const float* p, float (&x)[4]) {
        const float4 v = __ldcv(reinterpret_cast<const float4*>(p));
        x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
    }
    static __device__ __forceinline__ void ldcv(const long long* p, long long (&x)[4]) {
        const longlong2 a = __ldcv(reinterpret_cast<const longlong2*>(p));
        const longlong2 b = __ldcv(reinterpret_cast<const longlong2*>(p) + 1);
        x[0] = a.x; x[1] = a.y; x[2] = b.x; x[3] = b.y;
    }
    static __device__ __forceinline__ void st(float* p, const float (&x)[4]) {
        *reinterpret_cast<float4*>(p) = make_float4(x[0], x[1], x[2], x[3]);
    }
    static __device__ __forceinline__ void st(long long* p, const unsigned long long (&x)[4]) {
        reinterpret_cast<longlong2*>(p)[0] = make_longlong2((long long)x[0], (long long)x[1]);
        reinterpret_cast<longlong2*>(p)[1] = make_longlong2((long long)x[2], (long long)x[3]);
    }
};

// VEC: floats per access (4 needs dim % 4 == 0 and 16-byte aligned bases).  UNR: independent accesses per
// thread per trip (with the one-client-ahead prefetch in phase 2: 2 x UNR x VEC x 4 bytes in flight).
template <int VEC, int UNR, bool FIXED>
__device__ __forceinline__ void fedavg_fused_body(const AggArgs& a, const int mask_keys_in_smem) {
    using IO = VecIO<VEC>;
    cg::grid_group grid = cg::this_grid();
    extern __shared__ __align__(16) unsigned char agg_smem[];
    uint64_t* mkey = reinterpret_cast<uint64_t*>(agg_smem);
    uint64_t* nkey = mkey + (mask_keys_in_smem ? (size_t)a.k_local * a.k_global : 0);
    float* coef = reinterpret_cast<float*>(nkey + a.k_local);   // w_k/sum(w) * clip scale, 0 for non-participants
    float* wnoise = coef + a.k_local;                            // w_k/sum(w) * sigma * C
    int* gid = reinterpret_cast<int*>(wnoise + a.k_local);
    unsigned char* gpart = reinterpret_cast<unsigned char*>(gid + a.k_local);
    __shared__ float red[AGG_KCHUNK][AGG_THREADS / 32];
    const int tid = threadIdx.x;
    const int nvec = a.dim / VEC;                                // VEC == 4 only when dim % 4 == 0
    const int gstride = gridDim.x * AGG_THREADS;                 // in vectors
    const int g0 = blockIdx.x * AGG_THREADS + tid;
    const bool need_norm = a.mode == 1 && (a.clip > 0.f || a.norms_out != nullptr);
    const float noise_std = (a.mode == 1 && a.clip > 0.f) ? a.sigma * a.clip : 0.f;

    // ---- per-block tables that do not depend on the norms: participation, ids, RNG keys ----------
    for (int j = tid; j < a.k_global; j += AGG_THREADS) gpart[j] = (a.participate == nullptr || a.participate[j]) ? 1 : 0;
    for (int k = tid; k < a.k_local; k += AGG_THREADS) {
        const int gk = a.client_gids ? a.client_gids[k] : k;
        gid[k] = gk;
        nkey[k] = key3(a.seed ^ TAG_NOISE, (uint64_t)a.round, (uint64_t)gk);
    }
    if (FIXED && mask_keys_in_smem) {
        for (int i = tid; i < a.k_local * a.k_global; i += AGG_THREADS) {
            const int k = i / a.k_global, j = i - k * a.k_global;
            const int gk = a.client_gids ? a.client_gids[k] : k;
            const int lo = min(gk, j), hi = max(gk, j);
            const uint64_t pid = (uint64_t)lo * (uint64_t)a.k_global + (uint64_t)hi;
            mkey[i] = key3((a.pair_keys ? a.pair_keys[pid] : a.seed) ^ TAG_MASK, (uint64_t)a.round, pid);
        }
    }

    // ---- phase 1: squared norms, AGG_KCHUNK clients per sweep over the vector ---------------------
    if (need_norm) {
        for (int kc = 0; kc < a.k_local; kc += AGG_KCHUNK) {
            const int kn = min(AGG_KCHUNK, a.k_local - kc);
            float acc[AGG_KCHUNK];
#pragma unroll
            for (int kk = 0; kk < AGG_KCHUNK; ++kk) acc[kk] = 0.f;
            for (int v0 = g0; v0 < nvec; v0 += UNR * gstride) {
                float g[UNR][VEC];
#pragma unroll
                for (int u = 0; u < UNR; ++u) {
                    const int v = min(v0 + u * gstride, nvec - 1);   // clamped lanes re-read a valid vector
                    IO::ld(a.global_in + (size_t)v * VEC, g[u]);
                }
#pragma unroll
                for (int kk = 0; kk < AGG_KCHUNK; ++kk) {
                    if (kk < kn) {
                        const float* row = a.client_params + (size_t)(kc + kk) * a.dim;
                        float x[UNR][VEC];
#pragma unroll
                        for (int u = 0; u < UNR; ++u) {
                            const int v = min(v0 + u * gstride, nvec - 1);
                            IO::ldg(row + (size_t)v * VEC, x[u]);
                        }
#pragma unroll
                        for (int u = 0; u < UNR; ++u) {
                            const int v = v0 + u * gstride;
                            if (v < nvec) {
#pragma unroll
                                for (int e = 0; e < VEC; ++e) {
                                    const float dl = delta_of(a, x[u][e], v * VEC + e, g[u][e]);
                                    acc[kk] += dl * dl;
                                }
                            }
                        }
                    }
                }
            }
#pragma unroll
            for (int kk = 0; kk < AGG_KCHUNK; ++kk) {
                const float v = warp_sum(acc[kk]);
                if ((tid & 31) == 0) red[kk][tid >> 5] = v;
            }
            __syncthreads();
            if (tid < 32) {
                for (int kk = 0; kk < kn; ++kk) {
                    float v = tid < AGG_THREADS / 32 ? red[kk][tid] : 0.f;
                    v = warp_sum(v);
                    if (tid == 0 && v != 0.f) atomicAdd(a.scratch + kc + kk, v);
                }
            }
            __syncthreads();
        }
        grid.sync();
    }

    // ---- per-client coefficients -----------------------------------------------------------
    __syncthreads();
    float wsum = 0.f;
    for (int j = 0; j < a.k_global; ++j)
        if (gpart[j]) wsum += a.weights[j];
    const float inv_wsum = wsum > 0.f ? 1.f / wsum : 0.f;
    for (int k = tid; k < a.k_local; k += AGG_THREADS) {
        const int gk = gid[k];
        float c = gpart[gk] ? a.weights[gk] * inv_wsum : 0.f;
        wnoise[k] = a.weights[gk] * inv_wsum * noise_std;
        if (need_norm) {
            const float nrm = sqrtf(a.scratch[k]);
            if (blockIdx.x == 0 && a.norms_out) a.norms_out[k] = nrm;
            if (a.clip > 0.f && nrm > a.clip) c *= a.clip / nrm;
        }
        coef[k] = c;
    }
    __syncthreads();

    // ---- phase 2: weighted, noised, masked sum per element -----------------------------------
    // two-shot exchange: every rank starts its sweep at the slice of the NEXT rank, so that at any moment the W ranks
    // push to W different owners instead of all hammering rank 0's links first
    const int rot = (a.world > 1 && a.two_shot == 1) ? min(nvec - 1, ((a.rank + 1) % a.world) * ((nvec + a.world - 1) / a.world)) : 0;
    for (int v0 = g0; v0 < nvec; v0 += UNR * gstride) {
        float g[UNR][VEC], facc[UNR][VEC];
        unsigned long long iacc[UNR][VEC];
        int vi[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            vi[u] = min(v0 + u * gstride, nvec - 1);
            if (rot) { vi[u] += rot; if (vi[u] >= nvec) vi[u] -= nvec; }
            IO::ld(a.global_in + (size_t)vi[u] * VEC, g[u]);
#pragma unroll
            for (int e = 0; e < VEC; ++e) {
                facc[u][e] = 0.f;
                iacc[u][e] = 0ull;
            }
        }
        // first participating client, then a one-client-ahead prefetch of the rows
        int k = 0;
        while (k < a.k_local && !gpart[gid[k]]) ++k;
        float xn[UNR][VEC];
        if (k < a.k_local) {
#pragma unroll
            for (int u = 0; u < UNR; ++u) IO::ldg(a.client_params + (size_t)k * a.dim + (size_t)vi[u] * VEC, xn[u]);
        }
        while (k < a.k_local) {
            float x[UNR][VEC];
#pragma unroll
            for (int u = 0; u < UNR; ++u)
#pragma unroll
                for (int e = 0; e < VEC; ++e) x[u][e] = xn[u][e];
            int kn = k + 1;
            while (kn < a.k_local && !gpart[gid[kn]]) ++kn;
            if (kn < a.k_local) {
#pragma unroll
                for (int u = 0; u < UNR; ++u) IO::ldg(a.client_params + (size_t)kn * a.dim + (size_t)vi[u] * VEC, xn[u]);
            }
            const int gk = gid[k];
            const float ck = coef[k], wn = wnoise[k];
            const uint64_t nk = nkey[k];
#pragma unroll
            for (int u = 0; u < UNR; ++u) {
                if (v0 + u * gstride >= nvec) continue;       // clamped duplicate of the last vector: no work
#pragma unroll
                for (int e = 0; e < VEC; ++e) {
                    const int d = vi[u] * VEC + e;
                    float v = ck * delta_of(a, x[u][e], d, g[u][e]);
                    if (noise_std > 0.f) v += wn * normal_from_bits(stream_at(nk, (uint64_t)d));
                    if constexpr (!FIXED) {
                        facc[u][e] += v;
                    } else {
                        unsigned long long y = (unsigned long long)__double2ll_rn((double)v * FIXED_SCALE);
                        for (int j = 0; j < a.k_global; ++j) {
                            if (j == gk || !gpart[j]) continue;
                            uint64_t mk;
                            if (mask_keys_in_smem) {
                                mk = mkey[(size_t)k * a.k_global + j];
                            } else {
                                const int lo = min(gk, j), hi = max(gk, j);
                                const uint64_t pid = (uint64_t)lo * (uint64_t)a.k_global + (uint64_t)hi;
                                mk = key3((a.pair_keys ? a.pair_keys[pid] : a.seed) ^ TAG_MASK, (uint64_t)a.round, pid);
                            }
                            const uint64_t m = stream_at(mk, (uint64_t)d);
                            y += (gk < j) ? m : (0ull - m);
                        }
                        if (a.masked_out) a.masked_out[(size_t)k * a.dim + d] = (long long)y;
                        iacc[u][e] += y;
                    }
                }
            }
            k = kn;
        }
#pragma unroll
        for (int u = 0; u < UNR; ++u) {
            if (v0 + u * gstride >= nvec) continue;           // clamped duplicate: nothing to write
            const size_t d = (size_t)vi[u] * VEC;
            if (a.world == 1) {
                float o[VEC];
#pragma unroll
                for (int e = 0; e < VEC; ++e) {
                    const float upd = FIXED ? (float)((double)(long long)iacc[u][e] / FIXED_SCALE)
                                            : facc[u][e];
                    o[e] = (a.mode == 1) ? g[u][e] + a.server_lr * upd : upd;
                }
                IO::st(a.global_out + d, o);
            } else {
                if (a.two_shot == 2) {
                    // NVLS: the partial stays at home; the switch reads it when the owner of the slice reduces
                    if constexpr (FIXED) IO::st(reinterpret_cast<long long*>(a.nvls_partial) + d, iacc[u]);
                    else IO::st(reinterpret_cast<float*>(a.nvls_partial) + d, facc[u]);
                } else if (a.two_shot) {
                    // reduce-scatter by peer stores, overlapped with this sweep over the client rows: vector v belongs to
                    // rank o = v / chunk, whose buffer holds one row per source rank: [W][chunk] vectors (region 0)
                    const int chunk = (nvec + a.world - 1) / a.world;
                    const int o = vi[u] / chunk;
                    const size_t slot = ((size_t)a.rank * chunk + (size_t)(vi[u] - o * chunk)) * VEC;
                    if constexpr (FIXED) IO::st(a.recv_slots[o] + slot, iacc[u]);
                    else IO::st(reinterpret_cast<float*>(a.recv_slots[o]) + slot, facc[u]);
                } else {
                    // push this rank's partial into slot [rank] of every rank (self included)
                    for (int r = 0; r < a.world; ++r) {
                        if constexpr (FIXED) IO::st(a.recv_slots[r] + (size_t)a.rank * a.dim + d, iacc[u]);
                        else IO::st(reinterpret_cast<float*>(a.recv_slots[r]) + (size_t)a.rank * a.dim + d, facc[u]);
                    }
                }
            }
        }
    }
    if (a.world == 1) return;

    // ---- phase 3: publish, wait for every peer, reduce the slots -------------------------------
    __threadfence_system();
    grid.sync();
    if (blockIdx.x == 0 && tid < a.world) st_release_sys(a.flags[tid] + a.rank, a.epoch);
    if (blockIdx.x == 0 && tid < a.world) {
        const unsigned int* f = a.flags[a.rank] + tid;
        while ((int)(ld_acquire_sys(f) - a.epoch) < 0) { __nanosleep(64); }
    }
    grid.sync();
    if (a.two_shot == 2) {
        // ---- NVLS: in-switch sum of the W partials of this rank's slice, update, broadcast store -------------------
        const int chunk = (nvec + a.world - 1) / a.world;
        const int lo = min(nvec, a.rank * chunk), hi = min(nvec, lo + chunk);
        constexpr int NU = 4;                                  // independent multicast loads in flight per thread
        for (int v0 = lo + g0; v0 < hi; v0 += NU * gstride) {
            float upd[NU][VEC], g[NU][VEC];
#pragma unroll
            for (int u = 0; u < NU; ++u) {
                const int v = min(v0 + u * gstride, hi - 1);
                const size_t d = (size_t)v * VEC;
                if constexpr (FIXED) {
#pragma unroll
                    for (int e = 0; e < VEC; ++e) {
                        const unsigned long long sacc = mc_ld_reduce_u64(reinterpret_cast<const long long*>(a.nvls_partial_mc) + d + e);
                        upd[u][e] = (float)((double)(long long)sacc / FIXED_SCALE);
                    }
                } else if constexpr (VEC == 4) {
                    mc_ld_reduce_f32x4(reinterpret_cast<const float*>(a.nvls_partial_mc) + d, upd[u]);
                } else {
                    upd[u][0] = mc_ld_reduce_f32(reinterpret_cast<const float*>(a.nvls_partial_mc) + d);
                }
                IO::ld(a.global_in + d, g[u]);
            }
#pragma unroll
            for (int u = 0; u < NU; ++u) {
                const int v = v0 + u * gstride;
                if (v >= hi) continue;
                float o[VEC];
#pragma unroll
                for (int e = 0; e < VEC; ++e) o[e] = (a.mode == 1) ? g[u][e] + a.server_lr * upd[u][e] : upd[u][e];
                if constexpr (VEC == 4) mc_st_f32x4(a.nvls_out_mc + (size_t)v * VEC, o);
                else mc_st_f32(a.nvls_out_mc + (size_t)v * VEC, o[0]);
            }
        }
        __threadfence_system();
        grid.sync();
        if (blockIdx.x == 0 && tid < a.world) st_release_sys(a.flags[tid] + AGG_MAX_WORLD + a.rank, a.epoch);
        if (blockIdx.x == 0 && tid < a.world) {
            const unsigned int* f = a.flags[a.rank] + AGG_MAX_WORLD + tid;
            while ((int)(ld_acquire_sys(f) - a.epoch) < 0) { __nanosleep(64); }
        }
        grid.sync();
        if (a.global_out != a.nvls_out) {                      // callers that read the result row in place skip the copy
            for (int v = g0; v < nvec; v += gstride) {
                float t[VEC];
                IO::ldcv(a.nvls_out + (size_t)v * VEC, t);
                IO::st(a.global_out + (size_t)v * VEC, t);
            }
        }
        return;
    }
    if (a.two_shot) {
        // ---- the W partial rows of this rank's slice [lo, hi) arrived by peer stores: reduce them locally --------
        const int chunk = (nvec + a.world - 1) / a.world;
        const int lo = min(nvec, a.rank * chunk), hi = min(nvec, lo + chunk);
        const size_t row_words = (size_t)chunk * VEC;
        float* result = reinterpret_cast<float*>(a.recv_slots[a.rank] + (size_t)a.world * row_words);   // region 1
        for (int v = lo + g0; v < hi; v += gstride) {
            const size_t d = (size_t)v * VEC, loc = (size_t)(v - lo) * VEC;
            float upd[VEC];
            if constexpr (FIXED) {
                unsigned long long sacc[VEC];
#pragma unroll
                for (int e = 0; e < VEC; ++e) sacc[e] = 0ull;
                for (int r = 0; r < a.world; ++r) {
                    long long t[VEC];
                    IO::ldcv(a.recv_slots[a.rank] + (size_t)r * row_words + loc, t);
#pragma unroll
                    for (int e = 0; e < VEC; ++e) sacc[e] += (unsigned long long)t[e];
                }
#pragma unroll
                for (int e = 0; e < VEC; ++e) upd[e] = (float)((double)(long long)sacc[e] / FIXED_SCALE);
            } else {
#pragma unroll
                for (int e = 0; e < VEC; ++e) upd[e] = 0.f;
                for (int r = 0; r < a.world; ++r) {           // fixed rank order: the sum is the same on every run
                    float t[VEC];
                    IO::ldcv(reinterpret_cast<const float*>(a.recv_slots[a.rank]) + (size_t)r * row_words + loc, t);
#pragma unroll
                    for (int e = 0; e < VEC; ++e) upd[e] += t[e];
                }
            }
            float g[VEC], o[VEC];
            IO::ld(a.global_in + d, g);
#pragma unroll
            for (int e = 0; e < VEC; ++e) o[e] = (a.mode == 1) ? g[e] + a.server_lr * upd[e] : upd[e];
            IO::st(a.global_out + d, o);
            IO::st(result + loc, o);
        }
        // ---- second barrier (its own flag words), then gather the other ranks' slices ---------------------
        __threadfence_system();
        grid.sync();
        if (blockIdx.x == 0 && tid < a.world) st_release_sys(a.flags[tid] + AGG_MAX_WORLD + a.rank, a.epoch);
        if (blockIdx.x == 0 && tid < a.world) {
            const unsigned int* f = a.flags[a.rank] + AGG_MAX_WORLD + tid;
            while ((int)(ld_acquire_sys(f) - a.epoch) < 0) { __nanosleep(64); }
        }
        grid.sync();
        for (int q = 1; q < a.world; ++q) {
            const int r = (a.rank + q) % a.world;             // start with the next rank: W ranks, W different targets
            const int rlo = min(nvec, r * chunk), rhi = min(nvec, rlo + chunk);
            const float* src = reinterpret_cast<const float*>(a.recv_slots[r] + (size_t)a.world * row_words);
            for (int v = rlo + g0; v < rhi; v += gstride) {
                float t[VEC];
                IO::ldcv(src + (size_t)(v - rlo) * VEC, t);
                IO::st(a.global_out + (size_t)v * VEC, t);
            }
        }
        return;
    }
    const long long* mine = a.recv_slots[a.rank];
    for (int v = g0; v < nvec; v += gstride) {
        const size_t d = (size_t)v * VEC;
        float upd[VEC];
        if constexpr (FIXED) {
            unsigned long long s[VEC];
#pragma unroll
            for (int e = 0; e < VEC; ++e) s[e] = 0ull;
            for (int r = 0; r < a.world; ++r) {
                long long t[VEC];
                IO::ldcv(mine + (size_t)r * a.dim + d, t);
#pragma unroll
                for (int e = 0; e < VEC; ++e) s[e] += (unsigned long long)t[e];
            }
#pragma unroll
            for (int e = 0; e < VEC; ++e) upd[e] = (float)((double)(long long)s[e] / FIXED_SCALE);
        } else {
            const float* mf = reinterpret_cast<const float*>(mine);
#pragma unroll
            for (int e = 0; e < VEC; ++e) upd[e] = 0.f;
            for (int r = 0; r < a.world; ++r) {
                float t[VEC];
                IO::ldcv(mf + (size_t)r * a.dim + d, t);
#pragma unroll
                for (int e = 0; e < VEC; ++e) upd[e] += t[e];
            }
        }
        float g[VEC], o[VEC];
        IO::ld(a.global_in + d, g);
#pragma unroll
        for (int e = 0; e < VEC; ++e) o[e] = (a.mode == 1) ? g[e] + a.server_lr * upd[e] : upd[e];
        IO::st(a.global_out + d, o);
    }
}

// One GPU: four CTAs per SM (64 registers) - the sweep over the client rows is bandwidth-bound and lives on loads in
// flight; left alone, the compiler takes 78 registers for the body with its cross-GPU branches and the 1e8-float sweep
// loses 15 %.  Across GPUs (MULTI) the exchange loops want the registers instead (NVLS at 1e8 floats: 1.17 vs 1.21 ms).
template <int VEC, int UNR, bool FIXED, bool MULTI>
__global__ void __launch_bounds__(AGG_THREADS, MULTI ? 3 : 4)
fedavg_fused_kernel(const AggArgs a, const int mask_keys_in_smem) {
    fedavg_fused_body<VEC, UNR, FIXED>(a, mask_keys_in_smem);
}
// Loopback world: ONE cooperative launch plays every rank of a W-rank job on one GPU - blockIdx.y is the rank, the
// receive slots and flag words of all "ranks" live in this device's memory.  The blocks of all ranks are co-resident
// (cooperative launch), so the flag waits of the cross-rank barriers are safe; grid.sync() then spans all ranks, which
// is stronger than the per-rank barrier the real job has and changes nothing else.  This is how the multi-GPU phases
// (publish, flag wait, slot reduce / reduce-scatter + all-gather) are exercised on a single-GPU test box.
template <int VEC, int UNR, bool FIXED>
__global__ void __launch_bounds__(AGG_THREADS, 4)
fedavg_fused_loopback_kernel(const AggArgs* __restrict__ ranks, const int mask_keys_in_smem) {
    const AggArgs a = ranks[blockIdx.y];
    fedavg_fused_body<VEC, UNR, FIXED>(a, mask_keys_in_smem);
}

using AggKernel = void (*)(const AggArgs, const int);
using AggLoopKernel = void (*)(const AggArgs*, const int);

static int launch_fedavg(AggArgs& a, cudaStream_t stream) {
    QFX_ARG_CHECK(a.k_local >= 0 && a.k_global >= 1 && a.k_global >= a.k_local && a.dim >= 1, QFX_ERR_BAD_ARG, "bad K/D");
    QFX_ARG_CHECK(a.k_local >= 1 || a.world > 1, QFX_ERR_BAD_ARG, "no clients to aggregate");
    QFX_ARG_CHECK(a.mode == 0 || a.mode == 1, QFX_ERR_BAD_ARG, "mode must be 0 (weights) or 1 (delta)");
    QFX_ARG_CHECK(!(a.mode == 0 && (a.clip > 0.f || a.sigma > 0.f)), QFX_ERR_BAD_ARG,
                  "clip/noise need delta mode");
    QFX_ARG_CHECK(a.world >= 1 && a.world <= AGG_MAX_WORLD && a.rank >= 0 && a.rank < a.world, QFX_ERR_BAD_ARG,
                  "bad rank/world");
    if (a.world > 1 && a.two_shot == 2)
        QFX_ARG_CHECK(a.flags && a.nvls_partial && a.nvls_partial_mc && a.nvls_out && a.nvls_out_mc, QFX_ERR_BAD_ARG,
                      "the NVLS exchange needs flags and the unicast + multicast addresses of the partial and result rows");
    else
        QFX_ARG_CHECK(a.world == 1 || (a.recv_slots && a.flags), QFX_ERR_BAD_ARG, "multi-GPU needs slots and flags");
    QFX_ARG_CHECK(a.scratch != nullptr, QFX_ERR_BAD_ARG, "scratch is required");
    const int num_sms = device_sm_count();
    QFX_ARG_CHECK(num_sms > 0, QFX_ERR_UNSUPPORTED, "cannot query the SM count of the current device");
    // pairwise mask keys are hashed once per block into shared memory when they fit
    const size_t kSmemCap = 96 * 1024;
    int mask_keys = a.secagg && agg_smem_bytes(a.k_local, a.k_global, true) <= 32 * 1024 ? 1 : 0;
    const size_t smem = agg_smem_bytes(a.k_local, a.k_global, mask_keys != 0);
    QFX_ARG_CHECK(smem <= kSmemCap, QFX_ERR_TOO_LARGE, "too many clients for one aggregation launch (%zu bytes of tables)", smem);
    if (a.k_local > 0) QFX_CUDA_TRY(cudaMemsetAsync(a.scratch, 0, sizeof(float) * a.k_local, stream));
    // 16-byte accesses when every row starts on a 16-byte boundary
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    const bool nvls = a.world > 1 && a.two_shot == 2;
    const bool vec4 = a.dim % 4 == 0 && al16(a.client_params) && al16(a.global_in) && al16(a.global_out) &&
                      (!nvls || (al16(a.nvls_partial) && al16(a.nvls_partial_mc) && al16(a.nvls_out) && al16(a.nvls_out_mc)));
    const bool fixed = a.secagg != 0;
    const bool multi = a.world > 1;
    AggKernel kern;
    if (multi)
        kern = vec4 ? (fixed ? (AggKernel)fedavg_fused_kernel<4, 2, true, true> : (AggKernel)fedavg_fused_kernel<4, 2, false, true>)
                    : (fixed ? (AggKernel)fedavg_fused_kernel<1, 4, true, true> : (AggKernel)fedavg_fused_kernel<1, 4, false, true>);
    else
        kern = vec4 ? (fixed ? (AggKernel)fedavg_fused_kernel<4, 2, true, false> : (AggKernel)fedavg_fused_kernel<4, 2, false, false>)
                    : (fixed ? (AggKernel)fedavg_fused_kernel<1, 4, true, false> : (AggKernel)fedavg_fused_kernel<1, 4, false, false>);
    static PerDeviceFlag attr_set[8];
    const int which = (multi ? 4 : 0) + (vec4 ? 2 : 0) + (fixed ? 1 : 0);
    if (!attr_set[which].here()) {
        QFX_CUDA_TRY(cudaFuncSetAttribute((const void*)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemCap));
        attr_set[which].here() = true;
    }
    int occ = 0;
    QFX_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void*)kern, AGG_THREADS, smem));
    QFX_ARG_CHECK(occ >= 1, QFX_ERR_TOO_LARGE, "aggregation kernel does not fit on an SM");
    // small vectors: as few CTAs as cover the vector (grid barriers get cheaper); large: fill the GPU
    const int per_thread = vec4 ? 4 : 1;
    long long blocks = ((long long)a.dim / per_thread + AGG_THREADS - 1) / AGG_THREADS;
    const long long cap = (long long)num_sms * (occ > 6 ? 6 : occ);
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    void* params[] = {(void*)&a, (void*)&mask_keys};
    QFX_CUDA_TRY(cudaLaunchCooperativeKernel((const void*)kern, dim3((unsigned)blocks), dim3(AGG_THREADS),
                                             params, smem, stream));
    return 0;
}

}  // namespace qfx

using namespace qfx;

extern "C" int qfx_fedavg_fused(const float* client_params, const float* global_in, float* global_out,
                                const float* weights, const int* participate, int n_clients, int dim, int mode,
                                int n_wrap, float clip, float sigma, float server_lr, unsigned long long seed,
                                int round, int secagg, float* norms_out, long long* masked_out, float* scratch,
                                cudaStream_t stream) {
    AggArgs a = {};
    a.client_params = client_params; a.global_in = global_in; a.global_out = global_out;
    a.weights = weights; a.participate = participate; a.client_gids = nullptr;
    a.k_local = n_clients; a.k_global = n_clients; a.dim = dim; a.mode = mode; a.n_wrap = n_wrap;
    a.clip = clip; a.sigma = sigma; a.server_lr = server_lr; a.seed = seed; a.round = round; a.secagg = secagg;
    a.norms_out = norms_out; a.scratch = scratch; a.masked_out = masked_out;
    a.rank = 0; a.world = 1; a.recv_slots = nullptr; a.flags = nullptr; a.epoch = 0;
    return launch_fedavg(a, stream);
}

extern "C" int qfx_fedavg_fused_multi(const float* client_params, const float* global_in, float* global_out,
                                      const float* weights, const int* participate, const int* client_gids,
                                      int k_local, int k_global, int dim, int mode, int n_wrap, float clip,
                                      float sigma, float server_lr, unsigned long long seed, int round, int secagg,
                                      float* norms_out, float* scratch, int rank, int world,
                                      long long* const* recv_slots, unsigned int* const* flags, unsigned int epoch,
                                      cudaStream_t stream) {
    AggArgs a = {};
    a.client_params = client_params; a.global_in = global_in; a.global_out = global_out;
    a.weights = weights; a.participate = participate; a.client_gids = client_gids;
    a.k_local = k_local; a.k_global = k_global; a.dim = dim; a.mode = mode; a.n_wrap = n_wrap;
    a.clip = clip; a.sigma = sigma; a.server_lr = server_lr; a.seed = seed; a.round = round; a.secagg = secagg;
    a.norms_out = norms_out; a.scratch = scratch; a.masked_out = nullptr;
    a.rank = rank; a.world = world; a.recv_slots = recv_slots; a.flags = flags; a.epoch = epoch;
    return launch_fedavg(a, stream);
}

// Loopback test entry: `world` logical ranks on the current GPU in one cooperative launch (see the kernel above).
// Per-rank arrays are HOST arrays of length world; everything else is shared.  args_dev: world * 512 bytes of device
// scratch for the per-rank argument blocks.  Requires dim % 4 == 0 alignment rules to agree across ranks (the test
// allocates every buffer with torch, 256-byte aligned).
extern "C" int qfx_fedavg_fused_loopback(int world, const float* const* client_params, const int* const* client_gids,
                                         const int* k_local, float* const* global_out, float* const* scratch,
                                         const float* global_in, const float* weights, const int* participate,
                                         int k_global, int dim, int mode, int n_wrap, float clip, float sigma,
                                         float server_lr, unsigned long long seed, int round, int secagg,
                                         const unsigned long long* pair_keys, long long* const* recv_slots,
                                         unsigned int* const* flags, unsigned int epoch, int two_shot, void* args_dev,
                                         cudaStream_t stream) {
    QFX_ARG_CHECK(world >= 2 && world <= AGG_MAX_WORLD && args_dev && recv_slots && flags, QFX_ERR_BAD_ARG, "bad loopback arguments");
    static_assert(sizeof(AggArgs) <= 512, "argument block");
    AggArgs host[AGG_MAX_WORLD];
    int kmax = 0;
    bool vec4 = dim % 4 == 0;
    auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    for (int r = 0; r < world; ++r) {
        AggArgs a = {};
        a.client_params = client_params[r] ? client_params[r] : global_in; a.global_in = global_in; a.global_out = global_out[r];
        a.weights = weights; a.participate = participate; a.client_gids = client_gids[r];
        a.k_local = k_local[r]; a.k_global = k_global; a.dim = dim; a.mode = mode; a.n_wrap = n_wrap;
        a.clip = clip; a.sigma = sigma; a.server_lr = server_lr; a.seed = seed; a.round = round; a.secagg = secagg;
        a.norms_out = nullptr; a.scratch = scratch[r]; a.masked_out = nullptr; a.pair_keys = pair_keys;
        a.rank = r; a.world = world; a.recv_slots = recv_slots; a.flags = flags; a.epoch = epoch; a.two_shot = two_shot ? 1 : 0;
        QFX_ARG_CHECK(a.k_local >= 0 && a.scratch && a.global_out, QFX_ERR_BAD_ARG, "bad per-rank arguments");
        vec4 = vec4 && al16(a.client_params) && al16(a.global_in) && al16(a.global_out);
        if (a.k_local > kmax) kmax = a.k_local;
        if (a.k_local > 0) QFX_CUDA_TRY(cudaMemsetAsync(a.scratch, 0, sizeof(float) * a.k_local, stream));
        host[r] = a;
    }
    char* dev_args = (char*)args_dev;
    // one argument block per rank at a 512-byte pitch is awkward for indexing: pack them contiguously instead
    QFX_CUDA_TRY(cudaMemcpyAsync(dev_args, host, sizeof(AggArgs) * world, cudaMemcpyHostToDevice, stream));
    QFX_CUDA_TRY(cudaStreamSynchronize(stream));     // `host` is a stack array
    const bool fixed = secagg != 0;
    int mask_keys = fixed && agg_smem_bytes(kmax, k_global, true) <= 32 * 1024 ? 1 : 0;
    const size_t smem = agg_smem_bytes(kmax, k_global, mask_keys != 0);
    AggLoopKernel kern = vec4 ? (fixed ? (AggLoopKernel)fedavg_fused_loopback_kernel<4, 2, true> : (AggLoopKernel)fedavg_fused_loopback_kernel<4, 2, false>)
                              : (fixed ? (AggLoopKernel)fedavg_fused_loopback_kernel<1, 4, true> : (AggLoopKernel)fedavg_fused_loopback_kernel<1, 4, false>);
    QFX_CUDA_TRY(cudaFuncSetAttribute((const void*)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
    int occ = 0;
    QFX_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void*)kern, AGG_THREADS, smem));
    QFX_ARG_CHECK(occ >= 1, QFX_ERR_TOO_LARGE, "aggregation kernel does not fit on an SM");
    const int per_thread = vec4 ? 4 : 1;
    long long blocks = ((long long)dim / per_thread + AGG_THREADS - 1) / AGG_THREADS;
    const long long cap = (long long)device_sm_count() * (occ > 6 ? 6 : occ) / world;     // all ranks co-resident
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    const AggArgs* ranks = reinterpret_cast<const AggArgs*>(dev_args);
    void* params[] = {(void*)&ranks, (void*)&mask_keys};
    QFX_CUDA_TRY(cudaLaunchCooperativeKernel((const void*)kern, dim3((unsigned)blocks, (unsigned)world), dim3(AGG_THREADS),
                                             params, smem, stream));
    return 0;
}

// qfx_fedavg_fused_multi with an explicit exchange algorithm: two_shot = 0 is the one-barrier all-gather of partials
// (recv_slots: world x dim words per rank), two_shot = 1 the reduce-scatter by peer stores + all-gather by peer loads
// (recv_slots: (world + 1) * (ceil(dim / world) rounded up to 4) words per rank) - (W-1) D versus 2 (W-1)/W D words over
// NVLink per rank, the first half overlapped with the sweep over the client rows.
// pair_keys may be null (masks from the job seed).  flags: at least 2 * 16 words per rank.
extern "C" int qfx_fedavg_fused_multi_ex(const float* client_params, const float* global_in, float* global_out,
                                         const float* weights, const int* participate, const int* client_gids,
                                         int k_local, int k_global, int dim, int mode, int n_wrap, float clip,
                                         float sigma, float server_lr, unsigned long long seed, int round, int secagg,
                                         const unsigned long long* pair_keys, float* norms_out, float* scratch,
                                         int rank, int world, long long* const* recv_slots,
                                         unsigned int* const* flags, unsigned int epoch, int two_shot,
                                         cudaStream_t stream) {
    AggArgs a = {};
    a.client_params = client_params; a.global_in = global_in; a.global_out = global_out;
    a.weights = weights; a.participate = participate; a.client_gids = client_gids;
    a.k_local = k_local; a.k_global = k_global; a.dim = dim; a.mode = mode; a.n_wrap = n_wrap;
    a.clip = clip; a.sigma = sigma; a.server_lr = server_lr; a.seed = seed; a.round = round; a.secagg = secagg;
    a.norms_out = norms_out; a.scratch = scratch; a.masked_out = nullptr; a.pair_keys = pair_keys;
    a.rank = rank; a.world = world; a.recv_slots = recv_slots; a.flags = flags; a.epoch = epoch;
    a.two_shot = two_shot ? 1 : 0;
    return launch_fedavg(a, stream);
}

// K7 across GPUs through the NVSwitch reduction (NVLS).  partial / out are rows of dim words (partial: 8 bytes per word
// with secagg, else 4) in buffers bound to a multicast object that spans the job's GPUs; *_mc is the multicast address
// of the same row (parallel/runtime.py takes both from torch's symmetric memory rendezvous).  The result lands in
// out on every rank; global_out may be out itself (no copy) or any other row (copied at the end of the launch).
extern "C" int qfx_fedavg_fused_multi_nvls(const float* client_params, const float* global_in, float* global_out,
                                           const float* weights, const int* participate, const int* client_gids,
                                           int k_local, int k_global, int dim, int mode, int n_wrap, float clip,
                                           float sigma, float server_lr, unsigned long long seed, int round, int secagg,
                                           const unsigned long long* pair_keys, float* norms_out, float* scratch,
                                           int rank, int world, void* partial, const void* partial_mc, float* out,
                                           float* out_mc, unsigned int* const* flags, unsigned int epoch,
                                           cudaStream_t stream) {
    AggArgs a = {};
    a.client_params = client_params; a.global_in = global_in; a.global_out = global_out;
    a.weights = weights; a.participate = participate; a.client_gids = client_gids;
    a.k_local = k_local; a.k_global = k_global; a.dim = dim; a.mode = mode; a.n_wrap = n_wrap;
    a.clip = clip; a.sigma = sigma; a.server_lr = server_lr; a.seed = seed; a.round = round; a.secagg = secagg;
    a.norms_out = norms_out; a.scratch = scratch; a.masked_out = nullptr; a.pair_keys = pair_keys;
    a.rank = rank; a.world = world; a.recv_slots = nullptr; a.flags = flags; a.epoch = epoch;
    a.two_shot = 2;
    a.nvls_partial = partial; a.nvls_partial_mc = partial_mc; a.nvls_out = out; a.nvls_out_mc = out_mc;
    QFX_ARG_CHECK(world >= 2, QFX_ERR_BAD_ARG, "the NVLS exchange is a multi-GPU path");
    return launch_fedavg(a, stream);
}

// Same as qfx_fedavg_fused / qfx_fedavg_fused_multi with per-pair mask secrets: pair_keys [k_global, k_global] holds, at
// [lo * k_global + hi], the 64-bit secret clients lo < hi agreed on at registration (fl/secagg.py: KeyAgreement, the
// "seed exchange API" of /root/reference/ROADMAP.md:52-55); the mask stream of a pair is keyed by it instead of the job seed.
extern "C" int qfx_fedavg_fused_pairkeys(const float* client_params, const float* global_in, float* global_out,
                                         const float* weights, const int* participate, int n_clients, int dim, int mode,
                                         int n_wrap, float clip, float sigma, float server_lr, unsigned long long seed,
                                         int round, const unsigned long long* pair_keys, float* norms_out,
                                         long long* masked_out, float* scratch, cudaStream_t stream) {
    QFX_ARG_CHECK(pair_keys != nullptr, QFX_ERR_BAD_ARG, "pair_keys is required");
    AggArgs a = {};
    a.client_params = client_params; a.global_in = global_in; a.global_out = global_out;
    a.weights = weights; a.participate = participate; a.client_gids = nullptr;
    a.k_local = n_clients; a.k_global = n_clients; a.dim = dim; a.mode = mode; a.n_wrap = n_wrap;
    a.clip = clip; a.sigma = sigma; a.server_lr = server_lr; a.seed = seed; a.round = round; a.secagg = 1;
    a.norms_out = norms_out; a.scratch = scratch; a.masked_out = masked_out; a.pair_keys = pair_keys;
    a.rank = 0; a.world = 1; a.recv_slots = nullptr; a.flags = nullptr; a.epoch = 0;
    return launch_fedavg(a, stream);
}
extern "C" int qfx_fedavg_fused_multi_pairkeys(const float* client_params, const float* global_in, float* global_out,
                                               const float* weights, const int* participate, const int* client_gids,
                                               int k_local, int k_global, int dim, int mode, int n_wrap, float clip,
                                               float sigma, float server_lr, unsigned long long seed, int round,
                                               const unsigned long long* pair_keys, float* norms_out, float* scratch,
                                               int rank, int world, long long* const* recv_slots,
                                               unsigned int* const* flags, unsigned int epoch, cudaStream_t stream) {
    QFX_ARG_CHECK(pair_keys != nullptr, QFX_ERR_BAD_ARG, "pair_keys is required");
    AggArgs a = {};
    a.client_params = client_params; a.global_in = global_in; a.global_out = global_out;
    a.weights = weights; a.participate = participate; a.client_gids = client_gids;
    a.k_local = k_local; a.k_global = k_global; a.dim = dim; a.mode = mode; a.n_wrap = n_wrap;
    a.clip = clip; a.sigma = sigma; a.server_lr = server_lr; a.seed = seed; a.round = round; a.secagg = 1;
    a.norms_out = norms_out; a.scratch = scratch; a.masked_out = nullptr; a.pair_keys = pair_keys;
    a.rank = rank; a.world = world; a.recv_slots = recv_slots; a.flags = flags; a.epoch = epoch;
    return launch_fedavg(a, stream);
}

// ---------------------------------------------------------------------------------------------
// Aggregation for callers that must not allocate.
//
// qfx_fedavg_fused needs n_clients floats of workspace for the per-client squared norms and expects
// the caller to bring it.  Some callers have nowhere to get it from: a federated server written
// against this C ABI without PyTorch has no caching allocator to borrow from; a process whose device
// memory has been handed over wholesale to another framework's arena has been told not to call
// cudaMalloc at all; and a serving loop that aggregates between client launches cannot afford
// cudaMalloc/cudaFree there, because both synchronise the whole device and would stall the clients'
// kernels for the duration of the call.  For them the workspace lives in the library image itself:
// the array below is part of every GPU's copy of the module from the moment the library is loaded
// until the process exits (QFX_FEDAVG_STATIC_MAX_CLIENTS * 4 bytes = 16 KiB per GPU).
//
// `g_static_norm_scratch` is a device symbol.  In host code that name is only a handle for the
// runtime - handing it to a kernel as if it were a pointer would make the kernel read host-side
// bytes or fault - so the host asks the runtime for the device address once per device
// (cudaGetSymbolAddress, with that device current) and passes that address on.
static __device__ __align__(256) float g_static_norm_scratch[QFX_FEDAVG_STATIC_MAX_CLIENTS];

namespace qfx {
constexpr int STATIC_MAX_DEVICES = 64;
static float* g_static_scratch_addr[STATIC_MAX_DEVICES] = {};

// Device address of the static workspace on the current device; looked up on the first call per device.
static int static_scratch_for_current_device(float** out) {
    int dev = 0;
    QFX_CUDA_TRY(cudaGetDevice(&dev));
    QFX_ARG_CHECK(dev >= 0 && dev < STATIC_MAX_DEVICES, QFX_ERR_UNSUPPORTED, "device ordinal %d not supported", dev);
    if (g_static_scratch_addr[dev] == nullptr) {
        float* addr = nullptr;
        QFX_CUDA_TRY(cudaGetSymbolAddress((void**)&addr, g_static_norm_scratch));
        QFX_ARG_CHECK(addr != nullptr, QFX_ERR_UNSUPPORTED, "cudaGetSymbolAddress returned a null address");
        g_static_scratch_addr[dev] = addr;
    }
    *out = g_static_scratch_addr[dev];
    return 0;
}
}  // namespace qfx

// Same computation and arguments as qfx_fedavg_fused, minus the workspace pointer: inputs and outputs
// still come from the caller, the norm accumulators come from the static array above.  Nothing in
// this call allocates or frees memory, on the device or through any library.
//   * n_clients above QFX_FEDAVG_STATIC_MAX_CLIENTS is refused with QFX_ERR_TOO_LARGE and a message
//     naming the limit; use qfx_fedavg_fused with your own workspace for larger federations.
//   * The first call on each device looks up the array's device address.  Make that first call during
//     start-up, with the device current and before any stream capture begins on it.
//   * All calls on one GPU share the one array and it keeps whatever the previous call left in it.
//     The accumulators this call uses are cleared on `stream` before the kernel runs, so back-to-back
//     calls on one stream are safe; two calls that can overlap on the same GPU (different streams,
//     different host threads) would accumulate into the same floats and corrupt each other's clip
//     factors - serialise them or give one of them its own workspace through qfx_fedavg_fused.
extern "C" int qfx_fedavg_fused_static(const float* client_params, const float* global_in, float* global_out,
                                       const float* weights, const int* participate, int n_clients, int dim,
                                       int mode, int n_wrap, float clip, float sigma, float server_lr,
                                       unsigned long long seed, int round, int secagg, float* norms_out,
                                       long long* masked_out, cudaStream_t stream) {
    QFX_ARG_CHECK(n_clients <= QFX_FEDAVG_STATIC_MAX_CLIENTS, QFX_ERR_TOO_LARGE,
                  "qfx_fedavg_fused_static handles at most %d clients per call (got %d); "
                  "call qfx_fedavg_fused with a caller-provided workspace instead",
                  QFX_FEDAVG_STATIC_MAX_CLIENTS, n_clients);
    float* scratch = nullptr;
    if (int e = static_scratch_for_current_device(&scratch)) return e;
    AggArgs a = {};
    a.client_params = client_params; a.global_in = global_in; a.global_out = global_out;
    a.weights = weights; a.participate = participate; a.client_gids = nullptr;
    a.k_local = n_clients; a.k_global = n_clients; a.dim = dim; a.mode = mode; a.n_wrap = n_wrap;
    a.clip = clip; a.sigma = sigma; a.server_lr = server_lr; a.seed = seed; a.round = round; a.secagg = secagg;
    a.norms_out = norms_out; a.scratch = scratch; a.masked_out = masked_out;
    a.rank = 0; a.world = 1; a.recv_slots = nullptr; a.flags = nullptr; a.epoch = 0;
    return launch_fedavg(a, stream);  // clears the accumulators it uses (cudaMemsetAsync), then launches
}
