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
    // multi-GPU
    int rank, world;
    long long* const* recv_slots;  // [world] device pointers: recv_slots[r] = base of rank r's slot array
                                   // [world][dim] int64 (secagg) or viewed as float (plain)
    unsigned int* const* flags;    // [world] device pointers: flags[r] = rank r's flag array [world]
    unsigned int epoch;            // value written into flags this round (monotonic)
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ float delta_at(const AggArgs& a, int k, int d, float g) {
    float x = a.client_params[(size_t)k * a.dim + d];
    if (a.mode == 0) return x;
    x -= g;
    return d < a.n_wrap ? wrap_pi(x) : x;
}

__global__ void __launch_bounds__(AGG_THREADS)
fedavg_fused_kernel(const AggArgs a) {
    cg::grid_group grid = cg::this_grid();
    extern __shared__ float coef[];  // [k_local] w_k/sum(w) * clip scale, 0 for non-participants
    __shared__ float red[AGG_THREADS / 32];
    const int tid = threadIdx.x;
    const int gstride = gridDim.x * AGG_THREADS;
    const int g0 = blockIdx.x * AGG_THREADS + tid;
    const bool need_norm = a.mode == 1 && (a.clip > 0.f || a.norms_out != nullptr);

    // ---- phase 1: squared norms ----------------------------------------------------------
    if (need_norm) {
        for (int k = 0; k < a.k_local; ++k) {
            float acc = 0.f;
            for (int d = g0; d < a.dim; d += gstride) {
                const float v = delta_at(a, k, d, a.global_in[d]);
                acc += v * v;
            }
            acc = warp_sum(acc);
            if ((tid & 31) == 0) red[tid >> 5] = acc;
            __syncthreads();
            if (tid < 32) {
                float v = tid < AGG_THREADS / 32 ? red[tid] : 0.f;
                v = warp_sum(v);
                if (tid == 0 && v != 0.f) atomicAdd(a.scratch + k, v);
            }
            __syncthreads();
        }
        grid.sync();
    }

    // ---- per-client coefficients -----------------------------------------------------------
    float wsum = 0.f;
    for (int j = 0; j < a.k_global; ++j)
        if (a.participate == nullptr || a.participate[j]) wsum += a.weights[j];
    const float inv_wsum = wsum > 0.f ? 1.f / wsum : 0.f;
    for (int k = tid; k < a.k_local; k += AGG_THREADS) {
        const int gk = a.client_gids ? a.client_gids[k] : k;
        float c = (a.participate == nullptr || a.participate[gk]) ? a.weights[gk] * inv_wsum : 0.f;
        if (need_norm) {
            const float nrm = sqrtf(a.scratch[k]);
            if (blockIdx.x == 0 && a.norms_out) a.norms_out[k] = nrm;
            if (a.clip > 0.f && nrm > a.clip) c *= a.clip / nrm;
        }
        coef[k] = c;
    }
    __syncthreads();

    const float noise_std = (a.mode == 1 && a.clip > 0.f) ? a.sigma * a.clip : 0.f;
    const bool fixed = a.secagg != 0;

    // ---- phase 2: weighted, noised, masked sum per element -----------------------------------
    for (int d = g0; d < a.dim; d += gstride) {
        const float g = a.global_in[d];
        float facc = 0.f;
        unsigned long long iacc = 0ull;
        for (int k = 0; k < a.k_local; ++k) {
            const int gk = a.client_gids ? a.client_gids[k] : k;
            const bool part = a.participate == nullptr || a.participate[gk];
            if (!part) continue;
            float v = coef[k] * delta_at(a, k, d, g);
            if (noise_std > 0.f) {
                const uint64_t nk = key3(a.seed ^ TAG_NOISE, (uint64_t)a.round, (uint64_t)gk);
                v += a.weights[gk] * inv_wsum * noise_std * normal_from_bits(stream_at(nk, (uint64_t)d));
            }
            if (!fixed) {
                facc += v;
            } else {
                unsigned long long y = (unsigned long long)__double2ll_rn((double)v * FIXED_SCALE);
                for (int j = 0; j < a.k_global; ++j) {
                    if (j == gk || !(a.participate == nullptr || a.participate[j])) continue;
                    const int lo = min(gk, j), hi = max(gk, j);
                    const uint64_t mk = key3(a.seed ^ TAG_MASK, (uint64_t)a.round,
                                             (uint64_t)lo * (uint64_t)a.k_global + (uint64_t)hi);
                    const uint64_t m = stream_at(mk, (uint64_t)d);
                    y += (gk < j) ? m : (0ull - m);
                }
                if (a.masked_out) a.masked_out[(size_t)k * a.dim + d] = (long long)y;
                iacc += y;
            }
        }
        if (a.world == 1) {
            const float upd = fixed ? (float)((double)(long long)iacc / FIXED_SCALE) : facc;
            a.global_out[d] = (a.mode == 1) ? g + a.server_lr * upd : upd;
        } else {
            // push this rank's partial into slot [rank] of every rank (self included)
            for (int r = 0; r < a.world; ++r) {
                if (fixed) a.recv_slots[r][(size_t)a.rank * a.dim + d] = (long long)iacc;
                else reinterpret_cast<float*>(a.recv_slots[r])[(size_t)a.rank * a.dim + d] = facc;
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
    const long long* mine = a.recv_slots[a.rank];
    for (int d = g0; d < a.dim; d += gstride) {
        float upd;
        if (fixed) {
            unsigned long long s = 0ull;
            for (int r = 0; r < a.world; ++r)
                s += (unsigned long long)__ldcv(mine + (size_t)r * a.dim + d);
            upd = (float)((double)(long long)s / FIXED_SCALE);
        } else {
            const float* mf = reinterpret_cast<const float*>(mine);
            float s = 0.f;
            for (int r = 0; r < a.world; ++r) s += __ldcv(mf + (size_t)r * a.dim + d);
            upd = s;
        }
        const float g = a.global_in[d];
        a.global_out[d] = (a.mode == 1) ? g + a.server_lr * upd : upd;
    }
}

static int g_num_sms = 0;

static int launch_fedavg(AggArgs& a, cudaStream_t stream) {
    QFX_ARG_CHECK(a.k_local >= 0 && a.k_global >= 1 && a.k_global >= a.k_local && a.dim >= 1, QFX_ERR_BAD_ARG, "bad K/D");
    QFX_ARG_CHECK(a.k_local >= 1 || a.world > 1, QFX_ERR_BAD_ARG, "no clients to aggregate");
    QFX_ARG_CHECK(a.mode == 0 || a.mode == 1, QFX_ERR_BAD_ARG, "mode must be 0 (weights) or 1 (delta)");
    QFX_ARG_CHECK(!(a.mode == 0 && (a.clip > 0.f || a.sigma > 0.f)), QFX_ERR_BAD_ARG,
                  "clip/noise need delta mode");
    QFX_ARG_CHECK(a.world >= 1 && a.world <= AGG_MAX_WORLD && a.rank >= 0 && a.rank < a.world, QFX_ERR_BAD_ARG,
                  "bad rank/world");
    QFX_ARG_CHECK(a.world == 1 || (a.recv_slots && a.flags), QFX_ERR_BAD_ARG, "multi-GPU needs slots and flags");
    QFX_ARG_CHECK(a.scratch != nullptr, QFX_ERR_BAD_ARG, "scratch is required");
    QFX_ARG_CHECK((size_t)a.k_local * sizeof(float) <= 160 * 1024, QFX_ERR_TOO_LARGE, "too many local clients");
    if (g_num_sms == 0) {
        int dev = 0;
        QFX_CUDA_TRY(cudaGetDevice(&dev));
        QFX_CUDA_TRY(cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev));
    }
    if (a.k_local > 0) QFX_CUDA_TRY(cudaMemsetAsync(a.scratch, 0, sizeof(float) * a.k_local, stream));
    const size_t smem = sizeof(float) * a.k_local;
    if (smem > 48 * 1024)
        QFX_CUDA_TRY(cudaFuncSetAttribute(fedavg_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int blocks = (a.dim + AGG_THREADS - 1) / AGG_THREADS;
    const int cap = g_num_sms * 2;
    if (blocks > cap) blocks = cap;
    void* params[] = {(void*)&a};
    QFX_CUDA_TRY(cudaLaunchCooperativeKernel((const void*)fedavg_fused_kernel, dim3(blocks), dim3(AGG_THREADS),
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
