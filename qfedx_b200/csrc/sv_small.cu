// K1 (fused feature map + ansatz), K3 (<Z> readout), K4a (adjoint backward), K4b (parameter shift)
// for n <= 10 qubits, batched over circuits.  See sv_regs.cuh for the register tiling.
//
// Batch axis: one circuit per group of LANES lanes; 128 threads per CTA.  Parameters are
// theta[G][L][n][2] with an optional per-circuit group id (clients share a group).
#include "sv_regs.cuh"
#include "../../include/qfedx_b200.h"

namespace qfx {

constexpr int SV_THREADS = 128;
constexpr int MAX_PARAMS_SMEM = 2 * 10 * 64;  // n <= 10, L <= 64

template <int NQ, int LB_>
__device__ __forceinline__ void load_initial(Sv<NQ, LB_>& psi, const float* angles, const float* init_state, int basis,
                                             int b, int lg) {
    if (init_state != nullptr) psi.set_real(init_state + (size_t)b * SvCfg<NQ>::DIM, lg);
    else psi.set_product(angles + (size_t)b * NQ, basis, lg);
}

__device__ __forceinline__ void fill_trig_table(float2* tab, const float* theta, int n_params) {
    for (int p = threadIdx.x; p < n_params; p += blockDim.x) {
        float s, c;
        sincosf(0.5f * theta[p], &s, &c);
        tab[p] = make_float2(c, s);
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------------------
template <int NQ>
__global__ void __launch_bounds__(SV_THREADS)
sv_forward_kernel(int n_layers, int basis, int ent, const float* __restrict__ angles,
                  const float* __restrict__ init_state, const float* __restrict__ theta,
                  const int* __restrict__ gid, int B, float* __restrict__ expvals, float2* __restrict__ state_out) {
    using C = SvCfg<NQ>;
    __shared__ float2 tab[MAX_PARAMS_SMEM];
    extern __shared__ __align__(16) float2 dyn[];   // merged path: diag tables | per-circuit ring scratch
    const int n_params = 2 * NQ * n_layers;
    if (gid == nullptr) fill_trig_table(tab, theta, n_params);
    const int lg = threadIdx.x & (C::LANES - 1);
    const int circ = (blockIdx.x * SV_THREADS + threadIdx.x) >> C::LB;
    const bool live = circ < B;
    const int b = live ? circ : B - 1;  // dead lanes shadow a real circuit so shuffles stay converged

    Sv<NQ> psi;
    load_initial(psi, angles, init_state, basis, b, lg);
    if (gid == nullptr) {
        // one parameter group per launch: merged RZ diagonal and the CNOT ring as one permutation
        float2* dloc = dyn;
        float2* dlane = dloc + n_layers * C::A;
        float2* ring = dlane + n_layers * C::LANES + (size_t)(threadIdx.x >> C::LB) * C::DIM;
        build_diag_tables<NQ, C::LB>(dloc, dlane, tab, n_layers, threadIdx.x, SV_THREADS);
        __syncthreads();
        const DiagTables dt{dloc, dlane};
        ansatz_forward_merged(psi, tab, dt, n_layers, ent, lg, ring);
    } else {
        TrigDirect trig{theta + (size_t)gid[b] * n_params, -1, 0.f};
        ansatz_forward(psi, trig, n_layers, ent, lg);
    }
    float ez[NQ];
    psi.expval_z_all(ez, lg);
    if (live && lg == 0 && expvals != nullptr) {
#pragma unroll
        for (int q = 0; q < NQ; ++q) expvals[(size_t)b * NQ + q] = ez[q];
    }
    if (live && state_out != nullptr) {
#pragma unroll
        for (int i = 0; i < C::A; ++i)
            state_out[(size_t)b * C::DIM + ((lg << C::NL) | i)] = make_float2(psi.re[i], psi.im[i]);
    }
}

// ---------------------------------------------------------------------------------------
// Gradient sink: reduce over the circuit's lanes, then one atomic per circuit per parameter
// (shared-memory accumulator when the whole CTA works for one group, global otherwise).
template <int LANES>
struct AtomicSink {
    float* dst;  // [n_params] accumulator (shared or global)
    bool live;
    int lg;
    __device__ __forceinline__ void operator()(int p, float v) {
        v = group_sum<LANES>(v);
        if (live && lg == 0) atomicAdd(dst + p, v);
    }
};

// Sink of the merged backward sweep: one layer's 2 NQ gradient shares are reduced over the warp together and
// added to the CTA's shared-memory accumulator (one atomic per parameter and warp).
struct WarpLayerSink {
    float* dst;
    bool live;
    int lane;
    template <int NQ, int NG>
    __device__ __forceinline__ void layer(int l, float (&g)[NG]) {
        if (!live) {
#pragma unroll
            for (int i = 0; i < NG; ++i) g[i] = 0.f;
        }
        warp_reduce_many<NG>(g, lane);
        const int slot = slot_of_lane<NG>(lane);
        if (slot < 2 * NQ && (NG == 32 || (lane & 1) == 0))
            atomicAdd(dst + (slot < NQ ? (l * NQ + slot) * 2 : (l * NQ + slot - NQ) * 2 + 1), g[0]);
    }
};

template <int NQ>
__global__ void __launch_bounds__(SV_THREADS)
sv_adjoint_kernel(int n_layers, int basis, int ent, const float* __restrict__ angles,
                  const float* __restrict__ init_state, const float* __restrict__ theta,
                  const int* __restrict__ gid, int B, const float* __restrict__ cot,
                  float* __restrict__ grad_theta, float* __restrict__ grad_angles) {
    using C = SvCfg<NQ>;
    __shared__ float2 tab[MAX_PARAMS_SMEM];
    __shared__ float sgrad[MAX_PARAMS_SMEM];
    extern __shared__ __align__(16) float2 dyn[];
    const int n_params = 2 * NQ * n_layers;
    if (gid == nullptr) {
        for (int p = threadIdx.x; p < n_params; p += blockDim.x) sgrad[p] = 0.f;
        fill_trig_table(tab, theta, n_params);
    }
    const int lg = threadIdx.x & (C::LANES - 1);
    const int circ = (blockIdx.x * SV_THREADS + threadIdx.x) >> C::LB;
    const bool live = circ < B;
    const int b = live ? circ : B - 1;

    Sv<NQ> psi, lam;
    load_initial(psi, angles, init_state, basis, b, lg);
    float cq[NQ];
#pragma unroll
    for (int q = 0; q < NQ; ++q) cq[q] = cot[(size_t)b * NQ + q];

    if (gid == nullptr) {
        float2* dloc = dyn;
        float2* dlane = dloc + n_layers * C::A;
        float2* ring = dlane + n_layers * C::LANES + (size_t)(threadIdx.x >> C::LB) * C::DIM;
        build_diag_tables<NQ, C::LB>(dloc, dlane, tab, n_layers, threadIdx.x, SV_THREADS);
        __syncthreads();
        const DiagTables dt{dloc, dlane};
        ansatz_forward_merged(psi, tab, dt, n_layers, ent, lg, ring);
        lam.set_obs_times(psi, cq, lg);
        WarpLayerSink sink{sgrad, live, (int)(threadIdx.x & 31)};
        ansatz_backward_merged(psi, lam, tab, dt, n_layers, ent, lg, sink, ring);
    } else {
        const size_t off = (size_t)gid[b] * n_params;
        TrigDirect trig{theta + off, -1, 0.f};
        ansatz_forward(psi, trig, n_layers, ent, lg);
        lam.set_obs_times(psi, cq, lg);
        AtomicSink<C::LANES> sink{grad_theta + off, live, lg};
        ansatz_backward(psi, lam, trig, n_layers, ent, lg, sink);
    }
    if (grad_angles != nullptr && init_state == nullptr) {
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            float v = group_sum<C::LANES>(im_lam_p_psi(lam, psi, basis, q, lg));
            if (live && lg == 0) grad_angles[(size_t)b * NQ + q] = v;
        }
    }
    if (gid == nullptr) {
        __syncthreads();
        for (int p = threadIdx.x; p < n_params; p += blockDim.x) atomicAdd(grad_theta + p, sgrad[p]);
    }
}

// ---------------------------------------------------------------------------------------
// Parameter shift: work item = (sample b, parameter p, sign).  The 2P shifted circuits of one
// sample are an extra batch axis; each evaluates sum_q cot[b][q] <Z_q> and
// contributes +-0.5 of it to grad[p].
template <int NQ>
__global__ void __launch_bounds__(SV_THREADS)
sv_param_shift_kernel(int n_layers, int basis, int ent, const float* __restrict__ angles,
                      const float* __restrict__ init_state, const float* __restrict__ theta,
                      const int* __restrict__ gid, int B, const float* __restrict__ cot,
                      float* __restrict__ grad_theta) {
    using C = SvCfg<NQ>;
    __shared__ float2 tab[MAX_PARAMS_SMEM];
    const int n_params = 2 * NQ * n_layers;
    if (gid == nullptr) fill_trig_table(tab, theta, n_params);
    const int lg = threadIdx.x & (C::LANES - 1);
    const long long total = (long long)B * n_params * 2;
    const long long item = ((long long)blockIdx.x * SV_THREADS + threadIdx.x) >> C::LB;
    const bool live = item < total;
    const long long it = live ? item : total - 1;
    const int sign_i = (int)(it & 1);
    const int p = (int)((it >> 1) % n_params);
    const int b = (int)((it >> 1) / n_params);
    const float sgn = sign_i ? -1.f : 1.f;

    Sv<NQ> psi;
    load_initial(psi, angles, init_state, basis, b, lg);
    size_t off = 0;
    if (gid == nullptr) {
        TrigTable trig{tab, p, sgn};
        ansatz_forward(psi, trig, n_layers, ent, lg);
    } else {
        off = (size_t)gid[b] * n_params;
        TrigDirect trig{theta + off, p, sgn};
        ansatz_forward(psi, trig, n_layers, ent, lg);
    }
    float ez[NQ];
    psi.expval_z_all(ez, lg);
    float v = 0.f;
#pragma unroll
    for (int q = 0; q < NQ; ++q) v += cot[(size_t)b * NQ + q] * ez[q];
    if (live && lg == 0) atomicAdd(grad_theta + off + p, 0.5f * sgn * v);
}

// ---------------------------------------------------------------------------------------
template <template <int> class Launcher, typename... Args>
static int dispatch_nq(int n_qubits, Args&&... args) {
    switch (n_qubits) {
        case 1: return Launcher<1>::run(args...);
        case 2: return Launcher<2>::run(args...);
        case 3: return Launcher<3>::run(args...);
        case 4: return Launcher<4>::run(args...);
        case 5: return Launcher<5>::run(args...);
        case 6: return Launcher<6>::run(args...);
        case 7: return Launcher<7>::run(args...);
        case 8: return Launcher<8>::run(args...);
        case 9: return Launcher<9>::run(args...);
        case 10: return Launcher<10>::run(args...);
    }
    set_last_error("register-tiled kernels support 1..10 qubits, got %d", n_qubits);
    return QFX_ERR_UNSUPPORTED;
}

static inline unsigned blocks_for(long long circuits, int lanes) {
    long long threads = circuits * lanes;
    return (unsigned)((threads + SV_THREADS - 1) / SV_THREADS);
}

// dynamic shared memory of the merged path: diag tables + one 2^n float2 ring buffer per circuit of the CTA
template <int NQ>
static size_t merged_smem(int L, const int* gid) {
    using C = SvCfg<NQ>;
    if (gid != nullptr) return 0;
    return sizeof(float2) * ((size_t)L * (C::A + C::LANES) + (size_t)(SV_THREADS >> C::LB) * C::DIM);
}

template <int NQ>
struct FwdLaunch {
    static int run(int L, int basis, int ent, const float* angles, const float* init, const float* theta,
                   const int* gid, int B, float* expvals, float2* state_out, cudaStream_t st) {
        const size_t smem = merged_smem<NQ>(L, gid);
        if (smem > 24 * 1024)
            QFX_CUDA_TRY(cudaFuncSetAttribute(sv_forward_kernel<NQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        sv_forward_kernel<NQ><<<blocks_for(B, SvCfg<NQ>::LANES), SV_THREADS, smem, st>>>(
            L, basis, ent, angles, init, theta, gid, B, expvals, state_out);
        return check_launch("sv_forward_kernel");
    }
};
template <int NQ>
struct AdjLaunch {
    static int run(int L, int basis, int ent, const float* angles, const float* init, const float* theta,
                   const int* gid, int B, const float* cot, float* gth, float* gang, cudaStream_t st) {
        const size_t smem = merged_smem<NQ>(L, gid);
        if (smem > 24 * 1024)
            QFX_CUDA_TRY(cudaFuncSetAttribute(sv_adjoint_kernel<NQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        sv_adjoint_kernel<NQ><<<blocks_for(B, SvCfg<NQ>::LANES), SV_THREADS, smem, st>>>(
            L, basis, ent, angles, init, theta, gid, B, cot, gth, gang);
        return check_launch("sv_adjoint_kernel");
    }
};
template <int NQ>
struct ShiftLaunch {
    static int run(int L, int basis, int ent, const float* angles, const float* init, const float* theta,
                   const int* gid, int B, const float* cot, float* gth, cudaStream_t st) {
        const long long items = (long long)B * 2 * NQ * L * 2;
        sv_param_shift_kernel<NQ><<<blocks_for(items, SvCfg<NQ>::LANES), SV_THREADS, 0, st>>>(
            L, basis, ent, angles, init, theta, gid, B, cot, gth);
        return check_launch("sv_param_shift_kernel");
    }
};

static int check_common(int n_qubits, int n_layers, int basis, int ent, const void* angles, const void* init,
                        const void* theta, int B) {
    QFX_ARG_CHECK(n_qubits >= 1 && n_qubits <= 10, QFX_ERR_UNSUPPORTED,
                  "register-tiled kernels support 1..10 qubits, got %d", n_qubits);
    QFX_ARG_CHECK(n_layers >= 0 && 2 * n_qubits * n_layers <= MAX_PARAMS_SMEM, QFX_ERR_TOO_LARGE,
                  "2*n*L = %d exceeds the %d-parameter table", 2 * n_qubits * n_layers, MAX_PARAMS_SMEM);
    QFX_ARG_CHECK(basis >= 0 && basis <= 2 && ent >= 0 && ent <= 2, QFX_ERR_BAD_ARG, "bad basis/entangler code");
    QFX_ARG_CHECK((angles != nullptr) != (init != nullptr), QFX_ERR_BAD_ARG,
                  "give exactly one of angles / init_state");
    QFX_ARG_CHECK(theta != nullptr && B > 0, QFX_ERR_BAD_ARG, "null theta or empty batch");
    return 0;
}

}  // namespace qfx

using namespace qfx;

extern "C" int qfx_sv_forward(int n_qubits, int n_layers, int basis, int entangler, const float* angles,
                              const float* init_state, const float* theta, const int* gid, int batch,
                              float* expvals, float* state_out, cudaStream_t stream) {
    if (int e = check_common(n_qubits, n_layers, basis, entangler, angles, init_state, theta, batch)) return e;
    return dispatch_nq<FwdLaunch>(n_qubits, n_layers, basis, entangler, angles, init_state, theta, gid, batch,
                                  expvals, reinterpret_cast<float2*>(state_out), stream);
}

extern "C" int qfx_sv_adjoint_backward(int n_qubits, int n_layers, int basis, int entangler, const float* angles,
                                       const float* init_state, const float* theta, const int* gid, int batch,
                                       const float* cot, float* grad_theta, float* grad_angles,
                                       cudaStream_t stream) {
    if (int e = check_common(n_qubits, n_layers, basis, entangler, angles, init_state, theta, batch)) return e;
    QFX_ARG_CHECK(cot != nullptr && grad_theta != nullptr, QFX_ERR_BAD_ARG, "null cot / grad_theta");
    return dispatch_nq<AdjLaunch>(n_qubits, n_layers, basis, entangler, angles, init_state, theta, gid, batch, cot,
                                  grad_theta, grad_angles, stream);
}

extern "C" int qfx_sv_param_shift(int n_qubits, int n_layers, int basis, int entangler, const float* angles,
                                  const float* init_state, const float* theta, const int* gid, int batch,
                                  const float* cot, float* grad_theta, cudaStream_t stream) {
    if (int e = check_common(n_qubits, n_layers, basis, entangler, angles, init_state, theta, batch)) return e;
    QFX_ARG_CHECK(cot != nullptr && grad_theta != nullptr, QFX_ERR_BAD_ARG, "null cot / grad_theta");
    QFX_ARG_CHECK(n_layers > 0, QFX_ERR_BAD_ARG, "parameter shift needs at least one layer");
    return dispatch_nq<ShiftLaunch>(n_qubits, n_layers, basis, entangler, angles, init_state, theta, gid, batch,
                                    cot, grad_theta, stream);
}
