// Whole-client local training in ONE launch (the quantum counterpart of the reference's
// client_update, /root/reference/src/CFed/Classical_FL.py:40-64, for the VQC planned in
// /root/reference/ROADMAP.md:29-43,126-135).
//
// One thread-block cluster per selected client.  The cluster walks E epochs x ceil(n_k/bs)
// minibatches; per minibatch every circuit slot runs feature map -> ansatz -> <Z> -> logits ->
// loss -> adjoint reverse sweep (or 2P parameter-shift evaluations) entirely in registers,
// gradients are reduced in shared memory and across the cluster through DSMEM, and every CTA
// applies the identical Adam / SGD-momentum update to its own shared-memory copy of the
// parameters.  Nothing but the final parameters and two statistics leaves the SM, so a client
// costs one launch per round instead of (steps x kernels-per-step) launches.
#include <cooperative_groups.h>
#include "sv_regs.cuh"
#include "../../include/qfedx_b200.h"

namespace cg = cooperative_groups;

namespace qfx {

constexpr int TR_THREADS = 256;
constexpr int TR_MAX_P = 2 * 10 * 64;
constexpr int TR_MAX_C = 10;
constexpr int TR_MAX_D = TR_MAX_P + 2 * TR_MAX_C;
constexpr int TR_MAX_BS_SHIFT = 256;

struct TrainArgs {
    int n_layers, basis, ent, n_classes;
    const float* angles;      // [N, n]
    const int* labels;        // [N]
    const int* offsets;       // [K+1] client segment bounds into the sample axis
    const int* perm;          // [E, N] sample order per epoch (global indices) or null
    int n_total;
    const float* params_in;   // [D] global parameters (theta | a | b)
    const int* client_ids;    // [K_sel]
    float* params_out;        // [K_sel, D]
    float* stats_out;         // [K_sel, 2] mean loss / accuracy over the last epoch
    int epochs, batch_size;
    int opt;                  // 0 SGD-momentum, 1 Adam
    float lr, beta1, beta2, eps;
    int grad_mode;            // 0 adjoint, 1 parameter shift
    int ring_offset_floats;   // where the CNOT-ring scratch starts in dynamic shared memory (set by the launcher)
    unsigned long long shuffle_key;   // non-zero and perm == null: samples are visited in a keyed Feistel order
    ReadoutNoise rn;          // readout-level noise applied to <Z> before the logits (alpha 1, beta 0, shots 0 = exact)
    long long* prof;          // [n_selected][12] cycle counters per phase (thread 0 of each cluster's CTA 0), or null
};

static long long* g_train_prof[QFX_MAX_DEVICES] = {};

// Gradient sink: reduce over the whole warp (dead circuits contribute zero) and add into the warp's private
// row of the shared-memory gradient matrix - no atomics, no contention.  `first` marks the first pass of a
// minibatch: the row is then stored instead of read-modified-written, which takes the shared-memory load
// latency off the warp's in-order critical path.
struct WarpRowSink {
    float* row;
    bool live;
    int lane;
    bool first;
    __device__ __forceinline__ void operator()(int p, float v) {   // single value (readout / statistics)
        v = warp_sum(live ? v : 0.f);
        if (lane == 0) { if (first) row[p] = v; else row[p] += v; }
    }
    // the same when the warp is ONE circuit and `live` is set on its lane 0 only: nothing to reduce
    __device__ __forceinline__ void single(int p, float v) {
        if (live) { if (first) row[p] = v; else row[p] += v; }
    }
    // one layer's 2 NQ gradient shares: slot q -> parameter (l NQ + q) 2, slot NQ + q -> (l NQ + q) 2 + 1
    template <int NQ, int NG>
    __device__ __forceinline__ void layer(int l, float (&g)[NG]) {
        if (!live) {
#pragma unroll
            for (int i = 0; i < NG; ++i) g[i] = 0.f;
        }
        warp_reduce_many<NG>(g, lane);
        const int slot = slot_of_lane<NG>(lane);
        if (slot < 2 * NQ && (NG == 32 || (lane & 1) == 0)) {
            const int p = slot < NQ ? (l * NQ + slot) * 2 : (l * NQ + slot - NQ) * 2 + 1;
            if (first) row[p] = g[0]; else row[p] += g[0];
        }
    }
};

// loss and d loss / d logits for one sample; z has n_classes entries (1 => BCE-with-logits)
__device__ __forceinline__ float loss_and_grad(const float* z, float* dz, int n_classes, int y, bool* correct) {
    if (n_classes == 1) {
        const float x = z[0];
        const float sig = 1.f / (1.f + expf(-x));
        dz[0] = sig - (float)y;
        *correct = ((x > 0.f) ? 1 : 0) == y;
        return fmaxf(x, 0.f) - x * (float)y + log1pf(expf(-fabsf(x)));
    }
    float m = z[0];
    int arg = 0;
    for (int j = 1; j < n_classes; ++j)
        if (z[j] > m) { m = z[j]; arg = j; }
    float sum = 0.f;
    for (int j = 0; j < n_classes; ++j) { dz[j] = expf(z[j] - m); sum += dz[j]; }
    const float inv = 1.f / sum;
    float zy = 0.f;
    for (int j = 0; j < n_classes; ++j) {
        dz[j] *= inv;
        if (j == y) { zy = z[j]; dz[j] -= 1.f; }
    }
    *correct = arg == y;
    return m + logf(sum) - zy;
}

template <int NQ, int LB_>
__global__ void __launch_bounds__(TR_THREADS)
vqc_local_train_kernel(const TrainArgs a) {
    using C = SvCfg<NQ, LB_>;
    cg::cluster_group cluster = cg::this_cluster();
    const int cs = (int)cluster.num_blocks();
    const int crank = (int)cluster.block_rank();
    const int sel = blockIdx.x / cs;
    const int client = a.client_ids[sel];
    const int P = 2 * NQ * a.n_layers;
    const int NC = a.n_classes;
    const int D = P + 2 * NC;

    // dynamic shared memory: tab float2[P] | dloc float2[L*A] | dlane float2[L*LANES] | prm[D] | m1[D] |
    // m2[D] | sgrad[NW][DS] | scot[bs*NQ]
    extern __shared__ __align__(16) float smem[];
    constexpr int NW = TR_THREADS / 32;
    const int DS = (D + 2 + 3) & ~3;  // row stride; [D] loss sum, [D+1] correct count
    float2* tab = reinterpret_cast<float2*>(smem);
    float2* dloc = tab + P;
    float2* dlane = dloc + a.n_layers * C::A;
    float* prm = reinterpret_cast<float*>(dlane + a.n_layers * C::LANES);
    const DiagTables dt{dloc, dlane};
    float* m1 = prm + D;
    float* m2 = m1 + D;
    float* sgrad = m2 + D;
    float* scot = sgrad + NW * DS;
    // per-circuit scratch for the permutation form of the CNOT ring (8-byte aligned, after everything else)
    float2* ring_all = reinterpret_cast<float2*>(smem + a.ring_offset_floats);
    float2* ring_buf = ring_all + (size_t)(threadIdx.x >> C::LB) * C::DIM;

    const int tid = threadIdx.x;
    // phase timers (scripts/profile_vqc_train.py): 0 minibatch set-up, 1 forward + readout, 2 loss + readout sinks,
    // 3 backward, 4 wait for the cluster, 5 gradient reduction (DSMEM), 6 second cluster barrier, 7 optimizer,
    // 8 diagonal tables
    long long* prof = (a.prof && tid == 0 && crank == 0) ? a.prof + (size_t)sel * 12 : nullptr;
    long long t_mark = prof ? clock64() : 0;
#define VQC_MARK(id)                                   \
    do {                                               \
        if (prof) {                                    \
            const long long _t = clock64();            \
            prof[id] += _t - t_mark;                   \
            t_mark = _t;                               \
        }                                              \
    } while (0)
    for (int p = tid; p < D; p += TR_THREADS) {
        const float v = a.params_in[p];
        prm[p] = v; m1[p] = 0.f; m2[p] = 0.f;
        if (p < P) { float s, c; sincosf(0.5f * v, &s, &c); tab[p] = make_float2(c, s); }
    }
    __syncthreads();
    build_diag_tables<NQ, LB_>(dloc, dlane, tab, a.n_layers, tid, TR_THREADS);
    __syncthreads();

    const int lg = tid & (C::LANES - 1);
    const int slots_cta = TR_THREADS >> C::LB;
    const int slots_all = slots_cta * cs;
    // circuit slots are dealt round-robin over the cluster's CTAs, so a minibatch smaller than the cluster's slot count
    // leaves every CTA the same number of idle warps instead of leaving whole CTAs idle: a cluster wider than one
    // pass needs then shortens the step (fewer circuits share an SM's issue slots)
    const int slot = (tid >> C::LB) * cs + crank;

    const int seg_lo = a.offsets[client];
    const int n_k = a.offsets[client + 1] - seg_lo;
    const int n_batches = (n_k + a.batch_size - 1) / a.batch_size;
    float epoch_loss = 0.f, epoch_correct = 0.f;
    float b1t = 1.f, b2t = 1.f;  // beta^t for the Adam bias correction

    for (int e = 0; e < a.epochs; ++e) {
        epoch_loss = 0.f; epoch_correct = 0.f;
        const int* perm_e = a.perm ? a.perm + (size_t)e * a.n_total + seg_lo : nullptr;
        const unsigned long long fkey = a.shuffle_key ? key3(a.shuffle_key, (unsigned long long)client, (unsigned long long)e) : 0ull;
        for (int bi = 0; bi < n_batches; ++bi) {
            const int b_lo = bi * a.batch_size;
            const int nb = min(a.batch_size, n_k - b_lo);
            for (int p = tid; p < NW * DS; p += TR_THREADS) sgrad[p] = 0.f;
            __syncthreads();
            float* my_row = sgrad + (tid >> 5) * DS;
            const int lane = tid & 31;
            VQC_MARK(0);

            const int n_pass = (nb + slots_all - 1) / slots_all;
            for (int ps = 0; ps < n_pass; ++ps) {
                const int j = ps * slots_all + slot;
                const bool live = j < nb;
                // a warp that IS one circuit and has no sample left skips the simulation altogether: its gradient row
                // keeps the zeros written above (nothing below is a block-wide barrier)
                if (C::LANES == 32 && !live && a.grad_mode == 0) continue;
                const int jj = live ? j : nb - 1;
                const int sidx = perm_e ? perm_e[b_lo + jj]
                                        : seg_lo + (fkey ? (int)feistel_perm((uint32_t)(b_lo + jj), (uint32_t)n_k, fkey) : b_lo + jj);
                const int y = a.labels[sidx];

                Sv<NQ, LB_> psi;
                psi.set_product(a.angles + (size_t)sidx * NQ, a.basis, lg);
                ansatz_forward_merged(psi, tab, dt, a.n_layers, a.ent, lg, ring_buf);
                float ez[NQ];
                psi.expval_z_all(ez, lg);
                VQC_MARK(1);
                if (a.rn.alpha != 1.f || a.rn.beta != 0.f || a.rn.shots > 0) {
                    // noisy readout of the logit qubits; shot noise is keyed by (epoch, sample) so that every lane of the
                    // circuit draws the same outcome
#pragma unroll
                    for (int q = 0; q < NQ; ++q)
                        if (q < NC) ez[q] = readout_noise(ez[q], a.rn, ((unsigned long long)e << 32) | (unsigned)sidx, q);
                }

                float z[TR_MAX_C], dz[TR_MAX_C];
#pragma unroll
                for (int q = 0; q < NQ; ++q)
                    if (q < NC) z[q] = prm[P + q] * ez[q] + prm[P + NC + q];
                bool correct;
                const float loss = loss_and_grad(z, dz, NC, y, &correct);
                float cot[NQ];   // d loss / d <Z_q> of the EXACT expectation: through the affine part of the noise
#pragma unroll
                for (int q = 0; q < NQ; ++q) cot[q] = (q < NC) ? dz[q] * prm[P + q] * a.rn.alpha : 0.f;
                {
                    // readout and statistics: one contribution per circuit (its lane 0)
                    WarpRowSink once{my_row, live && lg == 0, lane, ps == 0};
                    if (C::LANES == 32) {
#pragma unroll
                        for (int q = 0; q < NQ; ++q)
                            if (q < NC) {
                                once.single(P + q, dz[q] * ez[q]);
                                once.single(P + NC + q, dz[q]);
                            }
                        once.single(D, loss);
                        once.single(D + 1, correct ? 1.f : 0.f);
                    } else {
#pragma unroll
                        for (int q = 0; q < NQ; ++q)
                            if (q < NC) {
                                once(P + q, dz[q] * ez[q]);
                                once(P + NC + q, dz[q]);
                            }
                        once(D, loss);
                        once(D + 1, correct ? 1.f : 0.f);
                    }
                }
                VQC_MARK(2);
                if (a.grad_mode == 0) {
                    Sv<NQ, LB_> lam;
                    lam.set_obs_times(psi, cot, lg);
                    WarpRowSink sink{my_row, live, lane, ps == 0};
                    ansatz_backward_merged(psi, lam, tab, dt, a.n_layers, a.ent, lg, sink, ring_buf);
                    VQC_MARK(3);
                } else if (live && lg == 0) {
#pragma unroll
                    for (int q = 0; q < NQ; ++q) scot[j * NQ + q] = cot[q];
                }
            }
            if (a.grad_mode == 1) {
                // 2P shifted circuits per sample, spread over the cluster's circuit slots.
                // Each CTA needs every sample's cotangent: with cs > 1 a CTA only evaluated its
                // own slots above, so it re-derives missing ones by reading the owner's scot.
                cluster.sync();
                const int items = nb * P * 2;
                const int n_pass2 = (items + slots_all - 1) / slots_all;
                for (int ps = 0; ps < n_pass2; ++ps) {
                    const int it0 = ps * slots_all + slot;
                    const bool live = it0 < items;
                    const int it = live ? it0 : items - 1;
                    const float sgn = (it & 1) ? -1.f : 1.f;
                    const int p = (it >> 1) % P;
                    const int j = (it >> 1) / P;
                    const int sidx = perm_e ? perm_e[b_lo + j]
                                            : seg_lo + (fkey ? (int)feistel_perm((uint32_t)(b_lo + j), (uint32_t)n_k, fkey) : b_lo + j);
                    const int owner = (j % slots_all) % cs;
                    const float* cot_src = cluster.map_shared_rank(scot, owner);
                    Sv<NQ, LB_> psi;
                    psi.set_product(a.angles + (size_t)sidx * NQ, a.basis, lg);
                    TrigTable trig{tab, p, sgn};
                    ansatz_forward(psi, trig, a.n_layers, a.ent, lg);
                    float ez[NQ];
                    psi.expval_z_all(ez, lg);
                    float v = 0.f;
#pragma unroll
                    for (int q = 0; q < NQ; ++q) v += cot_src[j * NQ + q] * ez[q];
                    // p differs between the circuits of a warp here, so this one stays an atomic
                    if (live && lg == 0) atomicAdd(&sgrad[p], 0.5f * sgn * v);
                }
            }

            // ---- reduce over the cluster and update -------------------------------------
            // the CTA's warp rows are summed into row 0 locally first, so that only one row per CTA crosses the cluster
            __syncthreads();
            for (int p = tid; p < D + 2; p += TR_THREADS) {
                float acc = 0.f;
#pragma unroll
                for (int w = 0; w < NW; ++w) acc += sgrad[w * DS + p];
                sgrad[p] = acc;
            }
            if (cs > 1) cluster.sync(); else __syncthreads();
            VQC_MARK(4);
            float gsum[(TR_MAX_D + 2 + TR_THREADS - 1) / TR_THREADS];
#pragma unroll
            for (int r = 0; r < (TR_MAX_D + 2 + TR_THREADS - 1) / TR_THREADS; ++r) {
                const int p = tid + r * TR_THREADS;
                float acc = 0.f;
                if (p < D + 2) {
                    if (cs > 1) {
                        for (int c = 0; c < cs; ++c) acc += cluster.map_shared_rank(sgrad, c)[p];
                    } else {
                        acc = sgrad[p];
                    }
                }
                gsum[r] = acc;
            }
            VQC_MARK(5);
            if (cs > 1) cluster.sync(); else __syncthreads();
            VQC_MARK(6);
            b1t *= a.beta1; b2t *= a.beta2;
            const float inv_nb = 1.f / (float)nb;
#pragma unroll
            for (int r = 0; r < (TR_MAX_D + 2 + TR_THREADS - 1) / TR_THREADS; ++r) {
                const int p = tid + r * TR_THREADS;
                if (p < D) {
                    const float g = gsum[r] * inv_nb;
                    float w = prm[p];
                    if (a.opt == 0) {
                        const float mom = a.beta1 * m1[p] + g;
                        m1[p] = mom;
                        w -= a.lr * mom;
                    } else {
                        const float mm = a.beta1 * m1[p] + (1.f - a.beta1) * g;
                        const float vv = a.beta2 * m2[p] + (1.f - a.beta2) * g * g;
                        m1[p] = mm; m2[p] = vv;
                        w -= a.lr * (mm / (1.f - b1t)) / (sqrtf(vv / (1.f - b2t)) + a.eps);
                    }
                    prm[p] = w;
                    if (p < P) { float s, c; sincosf(0.5f * w, &s, &c); tab[p] = make_float2(c, s); }
                } else if (p == D) {
                    epoch_loss += gsum[r];
                } else if (p == D + 1) {
                    epoch_correct += gsum[r];
                }
            }
            VQC_MARK(7);
            __syncthreads();
            build_diag_tables<NQ, LB_>(dloc, dlane, tab, a.n_layers, tid, TR_THREADS);
            __syncthreads();
            VQC_MARK(8);
        }
    }
#undef VQC_MARK
    if (crank == 0) {
        for (int p = tid; p < D; p += TR_THREADS) a.params_out[(size_t)sel * D + p] = prm[p];
        // the threads that own slots D and D+1 hold the epoch statistics
        const int pl = D % TR_THREADS, pc = (D + 1) % TR_THREADS;
        if (tid == pl) a.stats_out[sel * 2 + 0] = n_k > 0 ? epoch_loss / (float)n_k : 0.f;
        if (tid == pc) a.stats_out[sel * 2 + 1] = n_k > 0 ? epoch_correct / (float)n_k : 0.f;
    }
    if (cs > 1) cluster.sync();  // keep every CTA's shared memory alive until peers are done reading
}

// ---------------------------------------------------------------------------------------
// Evaluation in one launch: forward over a dataset, loss sum and correct count.
template <int NQ>
__global__ void __launch_bounds__(TR_THREADS)
vqc_eval_kernel(int n_layers, int basis, int ent, int n_classes, const float* __restrict__ angles,
                const int* __restrict__ labels, const float* __restrict__ params, int N, float* __restrict__ out,
                const ReadoutNoise rn) {
    using C = SvCfg<NQ>;
    __shared__ float2 tab[TR_MAX_P];
    __shared__ float red[2];
    const int P = 2 * NQ * n_layers;
    const int tid = threadIdx.x;
    if (tid < 2) red[tid] = 0.f;
    for (int p = tid; p < P; p += TR_THREADS) {
        float s, c; sincosf(0.5f * params[p], &s, &c); tab[p] = make_float2(c, s);
    }
    __syncthreads();
    const int lg = tid & (C::LANES - 1);
    const int slots_cta = TR_THREADS >> C::LB;
    const int slots_all = slots_cta * gridDim.x;
    const int slot = blockIdx.x * slots_cta + (tid >> C::LB);
    const int n_pass = (N + slots_all - 1) / slots_all;
    float loss_acc = 0.f, corr_acc = 0.f;
    for (int ps = 0; ps < n_pass; ++ps) {
        const int j = ps * slots_all + slot;
        const bool live = j < N;
        const int jj = live ? j : N - 1;
        Sv<NQ> psi;
        psi.set_product(angles + (size_t)jj * NQ, basis, lg);
        TrigTable trig{tab, -1, 0.f};
        ansatz_forward(psi, trig, n_layers, ent, lg);
        float ez[NQ];
        psi.expval_z_all(ez, lg);
        if (rn.alpha != 1.f || rn.beta != 0.f || rn.shots > 0) {
#pragma unroll
            for (int q = 0; q < NQ; ++q)
                if (q < n_classes) ez[q] = readout_noise(ez[q], rn, (unsigned long long)jj, q);
        }
        float z[TR_MAX_C], dz[TR_MAX_C];
#pragma unroll
        for (int q = 0; q < NQ; ++q)
            if (q < n_classes) z[q] = params[P + q] * ez[q] + params[P + n_classes + q];
        bool correct;
        const float loss = loss_and_grad(z, dz, n_classes, labels[jj], &correct);
        if (live && lg == 0) { loss_acc += loss; corr_acc += correct ? 1.f : 0.f; }
    }
    loss_acc = warp_sum(loss_acc);
    corr_acc = warp_sum(corr_acc);
    if ((tid & 31) == 0) { atomicAdd(&red[0], loss_acc); atomicAdd(&red[1], corr_acc); }
    __syncthreads();
    if (tid < 2) atomicAdd(out + tid, red[tid]);
}

template <int NQ, int LB_>
static int launch_train(const TrainArgs& a, int n_sel, int cluster_size, cudaStream_t st) {
    using C = SvCfg<NQ, LB_>;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(n_sel * cluster_size));
    cfg.blockDim = dim3(TR_THREADS);
    const int P = 2 * NQ * a.n_layers, D = P + 2 * a.n_classes;
    const int DS = (D + 2 + 3) & ~3;
    size_t floats = (size_t)(2 * P + 2 * a.n_layers * (C::A + C::LANES) + 3 * D + (TR_THREADS / 32) * DS +
                             (a.grad_mode == 1 ? a.batch_size * NQ : 0));
    floats = (floats + 1) & ~(size_t)1;                       // float2 alignment of the ring scratch
    TrainArgs args = a;
    args.ring_offset_floats = (int)floats;
    floats += (size_t)2 * (TR_THREADS >> C::LB) * C::DIM;     // one 2^n float2 buffer per circuit slot of the CTA
    const size_t smem = sizeof(float) * floats;
    QFX_ARG_CHECK(smem <= 200 * 1024, QFX_ERR_TOO_LARGE, "training state needs %zu bytes of shared memory", smem);
    if (smem > 48 * 1024)
        QFX_CUDA_TRY(cudaFuncSetAttribute(vqc_local_train_kernel<NQ, LB_>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)cluster_size;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    QFX_CUDA_TRY(cudaLaunchKernelEx(&cfg, vqc_local_train_kernel<NQ, LB_>, args));
    return 0;
}

// lane_mode 0: throughput split (most amplitudes per lane); 1: latency split (most lanes per circuit)
template <int NQ>
struct TrainLaunch {
    static int run(const TrainArgs& a, int n_sel, int cluster_size, int lane_mode, cudaStream_t st) {
        constexpr int LT = default_lane_bits(NQ), LL = latency_lane_bits(NQ);
        if (lane_mode == 1 && LL != LT) return launch_train<NQ, LL>(a, n_sel, cluster_size, st);
        return launch_train<NQ, LT>(a, n_sel, cluster_size, st);
    }
};
template <int NQ>
struct EvalLaunch {
    static int run(int L, int basis, int ent, int nc, const float* angles, const int* labels, const float* params,
                   int N, float* out, ReadoutNoise rn, cudaStream_t st) {
        QFX_CUDA_TRY(cudaMemsetAsync(out, 0, 2 * sizeof(float), st));
        const int slots_cta = TR_THREADS >> SvCfg<NQ>::LB;
        int blocks = (N + slots_cta - 1) / slots_cta;
        if (blocks > 4 * 148) blocks = 4 * 148;
        vqc_eval_kernel<NQ><<<blocks, TR_THREADS, 0, st>>>(L, basis, ent, nc, angles, labels, params, N, out, rn);
        return check_launch("vqc_eval_kernel");
    }
};

template <template <int> class Launcher, typename... Args>
static int dispatch_nq2(int n_qubits, Args&&... args) {
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

}  // namespace qfx

using namespace qfx;

extern "C" int qfx_vqc_local_train_noisy(int n_qubits, int n_layers, int basis, int entangler, int n_classes,
                                   const float* angles, const int* labels, const int* offsets, const int* perm,
                                   int n_total, const float* params_in, const int* client_ids, int n_selected,
                                   float* params_out, float* stats_out, int epochs, int batch_size, int optimizer,
                                   float lr, float beta1, float beta2, float eps, int grad_mode, int cluster_size,
                                   int lane_mode, unsigned long long shuffle_key, float noise_alpha, float noise_beta,
                                   int shots, unsigned long long noise_key, cudaStream_t stream) {
    QFX_ARG_CHECK(n_qubits >= 1 && n_qubits <= 10, QFX_ERR_UNSUPPORTED, "n_qubits %d not in 1..10", n_qubits);
    QFX_ARG_CHECK(n_layers >= 1 && 2 * n_qubits * n_layers <= TR_MAX_P, QFX_ERR_TOO_LARGE, "too many parameters");
    QFX_ARG_CHECK(n_classes >= 1 && n_classes <= n_qubits && n_classes <= TR_MAX_C, QFX_ERR_BAD_ARG,
                  "n_classes %d must be in 1..n_qubits", n_classes);
    QFX_ARG_CHECK(cluster_size >= 1 && cluster_size <= 8, QFX_ERR_BAD_ARG, "cluster_size %d not in 1..8", cluster_size);
    QFX_ARG_CHECK(n_selected >= 1 && epochs >= 1 && batch_size >= 1, QFX_ERR_BAD_ARG, "empty work");
    QFX_ARG_CHECK(grad_mode == 0 || (grad_mode == 1 && batch_size <= TR_MAX_BS_SHIFT), QFX_ERR_TOO_LARGE,
                  "parameter-shift mode supports batch_size <= %d", TR_MAX_BS_SHIFT);
    QFX_ARG_CHECK(optimizer == 0 || optimizer == 1, QFX_ERR_BAD_ARG, "optimizer must be 0 (sgd) or 1 (adam)");
    TrainArgs a;
    a.n_layers = n_layers; a.basis = basis; a.ent = entangler; a.n_classes = n_classes;
    a.angles = angles; a.labels = labels; a.offsets = offsets; a.perm = perm; a.n_total = n_total;
    a.params_in = params_in; a.client_ids = client_ids; a.params_out = params_out; a.stats_out = stats_out;
    a.epochs = epochs; a.batch_size = batch_size; a.opt = optimizer; a.lr = lr; a.beta1 = beta1; a.beta2 = beta2;
    a.eps = eps; a.grad_mode = grad_mode; a.shuffle_key = shuffle_key;
    QFX_ARG_CHECK(shots >= 0, QFX_ERR_BAD_ARG, "shots must be >= 0");
    a.rn = ReadoutNoise{noise_alpha, noise_beta, shots, noise_key};
    a.prof = g_train_prof[current_device()];
    return dispatch_nq2<TrainLaunch>(n_qubits, a, n_selected, cluster_size, lane_mode, stream);
}

// Phase timers of the training kernel: prof = device buffer of n_selected x 12 int64 (zeroed by the caller) that the next
// launches on this device accumulate cycle counts into, or null to switch the timers off (the default).
extern "C" int qfx_vqc_train_profile(long long* prof) {
    g_train_prof[current_device()] = prof;
    return 0;
}

extern "C" int qfx_vqc_local_train(int n_qubits, int n_layers, int basis, int entangler, int n_classes,
                                   const float* angles, const int* labels, const int* offsets, const int* perm,
                                   int n_total, const float* params_in, const int* client_ids, int n_selected,
                                   float* params_out, float* stats_out, int epochs, int batch_size, int optimizer,
                                   float lr, float beta1, float beta2, float eps, int grad_mode, int cluster_size,
                                   int lane_mode, unsigned long long shuffle_key, cudaStream_t stream) {
    return qfx_vqc_local_train_noisy(n_qubits, n_layers, basis, entangler, n_classes, angles, labels, offsets, perm, n_total,
                                     params_in, client_ids, n_selected, params_out, stats_out, epochs, batch_size, optimizer,
                                     lr, beta1, beta2, eps, grad_mode, cluster_size, lane_mode, shuffle_key, 1.f, 0.f, 0, 0ull,
                                     stream);
}

extern "C" int qfx_vqc_eval_noisy(int n_qubits, int n_layers, int basis, int entangler, int n_classes,
                                  const float* angles, const int* labels, const float* params, int n_samples,
                                  float* out2, float noise_alpha, float noise_beta, int shots,
                                  unsigned long long noise_key, cudaStream_t stream) {
    QFX_ARG_CHECK(n_qubits >= 1 && n_qubits <= 10, QFX_ERR_UNSUPPORTED, "n_qubits %d not in 1..10", n_qubits);
    QFX_ARG_CHECK(n_layers >= 0 && 2 * n_qubits * n_layers <= TR_MAX_P, QFX_ERR_TOO_LARGE, "too many parameters");
    QFX_ARG_CHECK(n_classes >= 1 && n_classes <= n_qubits, QFX_ERR_BAD_ARG, "bad n_classes");
    QFX_ARG_CHECK(n_samples > 0, QFX_ERR_BAD_ARG, "empty dataset");
    QFX_ARG_CHECK(shots >= 0, QFX_ERR_BAD_ARG, "shots must be >= 0");
    return dispatch_nq2<EvalLaunch>(n_qubits, n_layers, basis, entangler, n_classes, angles, labels, params,
                                    n_samples, out2, ReadoutNoise{noise_alpha, noise_beta, shots, noise_key}, stream);
}

extern "C" int qfx_vqc_eval(int n_qubits, int n_layers, int basis, int entangler, int n_classes,
                            const float* angles, const int* labels, const float* params, int n_samples,
                            float* out2, cudaStream_t stream) {
    return qfx_vqc_eval_noisy(n_qubits, n_layers, basis, entangler, n_classes, angles, labels, params, n_samples, out2,
                              1.f, 0.f, 0, 0ull, stream);
}

// K3 epilogue as a stand-alone pass for the autograd path: E [B, n] exact <Z> -> noisy readout (in place allowed).
__global__ void __launch_bounds__(256) readout_noise_kernel(const float* __restrict__ E, float* __restrict__ out, int B, int n,
                                                            const qfx::ReadoutNoise rn) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= B * n) return;
    out[i] = qfx::readout_noise(E[i], rn, (unsigned long long)(i / n), i % n);
}
extern "C" int qfx_readout_noise(const float* expvals, float* out, int batch, int n_qubits, float noise_alpha,
                                 float noise_beta, int shots, unsigned long long noise_key, cudaStream_t stream) {
    QFX_ARG_CHECK(expvals && out && batch >= 1 && n_qubits >= 1 && shots >= 0, QFX_ERR_BAD_ARG, "bad arguments");
    readout_noise_kernel<<<(batch * n_qubits + 255) / 256, 256, 0, stream>>>(expvals, out, batch, n_qubits,
                                                                             ReadoutNoise{noise_alpha, noise_beta, shots, noise_key});
    return check_launch("readout_noise_kernel");
}
