// K8: TinyCNN local training batched across the clients resident on one GPU.
//
// The reference trains clients one after another, a few dozen small library kernels per SGD step
// (/root/reference/src/CFed/Classical_FL.py:40-64, model :21-38).  Here every kernel of a step
// covers all clients at once (grid.y = client, each with its own weights), layers are fused
// (conv+bias+ReLU+maxpool; fc2+softmax-CE+backward head; fc1 backward+SGD+unpool) and the whole
// round is replayed from one CUDA graph (qfedx_b200/fl/trainers_cnn.py).  fp32 storage and
// accumulation; direct convolution with register tiling (2x2 outputs x 8 channels per thread) so
// the inner loops are FMA-bound rather than shared-memory-bound.
//
// Flat parameter layout per client (state_dict order, matches models/tinycnn.py):
//   conv1.w [16,1,5,5] @0 | conv1.b [16] @400 | conv2.w [32,16,5,5] @416 | conv2.b [32] @13216 |
//   fc1.w [64,1568] @13248 | fc1.b [64] @113600 | fc2.w [C,64] @113664 | fc2.b [C] @113664+64C
#include <stdlib.h>
#include "tinycnn.cuh"
#include "../../include/qfedx_b200.h"

namespace qfx {
namespace cnn {

// ------------------------------------------------------------------------------------------
// conv1 5x5 (1->16, pad 2) + bias + ReLU + 2x2 maxpool.  One CTA per image, one thread per pooled
// position; the 6x6 input window sits in registers, weights are broadcast from shared memory.
__device__ __forceinline__ void conv1_image(const float* __restrict__ img, const float* __restrict__ prm,
                                            float* __restrict__ a1, unsigned char* __restrict__ i1) {
    __shared__ __align__(16) float xs[32 * 32];
    __shared__ __align__(16) float ws[16 * 25 + 16];
    const int tid = threadIdx.x;
    for (int i = tid; i < 32 * 32; i += blockDim.x) xs[i] = 0.f;
    __syncthreads();
    for (int i = tid; i < 28 * 14; i += blockDim.x) {  // 14 float2 per image row
        const int y = i / 14, x2 = i - y * 14;
        cp_async8(xs + (y + 2) * 32 + 2 + 2 * x2, img + y * 28 + 2 * x2);
    }
    for (int i = tid; i < 104; i += blockDim.x) cp_async16(ws + 4 * i, prm + C1W + 4 * i);
    cp_async_wait_all();
    __syncthreads();
    if (tid >= 196) return;
    const int py = tid / 14, px = tid % 14;
    float win[6][6];
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int c = 0; c < 6; ++c) win[r][c] = xs[(2 * py + r) * 32 + 2 * px + c];
    for (int o = 0; o < 16; ++o) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int ky = 0; ky < 5; ++ky)
#pragma unroll
            for (int kx = 0; kx < 5; ++kx) {
                const float w = ws[o * 25 + ky * 5 + kx];
                acc[0] += w * win[ky][kx];     acc[1] += w * win[ky][kx + 1];
                acc[2] += w * win[ky + 1][kx]; acc[3] += w * win[ky + 1][kx + 1];
            }
        float best = acc[0]; int arg = 0;
#pragma unroll
        for (int j = 1; j < 4; ++j) if (acc[j] > best) { best = acc[j]; arg = j; }
        best = fmaxf(best + ws[400 + o], 0.f);
        a1[o * 196 + tid] = best;
        i1[o * 196 + tid] = (unsigned char)arg;
    }
}

__global__ void __launch_bounds__(224) conv1_fwd_train(const Step s) {
    const int sel = blockIdx.y, b = blockIdx.x;
    int lo;
    if (b >= batch_rows(s, sel, &lo)) return;
    const size_t off = ((size_t)sel * BS + b) * 16 * 196;
    conv1_image(s.X + (size_t)sample_index(s, sel, lo, b) * 784, s.params + (size_t)sel * s.nd, s.a1 + off, s.i1 + off);
}
__global__ void __launch_bounds__(224) conv1_fwd_eval(const float* __restrict__ X, const float* __restrict__ prm,
                                                      float* a1, unsigned char* i1) {
    const size_t off = (size_t)blockIdx.x * 16 * 196;
    conv1_image(X + (size_t)blockIdx.x * 784, prm, a1 + off, i1 + off);
}

// ------------------------------------------------------------------------------------------
// Shared direct-convolution core for conv2 forward (16->32) and conv2 data-gradient (32->16):
// CIN input planes of 14x14 (padded to 18x18 in shared memory), weights in shared memory as
// [cin][tap][cout]; a thread owns a 2x2 output block for CO consecutive output channels.
template <int CIN, int CO>
__device__ __forceinline__ void conv5x5_block(const float* __restrict__ xs /*[CIN][18][18]*/,
                                              const float* __restrict__ ws /*[CIN][25][COUT_ALL]*/, int cout_all,
                                              int co0, int py, int px, float (&acc)[4][CO]) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int o = 0; o < CO; ++o) acc[j][o] = 0.f;
    for (int c = 0; c < CIN; ++c) {
        float win[6][6];
        const float* xp = xs + c * 324 + (2 * py) * 18 + 2 * px;
#pragma unroll
        for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int q = 0; q < 6; ++q) win[r][q] = xp[r * 18 + q];
        const float* wp = ws + (size_t)c * 25 * cout_all + co0;
#pragma unroll
        for (int ky = 0; ky < 5; ++ky)
#pragma unroll
            for (int kx = 0; kx < 5; ++kx) {
                float w[CO];
#pragma unroll
                for (int v = 0; v < CO / 4; ++v) {
                    const float4 t = *reinterpret_cast<const float4*>(wp + (ky * 5 + kx) * cout_all + 4 * v);
                    w[4 * v] = t.x; w[4 * v + 1] = t.y; w[4 * v + 2] = t.z; w[4 * v + 3] = t.w;
                }
#pragma unroll
                for (int o = 0; o < CO; ++o) {
                    acc[0][o] += w[o] * win[ky][kx];     acc[1][o] += w[o] * win[ky][kx + 1];
                    acc[2][o] += w[o] * win[ky + 1][kx]; acc[3][o] += w[o] * win[ky + 1][kx + 1];
                }
            }
    }
}

constexpr int CONV2_SMEM = (16 * 324 + 16 * 25 * 32) * 4;  // 71,936 B
// 256 threads: 4 channel groups x 64 lanes (49 active) so that every warp reads ONE weight address
// (a broadcast) per tap.  w2f is the pre-transposed [c][tap][o] copy kept current by sgd_conv_train.
__device__ __forceinline__ void conv2_image(const float* __restrict__ a1, const float* __restrict__ w2f,
                                            const float* __restrict__ bias, float* __restrict__ a2,
                                            unsigned char* __restrict__ i2, float* smem, int g0) {
    float* xs = smem;                 // [16][18][18]
    float* ws = smem + 16 * 324;      // [16][25][32]
    const int tid = threadIdx.x;
    copy_f4(ws, w2f, 12800);
    fill_padded_planes(xs, a1, 16);
    cp_async_wait_all();
    __syncthreads();
    const int g = g0 + (tid >> 6), pos = tid & 63;
    if (pos >= 49) return;
    const int py = pos / 7, px = pos % 7;
    float acc[4][8];
    conv5x5_block<16, 8>(xs, ws, 32, g * 8, py, px, acc);
#pragma unroll
    for (int o = 0; o < 8; ++o) {
        float best = acc[0][o]; int arg = 0;
#pragma unroll
        for (int j = 1; j < 4; ++j) if (acc[j][o] > best) { best = acc[j][o]; arg = j; }
        best = fmaxf(best + bias[g * 8 + o], 0.f);
        a2[(g * 8 + o) * 49 + pos] = best;
        i2[(g * 8 + o) * 49 + pos] = (unsigned char)arg;
    }
}
__global__ void __launch_bounds__(256) conv2_fwd_train(const Step s) {
    extern __shared__ __align__(16) float smem[];
    const int sel = blockIdx.y, b = blockIdx.x;
    int lo;
    if (b >= batch_rows(s, sel, &lo)) return;
    const size_t r = (size_t)sel * BS + b;
    conv2_image(s.a1 + r * 3136, s.w2f + (size_t)sel * 12800, s.params + (size_t)sel * s.nd + C2B, s.a2 + r * 1568,
                s.i2 + r * 1568, smem, blockIdx.z * (4 / s.split));
}
__global__ void __launch_bounds__(256) conv2_fwd_eval(const float* __restrict__ a1, const float* __restrict__ w2f,
                                                      const float* __restrict__ prm, float* a2, unsigned char* i2) {
    extern __shared__ __align__(16) float smem[];
    const size_t r = blockIdx.x;
    conv2_image(a1 + r * 3136, w2f, prm + C2B, a2 + r * 1568, i2 + r * 1568, smem, 0);
}

// [o][c][tap] -> forward copy [c][tap][o] and data-gradient copy [o][24-tap][c]
__device__ __forceinline__ void transpose_conv2_weight(const float* __restrict__ w, float* __restrict__ w2f,
                                                       float* __restrict__ w2b, int i) {
    const int o = i / 400, rem = i - o * 400, c = rem / 25, tap = rem - c * 25;
    const float v = w[i];
    w2f[rem * 32 + o] = v;
    w2b[(o * 25 + (24 - tap)) * 16 + c] = v;
}
__global__ void __launch_bounds__(256) conv2_weight_prep(const float* __restrict__ params, int nd, float* w2f, float* w2b) {
    const int sel = blockIdx.y, i = blockIdx.x * 256 + threadIdx.x;
    if (i < 12800) transpose_conv2_weight(params + (size_t)sel * nd + C2W, w2f + (size_t)sel * 12800,
                                          w2b + (size_t)sel * 12800, i);
}

// ------------------------------------------------------------------------------------------
// fc1 forward, split-K: CTA (slice, client) computes the partial [32 x 64] product over 98 inputs.
// 128 threads, each a 4x4 (rows x features) register tile.
__global__ void __launch_bounds__(128) fc1_fwd_train(const Step s) {
    __shared__ float xs[BS][FC1_KS + 1];       // [b][k]; odd row stride -> conflict-free column reads
    __shared__ float wsm[64][FC1_KS + 1];      // [f][k]
    const int sel = blockIdx.y, sl = blockIdx.x, tid = threadIdx.x;
    int lo;
    const int nb = batch_rows(s, sel, &lo);
    if (nb <= 0) return;
    const float* __restrict__ x = s.a2 + (size_t)sel * BS * 1568 + sl * FC1_KS;
    const float* __restrict__ w = s.params + (size_t)sel * s.nd + F1W + sl * FC1_KS;
    for (int i = tid; i < BS * FC1_KS; i += 128) {
        const int b = i / FC1_KS, k = i - b * FC1_KS;
        if (b < nb) cp_async4(&xs[b][k], x + (size_t)b * 1568 + k); else xs[b][k] = 0.f;
    }
    for (int i = tid; i < 64 * FC1_KS; i += 128) {
        const int f = i / FC1_KS, k = i - f * FC1_KS;
        cp_async4(&wsm[f][k], w + (size_t)f * 1568 + k);
    }
    cp_async_wait_all();
    __syncthreads();
    // 4 rows x 4 features per thread; features strided by 16 so a warp reads 16 consecutive rows of wsm
    const int bt = (tid >> 4) * 4, f0 = tid & 15;
    float acc[4][4] = {};
#pragma unroll 4
    for (int k = 0; k < FC1_KS; ++k) {
        float xa[4], wa[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { xa[i] = xs[bt + i][k]; wa[i] = wsm[f0 + 16 * i][k]; }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] += xa[i] * wa[j];
    }
    float* out = s.fc1_part + (((size_t)sel * FC1_SLICES + sl) * BS) * 64;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) out[(bt + i) * 64 + f0 + 16 * j] = acc[i][j];
}

// ------------------------------------------------------------------------------------------
// Head: fc1 bias + ReLU + dropout, fc2, softmax cross-entropy, and everything of the backward
// pass that lives above fc1: d logits, fc2 / fc1-bias SGD updates, dh1.  One CTA per client.
struct HeadSmem {
    float h[BS][65];
    float w2[MAX_CLASSES][64];
    float z[BS][MAX_CLASSES];
    float dz[BS][MAX_CLASSES];
    float red[2];
};
// Runs inside every fc1_bwd CTA of a client (redundantly - the work is tiny and it removes a kernel whose
// single CTA per client left the GPU idle): fills dh[b][f] in shared memory.  The owner CTA (chunk 0)
// also accumulates the epoch statistics and stores the fc2 / fc1-bias gradients for sgd_conv_train.
__device__ __forceinline__ void head_compute(const Step& s, int sel, int nb, int lo, HeadSmem& hs, float (*dh)[64],
                                             bool owner) {
    const int tid = threadIdx.x, NC = s.n_classes;
    const float* __restrict__ prm = s.params + (size_t)sel * s.nd;
    const int f2b = F2W + 64 * NC;
    if (tid < 2) hs.red[tid] = 0.f;
    for (int i = tid; i < NC * 64; i += 256) hs.w2[i / 64][i % 64] = prm[F2W + i];
    // h = dropout(relu(sum of split-K partials + b1))
    const int rnd = s.round_dev ? *s.round_dev : s.round;
    const unsigned long long dkey = key3(s.seed ^ 0x44524f50ull, (unsigned long long)rnd * 4096ull + s.client_ids[sel],
                                         (unsigned long long)s.epoch * 65536ull + s.batch_index);
    const float keep_scale = s.p_drop > 0.f ? 1.f / (1.f - s.p_drop) : 1.f;
    // 8 outputs per thread, 14 split-K partials each: the loads of four outputs (56) are issued before any is
    // used - the sum otherwise waits for L2 eight times in a row
    constexpr int HG = 4;
#pragma unroll
    for (int i0 = 0; i0 < BS * 64; i0 += 256 * HG) {
        float pv[HG][FC1_SLICES];
#pragma unroll
        for (int g = 0; g < HG; ++g) {
            const int i = i0 + g * 256 + tid, b = i >> 6, f = i & 63;
            const float* __restrict__ part = s.fc1_part + ((size_t)sel * FC1_SLICES * BS + b) * 64 + f;
#pragma unroll
            for (int sl = 0; sl < FC1_SLICES; ++sl) pv[g][sl] = b < nb ? __ldg(part + (size_t)sl * BS * 64) : 0.f;
        }
#pragma unroll
        for (int g = 0; g < HG; ++g) {
            const int i = i0 + g * 256 + tid, b = i >> 6, f = i & 63;
            float v = 0.f;
            if (b < nb) {
                v = prm[F1B + f];
#pragma unroll
                for (int sl = 0; sl < FC1_SLICES; ++sl) v += pv[g][sl];
                v = fmaxf(v, 0.f);
                if (s.p_drop > 0.f) {
                    const float u = uniform_from_bits(stream_at(dkey, (unsigned long long)i));
                    v = u < s.p_drop ? 0.f : v * keep_scale;
                }
            }
            hs.h[b][f] = v;
        }
    }
    __syncthreads();
    for (int i = tid; i < BS * NC; i += 256) {   // logits
        const int b = i / NC, c = i % NC;
        float v = prm[f2b + c];
#pragma unroll 8
        for (int f = 0; f < 64; ++f) v += hs.h[b][f] * hs.w2[c][f];
        hs.z[b][c] = v;
    }
    __syncthreads();
    if (tid < BS) {
        const int b = tid;
        if (b < nb) {
            const int y = s.labels[sample_index(s, sel, lo, b)];
            float m = hs.z[b][0]; int arg = 0;
            for (int c = 1; c < NC; ++c) if (hs.z[b][c] > m) { m = hs.z[b][c]; arg = c; }
            float sum = 0.f;
            for (int c = 0; c < NC; ++c) sum += expf(hs.z[b][c] - m);
            const float inv = 1.f / sum, inv_nb = 1.f / (float)nb;
            for (int c = 0; c < NC; ++c) hs.dz[b][c] = (expf(hs.z[b][c] - m) * inv - (c == y ? 1.f : 0.f)) * inv_nb;
            if (owner) {
                atomicAdd(&hs.red[0], m + logf(sum) - hs.z[b][y]);
                atomicAdd(&hs.red[1], arg == y ? 1.f : 0.f);
            }
        } else {
            for (int c = 0; c < NC; ++c) hs.dz[b][c] = 0.f;
        }
    }
    __syncthreads();
    // dh = (dz W2) masked by ReLU/dropout (h > 0 <=> kept and active)
    for (int i = tid; i < BS * 64; i += 256) {
        const int b = i >> 6, f = i & 63;
        float v = 0.f;
        if (hs.h[b][f] > 0.f) {
            for (int c = 0; c < NC; ++c) v += hs.dz[b][c] * hs.w2[c][f];
            v *= keep_scale;
        }
        dh[b][f] = v;
    }
    if (owner) {
        if (tid == 0) { s.stats[sel * 2] += hs.red[0]; s.stats[sel * 2 + 1] += hs.red[1]; }
        // gradients of fc2 weights, fc2 bias, fc1 bias: applied by sgd_conv_train at the end of the step
        float* hg = s.head_grad + (size_t)sel * HEAD_GRAD_STRIDE;
        for (int i = tid; i < NC * 64 + NC + 64; i += 256) {
            float g = 0.f;
            if (i < NC * 64) {
                const int c = i / 64, f = i % 64;
                for (int b = 0; b < nb; ++b) g += hs.dz[b][c] * hs.h[b][f];
            } else if (i < NC * 64 + NC) {
                const int c = i - NC * 64;
                for (int b = 0; b < nb; ++b) g += hs.dz[b][c];
            } else {
                const int f = i - NC * 64 - NC;
                for (int b = 0; b < nb; ++b) {
                    if (hs.h[b][f] > 0.f) {
                        float v = 0.f;
                        for (int c = 0; c < NC; ++c) v += hs.dz[b][c] * hs.w2[c][f];
                        g += v * keep_scale;
                    }
                }
            }
            hg[i] = g;
        }
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------
// fc1 backward for a 32-column chunk of the 1568 inputs: dW1 (+ SGD update in place), dx, and dx
// pushed through maxpool2 / ReLU into the dense conv2-output gradient g2.
__global__ void __launch_bounds__(256) fc1_bwd_train(const Step s) {
    __shared__ __align__(16) float dh[BS][64];
    __shared__ float xs[BS][33];
    __shared__ float wsm[64][33];
    __shared__ HeadSmem hs;
    const int sel = blockIdx.y, k0 = blockIdx.x * 32, tid = threadIdx.x;
    int lo;
    const int nb = batch_rows(s, sel, &lo);
    if (nb <= 0) return;
    float* prm = s.params + (size_t)sel * s.nd;
    float* mom = s.moment + (size_t)sel * s.nd;
    for (int i = tid; i < BS * 32; i += 256) {
        const int b = i >> 5, k = i & 31;
        if (b < nb) cp_async4(&xs[b][k], s.a2 + ((size_t)sel * BS + b) * 1568 + k0 + k); else xs[b][k] = 0.f;
    }
    for (int i = tid; i < 64 * 32; i += 256) cp_async4(&wsm[i >> 5][i & 31], prm + F1W + (size_t)(i >> 5) * 1568 + k0 + (i & 31));
    // momentum of this thread's eight weights: fetched now, needed only by the update at the very end
    float mprev[8];
    {
        const int k = tid & 31, fg = (tid >> 5) * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) mprev[j] = mom[F1W + (size_t)(fg + j) * 1568 + k0 + k];
    }
    head_compute(s, sel, nb, lo, hs, dh, blockIdx.x == 0);   // the async copies above land meanwhile
    cp_async_wait_all();
    __syncthreads();
    // dx[b][k] = sum_f dh[b][f] W1[f][k]; 1024 outputs, 4 per thread (b = tid/8 .., k = ..)
    {
        const int k = tid & 31, bq = tid >> 5;  // rows bq, bq+8, bq+16, bq+24
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 8
        for (int f = 0; f < 64; ++f) {
            const float w = wsm[f][k];
#pragma unroll
            for (int r = 0; r < 4; ++r) acc[r] += dh[bq + 8 * r][f] * w;
        }
        const int col = k0 + k, oc = col / 49, pos = col % 49, py = pos / 7, px = pos % 7;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int b = bq + 8 * r;
            if (b >= nb) continue;
            const size_t row = (size_t)sel * BS + b;
            const float act = s.a2[row * 1568 + col];
            const int arg = s.i2[row * 1568 + col];
            const float gv = act > 0.f ? acc[r] : 0.f;
            float* g = s.g2 + (row * 32 + oc) * 196 + (2 * py) * 14 + 2 * px;
            g[0] = arg == 0 ? gv : 0.f;  g[1] = arg == 1 ? gv : 0.f;
            g[14] = arg == 2 ? gv : 0.f; g[15] = arg == 3 ? gv : 0.f;
        }
    }
    // dW1[f][k] = sum_b dh[b][f] x[b][k]; 2048 outputs, 8 per thread; SGD-momentum in place
    {
        const int k = tid & 31, fg = (tid >> 5) * 8;
        float acc[8] = {};
        for (int b = 0; b < nb; ++b) {
            const float xv = xs[b][k];
            const float4 d0 = *reinterpret_cast<const float4*>(&dh[b][fg]);
            const float4 d1 = *reinterpret_cast<const float4*>(&dh[b][fg + 4]);
            acc[0] += d0.x * xv; acc[1] += d0.y * xv; acc[2] += d0.z * xv; acc[3] += d0.w * xv;
            acc[4] += d1.x * xv; acc[5] += d1.y * xv; acc[6] += d1.z * xv; acc[7] += d1.w * xv;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const size_t p = F1W + (size_t)(fg + j) * 1568 + k0 + k;
            const float m = s.momentum * mprev[j] + acc[j];
            mom[p] = m;
            prm[p] = wsm[fg + j][k] - s.lr * m;
        }
    }
}

// ------------------------------------------------------------------------------------------
// conv2 data gradient: da1[c][y][x] = sum_o sum_taps g2pad[o][y+ky][x+kx] W[o][c][4-ky][4-kx].
constexpr int DGRAD_SMEM = (32 * 324 + 32 * 25 * 16) * 4;  // 92,672 B
__global__ void __launch_bounds__(256) conv2_dgrad_train(const Step s) {
    extern __shared__ __align__(16) float smem[];
    float* xs = smem;                // [32][18][18]
    float* ws = smem + 32 * 324;     // [32 o][25 taps][16 c], taps already flipped (w2b)
    const int sel = blockIdx.y, b = blockIdx.x, tid = threadIdx.x;
    int lo;
    if (b >= batch_rows(s, sel, &lo)) return;
    const size_t row = (size_t)sel * BS + b;
    copy_f4(ws, s.w2b + (size_t)sel * 12800, 12800);
    fill_padded_planes(xs, s.g2 + row * 32 * 196, 32);
    cp_async_wait_all();
    __syncthreads();
    const int g = blockIdx.z * (4 / s.split) + (tid >> 6), pos = tid & 63;
    if (pos >= 49) return;
    const int py = pos / 7, px = pos % 7;
    float acc[4][4];
    conv5x5_block<32, 4>(xs, ws, 16, g * 4, py, px, acc);
    float* out = s.da1 + row * 3136;
#pragma unroll
    for (int o = 0; o < 4; ++o) {
        float* p = out + (g * 4 + o) * 196 + (2 * py) * 14 + 2 * px;
        p[0] = acc[0][o]; p[1] = acc[1][o]; p[14] = acc[2][o]; p[15] = acc[3][o];
    }
}

// ------------------------------------------------------------------------------------------
// conv2 weight gradient for one image (split-K over the batch): thread (og, c, ky) accumulates
// dW[og*8..+8][c][ky][0..5) over all 14x14 positions; partial sums go to conv_part[image].
constexpr int WGRAD_SMEM = (196 * 32 + 16 * 324) * 4;  // 45,824 B
__global__ void __launch_bounds__(320) conv2_wgrad_train(const Step s) {
    extern __shared__ __align__(16) float smem[];
    float* gs = smem;               // [o][y][x] 32*196 (planar, straight copy)
    float* xs = smem + 196 * 32;    // [16][18][18]
    const int sel = blockIdx.y, b = blockIdx.x, tid = threadIdx.x;
    int lo;
    const int nb = batch_rows(s, sel, &lo);
    float* part = s.conv_part + ((size_t)sel * BS + b) * NCONV;
    if (b >= nb) {  // keep the reduction in sgd_conv branch-free: absent images contribute zeros
        if (nb > 0 && blockIdx.z == 0) for (int i = tid; i < NCONV; i += blockDim.x) part[i] = 0.f;
        return;
    }
    const size_t row = (size_t)sel * BS + b;
    copy_f4(gs, s.g2 + row * 32 * 196, 32 * 196);
    fill_padded_planes(xs, s.a1 + row * 3136, 16);
    cp_async_wait_all();
    __syncthreads();
    const int ky = tid % 5, c = (tid / 5) % 16, og = blockIdx.z * (4 / s.split) + tid / 80;  // og in 0..3
    float acc[8][5] = {};
    for (int y = 0; y < 14; ++y) {
        float xr[18];
        const float* xp = xs + c * 324 + (y + ky) * 18;
#pragma unroll
        for (int q = 0; q < 18; ++q) xr[q] = xp[q];
#pragma unroll
        for (int x = 0; x < 14; ++x) {
            float gv[8];  // same address for every lane of a warp: broadcast reads
#pragma unroll
            for (int o = 0; o < 8; ++o) gv[o] = gs[(og * 8 + o) * 196 + y * 14 + x];
#pragma unroll
            for (int o = 0; o < 8; ++o)
#pragma unroll
                for (int kx = 0; kx < 5; ++kx) acc[o][kx] += gv[o] * xr[x + kx];
        }
    }
#pragma unroll
    for (int o = 0; o < 8; ++o)
#pragma unroll
        for (int kx = 0; kx < 5; ++kx) part[C2W + (og * 8 + o) * 400 + c * 25 + ky * 5 + kx] = acc[o][kx];
    // conv2 bias gradient: sum of g2 over positions, one thread per output channel
    if (blockIdx.z == 0 && tid < 32) {
        float sum = 0.f;
        for (int i = 0; i < 196; ++i) sum += gs[tid * 196 + i];
        part[C2B + tid] = sum;
    }
}

// ------------------------------------------------------------------------------------------
// conv1 weight gradient for one image.  The gradient reaching conv1's output is sparse (one
// position per pooled window, and only where the pooled activation is positive), so the sum runs
// over the 16x196 pooled elements: thread (o, tap) adds da1 * x[argmax position + tap].
constexpr int XS1 = 37;  // row stride: bank = 5*ky + kx is distinct for the 25 taps of a warp
__global__ void __launch_bounds__(416) conv1_wgrad_train(const Step s) {
    __shared__ float xs[32 * XS1];
    __shared__ float gsm[16 * 196];
    __shared__ unsigned short osm[16 * 196];   // offset of the arg-max pixel of every pooled element in xs
    const int sel = blockIdx.y, b = blockIdx.x, tid = threadIdx.x;
    int lo;
    const int nb = batch_rows(s, sel, &lo);
    if (b >= nb) {  // absent image: its row of conv1 partials must read as zero in sgd_conv_train
        if (nb > 0) s.conv_part[((size_t)sel * BS + b) * NCONV + C1W + tid] = 0.f;   // 416 threads = 400 weights + 16 biases
        return;
    }
    const size_t row = (size_t)sel * BS + b;
    const float* img = s.X + (size_t)sample_index(s, sel, lo, b) * 784;
    for (int i = tid; i < 32 * XS1; i += 416) xs[i] = 0.f;
    __syncthreads();
    for (int i = tid; i < 784; i += 416) {
        const int y = i / 28, x = i - y * 28;
        cp_async4(xs + (y + 2) * XS1 + x + 2, img + i);
    }
    {
        const float4* __restrict__ a4 = reinterpret_cast<const float4*>(s.a1 + row * 3136);
        const float4* __restrict__ d4 = reinterpret_cast<const float4*>(s.da1 + row * 3136);
        const uchar4* __restrict__ i4 = reinterpret_cast<const uchar4*>(s.i1 + row * 3136);
        float4 av[2], dv[2]; uchar4 iv[2];
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            const int i = tid + r * 416;
            if (i < 784) { av[r] = __ldg(a4 + i); dv[r] = __ldg(d4 + i); iv[r] = __ldg(i4 + i); }
        }
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            const int i = tid + r * 416;
            if (i < 784) {
                gsm[4 * i] = av[r].x > 0.f ? dv[r].x : 0.f;     gsm[4 * i + 1] = av[r].y > 0.f ? dv[r].y : 0.f;
                gsm[4 * i + 2] = av[r].z > 0.f ? dv[r].z : 0.f; gsm[4 * i + 3] = av[r].w > 0.f ? dv[r].w : 0.f;
                const unsigned char args[4] = {iv[r].x, iv[r].y, iv[r].z, iv[r].w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int p = (4 * i + j) % 196, arg = args[j];
                    osm[4 * i + j] = (unsigned short)((2 * (p / 14) + (arg >> 1)) * XS1 + 2 * (p % 14) + (arg & 1));
                }
            }
        }
    }
    cp_async_wait_all();
    __syncthreads();
    float* part = s.conv_part + row * NCONV;
    if (tid < 400) {
        const int o = tid / 25, ky = (tid % 25) / 5, kx = tid % 5;
        float acc = 0.f;
        const float* xt = xs + ky * XS1 + kx;
#pragma unroll 4
        for (int p = 0; p < 196; ++p) acc += gsm[o * 196 + p] * xt[osm[o * 196 + p]];
        part[C1W + tid] = acc;
    } else {
        const int o = tid - 400;
        float acc = 0.f;
        for (int p = 0; p < 196; ++p) acc += gsm[o * 196 + p];
        part[C1B + o] = acc;
    }
}

// reduce the per-image partials and apply SGD-momentum to the 13,248 conv parameters; the same kernel applies
// the fc2 / fc1-bias gradients that the head left in head_grad
__global__ void __launch_bounds__(256) sgd_conv_train(const Step s) {
    const int sel = blockIdx.y, p0 = blockIdx.x * 256 + threadIdx.x;
    int lo;
    if (batch_rows(s, sel, &lo) <= 0) return;
    float* prm = s.params + (size_t)sel * s.nd;
    float* mom = s.moment + (size_t)sel * s.nd;
    const int NC = s.n_classes;
    if (p0 >= NCONV) {
        const int i = p0 - NCONV;
        if (i >= NC * 64 + NC + 64 || s.sgd_part == 2) return;
        const int p = i < NC * 64 + NC ? F2W + i : F1B + (i - NC * 64 - NC);
        const float m = s.momentum * mom[p] + s.head_grad[(size_t)sel * HEAD_GRAD_STRIDE + i];
        mom[p] = m;
        prm[p] -= s.lr * m;
        return;
    }
    const int p = p0;
    if ((s.sgd_part == 1 && p >= C2W) || (s.sgd_part == 2 && p < C2W)) return;
    const float* part = s.conv_part + (size_t)sel * BS * NCONV + p;
    float g = 0.f;
    // conv1 partials always have one row per image; conv2 partials have BS / wgrad_ipc rows
    const int rows = (p >= C2W && s.conv2_mode == 1) ? BS / s.wgrad_ipc : BS;
#pragma unroll 8
    for (int b = 0; b < rows; ++b) g += part[(size_t)b * NCONV];
    const float m = s.momentum * mom[p] + g;
    mom[p] = m;
    const float w = prm[p] - s.lr * m;
    prm[p] = w;
    if (p >= C2W && p < C2B) {  // keep the operand copies of the conv2 weights current
        const int i = p - C2W, o = i / 400, rem = i - o * 400, c = rem / 25, tap = rem - c * 25;
        s.w2f[(size_t)sel * 12800 + rem * 32 + o] = w;
        s.w2b[(size_t)sel * 12800 + (o * 25 + (24 - tap)) * 16 + c] = w;
        if (s.conv2_mode == 1) store_w2u(s.w2uf + (size_t)sel * 12800, s.w2ub + (size_t)sel * 12800, o, c, tap, w);
    }
}

// ------------------------------------------------------------------------------------------
// Evaluation head: fc1 + ReLU + fc2 + loss/argmax for 4 images per CTA (no dropout).
constexpr int EVB = 4;
__global__ void __launch_bounds__(256) head_eval(const float* __restrict__ a2, const int* __restrict__ labels,
                                                 const float* __restrict__ prm, int n_classes, int N,
                                                 float* __restrict__ out2) {
    __shared__ float xs[EVB][1568];
    __shared__ float h[EVB][64];
    __shared__ float z[EVB][MAX_CLASSES];
    const int tid = threadIdx.x, i0 = blockIdx.x * EVB;
    for (int i = tid; i < EVB * 1568; i += 256) {
        const int b = i / 1568;
        xs[b][i % 1568] = (i0 + b) < N ? a2[(size_t)(i0 + b) * 1568 + i % 1568] : 0.f;
    }
    __syncthreads();
    // 512 hidden outputs, 2 per thread: warp w handles feature pairs; coalesced W1 reads along k
    const int warp = tid >> 5, lane = tid & 31;
    for (int f = warp; f < 64; f += 8) {
        float acc[EVB] = {};
        const float* w = prm + F1W + (size_t)f * 1568;
        for (int k = lane; k < 1568; k += 32) {
            const float wv = w[k];
#pragma unroll
            for (int b = 0; b < EVB; ++b) acc[b] += wv * xs[b][k];
        }
#pragma unroll
        for (int b = 0; b < EVB; ++b) {
            const float v = warp_sum(acc[b]);
            if (lane == 0) h[b][f] = fmaxf(v + prm[F1B + f], 0.f);
        }
    }
    __syncthreads();
    if (tid < EVB * n_classes) {
        const int b = tid / n_classes, c = tid % n_classes;
        float v = prm[F2W + 64 * n_classes + c];
        for (int f = 0; f < 64; ++f) v += h[b][f] * prm[F2W + c * 64 + f];
        z[b][c] = v;
    }
    __syncthreads();
    if (tid < EVB && i0 + tid < N) {
        const int b = tid, y = labels[i0 + b];
        float m = z[b][0]; int arg = 0;
        for (int c = 1; c < n_classes; ++c) if (z[b][c] > m) { m = z[b][c]; arg = c; }
        float sum = 0.f;
        for (int c = 0; c < n_classes; ++c) sum += expf(z[b][c] - m);
        atomicAdd(out2, m + logf(sum) - z[b][y]);
        atomicAdd(out2 + 1, arg == y ? 1.f : 0.f);
    }
}

int launch_head_eval(const float* a2, const int* labels, const float* prm, int n_classes, int n, float* out2,
                     cudaStream_t stream) {
    head_eval<<<(n + EVB - 1) / EVB, 256, 0, stream>>>(a2, labels, prm, n_classes, n, out2);
    return check_launch("head_eval");
}

// Per-device host state: the second stream + events for the fork/join inside a step (works eagerly and under
// stream capture, where the cross-stream waits become graph dependencies), the pipelining bookkeeping of the
// overlapped steps and the "attributes set" flags.  Indexed by the current device so that one process can drive
// several GPUs.
struct DeviceState {
    cudaStream_t side_stream = nullptr;
    cudaEvent_t fork_event = nullptr, join_event = nullptr, dgrad_event = nullptr;
    bool side_pending = false;   // the side stream still owes the conv2 update of the previous step
    int a1_parity = 0;
    bool attr_set = false;
};
static DeviceState g_dev[QFX_MAX_DEVICES];
static int device_state(DeviceState** out) {
    int dev = 0;
    QFX_CUDA_TRY(cudaGetDevice(&dev));
    QFX_ARG_CHECK(dev >= 0 && dev < QFX_MAX_DEVICES, QFX_ERR_UNSUPPORTED, "device index %d beyond %d", dev, QFX_MAX_DEVICES);
    *out = &g_dev[dev];
    return 0;
}
static int side_stream_init(DeviceState& d) {
    if (d.side_stream) return 0;
    QFX_CUDA_TRY(cudaStreamCreateWithFlags(&d.side_stream, cudaStreamNonBlocking));
    QFX_CUDA_TRY(cudaEventCreateWithFlags(&d.fork_event, cudaEventDisableTiming));
    QFX_CUDA_TRY(cudaEventCreateWithFlags(&d.join_event, cudaEventDisableTiming));
    QFX_CUDA_TRY(cudaEventCreateWithFlags(&d.dgrad_event, cudaEventDisableTiming));
    return 0;
}

static int set_attrs(DeviceState& d) {
    if (d.attr_set) return 0;
    QFX_CUDA_TRY(cudaFuncSetAttribute(conv2_fwd_train, cudaFuncAttributeMaxDynamicSharedMemorySize, CONV2_SMEM));
    QFX_CUDA_TRY(cudaFuncSetAttribute(conv2_fwd_eval, cudaFuncAttributeMaxDynamicSharedMemorySize, CONV2_SMEM));
    QFX_CUDA_TRY(cudaFuncSetAttribute(conv2_dgrad_train, cudaFuncAttributeMaxDynamicSharedMemorySize, DGRAD_SMEM));
    QFX_CUDA_TRY(cudaFuncSetAttribute(conv2_wgrad_train, cudaFuncAttributeMaxDynamicSharedMemorySize, WGRAD_SMEM));
    d.attr_set = true;
    return 0;
}

}  // namespace cnn
}  // namespace qfx

using namespace qfx;
using namespace qfx::cnn;

extern "C" int qfx_tinycnn_workspace_bytes(int n_selected, size_t* bytes) {
    // a1,i1,a2,i2,fc1_part,g2,da1,conv_part,operand copies,head_grad,a1_alt,stats  (256-byte aligned pieces)
    const size_t K = (size_t)n_selected;
    size_t tot = 0;
    auto add = [&](size_t n) { tot += (n + 255) & ~(size_t)255; };
    add(K * BS * 3136 * 4); add(K * BS * 3136); add(K * BS * 1568 * 4); add(K * BS * 1568);
    add(K * FC1_SLICES * BS * 64 * 4);
    add(K * BS * 32 * 196 * 4); add(K * BS * 3136 * 4); add(K * BS * NCONV * 4); add(K * 12800 * 4); add(K * 12800 * 4); add(K * 12800 * 2); add(K * 12800 * 2); add(K * HEAD_GRAD_STRIDE * 4);
    add(K * BS * 3136 * 4);   // a1_alt
    add(K * 2 * 4);
    *bytes = tot;
    return 0;
}

static void carve(Step& s, void* workspace, size_t K) {
    char* w = (char*)workspace;
    auto take = [&](size_t n) { char* p = w; w += (n + 255) & ~(size_t)255; return p; };
    s.a1 = (float*)take(K * BS * 3136 * 4); s.i1 = (unsigned char*)take(K * BS * 3136);
    s.a2 = (float*)take(K * BS * 1568 * 4); s.i2 = (unsigned char*)take(K * BS * 1568);
    s.fc1_part = (float*)take(K * FC1_SLICES * BS * 64 * 4);
    s.g2 = (float*)take(K * BS * 32 * 196 * 4); s.da1 = (float*)take(K * BS * 3136 * 4);
    s.conv_part = (float*)take(K * BS * NCONV * 4);
    s.w2f = (float*)take(K * 12800 * 4); s.w2b = (float*)take(K * 12800 * 4);
    s.w2uf = (__nv_bfloat16*)take(K * 12800 * 2); s.w2ub = (__nv_bfloat16*)take(K * 12800 * 2);
    s.head_grad = (float*)take(K * HEAD_GRAD_STRIDE * 4);
    s.a1_alt = (float*)take(K * BS * 3136 * 4);
    s.stats = (float*)take(K * 2 * 4);
    s.sgd_part = 0;
}

// Call once per round after the global parameters were broadcast into `params`: builds the
// operand-layout copies of the conv2 weights that the step kernels keep current afterwards.
extern "C" int qfx_tinycnn_begin_round(const float* params, int nd, int n_selected, void* workspace,
                                       cudaStream_t stream) {
    QFX_ARG_CHECK(n_selected >= 1 && params && workspace, QFX_ERR_BAD_ARG, "bad arguments");
    Step s;
    carve(s, workspace, (size_t)n_selected);
    conv2_weight_prep<<<dim3(50, n_selected), 256, 0, stream>>>(params, nd, s.w2f, s.w2b);
    if (int e = check_launch("conv2_weight_prep")) return e;
    return launch_conv2_umma_prep(params, nd, n_selected, s.w2uf, s.w2ub, stream);
}

// One SGD step for every selected client: 8 kernels on `stream` (capture them in a CUDA graph).
// `workspace` comes from qfx_tinycnn_workspace_bytes; stats (last 2K floats of it) accumulate the
// epoch's loss sum and correct count - the caller zeroes them per epoch.
// Overlapped steps: the side stream still owes the conv2 update of the previous step; the owed work is joined in
// front of the next conv2 forward (or by qfx_tinycnn_finish_steps).  a1 alternates between two buffers.
static int train_step_impl(const float* X, const int* labels, const int* offsets, const int* perm,
                           int n_total, const int* client_ids, int n_selected, int epoch,
                           int batch_index, int batch_size, float* params, float* moment, int nd,
                           int n_classes, float lr, float momentum, float p_drop,
                           unsigned long long seed, int round, const int* round_dev, void* workspace,
                           int conv2_mode, unsigned long long shuffle_key, cudaStream_t stream, bool overlap) {
    QFX_ARG_CHECK(conv2_mode == 0 || conv2_mode == 1, QFX_ERR_BAD_ARG, "conv2_mode must be 0 (fp32) or 1 (tcgen05 bf16)");
    QFX_ARG_CHECK(batch_size >= 1 && batch_size <= BS, QFX_ERR_UNSUPPORTED,
                  "client-batched TinyCNN supports batch_size <= %d, got %d", BS, batch_size);
    QFX_ARG_CHECK(n_classes >= 2 && n_classes <= MAX_CLASSES, QFX_ERR_UNSUPPORTED, "n_classes %d not in 2..10", n_classes);
    QFX_ARG_CHECK(nd >= F2W + 65 * n_classes, QFX_ERR_BAD_ARG, "parameter row stride too small");
    QFX_ARG_CHECK(n_selected >= 1 && p_drop >= 0.f && p_drop < 1.f, QFX_ERR_BAD_ARG, "bad n_selected / p_drop");
    DeviceState* dsp = nullptr;
    if (int e = device_state(&dsp)) return e;
    DeviceState& d = *dsp;
    if (int e = set_attrs(d)) return e;
    Step s;
    s.X = X; s.labels = labels; s.offsets = offsets; s.perm = perm; s.client_ids = client_ids;
    s.n_total = n_total; s.epoch = epoch; s.batch_index = batch_index; s.batch_size = batch_size;
    s.params = params; s.moment = moment; s.nd = nd; s.n_classes = n_classes;
    s.lr = lr; s.momentum = momentum; s.p_drop = p_drop; s.seed = seed; s.round = round; s.round_dev = round_dev; s.conv2_mode = conv2_mode; s.shuffle_key = shuffle_key;
    // many clients: two images per weight-gradient CTA (one wave of CTAs, half the partial sums)
    s.wgrad_ipc = (conv2_mode == 1 && n_selected >= 5) ? 2 : 1;
    carve(s, workspace, (size_t)n_selected);
    const dim3 gimg(batch_size, n_selected);
    // few clients -> few CTAs: spread the channel groups of the conv2 kernels over up to 4 CTAs per image
    s.split = n_selected >= 4 ? 1 : (n_selected >= 2 ? 2 : 4);
    const dim3 gsplit(batch_size, n_selected, s.split);
    if (int e = side_stream_init(d)) return e;
    cudaStream_t side = d.side_stream;
#ifdef QFX_PROFILE_SKIP
    // profiling builds only (results are wrong when set; never compiled into the shipped library): QFX_CNN_SKIP is a
    // bit mask of kernels to leave out, to find the step's critical path by elimination - 1 conv1 fwd, 2 conv2 fwd,
    // 4 fc1 fwd, 8 fc1 bwd, 16 conv2 wgrad, 32 conv2 dgrad, 64 conv1 wgrad, 128 updates
    static const int skip = getenv("QFX_CNN_SKIP") ? atoi(getenv("QFX_CNN_SKIP")) : 0;
#else
    constexpr int skip = 0;
#endif
    if (overlap) {
        d.a1_parity ^= 1;
        if (d.a1_parity) { float* t = s.a1; s.a1 = s.a1_alt; s.a1_alt = t; }
    }
    if (!(skip & 1)) conv1_fwd_train<<<gimg, 224, 0, stream>>>(s);   // overlapped mode: beside the previous step's conv2 wgrad + update
    if (d.side_pending) {
        QFX_CUDA_TRY(cudaStreamWaitEvent(stream, d.join_event, 0));
        d.side_pending = false;
    }
    if (skip & 2) {}
    else if (conv2_mode == 1) { if (int e = launch_conv2_umma_fwd(s, n_selected, stream)) return e; }
    else conv2_fwd_train<<<gsplit, 256 / s.split, CONV2_SMEM, stream>>>(s);
    if (!(skip & 4)) fc1_fwd_train<<<dim3(FC1_SLICES, n_selected), 128, 0, stream>>>(s);
    if (!(skip & 8)) fc1_bwd_train<<<dim3(49, n_selected), 256, 0, stream>>>(s);
    // fork: the conv2 weight gradient only needs g2 and a1; it runs beside conv2 dgrad -> conv1 wgrad
    QFX_CUDA_TRY(cudaEventRecord(d.fork_event, stream));
    QFX_CUDA_TRY(cudaStreamWaitEvent(side, d.fork_event, 0));
    if (conv2_mode == 1) {
        if (!(skip & 16)) { if (int e = launch_conv2_umma_wgrad(s, n_selected, side)) return e; }
        if (!(skip & 32)) { if (int e = launch_conv2_umma_dgrad(s, n_selected, stream)) return e; }
    } else {
        conv2_wgrad_train<<<dim3(BS, n_selected, s.split), 320 / s.split, WGRAD_SMEM, side>>>(s);
        conv2_dgrad_train<<<gsplit, 256 / s.split, DGRAD_SMEM, stream>>>(s);
    }
    if (overlap) QFX_CUDA_TRY(cudaEventRecord(d.dgrad_event, stream));   // the data gradient has read the conv2 weights
    if (!(skip & 64)) conv1_wgrad_train<<<gimg, 416, 0, stream>>>(s);
    const dim3 gsgd((NCONV + 64 + 65 * n_classes + 255) / 256, n_selected);
    if (!overlap) {
        QFX_CUDA_TRY(cudaEventRecord(d.join_event, side));
        QFX_CUDA_TRY(cudaStreamWaitEvent(stream, d.join_event, 0));
        sgd_conv_train<<<gsgd, 256, 0, stream>>>(s);
    } else {
        // the conv2 parameters are updated on the side stream right behind their gradient; this stream only
        // updates conv1 + head and goes on to the next step's conv1 forward.  The update rewrites the operand
        // copies (w2b / w2ub) that this step's data gradient reads on the main stream, so it waits for it.
        Step sb = s;
        sb.sgd_part = 2;
        QFX_CUDA_TRY(cudaStreamWaitEvent(side, d.dgrad_event, 0));
        if (!(skip & 128)) sgd_conv_train<<<gsgd, 256, 0, side>>>(sb);
        QFX_CUDA_TRY(cudaEventRecord(d.join_event, side));
        d.side_pending = true;
        s.sgd_part = 1;
        if (!(skip & 128)) sgd_conv_train<<<gsgd, 256, 0, stream>>>(s);
    }
    return check_launch("tinycnn_train_step");
}

extern "C" int qfx_tinycnn_train_step(const float* X, const int* labels, const int* offsets, const int* perm,
                                      int n_total, const int* client_ids, int n_selected, int epoch,
                                      int batch_index, int batch_size, float* params, float* moment, int nd,
                                      int n_classes, float lr, float momentum, float p_drop,
                                      unsigned long long seed, int round, const int* round_dev, void* workspace,
                                      int conv2_mode, unsigned long long shuffle_key, cudaStream_t stream) {
    return train_step_impl(X, labels, offsets, perm, n_total, client_ids, n_selected, epoch, batch_index, batch_size,
                           params, moment, nd, n_classes, lr, momentum, p_drop, seed, round, round_dev, workspace,
                           conv2_mode, shuffle_key, stream, false);
}

extern "C" int qfx_tinycnn_train_step_overlapped(const float* X, const int* labels, const int* offsets,
                                                 const int* perm, int n_total, const int* client_ids, int n_selected,
                                                 int epoch, int batch_index, int batch_size, float* params,
                                                 float* moment, int nd, int n_classes, float lr, float momentum,
                                                 float p_drop, unsigned long long seed, int round,
                                                 const int* round_dev, void* workspace, int conv2_mode,
                                                 unsigned long long shuffle_key, cudaStream_t stream) {
    return train_step_impl(X, labels, offsets, perm, n_total, client_ids, n_selected, epoch, batch_index, batch_size,
                           params, moment, nd, n_classes, lr, momentum, p_drop, seed, round, round_dev, workspace,
                           conv2_mode, shuffle_key, stream, true);
}

extern "C" int qfx_tinycnn_finish_steps(cudaStream_t stream) {
    DeviceState* d = nullptr;
    if (int e = device_state(&d)) return e;
    if (d->side_pending) {
        QFX_CUDA_TRY(cudaStreamWaitEvent(stream, d->join_event, 0));
        d->side_pending = false;
    }
    return 0;
}

extern "C" int qfx_tinycnn_stats_offset(int n_selected, size_t* offset) {
    size_t tot = 0;
    qfx_tinycnn_workspace_bytes(n_selected, &tot);
    *offset = tot - (((size_t)n_selected * 8 + 255) & ~(size_t)255);
    return 0;
}

// Evaluation: forward over N images with one parameter vector; out2 = (loss sum, #correct).
// scratch: n * 23,520 + 102,400 bytes.
extern "C" int qfx_tinycnn_eval(const float* X, const int* labels, int n, const float* params, int n_classes,
                                void* scratch, float* out2, cudaStream_t stream) {
    QFX_ARG_CHECK(n >= 1 && n_classes >= 2 && n_classes <= MAX_CLASSES, QFX_ERR_BAD_ARG, "bad n / n_classes");
    DeviceState* d = nullptr;
    if (int e = device_state(&d)) return e;
    if (int e = set_attrs(*d)) return e;
    char* w = (char*)scratch;
    float* w2f = (float*)w; w += 12800 * 4;
    float* w2b = (float*)w; w += 12800 * 4;
    float* a1 = (float*)w; w += (size_t)n * 3136 * 4;
    float* a2 = (float*)w; w += (size_t)n * 1568 * 4;
    unsigned char* i1 = (unsigned char*)w; w += (size_t)n * 3136;
    unsigned char* i2 = (unsigned char*)w;
    QFX_CUDA_TRY(cudaMemsetAsync(out2, 0, 2 * sizeof(float), stream));
    conv2_weight_prep<<<dim3(50, 1), 256, 0, stream>>>(params, 0, w2f, w2b);
    conv1_fwd_eval<<<n, 224, 0, stream>>>(X, params, a1, i1);
    conv2_fwd_eval<<<n, 256, CONV2_SMEM, stream>>>(a1, w2f, params, a2, i2);
    head_eval<<<(n + EVB - 1) / EVB, 256, 0, stream>>>(a2, labels, params, n_classes, n, out2);
    return check_launch("tinycnn_eval");
}
