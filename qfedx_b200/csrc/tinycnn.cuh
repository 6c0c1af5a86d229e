// Shared definitions of the client-batched TinyCNN kernels (tinycnn.cu: fp32 CUDA-core kernels,
// tinycnn_umma.cu: tcgen05 implicit-GEMM conv2 kernels).
#pragma once
#include <cuda_bf16.h>
#include "common.cuh"

namespace qfx {
namespace cnn {

constexpr int BS = 32;           // max minibatch rows handled by the fc kernels
constexpr int C1W = 0, C1B = 400, C2W = 416, C2B = 13216, F1W = 13248, F1B = 113600, F2W = 113664;
constexpr int NCONV = 13248;     // conv parameters (reduced from per-image partials)
constexpr int FC1_SLICES = 14;   // split-K slices of the 1568-long fc1 reduction
constexpr int FC1_KS = 112;      // 1568 / 14
constexpr int MAX_CLASSES = 10;
constexpr int HEAD_GRAD_STRIDE = 64 + 65 * MAX_CLASSES + 6;   // fc2 weights, fc2 bias, fc1 bias gradients (padded)

struct Step {
    const float* X;          // [N,1,28,28] dataset
    const int* labels;       // [N]
    const int* offsets;      // [K+1]
    const int* perm;         // [E,N] or null
    const int* client_ids;   // [K_sel] rows -> client segment
    int n_total, epoch, batch_index, batch_size;
    float* params;           // [K_sel, ND]
    float* moment;           // [K_sel, ND]
    int nd;                  // row stride of params/moment
    int n_classes;
    int split;               // conv2 kernels: channel groups are spread over `split` CTAs (1, 2 or 4)
    float lr, momentum, p_drop;
    unsigned long long seed;
    int round;
    const int* round_dev;    // when non-null the round is read from device memory (CUDA-graph replays)
    unsigned long long shuffle_key;  // non-zero and perm == null: keyed Feistel sample order
    // activations / workspaces, all [K_sel, ...]
    float* a1; unsigned char* i1;    // [K,32,16,14,14]
    float* a1_alt;                   // second a1 buffer: overlapped steps alternate, so the next step's conv1 forward
                                     // may run while this step's conv2 weight gradient still reads a1
    int sgd_part;                    // sgd_conv_train: 0 everything, 1 conv1 + head only, 2 conv2 only
    float* a2; unsigned char* i2;    // [K,32,32,7,7]
    float* fc1_part;                 // [K,16,32,64]
    float* g2;                       // [K,32,32,14,14]
    float* da1;                      // [K,32,16,14,14]
    float* conv_part;                // [K,32,NCONV]
    float* head_grad;                // [K,HEAD_GRAD_STRIDE] fc2 / fc1-bias gradients of the step
    float* w2f;                      // [K,12800] conv2 weights as [c][tap][o]      (forward operand)
    float* w2b;                      // [K,12800] conv2 weights as [o][24-tap][c]   (data-gradient operand)
    __nv_bfloat16* w2uf;             // [K,12800] bf16 UMMA forward operand  [tap][c/8][oc/8][oc%8][c%8]
    __nv_bfloat16* w2ub;             // [K,12800] bf16 UMMA dgrad operand    [24-tap][oc/8][c/8][c%8][oc%8]
    int conv2_mode;                  // 0: fp32 CUDA-core conv2 kernels, 1: tcgen05 bf16 implicit-GEMM kernels
    int wgrad_ipc;                   // tcgen05 weight gradient: images accumulated per CTA; conv_part then has BS / ipc rows
    float* stats;                    // [K,2] running loss sum / correct count of the epoch
};

// Asynchronous global->shared copies (LDGSTS): every copy of a staging loop is in flight at once, so a
// CTA pays one memory latency for its whole working set instead of one per loop iteration.
__device__ __forceinline__ void cp_async4(void* smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async8(void* smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// padded 18x18 planes in shared memory from dense 14x14 planes in global memory (async; the caller
// waits with cp_async_wait_all + __syncthreads)
__device__ __forceinline__ void fill_padded_planes(float* __restrict__ xs, const float* __restrict__ src, int planes) {
    for (int i = threadIdx.x; i < planes * 324; i += blockDim.x) xs[i] = 0.f;
    __syncthreads();
    for (int i = threadIdx.x; i < planes * 98; i += blockDim.x) {  // 7 float2 per 14-float row
        const int c = i / 98, rem = i - c * 98, r = rem / 7, q2 = rem - r * 7;
        cp_async8(xs + c * 324 + (r + 2) * 18 + 2 + 2 * q2, src + c * 196 + r * 14 + 2 * q2);
    }
}
__device__ __forceinline__ void copy_f4(float* __restrict__ dst, const float* __restrict__ src, int n) {
    for (int i = threadIdx.x * 4; i < n; i += blockDim.x * 4) cp_async16(dst + i, src + i);
}

__device__ __forceinline__ int batch_rows(const Step& s, int sel, int* seg_lo) {
    const int client = s.client_ids[sel];
    const int lo = s.offsets[client], n_k = s.offsets[client + 1] - lo;
    *seg_lo = lo;
    const int b0 = s.batch_index * s.batch_size;
    return max(0, min(s.batch_size, n_k - b0));
}
__device__ __forceinline__ int sample_index(const Step& s, int sel, int seg_lo, int b) {
    const int j = s.batch_index * s.batch_size + b;
    if (s.perm) return s.perm[(size_t)s.epoch * s.n_total + seg_lo + j];
    if (s.shuffle_key == 0ull) return seg_lo + j;
    // keyed Feistel order, fresh per (round, client, epoch); the round comes from device memory under graph replay
    const int client = s.client_ids[sel];
    const int n_k = s.offsets[client + 1] - seg_lo;
    const unsigned long long rnd = (unsigned long long)(s.round_dev ? *s.round_dev : s.round);
    return seg_lo + (int)feistel_perm((uint32_t)j, (uint32_t)n_k, key3(s.shuffle_key ^ (rnd << 20), (unsigned long long)client,
                                                                      (unsigned long long)s.epoch));
}


// operand copies of the conv2 weights in UMMA layouts (bf16), kept current by the SGD kernel:
//   w2uf (forward):  [tap][k-chunk h (2)][oc group (4)][8 oc][8 c]     B is [N = oc][K = c], K-major
//   w2ub (dgrad):    [24-tap][k-chunk h (4)][c group (2)][8 c][8 oc]   B is [N = c][K = oc], K-major
__device__ __forceinline__ void store_w2u(__nv_bfloat16* w2uf, __nv_bfloat16* w2ub, int o, int c, int tap, float v) {
    const __nv_bfloat16 bv = __float2bfloat16_rn(v);
    w2uf[(((tap * 2 + (c >> 3)) * 4 + (o >> 3)) * 8 + (o & 7)) * 8 + (c & 7)] = bv;
    w2ub[((((24 - tap) * 4 + (o >> 3)) * 2 + (c >> 3)) * 8 + (c & 7)) * 8 + (o & 7)] = bv;
}

// tcgen05 conv2 kernels for the training step (tinycnn_umma.cu)
int launch_conv2_umma_prep(const float* params, int nd, int n_selected, __nv_bfloat16* w2uf, __nv_bfloat16* w2ub,
                           cudaStream_t stream);
int launch_conv2_umma_fwd(const Step& s, int n_selected, cudaStream_t stream);
int launch_conv2_umma_dgrad(const Step& s, int n_selected, cudaStream_t stream);
int launch_conv2_umma_wgrad(const Step& s, int n_selected, cudaStream_t stream);

// evaluation head (tinycnn.cu): fc1 + ReLU + fc2 + loss / arg-max over pooled conv2 outputs a2 [n,1568]
int launch_head_eval(const float* a2, const int* labels, const float* prm, int n_classes, int n, float* out2,
                     cudaStream_t stream);

}  // namespace cnn
}  // namespace qfx
