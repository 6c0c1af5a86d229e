// TinyCNN conv2 (5x5, 16 -> 32 channels, pad 2) on the tcgen05 tensor cores as a ZERO-COPY implicit GEMM.
//
// The activation of one image is kept in shared memory as bf16 "half planes": plane h holds channels
// 8h..8h+7 of every pixel, 16 bytes per pixel, pixels laid out row-major with a pitch of 32 and a
// 2-pixel zero border.  Eight consecutive pixels of a plane are 128 contiguous bytes - exactly one
// UMMA core matrix (8 rows x 16 bytes) of the un-swizzled K-major operand layout.  The im2col matrix of
// filter tap (ky, kx) is therefore nothing but the same planes viewed through a descriptor whose start
// address is shifted by (ky * 32 + kx) pixels: no im2col buffer is ever built.
//
//   forward     D[pixel, oc]  = sum_tap  A_tap[pixel, c]   W_tap[oc, c]^T     M = 4 x 128 pixels, N = 32, K = 16/tap
//   data grad   D[pixel, c]   = sum_tap  G_tap[pixel, oc]  Wflip_tap[c, oc]^T M = 4 x 128, N = 16, K = 32/tap
//   weight grad D_tap[oc, c]  = sum_pix  G[pix, oc]^T      A_tap[pix, c]       M = 64 (32 used), N = 16, K = 448 pixels
//
// (for the weight gradient both operands are read "MN-major": the 8 channels of a pixel are the
// contiguous 16 bytes, 8 consecutive pixels form the K direction of a core matrix.)
// Accumulators live in TMEM; one elected thread issues the MMAs; four warps run the epilogue
// (tcgen05.ld -> bias/ReLU/2x2 max-pool, or plain stores).  Operands are bf16, accumulation fp32,
// master weights and all other layers stay fp32 (standard mixed precision).
//
// Restates the conv2 part of the reference model (/root/reference/src/CFed/Classical_FL.py:25,33-34).
#include <cuda_bf16.h>
#include "tinycnn.cuh"
#include "umma.cuh"
#include "../../include/qfedx_b200.h"

namespace qfx {
namespace cnn {

constexpr int PITCH = 32;                 // pixels per stored row
constexpr int PLANE_PIX = 672;            // (14 + 4) rows x 32 + slack for the garbage rows of the last M block
constexpr int PLANE_BYTES = PLANE_PIX * 16;
constexpr int W2U_BYTES = 25 * 1024;      // 25 taps x (2 k-chunks x 4 oc-groups | 4 k-chunks x 2 c-groups) x 128 B

// planar fp32 [C][14][14] (global) -> bf16 half planes with pitch 32 and a 2-pixel zero border (shared).
// The planes must have been zero-filled before.  One item = (pixel, half plane): 8 channel values.
template <int C>
__device__ __forceinline__ void fill_half_planes(uint8_t* planes, const float* __restrict__ src) {
    constexpr int ITEMS = 196 * (C / 8);
    for (int it0 = threadIdx.x; it0 < ITEMS; it0 += 2 * blockDim.x) {
        float v[2][8];
        int dst[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {   // issue both items' loads before converting: two L2 latencies overlap
            const int it = it0 + u * blockDim.x;
            dst[u] = -1;
            if (it < ITEMS) {
                const int h = it / 196, pix = it - h * 196, y = pix / 14, x = pix - y * 14;
#pragma unroll
                for (int j = 0; j < 8; ++j) v[u][j] = __ldg(src + (8 * h + j) * 196 + pix);
                dst[u] = h * PLANE_BYTES + ((y + 2) * PITCH + x + 2) * 16;
            }
        }
#pragma unroll
        for (int u = 0; u < 2; ++u)
            if (dst[u] >= 0)
                *reinterpret_cast<uint4*>(planes + dst[u]) = make_uint4(pack_bf16(v[u][0], v[u][1]), pack_bf16(v[u][2], v[u][3]),
                                                                        pack_bf16(v[u][4], v[u][5]), pack_bf16(v[u][6], v[u][7]));
    }
}
__device__ __forceinline__ void zero_bytes16(uint8_t* p, int bytes) {
    for (int i = threadIdx.x * 16; i < bytes; i += blockDim.x * 16) *reinterpret_cast<uint4*>(p + i) = make_uint4(0, 0, 0, 0);
}

// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) conv2_weight_prep_umma(const float* __restrict__ params, int nd, __nv_bfloat16* w2uf,
                                                              __nv_bfloat16* w2ub) {
    const int sel = blockIdx.y, i = blockIdx.x * 256 + threadIdx.x;
    if (i >= 12800) return;
    const int o = i / 400, rem = i - o * 400, c = rem / 25, tap = rem - c * 25;
    store_w2u(w2uf + (size_t)sel * 12800, w2ub + (size_t)sel * 12800, o, c, tap, params[(size_t)sel * nd + C2W + i]);
}

// ------------------------------------------------------------------------------------------
// forward: conv + bias + ReLU + 2x2 max-pool for one image per CTA.  192 threads: warp 1 issues the MMAs,
// warps 2..5 run the epilogue, everybody stages the operands.
constexpr int FWD_SMEM = 2 * PLANE_BYTES + W2U_BYTES + 14 * 16 * 33 * 4 + 128;   // planes | weights | row maxima
__device__ __forceinline__ void conv2_fwd_umma_image(const float* __restrict__ a1, const __nv_bfloat16* __restrict__ w2uf,
                                                     const float* __restrict__ bias, float* __restrict__ a2,
                                                     unsigned char* __restrict__ i2, uint8_t* smem) {
    __shared__ __align__(8) uint64_t accum_bar;
    __shared__ uint32_t tmem_slot;
    uint8_t* planes = smem;                                   // 2 half planes
    uint8_t* wts = smem + 2 * PLANE_BYTES;                    // 25 KB
    float* hmax = reinterpret_cast<float*>(wts + W2U_BYTES);  // [14 rows][16 x-pairs][33]: horizontal maxima (+arg in the sign-free low bit)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (warp == 1 && lane == 0) {
        mbar_init1(&accum_bar);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" ::"r"(smem_addr(&tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    zero_bytes16(planes, 2 * PLANE_BYTES);
    for (int i = tid; i < W2U_BYTES / 16; i += blockDim.x) cp_async16(wts + 16 * i, reinterpret_cast<const uint8_t*>(w2uf) + 16 * i);
    __syncthreads();
    fill_half_planes<16>(planes, a1);
    cp_async_wait_all();
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the tensor core
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_slot;

    if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = idesc_bf16(128, 32, 0, 0);
            const uint32_t pa = smem_addr(planes), wa = smem_addr(wts);
#pragma unroll 1
            for (int blk = 0; blk < 4; ++blk) {
#pragma unroll 1
                for (int tap = 0; tap < 25; ++tap) {
                    const int ky = tap / 5, kx = tap - ky * 5;
                    // A: rows = 128 consecutive pixels starting at (128 blk + ky*32 + kx); 8-row groups 128 B apart;
                    //    second k-chunk (channels 8..15) one plane further
                    const uint64_t ad = desc_noswz(pa + (uint32_t)(128 * blk + ky * PITCH + kx) * 16, PLANE_BYTES, 128);
                    // B: rows = 32 oc; oc groups 128 B apart; second k-chunk 512 B further
                    const uint64_t bd = desc_noswz(wa + (uint32_t)tap * 1024, 512, 128);
                    umma_bf16(tmem + 32 * blk, ad, bd, idesc, tap > 0 ? 1u : 0u);
                }
            }
            umma_commit1(&accum_bar);
        }
        __syncwarp();
    } else if (warp >= 2) {
        const int q = warp & 3;                  // TMEM lane quarter = image row within the block
        mbar_wait0(&accum_bar);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll 1
        for (int blk = 0; blk < 4; ++blk) {
            const int y = 4 * blk + q;           // image row; lane = x
            uint32_t r[32];
            tmem_ld<32>(tmem + ((uint32_t)(q * 32) << 16) + 32 * blk, r);
            if (y < 14) {
#pragma unroll
                for (int o = 0; o < 32; ++o) {
                    const float v = __uint_as_float(r[o]);
                    const float p = __shfl_xor_sync(0xffffffffu, v, 1);
                    if (!(lane & 1) && lane < 14) {
                        // horizontal maximum of (x, x+1); the winner's column goes into bit 0 of the payload
                        const bool right = p > v;
                        const float best = right ? p : v;
                        hmax[(y * 16 + (lane >> 1)) * 33 + o] = __uint_as_float((__float_as_uint(best) & ~1u) | (right ? 1u : 0u));
                    }
                }
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        asm volatile("bar.sync 1, 128;" ::: "memory");
        // vertical maximum + bias + ReLU: 32 oc x 49 pooled positions
        for (int e = tid - 64; e < 32 * 49; e += 128) {
            const int o = e / 49, pos = e - o * 49, py = pos / 7, px = pos - py * 7;
            const uint32_t t0 = __float_as_uint(hmax[((2 * py) * 16 + px) * 33 + o]);
            const uint32_t t1 = __float_as_uint(hmax[((2 * py + 1) * 16 + px) * 33 + o]);
            const float v0 = __uint_as_float(t0 & ~1u), v1 = __uint_as_float(t1 & ~1u);
            const bool lower = v1 > v0;
            const float best = fmaxf((lower ? v1 : v0) + bias[o], 0.f);
            a2[o * 49 + pos] = best;
            i2[o * 49 + pos] = (unsigned char)(lower ? (2 + (t1 & 1u)) : (t0 & 1u));
        }
    }
    __syncthreads();
    if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" ::"r"(tmem) : "memory");
}

__global__ void __launch_bounds__(192) conv2_fwd_umma_debug(const float* __restrict__ a1, const __nv_bfloat16* __restrict__ w2uf,
                                                            const float* __restrict__ bias, float* a2, unsigned char* i2) {
    extern __shared__ __align__(128) uint8_t smem_dyn[];
    const size_t r = blockIdx.x;
    conv2_fwd_umma_image(a1 + r * 3136, w2uf, bias, a2 + r * 1568, i2 + r * 1568, smem_dyn);
}

// ------------------------------------------------------------------------------------------
// data gradient: da1[c][y][x] = sum_o sum_tap g2pad[o][y+ky][x+kx] W[o][c][4-ky][4-kx]; K = 32 oc per tap
constexpr int DGRAD_UMMA_SMEM = 4 * PLANE_BYTES + W2U_BYTES + 128;
__device__ __forceinline__ void conv2_dgrad_umma_image(const float* __restrict__ g2, const __nv_bfloat16* __restrict__ w2ub,
                                                       float* __restrict__ da1, uint8_t* smem) {
    __shared__ __align__(8) uint64_t accum_bar;
    __shared__ uint32_t tmem_slot;
    uint8_t* planes = smem;                       // 4 half planes of g2 (32 oc)
    uint8_t* wts = smem + 4 * PLANE_BYTES;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (warp == 1 && lane == 0) {
        mbar_init1(&accum_bar);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_addr(&tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    zero_bytes16(planes, 4 * PLANE_BYTES);
    for (int i = tid; i < W2U_BYTES / 16; i += blockDim.x) cp_async16(wts + 16 * i, reinterpret_cast<const uint8_t*>(w2ub) + 16 * i);
    __syncthreads();
    fill_half_planes<32>(planes, g2);
    cp_async_wait_all();
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_slot;
    if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = idesc_bf16(128, 16, 0, 0);
            const uint32_t pa = smem_addr(planes), wa = smem_addr(wts);
#pragma unroll 1
            for (int blk = 0; blk < 4; ++blk) {
#pragma unroll 1
                for (int tap = 0; tap < 25; ++tap) {
                    const int ky = tap / 5, kx = tap - ky * 5;
#pragma unroll
                    for (int j = 0; j < 2; ++j) {   // K = 32 oc = two MMAs of 16 (k-chunks 2j, 2j+1)
                        const uint64_t ad = desc_noswz(pa + (uint32_t)(2 * j) * PLANE_BYTES +
                                                       (uint32_t)(128 * blk + ky * PITCH + kx) * 16, PLANE_BYTES, 128);
                        const uint64_t bd = desc_noswz(wa + (uint32_t)tap * 1024 + (uint32_t)(2 * j) * 256, 256, 128);
                        umma_bf16(tmem + 16 * blk, ad, bd, idesc, (tap | j) ? 1u : 0u);
                    }
                }
            }
            umma_commit1(&accum_bar);
        }
        __syncwarp();
    } else if (warp >= 2) {
        const int q = warp & 3;
        mbar_wait0(&accum_bar);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll 1
        for (int blk = 0; blk < 4; ++blk) {
            const int y = 4 * blk + q;
            uint32_t r[16];
            tmem_ld<16>(tmem + ((uint32_t)(q * 32) << 16) + 16 * blk, r);
            if (y < 14 && lane < 14) {
#pragma unroll
                for (int c = 0; c < 16; ++c) da1[c * 196 + y * 14 + lane] = __uint_as_float(r[c]);
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }
    __syncthreads();
    if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem) : "memory");
}

// ------------------------------------------------------------------------------------------
// weight gradient of one image: D[(kx, c), oc] = sum_pixels a1pad[pixel + tap shift, c] * g2[pixel, oc] per filter row.
// Both operands are read MN-major straight from the half planes.  The A operand is the activation half plane with
// its 16-byte M chunks only ONE PIXEL (16 B) apart, so M chunk j is the plane shifted by j pixels: rows 0..39 of the
// accumulator are the five kx taps x 8 channels, rows 40..127 (further shifts) are never read.  B is g2 with its four
// 8-oc planes as N chunks (N = 32, no padding).  Ten accumulators (5 ky x 2 channel halves) of 32 columns each sit
// in two TMEM allocations (256 + 64 columns), which leaves 192 columns for the data-gradient CTAs that share the SM.
// The dynamic shared memory is sized so that exactly ONE weight-gradient CTA fits on an SM (two would deadlock on
// the second TMEM allocation) while a data-gradient CTA still fits beside it.
constexpr int WG_PLANES = 6;
constexpr int WG_STAGE_BYTES = 32 * 401 * 4;
constexpr int WGRAD_UMMA_SMEM = 120 * 1024;
static_assert(WG_PLANES * PLANE_BYTES + WG_STAGE_BYTES + 128 <= WGRAD_UMMA_SMEM, "wgrad shared memory layout");
static_assert(2 * WGRAD_UMMA_SMEM > 227 * 1024, "two weight-gradient CTAs must not fit on one SM");
__device__ __forceinline__ void conv2_wgrad_umma_image(const float* __restrict__ g2, const float* __restrict__ a1,
                                                       float* __restrict__ part, uint8_t* smem, int n_img) {
    __shared__ __align__(8) uint64_t accum_bar;
    __shared__ uint32_t tmem_slot[2];
    __shared__ float bias_sm[32];
    uint8_t* gpl = smem;                          // planes 0..3: g2 (32 oc); 4..5: a1 (16 c)
    uint8_t* apl = smem + 4 * PLANE_BYTES;
    float* stage = reinterpret_cast<float*>(smem + WG_PLANES * PLANE_BYTES);   // [32 oc][401] fp32
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (warp == 1 && lane == 0) {
        mbar_init1(&accum_bar);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_addr(&tmem_slot[0])) : "memory");
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_addr(&tmem_slot[1])) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    zero_bytes16(smem, WG_PLANES * PLANE_BYTES);
    uint32_t tmem0 = 0, tmem1 = 0;
#pragma unroll 1
    for (int img = 0; img < n_img; ++img) {
        // the previous image's MMAs must have consumed the planes before they are overwritten
        if (img > 0) { mbar_wait_parity(&accum_bar, (img - 1) & 1); }
        __syncthreads();
        fill_half_planes<32>(gpl, g2 + (size_t)img * 32 * 196);
        fill_half_planes<16>(apl, a1 + (size_t)img * 3136);
        // conv2 bias gradient in fp32 from the dense g2: every warp sums whole channels with coalesced reads
        for (int oc = warp; oc < 32; oc += 6) {
            const float* row = g2 + (size_t)img * 32 * 196 + oc * 196;
            float v = 0.f;
#pragma unroll
            for (int i = 0; i < 7; ++i) { const int p = lane + 32 * i; v += p < 196 ? __ldg(row + p) : 0.f; }
            v = warp_sum(v);
            if (lane == 0) bias_sm[oc] = (img ? bias_sm[oc] : 0.f) + v;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        tmem0 = tmem_slot[0]; tmem1 = tmem_slot[1];
        if (warp == 1 && lane == 0) {
            constexpr uint32_t idesc = idesc_bf16(128, 32, 1, 1);
            const uint32_t ga = smem_addr(gpl), aa = smem_addr(apl);
            // K runs over the stored pixels 64 .. 511 (rows 2..15 of the padded g2 planes), 16 per MMA;
            // 8-pixel K groups are LBO = 128 B apart.  A: M chunks one pixel (16 B) apart.  B: oc chunks one plane apart.
            const uint64_t a_hi = desc_noswz(0, 128, 16) >> 32, b_hi = desc_noswz(0, 128, PLANE_BYTES) >> 32;
            const uint32_t lbo_lo = (uint32_t)((128u >> 4) << 16);
#pragma unroll 1
            for (int kyh = 0; kyh < 10; ++kyh) {
                const int ky = kyh >> 1, h = kyh & 1;
                const uint32_t acc_addr = kyh < 8 ? tmem0 + 32 * kyh : tmem1 + 32 * (kyh - 8);
                // may start 2 pixels before the plane: g2 is zero for those K rows
                uint32_t a_lo = (((aa + (uint32_t)h * PLANE_BYTES + (uint32_t)(64 - 66 + ky * PITCH) * 16) >> 4) & 0x3FFF) | lbo_lo;
                uint32_t b_lo = (((ga + 64 * 16) >> 4) & 0x3FFF) | lbo_lo;
#pragma unroll 4
                for (int kk = 0; kk < 28; ++kk) {
                    umma_bf16(acc_addr, (a_hi << 32) | a_lo, (b_hi << 32) | b_lo, idesc, (img | kk) ? 1u : 0u);
                    a_lo += 16; b_lo += 16;   // 16 pixels = 256 B
                }
            }
            umma_commit1(&accum_bar);
        }
    }
    if (warp == 0) part[C2B + lane] = bias_sm[lane];
    if ((warp & 3) < 2) {
        // accumulator row = TMEM lane: rows 0..31 sit in lane quarter 0 (warps 0 and 4), rows 32..39 in quarter 1
        // (warps 1 and 5); the two warps of a quarter take five accumulators each
        mbar_wait_parity(&accum_bar, (n_img - 1) & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int quarter = warp & 3, row = quarter * 32 + lane, kx = row >> 3, c8 = row & 7;
        const uint32_t lane_field = (uint32_t)(quarter * 32) << 16;
        const int k0 = (warp >> 2) * 5;
#pragma unroll 1
        for (int kyh = k0; kyh < k0 + 5; ++kyh) {
            const int ky = kyh >> 1, h = kyh & 1;
            uint32_t r[32];
            tmem_ld<32>((kyh < 8 ? tmem0 + 32 * kyh : tmem1 + 32 * (kyh - 8)) + lane_field, r);
            if (kx < 5) {
                const int c = 8 * h + c8;
#pragma unroll
                for (int oc = 0; oc < 32; ++oc) stage[oc * 401 + c * 25 + ky * 5 + kx] = __uint_as_float(r[oc]);
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }
    __syncthreads();
    for (int e = tid; e < 12800; e += blockDim.x) part[C2W + e] = stage[(e / 400) * 401 + (e % 400)];
    if (warp == 2) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem0) : "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem1) : "memory");
    }
}

__global__ void __launch_bounds__(192) conv2_dgrad_umma_debug(const float* __restrict__ g2, const __nv_bfloat16* __restrict__ w2ub,
                                                              float* da1) {
    extern __shared__ __align__(128) uint8_t smem_dyn[];
    const size_t r = blockIdx.x;
    conv2_dgrad_umma_image(g2 + r * 32 * 196, w2ub, da1 + r * 3136, smem_dyn);
}
__global__ void __launch_bounds__(192) conv2_wgrad_umma_debug(const float* __restrict__ g2, const float* __restrict__ a1,
                                                              float* part) {
    extern __shared__ __align__(128) uint8_t smem_dyn[];
    const size_t r = blockIdx.x;
    conv2_wgrad_umma_image(g2 + r * 32 * 196, a1 + r * 3136, part + r * NCONV, smem_dyn, 1);
}

// ------------------------------------------------------------------------------------------
// training-step wrappers: grid (image, client)
__global__ void __launch_bounds__(192) conv2_fwd_umma_train(const Step s) {
    extern __shared__ __align__(128) uint8_t smem_dyn[];
    const int sel = blockIdx.y, b = blockIdx.x;
    int lo;
    if (b >= batch_rows(s, sel, &lo)) return;
    const size_t r = (size_t)sel * BS + b;
    conv2_fwd_umma_image(s.a1 + r * 3136, s.w2uf + (size_t)sel * 12800, s.params + (size_t)sel * s.nd + C2B,
                         s.a2 + r * 1568, s.i2 + r * 1568, smem_dyn);
}
__global__ void __launch_bounds__(192) conv2_dgrad_umma_train(const Step s) {
    extern __shared__ __align__(128) uint8_t smem_dyn[];
    const int sel = blockIdx.y, b = blockIdx.x;
    int lo;
    if (b >= batch_rows(s, sel, &lo)) return;
    const size_t r = (size_t)sel * BS + b;
    conv2_dgrad_umma_image(s.g2 + r * 32 * 196, s.w2ub + (size_t)sel * 12800, s.da1 + r * 3136, smem_dyn);
}
__global__ void __launch_bounds__(192) conv2_wgrad_umma_train(const Step s) {
    extern __shared__ __align__(128) uint8_t smem_dyn[];
    const int sel = blockIdx.y, ipc = s.wgrad_ipc, b0 = blockIdx.x * ipc;   // this CTA sums images b0 .. b0+ipc-1
    int lo;
    const int nb = batch_rows(s, sel, &lo);
    float* part = s.conv_part + ((size_t)sel * BS + blockIdx.x) * NCONV;
    if (b0 >= nb) {  // absent images contribute zeros to the reduction in sgd_conv_train
        if (nb > 0) for (int i = threadIdx.x; i < NCONV; i += blockDim.x) part[i] = 0.f;
        return;
    }
    const size_t r = (size_t)sel * BS + b0;
    conv2_wgrad_umma_image(s.g2 + r * 32 * 196, s.a1 + r * 3136, part, smem_dyn, min(ipc, nb - b0));
}

static PerDeviceFlag g_umma_attr;
static int umma_attrs() {
    if (g_umma_attr.here()) return 0;
    QFX_CUDA_TRY(cudaFuncSetAttribute(conv2_fwd_umma_train, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM));
    QFX_CUDA_TRY(cudaFuncSetAttribute(conv2_dgrad_umma_train, cudaFuncAttributeMaxDynamicSharedMemorySize, DGRAD_UMMA_SMEM));
    QFX_CUDA_TRY(cudaFuncSetAttribute(conv2_wgrad_umma_train, cudaFuncAttributeMaxDynamicSharedMemorySize, WGRAD_UMMA_SMEM));
    g_umma_attr.here() = true;
    return 0;
}
int launch_conv2_umma_prep(const float* params, int nd, int n_selected, __nv_bfloat16* w2uf, __nv_bfloat16* w2ub,
                           cudaStream_t stream) {
    conv2_weight_prep_umma<<<dim3(50, n_selected), 256, 0, stream>>>(params, nd, w2uf, w2ub);
    return check_launch("conv2_weight_prep_umma");
}
int launch_conv2_umma_fwd(const Step& s, int n_selected, cudaStream_t stream) {
    if (int e = umma_attrs()) return e;
    conv2_fwd_umma_train<<<dim3(s.batch_size, n_selected), 192, FWD_SMEM, stream>>>(s);
    return check_launch("conv2_fwd_umma_train");
}
int launch_conv2_umma_dgrad(const Step& s, int n_selected, cudaStream_t stream) {
    if (int e = umma_attrs()) return e;
    conv2_dgrad_umma_train<<<dim3(s.batch_size, n_selected), 192, DGRAD_UMMA_SMEM, stream>>>(s);
    return check_launch("conv2_dgrad_umma_train");
}
int launch_conv2_umma_wgrad(const Step& s, int n_selected, cudaStream_t stream) {
    if (int e = umma_attrs()) return e;
    conv2_wgrad_umma_train<<<dim3(BS / s.wgrad_ipc, n_selected), 192, WGRAD_UMMA_SMEM, stream>>>(s);
    return check_launch("conv2_wgrad_umma_train");
}

}  // namespace cnn
}  // namespace qfx

using namespace qfx;
using namespace qfx::cnn;

// Debug / test entry: conv2 forward (+bias, ReLU, pool) of n images with the tcgen05 kernel.
// a1 [n,16,14,14] fp32, weights [32,16,5,5] fp32, bias [32]; a2 [n,32,7,7], i2 [n,32,7,7] u8;
// scratch: 51,200 bytes for the operand-layout weight copies.
extern "C" int qfx_tinycnn_conv2_fwd_umma(const float* a1, const float* weight, const float* bias, int n, float* a2,
                                          unsigned char* i2, void* scratch, cudaStream_t stream) {
    QFX_ARG_CHECK(n >= 1 && a1 && weight && bias && a2 && i2 && scratch, QFX_ERR_BAD_ARG, "bad arguments");
    __nv_bfloat16* w2uf = (__nv_bfloat16*)scratch;
    __nv_bfloat16* w2ub = w2uf + 12800;
    // weight points at a bare [32,16,5,5] tensor: use nd = 0 and shift the base so that C2W + i lands on it
    conv2_weight_prep_umma<<<dim3(50, 1), 256, 0, stream>>>(weight - C2W, 0, w2uf, w2ub);
    QFX_CUDA_TRY(cudaFuncSetAttribute(conv2_fwd_umma_debug, cudaFuncAttributeMaxDynamicSharedMemorySize, FWD_SMEM));
    conv2_fwd_umma_debug<<<n, 192, FWD_SMEM, stream>>>(a1, w2uf, bias, a2, i2);
    return check_launch("conv2_fwd_umma");
}

// Debug / test entries for the backward kernels.  g2 [n,32,14,14] (gradient w.r.t. the conv2 output before
// pooling), a1 [n,16,14,14]; da1 [n,16,14,14]; part [n,13248] per-image partials (conv2 weight gradient at
// offset 416 in [o][c][tap] order, conv2 bias gradient at 13216).
extern "C" int qfx_tinycnn_conv2_dgrad_umma(const float* g2, const float* weight, int n, float* da1, void* scratch,
                                            cudaStream_t stream) {
    QFX_ARG_CHECK(n >= 1 && g2 && weight && da1 && scratch, QFX_ERR_BAD_ARG, "bad arguments");
    __nv_bfloat16* w2uf = (__nv_bfloat16*)scratch;
    __nv_bfloat16* w2ub = w2uf + 12800;
    conv2_weight_prep_umma<<<dim3(50, 1), 256, 0, stream>>>(weight - C2W, 0, w2uf, w2ub);
    QFX_CUDA_TRY(cudaFuncSetAttribute(conv2_dgrad_umma_debug, cudaFuncAttributeMaxDynamicSharedMemorySize, DGRAD_UMMA_SMEM));
    conv2_dgrad_umma_debug<<<n, 192, DGRAD_UMMA_SMEM, stream>>>(g2, w2ub, da1);
    return check_launch("conv2_dgrad_umma");
}

extern "C" int qfx_tinycnn_conv2_wgrad_umma(const float* g2, const float* a1, int n, float* part, cudaStream_t stream) {
    QFX_ARG_CHECK(n >= 1 && g2 && a1 && part, QFX_ERR_BAD_ARG, "bad arguments");
    QFX_CUDA_TRY(cudaFuncSetAttribute(conv2_wgrad_umma_debug, cudaFuncAttributeMaxDynamicSharedMemorySize, WGRAD_UMMA_SMEM));
    conv2_wgrad_umma_debug<<<n, 192, WGRAD_UMMA_SMEM, stream>>>(g2, a1, part);
    return check_launch("conv2_wgrad_umma");
}
