// K8p: persistent per-client TinyCNN local training - ONE launch per federated round.
//
// The reference walks a client's minibatches with a few dozen library kernels per SGD step
// (/root/reference/src/CFed/Classical_FL.py:40-64; model :21-38).  The step-per-graph-node version of K8
// (tinycnn.cu) still pays nine kernel boundaries per step.  Here a group of G CTAs per client stays resident for
// the whole round (cooperative launch, one CTA per SM) and walks epochs x minibatches itself; the kernel
// boundaries become five group barriers per step (a counter in global memory, release/acquire at gpu scope):
//
//   PA  per image   conv1+bias+ReLU+pool on CUDA cores -> bf16 hi/lo half planes in shared memory ->
//                   conv2 forward on tcgen05 (zero-copy implicit GEMM) -> bias+ReLU+pool from TMEM -> a2
//   PB  per K slice fc1 forward partial products (split-K over the 1568 inputs, weights stationary)
//   PC  per CTA     head (fc1 bias+ReLU+dropout, fc2, softmax-CE, d logits, dh), then per 32-column chunk
//                   dW1 + SGD-momentum in place and dx (masked by ReLU) for the image owners
//   PD  per image   unpool dx into bf16 hi/lo planes -> conv2 data gradient and weight gradient on tcgen05
//                   (weight gradient accumulated in TMEM over the CTA's images) -> conv1 weight gradient
//   PE  per slice   reduce the per-CTA gradient rows, SGD-momentum for conv + head parameters, refresh the
//                   operand copies of the conv2 weights
//
// Precision: conv2 runs on the tensor cores with a two-term bf16 split of BOTH operands (x = hi + lo, three MMAs
// hi*hi + hi*lo + lo*hi, fp32 accumulation in TMEM): products are exact to ~2^-16 relative - 30x tighter than the
// TF32 that cuDNN uses for the reference's fp32 convolution by default - everything else is fp32.
//
// Planes use a pitch of 16 pixels (14 valid + 2 zero columns; the zero columns double as left/right padding of the
// neighbouring rows), so one image is two 128-row MMA blocks and the weight gradient's K dimension is 224 pixels.
#include "tinycnn.cuh"
#include "umma.cuh"
#include "../../include/qfedx_b200.h"

namespace qfx {
namespace cnn {

constexpr int PNT = 512;                      // worker threads per CTA (16 warps)
constexpr int PNT_ALL = PNT + 32;             // + one warp whose lane 0 only issues the tensor-core work
// Workers synchronise among themselves on named barrier 1; the MMA issuer never joins it, so a full MMA queue
// cannot hold the workers up (they meet the issuer only through mbarriers).
#define WSYNC() asm volatile("bar.sync 1, 512;" ::: "memory")
constexpr int PPLANE = 352 * 16;              // bytes per half plane: 18 rows x 16 pixels x 16 B + zero slack
constexpr int WOP = 25600;                    // bytes of one bf16 operand copy of the conv2 weights
constexpr int OFF_A1 = 128;                   // [hi0 hi1 lo0 lo1] activation planes (behind 128 B of zeros), set 0
constexpr int OFF_A1B = OFF_A1 + 4 * PPLANE;  // set 1: PA double-buffers the planes so conv1 of the next image runs
                                              // on the CUDA cores while the tensor core works on the previous one
constexpr int OFF_G2 = OFF_A1B + 4 * PPLANE;  // [hi0..3 lo0..3] conv2-output-gradient planes
constexpr int OFF_WF = OFF_G2 + 8 * PPLANE;   // forward operand hi | lo; scratch of PB / PC / PD once consumed
constexpr int OFF_WB = OFF_WF + 2 * WOP;      // data-gradient operand hi | lo
constexpr int OFF_X32 = OFF_WB + 2 * WOP;     // input image, 32 x 32 with a 2-pixel zero border (conv1 forward)
constexpr int OFF_X37 = OFF_X32 + 32 * 32 * 4;   // same image with row stride 37 (conv1 weight gradient)
constexpr int OFF_W1C = OFF_X37 + 32 * 37 * 4;   // conv1 weights + bias (416 floats)
constexpr int OFF_B2 = OFF_W1C + 416 * 4;        // conv2 bias (32) | conv2 bias gradient accumulator (32)
constexpr int PSMEM = OFF_B2 + 256;
static_assert(OFF_G2 % 128 == 0 && OFF_WF % 128 == 0 && OFF_WB % 128 == 0 && OFF_X32 % 16 == 0 && OFF_W1C % 16 == 0, "alignment");
static_assert(PSMEM <= 227 * 1024, "shared memory");
// scratch inside [OFF_WF, OFF_WB + ...): sizes checked against the 2 * WOP bytes of the forward operand
constexpr int FCX = 116;                      // row stride of the fc1 forward tiles (16-byte aligned rows)
static_assert((32 + 64) * FCX * 4 <= 2 * WOP, "PB tiles");
// TMEM columns.  Forward (PA) and backward (PD) accumulators are never live together and share columns:
// forward 2 image buffers x 2 blocks x 64 (hi | lo weight halves); data gradient 2 x 32; weight gradient 10 x 40 behind it.
constexpr int TM_FWD = 0, TM_FWD_SET = 128, TM_DG = 0, TM_WG = 64, WG_N = 40;
static_assert(TM_WG + 10 * WG_N <= 512, "TMEM columns");
constexpr int XS1P = 37;

struct Round {
    Step s;                  // dataset, parameters, hyper-parameters (epoch / batch_index are walked by the kernel)
    int epochs, n_batches;
    float* dxm;              // [K,32,1568] ReLU-masked gradient of the pooled conv2 output
    __nv_bfloat16* wop;      // [K,4,12800] conv2 operand copies: forward hi, forward lo, dgrad hi, dgrad lo
    unsigned* bar;           // [K] group barrier counters (zeroed by the host before the launch)
    int* err;                // set when a barrier gave up waiting (a bug elsewhere must not hang the GPU)
    long long* prof;         // optional [K, G, 24] per-CTA cycle counters per phase (null: no timing)
};
// phase timer: thread 0 adds the cycles since the previous mark to slot `id`
#define QFX_MARK(id)                                                              \
    do {                                                                          \
        if (R.prof && tid == 0) {                                                 \
            const long long _t = clock64();                                       \
            R.prof[((size_t)sel * gridDim.x + cta) * 24 + (id)] += _t - t_mark;   \
            t_mark = _t;                                                          \
        }                                                                         \
    } while (0)

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// All G CTAs of a client meet here.  Writes before the barrier are visible to every CTA of the group after it: the
// worker barrier orders the CTA's writes before thread 0's release-add on the counter (cumulativity), the counter
// is polled with gpu-scope acquire loads (which also drop this SM's stale L1 lines), and the second worker barrier
// orders every thread's later loads behind thread 0's acquire.
__device__ __forceinline__ void group_barrier(unsigned* ctr, unsigned& target, unsigned G, int* err) {
    target += G;
    WSYNC();
    if (threadIdx.x == 0) {
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
        unsigned spins = 0;
        while (ld_acquire_u32(ctr) < target) {
            if (++spins > (1u << 26)) { atomicExch(err, 1); break; }   // ~seconds: give up instead of hanging
        }
    }
    WSYNC();
}

__device__ __forceinline__ void mbar_wait_all(uint64_t* bar, uint32_t& parity) {
    mbar_wait_parity(bar, parity);
    parity ^= 1u;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
// workers: operands written -> visible to the tensor core -> tell the issuer
__device__ __forceinline__ void hand_to_issuer(uint64_t* ready) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    WSYNC();
    if (threadIdx.x == 0) mbar_arrive(ready);
}
// generic-proxy writes to shared memory -> visible to the tensor core, then a CTA-wide sync
__device__ __forceinline__ void publish_operands() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// Operand copies of the conv2 weights (bf16, written by PE / the prep kernel, 2048 B per filter tap):
//   forward        [tap][k-chunk c/8 (2)][hi|lo][oc group (4)][8 oc][8 c]          B = [N = (hi|lo, oc)][K = c]
//   data gradient  [24-tap][k-chunk oc/8 (4)][hi|lo][c group (2)][8 c][8 oc]       B = [N = (hi|lo, c)][K = oc]
// Putting the hi and lo halves of the weights side by side in N lets ONE MMA multiply an activation half plane by
// both (the epilogue adds the two column halves), so a split product costs two activation reads instead of three.
__host__ __device__ __forceinline__ int wop_fwd_index(int tap, int c, int o, int hl) {
    return ((((tap * 2 + (c >> 3)) * 2 + hl) * 4 + (o >> 3)) * 8 + (o & 7)) * 8 + (c & 7);
}
__host__ __device__ __forceinline__ int wop_dgrad_index(int tap, int c, int o, int hl) {
    return 25600 + (((((24 - tap) * 4 + (o >> 3)) * 2 + hl) * 2 + (c >> 3)) * 8 + (c & 7)) * 8 + (o & 7);
}
__device__ __forceinline__ void store_wop(__nv_bfloat16* wo, int o, int c, int tap, float w) {
    float hi, lo;
    split_bf16(w, hi, lo);
    const __nv_bfloat16 bh = __float2bfloat16_rn(hi), bl = __float2bfloat16_rn(lo);
    wo[wop_fwd_index(tap, c, o, 0)] = bh;   wo[wop_fwd_index(tap, c, o, 1)] = bl;
    wo[wop_dgrad_index(tap, c, o, 0)] = bh; wo[wop_dgrad_index(tap, c, o, 1)] = bl;
}

// ------------------------------------------------------------------------------------------
// PA, first half: conv1 (5x5, 1 -> 16, pad 2) + bias + ReLU + 2x2 max-pool of the image in xs32.  Thread
// (half, position) computes 8 channels of one pooled position and writes them straight into the bf16 hi/lo planes
// (16 bytes per pixel and half plane); the fp32 activation and the arg-max / sign byte go to global memory for PD.
__device__ __forceinline__ void conv1_to_planes(uint8_t* smem, float* __restrict__ a1g, unsigned char* __restrict__ i1g,
                                                int planes_off = OFF_A1) {
    const int tid = threadIdx.x;
    if (tid >= 392) return;
    const float* xs = reinterpret_cast<const float*>(smem + OFF_X32);
    const float* ws = reinterpret_cast<const float*>(smem + OFF_W1C);
    const int half = tid >= 196 ? 1 : 0, pos = tid - half * 196, py = pos / 14, px = pos - py * 14;
    float win[6][6];
#pragma unroll
    for (int r = 0; r < 6; ++r)
#pragma unroll
        for (int c = 0; c < 6; ++c) win[r][c] = xs[(2 * py + r) * 32 + 2 * px + c];
    float v8[8];
#pragma unroll
    for (int oo = 0; oo < 8; ++oo) {
        const int o = half * 8 + oo;
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
        v8[oo] = best;
        if (a1g) {
            a1g[o * 196 + pos] = best;
            i1g[o * 196 + pos] = (unsigned char)(arg | (best > 0.f ? 4 : 0));
        }
    }
    uint4 hi, lo;
    pack8_hi_lo(v8, hi, lo);
    const int slot = (py + 2) * 16 + px;
    *reinterpret_cast<uint4*>(smem + planes_off + half * PPLANE + slot * 16) = hi;
    *reinterpret_cast<uint4*>(smem + planes_off + (2 + half) * PPLANE + slot * 16) = lo;
}

// activation planes of an image from its fp32 copy in global memory (PD, when the CTA owns several images)
__device__ __forceinline__ void rebuild_a1_planes(uint8_t* smem, const float* a1g) {
    const int tid = threadIdx.x;
    if (tid >= 392) return;
    const int half = tid >= 196 ? 1 : 0, pos = tid - half * 196, py = pos / 14, px = pos - py * 14;
    float v8[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v8[j] = __ldcg(a1g + (half * 8 + j) * 196 + pos);
    uint4 hi, lo;
    pack8_hi_lo(v8, hi, lo);
    const int slot = (py + 2) * 16 + px;
    *reinterpret_cast<uint4*>(smem + OFF_A1 + half * PPLANE + slot * 16) = hi;
    *reinterpret_cast<uint4*>(smem + OFF_A1 + (2 + half) * PPLANE + slot * 16) = lo;
}

// one thread: conv2 forward, D[pixel slot, (hi|lo, oc)] += sum over taps A_tap[pixel slot, c] [W_hi | W_lo]_tap[oc, c]^T,
// once for the hi and once for the lo activation planes (all four partial products; the epilogue adds the halves)
template <bool ELECT = false>
__device__ __forceinline__ void issue_conv2_fwd(uint32_t sbase, uint32_t tmem, int planes_off = OFF_A1) {
    constexpr uint32_t idesc = idesc_bf16(128, 64, 0, 0);
#pragma unroll 1
    for (int blk = 0; blk < 2; ++blk) {
#pragma unroll 5
        for (int tap = 0; tap < 25; ++tap) {
            const int ky = tap / 5, kx = tap - ky * 5;
            // rows = 128 consecutive pixel slots; output slot p reads input slot p + ky*16 + kx - 2 (the two slots in front
            // of the first plane are zeros); 8-row groups 128 B apart, the second k-chunk (channels 8..15) one plane on
            const uint32_t a_hi = sbase + (uint32_t)planes_off + (uint32_t)((128 * blk + ky * 16 + kx - 2) * 16);
            const uint64_t ah = desc_noswz(a_hi, PPLANE, 128), al = desc_noswz(a_hi + 2 * PPLANE, PPLANE, 128);
            const uint64_t b = desc_noswz(sbase + OFF_WF + (uint32_t)tap * 2048, 1024, 128);
            if constexpr (ELECT) {
                umma_bf16_elect(tmem + TM_FWD + 64 * blk, al, b, idesc, tap > 0 ? 1u : 0u);
                umma_bf16_elect(tmem + TM_FWD + 64 * blk, ah, b, idesc, 1u);
            } else {
                umma_bf16(tmem + TM_FWD + 64 * blk, al, b, idesc, tap > 0 ? 1u : 0u);
                umma_bf16(tmem + TM_FWD + 64 * blk, ah, b, idesc, 1u);
            }
        }
    }
}

// conv2 forward epilogue: bias + ReLU + 2x2 max-pool straight from TMEM.  TMEM lane = pixel slot within the block:
// a lane quarter holds two image rows (lanes 0-15 / 16-31), so both pooling partners are one shuffle away.
__device__ __forceinline__ void conv2_fwd_epilogue(uint32_t tmem, const float* __restrict__ bias /*smem*/,
                                                   float* __restrict__ a2, unsigned char* __restrict__ i2) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, q = warp & 3, cg = warp >> 2;
#pragma unroll
    for (int blk = 0; blk < 2; ++blk) {
        uint32_t r[8], r2[8];
        tmem_ld<8>(tmem + ((uint32_t)(q * 32) << 16) + TM_FWD + 64 * blk + 8 * cg, r);
        tmem_ld<8>(tmem + ((uint32_t)(q * 32) << 16) + TM_FWD + 64 * blk + 32 + 8 * cg, r2);
        const int x = lane & 15, py = 4 * blk + q, px = x >> 1;
#pragma unroll
        for (int o = 0; o < 8; ++o) {
            const float v = __uint_as_float(r[o]) + __uint_as_float(r2[o]);
            const float p = __shfl_xor_sync(0xffffffffu, v, 1);
            const bool right = p > v;                       // meaningful on even-x lanes
            const float hbest = right ? p : v;
            const float vb = __shfl_xor_sync(0xffffffffu, hbest, 16);
            const int va = __shfl_xor_sync(0xffffffffu, right ? 1 : 0, 16);
            if (lane < 16 && !(lane & 1) && x < 14 && py < 7) {
                const bool lower = vb > hbest;
                const int oc = 8 * cg + o;
                const float best = fmaxf((lower ? vb : hbest) + bias[oc], 0.f);
                a2[oc * 49 + py * 7 + px] = best;
                i2[oc * 49 + py * 7 + px] = (unsigned char)(lower ? 2 + va : (right ? 1 : 0));
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}

// ------------------------------------------------------------------------------------------
// PD: planes of g2 (gradient w.r.t. the conv2 output before pooling) from the pooled, ReLU-masked gradient dxm and
// the arg-max bytes: every 2x2 window gets its value at the arg-max position and zeros elsewhere.
__device__ __forceinline__ void build_g2_planes(uint8_t* smem, const float* __restrict__ dxm, const unsigned char* __restrict__ i2) {
    const int tid = threadIdx.x;
    if (tid >= 392) return;
    const int item = tid >> 1, dy = tid & 1, h = item / 49, pos = item - h * 49, py = pos / 7, px = pos - py * 7;
    float v[8]; int a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        v[j] = __ldcg(dxm + (8 * h + j) * 49 + pos);
        a[j] = __ldcg(i2 + (8 * h + j) * 49 + pos);
    }
#pragma unroll
    for (int dx = 0; dx < 2; ++dx) {
        float w[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) w[j] = a[j] == dy * 2 + dx ? v[j] : 0.f;
        uint4 hi, lo;
        pack8_hi_lo(w, hi, lo);
        const int slot = (2 * py + dy + 2) * 16 + 2 * px + dx;
        *reinterpret_cast<uint4*>(smem + OFF_G2 + h * PPLANE + slot * 16) = hi;
        *reinterpret_cast<uint4*>(smem + OFF_G2 + (4 + h) * PPLANE + slot * 16) = lo;
    }
}

// one thread: data gradient D[pixel slot, (hi|lo, c)] += sum_tap G_tap[pixel slot, oc] [Wflip_hi | Wflip_lo]_tap[c, oc]^T
// (K = 32 oc = 2 MMAs), for the hi and the lo planes of g2
__device__ __forceinline__ void issue_conv2_dgrad(uint32_t sbase, uint32_t tmem) {   // convergent warp, one lane issues
    constexpr uint32_t idesc = idesc_bf16(128, 32, 0, 0);
#pragma unroll 1
    for (int blk = 0; blk < 2; ++blk) {
#pragma unroll 5
        for (int tap = 0; tap < 25; ++tap) {
            const int ky = tap / 5, kx = tap - ky * 5;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const uint32_t a_hi = sbase + OFF_G2 + (uint32_t)(2 * j) * PPLANE + (uint32_t)((128 * blk + ky * 16 + kx - 2) * 16);
                const uint64_t ah = desc_noswz(a_hi, PPLANE, 128), al = desc_noswz(a_hi + 4 * PPLANE, PPLANE, 128);
                const uint64_t b = desc_noswz(sbase + OFF_WB + (uint32_t)tap * 2048 + (uint32_t)(2 * j) * 512, 512, 128);
                umma_bf16_elect(tmem + TM_DG + 32 * blk, al, b, idesc, (tap | j) ? 1u : 0u);
                umma_bf16_elect(tmem + TM_DG + 32 * blk, ah, b, idesc, 1u);
            }
        }
    }
}

// one thread: weight gradient.  Accumulator (ky, h): D[(hi|lo, oc), (kx, c8)] += sum over pixel slots
// g2[slot, oc] a1[slot + shift(ky) + kx, 8h + c8].  Both operands are read MN-major straight from the planes: A = g2 with
// M = 64 - the eight contiguous planes [hi oc groups | lo oc groups] are the eight M chunks - and B = an activation half
// plane whose N chunks are one pixel apart (= the five kx shifts, N = 40).  Issued for the hi and the lo activation plane,
// so all four partial products land in the accumulator; the epilogue adds rows oc and 32 + oc.  K = pixel slots 32 .. 255.
__device__ __forceinline__ void issue_conv2_wgrad(uint32_t sbase, uint32_t tmem, bool accumulate) {
    constexpr uint32_t idesc = idesc_bf16(64, WG_N, 1, 1);
    const uint64_t a_hi_bits = desc_noswz(0, 128, PPLANE) >> 32, b_hi_bits = desc_noswz(0, 128, 16) >> 32;
    const uint32_t lbo_lo = (uint32_t)((128u >> 4) << 16);
#pragma unroll 1
    for (int kyh = 0; kyh < 10; ++kyh) {
        const int ky = kyh >> 1, h = kyh & 1;
        const uint32_t acc = tmem + TM_WG + WG_N * kyh;
        uint32_t a = (((sbase + OFF_G2 + 32 * 16) >> 4) & 0x3FFF) | lbo_lo;
        uint32_t b_h = (((sbase + OFF_A1 + (uint32_t)h * PPLANE + (uint32_t)((16 * ky - 2) * 16)) >> 4) & 0x3FFF) | lbo_lo;
        uint32_t b_l = (((sbase + OFF_A1 + (uint32_t)(2 + h) * PPLANE + (uint32_t)((16 * ky - 2) * 16)) >> 4) & 0x3FFF) | lbo_lo;
#pragma unroll 2
        for (int kk = 0; kk < 14; ++kk) {
            umma_bf16_elect(acc, (a_hi_bits << 32) | a, (b_hi_bits << 32) | b_l, idesc, (accumulate || kk) ? 1u : 0u);
            umma_bf16_elect(acc, (a_hi_bits << 32) | a, (b_hi_bits << 32) | b_h, idesc, 1u);
            a += 16; b_h += 16; b_l += 16;    // 16 pixel slots = 256 B
        }
    }
}

// ------------------------------------------------------------------------------------------
struct HeadSm {
    float h[BS][65];
    float w2[MAX_CLASSES][64];
    float z[BS][MAX_CLASSES];
    float dz[BS][MAX_CLASSES];
    float red[2];
};
constexpr int PCX = 36;      // row stride of the fc1-backward tiles: 16-byte aligned rows for cp.async; lanes walk along k
struct PcSm {
    float dh[BS][64];
    float xs[2][BS][PCX];    // double-buffered: the next chunk's tiles land while this chunk is computed
    float wsm[2][64][PCX];
    HeadSm hs;
};
static_assert(sizeof(PcSm) <= 2 * WOP, "PC scratch");

// Head for all rows of the minibatch (every CTA of the group computes it: the work is tiny and it saves a barrier):
// h = dropout(relu(sum of split-K partials + b1)), logits, softmax-CE, d logits, dh.  The owner CTA also accumulates
// the epoch statistics and leaves the fc2 / fc1-bias gradients for PE.
__device__ __forceinline__ void head_phase(const Step& s, int sel, int nb, int lo, PcSm& sm, bool owner, int cta, int G,
                                           long long* prof /*this CTA's counters or null*/, long long& t_mark) {
#define HEAD_MARK(id)                                                      \
    do {                                                                   \
        if (prof && threadIdx.x == 0) {                                    \
            const long long _t = clock64();                                \
            prof[id] += _t - t_mark;                                       \
            t_mark = _t;                                                   \
        }                                                                  \
    } while (0)
    HeadSm& hs = sm.hs;
    const int tid = threadIdx.x, NC = s.n_classes;
    const float* __restrict__ prm = s.params + (size_t)sel * s.nd;
    const int f2b = F2W + 64 * NC;
    if (tid < 2) hs.red[tid] = 0.f;
    // the row's label (two dependent global loads: sample order, then label) is fetched now, under the partial sums, and
    // not when the softmax needs it
    int y_row = 0;
    if (tid < BS && tid < nb) y_row = s.labels[sample_index(s, sel, lo, tid)];
    for (int i = tid; i < NC * 64; i += PNT) hs.w2[i / 64][i % 64] = __ldcg(prm + F2W + i);
    const int rnd = s.round_dev ? *s.round_dev : s.round;
    const unsigned long long dkey = key3(s.seed ^ 0x44524f50ull, (unsigned long long)rnd * 4096ull + s.client_ids[sel],
                                         (unsigned long long)s.epoch * 65536ull + s.batch_index);
    const float keep_scale = s.p_drop > 0.f ? 1.f / (1.f - s.p_drop) : 1.f;
    {   // 2048 outputs: thread = (row b, four consecutive features); its 14 split-K partials arrive as 14 independent
        // 16-byte loads issued before any is used
        const int b = tid >> 4, f4 = (tid & 15) * 4;
        float4 pv[FC1_SLICES];
        const float* part = s.fc1_part + ((size_t)sel * FC1_SLICES * BS + b) * 64 + f4;
#pragma unroll
        for (int sl = 0; sl < FC1_SLICES; ++sl)
            pv[sl] = b < nb ? __ldcg(reinterpret_cast<const float4*>(part + (size_t)sl * BS * 64)) : make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 bias = __ldcg(reinterpret_cast<const float4*>(prm + F1B + f4));
        float v[4] = {bias.x, bias.y, bias.z, bias.w};
#pragma unroll
        for (int sl = 0; sl < FC1_SLICES; ++sl) { v[0] += pv[sl].x; v[1] += pv[sl].y; v[2] += pv[sl].z; v[3] += pv[sl].w; }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float h = 0.f;
            if (b < nb) {
                h = fmaxf(v[j], 0.f);
                if (s.p_drop > 0.f) {
                    const float u = uniform_from_bits(stream_at(dkey, (unsigned long long)(b * 64 + f4 + j)));
                    h = u < s.p_drop ? 0.f : h * keep_scale;
                }
            }
            hs.h[b][f4 + j] = h;
        }
    }
    WSYNC();
    HEAD_MARK(19);
    for (int i = tid; i < BS * NC; i += PNT) {
        const int b = i / NC, c = i % NC;
        float v = __ldcg(prm + f2b + c);
#pragma unroll 8
        for (int f = 0; f < 64; ++f) v += hs.h[b][f] * hs.w2[c][f];
        hs.z[b][c] = v;
    }
    WSYNC();
    if (tid < BS) {
        const int b = tid;
        if (b < nb) {
            const int y = y_row;
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
    WSYNC();
    HEAD_MARK(20);
    for (int i = tid; i < BS * 64; i += PNT) {
        const int b = i >> 6, f = i & 63;
        float v = 0.f;
        if (hs.h[b][f] > 0.f) {
            for (int c = 0; c < NC; ++c) v += hs.dz[b][c] * hs.w2[c][f];
            v *= keep_scale;
        }
        sm.dh[b][f] = v;
    }
    if (owner && tid == 0) { s.stats[sel * 2] += hs.red[0]; s.stats[sel * 2 + 1] += hs.red[1]; }
    WSYNC();
    HEAD_MARK(21);
    {
        // gradients of fc2 weights, fc2 bias, fc1 bias (applied in PE): every CTA has the whole head, so the
        // NC*64 + NC + 64 sums over the minibatch are dealt round-robin to the warps of all G CTAs; lane = row
        float* hg = s.head_grad + (size_t)sel * HEAD_GRAD_STRIDE;
        const int warp = tid >> 5, b = tid & 31;
        for (int i = warp * G + cta; i < NC * 64 + NC + 64; i += (PNT / 32) * G) {
            float t;
            if (i < NC * 64) t = hs.dz[b][i / 64] * hs.h[b][i % 64];
            else if (i < NC * 64 + NC) t = hs.dz[b][i - NC * 64];
            else t = sm.dh[b][i - NC * 64 - NC];
            t = warp_sum(b < nb ? t : 0.f);
            if (b == 0) hg[i] = t;
        }
    }
}

// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(PNT_ALL, 1) tinycnn_round_kernel(const Round R) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar_f[2], bar_d, bar_w, rdy_f[2], rdy_b;   // forward: one pair per image buffer
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int sel = blockIdx.y, cta = blockIdx.x, G = gridDim.x;
    Step s = R.s;
    const uint32_t sbase = smem_addr(smem);
    float* prm = s.params + (size_t)sel * s.nd;
    float* mom = s.moment + (size_t)sel * s.nd;
    const __nv_bfloat16* wop = R.wop + (size_t)sel * 4 * 12800;
    unsigned* bar = R.bar + sel;
    unsigned bar_target = 0;
    uint32_t pf[2] = {0, 0}, pd = 0, pw = 0;  // mbarrier phase parities
    long long t_mark = clock64();

    // ---- one-time set-up: zero planes and image borders, TMEM, mbarriers
    for (int i = tid * 16; i < OFF_WF; i += PNT_ALL * 16) *reinterpret_cast<uint4*>(smem + i) = make_uint4(0, 0, 0, 0);
    for (int i = OFF_X32 + tid * 16; i < OFF_W1C; i += PNT_ALL * 16) *reinterpret_cast<uint4*>(smem + i) = make_uint4(0, 0, 0, 0);
    if (warp == 1 && lane == 0) {
        mbar_init1(&bar_f[0]); mbar_init1(&bar_f[1]); mbar_init1(&bar_d); mbar_init1(&bar_w);
        mbar_init1(&rdy_f[0]); mbar_init1(&rdy_f[1]); mbar_init1(&rdy_b);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_addr(&tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_slot;
    float* xs32 = reinterpret_cast<float*>(smem + OFF_X32);
    float* xs37 = reinterpret_cast<float*>(smem + OFF_X37);
    float* b2s = reinterpret_cast<float*>(smem + OFF_B2);       // conv2 bias
    float* b2g = b2s + 32;                                       // conv2 bias gradient of this CTA's images
    const int NC = s.n_classes;

    if (warp == PNT / 32) {
        // ---- MMA issuer: lane 0 walks the same steps as the workers and issues each image's tensor-core work as soon as
        // the workers have published its operands; completion goes back through bar_f / bar_d / bar_w
        // all 32 lanes walk the loops together (uniform control flow, descriptors in uniform registers); one elected lane
        // issues each MMA / commit.  A full 512-column allocation starts at TMEM address 0: using the constant keeps the
        // accumulator addresses out of vector registers (checked once below).
        {
            if (tmem != 0u && lane == 0) atomicExch(R.err, 2);
            constexpr uint32_t tm0 = 0u;
            uint32_t prf[2] = {0, 0}, prb = 0;
#pragma unroll 1
            for (int epoch = 0; epoch < R.epochs; ++epoch) {
                s.epoch = epoch;
#pragma unroll 1
                for (int bi = 0; bi < R.n_batches; ++bi) {
                    s.batch_index = bi;
                    int lo;
                    const int nb = batch_rows(s, sel, &lo);
                    if (nb <= 0) continue;
                    const int my_imgs = cta < nb ? (nb - cta + G - 1) / G : 0;
#pragma unroll 1
                    for (int i = 0; i < my_imgs; ++i) {
                        if (i & 1) {
                            mbar_wait_all(&rdy_f[1], prf[1]);
                            issue_conv2_fwd<true>(sbase, tm0 + TM_FWD_SET, OFF_A1B);
                            umma_commit1_elect(&bar_f[1]);
                        } else {
                            mbar_wait_all(&rdy_f[0], prf[0]);
                            issue_conv2_fwd<true>(sbase, tm0, OFF_A1);
                            umma_commit1_elect(&bar_f[0]);
                        }
                    }
#pragma unroll 1
                    for (int i = 0; i < my_imgs; ++i) {
                        mbar_wait_all(&rdy_b, prb);
                        issue_conv2_dgrad(sbase, tm0);
                        umma_commit1_elect(&bar_d);
                        issue_conv2_wgrad(sbase, tm0, i > 0);
                        umma_commit1_elect(&bar_w);
                    }
                }
            }
        }
    } else {
#pragma unroll 1
    for (int epoch = 0; epoch < R.epochs; ++epoch) {
        s.epoch = epoch;
        if (cta == 0 && tid < 2) s.stats[sel * 2 + tid] = 0.f;
#pragma unroll 1
        for (int bi = 0; bi < R.n_batches; ++bi) {
            s.batch_index = bi;
            int lo;
            const int nb = batch_rows(s, sel, &lo);
            if (nb <= 0) continue;                              // uniform over the client's CTAs
            const int my_imgs = cta < nb ? (nb - cta + G - 1) / G : 0;

            // ================= PA: forward through the convolutions, per image =================
            // this step's weights: conv1 + conv2 bias, forward and data-gradient operand copies (written by PE)
            for (int i = tid; i < 104; i += PNT) cp_async16(smem + OFF_W1C + 16 * i, prm + C1W + 4 * i);
            for (int i = tid; i < 8; i += PNT) cp_async16(smem + OFF_B2 + 16 * i, prm + C2B + 4 * i);
            if (my_imgs > 0) {
                for (int i = tid; i < 4 * WOP / 16; i += PNT)
                    cp_async16(smem + OFF_WF + 16 * i, reinterpret_cast<const uint8_t*>(wop) + 16 * i);
            }
            if (tid < 32) b2g[tid] = 0.f;
            QFX_MARK(15);
            // software pipeline over this CTA's images: conv1 of image i (CUDA cores) runs while the tensor core still
            // works on image i-1; the epilogue of image i-1 runs while the tensor core works on image i
            {
                int n_img = 0;
                size_t prev_row = 0;
#pragma unroll 1
                for (int img = cta; img < nb; img += G, ++n_img) {
                    const int buf = n_img & 1;
                    const size_t row = (size_t)sel * BS + img;
                    const float* src = s.X + (size_t)sample_index(s, sel, lo, img) * 784;
                    for (int i = tid; i < 28 * 14; i += PNT) {
                        const int y = i / 14, x2 = i - y * 14;
                        cp_async8(xs32 + (y + 2) * 32 + 2 + 2 * x2, src + y * 28 + 2 * x2);
                    }
                    cp_async_wait_all();
                    WSYNC();
                    QFX_MARK(0);
                    conv1_to_planes(smem, s.a1 + row * 3136, s.i1 + row * 3136, buf ? OFF_A1B : OFF_A1);
                    hand_to_issuer(&rdy_f[buf]);
                    QFX_MARK(1);
                    if (n_img > 0) {
                        mbar_wait_all(&bar_f[buf ^ 1], pf[buf ^ 1]);
                        QFX_MARK(2);
                        conv2_fwd_epilogue(tmem + (buf ^ 1) * TM_FWD_SET, b2s, s.a2 + prev_row * 1568, s.i2 + prev_row * 1568);
                        QFX_MARK(3);
                    }
                    prev_row = row;
                }
                if (n_img > 0) {
                    const int buf = (n_img - 1) & 1;
                    mbar_wait_all(&bar_f[buf], pf[buf]);
                    QFX_MARK(2);
                    conv2_fwd_epilogue(tmem + buf * TM_FWD_SET, b2s, s.a2 + prev_row * 1568, s.i2 + prev_row * 1568);
                    QFX_MARK(3);
                }
            }
            cp_async_wait_all();
            group_barrier(bar, bar_target, G, R.err);
            QFX_MARK(4);

            // ================= PB: fc1 forward, split-K partial products per 112-column slice =================
#pragma unroll 1
            for (int sl = cta; sl < FC1_SLICES; sl += G) {
                float* xs = reinterpret_cast<float*>(smem + OFF_WF);       // [32][FCX]
                float* wsm = xs + BS * FCX;                                 // [64][FCX]
                const float* x = s.a2 + (size_t)sel * BS * 1568 + sl * FC1_KS;
                const float* w = prm + F1W + sl * FC1_KS;
                WSYNC();
                for (int i = tid; i < (BS + 64) * (FC1_KS / 4); i += PNT) {
                    const int r = i / (FC1_KS / 4), c = i - r * (FC1_KS / 4);
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (r < BS) { if (r < nb) v = __ldcg(reinterpret_cast<const float4*>(x + (size_t)r * 1568) + c); }
                    else v = __ldcg(reinterpret_cast<const float4*>(w + (size_t)(r - BS) * 1568) + c);
                    *reinterpret_cast<float4*>(xs + r * FCX + 4 * c) = v;
                }
                WSYNC();
                if (tid < 128) {
                    const int bt = (tid >> 4) * 4, f0 = tid & 15;
                    float acc[4][4] = {};
#pragma unroll 2
                    for (int k4 = 0; k4 < FC1_KS / 4; ++k4) {
                        float4 xa[4], wa[4];
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            xa[i] = *reinterpret_cast<const float4*>(xs + (bt + i) * FCX + 4 * k4);
                            wa[i] = *reinterpret_cast<const float4*>(wsm + (f0 + 16 * i) * FCX + 4 * k4);
                        }
#pragma unroll
                        for (int i = 0; i < 4; ++i)
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                acc[i][j] += xa[i].x * wa[j].x + xa[i].y * wa[j].y + xa[i].z * wa[j].z + xa[i].w * wa[j].w;
                    }
                    float* out = s.fc1_part + (((size_t)sel * FC1_SLICES + sl) * BS) * 64;
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) out[(bt + i) * 64 + f0 + 16 * j] = acc[i][j];
                }
            }
            QFX_MARK(5);
            group_barrier(bar, bar_target, G, R.err);
            QFX_MARK(6);

            // ================= PC: head, then fc1 backward per 32-column chunk =================
            {
                PcSm& sm = *reinterpret_cast<PcSm*>(smem + OFF_WF);
                // tiles of a 32-column chunk (a2 rows and W1 rows) straight from L2 into shared memory; rows beyond the
                // minibatch were zeroed once below and are never overwritten
                auto fetch_chunk = [&](int chunk, int buf) {
                    const int k0 = chunk * 32;
                    for (int i = tid; i < (BS + 64) * 8; i += PNT) {
                        const int r = i >> 3, c = i & 7;
                        if (r < BS) {
                            if (r < nb) cp_async16(&sm.xs[buf][r][4 * c], s.a2 + ((size_t)sel * BS + r) * 1568 + k0 + 4 * c);
                        } else {
                            cp_async16(&sm.wsm[buf][r - BS][4 * c], prm + F1W + (size_t)(r - BS) * 1568 + k0 + 4 * c);
                        }
                    }
                };
                if (cta < 49) {
                    for (int i = tid; i < 2 * BS * PCX; i += PNT) (&sm.xs[0][0][0])[i] = 0.f;
                    WSYNC();
                    fetch_chunk(cta, 0);      // in flight while the head is computed
                }
                head_phase(s, sel, nb, lo, sm, cta == 0, cta, G,
                           R.prof ? R.prof + ((size_t)sel * gridDim.x + cta) * 24 : nullptr, t_mark);
                QFX_MARK(7);
                int buf = 0;
#pragma unroll 1
                for (int chunk = cta; chunk < 49; chunk += G, buf ^= 1) {
                    const int k0 = chunk * 32;
                    float mprev[8];
                    const int t = tid & 255, k = t & 31;
                    if (tid >= 256) {
                        const int fg = (t >> 5) * 8;
#pragma unroll
                        for (int j = 0; j < 8; ++j) mprev[j] = __ldcg(mom + F1W + (size_t)(fg + j) * 1568 + k0 + k);
                    }
                    cp_async_wait_all();
                    WSYNC();                                  // this chunk's tiles are in; the other buffer is free again
                    if (chunk + G < 49) fetch_chunk(chunk + G, buf ^ 1);
                    if (tid < 256) {
                        // dx[b][k] = sum_f dh[b][f] W1[f][k], masked by the ReLU of a2: rows bq, bq+8, bq+16, bq+24
                        const int bq = t >> 5;
                        float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 8
                        for (int f = 0; f < 64; ++f) {
                            const float w = sm.wsm[buf][f][k];
#pragma unroll
                            for (int r = 0; r < 4; ++r) acc[r] += sm.dh[bq + 8 * r][f] * w;
                        }
#pragma unroll
                        for (int r = 0; r < 4; ++r) {
                            const int b = bq + 8 * r;
                            if (b < nb) R.dxm[((size_t)sel * BS + b) * 1568 + k0 + k] = sm.xs[buf][b][k] > 0.f ? acc[r] : 0.f;
                        }
                    } else {
                        // dW1[f][k] = sum_b dh[b][f] x[b][k]; SGD-momentum in place
                        const int fg = (t >> 5) * 8;
                        float acc[8] = {};
                        for (int b = 0; b < nb; ++b) {
                            const float xv = sm.xs[buf][b][k];
                            const float4 d0 = *reinterpret_cast<const float4*>(&sm.dh[b][fg]);
                            const float4 d1 = *reinterpret_cast<const float4*>(&sm.dh[b][fg + 4]);
                            acc[0] += d0.x * xv; acc[1] += d0.y * xv; acc[2] += d0.z * xv; acc[3] += d0.w * xv;
                            acc[4] += d1.x * xv; acc[5] += d1.y * xv; acc[6] += d1.z * xv; acc[7] += d1.w * xv;
                        }
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const size_t p = F1W + (size_t)(fg + j) * 1568 + k0 + k;
                            const float m = s.momentum * mprev[j] + acc[j];
                            mom[p] = m;
                            prm[p] = sm.wsm[buf][fg + j][k] - s.lr * m;
                        }
                    }
                }
                cp_async_wait_all();
            }
            QFX_MARK(8);
            group_barrier(bar, bar_target, G, R.err);
            QFX_MARK(9);

            // ================= PD: backward through the convolutions, per image =================
            {
                float* gsm = reinterpret_cast<float*>(smem + OFF_WF);                        // [16][196]
                unsigned short* osm = reinterpret_cast<unsigned short*>(smem + OFF_WF + 12544);   // [16][196]
                unsigned char* i1s = smem + OFF_WF + 12544 + 6272;                            // [16][196]
                float c1acc = 0.f;                    // warp o, lane t: conv1 weight (o, tap t < 25) / bias (t == 25) gradient of my images
                int n_done = 0;
#pragma unroll 1
                for (int img = cta; img < nb; img += G, ++n_done) {
                    const size_t row = (size_t)sel * BS + img;
                    WSYNC();
                    const float* src = s.X + (size_t)sample_index(s, sel, lo, img) * 784;
                    for (int i = tid; i < 784; i += PNT) {
                        const int y = i / 28, x = i - y * 28;
                        cp_async4(xs37 + (y + 2) * XS1P + x + 2, src + i);
                    }
                    for (int i = tid; i < 196; i += PNT) cp_async16(i1s + 16 * i, s.i1 + row * 3136 + 16 * i);
                    if (my_imgs > 1) rebuild_a1_planes(smem, s.a1 + row * 3136);
                    build_g2_planes(smem, R.dxm + row * 1568, s.i2 + row * 1568);
                    {   // conv2 bias gradient: every pooled element carries exactly one non-zero of g2
                        for (int oc = warp; oc < 32; oc += PNT / 32) {
                            const float* d = R.dxm + row * 1568 + oc * 49;
                            float v = __ldcg(d + lane) + (lane + 32 < 49 ? __ldcg(d + lane + 32) : 0.f);
                            v = warp_sum(v);
                            if (lane == 0) b2g[oc] += v;
                        }
                    }
                    cp_async_wait_all();
                    hand_to_issuer(&rdy_b);
                    QFX_MARK(10);
                    mbar_wait_all(&bar_d, pd);
                    QFX_MARK(11);
                    {   // data-gradient epilogue: through conv1's ReLU into the sparse form the conv1 weight gradient reads
                        const int q = warp & 3, combo = warp >> 2, blk = combo >> 1, cg = combo & 1;
                        uint32_t r[8], r2[8];
                        tmem_ld<8>(tmem + ((uint32_t)(q * 32) << 16) + TM_DG + 32 * blk + 8 * cg, r);
                        tmem_ld<8>(tmem + ((uint32_t)(q * 32) << 16) + TM_DG + 32 * blk + 16 + 8 * cg, r2);
                        const int slot = 128 * blk + 32 * q + lane, y = slot >> 4, x = slot & 15;
                        if (y < 14 && x < 14) {
                            const int p = y * 14 + x;
#pragma unroll
                            for (int c = 0; c < 8; ++c) {
                                const int ch = 8 * cg + c;
                                const int byte = i1s[ch * 196 + p];
                                gsm[ch * 196 + p] = (byte & 4) ? __uint_as_float(r[c]) + __uint_as_float(r2[c]) : 0.f;
                                osm[ch * 196 + p] = (unsigned short)((2 * y + ((byte >> 1) & 1)) * XS1P + 2 * x + (byte & 1));
                            }
                        }
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    }
                    WSYNC();
                    {   // conv1 weight gradient: warp = output channel, lanes = pooled positions; every non-zero of the
                        // (sparse) gradient meets its 5x5 input window; the 25 + 1 sums are reduced over the warp and
                        // lane `tap` keeps the running total of (channel, tap) across this CTA's images
                        const int o = warp;
                        float acc[26];
#pragma unroll
                        for (int t = 0; t < 26; ++t) acc[t] = 0.f;
#pragma unroll 1
                        for (int p = lane; p < 196; p += 32) {
                            const float g = gsm[o * 196 + p];
                            const float* xt = xs37 + osm[o * 196 + p];
#pragma unroll
                            for (int ky = 0; ky < 5; ++ky)
#pragma unroll
                                for (int kx = 0; kx < 5; ++kx) acc[ky * 5 + kx] += g * xt[ky * XS1P + kx];
                            acc[25] += g;
                        }
#pragma unroll
                        for (int t = 0; t < 26; ++t) {
                            const float v = warp_sum(acc[t]);
                            if (lane == t) c1acc += v;
                        }
                    }
                    QFX_MARK(12);
                    mbar_wait_all(&bar_w, pw);       // the planes may be rebuilt only after the weight-gradient MMAs read them
                    QFX_MARK(13);
                }
                if (my_imgs > 0) {
                    // weight-gradient accumulators -> [oc][c][ky][kx] through shared memory -> this CTA's gradient row
                    WSYNC();
                    float* stage = reinterpret_cast<float*>(smem + OFF_WF);    // [32][401]
                    float* part = s.conv_part + ((size_t)sel * BS + cta) * NCONV;
                    {
                        // M = 64 accumulators: row r sits in TMEM lane 32 * (r / 16) + r % 16, so lane quarter q holds
                        // rows 16q .. 16q+15 in its first 16 lanes: quarters 0, 1 = hi part of oc 0..31, quarters 2, 3 = lo part
                        const int q = warp & 3, oc = 16 * (q & 1) + (lane & 15);
                        for (int pass = 0; pass < 2; ++pass) {        // hi rows store, lo rows add
                            if ((q >> 1) == pass) {
#pragma unroll 1
                                for (int kyh = warp >> 2; kyh < 10; kyh += 4) {
                                    const int ky = kyh >> 1, h = kyh & 1;
                                    uint32_t r[32], r8[8];
                                    const uint32_t ta = tmem + ((uint32_t)(q * 32) << 16) + TM_WG + WG_N * kyh;
                                    tmem_ld<32>(ta, r);
                                    tmem_ld<8>(ta + 32, r8);
                                    if (lane < 16) {
#pragma unroll
                                        for (int col = 0; col < 40; ++col) {
                                            const int kx = col >> 3, c = 8 * h + (col & 7);
                                            const float v = __uint_as_float(col < 32 ? r[col] : r8[col - 32]);
                                            float* d = stage + oc * 401 + c * 25 + ky * 5 + kx;
                                            *d = pass ? *d + v : v;
                                        }
                                    }
                                }
                            }
                            WSYNC();
                        }
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    }
                    if (warp < 16 && lane < 26) part[lane < 25 ? C1W + warp * 25 + lane : C1B + warp] = c1acc;
                    WSYNC();
                    if (tid < 32) part[C2B + tid] = b2g[tid];
                    for (int e = tid; e < 12800; e += PNT) part[C2W + e] = stage[(e / 400) * 401 + (e % 400)];
                }
            }
            QFX_MARK(14);
            group_barrier(bar, bar_target, G, R.err);
            QFX_MARK(16);

            // ================= PE: reduce the gradient rows, SGD-momentum, refresh the operand copies =================
            {
                const int rows = min(G, nb);
                const int n_head = NC * 64 + NC + 64;
                for (int p0 = cta * PNT + tid; p0 < NCONV + n_head; p0 += G * PNT) {
                    if (p0 >= NCONV) {
                        const int i = p0 - NCONV;
                        const int p = i < NC * 64 + NC ? F2W + i : F1B + (i - NC * 64 - NC);
                        const float m = s.momentum * __ldcg(mom + p) + __ldcg(s.head_grad + (size_t)sel * HEAD_GRAD_STRIDE + i);
                        mom[p] = m;
                        prm[p] = __ldcg(prm + p) - s.lr * m;
                        continue;
                    }
                    const float* part = s.conv_part + (size_t)sel * BS * NCONV + p0;
                    float pv[BS];
#pragma unroll
                    for (int r = 0; r < BS; ++r) pv[r] = r < rows ? __ldcg(part + (size_t)r * NCONV) : 0.f;   // all loads in flight
                    float g = 0.f;
#pragma unroll
                    for (int r = 0; r < BS; ++r) g += pv[r];
                    const float m = s.momentum * __ldcg(mom + p0) + g;
                    mom[p0] = m;
                    const float w = __ldcg(prm + p0) - s.lr * m;
                    prm[p0] = w;
                    if (p0 >= C2W && p0 < C2B) {
                        const int i = p0 - C2W, o = i / 400, rem = i - o * 400, c = rem / 25, tap = rem - c * 25;
                        store_wop(R.wop + (size_t)sel * 4 * 12800, o, c, tap, w);
                    }
                }
            }
            QFX_MARK(17);
            group_barrier(bar, bar_target, G, R.err);
            QFX_MARK(18);
        }
    }
    }   // workers
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}

// operand copies (bf16 hi / lo, forward and data-gradient layouts) of the conv2 weights of every selected client
__global__ void __launch_bounds__(256) conv2_weight_prep_split(const float* __restrict__ params, int nd, __nv_bfloat16* wop) {
    const int sel = blockIdx.y, i = blockIdx.x * 256 + threadIdx.x;
    if (i >= 12800) return;
    const int o = i / 400, rem = i - o * 400, c = rem / 25, tap = rem - c * 25;
    store_wop(wop + (size_t)sel * 4 * 12800, o, c, tap, params[(size_t)sel * nd + C2W + i]);
}

// ------------------------------------------------------------------------------------------
// Evaluation forward on the same path and precision as training: conv1 -> conv2 (tcgen05, split operands) per image,
// a persistent CTA per SM walking the images; the fc head runs in head_eval (tinycnn.cu).
__global__ void __launch_bounds__(PNT, 1) tinycnn_eval_conv_kernel(const float* __restrict__ X, int n, const float* __restrict__ prm,
                                                                    const __nv_bfloat16* __restrict__ wop, float* __restrict__ a2,
                                                                    unsigned char* __restrict__ i2) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar_f;
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t sbase = smem_addr(smem);
    uint32_t pf = 0;
    for (int i = tid * 16; i < OFF_WF; i += PNT * 16) *reinterpret_cast<uint4*>(smem + i) = make_uint4(0, 0, 0, 0);
    for (int i = OFF_X32 + tid * 16; i < OFF_X37; i += PNT * 16) *reinterpret_cast<uint4*>(smem + i) = make_uint4(0, 0, 0, 0);
    if (warp == 1 && lane == 0) {
        mbar_init1(&bar_f);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" ::"r"(smem_addr(&tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int i = tid; i < 104; i += PNT) cp_async16(smem + OFF_W1C + 16 * i, prm + C1W + 4 * i);
    for (int i = tid; i < 8; i += PNT) cp_async16(smem + OFF_B2 + 16 * i, prm + C2B + 4 * i);
    for (int i = tid; i < 2 * WOP / 16; i += PNT) cp_async16(smem + OFF_WF + 16 * i, reinterpret_cast<const uint8_t*>(wop) + 16 * i);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_slot;
    float* xs32 = reinterpret_cast<float*>(smem + OFF_X32);
    const float* b2s = reinterpret_cast<const float*>(smem + OFF_B2);
#pragma unroll 1
    for (int img = blockIdx.x; img < n; img += gridDim.x) {
        const float* src = X + (size_t)img * 784;
        for (int i = tid; i < 28 * 14; i += PNT) {
            const int y = i / 14, x2 = i - y * 14;
            cp_async8(xs32 + (y + 2) * 32 + 2 + 2 * x2, src + y * 28 + 2 * x2);
        }
        cp_async_wait_all();
        __syncthreads();
        conv1_to_planes(smem, nullptr, nullptr);
        publish_operands();
        if (warp == 1 && lane == 0) {
            issue_conv2_fwd(sbase, tmem);
            umma_commit1(&bar_f);
        }
        mbar_wait_all(&bar_f, pf);
        conv2_fwd_epilogue(tmem, b2s, a2 + (size_t)img * 1568, i2 + (size_t)img * 1568);
        __syncthreads();
    }
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" ::"r"(tmem) : "memory");
}

static long long* g_round_prof[QFX_MAX_DEVICES] = {};
static PerDeviceFlag g_round_attr;
static int round_attrs() {
    if (g_round_attr.here()) return 0;
    QFX_CUDA_TRY(cudaFuncSetAttribute(tinycnn_round_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PSMEM));
    QFX_CUDA_TRY(cudaFuncSetAttribute(tinycnn_eval_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PSMEM));
    g_round_attr.here() = true;
    return 0;
}

struct RoundWs {
    float* a1; unsigned char* i1; float* a2; unsigned char* i2; float* fc1_part; float* dxm; float* conv_part;
    float* head_grad; __nv_bfloat16* wop; unsigned* bar; int* err; float* stats;
    size_t total;
};
static RoundWs carve_round(void* workspace, size_t K) {
    RoundWs w;
    char* p = (char*)workspace;
    auto take = [&](size_t n) { char* q = p; p += (n + 255) & ~(size_t)255; return q; };
    w.a1 = (float*)take(K * BS * 3136 * 4); w.i1 = (unsigned char*)take(K * BS * 3136);
    w.a2 = (float*)take(K * BS * 1568 * 4); w.i2 = (unsigned char*)take(K * BS * 1568);
    w.fc1_part = (float*)take(K * FC1_SLICES * BS * 64 * 4);
    w.dxm = (float*)take(K * BS * 1568 * 4);
    w.conv_part = (float*)take(K * BS * NCONV * 4);
    w.head_grad = (float*)take(K * HEAD_GRAD_STRIDE * 4);
    w.wop = (__nv_bfloat16*)take(K * 4 * 12800 * 2);
    w.bar = (unsigned*)take(K * 4 + 4);
    w.err = (int*)(w.bar + K);
    w.stats = (float*)take(K * 2 * 4);
    w.total = (size_t)(p - (char*)workspace);
    return w;
}

}  // namespace cnn
}  // namespace qfx

using namespace qfx;
using namespace qfx::cnn;

extern "C" int qfx_tinycnn_round_workspace_bytes(int n_selected, size_t* bytes, size_t* stats_offset) {
    QFX_ARG_CHECK(n_selected >= 1 && bytes, QFX_ERR_BAD_ARG, "bad arguments");
    const RoundWs w = carve_round(nullptr, (size_t)n_selected);
    *bytes = w.total;
    if (stats_offset) *stats_offset = (size_t)((char*)w.stats - (char*)nullptr);
    return 0;
}

// One federated round of local training for every selected client in ONE launch (per group of at most #SM clients).
// params [K, nd] must hold the global model in every row and moment [K, nd] zeros; on return params holds the
// clients' trained models and the workspace tail (stats_offset) the last epoch's (loss sum, #correct) per client.
// ctas_per_client = 0 picks min(32, #SM / K).  Returns QFX_ERR_UNSUPPORTED (-2) if a group barrier timed out.
extern "C" int qfx_tinycnn_train_round(const float* X, const int* labels, const int* offsets, const int* perm, int n_total,
                                       const int* client_ids, int n_selected, int epochs, int n_batches, int batch_size,
                                       float* params, float* moment, int nd, int n_classes, float lr, float momentum,
                                       float p_drop, unsigned long long seed, int round, const int* round_dev,
                                       void* workspace, unsigned long long shuffle_key, int ctas_per_client,
                                       cudaStream_t stream) {
    QFX_ARG_CHECK(batch_size >= 1 && batch_size <= BS, QFX_ERR_UNSUPPORTED,
                  "client-batched TinyCNN supports batch_size <= %d, got %d", BS, batch_size);
    QFX_ARG_CHECK(n_classes >= 2 && n_classes <= MAX_CLASSES, QFX_ERR_UNSUPPORTED, "n_classes %d not in 2..10", n_classes);
    QFX_ARG_CHECK(nd >= F2W + 65 * n_classes && nd % 4 == 0, QFX_ERR_BAD_ARG, "parameter row stride too small or not a multiple of 4");
    QFX_ARG_CHECK(n_selected >= 1 && p_drop >= 0.f && p_drop < 1.f && epochs >= 1 && n_batches >= 1, QFX_ERR_BAD_ARG, "bad arguments");
    QFX_ARG_CHECK(X && labels && offsets && client_ids && params && moment && workspace, QFX_ERR_BAD_ARG, "null pointer");
    if (int e = round_attrs()) return e;
    const int sms = device_sm_count();
    QFX_ARG_CHECK(sms > 0, QFX_ERR_UNSUPPORTED, "cannot query the SM count");
    RoundWs w = carve_round(workspace, (size_t)n_selected);
    QFX_CUDA_TRY(cudaMemsetAsync(w.bar, 0, sizeof(unsigned) * (size_t)n_selected + sizeof(int), stream));
    conv2_weight_prep_split<<<dim3(50, n_selected), 256, 0, stream>>>(params, nd, w.wop);
    if (int e = check_launch("conv2_weight_prep_split")) return e;
    for (int k0 = 0; k0 < n_selected; k0 += sms) {      // more clients than SMs: groups of #SM clients, one CTA each
        const int kc = n_selected - k0 < sms ? n_selected - k0 : sms;
        int G = ctas_per_client > 0 ? ctas_per_client : sms / kc;
        if (G > 32) G = 32;
        if (G < 1) G = 1;
        QFX_ARG_CHECK(G * kc <= sms, QFX_ERR_TOO_LARGE, "%d CTAs per client x %d clients do not fit on %d SMs", G, kc, sms);
        Round R;
        Step& s = R.s;
        s.X = X; s.labels = labels; s.offsets = offsets; s.perm = perm; s.client_ids = client_ids + k0;
        s.n_total = n_total; s.epoch = 0; s.batch_index = 0; s.batch_size = batch_size;
        s.params = params + (size_t)k0 * nd; s.moment = moment + (size_t)k0 * nd; s.nd = nd; s.n_classes = n_classes;
        s.split = 1; s.lr = lr; s.momentum = momentum; s.p_drop = p_drop; s.seed = seed; s.round = round; s.round_dev = round_dev;
        s.shuffle_key = shuffle_key;
        s.a1 = w.a1 + (size_t)k0 * BS * 3136; s.i1 = w.i1 + (size_t)k0 * BS * 3136; s.a1_alt = nullptr; s.sgd_part = 0;
        s.a2 = w.a2 + (size_t)k0 * BS * 1568; s.i2 = w.i2 + (size_t)k0 * BS * 1568;
        s.fc1_part = w.fc1_part + (size_t)k0 * FC1_SLICES * BS * 64;
        s.g2 = nullptr; s.da1 = nullptr;
        s.conv_part = w.conv_part + (size_t)k0 * BS * NCONV;
        s.head_grad = w.head_grad + (size_t)k0 * HEAD_GRAD_STRIDE;
        s.w2f = nullptr; s.w2b = nullptr; s.w2uf = nullptr; s.w2ub = nullptr; s.conv2_mode = 2; s.wgrad_ipc = 1;
        s.stats = w.stats + (size_t)k0 * 2;
        R.epochs = epochs; R.n_batches = n_batches;
        R.dxm = w.dxm + (size_t)k0 * BS * 1568;
        R.wop = w.wop + (size_t)k0 * 4 * 12800;
        R.bar = w.bar + k0;
        R.err = w.err;
        R.prof = g_round_prof[current_device()] ? g_round_prof[current_device()] + (size_t)k0 * 32 * 24 : nullptr;
        void* args[] = {(void*)&R};
        QFX_CUDA_TRY(cudaLaunchCooperativeKernel((const void*)tinycnn_round_kernel, dim3((unsigned)G, (unsigned)kc), dim3(PNT_ALL),
                                                 args, (size_t)PSMEM, stream));
    }
    return 0;
}

// Phase timing for tuning: cycles per phase and CTA are added into prof [K, G, 24] (int64, caller-zeroed; G = CTAs per
// client of the launch) by every following qfx_tinycnn_train_round on the current device; null switches it off.
// Slots: 15 weight load issue, 0 image load, 1 conv1, 2 conv2 forward MMAs, 3 forward epilogue, 4 barrier 1, 5 fc1
// forward, 6 barrier 2, 7 head, 8 fc1 backward, 9 barrier 3, 10 plane build, 11 data-gradient MMAs, 12 its epilogue +
// conv1 weight gradient, 13 weight-gradient MMA wait, 14 weight-gradient epilogue, 16 barrier 4, 17 update, 18 barrier 5.
extern "C" int qfx_tinycnn_round_profile(long long* prof) {
    g_round_prof[current_device()] = prof;
    return 0;
}

// 1 if a group barrier of the last qfx_tinycnn_train_round on this workspace gave up waiting (synchronises the stream)
extern "C" int qfx_tinycnn_round_status(int n_selected, void* workspace, cudaStream_t stream, int* timed_out) {
    QFX_ARG_CHECK(n_selected >= 1 && workspace && timed_out, QFX_ERR_BAD_ARG, "bad arguments");
    RoundWs w = carve_round(workspace, (size_t)n_selected);
    QFX_CUDA_TRY(cudaMemcpyAsync(timed_out, w.err, sizeof(int), cudaMemcpyDeviceToHost, stream));
    QFX_CUDA_TRY(cudaStreamSynchronize(stream));
    return 0;
}

// Evaluation on the tensor-core path: out2 = (loss sum, #correct).  scratch: 102,400 + n * 7,840 bytes.
extern "C" int qfx_tinycnn_eval_tc(const float* X, const int* labels, int n, const float* params, int n_classes,
                                   void* scratch, float* out2, cudaStream_t stream) {
    QFX_ARG_CHECK(n >= 1 && n_classes >= 2 && n_classes <= MAX_CLASSES && X && labels && params && scratch && out2,
                  QFX_ERR_BAD_ARG, "bad arguments");
    if (int e = round_attrs()) return e;
    char* p = (char*)scratch;
    __nv_bfloat16* wop = (__nv_bfloat16*)p; p += 102400;
    float* a2 = (float*)p; p += (size_t)n * 1568 * 4;
    unsigned char* i2 = (unsigned char*)p;
    QFX_CUDA_TRY(cudaMemsetAsync(out2, 0, 2 * sizeof(float), stream));
    conv2_weight_prep_split<<<dim3(50, 1), 256, 0, stream>>>(params, 0, wop);
    const int sms = device_sm_count();
    tinycnn_eval_conv_kernel<<<n < sms ? n : sms, PNT, PSMEM, stream>>>(X, n, params, wop, a2, i2);
    if (int e = check_launch("tinycnn_eval_conv_kernel")) return e;
    return launch_head_eval(a2, labels, params, n_classes, n, out2, stream);
}
