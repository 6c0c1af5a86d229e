// K2: batched circuit unitaries as a complex GEMM on the 5th-generation tensor cores.
//
// All samples of a client share the ansatz parameters, so the ansatz is ONE unitary U (2^n x 2^n)
// applied to a batch of input states: Psi_out = U Psi_in.  With the real embedding
//     [out_re]   [ Ur  -Ui ] [in_re]
//     [out_im] = [ Ui   Ur ] [in_im]
// this is a real GEMM D[M x N] = A[M x K] * B[N x K]^T with M = K = 2 * 2^n and N = batch.
// A (the embedded unitary, row-major, K contiguous) and B (one row per state: [re | im], K
// contiguous) are both "K-major", which is exactly what TMA delivers and tcgen05 consumes.
//
// Kernel structure (one 128 x BN output tile per CTA, 192 threads):
//   warp 0   : TMA producer  - cp.async.bulk.tensor.2d loads of 128 x 32 fp32 boxes, SWIZZLE_128B,
//              into a 3-stage shared-memory ring guarded by full/empty mbarriers
//   warp 1   : MMA issuer    - one elected lane issues tcgen05.mma.cta_group::1.kind::tf32 with
//              shared-memory matrix descriptors; accumulators live in TMEM (BN columns x 128 lanes);
//              tcgen05.commit releases ring slots and finally signals the epilogue
//   warps 2-5: epilogue      - tcgen05.ld (32 lanes x 32 columns per warp per step) -> registers ->
//              coalesced global stores of the output states, optionally fused with the <Z_q>
//              reduction (K3) so the states never have to be re-read
// Precision: "3xTF32".  Each fp32 operand is split on the host side of the kernel into
// hi = tf32-truncated value and lo = x - hi; D += Ahi*Bhi + Ahi*Blo + Alo*Bhi restores ~2^-21
// relative accuracy (plain TF32 would be 2^-10 and cannot meet the 1e-5 tolerance against the
// float64 oracle).  `precision = 1` runs the single TF32 product.
//
// No reference counterpart: /root/reference has no simulator kernel; the op accelerates the
// planned VQC forward (/root/reference/ROADMAP.md:126-128).
#include <cuda.h>
#include "common.cuh"
#include "../../include/qfedx_b200.h"

namespace qfx {
namespace umma {

constexpr int BM = 128;          // output rows per CTA = UMMA M
constexpr int BK = 32;           // fp32 elements per 128-byte swizzled row
constexpr int UMMA_K = 8;        // tf32: 32 bytes per MMA along K
constexpr int STAGES = 3;
constexpr int THREADS = 192;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred = 0;
    asm volatile(
        "{\n"
        ".reg .b32 rx;\n"
        ".reg .pred px;\n"
        "elect.sync rx|px, 0xffffffff;\n"
        "selp.b32 %0, 1, 0, px;\n"
        "}\n" : "=r"(pred));
    return pred != 0;
}

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor, sm100):
// start address >> 4 in [0,14), LBO >> 4 in [16,30) (unused for swizzled K-major, set to 1),
// SBO >> 4 in [32,46) = 1024 B between 8-row groups, version 1 in [46,48), layout 2 in [61,64).
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFF) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): F32 accumulate, TF32 x TF32, both K-major.
__host__ __device__ constexpr uint32_t make_idesc(int m, int n) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct GemmArgs {
    int m_total, n_total, k_total;  // M = K = 2 * 2^n, N = batch (multiple of BN)
    int n_products;                 // 3 = 3xTF32, 1 = plain TF32
    float* out;                     // [N, M] output states (row = state), nullable
    float* expvals;                 // [N, n_qubits] accumulated <Z_q> (zeroed by the launcher), nullable
    int n_qubits;
};

template <int BN>
__global__ void __launch_bounds__(THREADS, 1)
unitary_gemm_kernel(const __grid_constant__ CUtensorMap map_a_hi, const __grid_constant__ CUtensorMap map_a_lo,
                    const __grid_constant__ CUtensorMap map_b_hi, const __grid_constant__ CUtensorMap map_b_lo,
                    const GemmArgs g) {
    constexpr uint32_t A_BYTES = BM * BK * 4, B_BYTES = BN * BK * 4;
    constexpr uint32_t STAGE_BYTES = 2 * A_BYTES + 2 * B_BYTES;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // swizzle-128B tiles need 1024-byte alignment: align the dynamic region by hand
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t full_bar[STAGES];
    __shared__ __align__(8) uint64_t empty_bar[STAGES];
    __shared__ __align__(8) uint64_t accum_bar;
    __shared__ uint32_t tmem_base_slot;
    __shared__ float ztile[4 * 32 * 33];   // per epilogue warp: transposed block of squared amplitudes
    __shared__ float zacc[BN * 16];        // <Z> partial sums of this CTA: [column][qubit (<= 12, padded to 16)]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const int k_blocks = g.k_total / BK;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a_hi) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a_lo) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b_hi) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b_lo) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&accum_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {  // TMEM allocation is warp-wide; the same warp frees it at the end
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32(&tmem_base_slot)), "n"(BN) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_base_slot;

    if (warp == 0) {
        // ================= TMA producer =================
        if (elect_one()) {
            const bool need_lo = g.n_products == 3;
            for (int kb = 0; kb < k_blocks; ++kb) {
                const int s = kb % STAGES;
                const uint32_t phase = (kb / STAGES) & 1;
                mbar_wait(&empty_bar[s], phase ^ 1);  // slot free (first pass falls through)
                uint8_t* st = smem + (size_t)s * STAGE_BYTES;
                mbar_expect_tx(&full_bar[s], need_lo ? STAGE_BYTES : (A_BYTES + B_BYTES));
                tma_load_2d(&map_a_hi, &full_bar[s], st, kb * BK, m0);
                tma_load_2d(&map_b_hi, &full_bar[s], st + 2 * A_BYTES, kb * BK, n0);
                if (need_lo) {
                    tma_load_2d(&map_a_lo, &full_bar[s], st + A_BYTES, kb * BK, m0);
                    tma_load_2d(&map_b_lo, &full_bar[s], st + 2 * A_BYTES + B_BYTES, kb * BK, n0);
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        constexpr uint32_t idesc = make_idesc(BM, BN);
        for (int kb = 0; kb < k_blocks; ++kb) {
            const int s = kb % STAGES;
            const uint32_t phase = (kb / STAGES) & 1;
            mbar_wait(&full_bar[s], phase);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (elect_one()) {
                const uint32_t a_hi = smem_u32(smem + (size_t)s * STAGE_BYTES);
                const uint32_t a_lo = a_hi + A_BYTES;
                const uint32_t b_hi = a_hi + 2 * A_BYTES;
                const uint32_t b_lo = b_hi + B_BYTES;
#pragma unroll
                for (int k = 0; k < BK / UMMA_K; ++k) {
                    const uint32_t koff = k * UMMA_K * 4;  // bytes along the swizzled 128-byte row
                    const uint32_t first = (kb | k) == 0 ? 0u : 1u;
                    umma_tf32(tmem_base, make_desc(a_hi + koff), make_desc(b_hi + koff), idesc, first);
                    if (g.n_products == 3) {
                        umma_tf32(tmem_base, make_desc(a_hi + koff), make_desc(b_lo + koff), idesc, 1u);
                        umma_tf32(tmem_base, make_desc(a_lo + koff), make_desc(b_hi + koff), idesc, 1u);
                    }
                }
                umma_commit(&empty_bar[s]);                       // frees the slot when these MMAs retire
                if (kb == k_blocks - 1) umma_commit(&accum_bar);  // accumulator complete
            }
            __syncwarp();
        }
    } else {
        // ================= epilogue (warps 2..5) =================
        // a warp may only touch the TMEM lane quarter (warp_id % 4): rows 32*q .. 32*q+31 of the tile
        const int q = warp & 3;
        mbar_wait(&accum_bar, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int row = m0 + q * 32 + lane;  // output component index m (0 .. M-1)
        // sign table of this row for the <Z> reduction: component m = part * dim + amplitude index
        const int dim = g.m_total >> 1;
        if (g.expvals != nullptr) {
            for (int e = threadIdx.x - 64; e < BN * 16; e += 128) zacc[e] = 0.f;
            asm volatile("bar.sync 1, 128;" ::: "memory");
        }
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, r);
            if (g.out != nullptr) {
#pragma unroll
                for (int j = 0; j < 32; ++j)  // lanes write consecutive m of one state: coalesced
                    g.out[(size_t)(n0 + c0 + j) * g.m_total + row] = __uint_as_float(r[j]);
            }
            if (g.expvals != nullptr) {
                // <Z_q> = sum_i sign_q(i) |amp_i|^2, and |amp|^2 = re^2 + im^2 is linear in the squares, so
                // every row contributes sign * v^2 on its own.  Transpose the 32x32 block of squares through
                // shared memory: lane j then owns column j and sums it over the warp's 32 rows, once
                // unsigned (for the qubits whose bit is the same for all 32 rows) and once per lane-bit qubit.
                float* tp = ztile + (warp - 2) * (32 * 33);
#pragma unroll
                for (int j = 0; j < 32; ++j) { const float v = __uint_as_float(r[j]); tp[lane * 33 + j] = v * v; }
                __syncwarp();
                float tot = 0.f, d0 = 0.f, d1 = 0.f, d2 = 0.f, d3 = 0.f, d4 = 0.f;
#pragma unroll
                for (int rr = 0; rr < 32; ++rr) {
                    const float pv = tp[rr * 33 + lane];
                    tot += pv;
                    d0 += (rr & 1) ? -pv : pv;  d1 += (rr & 2) ? -pv : pv;  d2 += (rr & 4) ? -pv : pv;
                    d3 += (rr & 8) ? -pv : pv;  d4 += (rr & 16) ? -pv : pv;
                }
                __syncwarp();
                // rows of this warp: amp = base + rr with base a multiple of 32 -> qubits 0..4 follow rr,
                // qubits >= 5 follow base
                const int base = (m0 + q * 32) % dim;
                float* zc = zacc + (c0 + lane) * 16;
                const float dl[5] = {d0, d1, d2, d3, d4};
                for (int qb = 0; qb < g.n_qubits; ++qb) {
                    const float v = qb < 5 ? dl[qb] : (((base >> qb) & 1) ? -tot : tot);
                    atomicAdd(zc + qb, v);
                }
            }
        }
        if (g.expvals != nullptr) {
            // combine the four epilogue warps in shared memory, then one global atomic per (column, qubit)
            asm volatile("bar.sync 1, 128;" ::: "memory");
            const int t = threadIdx.x - 64;  // 0..127
            for (int e = t; e < BN * g.n_qubits; e += 128) {
                const int col = e / g.n_qubits, qb = e - col * g.n_qubits;
                atomicAdd(g.expvals + (size_t)(n0 + col) * g.n_qubits + qb, zacc[col * 16 + qb]);
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }
    __syncthreads();
    if (warp == 2) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(BN) : "memory");
    }
}

// ------------------------------------------------------------------------------------------
// pack kernels
__device__ __forceinline__ void split_tf32(float x, float* hi, float* lo) {
    const float h = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);  // keep 10 mantissa bits
    *hi = h;
    *lo = x - h;
}

// U^T given as complex64 rows (row b = U|b>, what qfx_sv_forward writes for basis inputs) ->
// embedded real matrix A [2 dim, 2 dim] (hi / lo parts)
__global__ void pack_unitary_kernel(const float2* __restrict__ ut, int dim, float* __restrict__ a_hi,
                                    float* __restrict__ a_lo) {
    const int M = 2 * dim;
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < (size_t)M * M;
         idx += (size_t)gridDim.x * blockDim.x) {
        const int m = (int)(idx / M), k = (int)(idx % M);
        const int i = m % dim, j = k % dim;
        const float2 u = ut[(size_t)j * dim + i];  // U[i][j]
        float v;
        if (m < dim) v = (k < dim) ? u.x : -u.y;
        else v = (k < dim) ? u.y : u.x;
        split_tf32(v, a_hi + idx, a_lo + idx);
    }
}

// feature-map product states straight into the GEMM's B operand: row b = [re(2^n) | im(2^n)].
// One CTA handles PACK_SAMPLES samples: their n (cos, sin) pairs are computed once into shared
// memory, then every thread multiplies n factors per amplitude.
constexpr int PACK_SAMPLES = 8;
__global__ void __launch_bounds__(256)
pack_product_states_kernel(const float* __restrict__ angles, int n_qubits, int basis, int batch, int batch_padded,
                           float* __restrict__ b_hi, float* __restrict__ b_lo) {
    __shared__ float2 cs[PACK_SAMPLES][12];
    const int dim = 1 << n_qubits;
    for (int b0 = blockIdx.x * PACK_SAMPLES; b0 < batch_padded; b0 += gridDim.x * PACK_SAMPLES) {
        __syncthreads();
        if (threadIdx.x < PACK_SAMPLES * n_qubits) {
            const int sidx = threadIdx.x / n_qubits, q = threadIdx.x - sidx * n_qubits;
            float sn = 0.f, c = 0.f;
            if (b0 + sidx < batch) sincosf(0.5f * angles[(size_t)(b0 + sidx) * n_qubits + q], &sn, &c);
            cs[sidx][q] = make_float2(c, sn);
        }
        __syncthreads();
        for (int e = threadIdx.x; e < PACK_SAMPLES * dim; e += 256) {
            const int sidx = e / dim, i = e - sidx * dim, b = b0 + sidx;
            if (b >= batch_padded) break;
            float re = (b < batch) ? 1.f : 0.f, im = 0.f;
            for (int q = 0; q < n_qubits; ++q) {
                const float2 v = cs[sidx][q];
                const bool bit = (i >> q) & 1;
                float fr, fi;
                if (basis == 1) { fr = bit ? v.y : v.x; fi = 0.f; }
                else if (basis == 0) { fr = bit ? 0.f : v.x; fi = bit ? -v.y : 0.f; }
                else { fr = bit ? 0.f : v.x; fi = bit ? 0.f : -v.y; }
                const float nr = re * fr - im * fi, ni = re * fi + im * fr;
                re = nr; im = ni;
            }
            const size_t o = (size_t)b * 2 * dim;
            split_tf32(re, b_hi + o + i, b_lo + o + i);
            split_tf32(im, b_hi + o + dim + i, b_lo + o + dim + i);
        }
    }
}

__global__ void split_rows_kernel(const float* __restrict__ x, size_t n_valid, size_t n_total, float* __restrict__ hi,
                                  float* __restrict__ lo) {
    for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < n_total;
         idx += (size_t)gridDim.x * blockDim.x) {
        split_tf32(idx < n_valid ? x[idx] : 0.f, hi + idx, lo + idx);
    }
}

// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_encode = nullptr;

static int get_encoder() {
    if (g_encode) return 0;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    QFX_CUDA_TRY(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    QFX_ARG_CHECK(fn != nullptr && qres == cudaDriverEntryPointSuccess, QFX_ERR_UNSUPPORTED,
                  "cuTensorMapEncodeTiled is not available from this driver");
    g_encode = (EncodeTiledFn)fn;
    return 0;
}

// row-major fp32 [rows, cols] tensor, box = [box_rows, 32 floats], 128-byte swizzle
static int make_map(CUtensorMap* map, const float* base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
    const cuuint64_t dims[2] = {cols, rows};
    const cuuint64_t strides[1] = {cols * sizeof(float)};
    const cuuint32_t box[2] = {BK, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    QFX_ARG_CHECK(r == CUDA_SUCCESS, QFX_ERR_BAD_ARG, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return 0;
}

template <int BN>
static int launch_gemm(const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo, const GemmArgs& g,
                       cudaStream_t stream) {
    CUtensorMap ma_hi, ma_lo, mb_hi, mb_lo;
    if (int e = make_map(&ma_hi, a_hi, g.m_total, g.k_total, BM)) return e;
    if (int e = make_map(&ma_lo, a_lo, g.m_total, g.k_total, BM)) return e;
    if (int e = make_map(&mb_hi, b_hi, g.n_total, g.k_total, BN)) return e;
    if (int e = make_map(&mb_lo, b_lo, g.n_total, g.k_total, BN)) return e;
    constexpr size_t smem = (size_t)STAGES * (2 * BM * BK * 4 + 2 * BN * BK * 4) + 1024;
    QFX_CUDA_TRY(cudaFuncSetAttribute(unitary_gemm_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid(g.m_total / BM, g.n_total / BN);
    unitary_gemm_kernel<BN><<<grid, THREADS, smem, stream>>>(ma_hi, ma_lo, mb_hi, mb_lo, g);
    return check_launch("unitary_gemm_kernel");
}

}  // namespace umma
}  // namespace qfx

using namespace qfx;
using namespace qfx::umma;

extern "C" int qfx_unitary_pack(const float* u_transposed_c64, int n_qubits, float* a_hi, float* a_lo,
                                cudaStream_t stream) {
    QFX_ARG_CHECK(n_qubits >= 6 && n_qubits <= 12, QFX_ERR_UNSUPPORTED, "unitary GEMM supports 6..12 qubits, got %d", n_qubits);
    const int dim = 1 << n_qubits;
    pack_unitary_kernel<<<592, 256, 0, stream>>>(reinterpret_cast<const float2*>(u_transposed_c64), dim, a_hi, a_lo);
    return check_launch("pack_unitary_kernel");
}

extern "C" int qfx_unitary_pack_product_states(const float* angles, int n_qubits, int basis, int batch,
                                               int batch_padded, float* b_hi, float* b_lo, cudaStream_t stream) {
    QFX_ARG_CHECK(batch_padded >= batch && batch_padded % 64 == 0, QFX_ERR_BAD_ARG, "batch_padded must be a multiple of 64");
    pack_product_states_kernel<<<1184, 256, 0, stream>>>(angles, n_qubits, basis, batch, batch_padded, b_hi, b_lo);
    return check_launch("pack_product_states_kernel");
}

extern "C" int qfx_unitary_split_states(const float* states, int n_qubits, int batch, int batch_padded, float* b_hi,
                                        float* b_lo, cudaStream_t stream) {
    QFX_ARG_CHECK(batch_padded >= batch && batch_padded % 64 == 0, QFX_ERR_BAD_ARG, "batch_padded must be a multiple of 64");
    const size_t row = (size_t)2 << n_qubits;
    split_rows_kernel<<<1184, 256, 0, stream>>>(states, (size_t)batch * row, (size_t)batch_padded * row, b_hi, b_lo);
    return check_launch("split_rows_kernel");
}

extern "C" int qfx_unitary_gemm(const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo,
                                int n_qubits, int batch_padded, int precision, float* out_states, float* expvals,
                                cudaStream_t stream) {
    QFX_ARG_CHECK(n_qubits >= 6 && n_qubits <= 12, QFX_ERR_UNSUPPORTED, "unitary GEMM supports 6..12 qubits, got %d", n_qubits);
    QFX_ARG_CHECK(batch_padded > 0 && batch_padded % 64 == 0, QFX_ERR_BAD_ARG, "batch must be padded to a multiple of 64");
    QFX_ARG_CHECK(precision == 1 || precision == 3, QFX_ERR_BAD_ARG, "precision must be 1 (TF32) or 3 (3xTF32)");
    QFX_ARG_CHECK(out_states != nullptr || expvals != nullptr, QFX_ERR_BAD_ARG, "nothing to compute");
    if (int e = get_encoder()) return e;
    GemmArgs g;
    g.m_total = 2 << n_qubits; g.k_total = g.m_total; g.n_total = batch_padded;
    g.n_products = precision; g.out = out_states; g.expvals = expvals; g.n_qubits = n_qubits;
    if (expvals) QFX_CUDA_TRY(cudaMemsetAsync(expvals, 0, sizeof(float) * (size_t)batch_padded * n_qubits, stream));
    if (batch_padded % 128 == 0) return launch_gemm<128>(a_hi, a_lo, b_hi, b_lo, g, stream);
    return launch_gemm<64>(a_hi, a_lo, b_hi, b_lo, g, stream);
}
