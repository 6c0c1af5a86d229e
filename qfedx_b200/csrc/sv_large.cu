// K5 / K6: large statevectors (n > 10 qubits) in global memory.
//
// K5  sv_large_pass_kernel   one compiled pass (csrc/circuit_passes.cu): a CTA gathers the 2^T
//     amplitudes of a tile (T <= 12, 32 KB) into shared memory, applies every gate of the pass
//     there and scatters them back - several gates per trip through L2/HBM.  grid.y batches
//     independent states (clients, or the 2P shifted circuits of the parameter-shift rule, each
//     with its own angle shift), so a 20-qubit state (8 MiB, L2-resident on B200) is re-read from
//     L2 rather than HBM between passes.
// K3' sv_large_expz_masks_kernel   <Z-string> readouts (<Z_q> = one-bit masks), the shard's rank bits as high index bits.
// K6  sv_global_swap_fused_kernel   sharded statevector: the top g "global" qubits are the rank
//     index.  Exchanging them with the top g local qubits is an all-to-all of contiguous slices;
//     each thread PULLS its 2^g amplitudes from the 2^g peers with plain loads over NVLink (peer
//     pointers from the symmetric heap, no NCCL), applies the gates that follow on those qubits in
//     registers, and writes the result locally - the swap and the next gates cost one pass.
//
// No reference counterpart (the reference never simulates more than one 4-qubit state,
// /root/reference/src/QFed/qAmplitude.py:44-46); configs 4 and 5 of BASELINE.json.
#include "common.cuh"
#include "../../include/qfedx_b200.h"

namespace qfx {
namespace large {

constexpr int MAX_TILE_BITS = 12;
constexpr int MAX_OPS_PER_PASS = 160;
constexpr int PASS_THREADS = 256;

struct PassArgs {
    float2* states;            // [batch][2^n]
    int n_qubits, tile_bits, low_bits;
    int tile_q[MAX_TILE_BITS]; // qubit of each tile bit, ascending
    const int* ops;            // [n_ops][5] (kind, local bit 0, local bit 1, param 0, param 1)
    int n_ops;
    const float* theta;        // [theta_rows][n_params]; row = batch index % theta_rows
    int theta_rows, n_params;
    const float* angles;       // [angle_rows][n_feature]; feature-map angles (QFX_GATE_RY_FEATURE)
    int angle_rows, n_feature;
    const int* shift_param;    // [batch] parameter index to shift, -1 none (nullable)
    const float* shift_value;  // [batch] shift in radians
    int init_zero_state;       // 1: ignore memory contents, start from |0..0>
    // the qubits OUTSIDE the tile as ascending runs of consecutive positions (filled by the launcher): the CTA
    // index is deposited into them run by run instead of bit by bit
    int n_runs, run_pos[MAX_TILE_BITS + 1], run_width[MAX_TILE_BITS + 1];
};

__device__ __forceinline__ unsigned long long tile_base(const PassArgs& a, unsigned long long t) {
    unsigned long long base = 0ull;
    for (int r = 0; r < a.n_runs; ++r) {
        base |= (t & ((1ull << a.run_width[r]) - 1ull)) << a.run_pos[r];
        t >>= a.run_width[r];
    }
    return base;
}

__device__ __forceinline__ void cmul_acc(float2& out, float2 m, float2 v) {   // two FFMA per component
    out.x = fmaf(-m.y, v.y, fmaf(m.x, v.x, out.x));
    out.y = fmaf(m.y, v.x, fmaf(m.x, v.y, out.y));
}
// m0 * v0 + m1 * v1 without a zero accumulator: FMUL + 3 FFMA per component
__device__ __forceinline__ float2 cmul2(float2 m0, float2 v0, float2 m1, float2 v1) {
    float2 r;
    r.x = fmaf(-m1.y, v1.y, fmaf(m1.x, v1.x, fmaf(-m0.y, v0.y, m0.x * v0.x)));
    r.y = fmaf(m1.y, v1.x, fmaf(m1.x, v1.y, fmaf(m0.y, v0.x, m0.x * v0.y)));
    return r;
}

// Global offset of tile element j0 = tid + 256 k: the deposit of j0's bits into the tile's qubit positions
// splits into a per-thread part (the 8 bits of tid, computed once) and a per-k part (16 table entries in
// shared memory), so the gather / scatter loops cost two ORs per element instead of a 64-bit bit loop.
__device__ __forceinline__ unsigned long long deposit_bits(unsigned v, int first_bit, int n_bits, const int* tile_q) {
    unsigned long long off = 0ull;
    for (int b = 0; b < n_bits; ++b) off |= (unsigned long long)((v >> b) & 1u) << tile_q[first_bit + b];
    return off;
}

// 2x2 complex matrix (row-major m00 m01 m10 m11) of one op for this batch element
__device__ __forceinline__ void op_matrix(const PassArgs& a, const int* op, int batch, float2* m) {
    const int kind = op[0];
    const float* th = a.theta ? a.theta + (size_t)(batch % a.theta_rows) * a.n_params : nullptr;
    auto param = [&](int p) -> float {
        float v = th[p];
        if (a.shift_param && a.shift_param[batch] == p) v += a.shift_value[batch];
        return v;
    };
    float s, c;
    if (kind == QFX_GATE_RX) {
        sincosf(0.5f * param(op[3]), &s, &c);
        m[0] = make_float2(c, 0.f); m[1] = make_float2(0.f, -s); m[2] = make_float2(0.f, -s); m[3] = make_float2(c, 0.f);
    } else if (kind == QFX_GATE_RY) {
        sincosf(0.5f * param(op[3]), &s, &c);
        m[0] = make_float2(c, 0.f); m[1] = make_float2(-s, 0.f); m[2] = make_float2(s, 0.f); m[3] = make_float2(c, 0.f);
    } else if (kind == QFX_GATE_RY_FEATURE) {
        sincosf(0.5f * a.angles[(size_t)(batch % a.angle_rows) * a.n_feature + op[3]], &s, &c);
        m[0] = make_float2(c, 0.f); m[1] = make_float2(-s, 0.f); m[2] = make_float2(s, 0.f); m[3] = make_float2(c, 0.f);
    } else if (kind == QFX_GATE_RZ) {
        sincosf(0.5f * param(op[3]), &s, &c);
        m[0] = make_float2(c, -s); m[1] = make_float2(0.f, 0.f); m[2] = make_float2(0.f, 0.f); m[3] = make_float2(c, s);
    } else if (kind == QFX_GATE_RXRZ) {  // RZ(phi) * RX(theta)
        float sp, cp;
        sincosf(0.5f * param(op[3]), &s, &c);
        sincosf(0.5f * param(op[4]), &sp, &cp);
        m[0] = make_float2(cp * c, -sp * c);   // e^{-i phi/2} c
        m[1] = make_float2(-sp * s, -cp * s);  // -i e^{-i phi/2} s
        m[2] = make_float2(sp * s, -cp * s);   // -i e^{+i phi/2} s
        m[3] = make_float2(cp * c, sp * c);    // e^{+i phi/2} c
    } else {
        m[0] = make_float2(1.f, 0.f); m[1] = make_float2(0.f, 0.f); m[2] = make_float2(0.f, 0.f); m[3] = make_float2(1.f, 0.f);
    }
}

__global__ void __launch_bounds__(PASS_THREADS)
sv_large_pass_kernel(const PassArgs a) {
    extern __shared__ __align__(16) float2 tile[];        // [2^T]
    __shared__ float2 mats[MAX_OPS_PER_PASS][4];
    __shared__ int sops[MAX_OPS_PER_PASS][3];
    const int tid = threadIdx.x, batch = blockIdx.y;
    const int T = a.tile_bits, tile_n = 1 << T;
    float2* st = a.states + ((size_t)batch << a.n_qubits);

    for (int i = tid; i < a.n_ops; i += PASS_THREADS) {
        const int* op = a.ops + 5 * i;
        sops[i][0] = op[0]; sops[i][1] = op[1]; sops[i][2] = op[2];
        op_matrix(a, op, batch, mats[i]);
    }
    // base offset of this tile: spread blockIdx.x over the non-tile bit positions
    const unsigned long long base = tile_base(a, blockIdx.x);
    // gather: the low `low_bits` tile bits are qubits 0..low-1, so 2^low consecutive amplitudes move together
    const int low = a.low_bits, chunk = 1 << low;
    for (int j0 = tid * 1; j0 < tile_n; j0 += PASS_THREADS) {
        const int jl = j0 & (chunk - 1), jh = j0 >> low;
        unsigned long long off = base | (unsigned long long)jl;
        for (int b = low; b < T; ++b) off |= (unsigned long long)((jh >> (b - low)) & 1) << a.tile_q[b];
        float2 v;
        if (a.init_zero_state) v = make_float2((off == 0ull) ? 1.f : 0.f, 0.f);
        else v = st[off];
        tile[j0] = v;
    }
    __syncthreads();

    for (int i = 0; i < a.n_ops; ++i) {
        const int kind = sops[i][0], b0 = sops[i][1], b1 = sops[i][2];
        if (kind == QFX_GATE_CNOT) {
            for (int idx = tid; idx < (tile_n >> 2); idx += PASS_THREADS) {
                // insert zeros at bit positions lo < hi, then set control, toggle target
                const int lo = min(b0, b1), hi = max(b0, b1);
                int j = idx;
                j = ((j >> lo) << (lo + 1)) | (j & ((1 << lo) - 1));
                j = ((j >> hi) << (hi + 1)) | (j & ((1 << hi) - 1));
                const int ja = j | (1 << b0), jb = ja | (1 << b1);
                const float2 t = tile[ja]; tile[ja] = tile[jb]; tile[jb] = t;
            }
        } else if (kind == QFX_GATE_CZ) {
            for (int idx = tid; idx < (tile_n >> 2); idx += PASS_THREADS) {
                const int lo = min(b0, b1), hi = max(b0, b1);
                int j = idx;
                j = ((j >> lo) << (lo + 1)) | (j & ((1 << lo) - 1));
                j = ((j >> hi) << (hi + 1)) | (j & ((1 << hi) - 1));
                j |= (1 << b0) | (1 << b1);
                tile[j] = make_float2(-tile[j].x, -tile[j].y);
            }
        } else if (kind == QFX_GATE_RZ) {
            const float2 p0 = mats[i][0], p1 = mats[i][3];
            for (int j = tid; j < tile_n; j += PASS_THREADS) {
                const float2 v = tile[j], p = ((j >> b0) & 1) ? p1 : p0;
                tile[j] = make_float2(p.x * v.x - p.y * v.y, p.x * v.y + p.y * v.x);
            }
        } else {
            const float2 m00 = mats[i][0], m01 = mats[i][1], m10 = mats[i][2], m11 = mats[i][3];
            for (int idx = tid; idx < (tile_n >> 1); idx += PASS_THREADS) {
                const int j0 = ((idx >> b0) << (b0 + 1)) | (idx & ((1 << b0) - 1)), j1 = j0 | (1 << b0);
                const float2 v0 = tile[j0], v1 = tile[j1];
                tile[j0] = cmul2(m00, v0, m01, v1);
                tile[j1] = cmul2(m10, v0, m11, v1);
            }
        }
        __syncthreads();
    }

    for (int j0 = tid; j0 < tile_n; j0 += PASS_THREADS) {
        const int jl = j0 & (chunk - 1), jh = j0 >> low;
        unsigned long long off = base | (unsigned long long)jl;
        for (int b = low; b < T; ++b) off |= (unsigned long long)((jh >> (b - low)) & 1) << a.tile_q[b];
        st[off] = tile[j0];
    }
}

// ---------------------------------------------------------------------------------------------
// Pass kernel v2 (tile of exactly 12 qubits): the tile lives in REGISTERS, 16 amplitudes per thread.
// The 12 tile bits are assigned to positions: 5 lane bits (gates exchange amplitudes with shuffles),
// 4 register bits (gates are pure register arithmetic; static indexing through a switch over the bit),
// 3 warp bits (inactive).  A pass is split by the host planner (qfedx_b200/ops/large.py) into segments,
// each with its own assignment such that every gate of the segment only touches lane / register bits;
// between segments the tile is re-laid-out through shared memory.  The planner puts one-qubit gates and CNOT
// targets on register bits only (a lane-bit gate costs two shuffles per amplitude on top of the arithmetic);
// the shared-memory copy of the tile is XOR-swizzled so that any lane set covering the four bit classes
// mod 4 is bank-conflict free.  Real rotations take 4 instructions per amplitude, fused RX.RZ 6 (scalar
// phases collected into one carrier gate), a CNOT 2 - against ~40 for the shared-memory loop of v1.
constexpr int V2_T = 12;

template <int RB>
__device__ __forceinline__ void v2_gate_reg(float2 (&a)[16], const float2* m) {
    const float2 m00 = m[0], m01 = m[1], m10 = m[2], m11 = m[3];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if (!((i >> RB) & 1)) {
            const int j = i | (1 << RB);
            const float2 v0 = a[i], v1 = a[j];
            a[i] = cmul2(m00, v0, m01, v1);
            a[j] = cmul2(m10, v0, m11, v1);
        }
    }
}
// Real rotation (RY, feature map): v0' = c v0 - s v1, v1' = s v0 + c v1 - 4 instructions per amplitude
template <int RB>
__device__ __forceinline__ void v2_ry_reg(float2 (&a)[16], float c, float s) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if (!((i >> RB) & 1)) {
            const int j = i | (1 << RB);
            const float2 v0 = a[i], v1 = a[j];
            a[i] = make_float2(fmaf(-s, v1.x, c * v0.x), fmaf(-s, v1.y, c * v0.y));
            a[j] = make_float2(fmaf(s, v0.x, c * v1.x), fmaf(s, v0.y, c * v1.y));
        }
    }
}
__device__ __forceinline__ void v2_ry_lane(float2 (&a)[16], float c, float s, int lb, int lane) {
    const float so = ((lane >> lb) & 1) ? s : -s;    // partner coefficient: +s on the |1> side
    const int mask = 1 << lb;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const float px = __shfl_xor_sync(0xffffffffu, a[i].x, mask);
        const float py = __shfl_xor_sync(0xffffffffu, a[i].y, mask);
        a[i] = make_float2(fmaf(so, px, c * a[i].x), fmaf(so, py, c * a[i].y));
    }
}
// RZ(phi) RX(theta) = e^{-i phi/2} diag(1, e^{i phi}) RX(theta): the pair update without the scalar phase costs
// 12 instructions instead of 16; the scalar phases of all such gates of a pass are folded into one carrier
// gate's matrix (see the kernel preamble), so the state is exact.
template <int RB>
__device__ __forceinline__ void v2_rxrz_reg(float2 (&a)[16], const float2* m) {
    const float c = m[0].x, s = m[0].y, cp = m[1].x, sp = m[1].y;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if (!((i >> RB) & 1)) {
            const int j = i | (1 << RB);
            const float2 v0 = a[i], v1 = a[j];
            a[i] = make_float2(fmaf(s, v1.y, c * v0.x), fmaf(-s, v1.x, c * v0.y));      // c v0 - i s v1
            const float tx = fmaf(s, v0.y, c * v1.x), ty = fmaf(-s, v0.x, c * v1.y);    // c v1 - i s v0
            a[j] = make_float2(fmaf(-sp, ty, cp * tx), fmaf(sp, tx, cp * ty));          // e^{i phi} t
        }
    }
}
__device__ __forceinline__ void v2_gate_lane(float2 (&a)[16], const float2* m, int lb, int lane) {
    const bool hi = (lane >> lb) & 1;
    const float2 ms = hi ? m[3] : m[0], mo = hi ? m[2] : m[1];   // own / partner coefficient
    const int mask = 1 << lb;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        float2 p;
        p.x = __shfl_xor_sync(0xffffffffu, a[i].x, mask);
        p.y = __shfl_xor_sync(0xffffffffu, a[i].y, mask);
        a[i] = cmul2(ms, a[i], mo, p);
    }
}
// 16-bit mask: bit i set when position `pos` (lane 0..4 / register 5..8) is 1 for register i of this thread
__device__ __forceinline__ unsigned v2_mask(int pos, int lane) {
    if (pos < 5) return ((lane >> pos) & 1) ? 0xFFFFu : 0u;
    const unsigned pat[4] = {0xAAAAu, 0xCCCCu, 0xF0F0u, 0xFF00u};
    return pat[pos - 5];
}
__device__ __forceinline__ int v2_swz(int j) { return j ^ ((j >> 4) & 15) ^ ((j >> 8) & 15); }

// CNOT with the target on register bit RB: control on a lane bit -> one predicate per thread ...
template <int RB>
__device__ __forceinline__ void v2_cnot_lane_ctl(float2 (&a)[16], bool on) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if (!((i >> RB) & 1)) {
            const int j = i | (1 << RB);
            const float2 vi = a[i], vj = a[j];
            a[i] = on ? vj : vi;
            a[j] = on ? vi : vj;
        }
    }
}
// ... control on register bit CB -> a fixed exchange of four register pairs
template <int CB, int RB>
__device__ __forceinline__ void v2_cnot_reg_ctl(float2 (&a)[16]) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        if (((i >> CB) & 1) && !((i >> RB) & 1)) {
            const int j = i | (1 << RB);
            const float2 t = a[i]; a[i] = a[j]; a[j] = t;
        }
    }
}
template <int CB>
__device__ __forceinline__ void v2_cnot_reg_ctl_any(float2 (&a)[16], int rb) {
    switch (rb) {
        case 0: if (CB != 0) v2_cnot_reg_ctl<CB, CB == 0 ? 1 : 0>(a); break;
        case 1: if (CB != 1) v2_cnot_reg_ctl<CB, CB == 1 ? 0 : 1>(a); break;
        case 2: if (CB != 2) v2_cnot_reg_ctl<CB, CB == 2 ? 0 : 2>(a); break;
        default: if (CB != 3) v2_cnot_reg_ctl<CB, CB == 3 ? 0 : 3>(a); break;
    }
}

struct PassArgsV2 {
    PassArgs base;
    const int* plan;   // [n_seg] then per segment: pos_to_bit[12], n_ops, n_ops x (kind, posA, posB, op index)
    int plan_len;
};

// ints of plan kept in shared memory: header, segments, and the optional trailer (per-op mode + carrier index)
constexpr int V2_PLAN_MAX = 1 + 4 * (13 + 4 * MAX_OPS_PER_PASS / 2) + MAX_OPS_PER_PASS + 1;
constexpr int V2_KIND_RXRZ_STRIPPED = 7;   // plan-private op kind: RXRZ on a register bit without its scalar phase

// The kernel is persistent over the tiles of one state: a CTA sets up its gate matrices and op stream once, then walks
// tiles blockIdx.x, blockIdx.x + gridDim.x, ...; the NEXT tile is fetched with cp.async into the other half of a
// double-buffered shared-memory tile while the current one is worked on in registers, so global loads, gate math and
// the stores of the previous tile overlap inside one CTA instead of only across the three CTAs of an SM.
__device__ __forceinline__ void cp_async_8(void* smem_dst, const void* gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

__global__ void __launch_bounds__(PASS_THREADS, 3)
sv_large_pass_v2_kernel(const PassArgsV2 pa) {
    const PassArgs& a = pa.base;
    extern __shared__ __align__(16) float2 tiles2[];            // two tiles of 2^V2_T amplitudes
    __shared__ float2 mats[MAX_OPS_PER_PASS][4];
    __shared__ short splan[V2_PLAN_MAX];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, batch = blockIdx.y;
    float2* st = a.states + ((size_t)batch << a.n_qubits);
    const unsigned n_tiles = 1u << (a.n_qubits - V2_T);
    __shared__ unsigned long long koff[16];
    if (tid < 16) koff[tid] = deposit_bits((unsigned)tid, 8, V2_T - 8, a.tile_q);
    const unsigned long long tlow = deposit_bits((unsigned)tid, 0, 8, a.tile_q);
    __syncthreads();
    // first tile on its way before the set-up below touches global memory
    if (!a.init_zero_state && blockIdx.x < n_tiles) {
        const unsigned long long toff = tile_base(a, blockIdx.x) | tlow;
#pragma unroll
        for (int k = 0; k < 16; ++k) cp_async_8(&tiles2[v2_swz(tid + 256 * k)], st + (toff | koff[k]));
    }
    for (int i = tid; i < pa.plan_len; i += PASS_THREADS) splan[i] = (short)pa.plan[i];   // op stream out of global memory
    // plan[0] = n_seg | has_trailer << 16; trailer = n_ops op modes (0 general, 1 stripped RXRZ, 2 carrier) + carrier
    const int plan0 = pa.plan[0];
    const bool has_modes = (plan0 >> 16) & 1;
    const int* gmodes = pa.plan + pa.plan_len - a.n_ops - 1;
    for (int i = tid; i < a.n_ops; i += PASS_THREADS) {
        if (has_modes && gmodes[i] == 1) {
            const int* op = a.ops + 5 * i;
            const float* th = a.theta + (size_t)(batch % a.theta_rows) * a.n_params;
            float t0 = th[op[3]], t1 = th[op[4]];
            if (a.shift_param && a.shift_param[batch] == op[3]) t0 += a.shift_value[batch];
            if (a.shift_param && a.shift_param[batch] == op[4]) t1 += a.shift_value[batch];
            float sn, cs, sp, cp;
            sincosf(0.5f * t0, &sn, &cs);
            sincosf(t1, &sp, &cp);
            mats[i][0] = make_float2(cs, sn); mats[i][1] = make_float2(cp, sp);
            mats[i][2] = make_float2(0.5f * t1, 0.f); mats[i][3] = make_float2(0.f, 0.f);
        } else {
            op_matrix(a, a.ops + 5 * i, batch, mats[i]);
        }
    }
    __syncthreads();
    if (has_modes && tid < 32) {   // scalar phases of the stripped gates -> the carrier's matrix
        const int carrier = splan[pa.plan_len - 1];
        float ph = 0.f;
        for (int i = tid; i < a.n_ops; i += 32)
            if (splan[pa.plan_len - a.n_ops - 1 + i] == 1) ph -= mats[i][2].x;
        ph = warp_sum(ph);
        if (carrier >= 0 && tid < 4) {
            float sg, cg;
            sincosf(ph, &sg, &cg);
            const float2 v = mats[carrier][tid];
            mats[carrier][tid] = make_float2(cg * v.x - sg * v.y, cg * v.y + sg * v.x);
        }
    }
    // tile element j lives at v2_swz(j): low nibble XOR-ed with the two higher nibbles (linear over GF(2), so the
    // address of a union of disjoint bit sets is the XOR of their addresses).  Any segment whose lane bits cover
    // the four classes mod 4 then reads and writes its layout without bank conflicts.
    int parity = 0;
    for (unsigned t = blockIdx.x; t < n_tiles; t += gridDim.x, parity ^= 1) {
    float2* tile = tiles2 + (parity << V2_T);
    const unsigned long long toff = tile_base(a, t) | tlow;
    cp_async_commit_wait_all();      // this thread's pieces of tile t have landed
    __syncthreads();                 // ... everybody's; and every thread is done reading the other buffer (stores of t - 1)
    if (a.init_zero_state) {
#pragma unroll
        for (int k = 0; k < 16; ++k) tile[v2_swz(tid + 256 * k)] = make_float2(((toff | koff[k]) == 0ull) ? 1.f : 0.f, 0.f);
        __syncthreads();
    } else if (t + gridDim.x < n_tiles) {
        float2* nxt = tiles2 + ((parity ^ 1) << V2_T);
        const unsigned long long noff = tile_base(a, t + gridDim.x) | tlow;
#pragma unroll
        for (int k = 0; k < 16; ++k) cp_async_8(&nxt[v2_swz(tid + 256 * k)], st + (noff | koff[k]));
    }

    const short* plan = splan;
    const int n_seg = plan0 & 0xFFFF;
    int cur = 1;
    for (int sg = 0; sg < n_seg; ++sg) {
        const short* p2b = plan + cur;          // position -> tile bit
        const int n_ops = plan[cur + 12];
        const short* ops = plan + cur + 13;
        cur += 13 + 4 * n_ops;
        // tile index of (warp, lane, register r)
        int bwl = 0;
#pragma unroll
        for (int p = 0; p < 5; ++p) bwl |= ((lane >> p) & 1) << p2b[p];
#pragma unroll
        for (int p = 0; p < 3; ++p) bwl |= ((warp >> p) & 1) << p2b[9 + p];
        bwl = v2_swz(bwl);
        const int rb0 = v2_swz(1 << p2b[5]), rb1 = v2_swz(1 << p2b[6]), rb2 = v2_swz(1 << p2b[7]),
                  rb3 = v2_swz(1 << p2b[8]);
        float2 r[16];
#pragma unroll
        for (int i = 0; i < 16; ++i)
            r[i] = tile[bwl ^ ((i & 1) ? rb0 : 0) ^ ((i & 2) ? rb1 : 0) ^ ((i & 4) ? rb2 : 0) ^ ((i & 8) ? rb3 : 0)];
        for (int o = 0; o < n_ops; ++o) {
            const int kind = ops[4 * o], pA = ops[4 * o + 1], pB = ops[4 * o + 2];
            const float2* m = mats[ops[4 * o + 3]];
            if (kind == QFX_GATE_CNOT) {
                const unsigned cm = v2_mask(pA, lane);
                if (pB >= 5 && pA < 5) {
                    const bool on = (lane >> pA) & 1;
                    switch (pB - 5) {
                        case 0: v2_cnot_lane_ctl<0>(r, on); break;
                        case 1: v2_cnot_lane_ctl<1>(r, on); break;
                        case 2: v2_cnot_lane_ctl<2>(r, on); break;
                        default: v2_cnot_lane_ctl<3>(r, on); break;
                    }
                } else if (pB >= 5) {
                    switch (pA - 5) {
                        case 0: v2_cnot_reg_ctl_any<0>(r, pB - 5); break;
                        case 1: v2_cnot_reg_ctl_any<1>(r, pB - 5); break;
                        case 2: v2_cnot_reg_ctl_any<2>(r, pB - 5); break;
                        default: v2_cnot_reg_ctl_any<3>(r, pB - 5); break;
                    }
                } else {
                    const int mask = 1 << pB;
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const float px = __shfl_xor_sync(0xffffffffu, r[i].x, mask);
                        const float py = __shfl_xor_sync(0xffffffffu, r[i].y, mask);
                        if ((cm >> i) & 1) r[i] = make_float2(px, py);
                    }
                }
            } else if (kind == QFX_GATE_CZ) {
                const unsigned cm = v2_mask(pA, lane) & v2_mask(pB, lane);
#pragma unroll
                for (int i = 0; i < 16; ++i)
                    if ((cm >> i) & 1) r[i] = make_float2(-r[i].x, -r[i].y);
            } else if (kind == QFX_GATE_RY || kind == QFX_GATE_RY_FEATURE) {
                const float c = m[0].x, sn = m[2].x;
                if (pA >= 5) {
                    switch (pA - 5) {
                        case 0: v2_ry_reg<0>(r, c, sn); break;
                        case 1: v2_ry_reg<1>(r, c, sn); break;
                        case 2: v2_ry_reg<2>(r, c, sn); break;
                        default: v2_ry_reg<3>(r, c, sn); break;
                    }
                } else {
                    v2_ry_lane(r, c, sn, pA, lane);
                }
            } else if (kind == V2_KIND_RXRZ_STRIPPED) {
                switch (pA - 5) {
                    case 0: v2_rxrz_reg<0>(r, m); break;
                    case 1: v2_rxrz_reg<1>(r, m); break;
                    case 2: v2_rxrz_reg<2>(r, m); break;
                    default: v2_rxrz_reg<3>(r, m); break;
                }
            } else if (pA >= 5) {
                switch (pA - 5) {
                    case 0: v2_gate_reg<0>(r, m); break;
                    case 1: v2_gate_reg<1>(r, m); break;
                    case 2: v2_gate_reg<2>(r, m); break;
                    default: v2_gate_reg<3>(r, m); break;
                }
            } else {
                v2_gate_lane(r, m, pA, lane);
            }
        }
        __syncthreads();   // every thread has read its amplitudes of this layout
#pragma unroll
        for (int i = 0; i < 16; ++i)
            tile[bwl ^ ((i & 1) ? rb0 : 0) ^ ((i & 2) ? rb1 : 0) ^ ((i & 4) ? rb2 : 0) ^ ((i & 8) ? rb3 : 0)] = r[i];
        __syncthreads();
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) st[toff | koff[k]] = tile[v2_swz(tid + 256 * k)];
    }   // tiles of this CTA
}

// <Z-string> readout: out[q] = sum_i |psi_i|^2 (-1)^{parity(idx & masks[q])} with idx = high_bits << n_local | i.
// A trailing CNOT ring only permutes basis states, so <Z_q> after it equals a parity readout before it
// (Z_t -> Z_c Z_t under CNOT(c,t)): the ring's passes - and for a sharded state its swaps - are never run.
struct MaskArgs { unsigned long long m[32]; };
// The n_out parities of one basis index form a GF(2)-linear map of its bits: signs(idx) = XOR over the bytes of idx of
// a 256-entry table (bit q of an entry = parity(byte & byte of mask q)).  A thread's low index byte never changes,
// so per amplitude the sign word costs one shared-memory lookup per remaining byte instead of n_out 64-bit
// popcounts (which made the kernel XU-bound at 75 % before it was anywhere near the memory system).
__global__ void __launch_bounds__(256, 4)
sv_large_expz_masks_kernel(const float2* __restrict__ states, int n_local, int n_out, unsigned int high_bits,
                           const MaskArgs masks, float* __restrict__ out /*[batch][n_out]*/) {
    __shared__ float red[32];
    __shared__ unsigned int tab[4][256];                       // index bits 11..18, 19..26, 27..34, 35..42
    __shared__ unsigned int col[64];                           // col[b]: bit q set iff mask q contains index bit b
    __shared__ float sig[3][32];                               // +-1 per output for index bits 0, 9 and 10
    const int batch = blockIdx.y, tid = threadIdx.x;
    const float4* st = reinterpret_cast<const float4*>(states + ((size_t)batch << n_local));   // two amplitudes per load
    const size_t N2 = (size_t)1 << (n_local - 1);
    const unsigned long long hi = (unsigned long long)high_bits << n_local;
    if (tid < 64) {
        unsigned int w = 0u;
        for (int q = 0; q < n_out; ++q) w |= (unsigned)((masks.m[q] >> tid) & 1ull) << q;
        col[tid] = w;
    }
    if (tid < 96) {
        const int which = tid >> 5, q = tid & 31, bit = which == 0 ? 0 : 8 + which;
        sig[which][q] = (q < n_out && ((masks.m[q] >> bit) & 1ull)) ? -1.f : 1.f;
    }
    __syncthreads();
    for (int e = tid; e < 4 * 256; e += 256) {
        const int b0 = 11 + 8 * (e >> 8);
        unsigned int w = 0u;
#pragma unroll
        for (int b = 0; b < 8; ++b)
            if ((e >> b) & 1) w ^= col[b0 + b];
        tab[e >> 8][e & 255] = w;
    }
    // A thread's trip covers the eight amplitudes that differ in index bits 0 (the two halves of a 16-byte load), 9
    // and 10 (four loads 256 pairs apart): their |psi|^2 are folded with the outputs' +-1 for those three bits (seven
    // FMAs per output instead of eight sign flips and adds), and the sign of everything else - the thread's own bits
    // 1..8, the table bytes of the bits above 10, and the bits of `hi` the tables do not reach - is applied once.
    unsigned int s_fixed = 0u;
    for (unsigned long long fixed = (unsigned long long)(2 * tid) | (hi & 0x7ffull) | (hi & ~((1ull << 43) - 1ull)); fixed;
         fixed &= fixed - 1ull)
        s_fixed ^= col[__ffsll((long long)fixed) - 1];
    if (tid < 32) red[tid] = 0.f;
    __syncthreads();
    float acc[32];
#pragma unroll
    for (int q = 0; q < 32; ++q) acc[q] = 0.f;
    const size_t n_groups = (N2 + 1023) >> 10;                  // groups of 4 x 256 pairs (index bits 0..10)
    for (size_t grp = blockIdx.x; grp < n_groups; grp += gridDim.x) {
        float pa[4], pb[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const size_t i = (grp << 10) + (size_t)u * 256 + tid;
            const float4 v = i < N2 ? __ldcs(st + i) : make_float4(0.f, 0.f, 0.f, 0.f);
            pa[u] = v.x * v.x + v.y * v.y;
            pb[u] = v.z * v.z + v.w * v.w;
        }
        const unsigned long long up = (hi | (grp << 11)) >> 11;
        const unsigned int sg = s_fixed ^ tab[0][up & 255] ^ tab[1][(up >> 8) & 255] ^ tab[2][(up >> 16) & 255] ^
                                tab[3][(up >> 24) & 255];
#pragma unroll
        for (int q = 0; q < 32; ++q) {
            if (q < n_out) {                                     // uniform
                const float s0 = sig[0][q], s9 = sig[1][q], s10 = sig[2][q];
                const float a0 = fmaf(s0, pb[0], pa[0]), a1 = fmaf(s0, pb[1], pa[1]);
                const float a2 = fmaf(s0, pb[2], pa[2]), a3 = fmaf(s0, pb[3], pa[3]);
                const float r = fmaf(s10, fmaf(s9, a3, a2), fmaf(s9, a1, a0));
                acc[q] += __uint_as_float(__float_as_uint(r) ^ ((sg << (31 - q)) & 0x80000000u));
            }
        }
    }
#pragma unroll
    for (int q = 0; q < 32; ++q) {
        if (q < n_out) {
            const float v = warp_sum(acc[q]);
            if ((tid & 31) == 0) atomicAdd(&red[q], v);
        }
    }
    __syncthreads();
    if (tid < n_out) atomicAdd(out + (size_t)batch * n_out + tid, red[tid]);
}

// ------------------------------------------------------------------------------------------
struct SwapArgs {
    const float2* const* src;  // [world] peer pointers to every rank's current shard
    float2* dst;               // local destination shard
    int n_local, g_bits, rank, world;
    const int* ops;            // fused gates on the swapped-in qubits: local bit = 0..g-1 (slice index bits)
    int n_ops;
    const float* theta; int n_params;
    const float* angles; int n_feature;   // feature-map angles of the (single) sample
    unsigned int* const* flags;  // [world] peer pointers to flag arrays [world]
    unsigned int epoch;
    unsigned int* counter;       // local u32 scratch for the grid-wide arrival count (zero before launch)
};

__device__ __forceinline__ void st_release_sys_u32(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys_u32(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

template <int G>
__global__ void __launch_bounds__(256)
sv_global_swap_fused_kernel(const SwapArgs a) {
    constexpr int R = 1 << G;
    __shared__ float2 mats[32][4];
    __shared__ int sops[32][3];
    const int tid = threadIdx.x;
    // every rank's source shard must be complete before anyone reads it: one flag per (dst, src)
    if (a.flags != nullptr) {
        if (blockIdx.x == 0 && tid < a.world) st_release_sys_u32(a.flags[tid] + a.rank, a.epoch);
        if (tid < a.world) {
            const unsigned int* f = a.flags[a.rank] + tid;
            while ((int)(ld_acquire_sys_u32(f) - a.epoch) < 0) { __nanosleep(32); }
        }
    }
    for (int i = tid; i < a.n_ops; i += 256) {
        const int* op = a.ops + 5 * i;
        sops[i][0] = op[0]; sops[i][1] = op[1]; sops[i][2] = op[2];
        PassArgs pa = {};
        pa.theta = a.theta; pa.theta_rows = 1; pa.n_params = a.n_params;
        pa.angles = a.angles; pa.angle_rows = 1; pa.n_feature = a.n_feature;
        op_matrix(pa, op, 0, mats[i]);
    }
    __syncthreads();
    const int slice_bits = a.n_local - G;
    const size_t slice = (size_t)1 << slice_bits;
    const float2* srcs[R];
#pragma unroll
    for (int s = 0; s < R; ++s) srcs[s] = a.src[s] + ((size_t)a.rank << slice_bits);
    for (size_t j = ((size_t)blockIdx.x * 256 + tid) * 2; j < slice; j += (size_t)gridDim.x * 512) {
        float4 v[R];  // two consecutive amplitudes from every peer (16-byte loads over NVLink)
#pragma unroll
        for (int s = 0; s < R; ++s) v[s] = *reinterpret_cast<const float4*>(srcs[s] + j);
        for (int i = 0; i < a.n_ops; ++i) {
            const int kind = sops[i][0], b0 = sops[i][1], b1 = sops[i][2];
            if (kind == QFX_GATE_CNOT) {
#pragma unroll
                for (int s = 0; s < R; ++s)
                    if (((s >> b0) & 1) && !((s >> b1) & 1)) {
                        const int t = s | (1 << b1);
                        if (t < R) { const float4 tmp = v[s]; v[s] = v[t]; v[t] = tmp; }
                    }
            } else if (kind == QFX_GATE_CZ) {
#pragma unroll
                for (int s = 0; s < R; ++s)
                    if (((s >> b0) & 1) && ((s >> b1) & 1)) v[s] = make_float4(-v[s].x, -v[s].y, -v[s].z, -v[s].w);
            } else {
                const float2 m00 = mats[i][0], m01 = mats[i][1], m10 = mats[i][2], m11 = mats[i][3];
#pragma unroll
                for (int s = 0; s < R; ++s)
                    if (!((s >> b0) & 1)) {
                        const int t = s | (1 << b0);
                        if (t < R) {
                            const float4 x0 = v[s], x1 = v[t];
                            float2 r0 = make_float2(0.f, 0.f), r1 = r0, r2 = r0, r3 = r0;
                            cmul_acc(r0, m00, make_float2(x0.x, x0.y)); cmul_acc(r0, m01, make_float2(x1.x, x1.y));
                            cmul_acc(r1, m10, make_float2(x0.x, x0.y)); cmul_acc(r1, m11, make_float2(x1.x, x1.y));
                            cmul_acc(r2, m00, make_float2(x0.z, x0.w)); cmul_acc(r2, m01, make_float2(x1.z, x1.w));
                            cmul_acc(r3, m10, make_float2(x0.z, x0.w)); cmul_acc(r3, m11, make_float2(x1.z, x1.w));
                            v[s] = make_float4(r0.x, r0.y, r2.x, r2.y);
                            v[t] = make_float4(r1.x, r1.y, r3.x, r3.y);
                        }
                    }
            }
        }
#pragma unroll
        for (int s = 0; s < R; ++s) *reinterpret_cast<float4*>(a.dst + ((size_t)s << slice_bits) + j) = v[s];
    }
}

// ---------------------------------------------------------------------------------------------
// Adjoint differentiation on the pass engine (K4a for states that do not fit in registers; the "adjoint" option of
// /root/reference/ROADMAP.md:23,27 next to the parameter-shift rule).  After the forward passes, lambda = O psi with
// O = sum_q cot_q Z-string_q (diagonal, sv_large_parity_obs_kernel); then the passes are walked BACKWARDS with both
// states in shared memory: at every parametrised gate R(theta) = exp(-i theta P / 2) the gradient is Im <lambda|P|psi>
// taken just after the gate, after which R^dagger is applied to both states.  One forward and one backward sweep
// replace the 2P shifted circuits of the shift rule.
struct ObsArgs { unsigned long long m[32]; };
__global__ void __launch_bounds__(256)
sv_large_parity_obs_kernel(const float2* __restrict__ psi, float2* __restrict__ lam, int n_local, int n_out,
                           const ObsArgs masks, const float* __restrict__ cot /*[batch][n_out]*/) {
    const int batch = blockIdx.y;
    const size_t N = (size_t)1 << n_local;
    const float2* p = psi + ((size_t)batch << n_local);
    float2* l = lam + ((size_t)batch << n_local);
    float c[32];
#pragma unroll
    for (int q = 0; q < 32; ++q) c[q] = q < n_out ? cot[(size_t)batch * n_out + q] : 0.f;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < N; i += (size_t)gridDim.x * 256) {
        float w = 0.f;
#pragma unroll
        for (int q = 0; q < 32; ++q)
            if (q < n_out) w += (__popcll(i & masks.m[q]) & 1) ? -c[q] : c[q];
        const float2 v = p[i];
        l[i] = make_float2(w * v.x, w * v.y);
    }
}

struct AdjArgs {
    PassArgs base;             // states = psi; ops / theta / angles as in the forward pass
    float2* lam;               // [batch][2^n]
    float* grad;               // [theta_rows][n_params], accumulated with atomics
};

__device__ __forceinline__ float block_sum_256(float v, float* red /*[8]*/) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.f;
    if (threadIdx.x < 8) t = red[threadIdx.x];
    if (threadIdx.x < 32) t = group_sum<8>(t);
    return t;   // valid on thread 0
}
// apply the 2x2 matrix m to the pairs (j0, j1 = j0 | 1 << b) of BOTH tiles
__device__ __forceinline__ void apply_pair_both(float2* psi, float2* lam, int tile_n, int b, float2 m00, float2 m01,
                                                float2 m10, float2 m11) {
    for (int idx = threadIdx.x; idx < (tile_n >> 1); idx += PASS_THREADS) {
        const int j0 = ((idx >> b) << (b + 1)) | (idx & ((1 << b) - 1)), j1 = j0 | (1 << b);
        const float2 p0 = psi[j0], p1 = psi[j1], l0 = lam[j0], l1 = lam[j1];
        psi[j0] = cmul2(m00, p0, m01, p1); psi[j1] = cmul2(m10, p0, m11, p1);
        lam[j0] = cmul2(m00, l0, m01, l1); lam[j1] = cmul2(m10, l0, m11, l1);
    }
}
// Im <lambda| P |psi> over this tile for P = X (0), Y (1), Z (2) on tile bit b (partial sum of this thread)
__device__ __forceinline__ float pauli_im(const float2* psi, const float2* lam, int tile_n, int b, int pauli) {
    float acc = 0.f;
    for (int idx = threadIdx.x; idx < (tile_n >> 1); idx += PASS_THREADS) {
        const int j0 = ((idx >> b) << (b + 1)) | (idx & ((1 << b) - 1)), j1 = j0 | (1 << b);
        const float2 p0 = psi[j0], p1 = psi[j1], l0 = lam[j0], l1 = lam[j1];
        // Im(conj(l) * v) = l.x v.y - l.y v.x
        if (pauli == 0) acc += (l0.x * p1.y - l0.y * p1.x) + (l1.x * p0.y - l1.y * p0.x);
        else if (pauli == 1) acc += -(l0.x * p1.x + l0.y * p1.y) + (l1.x * p0.x + l1.y * p0.y);   // Y psi = (-i p1, i p0)
        else acc += (l0.x * p0.y - l0.y * p0.x) - (l1.x * p1.y - l1.y * p1.x);
    }
    return acc;
}

__global__ void __launch_bounds__(PASS_THREADS)
sv_large_adjoint_pass_kernel(const AdjArgs A) {
    const PassArgs& a = A.base;
    extern __shared__ __align__(16) float2 tiles[];        // psi [2^T] | lambda [2^T]
    __shared__ float trig[MAX_OPS_PER_PASS][4];            // cos, sin of param 0 / param 1 (half angles)
    __shared__ int sops[MAX_OPS_PER_PASS][5];
    __shared__ float red[8];
    const int tid = threadIdx.x, batch = blockIdx.y;
    const int T = a.tile_bits, tile_n = 1 << T;
    float2* psi = tiles;
    float2* lam = tiles + tile_n;
    float2* st = a.states + ((size_t)batch << a.n_qubits);
    float2* sl = A.lam + ((size_t)batch << a.n_qubits);
    const float* th = a.theta ? a.theta + (size_t)(batch % a.theta_rows) * a.n_params : nullptr;
    float* grad = A.grad + (size_t)(batch % a.theta_rows) * a.n_params;
    for (int i = tid; i < a.n_ops; i += PASS_THREADS) {
        const int* op = a.ops + 5 * i;
#pragma unroll
        for (int k = 0; k < 5; ++k) sops[i][k] = op[k];
        float s0 = 0.f, c0 = 1.f, s1 = 0.f, c1 = 1.f;
        const int kind = op[0];
        if (kind == QFX_GATE_RX || kind == QFX_GATE_RY || kind == QFX_GATE_RZ || kind == QFX_GATE_RXRZ) sincosf(0.5f * th[op[3]], &s0, &c0);
        if (kind == QFX_GATE_RXRZ) sincosf(0.5f * th[op[4]], &s1, &c1);
        if (kind == QFX_GATE_RY_FEATURE) sincosf(0.5f * a.angles[(size_t)(batch % a.angle_rows) * a.n_feature + op[3]], &s0, &c0);
        trig[i][0] = c0; trig[i][1] = s0; trig[i][2] = c1; trig[i][3] = s1;
    }
    const unsigned long long base = tile_base(a, blockIdx.x);
    const int low = a.low_bits, chunk = 1 << low;
    for (int j0 = tid; j0 < tile_n; j0 += PASS_THREADS) {
        const int jl = j0 & (chunk - 1), jh = j0 >> low;
        unsigned long long off = base | (unsigned long long)jl;
        for (int b = low; b < T; ++b) off |= (unsigned long long)((jh >> (b - low)) & 1) << a.tile_q[b];
        psi[j0] = st[off];
        lam[j0] = sl[off];
    }
    __syncthreads();
    const float2 zero = make_float2(0.f, 0.f);
    for (int i = a.n_ops - 1; i >= 0; --i) {
        const int kind = sops[i][0], b0 = sops[i][1], b1 = sops[i][2];
        const float c = trig[i][0], s = trig[i][1], cp = trig[i][2], sp = trig[i][3];
        if (kind == QFX_GATE_CNOT || kind == QFX_GATE_CZ) {         // self-inverse: the same move on both states
            for (int idx = tid; idx < (tile_n >> 2); idx += PASS_THREADS) {
                const int lo = min(b0, b1), hi = max(b0, b1);
                int j = idx;
                j = ((j >> lo) << (lo + 1)) | (j & ((1 << lo) - 1));
                j = ((j >> hi) << (hi + 1)) | (j & ((1 << hi) - 1));
                if (kind == QFX_GATE_CNOT) {
                    const int ja = j | (1 << b0), jb = ja | (1 << b1);
                    float2 t = psi[ja]; psi[ja] = psi[jb]; psi[jb] = t;
                    t = lam[ja]; lam[ja] = lam[jb]; lam[jb] = t;
                } else {
                    j |= (1 << b0) | (1 << b1);
                    psi[j] = make_float2(-psi[j].x, -psi[j].y);
                    lam[j] = make_float2(-lam[j].x, -lam[j].y);
                }
            }
            __syncthreads();
            continue;
        }
        if (kind == QFX_GATE_RXRZ) {
            // the gate is RZ(phi) RX(theta): walking backwards meets RZ first
            float g = block_sum_256(pauli_im(psi, lam, tile_n, b0, 2), red);
            if (tid == 0) atomicAdd(grad + sops[i][4], g);
            __syncthreads();
            apply_pair_both(psi, lam, tile_n, b0, make_float2(cp, sp), zero, zero, make_float2(cp, -sp));   // RZ(phi)^dagger
            __syncthreads();
            g = block_sum_256(pauli_im(psi, lam, tile_n, b0, 0), red);
            if (tid == 0) atomicAdd(grad + sops[i][3], g);
            __syncthreads();
            apply_pair_both(psi, lam, tile_n, b0, make_float2(c, 0.f), make_float2(0.f, s), make_float2(0.f, s), make_float2(c, 0.f));
            __syncthreads();
            continue;
        }
        if (kind == QFX_GATE_RX || kind == QFX_GATE_RY || kind == QFX_GATE_RZ) {
            const int pauli = kind == QFX_GATE_RX ? 0 : (kind == QFX_GATE_RY ? 1 : 2);
            const float g = block_sum_256(pauli_im(psi, lam, tile_n, b0, pauli), red);
            if (tid == 0) atomicAdd(grad + sops[i][3], g);
            __syncthreads();
        }
        // inverse rotation (RY_FEATURE carries no trainable parameter: it is only undone)
        if (kind == QFX_GATE_RX) apply_pair_both(psi, lam, tile_n, b0, make_float2(c, 0.f), make_float2(0.f, s), make_float2(0.f, s), make_float2(c, 0.f));
        else if (kind == QFX_GATE_RY || kind == QFX_GATE_RY_FEATURE)
            apply_pair_both(psi, lam, tile_n, b0, make_float2(c, 0.f), make_float2(s, 0.f), make_float2(-s, 0.f), make_float2(c, 0.f));
        else if (kind == QFX_GATE_RZ) apply_pair_both(psi, lam, tile_n, b0, make_float2(c, s), zero, zero, make_float2(c, -s));
        __syncthreads();
    }
    for (int j0 = tid; j0 < tile_n; j0 += PASS_THREADS) {
        const int jl = j0 & (chunk - 1), jh = j0 >> low;
        unsigned long long off = base | (unsigned long long)jl;
        for (int b = low; b < T; ++b) off |= (unsigned long long)((jh >> (b - low)) & 1) << a.tile_q[b];
        st[off] = psi[j0];
        sl[off] = lam[j0];
    }
}

static int fill_pass_args(PassArgs& a, float* states, int n_qubits, int tile_bits, int low_bits, const int* tile_qubits_host,
                          const int* ops_dev, int n_ops, const float* theta, int theta_rows, int n_params,
                          const float* angles, int angle_rows, int n_feature) {
    QFX_ARG_CHECK(n_qubits >= 2 && n_qubits <= 34, QFX_ERR_UNSUPPORTED, "n_qubits %d not in 2..34", n_qubits);
    QFX_ARG_CHECK(tile_bits >= 1 && tile_bits <= MAX_TILE_BITS && tile_bits <= n_qubits, QFX_ERR_BAD_ARG, "bad tile_bits %d", tile_bits);
    QFX_ARG_CHECK(low_bits >= 0 && low_bits <= tile_bits, QFX_ERR_BAD_ARG, "bad low_bits");
    QFX_ARG_CHECK(n_ops >= 0 && n_ops <= MAX_OPS_PER_PASS, QFX_ERR_TOO_LARGE, "pass has %d ops (max %d)", n_ops, MAX_OPS_PER_PASS);
    a = PassArgs{};
    a.states = reinterpret_cast<float2*>(states);
    a.n_qubits = n_qubits; a.tile_bits = tile_bits; a.low_bits = low_bits;
    unsigned long long in_tile = 0ull;
    for (int b = 0; b < tile_bits; ++b) { a.tile_q[b] = tile_qubits_host[b]; in_tile |= 1ull << a.tile_q[b]; }
    a.n_runs = 0;
    for (int q = 0; q < n_qubits;) {
        if ((in_tile >> q) & 1ull) { ++q; continue; }
        int w = 0;
        while (q + w < n_qubits && !((in_tile >> (q + w)) & 1ull)) ++w;
        a.run_pos[a.n_runs] = q; a.run_width[a.n_runs] = w; ++a.n_runs;
        q += w;
    }
    a.ops = ops_dev; a.n_ops = n_ops; a.theta = theta; a.theta_rows = theta_rows > 0 ? theta_rows : 1; a.n_params = n_params;
    a.angles = angles; a.angle_rows = angle_rows > 0 ? angle_rows : 1; a.n_feature = n_feature;
    return 0;
}

static int num_sms() { return device_sm_count(); }

}  // namespace large
}  // namespace qfx

using namespace qfx;
using namespace qfx::large;

extern "C" int qfx_sv_large_pass(float* states, int n_qubits, int batch, int tile_bits, int low_bits,
                                 const int* tile_qubits_host, const int* ops_dev, int n_ops, const float* theta,
                                 int theta_rows, int n_params, const float* angles, int angle_rows, int n_feature,
                                 const int* shift_param, const float* shift_value, int init_zero_state,
                                 const int* plan_dev, int plan_len, cudaStream_t stream) {
    QFX_ARG_CHECK(n_qubits >= 2 && n_qubits <= 34, QFX_ERR_UNSUPPORTED, "n_qubits %d not in 2..34", n_qubits);
    QFX_ARG_CHECK(tile_bits >= 1 && tile_bits <= MAX_TILE_BITS && tile_bits <= n_qubits, QFX_ERR_BAD_ARG, "bad tile_bits %d", tile_bits);
    QFX_ARG_CHECK(low_bits >= 0 && low_bits <= tile_bits, QFX_ERR_BAD_ARG, "bad low_bits");
    QFX_ARG_CHECK(n_ops >= 0 && n_ops <= MAX_OPS_PER_PASS, QFX_ERR_TOO_LARGE, "pass has %d ops (max %d)", n_ops, MAX_OPS_PER_PASS);
    QFX_ARG_CHECK(batch >= 1 && batch <= 65535, QFX_ERR_BAD_ARG, "bad batch");
    PassArgs a = {};
    a.states = reinterpret_cast<float2*>(states);
    a.n_qubits = n_qubits; a.tile_bits = tile_bits; a.low_bits = low_bits;
    for (int b = 0; b < tile_bits; ++b) {
        a.tile_q[b] = tile_qubits_host[b];
        QFX_ARG_CHECK(b < low_bits ? a.tile_q[b] == b : (a.tile_q[b] > a.tile_q[b - (b > 0)] || b == 0), QFX_ERR_BAD_ARG,
                      "tile qubits must be ascending and start with 0..low_bits-1");
    }
    {   // runs of consecutive qubits outside the tile (at most tile_bits + 1 of them)
        unsigned long long in_tile = 0ull;
        for (int b = 0; b < tile_bits; ++b) in_tile |= 1ull << a.tile_q[b];
        a.n_runs = 0;
        for (int q = 0; q < n_qubits;) {
            if ((in_tile >> q) & 1ull) { ++q; continue; }
            int w = 0;
            while (q + w < n_qubits && !((in_tile >> (q + w)) & 1ull)) ++w;
            a.run_pos[a.n_runs] = q; a.run_width[a.n_runs] = w; ++a.n_runs;
            q += w;
        }
    }
    a.ops = ops_dev; a.n_ops = n_ops; a.theta = theta; a.theta_rows = theta_rows > 0 ? theta_rows : 1; a.n_params = n_params;
    a.angles = angles; a.angle_rows = angle_rows > 0 ? angle_rows : 1; a.n_feature = n_feature;
    a.shift_param = shift_param; a.shift_value = shift_value; a.init_zero_state = init_zero_state;
    const size_t smem = sizeof(float2) << tile_bits;
    dim3 grid(1u << (n_qubits - tile_bits), batch);
    if (plan_dev != nullptr && tile_bits == V2_T && plan_len <= V2_PLAN_MAX) {   // register-resident engine
        PassArgsV2 pa;
        pa.base = a;
        pa.plan = plan_dev;
        pa.plan_len = plan_len;
        // persistent over tiles: as many CTAs as are resident at once (3 per SM), each walking its share of one
        // state's tiles with the next tile prefetched into the second half of the dynamic shared memory
        constexpr int kTilesSmem = 2 * (int)(sizeof(float2) << V2_T);
        static PerDeviceFlag attr_set;
        if (!attr_set.here()) {
            QFX_CUDA_TRY(cudaFuncSetAttribute((const void*)sv_large_pass_v2_kernel,
                                              cudaFuncAttributeMaxDynamicSharedMemorySize, kTilesSmem));
            attr_set.here() = true;
        }
        // CTAs per state: at least 16 tiles each (the set-up is paid once per CTA), at least one resident wave in
        // total, and - when one state's share of the GPU is more than one CTA - a whole number of waves
        const unsigned resident = (unsigned)num_sms() * 3u;
        const unsigned per = resident / (unsigned)batch > 0u ? resident / (unsigned)batch : 1u;
        unsigned gx = grid.x / 16u > per ? grid.x / 16u : per;
        if ((unsigned)batch <= resident) gx = (gx + per - 1u) / per * per;
        if (gx > grid.x) gx = grid.x;
        sv_large_pass_v2_kernel<<<dim3(gx, batch), PASS_THREADS, kTilesSmem, stream>>>(pa);
        return check_launch("sv_large_pass_v2_kernel");
    }
    sv_large_pass_kernel<<<grid, PASS_THREADS, smem, stream>>>(a);
    return check_launch("sv_large_pass_kernel");
}

static int launch_expz_masks(const float* states, int n_local, int n_out, unsigned int high_bits, const MaskArgs& m,
                             int batch, float* out, cudaStream_t stream) {
    QFX_CUDA_TRY(cudaMemsetAsync(out, 0, sizeof(float) * (size_t)batch * n_out, stream));
    // one resident wave: the per-CTA tables are paid once and no tail wave trails the others
    int occ = 0;
    QFX_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, (const void*)sv_large_expz_masks_kernel, 256, 0));
    size_t blocks = (((size_t)1 << n_local) + 256 * 8 - 1) / (256 * 8);     // groups: a thread takes 4 pairs per trip
    size_t cap = (size_t)num_sms() * (size_t)(occ > 0 ? occ : 1) / (size_t)batch;
    if (cap < 1) cap = 1;
    if (blocks > cap) blocks = cap;
    sv_large_expz_masks_kernel<<<dim3((unsigned)blocks, batch), 256, 0, stream>>>(
        reinterpret_cast<const float2*>(states), n_local, n_out, high_bits, m, out);
    return check_launch("sv_large_expz_masks_kernel");
}

// <Z_q> of every qubit = the Z-string readout with one-bit masks (q >= n_local: the bit comes from `high_bits`)
extern "C" int qfx_sv_large_expz(const float* states, int n_local, int n_total, unsigned int high_bits, int batch,
                                 float* out, cudaStream_t stream) {
    QFX_ARG_CHECK(n_local >= 1 && n_total >= n_local && n_total <= 32, QFX_ERR_BAD_ARG, "bad qubit counts");
    QFX_ARG_CHECK(batch >= 1 && batch <= 65535, QFX_ERR_BAD_ARG, "bad batch");
    MaskArgs m = {};
    for (int q = 0; q < n_total; ++q) m.m[q] = 1ull << q;
    return launch_expz_masks(states, n_local, n_total, high_bits, m, batch, out, stream);
}

extern "C" int qfx_sv_large_expz_masks(const float* states, int n_local, int n_out, unsigned int high_bits,
                                       const unsigned long long* masks_host, int batch, float* out,
                                       cudaStream_t stream) {
    QFX_ARG_CHECK(n_local >= 1 && n_local <= 34 && n_out >= 1 && n_out <= 32, QFX_ERR_BAD_ARG, "bad qubit counts");
    QFX_ARG_CHECK(masks_host != nullptr && batch >= 1 && batch <= 65535, QFX_ERR_BAD_ARG, "bad masks / batch");
    MaskArgs m = {};
    for (int q = 0; q < n_out; ++q) m.m[q] = masks_host[q];
    return launch_expz_masks(states, n_local, n_out, high_bits, m, batch, out, stream);
}

// CTAs per SM of the fused swap's grid (tuning knob, per device; results do not depend on it).  More CTAs keep more
// 16-byte peer loads in flight per SM; beyond what fits on an SM the extra CTAs simply run as a second wave.
static int g_swap_ctas_per_sm[QFX_MAX_DEVICES] = {};
extern "C" int qfx_sv_swap_set_ctas_per_sm(int ctas_per_sm) {
    QFX_ARG_CHECK(ctas_per_sm >= 0 && ctas_per_sm <= 32, QFX_ERR_BAD_ARG, "ctas_per_sm must be in 0..32 (0 = default)");
    g_swap_ctas_per_sm[current_device()] = ctas_per_sm;
    return 0;
}

// src_table: device array of `world` pointers (every rank's source shard, symmetric heap);
// flags_table: device array of `world` pointers to u32[world] flag arrays (nullable = no barrier,
// e.g. single-process emulation where the caller orders the launches).
extern "C" int qfx_sv_global_swap_fused(const float* const* src_table, float* dst, int n_local, int g_bits, int rank,
                                        int world, const int* ops_dev, int n_ops, const float* theta, int n_params,
                                        const float* angles, int n_feature, unsigned int* const* flags_table,
                                        unsigned int epoch, cudaStream_t stream) {
    QFX_ARG_CHECK(g_bits >= 1 && g_bits <= 3 && world == (1 << g_bits), QFX_ERR_UNSUPPORTED, "world must be 2, 4 or 8");
    QFX_ARG_CHECK(n_local - g_bits >= 1, QFX_ERR_BAD_ARG, "shard too small: a slice must hold at least two amplitudes");
    QFX_ARG_CHECK(n_ops >= 0 && n_ops <= 32, QFX_ERR_TOO_LARGE, "at most 32 fused ops");
    SwapArgs a = {};
    a.src = reinterpret_cast<const float2* const*>(src_table); a.dst = reinterpret_cast<float2*>(dst);
    a.n_local = n_local; a.g_bits = g_bits; a.rank = rank; a.world = world; a.ops = ops_dev; a.n_ops = n_ops;
    a.theta = theta; a.n_params = n_params; a.angles = angles; a.n_feature = n_feature;
    a.flags = flags_table; a.epoch = epoch;
    const size_t slice = (size_t)1 << (n_local - g_bits);
    size_t blocks = (slice / 2 + 255) / 256;
    const int cps = g_swap_ctas_per_sm[current_device()] > 0 ? g_swap_ctas_per_sm[current_device()] : 4;
    const size_t cap = (size_t)num_sms() * cps;  // block 0 (always in the first wave) publishes this rank's flag
    if (blocks > cap) blocks = cap;
    if (g_bits == 1) sv_global_swap_fused_kernel<1><<<(unsigned)blocks, 256, 0, stream>>>(a);
    else if (g_bits == 2) sv_global_swap_fused_kernel<2><<<(unsigned)blocks, 256, 0, stream>>>(a);
    else sv_global_swap_fused_kernel<3><<<(unsigned)blocks, 256, 0, stream>>>(a);
    return check_launch("sv_global_swap_fused_kernel");
}

// lambda = (sum_q cot[b][q] * Z-string(masks[q])) psi: the start of the adjoint sweep (lam may not alias psi)
extern "C" int qfx_sv_large_parity_obs(const float* psi, float* lam, int n_local, int n_out,
                                       const unsigned long long* masks_host, const float* cot, int batch,
                                       cudaStream_t stream) {
    QFX_ARG_CHECK(psi && lam && cot && masks_host && n_local >= 1 && n_local <= 34 && n_out >= 1 && n_out <= 32 && batch >= 1 &&
                  batch <= 65535, QFX_ERR_BAD_ARG, "bad arguments");
    ObsArgs m = {};
    for (int q = 0; q < n_out; ++q) m.m[q] = masks_host[q];
    size_t blocks = (((size_t)1 << n_local) + 256 * 8 - 1) / (256 * 8);
    const size_t cap = (size_t)num_sms() * 8 / (batch > 8 ? 8 : batch) + 1;
    if (blocks > cap) blocks = cap;
    sv_large_parity_obs_kernel<<<dim3((unsigned)blocks, batch), 256, 0, stream>>>(
        reinterpret_cast<const float2*>(psi), reinterpret_cast<float2*>(lam), n_local, n_out, m, cot);
    return check_launch("sv_large_parity_obs_kernel");
}

// One pass of the adjoint sweep: the ops of a compiled pass walked backwards over psi and lambda (both updated in
// place to their values BEFORE the pass); d/d theta accumulates into grad [theta_rows][n_params] (caller-zeroed).
extern "C" int qfx_sv_large_adjoint_pass(float* psi, float* lam, int n_qubits, int batch, int tile_bits, int low_bits,
                                         const int* tile_qubits_host, const int* ops_dev, int n_ops, const float* theta,
                                         int theta_rows, int n_params, const float* angles, int angle_rows, int n_feature,
                                         float* grad, cudaStream_t stream) {
    QFX_ARG_CHECK(psi && lam && grad && theta && batch >= 1 && batch <= 65535, QFX_ERR_BAD_ARG, "bad arguments");
    AdjArgs A;
    if (int e = fill_pass_args(A.base, psi, n_qubits, tile_bits, low_bits, tile_qubits_host, ops_dev, n_ops, theta, theta_rows,
                               n_params, angles, angle_rows, n_feature)) return e;
    A.lam = reinterpret_cast<float2*>(lam);
    A.grad = grad;
    const size_t smem = (sizeof(float2) << tile_bits) * 2;
    static PerDeviceFlag attr;
    if (!attr.here()) {
        QFX_CUDA_TRY(cudaFuncSetAttribute(sv_large_adjoint_pass_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 72 * 1024));
        attr.here() = true;
    }
    dim3 grid(1u << (n_qubits - tile_bits), batch);
    sv_large_adjoint_pass_kernel<<<grid, PASS_THREADS, smem, stream>>>(A);
    return check_launch("sv_large_adjoint_pass_kernel");
}
