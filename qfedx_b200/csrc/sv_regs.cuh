// Register-tiled statevector for n <= 10 qubits (kernels K1/K3/K4a/K4b of SURVEY.md 2.5).
//
// A circuit's 2^n complex amplitudes live entirely in registers, spread over a group of
// LANES = 2^LB consecutive lanes of one warp with A = 2^NL amplitudes per lane:
//   n <= 4      : LB = 0, one thread owns the whole state (16 amplitudes = 32 registers)
//   5 <= n <= 9 : LB = n-4, 16 amplitudes per lane, 2..32 lanes per circuit
//   n == 10     : LB = 5, 32 amplitudes per lane
// Amplitude index = (lane_in_group << NL) | local, so qubits [0, NL) are "local" (gates are
// pure register arithmetic, CNOTs between local qubits are compile-time register renames)
// and qubits [NL, n) are "lane" qubits (gates exchange amplitudes with __shfl_xor_sync).
// Every loop over amplitudes and qubits is fully unrolled so that all indexing is static
// and nothing spills to local memory (checked with -Xptxas -v).
//
// Gate conventions: qfedx_b200/oracle/statevector.py (little-endian, exp(-i a P/2)).
// No reference counterpart: /root/reference has no simulator; the ansatz is the one planned
// in /root/reference/ROADMAP.md:126-128.
#pragma once
#include "common.cuh"

namespace qfx {

constexpr unsigned FULL_MASK = 0xffffffffu;

// Lane-qubit loops of the merged sweeps: rolled (one copy of the gate code, small kernel) or unrolled
// (static qubit indices, more scheduling freedom).  Measured on B200 for the 8-qubit training kernel.
#ifndef QFX_ROLL_LANE_LOOPS
#define QFX_ROLL_LANE_LOOPS 0
#endif
#if QFX_ROLL_LANE_LOOPS
#define QFX_LANE_LOOP _Pragma("unroll 1")
#else
#define QFX_LANE_LOOP _Pragma("unroll")
#endif


// default lane split: favours throughput (most amplitudes per lane, fewest shuffles)
constexpr int default_lane_bits(int nq) { return (nq <= 4) ? 0 : ((nq <= 9) ? nq - 4 : 5); }
// latency split: as many lanes per circuit as possible (shortest instruction stream per warp)
constexpr int latency_lane_bits(int nq) { return nq <= 1 ? 0 : (nq - 1 < 5 ? nq - 1 : 5); }

template <int NQ, int LB_ = default_lane_bits(NQ)>
struct SvCfg {
    static_assert(NQ >= 1 && NQ <= 10, "register-tiled statevector supports 1..10 qubits");
    static_assert(LB_ >= 0 && LB_ <= 5 && LB_ <= NQ && NQ - LB_ <= 5, "bad lane split");
    static constexpr int LB = LB_;
    static constexpr int NL = NQ - LB;
    static constexpr int A = 1 << NL;
    static constexpr int LANES = 1 << LB;
    static constexpr int DIM = 1 << NQ;
};

enum : int { BASIS_RX = 0, BASIS_RY = 1, BASIS_RZ = 2 };
enum : int { ENT_CNOT = 0, ENT_CZ = 1, ENT_NONE = 2 };

template <int NQ, int LB_ = default_lane_bits(NQ)>
struct Sv {
    using C = SvCfg<NQ, LB_>;
    float re[C::A], im[C::A];

    // ---- initial states ------------------------------------------------------------
    __device__ __forceinline__ void set_basis(int b, int lg) {
#pragma unroll
        for (int i = 0; i < C::A; ++i) {
            re[i] = (((lg << C::NL) | i) == b) ? 1.0f : 0.0f;
            im[i] = 0.0f;
        }
    }
    // product state of one rotation per qubit on |0..0>; ang points at NQ floats
    __device__ __forceinline__ void set_product(const float* __restrict__ ang, int basis, int lg) {
#pragma unroll
        for (int i = 0; i < C::A; ++i) { re[i] = 1.0f; im[i] = 0.0f; }
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            float s, c;
            sincosf(0.5f * ang[q], &s, &c);
            // factor for bit 0: f0, for bit 1: f1 (complex)
            float f0r, f0i, f1r, f1i;
            if (basis == BASIS_RY)      { f0r = c; f0i = 0.f; f1r = s;   f1i = 0.f; }
            else if (basis == BASIS_RX) { f0r = c; f0i = 0.f; f1r = 0.f; f1i = -s; }
            else                        { f0r = c; f0i = -s;  f1r = 0.f; f1i = 0.f; }
#pragma unroll
            for (int i = 0; i < C::A; ++i) {
                const bool bit = (q < C::NL) ? ((i >> q) & 1) : ((lg >> (q - C::NL)) & 1);
                const float fr = bit ? f1r : f0r, fi = bit ? f1i : f0i;
                const float r = re[i], m = im[i];
                re[i] = r * fr - m * fi;
                im[i] = r * fi + m * fr;
            }
        }
    }
    // real amplitude vector (amplitude encoding); src points at DIM floats, already normalised
    __device__ __forceinline__ void set_real(const float* __restrict__ src, int lg) {
#pragma unroll
        for (int i = 0; i < C::A; ++i) { re[i] = src[(lg << C::NL) | i]; im[i] = 0.0f; }
    }

    // ---- single-qubit rotations (q is a compile-time constant after unrolling) -------
    __device__ __forceinline__ void rx(int q, float c, float s, int lg) {
        if (q < C::NL) {
#pragma unroll
            for (int i = 0; i < C::A; ++i) {
                if (!((i >> q) & 1)) {
                    const int j = i | (1 << q);
                    const float r0 = re[i], i0 = im[i], r1 = re[j], i1 = im[j];
                    re[i] = c * r0 + s * i1;  im[i] = c * i0 - s * r1;
                    re[j] = c * r1 + s * i0;  im[j] = c * i1 - s * r0;
                }
            }
        } else {
            const int m = 1 << (q - C::NL);
#pragma unroll
            for (int i = 0; i < C::A; ++i) {
                const float pr = __shfl_xor_sync(FULL_MASK, re[i], m);
                const float pi = __shfl_xor_sync(FULL_MASK, im[i], m);
                const float r0 = re[i], i0 = im[i];
                re[i] = c * r0 + s * pi;
                im[i] = c * i0 - s * pr;
            }
        }
    }
    __device__ __forceinline__ void ry(int q, float c, float s, int lg) {
        if (q < C::NL) {
#pragma unroll
            for (int i = 0; i < C::A; ++i) {
                if (!((i >> q) & 1)) {
                    const int j = i | (1 << q);
                    const float r0 = re[i], i0 = im[i], r1 = re[j], i1 = im[j];
                    re[i] = c * r0 - s * r1;  im[i] = c * i0 - s * i1;
                    re[j] = s * r0 + c * r1;  im[j] = s * i0 + c * i1;
                }
            }
        } else {
            const int m = 1 << (q - C::NL);
            const float sg = ((lg >> (q - C::NL)) & 1) ? s : -s;
#pragma unroll
            for (int i = 0; i < C::A; ++i) {
                const float pr = __shfl_xor_sync(FULL_MASK, re[i], m);
                const float pi = __shfl_xor_sync(FULL_MASK, im[i], m);
                re[i] = c * re[i] + sg * pr;
                im[i] = c * im[i] + sg * pi;
            }
        }
    }
    __device__ __forceinline__ void rz(int q, float c, float s, int lg) {
        // bit 0: *(c - i s); bit 1: *(c + i s)
        const float s_lane = (q >= C::NL && ((lg >> (q - C::NL)) & 1)) ? -s : s;
#pragma unroll
        for (int i = 0; i < C::A; ++i) {
            const float sg = (q < C::NL) ? (((i >> q) & 1) ? -s : s) : s_lane;
            const float r = re[i], m = im[i];
            re[i] = c * r + sg * m;
            im[i] = c * m - sg * r;
        }
    }

    // ---- entanglers -------------------------------------------------------------------
    __device__ __forceinline__ void cnot(int ctl, int tgt, int lg) {
        if (tgt < C::NL) {
            const bool lane_ctl_on = (ctl >= C::NL) ? ((lg >> (ctl - C::NL)) & 1) : true;
#pragma unroll
            for (int i = 0; i < C::A; ++i) {
                const bool ctl_local_on = (ctl < C::NL) ? ((i >> ctl) & 1) : true;
                if (ctl_local_on && !((i >> tgt) & 1)) {
                    const int j = i | (1 << tgt);
                    if (ctl < C::NL) {  // static register rename
                        float t = re[i]; re[i] = re[j]; re[j] = t;
                        t = im[i]; im[i] = im[j]; im[j] = t;
                    } else {
                        const float ri = re[i], rj = re[j], ii = im[i], ij = im[j];
                        re[i] = lane_ctl_on ? rj : ri;  re[j] = lane_ctl_on ? ri : rj;
                        im[i] = lane_ctl_on ? ij : ii;  im[j] = lane_ctl_on ? ii : ij;
                    }
                }
            }
        } else {
            const int m = 1 << (tgt - C::NL);
            if (ctl < C::NL) {
#pragma unroll
                for (int i = 0; i < C::A; ++i) {
                    if ((i >> ctl) & 1) {
                        re[i] = __shfl_xor_sync(FULL_MASK, re[i], m);
                        im[i] = __shfl_xor_sync(FULL_MASK, im[i], m);
                    }
                }
            } else {
                const bool on = (lg >> (ctl - C::NL)) & 1;
#pragma unroll
                for (int i = 0; i < C::A; ++i) {
                    const float pr = __shfl_xor_sync(FULL_MASK, re[i], m);
                    const float pi = __shfl_xor_sync(FULL_MASK, im[i], m);
                    re[i] = on ? pr : re[i];
                    im[i] = on ? pi : im[i];
                }
            }
        }
    }
    __device__ __forceinline__ void cz(int a, int b, int lg) {
        const bool lane_a = (a >= C::NL) ? ((lg >> (a - C::NL)) & 1) : true;
        const bool lane_b = (b >= C::NL) ? ((lg >> (b - C::NL)) & 1) : true;
#pragma unroll
        for (int i = 0; i < C::A; ++i) {
            const bool on_a = (a < C::NL) ? ((i >> a) & 1) : lane_a;
            const bool on_b = (b < C::NL) ? ((i >> b) & 1) : lane_b;
            if (on_a && on_b) { re[i] = -re[i]; im[i] = -im[i]; }
        }
    }
    __device__ __forceinline__ void entangle(int ent, bool reverse, int lg) {
        if (ent == ENT_NONE || NQ < 2) return;
        if (NQ == 2) {
            if (ent == ENT_CNOT) cnot(0, 1, lg); else cz(0, 1, lg);
            return;
        }
#pragma unroll
        for (int k = 0; k < NQ; ++k) {
            const int q = reverse ? (NQ - 1 - k) : k;
            const int t = (q + 1) % NQ;
            if (ent == ENT_CNOT) cnot(q, t, lg); else cz(q, t, lg);
        }
    }

    // ---- lane-qubit gates with a RUNTIME lane index k (qubit NL + k) --------------------------
    // Same arithmetic as rx()/cnot() on lane qubits, but callable from a rolled loop: one copy of
    // the code serves all lane qubits, which keeps the training kernel inside the instruction cache.
    __device__ __forceinline__ void rx_lane_rt(int k, float c, float s) {
        const int m = 1 << k;
#pragma unroll
        for (int i = 0; i < C::A; ++i) {
            const float pr = __shfl_xor_sync(FULL_MASK, re[i], m);
            const float pi = __shfl_xor_sync(FULL_MASK, im[i], m);
            const float r0 = re[i], i0 = im[i];
            re[i] = c * r0 + s * pi;
            im[i] = c * i0 - s * pr;
        }
    }
    __device__ __forceinline__ void cnot_lanes_rt(int kc, int kt, int lg) {
        const int m = 1 << kt;
        const bool on = (lg >> kc) & 1;
#pragma unroll
        for (int i = 0; i < C::A; ++i) {
            const float pr = __shfl_xor_sync(FULL_MASK, re[i], m);
            const float pi = __shfl_xor_sync(FULL_MASK, im[i], m);
            re[i] = on ? pr : re[i];
            im[i] = on ? pi : im[i];
        }
    }
    // ring entangler with the lane-lane links in a rolled loop
    __device__ __forceinline__ void entangle_rolled(int ent, bool reverse, int lg) {
        if (ent != ENT_CNOT || NQ < 3 || C::LB < 2) { entangle(ent, reverse, lg); return; }
        if (!reverse) {
#pragma unroll
            for (int q = 0; q + 1 < C::NL; ++q) cnot(q, q + 1, lg);
            cnot(C::NL - 1, C::NL, lg);
            QFX_LANE_LOOP
            for (int k = 0; k + 1 < C::LB; ++k) cnot_lanes_rt(k, k + 1, lg);
            cnot(NQ - 1, 0, lg);
        } else {
            cnot(NQ - 1, 0, lg);
            QFX_LANE_LOOP
            for (int k = C::LB - 2; k >= 0; --k) cnot_lanes_rt(k, k + 1, lg);
            cnot(C::NL - 1, C::NL, lg);
#pragma unroll
            for (int q = C::NL - 2; q >= 0; --q) cnot(q, q + 1, lg);
        }
    }

    // ---- the whole CNOT ring as ONE index permutation through shared memory ---------------------
    // CNOT(0,1) CNOT(1,2) ... CNOT(n-2,n-1) CNOT(n-1,0) maps basis index i to pi(i) with
    //   bit_k(pi) = b_0 ^ ... ^ b_k  (k >= 1),   bit_0(pi) = b_0 ^ bit_{n-1}(pi)
    // - a linear map over GF(2), so pi(lane bits | local bits) = pi(lane part) ^ pi(local part): the local
    // parts are compile-time constants and the lane part is computed once per thread.  The ring then costs
    // one shared-memory round trip (A stores + A loads of 8 bytes) instead of n dependent CNOT steps, five
    // of which need two shuffles and two selects per amplitude.  `buf` holds the circuit's 2^n amplitudes.
    static __device__ __forceinline__ constexpr unsigned ring_perm(unsigned i) {
        unsigned p = i;
        p ^= p << 1; p ^= p << 2; p ^= p << 4; p ^= p << 8;
        p &= (1u << NQ) - 1u;
        return p ^ ((p >> (NQ - 1)) & 1u);
    }
    static __device__ __forceinline__ constexpr unsigned ring_perm_inv(unsigned j) {
        unsigned p = j ^ ((j >> (NQ - 1)) & 1u);          // undo CNOT(n-1, 0)
        return (p ^ (p << 1)) & ((1u << NQ) - 1u);          // undo the prefix XOR
    }
    // physical slot of amplitude index d: XOR-fold the high nibble into the low one so that the 16 lanes of a
    // half-warp, whose indices differ in high bits, land in different 8-byte banks (a bijection on indices)
    static __device__ __forceinline__ constexpr unsigned ring_slot(unsigned d) { return d ^ ((d >> 4) & 0xFu); }
    __device__ __forceinline__ void ring_via_smem(float2* buf, bool reverse, int lg) {
        const unsigned lane_part = reverse ? ring_perm_inv((unsigned)lg << C::NL) : ring_perm((unsigned)lg << C::NL);
#pragma unroll
        for (int i = 0; i < C::A; ++i) {
            const unsigned dst = lane_part ^ (reverse ? ring_perm_inv((unsigned)i) : ring_perm((unsigned)i));
            buf[ring_slot(dst)] = make_float2(re[i], im[i]);
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < C::A; ++i) {
            const float2 v = buf[ring_slot(((unsigned)lg << C::NL) | (unsigned)i)];
            re[i] = v.x; im[i] = v.y;
        }
        __syncwarp();
    }

    // ---- reductions --------------------------------------------------------------------
    // <Z_q> for all qubits, reduced over the circuit's lanes; every lane gets the result
    __device__ __forceinline__ void expval_z_all(float* __restrict__ ez /*[NQ]*/, int lg) const {
        float p[C::A];
        float tot = 0.f;
#pragma unroll
        for (int i = 0; i < C::A; ++i) { p[i] = re[i] * re[i] + im[i] * im[i]; tot += p[i]; }
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            float v;
            if (q < C::NL) {
                v = 0.f;
#pragma unroll
                for (int i = 0; i < C::A; ++i) v += ((i >> q) & 1) ? -p[i] : p[i];
            } else {
                v = ((lg >> (q - C::NL)) & 1) ? -tot : tot;
            }
            ez[q] = group_sum<C::LANES>(v);
        }
    }
    // lam = diag(sum_q cot[q] Z_q) * psi
    __device__ __forceinline__ void set_obs_times(const Sv& psi, const float* __restrict__ cot, int lg) {
        float lane_part = 0.f;
#pragma unroll
        for (int q = C::NL; q < NQ; ++q) lane_part += ((lg >> (q - C::NL)) & 1) ? -cot[q] : cot[q];
#pragma unroll
        for (int i = 0; i < C::A; ++i) {
            float d = lane_part;
#pragma unroll
            for (int q = 0; q < C::NL; ++q) d += ((i >> q) & 1) ? -cot[q] : cot[q];
            re[i] = d * psi.re[i];
            im[i] = d * psi.im[i];
        }
    }
};

// Im <lam| Z_q |psi>, this lane's share
template <int NQ, int LB_>
__device__ __forceinline__ float im_lam_z_psi(const Sv<NQ, LB_>& lam, const Sv<NQ, LB_>& psi, int q, int lg) {
    using C = SvCfg<NQ, LB_>;
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < C::A; ++i) {
        const float t = lam.re[i] * psi.im[i] - lam.im[i] * psi.re[i];
        const bool bit = (q < C::NL) ? ((i >> q) & 1) : ((lg >> (q - C::NL)) & 1);
        acc += bit ? -t : t;
    }
    return acc;
}

// One reverse step through RX(theta) on qubit q: returns this lane's share of
// Im <lam| X_q |psi> (evaluated after the gate) and replaces psi, lam by RX^dagger applied
// to them.  For lane qubits the partner amplitudes are fetched once and used for both.
template <int NQ, int LB_>
__device__ __forceinline__ float grad_and_undo_rx(Sv<NQ, LB_>& psi, Sv<NQ, LB_>& lam, int q, float c, float s, int lg) {
    using C = SvCfg<NQ, LB_>;
    float acc = 0.f;
    if (q < C::NL) {
#pragma unroll
        for (int i = 0; i < C::A; ++i) {
            if (!((i >> q) & 1)) {
                const int j = i | (1 << q);
                acc += lam.re[i] * psi.im[j] - lam.im[i] * psi.re[j];
                acc += lam.re[j] * psi.im[i] - lam.im[j] * psi.re[i];
            }
        }
        psi.rx(q, c, -s, lg);
        lam.rx(q, c, -s, lg);
    } else {
        const int m = 1 << (q - C::NL);
#pragma unroll
        for (int i = 0; i < C::A; ++i) {
            const float pr = __shfl_xor_sync(FULL_MASK, psi.re[i], m);
            const float pi = __shfl_xor_sync(FULL_MASK, psi.im[i], m);
            const float qr = __shfl_xor_sync(FULL_MASK, lam.re[i], m);
            const float qi = __shfl_xor_sync(FULL_MASK, lam.im[i], m);
            acc += lam.re[i] * pi - lam.im[i] * pr;
            const float r0 = psi.re[i], i0 = psi.im[i];
            psi.re[i] = c * r0 - s * pi;  psi.im[i] = c * i0 + s * pr;
            const float l0 = lam.re[i], m0 = lam.im[i];
            lam.re[i] = c * l0 - s * qi;  lam.im[i] = c * m0 + s * qr;
        }
    }
    return acc;
}

// grad_and_undo_rx for lane qubit NL + k with a runtime k (rolled loops)
template <int NQ, int LB_>
__device__ __forceinline__ float grad_and_undo_rx_lane_rt(Sv<NQ, LB_>& psi, Sv<NQ, LB_>& lam, int k, float c, float s) {
    using C = SvCfg<NQ, LB_>;
    const int m = 1 << k;
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < C::A; ++i) {
        const float pr = __shfl_xor_sync(FULL_MASK, psi.re[i], m);
        const float pi = __shfl_xor_sync(FULL_MASK, psi.im[i], m);
        const float qr = __shfl_xor_sync(FULL_MASK, lam.re[i], m);
        const float qi = __shfl_xor_sync(FULL_MASK, lam.im[i], m);
        acc += lam.re[i] * pi - lam.im[i] * pr;
        const float r0 = psi.re[i], i0 = psi.im[i];
        psi.re[i] = c * r0 - s * pi;  psi.im[i] = c * i0 + s * pr;
        const float l0 = lam.re[i], m0 = lam.im[i];
        lam.re[i] = c * l0 - s * qi;  lam.im[i] = c * m0 + s * qr;
    }
    return acc;
}

// Im <lam| P_q |psi> for the feature-map generator (no undo needed: it is the first layer)
template <int NQ, int LB_>
__device__ __forceinline__ float im_lam_p_psi(const Sv<NQ, LB_>& lam, const Sv<NQ, LB_>& psi, int basis, int q, int lg) {
    using C = SvCfg<NQ, LB_>;
    if (basis == BASIS_RZ) return im_lam_z_psi(lam, psi, q, lg);
    float acc = 0.f;
    if (q < C::NL) {
#pragma unroll
        for (int i = 0; i < C::A; ++i) {
            if (!((i >> q) & 1)) {
                const int j = i | (1 << q);
                if (basis == BASIS_RX) {
                    acc += lam.re[i] * psi.im[j] - lam.im[i] * psi.re[j];
                    acc += lam.re[j] * psi.im[i] - lam.im[j] * psi.re[i];
                } else {  // Y
                    acc -= lam.re[i] * psi.re[j] + lam.im[i] * psi.im[j];
                    acc += lam.re[j] * psi.re[i] + lam.im[j] * psi.im[i];
                }
            }
        }
    } else {
        const int m = 1 << (q - C::NL);
        const bool bit = (lg >> (q - C::NL)) & 1;
#pragma unroll
        for (int i = 0; i < C::A; ++i) {
            const float pr = __shfl_xor_sync(FULL_MASK, psi.re[i], m);
            const float pi = __shfl_xor_sync(FULL_MASK, psi.im[i], m);
            if (basis == BASIS_RX) {
                acc += lam.re[i] * pi - lam.im[i] * pr;
            } else {
                const float t = lam.re[i] * pr + lam.im[i] * pi;
                acc += bit ? t : -t;
            }
        }
    }
    return acc;
}

// ---- trig sources: where cos/sin(theta/2) come from ------------------------------------
// Table in shared memory (one parameter group per CTA), with an optional +-pi/2 shift of one
// parameter for the parameter-shift rule: theta/2 +- pi/4.
struct TrigTable {
    const float2* __restrict__ t;
    int shift_p;       // -1: none
    float shift_sign;  // +1 / -1
    __device__ __forceinline__ float2 get(int p) const {
        float2 v = t[p];
        if (p == shift_p) {
            const float k = 0.70710678118654752f;
            const float c = (v.x - shift_sign * v.y) * k;
            const float s = (v.y + shift_sign * v.x) * k;
            v = make_float2(c, s);
        }
        return v;
    }
};
// Straight from a parameter vector in global memory (circuits of mixed groups in one warp).
struct TrigDirect {
    const float* __restrict__ theta;
    int shift_p;
    float shift_sign;
    __device__ __forceinline__ float2 get(int p) const {
        float a = 0.5f * theta[p];
        if (p == shift_p) a += shift_sign * 0.78539816339744831f;
        float s, c;
        sincosf(a, &s, &c);
        return make_float2(c, s);
    }
};

// ---- ansatz sweeps ---------------------------------------------------------------------
template <int NQ, int LB_, class Trig>
__device__ __forceinline__ void ansatz_forward(Sv<NQ, LB_>& psi, const Trig& trig, int n_layers, int ent, int lg) {
    for (int l = 0; l < n_layers; ++l) {
#pragma unroll
        for (int q = 0; q < NQ; ++q) {
            const float2 a = trig.get((l * NQ + q) * 2 + 0);
            psi.rx(q, a.x, a.y, lg);
            const float2 b = trig.get((l * NQ + q) * 2 + 1);
            psi.rz(q, b.x, b.y, lg);
        }
        psi.entangle(ent, false, lg);
    }
}

// Reverse sweep.  sink(p, v) receives this lane's share of d<O>/d theta_p.
template <int NQ, int LB_, class Trig, class Sink>
__device__ __forceinline__ void ansatz_backward(Sv<NQ, LB_>& psi, Sv<NQ, LB_>& lam, const Trig& trig, int n_layers, int ent,
                                                int lg, Sink& sink) {
    for (int l = n_layers - 1; l >= 0; --l) {
        psi.entangle(ent, true, lg);
        lam.entangle(ent, true, lg);
#pragma unroll
        for (int k = 0; k < NQ; ++k) {
            const int q = NQ - 1 - k;
            const int p = (l * NQ + q) * 2;
            const float2 b = trig.get(p + 1);
            sink(p + 1, im_lam_z_psi(lam, psi, q, lg));
            psi.rz(q, b.x, -b.y, lg);
            lam.rz(q, b.x, -b.y, lg);
            const float2 a = trig.get(p);
            sink(p, grad_and_undo_rx(psi, lam, q, a.x, a.y, lg));
        }
    }
}

// ---- many-values warp reduction ---------------------------------------------------------------
// Sums N per-lane values over the 32 lanes of a warp with N-1 (+1) shuffles instead of 5 N: every
// butterfly step halves the number of values a lane carries (it keeps the half selected by one of its
// lane bits and sends the other half to its partner).  Afterwards lane L holds the total of value
// index slot_of_lane(L) in v[0]; for N = 16 the two lanes that differ in bit 0 hold the same total.
template <int N>
__device__ __forceinline__ void warp_reduce_many(float (&v)[N], int lane) {
    static_assert(N == 16 || N == 32, "warp_reduce_many handles 16 or 32 values");
#pragma unroll
    for (int cnt = N / 2, mask = 16; cnt >= 1; cnt >>= 1, mask >>= 1) {
        const bool up = (lane & mask) != 0;
#pragma unroll
        for (int i = 0; i < cnt; ++i) {
            const float keep = up ? v[i + cnt] : v[i];
            const float send = up ? v[i] : v[i + cnt];
            v[i] = keep + __shfl_xor_sync(FULL_MASK, send, mask);
        }
    }
    if (N == 16) v[0] += __shfl_xor_sync(FULL_MASK, v[0], 1);
}
template <int N>
__device__ __forceinline__ int slot_of_lane(int lane) {
    // step with mask m keeps the upper half when (lane & m): the kept index gains cnt = N/2, N/4, ...
    if (N == 32) return ((lane >> 4) & 1) * 16 + ((lane >> 3) & 1) * 8 + ((lane >> 2) & 1) * 4 + ((lane >> 1) & 1) * 2 + (lane & 1);
    return ((lane >> 4) & 1) * 8 + ((lane >> 3) & 1) * 4 + ((lane >> 2) & 1) * 2 + ((lane >> 1) & 1);
}

// ---- merged-diagonal sweeps (all circuits of a CTA share theta) ---------------------------
// Gates on different qubits commute, so a layer is [RX on every qubit][RZ on every qubit][ring].
// The RZ block is one diagonal; it factors into a table over the local bits and a table over the
// lane bits, both precomputed per optimizer step in shared memory (build_diag_tables).  Applying
// it is two complex multiplies per amplitude instead of one per amplitude per qubit.
struct DiagTables {
    const float2* __restrict__ loc;   // [L][A]     product over local qubits
    const float2* __restrict__ lane;  // [L][LANES] product over lane qubits
};

template <int NQ, int LB_>
__device__ __forceinline__ void build_diag_tables(float2* loc, float2* lane, const float2* tab, int n_layers,
                                                  int tid, int nthreads) {
    using C = SvCfg<NQ, LB_>;
    for (int e = tid; e < n_layers * (C::A + C::LANES); e += nthreads) {
        const int l = e / (C::A + C::LANES), r = e % (C::A + C::LANES);
        const bool is_lane = r >= C::A;
        const int idx = is_lane ? r - C::A : r;
        float fr = 1.f, fi = 0.f;
        for (int k = 0; k < (is_lane ? C::LB : C::NL); ++k) {
            const int q = is_lane ? C::NL + k : k;
            const float2 cs = tab[(l * NQ + q) * 2 + 1];
            const float si = ((idx >> k) & 1) ? cs.y : -cs.y;  // e^{-+ i phi/2}
            const float nr = fr * cs.x - fi * si, ni = fr * si + fi * cs.x;
            fr = nr; fi = ni;
        }
        if (is_lane) lane[l * C::LANES + idx] = make_float2(fr, fi);
        else loc[l * C::A + idx] = make_float2(fr, fi);
    }
}

template <int NQ, int LB_>
__device__ __forceinline__ void apply_diag(Sv<NQ, LB_>& psi, const DiagTables& dt, int l, int lg, bool conj) {
    using C = SvCfg<NQ, LB_>;
    const float2 dl = dt.lane[l * C::LANES + lg];
#pragma unroll
    for (int i = 0; i < C::A; ++i) {
        const float2 d0 = dt.loc[l * C::A + i];
        const float fr = d0.x * dl.x - d0.y * dl.y;
        const float fi = conj ? -(d0.x * dl.y + d0.y * dl.x) : (d0.x * dl.y + d0.y * dl.x);
        const float r = psi.re[i], m = psi.im[i];
        psi.re[i] = r * fr - m * fi;
        psi.im[i] = r * fi + m * fr;
    }
}

// `ring_buf`: per-circuit shared-memory scratch of 2^n float2 for the permutation form of the CNOT ring
// (nullptr: apply the CNOTs one by one).
template <int NQ, int LB_>
__device__ __forceinline__ void ansatz_forward_merged(Sv<NQ, LB_>& psi, const float2* tab, const DiagTables& dt,
                                                      int n_layers, int ent, int lg, float2* ring_buf = nullptr) {
    using C = SvCfg<NQ, LB_>;
#pragma unroll 1
    for (int l = 0; l < n_layers; ++l) {
#pragma unroll
        for (int q = 0; q < C::NL; ++q) {
            const float2 a = tab[(l * NQ + q) * 2];
            psi.rx(q, a.x, a.y, lg);
        }
        if (C::LB > 0) {
            float2 a_next = tab[(l * NQ + C::NL) * 2];
            QFX_LANE_LOOP
            for (int k = 0; k < C::LB; ++k) {
                const float2 a = a_next;
                if (k + 1 < C::LB) a_next = tab[(l * NQ + C::NL + k + 1) * 2];   // prefetch the next gate's (cos, sin)
                psi.rx_lane_rt(k, a.x, a.y);
            }
        }
        apply_diag(psi, dt, l, lg, false);
        if (ring_buf != nullptr && ent == ENT_CNOT && NQ >= 3 && SvCfg<NQ, LB_>::LB >= 2) psi.ring_via_smem(ring_buf, false, lg);
        else psi.entangle_rolled(ent, false, lg);
    }
}

// Reverse sweep with the merged diagonal.  The 2 NQ gradient shares of a layer are collected in registers
// and handed to sink.layer(l, g) in one block (slot q = RX of qubit q, slot NQ + q = RZ of qubit q), so the
// sink can reduce them over the warp together instead of one 5-step butterfly per parameter.
template <int NQ, int LB_, class Sink>
__device__ __forceinline__ void ansatz_backward_merged(Sv<NQ, LB_>& psi, Sv<NQ, LB_>& lam, const float2* tab,
                                                       const DiagTables& dt, int n_layers, int ent, int lg,
                                                       Sink& sink, float2* ring_buf = nullptr) {
    using C = SvCfg<NQ, LB_>;
    constexpr int NG = (2 * NQ <= 16) ? 16 : 32;
#pragma unroll 1
    for (int l = n_layers - 1; l >= 0; --l) {
        float g[NG];
#pragma unroll
        for (int i = 0; i < NG; ++i) g[i] = 0.f;
        if (ring_buf != nullptr && ent == ENT_CNOT && NQ >= 3 && SvCfg<NQ, LB_>::LB >= 2) {
            psi.ring_via_smem(ring_buf, true, lg);
            lam.ring_via_smem(ring_buf, true, lg);
        } else {
            psi.entangle_rolled(ent, true, lg);
            lam.entangle_rolled(ent, true, lg);
        }
        // all RZ gradients from one pass: t_i = Im(conj(lam_i) psi_i), signed sums per qubit
        {
            float t[C::A];
            float tot = 0.f;
#pragma unroll
            for (int i = 0; i < C::A; ++i) { t[i] = lam.re[i] * psi.im[i] - lam.im[i] * psi.re[i]; tot += t[i]; }
#pragma unroll
            for (int q = 0; q < NQ; ++q) {
                float v;
                if (q < C::NL) {
                    v = 0.f;
#pragma unroll
                    for (int i = 0; i < C::A; ++i) v += ((i >> q) & 1) ? -t[i] : t[i];
                } else {
                    v = ((lg >> (q - C::NL)) & 1) ? -tot : tot;
                }
                g[NQ + q] = v;
            }
        }
        apply_diag(psi, dt, l, lg, true);
        apply_diag(lam, dt, l, lg, true);
        {
            float2 a_next = tab[(l * NQ + NQ - 1) * 2];
            QFX_LANE_LOOP
            for (int k = C::LB - 1; k >= 0; --k) {
                const float2 a = a_next;
                if (k > 0) a_next = tab[(l * NQ + C::NL + k - 1) * 2];   // prefetch the next gate's (cos, sin)
                const float v = grad_and_undo_rx_lane_rt(psi, lam, k, a.x, a.y);
#pragma unroll
                for (int j = 0; j < C::LB; ++j) g[C::NL + j] += (j == k) ? v : 0.f;   // static slots, rolled loop
            }
        }
#pragma unroll
        for (int k = 0; k < C::NL; ++k) {
            const int q = C::NL - 1 - k;
            const float2 a = tab[(l * NQ + q) * 2];
            g[q] = grad_and_undo_rx(psi, lam, q, a.x, a.y, lg);
        }
        sink.template layer<NQ, NG>(l, g);
    }
}

}  // namespace qfx
