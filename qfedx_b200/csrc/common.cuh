// Shared device/host helpers for the qfedx_b200 sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace qfx {

// ---- error plumbing for the C ABI -------------------------------------------------
// Every extern "C" entry point returns 0 on success or a cudaError_t / negative
// argument-error code; qfx_last_error() returns the text of the last failure.
void set_last_error(const char* fmt, ...);
inline int check_launch(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        set_last_error("%s: %s", what, cudaGetErrorString(e));
        return (int)e;
    }
    return 0;
}
#define QFX_CUDA_TRY(expr)                                                     \
    do {                                                                       \
        cudaError_t _e = (expr);                                               \
        if (_e != cudaSuccess) {                                               \
            ::qfx::set_last_error("%s: %s", #expr, cudaGetErrorString(_e));    \
            return (int)_e;                                                    \
        }                                                                      \
    } while (0)
#define QFX_ARG_CHECK(cond, code, ...)                                         \
    do {                                                                       \
        if (!(cond)) {                                                         \
            ::qfx::set_last_error(__VA_ARGS__);                                \
            return (code);                                                     \
        }                                                                      \
    } while (0)

constexpr int QFX_MAX_DEVICES = 64;   // per-device host state tables (streams, events, attribute flags)
enum : int { QFX_ERR_BAD_ARG = -1, QFX_ERR_UNSUPPORTED = -2, QFX_ERR_TOO_LARGE = -3 };

// ---- per-device host state ------------------------------------------------------------
// Function attributes and SM counts belong to a device, not to the process: tables are indexed by the
// current device so that one process can drive several GPUs.
inline int current_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    return (dev >= 0 && dev < QFX_MAX_DEVICES) ? dev : 0;
}
inline int device_sm_count() {
    static int sms[QFX_MAX_DEVICES] = {};
    const int dev = current_device();
    if (sms[dev] == 0) cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev);
    return sms[dev];
}
struct PerDeviceFlag {   // "done once on this device" marker
    bool done[QFX_MAX_DEVICES] = {};
    bool& here() { return done[current_device()]; }
};

// ---- counter-based RNG (shared bit-for-bit with qfedx_b200/fl/prng.py) --------------
// splitmix64 finaliser as a stateless hash; streams are keyed by mixing words in.
__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ uint64_t key3(uint64_t seed, uint64_t a, uint64_t b) {
    return mix64(mix64(mix64(seed) ^ a) ^ (b * 0xD1342543DE82EF95ull));
}
__host__ __device__ __forceinline__ uint64_t stream_at(uint64_t key, uint64_t idx) {
    return mix64(key ^ (idx * 0x9E3779B97F4A7C15ull));
}
// two uniforms in (0,1] from one 64-bit word -> one standard normal (Box-Muller)
__device__ __forceinline__ float normal_from_bits(uint64_t r) {
    float u1 = ((uint32_t)(r >> 40) + 1u) * (1.0f / 16777216.0f);          // (0,1]
    float u2 = ((uint32_t)(r & 0xFFFFFFu)) * (1.0f / 16777216.0f);          // [0,1)
    float rad = sqrtf(-2.0f * __logf(u1));
    float s, c;
    __sincosf(6.283185307179586f * u2, &s, &c);
    return rad * c;
}
__device__ __forceinline__ float uniform_from_bits(uint64_t r) {
    return ((uint32_t)(r >> 40)) * (1.0f / 16777216.0f);                     // [0,1)
}

// ---- readout-level noise on <Z> (noise models of /root/reference/ROADMAP.md:64-73) ------------------
// Depolarizing noise on the measured qubit, amplitude damping towards |0> and a readout confusion matrix all act on
// an exact <Z> as one affine map E' = alpha E + beta; `shots` > 0 then replaces E' by the mean of that many +-1
// measurement outcomes (binomial sample, keyed counter RNG; gradients flow through the affine part only).
// Mirrors qfedx_b200/noise: readout_affine / apply_to_expvals.
struct ReadoutNoise {
    float alpha, beta;
    int shots;
    unsigned long long key;
};
__device__ __forceinline__ float sample_shots(float e, int shots, unsigned long long key) {
    const float p0 = fminf(fmaxf(0.5f * (1.f + e), 0.f), 1.f);
    float k;
    if (shots <= 128) {          // exact: one Bernoulli draw per shot
        int c = 0;
        for (int i = 0; i < shots; ++i) c += uniform_from_bits(stream_at(key, (unsigned long long)i)) < p0 ? 1 : 0;
        k = (float)c;
    } else {                     // normal approximation of Binomial(shots, p0), rounded and clamped
        const float z = normal_from_bits(stream_at(key, 0ull));
        k = rintf((float)shots * p0 + sqrtf((float)shots * p0 * (1.f - p0)) * z);
        k = fminf(fmaxf(k, 0.f), (float)shots);
    }
    return 2.f * k / (float)shots - 1.f;
}
__device__ __forceinline__ float readout_noise(float e, const ReadoutNoise& rn, unsigned long long sample, int qubit) {
    float v = rn.alpha * e + rn.beta;
    if (rn.shots > 0) v = sample_shots(v, rn.shots, key3(rn.key, sample, (unsigned long long)qubit));
    return v;
}

// ---- keyed permutation of [0, n) without storage -----------------------------------------
// Four-round (unbalanced) Feistel network on exactly ceil(log2 n) bits with cycle walking: a bijection of
// [0, n) evaluated in O(1) per index (fewer than two walks on average), so a kernel can visit a client's
// samples in a fresh pseudo-random order every epoch without a permutation array.  32-bit round function.
// Mirrored bit for bit in qfedx_b200/fl/prng.py: feistel_permutation.
__host__ __device__ __forceinline__ uint32_t feistel_round(uint32_t x, uint32_t k) {
    x ^= k;
    x ^= x >> 16; x *= 0x85EBCA6Bu;
    x ^= x >> 13; x *= 0xC2B2AE35u;
    return x ^ (x >> 16);
}
__host__ __device__ __forceinline__ uint32_t feistel_perm(uint32_t j, uint32_t n, uint64_t key) {
    if (n <= 1) return 0;
    int bits = 1;
    while ((1u << bits) < n) ++bits;
    const int a = bits >> 1, b = bits - a;             // |l| = a, |r| = b (b >= a, b - a <= 1)
    const uint32_t k0 = (uint32_t)key, k1 = (uint32_t)(key >> 32);
    uint32_t x = j;
    do {
        uint32_t l = x & ((1u << a) - 1u), r = x >> a;
        int la = a, lb = b;                            // current widths of l and r
#pragma unroll
        for (int rnd = 0; rnd < 4; ++rnd) {
            const uint32_t f = feistel_round(r, (rnd & 1 ? k1 : k0) + 0x9E3779B9u * (uint32_t)rnd) & ((1u << la) - 1u);
            const uint32_t nl = r;
            r = l ^ f;
            l = nl;
            const int t = la; la = lb; lb = t;
        }
        x = (r << la) | l;                             // after an even number of rounds the widths are (a, b) again
    } while (x >= n);
    return x;
}

// ---- warp helpers -----------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
template <int WIDTH>
__device__ __forceinline__ float group_sum(float v) {  // butterfly over WIDTH consecutive lanes
#pragma unroll
    for (int o = WIDTH / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float wrap_pi(float x) {  // to [-pi, pi)
    const float two_pi = 6.283185307179586f;
    return x - two_pi * floorf((x + 3.14159265358979323846f) / two_pi);
}

}  // namespace qfx
