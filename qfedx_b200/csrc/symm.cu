// Symmetric-heap plumbing for the fused compute+collective kernels: every rank allocates the same
// buffers, exports CUDA IPC handles, and maps every peer's buffer so kernels can issue plain
// ld/st (and system-scope flag operations) on peer memory over NVLink / NVSwitch.
// torch.distributed is used only to exchange the 64-byte handles (qfedx_b200/parallel/symm.py).
// No reference counterpart: /root/reference has no communication layer.
#include "common.cuh"
#include "../../include/qfedx_b200.h"
#include <string.h>

using namespace qfx;

extern "C" int qfx_symm_alloc(size_t bytes, void** ptr, unsigned char* handle64) {
    QFX_ARG_CHECK(bytes > 0 && ptr && handle64, QFX_ERR_BAD_ARG, "qfx_symm_alloc: bad arguments");
    QFX_CUDA_TRY(cudaMalloc(ptr, bytes));
    QFX_CUDA_TRY(cudaMemset(*ptr, 0, bytes));
    cudaIpcMemHandle_t h;
    QFX_CUDA_TRY(cudaIpcGetMemHandle(&h, *ptr));
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle64, &h, 64);
    return 0;
}

extern "C" int qfx_symm_open(const unsigned char* handle64, void** ptr) {
    QFX_ARG_CHECK(ptr && handle64, QFX_ERR_BAD_ARG, "qfx_symm_open: bad arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    QFX_CUDA_TRY(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}

extern "C" int qfx_symm_close(void* ptr) {
    QFX_CUDA_TRY(cudaIpcCloseMemHandle(ptr));
    return 0;
}

extern "C" int qfx_symm_free(void* ptr) {
    QFX_CUDA_TRY(cudaFree(ptr));
    return 0;
}

// Same-process multi-GPU (tests, single-process launchers): enable peer access both ways.
extern "C" int qfx_enable_peer(int dev, int peer) {
    int can = 0;
    QFX_CUDA_TRY(cudaDeviceCanAccessPeer(&can, dev, peer));
    QFX_ARG_CHECK(can, QFX_ERR_UNSUPPORTED, "device %d cannot access peer %d", dev, peer);
    int cur = 0;
    QFX_CUDA_TRY(cudaGetDevice(&cur));
    QFX_CUDA_TRY(cudaSetDevice(dev));
    cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
    cudaSetDevice(cur);
    QFX_CUDA_TRY(e);
    return 0;
}
