// Set-up of the peer exchange boxes (CUDA IPC): see include/mercury_b200.h, "exchange over peer memory".
// The kernels that use the boxes live in cdf.cu (sum_exchange_kernel, select_exchange_kernel).
#include <cuda_runtime.h>
#include <string.h>

#include "../../include/mercury_b200.h"
#include "common.cuh"

#define MB_CHECK_CUDA(call)                                               \
    do {                                                                  \
        cudaError_t e__ = (call);                                         \
        if (e__ != cudaSuccess) return ::mercury::cuda_fail(e__, #call);  \
    } while (0)

static_assert(sizeof(cudaIpcMemHandle_t) == MERCURY_PEER_HANDLE_BYTES, "IPC handle size");

extern "C" int mercury_peer_create(void** box, unsigned char* handle) {
    MB_CHECK_ARG(box && handle, "mercury_peer_create: null pointer");
    void* p = nullptr;
    MB_CHECK_CUDA(cudaMalloc(&p, MERCURY_PEER_BOX_BYTES));
    cudaError_t e = cudaMemset(p, 0, MERCURY_PEER_BOX_BYTES);
    if (e == cudaSuccess && mercury_peer_set_wait(p, ::mercury::env_knob("MERCURY_PEER_WAIT_MS", 4000, 1, 3600000), nullptr) != MERCURY_OK)
        e = cudaErrorUnknown;
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        cudaFree(p);
        MB_CHECK_CUDA(e);
    }
    memcpy(handle, &h, sizeof(h));
    *box = p;
    return MERCURY_OK;
}

extern "C" int mercury_peer_open(const unsigned char* handle, void** peer_box) {
    MB_CHECK_ARG(handle && peer_box, "mercury_peer_open: null pointer");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    MB_CHECK_CUDA(cudaIpcOpenMemHandle(peer_box, h, cudaIpcMemLazyEnablePeerAccess));
    return MERCURY_OK;
}

extern "C" int mercury_peer_close(void* peer_box) {
    MB_CHECK_ARG(peer_box, "mercury_peer_close: null pointer");
    MB_CHECK_CUDA(cudaIpcCloseMemHandle(peer_box));
    return MERCURY_OK;
}

extern "C" int mercury_peer_destroy(void* box) {
    MB_CHECK_ARG(box, "mercury_peer_destroy: null pointer");
    MB_CHECK_CUDA(cudaFree(box));
    return MERCURY_OK;
}
