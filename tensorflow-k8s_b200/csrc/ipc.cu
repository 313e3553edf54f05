// ipc.cu -- device buffers that other ranks of the box can address (include/b200ps.h
// "Transport group").
//
// Reference: none -- between PS and WORKER tasks (k8s-operator.md:6) a TFJob moves tensors with
// TensorFlow's gRPC runtime over the pod network.  On one NVSwitch box the transport is the
// memory system itself: every rank (one process per GPU) exports its gradient and parameter
// buffers as CUDA IPC handles, every other rank maps them, and the step kernel (dense.cu) loads
// gradients from and stores parameters to the mapped peer pointers over NVLink.
#include <cuda_runtime.h>

#include <cstring>

#include "common.h"
#include "cuda_util.h"

static_assert(sizeof(cudaIpcMemHandle_t) == B200PS_IPC_HANDLE_BYTES, "IPC handle size");

extern "C" int b200ps_dev_alloc(size_t bytes, void** d_ptr) {
    B200PS_REQUIRE(d_ptr && bytes > 0, B200PS_ERR_INVALID, "dev_alloc needs a destination and bytes > 0");
    B200PS_CUDA(cudaMalloc(d_ptr, bytes));
    return B200PS_OK;
}

extern "C" int b200ps_dev_free(void* d_ptr) {
    if (d_ptr) B200PS_CUDA(cudaFree(d_ptr));
    return B200PS_OK;
}

extern "C" int b200ps_ipc_export(const void* d_ptr, void* handle_out) {
    B200PS_REQUIRE(d_ptr && handle_out, B200PS_ERR_INVALID, "null argument");
    cudaIpcMemHandle_t h;
    B200PS_CUDA(cudaIpcGetMemHandle(&h, const_cast<void*>(d_ptr)));
    memcpy(handle_out, &h, sizeof h);
    return B200PS_OK;
}

extern "C" int b200ps_ipc_open(const void* handle, void** d_ptr) {
    B200PS_REQUIRE(handle && d_ptr, B200PS_ERR_INVALID, "null argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    B200PS_CUDA(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return B200PS_OK;
}

extern "C" int b200ps_ipc_close(void* d_ptr) {
    if (d_ptr) B200PS_CUDA(cudaIpcCloseMemHandle(d_ptr));
    return B200PS_OK;
}
