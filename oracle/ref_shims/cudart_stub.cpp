// oracle/ref_shims/cudart_stub.cpp -- TEST INFRASTRUCTURE ONLY.
// The host-compiled reference calls a handful of CUDA runtime entry points for "managed" memory
// (Texture.cu:14,25; Managed.cuh:6-16; CudaCommon.cu:10; Scene.cu:321). The oracle runs on the CPU
// with no CUDA runtime linked, so they are satisfied by malloc/memcpy/free here.
#include <cuda_runtime_api.h>
#include <cstdlib>
#include <cstring>

extern "C" {
cudaError_t cudaMallocManaged(void** p, size_t n, unsigned int) { *p = malloc(n ? n : 1); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
cudaError_t cudaMalloc(void** p, size_t n) { *p = malloc(n ? n : 1); return *p ? cudaSuccess : cudaErrorMemoryAllocation; }
cudaError_t cudaFree(void* p) { free(p); return cudaSuccess; }
cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { memcpy(d, s, n); return cudaSuccess; }
cudaError_t cudaGetLastError(void) { return cudaSuccess; }
cudaError_t cudaPeekAtLastError(void) { return cudaSuccess; }
cudaError_t cudaDeviceSynchronize(void) { return cudaSuccess; }
cudaError_t cudaDeviceReset(void) { return cudaSuccess; }
const char* cudaGetErrorString(cudaError_t) { return "cudart stub"; }
const char* cudaGetErrorName(cudaError_t) { return "cudart stub"; }
}
