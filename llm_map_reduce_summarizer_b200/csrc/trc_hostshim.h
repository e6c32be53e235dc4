// TEST INFRASTRUCTURE ONLY.  Lets g++ compile trchunk.cu (-DTRC_HOSTSIM -x c++) into
// tests/hostsim/libtrchunk_hostsim.so so that the CPU test-suite can run every stage's logic and the
// host orchestration against the oracle in a container without a GPU.  The package never loads that
// library; libtrchunk.so (nvcc, sm_100a) is the only product binary.
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
typedef void* cudaStream_t;
typedef void* cudaEvent_t;
typedef int cudaError_t;
enum { cudaSuccess = 0 };
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice };
enum { cudaStreamNonBlocking = 1 };
struct cudaDeviceProp { int major; };
static inline const char* cudaGetErrorString(cudaError_t) { return "hostsim error"; }
static inline cudaError_t cudaMalloc(void* pp, size_t n) { *(void**)pp = malloc(n ? n : 1); return *(void**)pp ? 0 : 2; }
template <class T> static inline cudaError_t cudaMalloc(T** pp, size_t n) { *pp = (T*)malloc(n ? n : 1); return *pp ? 0 : 2; }
static inline cudaError_t cudaFree(void* p) { free(p); return 0; }
template <class T> static inline cudaError_t cudaMallocHost(T** pp, size_t n) { *pp = (T*)malloc(n); return 0; }
enum { cudaHostAllocMapped = 2 };
static inline cudaError_t cudaHostAlloc(void** pp, size_t n, int) { *pp = malloc(n); return 0; }
static inline cudaError_t cudaHostGetDevicePointer(void** d, void* h, int) { *d = h; return 0; }
static inline cudaError_t cudaFreeHost(void* p) { free(p); return 0; }
static inline cudaError_t cudaMemcpy(void* d, const void* s, size_t n, cudaMemcpyKind) { memcpy(d, s, n); return 0; }
static inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t) { memcpy(d, s, n); return 0; }
static inline cudaError_t cudaMemsetAsync(void* d, int v, size_t n, cudaStream_t) { memset(d, v, n); return 0; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return 0; }
static inline cudaError_t cudaDeviceSynchronize() { return 0; }
static inline cudaError_t cudaSetDevice(int) { return 0; }
static inline cudaError_t cudaGetDeviceCount(int* n) { *n = 1; return 0; }
static inline cudaError_t cudaGetDeviceProperties(cudaDeviceProp* p, int) { p->major = 10; return 0; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, int) { *s = nullptr; return 0; }
static inline cudaError_t cudaStreamDestroy(cudaStream_t) { return 0; }
static inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = nullptr; return 0; }
enum { cudaEventDisableTiming = 2 };
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, int) { *e = nullptr; return 0; }
static inline cudaError_t cudaEventDestroy(cudaEvent_t) { return 0; }
static inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t) { return 0; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return 0; }
static inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t, cudaEvent_t) { *ms = 0; return 0; }
static inline cudaError_t cudaGetLastError() { return 0; }
template <class T, class U> static inline T atomicAdd(T* p, U v) { T o = *p; *p = (T)(o + v); return o; }
template <class T> static inline T atomicMin(T* p, T v) { T o = *p; if (v < o) *p = v; return o; }
#define __global__ static
#define __restrict__
#define __launch_bounds__(x)
