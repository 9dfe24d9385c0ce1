// ref_shim.cu -- extern "C" doorway to the reference's own launchers, for parity runs on the GPU box.
// TEST INFRASTRUCTURE: compiled together with /root/reference/src/{DirectComputeKernels,BarnesHutKernels}.cu
// (unmodified, from where they lie) into oracle/_ref/libnbody_ref.so by oracle/Makefile.  Nothing from the
// reference is copied into this repository.
#include <cuda_runtime.h>
#include "DirectComputeKernels.cuh"   // direct2DCompute / direct3DCompute   (src/DirectComputeKernels.cuh:7,12)
#include "BarnesHutKernels.cuh"       // barnesHutCompute                     (src/BarnesHutKernels.cuh:5)

extern "C" {

// All pointers are device pointers owned by the caller, exactly as ComputeMethods.cpp:21,23,48 passes them.
int ref_direct(float* pos, float* acc, int n, int dim, float softening) {
    if (dim == 2) direct2DCompute((float4*)pos, (float4*)acc, n, softening);
    else          direct3DCompute((float4*)pos, (float4*)acc, n, softening);
    return (int)cudaGetLastError();
}

// pos must hold m = knodes*n float4 (bodies then node storage); nodes 4*m ints; bounds one float4 that the
// CALLER initialises (the reference never does).
int ref_barnes_hut(float* pos, float* acc, float* bounds, int* index, int* nodes, int* sortedIdx, int* keys,
                   int n, int m, float theta, float softening) {
    barnesHutCompute((float4*)pos, (float4*)acc, (float4*)bounds, index, nodes, sortedIdx, keys, n, m, theta,
                     softening);
    return (int)cudaDeviceSynchronize();
}

}
