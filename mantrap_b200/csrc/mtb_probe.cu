// Peak probes used by bench.py for the roofline denominators MEASURED_PEAKS.json does not carry: dependent-chain
// FP32 FMA throughput and MUFU (ex2 / rsqrt) throughput of the device the benchmark runs on.
#include <cuda_runtime.h>

#include "mantrap_b200.h"

namespace mtb {

__global__ void __launch_bounds__(256) fma_probe_kernel(float* out, int iters) {
    float a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = 1.0f + 1e-3f * (float)(threadIdx.x + j);
    const float b = 0.999f, c = 1e-3f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] = fmaf(a[j], b, c);
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += a[j];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) mufu_probe_kernel(float* out, int iters) {
    float a[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = 1.0f + 1e-3f * (float)(threadIdx.x + j);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] = rsqrtf(a[j]);
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += a[j];
    out[(size_t)blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace mtb

extern "C" int mtb_probe_launch(int kind, float* out, int grid, int iters, void* stream) {
    if (out == nullptr || grid < 1 || iters < 1) return MTB_ERR_BAD_ARGUMENT;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (kind == 0) mtb::fma_probe_kernel<<<grid, 256, 0, st>>>(out, iters);
    else mtb::mufu_probe_kernel<<<grid, 256, 0, st>>>(out, iters);
    return (int)cudaGetLastError();
}
