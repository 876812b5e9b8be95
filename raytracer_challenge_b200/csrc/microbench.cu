// Roofline denominators that MEASURED_PEAKS.json does not carry: FP64 / FP32 CUDA-core issue
// rates with and without multiply-add fusion, and the L2-resident read bandwidth.  Timed with CUDA events.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/rtc.h"

namespace {

__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }  // never contracted
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }

template <typename T, bool FUSED>
__global__ void flops_kernel(T* out, int iters, T a, T b) {
    T x0 = (T)threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
        if (FUSED) {
            x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
            x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
        } else {  // separately rounded multiply and add, as the parity-critical code is compiled (-fmad=false)
            x0 = add_rn(mul_rn(x0, a), b); x1 = add_rn(mul_rn(x1, a), b); x2 = add_rn(mul_rn(x2, a), b); x3 = add_rn(mul_rn(x3, a), b);
            x4 = add_rn(mul_rn(x4, a), b); x5 = add_rn(mul_rn(x5, a), b); x6 = add_rn(mul_rn(x6, a), b); x7 = add_rn(mul_rn(x7, a), b);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}

__global__ void l2_read_kernel(const float4* __restrict__ src, size_t n, int reps, float* sink) {
    float acc = 0.f;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int r = 0; r < reps; ++r)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
            const float4 v = __ldcg(src + i);
            acc += v.x + v.y + v.z + v.w;
        }
    if (acc == 123.456f) *sink = acc;
}

template <typename T, bool FUSED>
int run_flops(double* result) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int block = 256, grid = sms * 8, iters = 1 << 14;
    T* out = nullptr;
    if (cudaMalloc((void**)&out, sizeof(T) * (size_t)block * grid) != cudaSuccess) return 1;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    flops_kernel<T, FUSED><<<grid, block>>>(out, 64, (T)1.0000001, (T)1e-7);
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        flops_kernel<T, FUSED><<<grid, block>>>(out, iters, (T)1.0000001, (T)1e-7);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    if (cudaGetLastError() != cudaSuccess) return 1;
    // 8 chains x (1 mul + 1 add) per iteration per thread
    *result = 16.0 * iters * (double)block * grid / (best * 1e-3) / 1e12;  // TFLOP/s
    return 0;
}

}  // namespace

extern "C" int rtc_microbench(int what, double* result) {
    if (!result) return 1;
    switch (what) {
        case RTC_MB_FP64_FMA: return run_flops<double, true>(result);
        case RTC_MB_FP64_MUL_ADD: return run_flops<double, false>(result);
        case RTC_MB_FP32_FMA: return run_flops<float, true>(result);
        case RTC_MB_FP32_MUL_ADD: return run_flops<float, false>(result);
        case RTC_MB_L2_READ: {
            const size_t bytes = (size_t)48 << 20;  // well inside the 126 MB L2
            float4* buf = nullptr;
            float* sink = nullptr;
            if (cudaMalloc((void**)&buf, bytes) != cudaSuccess || cudaMalloc((void**)&sink, 4) != cudaSuccess) return 1;
            cudaMemset(buf, 0, bytes);
            int dev = 0, sms = 0;
            cudaGetDevice(&dev);
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0);
            cudaEventCreate(&e1);
            l2_read_kernel<<<sms * 8, 512>>>(buf, bytes / 16, 2, sink);
            float best = 1e30f;
            const int reps = 20;
            for (int rep = 0; rep < 3; ++rep) {
                cudaEventRecord(e0);
                l2_read_kernel<<<sms * 8, 512>>>(buf, bytes / 16, reps, sink);
                cudaEventRecord(e1);
                cudaEventSynchronize(e1);
                float ms = 0.f;
                cudaEventElapsedTime(&ms, e0, e1);
                if (ms < best) best = ms;
            }
            cudaEventDestroy(e0);
            cudaEventDestroy(e1);
            cudaFree(buf);
            cudaFree(sink);
            if (cudaGetLastError() != cudaSuccess) return 1;
            *result = (double)bytes * reps / (best * 1e-3) / 1e9;  // GB/s
            return 0;
        }
        default: return 1;
    }
}
