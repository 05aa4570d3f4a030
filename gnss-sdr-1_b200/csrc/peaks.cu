// FP32 FMA peak micro-benchmark (roofline denominator of the tracking kernel; MEASURED_PEAKS.json has no FP32 entry).
#include "gb200_internal.cuh"

namespace
{
__global__ void __launch_bounds__(512) gb_fma_peak_kernel(float* out, int iters, float a, float b)
{
    float x0 = threadIdx.x * 1e-3f, x1 = x0 + 1.f, x2 = x0 + 2.f, x3 = x0 + 3.f, x4 = x0 + 4.f, x5 = x0 + 5.f, x6 = x0 + 6.f, x7 = x0 + 7.f;
#pragma unroll 1
    for (int i = 0; i < iters; i++)
        {
#pragma unroll
            for (int k = 0; k < 8; k++)
                {
                    x0 = fmaf(x0, a, b);
                    x1 = fmaf(x1, a, b);
                    x2 = fmaf(x2, a, b);
                    x3 = fmaf(x3, a, b);
                    x4 = fmaf(x4, a, b);
                    x5 = fmaf(x5, a, b);
                    x6 = fmaf(x6, a, b);
                    x7 = fmaf(x7, a, b);
                }
        }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}
}  // namespace

// returns flops executed; caller times the launch
double gb_fma_peak_launch(float* d_out, int blocks, int iters, cudaStream_t st)
{
    gb_fma_peak_kernel<<<blocks, 512, 0, st>>>(d_out, iters, 0.999f, 1e-3f);
    return (double)blocks * 512.0 * (double)iters * 64.0 * 2.0;
}
