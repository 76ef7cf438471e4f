// ubench_fma.cu -- FP32 pipe micro-benchmark for sm_100a: scalar FFMA vs packed FFMA2/FADD2/FMUL2
// issue rates, plus shared-memory LDS.64 vs LDS.128 bandwidth.  Prints per-SM lanes/clk so the
// FFT kernels' floors in DESIGN.md are measured, not assumed.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_fma ubench_fma.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(1024) k_fma(float *out, int iters, float seed) {
    float2 a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = make_float2(seed + i + threadIdx.x, seed - i);
    const float2 m = make_float2(1.000001f, 0.999999f), c = make_float2(1e-7f, -1e-7f);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (MODE == 0) { a[i].x = fmaf(a[i].x, m.x, c.x); a[i].y = fmaf(a[i].y, m.y, c.y); }
                if (MODE == 1) a[i] = __ffma2_rn(a[i], m, c);
                if (MODE == 2) a[i] = __fadd2_rn(a[i], c);
                if (MODE == 3) a[i] = __fmul2_rn(a[i], m);
                if (MODE == 4) { a[i].x = a[i].x + c.x; a[i].y = a[i].y + c.y; }
                if (MODE == 5) { a[i].x = fmaf(a[i].x, m.x, a[(i + 1) & 7].y); a[i].y = fmaf(a[i].y, m.y, a[(i + 3) & 7].x); }
                if (MODE == 6) a[i] = __ffma2_rn(a[i], m, a[(i + 1) & 7]);
            }
        }
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += a[i].x + a[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int VEC>
__global__ void __launch_bounds__(512) k_lds(float *out, int iters) {
    extern __shared__ float4 sm[];
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) sm[i] = make_float4(i, 1, 2, 3);
    __syncthreads();
    float acc = 0;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < 16; k++) {
            int idx = (threadIdx.x + k * 512 + it) & 8191;
            if (VEC == 4) { float4 v = sm[idx]; acc += v.x + v.y + v.z + v.w; }
            else { float2 v = reinterpret_cast<float2 *>(sm)[idx * 2 + (k & 1)]; acc += v.x + v.y; }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    float *out; cudaMalloc(&out, 148 * 2048 * 4 * 2);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const char *names[] = { "FFMA scalar (imm-free)", "FFMA2 packed", "FADD2 packed", "FMUL2 packed", "FADD scalar", "FFMA scalar 3-reg", "FFMA2 3-reg" };
    const int iters = 4096;
    for (int threads : { 512, 1024 }) {
        for (int mode = 0; mode < 7; mode++) {
            for (int rep = 0; rep < 2; rep++) {
                cudaEventRecord(e0);
                switch (mode) {
                    case 0: k_fma<0><<<p.multiProcessorCount, threads>>>(out, iters, 1.f); break;
                    case 1: k_fma<1><<<p.multiProcessorCount, threads>>>(out, iters, 1.f); break;
                    case 2: k_fma<2><<<p.multiProcessorCount, threads>>>(out, iters, 1.f); break;
                    case 3: k_fma<3><<<p.multiProcessorCount, threads>>>(out, iters, 1.f); break;
                    case 4: k_fma<4><<<p.multiProcessorCount, threads>>>(out, iters, 1.f); break;
                    case 5: k_fma<5><<<p.multiProcessorCount, threads>>>(out, iters, 1.f); break;
                    case 6: k_fma<6><<<p.multiProcessorCount, threads>>>(out, iters, 1.f); break;
                }
                cudaEventRecord(e1); cudaEventSynchronize(e1);
            }
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            double lane_ops = (double)p.multiProcessorCount * threads * iters * 4 * 8 * 2;   // fp32 lane-ops
            double per_sm_clk = lane_ops / (ms * 1e-3) / p.multiProcessorCount / (clk_khz * 1e3);
            printf("threads/SM %4d  %-24s %8.3f ms  %7.1f fp32 lane-ops/clk/SM (at %d MHz nominal)  %.1f Tlane-op/s\n",
                   threads, names[mode], ms, per_sm_clk, clk_khz / 1000, lane_ops / (ms * 1e-3) / 1e12);
        }
    }
    cudaFuncSetAttribute(k_lds<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072);
    cudaFuncSetAttribute(k_lds<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 131072);
    for (int vec : { 2, 4 }) {
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            if (vec == 4) k_lds<4><<<p.multiProcessorCount, 512, 131072>>>(out, 2048);
            else k_lds<2><<<p.multiProcessorCount, 512, 131072>>>(out, 2048);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
        }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        double bytes = (double)p.multiProcessorCount * 512 * 2048 * 16 * (vec * 4);
        printf("LDS.%d  %8.3f ms  %.1f B/clk/SM\n", vec * 32, ms, bytes / (ms * 1e-3) / p.multiProcessorCount / (clk_khz * 1e3));
    }
    printf("err: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
