// Host<->device copy ceilings for the e2e path: linear vs pitched (cudaMemcpy2DAsync) read-back,
// chunked, both directions at once.  nvcc -O2 -o tools/pcie_probe tools/pcie_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <chrono>
int main() {
    const size_t nb = 1024, plen = 104 + 22 * 16384, pitch = (8 + plen + 15) / 16 * 16, inrow = 22 * 16384;
    uint8_t *h_in, *h_out, *d_in, *d_out;
    cudaMallocHost(&h_in, nb * inrow); cudaMallocHost(&h_out, nb * plen);
    cudaMalloc(&d_in, nb * inrow); cudaMalloc(&d_out, nb * pitch + 16);
    cudaStream_t s1, s2; cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking); cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking);
    for (int mode = 0; mode < 4; mode++) {
        const size_t chunk = (mode & 1) ? 37 : 148;
        const bool twod = mode & 2;
        double best = 1e9;
        for (int rep = 0; rep < 6; rep++) {
            cudaDeviceSynchronize();
            auto t0 = std::chrono::steady_clock::now();
            for (size_t b0 = 0; b0 < nb; b0 += chunk) {
                size_t cb = b0 + chunk < nb ? chunk : nb - b0;
                cudaMemcpyAsync(d_in + b0 * inrow, h_in + b0 * inrow, cb * inrow, cudaMemcpyHostToDevice, s1);
                if (twod) cudaMemcpy2DAsync(h_out + b0 * plen, plen, d_out + b0 * pitch + 8, pitch, plen, cb, cudaMemcpyDeviceToHost, s2);
                else cudaMemcpyAsync(h_out + b0 * plen, d_out + b0 * plen, cb * plen, cudaMemcpyDeviceToHost, s2);
            }
            cudaDeviceSynchronize();
            double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            if (ms < best) best = ms;
        }
        printf("chunk %3zu %s read-back: %.2f ms (%.1f GB/s per direction)\n", chunk, twod ? "pitched 2D" : "linear", best, nb * inrow / best / 1e6);
    }
    return 0;
}
