// frac_kernel.cu -- fractional part of the delay by parabolic interpolation of the correlation peak
// (NOT on the parity path).
//
// The reference computes this in ccoherent::computelag (src/ccoherent.cc:206-218) from the three
// |c|^2 values around the peak, but with a sign slip in the curvature term (a = -y[-1] - 2 y[0] + y[1])
// and then leaves the result unused ("disabled fractional sync, obviously it's not doing what it's
// supposed to do").  Here it is enabled with the textbook three-point parabola
//     D = 0.5 (y[-1] - y[+1]) / (y[-1] - 2 y[0] + y[+1]),      y[k] = |c[lag + k]|^2,
// under the reference's own guard (peak index strictly inside the buffer, :206).
// The three correlation values are NOT taken from the float FFT: they are recomputed in the time
// domain, exactly, in integers (the samples are int8):
//     c[m] = sum_n x[n + m] conj(r[n]),   0 <= n < L,  0 <= n + m < L
// which is what the zero-padded circular correlation of src/crtlsdr.cc:205-218 equals for |m| < L.
// One warp per (block, channel), two samples per lane and step with dp4a; a side computation of
// SYNC mode (three length-L integer dot products per item against ~2.5 MFLOP of transforms).
#include "coherent_cuda.h"
#include "kernels.h"

namespace coh {

constexpr int kFracWarps = 8;

// iq [blocks][nch][B]; lag [blocks][nch] (float, integer-valued: what the batch reported);
// frac [blocks][nch] float; y3 [blocks][nch][3] double (optional): |c[lag-1]|^2, |c[lag]|^2, |c[lag+1]|^2
__global__ void __launch_bounds__(kFracWarps * 32)
k_fraclag(const uint8_t *__restrict__ iq, uint32_t nch, uint32_t B, uint32_t n_items, uint32_t xormask,
          const float *__restrict__ lag, float *__restrict__ frac, double *__restrict__ y3) {
    const uint32_t item = blockIdx.x * kFracWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (item >= n_items) return;
    const uint32_t b = item / nch, c = item - b * nch;
    const int L = (int)(B >> 1);
    if (c == 0) {                                                  // reference column: nothing to estimate
        if (lane == 0) { frac[item] = 0.0f; if (y3) { y3[3 * (size_t)item] = y3[3 * (size_t)item + 1] = y3[3 * (size_t)item + 2] = 0.0; } }
        return;
    }
    const uint32_t *xw = reinterpret_cast<const uint32_t *>(iq + (size_t)item * B);
    const uint32_t *rw = reinterpret_cast<const uint32_t *>(iq + (size_t)b * nch * B);
    const int m0 = (int)lag[item];
    const int W = L >> 1;                                          // 32-bit words (two samples) per row
    const uint32_t xm = xormask | (xormask << 16);
    // word w of the row in two's complement, zero outside the row (zero samples add nothing)
    auto word = [&](const uint32_t *row, int w) -> uint32_t { return (w >= 0 && w < W) ? (__ldg(row + w) ^ xm) : 0u; };
    // Two samples per lane and step.  A lag of s samples is s/2 words for even s; for odd s the pair
    // straddles two words (funnel shift by 16 bits).  With t = floor(m0 / 2):
    //   m0 even: lags m0-1, m0, m0+1 = (A>>16 | B<<16), B, (B>>16 | C<<16)   with A, B, C = words j+t-1, j+t, j+t+1
    //   m0 odd : lags m0-1, m0, m0+1 = B, (B>>16 | C<<16), C                  with    B, C = words j+t, j+t+1
    const int t = (m0 >= 0) ? (m0 >> 1) : -((-m0 + 1) >> 1);       // floor division
    const bool odd = (m0 - 2 * t) != 0;
    int re[3] = {0, 0, 0}, tt[3] = {0, 0, 0}, uu[3] = {0, 0, 0};
    for (int j = lane; j < W; j += 32) {
        const uint32_t r2 = __ldg(rw + j) ^ xm;
        const uint32_t Bw = word(xw, j + t), Cw = word(xw, j + t + 1);
        uint32_t x3[3];
        if (odd) { x3[0] = Bw; x3[1] = __funnelshift_r(Bw, Cw, 16); x3[2] = Cw; }
        else { const uint32_t Aw = word(xw, j + t - 1); x3[0] = __funnelshift_r(Aw, Bw, 16); x3[1] = Bw; x3[2] = __funnelshift_r(Bw, Cw, 16); }
#pragma unroll
        for (int k = 0; k < 3; k++) {
            // bytes [xr0 xi0 xr1 xi1] . [rr0 ri0 rr1 ri1]:  x conj(r) = (xr rr + xi ri) + j (xi rr - xr ri)
            re[k] = __dp4a((int)x3[k], (int)r2, re[k]);
            const uint32_t xs = __byte_perm(x3[k], 0, 0x2301);     // [xi0 xr0 xi1 xr1]
            tt[k] = __dp4a((int)xs, (int)r2, tt[k]);               // xi rr + xr ri
            uu[k] = __dp4a((int)(xs & 0xFF00FF00u), (int)r2, uu[k]);   // xr ri
        }
    }
    long long acc[3][2];
#pragma unroll
    for (int k = 0; k < 3; k++) { acc[k][0] = re[k]; acc[k][1] = (long long)tt[k] - 2ll * (long long)uu[k]; }
#pragma unroll
    for (int k = 0; k < 3; k++)
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            acc[k][0] += __shfl_down_sync(0xffffffffu, acc[k][0], off);
            acc[k][1] += __shfl_down_sync(0xffffffffu, acc[k][1], off);
        }
    if (lane == 0) {
        double y[3];
#pragma unroll
        for (int k = 0; k < 3; k++) y[k] = (double)acc[k][0] * (double)acc[k][0] + (double)acc[k][1] * (double)acc[k][1];
        const int idx = m0 + L;                                    // index into the 2L-point correlation
        double D = 0.0;
        if (idx > 1 && idx < 2 * L - 1) {                          // src/ccoherent.cc:206
            const double a = y[0] - 2.0 * y[1] + y[2];
            if (a != 0.0) D = 0.5 * (y[0] - y[2]) / a;
        }
        frac[item] = (float)D;
        if (y3) { y3[3 * (size_t)item] = y[0]; y3[3 * (size_t)item + 1] = y[1]; y3[3 * (size_t)item + 2] = y[2]; }
    }
}

cudaError_t launch_fraclag(const uint8_t *iq, uint32_t nch, uint32_t B, uint32_t n_blocks, uint32_t xormask,
                           const float *lag, float *frac, double *y3, cudaStream_t st) {
    const uint32_t n_items = n_blocks * nch;
    if (n_items == 0) return cudaSuccess;
    k_fraclag<<<(n_items + kFracWarps - 1) / kFracWarps, kFracWarps * 32, 0, st>>>(iq, nch, B, n_items, xormask, lag, frac, y3);
    return cudaGetLastError();
}

}  // namespace coh
