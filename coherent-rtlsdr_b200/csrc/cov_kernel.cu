// cov_kernel.cu -- sample covariance of the coherent matrix, the first thing the reference's clients
// compute from a packet (NOT on the parity path):
//     Rxx = X^H X / rows      beamformclient/heatmap2d2.cpp:199,  matlabclient/functions/pmusic.m:12-14
// The packet columns are int8 (I,Q) pairs, so the sums are formed exactly in integers with dp4a:
//     S[i][j] = sum_n conj(x_i[n]) x_j[n]   (int64 re, im; i <= j computed, the rest by symmetry on the host)
//     m[i]    = sum_n x_i[n]                 (int32 re, im; the clients subtract the column mean first)
// and the host turns them into the clients' float Rxx (mean removal, 1/128 scaling, reference
// column dropped).  Two kernels, one CTA per block each:
//   k_covariance_mma  up to 22 columns (the reference's URA21 packet): int8 tensor-core MMA, below
//   k_covariance      any column count: columns staged tile by tile in shared memory, each thread owns
//                     (i, j) pairs and walks the tile with LDS.128 + dp4a (INT-pipe bound: 253 pairs x
//                     3 dp4a per 2 samples, 0.43 ms per 1024 URA21 blocks against 0.06 ms of HBM time)
#include "coherent_cuda.h"
#include "kernels.h"
#include <cstdlib>

namespace coh {

constexpr int kCovThreads = 256;

__device__ __forceinline__ void cov_word(uint32_t a, uint32_t b, int &re, int &t, int &u) {
    // a, b = [I0 Q0 I1 Q1] of columns i, j:  conj(a) b = (Ii Ij + Qi Qj) + j (Ii Qj - Qi Ij)
    re = __dp4a((int)a, (int)b, re);
    const uint32_t asw = __byte_perm(a, 0, 0x2301);               // [Q0 I0 Q1 I1]
    t = __dp4a((int)asw, (int)b, t);                              // sum Qi Ij + Ii Qj
    u = __dp4a((int)(asw & 0x00FF00FFu), (int)b, u);              // sum Qi Ij          -> im = t - 2u
}

// packets: device, block b at packets + b*pitch + lead, columns at + hdr + c*B.  S: [blocks][nch][nch][2]
// int64 (zeroed by the launcher), m: [blocks][nch][2] int32.
__global__ void __launch_bounds__(kCovThreads)
k_covariance(const int8_t *__restrict__ packets, size_t pitch, size_t lead, size_t hdr, uint32_t nch, uint32_t B,
             uint32_t tile_bytes, long long *__restrict__ S, int *__restrict__ m) {
    extern __shared__ __align__(16) uint4 tile[];                 // [nch][tile_bytes/16 + 1]
    const uint32_t b = blockIdx.x, tid = threadIdx.x;
    const int8_t *pkt = packets + (size_t)b * pitch + lead + hdr;
    const uint32_t tw = tile_bytes / 16, row = tw + 1;            // uint4 per column per tile, padded row
    const uint32_t n_pairs = nch * (nch + 1) / 2;
    long long *Sb = S + (size_t)b * nch * nch * 2;

    for (uint32_t t0 = 0; t0 < B; t0 += tile_bytes) {
        __syncthreads();                                          // previous tile fully consumed
        for (uint32_t q = tid; q < nch * tw; q += kCovThreads) {
            const uint32_t c = q / tw, k = q - c * tw;
            tile[c * row + k] = __ldg(reinterpret_cast<const uint4 *>(pkt + (size_t)c * B + t0) + k);
        }
        __syncthreads();
        for (uint32_t p = tid; p < n_pairs; p += kCovThreads) {
            // p -> (i, j), i <= j, row-major over the upper triangle
            uint32_t i = 0, rem = p;
            while (rem >= nch - i) { rem -= nch - i; i++; }
            const uint32_t j = i + rem;
            const uint4 *yi = tile + i * row, *yj = tile + j * row;
            int re = 0, t = 0, u = 0, si = 0, sq = 0;
            for (uint32_t k = 0; k < tw; k++) {
                const uint4 a = yi[k], c4 = yj[k];
                cov_word(a.x, c4.x, re, t, u); cov_word(a.y, c4.y, re, t, u);
                cov_word(a.z, c4.z, re, t, u); cov_word(a.w, c4.w, re, t, u);
                if (i == j) {
                    si = __dp4a((int)a.x, 0x00010001, si); sq = __dp4a((int)a.x, 0x01000100, sq);
                    si = __dp4a((int)a.y, 0x00010001, si); sq = __dp4a((int)a.y, 0x01000100, sq);
                    si = __dp4a((int)a.z, 0x00010001, si); sq = __dp4a((int)a.z, 0x01000100, sq);
                    si = __dp4a((int)a.w, 0x00010001, si); sq = __dp4a((int)a.w, 0x01000100, sq);
                }
            }
            // one owner per (block, pair): plain read-modify-write, tiles are sequential in the CTA
            long long *o = Sb + ((size_t)i * nch + j) * 2;
            o[0] += (long long)re;
            o[1] += (long long)t - 2ll * (long long)u;
            if (i == j) { int *mo = m + ((size_t)b * nch + i) * 2; mo[0] += si; mo[1] += sq; }
        }
    }
}


// ---- tensor-core version for up to 22 columns (the reference's URA21 packet) --------------------
// G = Y Y^T with Y the 22 x B int8 matrix of packet columns is three int8 GEMMs over K = B bytes:
//   re = Y Y^T,   t = Ysw Y^T,   u = (Ysw & 0x00FF00FF) Y^T,   im = t - 2u      (Ysw = I/Q swapped)
// done with mma.sync.m16n8k32.s8 on a 32 x 24 padded tile.  Rows of A and columns of B are the same
// bytes, so the B fragments are reused A fragments, and only the four 16 x 8 tiles that touch the
// upper triangle are computed: three 16-byte loads feed 24 MMAs per 64 bytes of K.  The two padding columns 22 and 23 carry the constant patterns [1 0 1 0] / [0 1 0 1], which
// makes re[i][22], re[i][23] the column sums m_i for free.  Eight warps split K, accumulate in
// int32 registers and merge through 64-bit shared-memory atomics.  Exact, like the dp4a kernel.
__device__ __forceinline__ void mma_s8(int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

constexpr int kMmaWarps = 8;
__global__ void __launch_bounds__(kMmaWarps * 32, 2)
k_covariance_mma(const int8_t *__restrict__ packets, size_t pitch, size_t lead, size_t hdr, uint32_t nch, uint32_t B,
                 long long *__restrict__ S, int *__restrict__ m) {
    __shared__ unsigned long long acc64[3][32][24];               // re, t, u
    const uint32_t b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t g = lane >> 2, t4 = lane & 3;
    for (uint32_t q = tid; q < 3 * 32 * 24; q += kMmaWarps * 32) (&acc64[0][0][0])[q] = 0ull;
    __syncthreads();
    const int8_t *pkt = packets + (size_t)b * pitch + lead + hdr;
    // rows this thread loads: g, g+8 (tile 0) and 16+g (tile 1; rows >= nch are zero, 24..31 never loaded)
    const uint32_t r0 = g, r1 = g + 8, r2 = g + 16;
    const bool v0 = r0 < nch, v1 = r1 < nch, v2 = r2 < nch;
    // K is a summation index: any assignment of bytes to k positions is valid as long as A and B
    // agree, and they do because both come from the same registers.  So each thread fetches 16
    // contiguous bytes of its rows per 64-byte chunk (one LDG.128, full sectors) and feeds words
    // (x, y) to one MMA k-step and (z, w) to the next; only the I/Q pairing inside a word matters.
    const uint4 *p0 = reinterpret_cast<const uint4 *>(pkt + (size_t)(v0 ? r0 : 0) * B) + t4;
    const uint4 *p1 = reinterpret_cast<const uint4 *>(pkt + (size_t)(v1 ? r1 : 0) * B) + t4;
    const uint4 *p2 = reinterpret_cast<const uint4 *>(pkt + (size_t)(v2 ? r2 : 0) * B) + t4;

    const uint32_t chunks = B / 64, per_warp = (chunks + kMmaWarps - 1) / kMmaWarps;
    const uint32_t k_begin = warp * per_warp, k_end = (k_begin + per_warp < chunks) ? k_begin + per_warp : chunks;
    // flush to 64 bits before an int32 accumulator can overflow: 64 bytes * 128 * 128 per chunk
    constexpr uint32_t kFlush = 1024;
    const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);

    for (uint32_t kb = k_begin; kb < k_end; kb += kFlush) {
        int c[3][2][3][4] = {};                                    // [variant][m tile][n tile][frag]
        const uint32_t ke = (kb + kFlush < k_end) ? kb + kFlush : k_end;
#pragma unroll 2
        for (uint32_t k = kb; k < ke; k++) {
            const uint4 q0 = v0 ? __ldg(p0 + k * 4) : z4, q1 = v1 ? __ldg(p1 + k * 4) : z4, q2 = v2 ? __ldg(p2 + k * 4) : z4;
#pragma unroll
            for (int h = 0; h < 2; h++) {
                uint32_t a[2][4];
                a[0][0] = h ? q0.z : q0.x; a[0][2] = h ? q0.w : q0.y;
                a[0][1] = h ? q1.z : q1.x; a[0][3] = h ? q1.w : q1.y;
                a[1][0] = h ? q2.z : q2.x; a[1][2] = h ? q2.w : q2.y;
                a[1][1] = 0u;              a[1][3] = 0u;
                // B fragments = the same bytes: columns g (n tile 0), 8 + g (n tile 1), 16 + g (n tile 2)
                uint32_t bq[3][2] = { { a[0][0], a[0][2] }, { a[0][1], a[0][3] }, { a[1][0], a[1][2] } };
                if (r2 == 22) { bq[2][0] = 0x00010001u; bq[2][1] = 0x00010001u; }     // sum I
                if (r2 == 23) { bq[2][0] = 0x01000100u; bq[2][1] = 0x01000100u; }     // sum Q
#pragma unroll
                for (int mt = 0; mt < 2; mt++) {
                    uint32_t asw[4], amk[4];
#pragma unroll
                    for (int r = 0; r < 4; r++) { asw[r] = __byte_perm(a[mt][r], 0, 0x2301); amk[r] = asw[r] & 0x00FF00FFu; }
#pragma unroll
                    for (int nt = 0; nt < 3; nt++) {
                        if (mt == 1 && nt < 2) continue;          // rows 16.. x columns ..15: below the diagonal
                        mma_s8(c[0][mt][nt], a[mt], bq[nt][0], bq[nt][1]);
                        mma_s8(c[1][mt][nt], asw, bq[nt][0], bq[nt][1]);
                        mma_s8(c[2][mt][nt], amk, bq[nt][0], bq[nt][1]);
                    }
                }
            }
        }
        // C fragment: c0 (row g, col 2 t4), c1 (row g, col 2 t4 + 1), c2 (row g + 8, col 2 t4), c3 (row g + 8, col 2 t4 + 1)
#pragma unroll
        for (int vtn = 0; vtn < 3; vtn++)
#pragma unroll
            for (int mt = 0; mt < 2; mt++)
#pragma unroll
                for (int nt = 0; nt < 3; nt++)
#pragma unroll
                    for (int f = 0; f < 4; f++) {
                        if (mt == 1 && nt < 2) continue;
                        const int row = mt * 16 + g + (f >> 1) * 8, col = nt * 8 + 2 * t4 + (f & 1);
                        const int v = c[vtn][mt][nt][f];
                        if (v != 0) atomicAdd(&acc64[vtn][row][col], (unsigned long long)(long long)v);
                    }
    }
    __syncthreads();
    long long *Sb = S + (size_t)b * nch * nch * 2;
    for (uint32_t q = tid; q < nch * nch; q += kMmaWarps * 32) {
        const uint32_t i = q / nch, j = q - i * nch;
        if (i > j) continue;
        Sb[(size_t)q * 2] = (long long)acc64[0][i][j];
        Sb[(size_t)q * 2 + 1] = (long long)acc64[1][i][j] - 2ll * (long long)acc64[2][i][j];
    }
    for (uint32_t i = tid; i < nch; i += kMmaWarps * 32) {
        m[((size_t)b * nch + i) * 2] = (int)(long long)acc64[0][i][22];
        m[((size_t)b * nch + i) * 2 + 1] = (int)(long long)acc64[0][i][23];
    }
}

cudaError_t launch_covariance(const int8_t *packets, size_t pitch, size_t lead, size_t hdr, uint32_t nch, uint32_t B,
                              uint32_t n_blocks, long long *S, int *m, cudaStream_t st) {
    if (n_blocks == 0) return cudaSuccess;
    if (nch <= 22 && B >= 64) {
        // every element is written by the kernel (i <= j and m); zero the rest for the host's symmetry fill
        cudaError_t e0 = cudaMemsetAsync(S, 0, (size_t)n_blocks * nch * nch * 2 * sizeof(long long), st);
        if (e0 != cudaSuccess) return e0;
        k_covariance_mma<<<n_blocks, kMmaWarps * 32, 0, st>>>(packets, pitch, lead, hdr, nch, B, S, m);
        return cudaGetLastError();
    }
    // largest power-of-two tile that keeps nch columns within ~160 KB of shared memory
    uint32_t tile = 2048;
    while (tile > 16 && (size_t)nch * (tile + 16) > 160 * 1024) tile >>= 1;
    if (tile > B) tile = B;
    const size_t smem = (size_t)nch * (tile + 16);
    static bool attr_done[64] = {};
    int dev = 0; cudaGetDevice(&dev);
    if (!attr_done[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(k_covariance, cudaFuncAttributeMaxDynamicSharedMemorySize, 164 * 1024);
        if (e != cudaSuccess) return e;
        attr_done[dev & 63] = true;
    }
    cudaError_t e = cudaMemsetAsync(S, 0, (size_t)n_blocks * nch * nch * 2 * sizeof(long long), st);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(m, 0, (size_t)n_blocks * nch * 2 * sizeof(int), st);
    if (e != cudaSuccess) return e;
    k_covariance<<<n_blocks, kCovThreads, smem, st>>>(packets, pitch, lead, hdr, nch, B, tile, S, m);
    return cudaGetLastError();
}

}  // namespace coh
