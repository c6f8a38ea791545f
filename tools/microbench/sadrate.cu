// sadrate.cu -- microbenchmark: peak issue rate of the packed-byte SAD instruction (VABSDIFF4.U8.ACC) on B200,
// alone and mixed with the funnel shifts / shared-memory loads a SAD search needs.  Gives the INT-roofline
// denominator P_sad.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o sadrate sadrate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t sad4(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

template <int MODE>
__global__ void __launch_bounds__(256) k(uint32_t *out, int iters, uint32_t seed) {
    __shared__ uint32_t sm[2048];
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm[i] = i * 2654435761u + seed;
    __syncthreads();
    uint32_t a[8], acc[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = seed * (i + 1) + threadIdx.x; acc[i] = 0; }
    uint32_t b = seed ^ threadIdx.x;
    const int lane = threadIdx.x & 31;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 4; u++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (MODE == 0) acc[i] = sad4(a[i], b, acc[i]);                       // pure SAD
                if (MODE == 1) { acc[i] = sad4(__funnelshift_r(a[i], a[(i + 1) & 7], 8), b, acc[i]); }  // SHF + SAD
                if (MODE == 2) acc[i] = acc[i] + (a[i] ^ b);                          // LOP3 + IADD (alu baseline)
                if (MODE == 3) acc[i] = sad4(sm[(lane + i * 32 + u * 256 + it) & 2047], b, acc[i]);  // LDS + SAD
                if (MODE == 4) { uint32_t w = sm[(lane + i * 32 + u * 256 + it) & 2047];
                                 acc[i] = sad4(w, b, acc[i]); acc[(i + 1) & 7] = sad4(w, a[i], acc[(i + 1) & 7]);
                                 acc[(i + 2) & 7] = sad4(w, a[(i + 3) & 7], acc[(i + 2) & 7]);
                                 acc[(i + 3) & 7] = sad4(w, a[(i + 5) & 7], acc[(i + 3) & 7]); }  // 1 LDS : 4 SAD
                if (MODE == 5) acc[i] = __sad(a[i], b, acc[i]);                      // scalar SAD (VABSDIFF)
                if (MODE == 6) acc[i] = __dp4a(a[i], b, acc[i]);                     // IDP4A for comparison
            }
            b += acc[0] & 1;
        }
    }
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) r += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// MODE 7: the inner body of the search kernels -- per window row 5 LDS, 12 SHF, 32 SAD (two candidate rows share it);
// MODE 8: the same plus 12 integer ops per row (address / key arithmetic), the measured in-kernel mix.
template <int MODE>
__global__ void __launch_bounds__(128) kmix(uint32_t *out, int iters, uint32_t seed) {
    __shared__ uint32_t sm[4096];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = i * 2654435761u + seed;
    __syncthreads();
    uint32_t sreg[8][4], acc[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}}, junk = seed;
#pragma unroll
    for (int r = 0; r < 8; r++)
#pragma unroll
        for (int c = 0; c < 4; c++) sreg[r][c] = seed * (r * 4 + c + 1) + threadIdx.x;
    const int lane = threadIdx.x & 31;
    for (int it = 0; it < iters; it++) {
        const uint32_t *rowp = sm + ((lane + it * 37) & 1023);
#pragma unroll
        for (int rr = 0; rr < 9; rr++) {
            const uint32_t *rp = rowp + rr * 33;
            uint32_t        w[5];
#pragma unroll
            for (int c = 0; c < 5; c++) w[c] = rp[c];
#pragma unroll
            for (int c = 0; c < 4; c++) {
                const uint32_t s0 = w[c], s1 = __funnelshift_r(w[c], w[c + 1], 8), s2 = __funnelshift_r(w[c], w[c + 1], 16),
                               s3 = __funnelshift_r(w[c], w[c + 1], 24);
                if (rr < 8) {
                    const uint32_t sw = sreg[rr][c];
                    acc[0][0] = sad4(s0, sw, acc[0][0]); acc[0][1] = sad4(s1, sw, acc[0][1]);
                    acc[0][2] = sad4(s2, sw, acc[0][2]); acc[0][3] = sad4(s3, sw, acc[0][3]);
                }
                if (rr >= 1) {
                    const uint32_t sw = sreg[rr - 1][c];
                    acc[1][0] = sad4(s0, sw, acc[1][0]); acc[1][1] = sad4(s1, sw, acc[1][1]);
                    acc[1][2] = sad4(s2, sw, acc[1][2]); acc[1][3] = sad4(s3, sw, acc[1][3]);
                }
                if (MODE == 8) { junk = junk * 5u + w[c]; junk ^= junk >> 3; junk += s1 & 7u; }
            }
        }
    }
    uint32_t r = junk;
#pragma unroll
    for (int i = 0; i < 4; i++) r += acc[0][i] + acc[1][i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// fsr by k bytes on the FMA pipe: (lo >> 8k) | (hi << (32 - 8k)) = mulhi(lo, 2^(32-8k)) + hi * 2^(32-8k)
template <int K>
__device__ __forceinline__ uint32_t fsr_mad(uint32_t lo, uint32_t hi) {
    constexpr uint32_t C = 1u << (32 - 8 * K);
    return hi * C + __umulhi(lo, C);
}

// MODE 9: the search body with the alignment shifts on the FMA pipe (2 IMAD per shifted word) instead of SHF
// MODE 10..13: pipe probes (8 independent chains): 10 IMAD, 11 IMAD.HI, 12 SAD + IMAD 1:1, 13 SHF (non-hoistable)
template <int MODE>
__global__ void __launch_bounds__(128) kmix2(uint32_t *out, int iters, uint32_t seed) {
    __shared__ uint32_t sm[4096];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = i * 2654435761u + seed;
    __syncthreads();
    uint32_t sreg[8][4], acc[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
#pragma unroll
    for (int r = 0; r < 8; r++)
#pragma unroll
        for (int c = 0; c < 4; c++) sreg[r][c] = seed * (r * 4 + c + 1) + threadIdx.x;
    const int lane = threadIdx.x & 31;
    if (MODE == 9) {
        for (int it = 0; it < iters; it++) {
            const uint32_t *rowp = sm + ((lane + it * 37) & 1023);
#pragma unroll
            for (int rr = 0; rr < 9; rr++) {
                const uint32_t *rp = rowp + rr * 33;
                uint32_t        w[5];
#pragma unroll
                for (int c = 0; c < 5; c++) w[c] = rp[c];
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const uint32_t s0 = w[c], s1 = fsr_mad<1>(w[c], w[c + 1]), s2 = fsr_mad<2>(w[c], w[c + 1]), s3 = fsr_mad<3>(w[c], w[c + 1]);
                    if (rr < 8) {
                        const uint32_t sw = sreg[rr][c];
                        acc[0][0] = sad4(s0, sw, acc[0][0]); acc[0][1] = sad4(s1, sw, acc[0][1]);
                        acc[0][2] = sad4(s2, sw, acc[0][2]); acc[0][3] = sad4(s3, sw, acc[0][3]);
                    }
                    if (rr >= 1) {
                        const uint32_t sw = sreg[rr - 1][c];
                        acc[1][0] = sad4(s0, sw, acc[1][0]); acc[1][1] = sad4(s1, sw, acc[1][1]);
                        acc[1][2] = sad4(s2, sw, acc[1][2]); acc[1][3] = sad4(s3, sw, acc[1][3]);
                    }
                }
            }
        }
    } else {
        uint32_t a[8], b = seed ^ threadIdx.x;
#pragma unroll
        for (int i = 0; i < 8; i++) a[i] = seed * (i + 3) + threadIdx.x;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int u = 0; u < 32; u++) {
                const int i = u & 7;
                if (MODE == 10) a[i] = a[i] * 3u + b;
                if (MODE == 11) a[i] = __umulhi(a[i], 0x01000000u) + b;
                if (MODE == 12) { a[i] = a[i] * 3u + b; acc[i >> 2][i & 3] = sad4(a[(i + 4) & 7], b, acc[i >> 2][i & 3]); }
                if (MODE == 13) a[i] = __funnelshift_r(a[i], a[(i + 1) & 7], 8);
            }
        }
#pragma unroll
        for (int i = 0; i < 8; i++) acc[0][0] += a[i];
    }
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) r += acc[0][i] + acc[1][i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE>
void runmix2(const char *name, int ctas_per_sm, double ops_per_iter) {
    int dev = 0, sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    uint32_t *out;
    const int grid = sms * ctas_per_sm, iters = 4000;
    cudaMalloc(&out, (size_t)grid * 128 * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    kmix2<MODE><<<grid, 128>>>(out, 100, 1);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        kmix2<MODE><<<grid, 128>>>(out, iters, r + 2);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double ops = (double)grid * 128 * iters * ops_per_iter;
    printf("{\"bench\":\"%s\",\"warps_per_sm\":%d,\"ms\":%.3f,\"Gops_per_s\":%.1f,\"warp_ops_per_clk_per_smsp_at_1965MHz\":%.3f}\n", name,
           ctas_per_sm * 4, best, ops / best / 1e6, ops / best / 1e3 / 32 / (sms * 4) / 1965e3);
    cudaFree(out);
}

template <int MODE>
void runmix(const char *name, int ctas_per_sm) {
    int dev = 0, sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    uint32_t *out;
    const int grid = sms * ctas_per_sm, iters = 4000;
    cudaMalloc(&out, (size_t)grid * 128 * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    kmix<MODE><<<grid, 128>>>(out, 100, 1);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        kmix<MODE><<<grid, 128>>>(out, iters, r + 2);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double sads = (double)grid * 128 * iters * 256.0;
    printf("{\"bench\":\"%s\",\"warps_per_sm\":%d,\"ms\":%.3f,\"Gsad4_per_s\":%.1f,\"Tpx_per_s\":%.2f,\"per_sm_per_clk_at_1965MHz\":%.2f}\n", name,
           ctas_per_sm * 4, best, sads / best / 1e6, sads * 4 / best / 1e9, sads / best / 1e3 / sms / 1965e3);
    cudaFree(out);
}

template <int MODE>
void run(const char *name, int ops_per_inner, int ctas_per_sm) {
    int dev = 0, sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    uint32_t *out;
    const int grid = sms * ctas_per_sm, iters = 20000;
    cudaMalloc(&out, (size_t)grid * 256 * 4);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<grid, 256>>>(out, 100, 1);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        k<MODE><<<grid, 256>>>(out, iters, r + 2);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double sads = (double)grid * 256 * iters * 32.0 * ops_per_inner;
    printf("{\"bench\":\"%s\",\"ctas_per_sm\":%d,\"ms\":%.3f,\"Gsad4_per_s\":%.1f,\"Tpx_per_s\":%.2f,\"per_sm_per_clk_at_1965MHz\":%.2f}\n", name,
           ctas_per_sm, best, sads / best / 1e6, sads * 4 / best / 1e9, sads / best / 1e3 / sms / 1965e3);
    cudaFree(out);
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    printf("{\"gpu\":\"%s\",\"sms\":%d,\"clock_khz\":%d}\n", p.name, p.multiProcessorCount, p.clockRate);
    for (int c = 1; c <= 4; c *= 2) {
        run<0>("vabsdiff4.acc", 1, c);
        run<1>("shf+vabsdiff4.acc", 1, c);
        run<2>("lop3+iadd", 1, c);
        run<3>("lds+vabsdiff4.acc", 1, c);
        run<4>("lds+4xvabsdiff4.acc", 4, c);
        run<5>("sad.u32", 1, c);
        run<6>("dp4a", 1, c);
    }
    for (int c = 4; c <= 8; c += 4) {
        runmix2<9>("search-body, shifts as IMAD pairs (SAD count)", c, 256.0);
        runmix2<10>("imad", c, 32.0);
        runmix2<11>("imad.hi", c, 32.0);
        runmix2<12>("imad+sad (pairs)", c, 32.0);
        runmix2<13>("shf", c, 32.0);
    }
    for (int c = 2; c <= 12; c += 2) {
        runmix<7>("search-body 5LDS+12SHF+32SAD", c);
        runmix<8>("search-body + 12 int ops/row", c);
    }
    return 0;
}
