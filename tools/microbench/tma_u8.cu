// tma_u8.cu -- stand-alone check of the TMA pattern st_sad16_kernel uses: a 2-D tiled u8 tensor map over a padded
// plane, boxes of 8 rows x {48, 80, 112, 144} bytes fetched at arbitrary byte coordinates by several lanes, one
// mbarrier.  nvcc -gencode arch=compute_100a,code=sm_100a -o tma_u8 tma_u8.cu && ./tma_u8
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

__device__ __forceinline__ void mbar_init(uint32_t mbar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\nfence.mbarrier_init.release.cluster;" ::"r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t mbar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t mbar, uint32_t parity) {
    uint32_t done;
    do {
        asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                     : "=r"(done) : "r"(mbar), "r"(parity) : "memory");
    } while (!done);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void *tmap, int x, int y, uint32_t mbar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(tmap), "r"(x), "r"(y), "r"(mbar) : "memory");
}

__global__ void k(const void *tmaps, int cls, int box_bytes, int x, int y, int nbox, uint8_t *out, int mode) {
    extern __shared__ __align__(128) uint32_t win[];
    __shared__ __align__(8) unsigned long long mb;
    const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(&mb), sbase = (uint32_t)__cvta_generic_to_shared(win);
    if (threadIdx.x == 0) { mbar_init(mbar, 1); mbar_expect_tx(mbar, (uint32_t)(nbox * 8 * box_bytes)); }
    __syncthreads();
    const uint8_t *tm = static_cast<const uint8_t *>(tmaps) + cls * 128;
    if (threadIdx.x == 0 && (mode & 2)) printf("smem base & 1023 = %u, mbar %u\n", sbase & 1023u, mbar);
    if (mode & 1) {
        if (threadIdx.x == 0)
            for (int b = 0; b < nbox; b++) tma_load_2d(sbase + (uint32_t)(b * 8 * box_bytes), tm, x, y + 8 * b, mbar);
    } else if (threadIdx.x < 32)
        for (int b = threadIdx.x; b < nbox; b += 32) tma_load_2d(sbase + (uint32_t)(b * 8 * box_bytes), tm, x, y + 8 * b, mbar);
    mbar_wait(mbar, 0);
    for (int i = threadIdx.x; i < nbox * 8 * box_bytes; i += blockDim.x) out[i] = reinterpret_cast<const uint8_t *>(win)[i];
}

int main(int argc, char **argv) {
    const int a_cls = argc > 1 ? atoi(argv[1]) : -1, a_x = argc > 2 ? atoi(argv[2]) : 0, a_y = argc > 3 ? atoi(argv[3]) : 0, a_nbox = argc > 4 ? atoi(argv[4]) : 1, a_mode = argc > 5 ? atoi(argv[5]) : 0;
    const int stride = 1024, rows = 604, boxes[4] = {48, 80, 112, 144};
    std::vector<uint8_t> h((size_t)stride * rows);
    for (size_t i = 0; i < h.size(); i++) h[i] = (uint8_t)((i * 2654435761u) >> 13);
    uint8_t *d, *d_out, *d_maps;
    cudaMalloc(&d, h.size() + 1024); cudaMalloc(&d_out, 64 * 1024); cudaMalloc(&d_maps, 512);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn) { printf("no entry point\n"); return 1; }
    typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                 const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    alignas(64) CUtensorMap maps[4];
    for (int c = 0; c < 4; c++) {
        const cuuint64_t gdim[2] = {(cuuint64_t)stride, (cuuint64_t)rows}, gstr[1] = {(cuuint64_t)stride};
        const cuuint32_t box[2] = {(cuuint32_t)boxes[c], 8}, estr[2] = {1, 1};
        CUresult r = ((EncodeFn)fn)(&maps[c], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        printf("encode class %d: %d\n", c, (int)r);
    }
    cudaMemcpy(d_maps, maps, sizeof(maps), cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    int bad = 0;
    for (int c = 0; c < 4; c++)
        for (int t = 0; t < 6; t++) {
            int x = (t * 37 + 5) % 700, y = (t * 53 + 3) % 560, nbox = 1 + (t * 7) % 40, bb = boxes[c];
            if (a_cls >= 0) { if (c != a_cls || t) continue; x = a_x; y = a_y; nbox = a_nbox; }
            if (nbox * 8 * bb > 60 * 1024) continue;
            k<<<1, 128, 64 * 1024>>>(d_maps, c, bb, x, y, nbox, d_out, a_mode);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("class %d x %d y %d nbox %d: %s\n", c, x, y, nbox, cudaGetErrorString(e)); return 2; }
            std::vector<uint8_t> o((size_t)nbox * 8 * bb);
            cudaMemcpy(o.data(), d_out, o.size(), cudaMemcpyDeviceToHost);
            int mism = 0;
            for (int r = 0; r < nbox * 8; r++)
                for (int cc = 0; cc < bb; cc++) {
                    const int gy = y + r, gx = x + cc;
                    const uint8_t want = (gy < rows && gx < stride) ? h[(size_t)gy * stride + gx] : 0;
                    if (o[(size_t)r * bb + cc] != want) mism++;
                }
            printf("class %d x %d y %d nbox %d: %d mismatches\n", c, x, y, nbox, mism);
            bad += mism != 0;
        }
    printf(bad ? "FAILED\n" : "OK\n");
    return bad ? 3 : 0;
}
