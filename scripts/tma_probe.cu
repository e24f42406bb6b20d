// tma_probe.cu -- does a 4-D tensor copy work with (a) the map at the start of the kernel parameter
// block, (b) the map 16 KB into it, (c) the map in global memory?  Same tensor shape as the
// column-cropped loads of pva_tiled.cu (32-bit words, rows of one residue class mod 4 of a 340-byte
// pitch plane).  One variant per process (a fault poisons the context):
//   nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a scripts/tma_probe.cu -o scripts/tma_probe
//   scripts/tma_probe 0|1|2
#include <cuda.h>
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

constexpr int W = 340, H = 256, T = 4, C = 3;
static int BW = 272, NG = 9, NC = 3, KRES = 1, DT = 2, PROMO = 0, PITCH4 = 4 * W;  // box: BW bytes x NG row groups x 1 frame x NC planes

template <int PAD>
struct Args {
    char pad[PAD > 0 ? PAD : 64];
    alignas(64) CUtensorMap map;
    const CUtensorMap* gmap;
    unsigned char* out;
    int x, g, t;
    uint32_t bytes;
};

template <int PAD, bool GLOBAL>
__global__ void probe(const __grid_constant__ Args<PAD> a) {
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) unsigned long long bar;
    const uint32_t b = (uint32_t)__cvta_generic_to_shared(&bar);
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(smem);
    const uint32_t bytes = a.bytes;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(b) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
        const CUtensorMap* m = GLOBAL ? a.gmap : &a.map;
        asm volatile(
            "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];" ::"r"(dst),
            "l"(reinterpret_cast<uint64_t>(m)), "r"(a.x), "r"(a.g), "r"(a.t), "r"(0), "r"(b)
            : "memory");
    }
    uint32_t ok = 0;
    while (!ok) {
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(b), "r"(0)
            : "memory");
    }
    for (uint32_t i = threadIdx.x; i < bytes; i += blockDim.x) a.out[i] = smem[i];
}

#define CK(x)                                                                      \
    do {                                                                           \
        cudaError_t e = (x);                                                       \
        if (e != cudaSuccess) {                                                    \
            printf("FAIL %s: %s\n", #x, cudaGetErrorString(e));                    \
            return 1;                                                              \
        }                                                                          \
    } while (0)

template <int PAD, bool GLOBAL>
int run(const char* name) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q));
    EncodeFn enc = (EncodeFn)sym;
    const size_t n = (size_t)C * T * H * W;
    unsigned char* h = (unsigned char*)malloc(n);
    for (size_t i = 0; i < n; ++i) h[i] = (unsigned char)((i * 2654435761u) >> 13);
    unsigned char *d, *out;
    CK(cudaMalloc(&d, n + 64));
    CK(cudaMemcpy(d, h, n, cudaMemcpyHostToDevice));
    const int bytes = NC * NG * BW;
    CK(cudaMalloc(&out, bytes));
    const int k = KRES, ca = 42, ra = 37, t = 2;  // residue class 1, window from column 42, rows from 37
    const uintptr_t row_k = (uintptr_t)d + (size_t)k * W;
    const uintptr_t delta = row_k & 15;
    static Args<PAD> a;
    memset(&a, 0, sizeof(a));
    const int es_ = DT == 2 ? 4 : 1;
    const cuuint64_t dims[4] = {(delta + W) / es_, (cuuint64_t)((H - k + 3) / 4), T, C};
    const cuuint64_t strides[3] = {(cuuint64_t)PITCH4, (cuuint64_t)H * W, (cuuint64_t)T * H * W};
    const cuuint32_t box[4] = {(cuuint32_t)(BW / es_), (cuuint32_t)NG, 1, (cuuint32_t)NC};
    const cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = enc(&a.map, DT == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT32 : CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, (void*)(row_k - delta), dims, strides, box, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)PROMO,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("%s: encode rc=%d delta=%d sizeof(Args)=%zu map offset=%zu\n", name, (int)r, (int)delta, sizeof(a),
           (size_t)((char*)&a.map - (char*)&a));
    if (r != CUDA_SUCCESS) return 1;
    printf("query result %d; map words:", (int)q);
    for (int i = 0; i < 16; ++i) printf(" %016llx", (unsigned long long)((const uint64_t*)&a.map)[i]);
    printf("\n");
    if (getenv("PROBE_DLSYM")) {
        void* lib = dlopen("libcuda.so.1", RTLD_NOW);
        EncodeFn e2 = lib ? (EncodeFn)dlsym(lib, "cuTensorMapEncodeTiled") : nullptr;
        printf("dlsym encoder %p vs entry point %p\n", (void*)e2, (void*)enc);
        if (e2) {
            CUtensorMap m2;
            memset(&m2, 0, sizeof(m2));
            CUresult r2 = e2(&m2, DT == 2 ? CU_TENSOR_MAP_DATA_TYPE_UINT32 : CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, (void*)(row_k - delta), dims, strides, box, es,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)PROMO,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            printf("dlsym encode rc=%d same=%d\n", (int)r2, memcmp(&m2, &a.map, sizeof(m2)) == 0);
            a.map = m2;
        }
    }
    CUtensorMap* gm;
    CK(cudaMalloc(&gm, sizeof(CUtensorMap)));
    CK(cudaMemcpy(gm, &a.map, sizeof(CUtensorMap), cudaMemcpyHostToDevice));
    a.gmap = gm;
    a.out = out;
    a.x = DT == 2 ? (int)((ca + delta) >> 2) : (int)(((ca + delta) & ~3));
    a.bytes = bytes;
    a.g = ra >> 2;
    a.t = t;
    CK(cudaFuncSetAttribute(probe<PAD, GLOBAL>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes + 128));
    probe<PAD, GLOBAL><<<1, 128, bytes + 128>>>(a);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    unsigned char* got = (unsigned char*)malloc(bytes);
    CK(cudaMemcpy(got, out, bytes, cudaMemcpyDeviceToHost));
    int bad = 0;
    for (int c = 0; c < NC; ++c)
        for (int g = 0; g < NG; ++g)
            for (int xb = 0; xb < BW; ++xb) {
                const int r_ = 4 * ((ra >> 2) + g) + k, col = (ca & ~3) + xb;
                const unsigned char want = (col < W && r_ < H) ? h[(((size_t)c * T + t) * H + r_) * W + col] : 0;
                if (got[(c * NG + g) * BW + xb] != want) ++bad;
            }
    printf("%s: %s (%d mismatching bytes of %d)\n", name, bad ? "WRONG DATA" : "OK", bad, bytes);
    return bad != 0;
}

int main(int argc, char** argv) {
    const int v = argc > 1 ? atoi(argv[1]) : 0;
    // optional: box bytes, row groups, planes, residue, dtype (0 u8 / 2 u32), L2 promotion
    if (argc > 2) BW = atoi(argv[2]);
    if (argc > 3) NG = atoi(argv[3]);
    if (argc > 4) NC = atoi(argv[4]);
    if (argc > 5) KRES = atoi(argv[5]);
    if (argc > 6) DT = atoi(argv[6]);
    if (argc > 7) PROMO = atoi(argv[7]);
    printf("box %d B x %d groups x %d planes, residue %d, dtype %d, promo %d\n", BW, NG, NC, KRES, DT, PROMO);
    if (v == 0) return run<0, false>("map at the start of the parameter block");
    if (v == 1) return run<16384, false>("map 16 KB into the parameter block");
    if (v == 3) return run<3072, false>("map 3 KB into the parameter block");
    if (v == 4) return run<8192, false>("map 8 KB into the parameter block");
    return run<0, true>("map in global memory");
}
