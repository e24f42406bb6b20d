// Micro-benchmark (not part of the product): what does HBM sustain for a pure output stream of the
// size this path writes (126 MB per launch), through (a) 128-bit st.global from registers and
// (b) cp.async.bulk shared->global in 6 KB / 12 KB pieces, as pva_tiled_kernel issues them?
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o scripts/write_bw scripts/write_bw.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include <stdint.h>

__global__ void k_stg(uint4* out, size_t n16) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    uint4 v = make_uint4(threadIdx.x, blockIdx.x, 3, 4);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) out[i] = v;
}
// contiguous per-CTA ranges, like the tiled kernel's tile ranges
__global__ void k_stg_blocked(uint4* out, size_t n16) {
    const size_t per = (n16 + gridDim.x - 1) / gridDim.x;
    const size_t b = per * blockIdx.x, e = min(b + per, n16);
    uint4 v = make_uint4(threadIdx.x, blockIdx.x, 3, 4);
    for (size_t i = b + threadIdx.x; i < e; i += blockDim.x) out[i] = v;
}
__global__ void k_bulk(char* out, size_t bytes, int piece) {
    extern __shared__ __align__(128) char smem[];
    for (int i = threadIdx.x; i < piece / 4; i += blockDim.x) ((int*)smem)[i] = i;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const size_t n_pieces = bytes / piece;
    const size_t per = (n_pieces + gridDim.x - 1) / gridDim.x;
    const size_t b = per * blockIdx.x, e = min(b + per, n_pieces);
    if (threadIdx.x < 8) {  // 8 issuing threads per CTA, like 8 strips
        const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
        for (size_t i = b + threadIdx.x; i < e; i += 8) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + i * piece), "r"(s), "r"(piece) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}
int main() {
    const size_t bytes = 125829120;  // fast + slow outputs of one headline launch
    const int nbuf = 6, reps = 300;
    char* buf[nbuf];
    for (auto& b : buf) cudaMalloc(&b, bytes);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0), cudaEventCreate(&e1);
    auto run = [&](const char* name, auto launch) {
        for (int i = 0; i < 20; ++i) launch(buf[i % nbuf]);
        cudaDeviceSynchronize();
        cudaEventRecord(e0);
        for (int i = 0; i < reps; ++i) launch(buf[i % nbuf]);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        printf("%-42s %7.1f us  %6.2f TB/s  (%s)\n", name, ms / reps * 1e3, bytes / (ms / reps * 1e-3) / 1e12, cudaGetErrorString(cudaGetLastError()));
    };
    run("st.global.v4 grid-stride, 148x8 CTAs", [&](char* b) { k_stg<<<148 * 8, 256>>>((uint4*)b, bytes / 16); });
    run("st.global.v4 grid-stride, 148x32 CTAs", [&](char* b) { k_stg<<<148 * 32, 256>>>((uint4*)b, bytes / 16); });
    run("st.global.v4 blocked ranges, 148 CTAs x 512", [&](char* b) { k_stg_blocked<<<148, 512>>>((uint4*)b, bytes / 16); });
    run("st.global.v4 blocked ranges, 148x4 CTAs", [&](char* b) { k_stg_blocked<<<148 * 4, 512>>>((uint4*)b, bytes / 16); });
    for (int piece : {6144, 12288, 49152}) {
        char name[64];
        snprintf(name, sizeof name, "cp.async.bulk %d B pieces, 148 CTAs", piece);
        cudaFuncSetAttribute(k_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, piece);
        run(name, [&](char* b) { k_bulk<<<148, 256, piece>>>(b, bytes, piece); });
    }
    return 0;
}
