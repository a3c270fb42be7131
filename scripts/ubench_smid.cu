// Scratch: is the block -> SM placement of a one-wave grid (292 blocks x 512 threads, 103 KB of shared memory, 2 per SM)
// the same from launch to launch?   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ubench_smid ubench_smid.cu
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(512, 2) k(unsigned* smid_out, long long* t_out, int spin) {
    extern __shared__ unsigned char smem[];
    unsigned smid; asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    if (threadIdx.x == 0) { smid_out[blockIdx.x] = smid; t_out[blockIdx.x] = clock64(); }
    // unequal work so that blocks finish at different times (as the real kernel's do)
    float a = threadIdx.x;
    for (int i = 0; i < spin * (1 + (int)(blockIdx.x % 7)); ++i) a = a * 1.0001f + 0.5f;
    if (a == 12345.f) smem[0] = 1;
}
int main() {
    const int nb = 292, nl = 200;
    unsigned* d; long long* t; cudaMalloc(&d, nb * 4); cudaMalloc(&t, nb * 8);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 105000);
    std::vector<unsigned> first(nb), cur(nb);
    int changed_launches = 0, max_changed = 0, pair_changed_launches = 0;
    for (int l = 0; l < nl; ++l) {
        k<<<nb, 512, 105000>>>(d, t, 2000);
        cudaMemcpy(cur.data(), d, nb * 4, cudaMemcpyDeviceToHost);
        if (l == 0) { first = cur; continue; }
        int ch = 0, pch = 0;
        for (int b = 0; b < nb; ++b) ch += cur[b] != first[b];
        // pairing: which other block shares my SM, and am I the lower blockIdx of the pair?
        for (int b = 0; b < nb; ++b) {
            int mate0 = -1, mate1 = -1;
            for (int c = 0; c < nb; ++c) { if (c != b && first[c] == first[b]) mate0 = c; if (c != b && cur[c] == cur[b]) mate1 = c; }
            const bool low0 = mate0 < 0 || b < mate0, low1 = mate1 < 0 || b < mate1;
            pch += low0 != low1;
        }
        changed_launches += ch > 0; pair_changed_launches += pch > 0; if (ch > max_changed) max_changed = ch;
    }
    printf("%d launches of %d blocks: placement differs from the first launch in %d launches (at most %d blocks moved); "
           "the 'lower blockIdx of its SM' role of some block differs in %d launches\n", nl, nb, changed_launches, max_changed, pair_changed_launches);
    printf("first launch, smid of blocks 0..15:"); for (int b = 0; b < 16; ++b) printf(" %u", first[b]); printf("\n");
    printf("first launch, smid of blocks 146..161:"); for (int b = 146; b < 162; ++b) printf(" %u", first[b]); printf("\n");
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
