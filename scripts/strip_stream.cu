// Scratch microbenchmark: HBM throughput of a 4-read / 3-write streaming kernel when every block walks down a COLUMN
// STRIP of a 4096 x 4096 fp32 image, as a function of the strip width (bytes of one row chunk).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o strip_stream strip_stream.cu && ./strip_stream
#include <cstdio>
#include <cuda_runtime.h>
template <int CHUNK_F4>   // float4 per row chunk: 32 -> 512 B, 64 -> 1 KB, 128 -> 2 KB
__global__ void __launch_bounds__(256, 2) k(const float4* a, const float4* b, const float4* c, const float4* d,
                                            float4* o0, float4* o1, float4* o2, int W4, int H, int nstrips, int seg_rows) {
    const int strip = blockIdx.x % nstrips, seg = blockIdx.x / nstrips;
    constexpr int ROWS = 256 / CHUNK_F4;          // rows covered by the block per iteration
    const int col = strip * CHUNK_F4 + threadIdx.x % CHUNK_F4, r0 = seg * seg_rows + threadIdx.x / CHUNK_F4;
    const int r1 = min(H, (seg + 1) * seg_rows);
    constexpr int U = 4;
    for (int r = r0; r < r1; r += ROWS * U) {
        float4 va[U], vb[U], vc[U], vd[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int rr = r + u * ROWS;
            if (rr < r1) {
                const size_t i = (size_t)rr * W4 + col;
                va[u] = a[i]; vb[u] = b[i]; vc[u] = c[i]; vd[u] = d[i];
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int rr = r + u * ROWS;
            if (rr < r1) {
                const size_t i = (size_t)rr * W4 + col;
                float4 s = make_float4(va[u].x + vb[u].x, va[u].y + vb[u].y, va[u].z + vb[u].z, va[u].w + vb[u].w);
                float4 t = make_float4(vc[u].x * vd[u].x, vc[u].y * vd[u].y, vc[u].z * vd[u].z, vc[u].w * vd[u].w);
                o0[i] = s; o1[i] = t; o2[i] = make_float4(s.x + t.x, s.y + t.y, s.z + t.z, s.w + t.w);
            }
        }
    }
}
template <int CHUNK_F4> void run(const char* name, float4** p, int W, int H, int nblocks_target, bool aligned) {
    const int W4 = W / 4, nstrips = W4 / CHUNK_F4;
    int nseg = nblocks_target / nstrips;
    int seg_rows = (H + nseg - 1) / nseg;
    if (!aligned) seg_rows += 7;                  // segments of different strips stay aligned anyway here
    nseg = (H + seg_rows - 1) / seg_rows;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int w = 0; w < 3; ++w) k<CHUNK_F4><<<nstrips * nseg, 256>>>(p[0], p[1], p[2], p[3], p[4], p[5], p[6], W4, H, nstrips, seg_rows);
    cudaEventRecord(e0);
    const int n = 50;
    for (int w = 0; w < n; ++w) k<CHUNK_F4><<<nstrips * nseg, 256>>>(p[0], p[1], p[2], p[3], p[4], p[5], p[6], W4, H, nstrips, seg_rows);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double bytes = 7.0 * W * H * 4;
    printf("%-28s blocks %4d seg_rows %4d: %.1f us  %.0f GB/s\n", name, nstrips * nseg, seg_rows, 1e3 * ms / n, bytes / (ms / n * 1e-3) / 1e9);
}
int main() {
    const int W = 4096, H = 4096;
    float4* p[7];
    for (int i = 0; i < 7; ++i) { cudaMalloc(&p[i], (size_t)W * H * 4); cudaMemset(p[i], 0, (size_t)W * H * 4); }
    for (int blocks : {296, 592, 1184, 4736}) {
        printf("-- target blocks %d\n", blocks);
        run<32>("strip 512 B", p, W, H, blocks, true);
        run<64>("strip 1 KB", p, W, H, blocks, true);
        run<128>("strip 2 KB", p, W, H, blocks, true);
        run<256>("strip 4 KB", p, W, H, blocks, true);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
