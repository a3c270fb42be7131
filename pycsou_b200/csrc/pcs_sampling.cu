// libpcsb200: gather / scatter kernels behind Masking, DownSampling and SubSampling (sm_100a).
//
// Replaces numpy boolean fancy indexing in pycsou/linop/sampling.py:190-196 (Masking / DownSampling) and
// pylops.Restriction (pycsou/linop/sampling.py:121) -- SURVEY.md rows a15-a17.  Masks are turned into a sorted
// int64 index list once at construction; the kernels move opaque 4/8/16-byte elements, so they serve
// float32, float64/complex64 and complex128 alike and are bit-exact by construction.
//
// (outer, n, inner) view for Restriction along an axis: out[o, j, k] = x[o, idx[j], k].
#include "pcs_common.cuh"
#include <mutex>

using namespace pcs;

constexpr int SG_BLOCK = 256;

template <typename E>
__global__ void __launch_bounds__(SG_BLOCK) take_kernel(E* __restrict__ out, const E* __restrict__ x,
                                                        const int64_t* __restrict__ idx, int64_t outer, int64_t n,
                                                        int64_t m, int64_t inner) {
    const int64_t total = outer * m * inner, stride = (int64_t)gridDim.x * SG_BLOCK;
    for (int64_t f = (int64_t)blockIdx.x * SG_BLOCK + threadIdx.x; f < total; f += stride) {
        const int64_t k = f % inner, j = (f / inner) % m, o = f / (inner * m);
        out[f] = x[(o * n + idx[j]) * inner + k];
    }
}

// out must be zero-filled by the caller (the C wrapper does it); duplicates: last writer wins is NOT guaranteed on a
// GPU, so the host wrapper rejects duplicate indices for the adjoint (pylops assignment semantics need an order).
template <typename E>
__global__ void __launch_bounds__(SG_BLOCK) put_kernel(E* __restrict__ out, const E* __restrict__ y,
                                                       const int64_t* __restrict__ idx, int64_t outer, int64_t n,
                                                       int64_t m, int64_t inner) {
    const int64_t total = outer * m * inner, stride = (int64_t)gridDim.x * SG_BLOCK;
    for (int64_t f = (int64_t)blockIdx.x * SG_BLOCK + threadIdx.x; f < total; f += stride) {
        const int64_t k = f % inner, j = (f / inner) % m, o = f / (inner * m);
        out[(o * n + idx[j]) * inner + k] = y[f];
    }
}

template <typename E>
static int take_launch(void* out, const void* x, const int64_t* idx, int64_t outer, int64_t n, int64_t m,
                       int64_t inner, cudaStream_t s) {
    const int64_t total = outer * m * inner;
    if (total == 0) return PCS_OK;
    take_kernel<E><<<grid_for(total, SG_BLOCK, 8), SG_BLOCK, 0, s>>>((E*)out, (const E*)x, idx, outer, n, m, inner);
    return check_launch("pcs_take");
}

template <typename E>
static int put_launch(void* out, const void* y, const int64_t* idx, int64_t outer, int64_t n, int64_t m, int64_t inner,
                      cudaStream_t s) {
    cudaError_t e = cudaMemsetAsync(out, 0, (size_t)(outer * n * inner) * sizeof(E), s);
    if (e != cudaSuccess) {
        set_error("pcs_put: memset failed: %s", cudaGetErrorString(e));
        return PCS_ERR_CUDA;
    }
    const int64_t total = outer * m * inner;
    if (total == 0) return PCS_OK;
    put_kernel<E><<<grid_for(total, SG_BLOCK, 8), SG_BLOCK, 0, s>>>((E*)out, (const E*)y, idx, outer, n, m, inner);
    return check_launch("pcs_put");
}

struct alignas(16) B16 { uint64_t a, b; };

// out[(o, j, k)] = x[(o, idx[j], k)], element size in bytes in {4, 8, 16}
extern "C" int pcs_take(void* out, const void* x, const void* idx_i64, int64_t outer, int64_t n, int64_t m,
                        int64_t inner, int elem_bytes, void* stream) {
    PCS_REQUIRE(outer >= 0 && n >= 0 && m >= 0 && inner >= 0, "pcs_take: negative dimension");
    PCS_REQUIRE(outer * m * inner == 0 || (out && x && idx_i64), "pcs_take: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t* idx = (const int64_t*)idx_i64;
    switch (elem_bytes) {
        case 4: return take_launch<uint32_t>(out, x, idx, outer, n, m, inner, s);
        case 8: return take_launch<uint64_t>(out, x, idx, outer, n, m, inner, s);
        case 16: return take_launch<B16>(out, x, idx, outer, n, m, inner, s);
    }
    set_error("pcs_take: element size must be 4, 8 or 16 bytes (got %d)", elem_bytes);
    return PCS_ERR_ARG;
}

// out = zeros(outer, n, inner); out[(o, idx[j], k)] = y[(o, j, k)]  (indices must be unique)
extern "C" int pcs_put(void* out, const void* y, const void* idx_i64, int64_t outer, int64_t n, int64_t m,
                       int64_t inner, int elem_bytes, void* stream) {
    PCS_REQUIRE(outer >= 0 && n >= 0 && m >= 0 && inner >= 0, "pcs_put: negative dimension");
    PCS_REQUIRE(outer * n * inner == 0 || out, "pcs_put: null output");
    PCS_REQUIRE(outer * m * inner == 0 || (y && idx_i64), "pcs_put: null pointer");
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t* idx = (const int64_t*)idx_i64;
    if (outer * n * inner == 0) return PCS_OK;
    switch (elem_bytes) {
        case 4: return put_launch<uint32_t>(out, y, idx, outer, n, m, inner, s);
        case 8: return put_launch<uint64_t>(out, y, idx, outer, n, m, inner, s);
        case 16: return put_launch<B16>(out, y, idx, outer, n, m, inner, s);
    }
    set_error("pcs_put: element size must be 4, 8 or 16 bytes (got %d)", elem_bytes);
    return PCS_ERR_ARG;
}

// Boolean mask -> number of set entries and their sorted positions.  Single-block-per-chunk two-pass compaction:
// pass 1 counts per chunk, host-free exclusive scan by one block, pass 2 writes indices.  Construction-time only.
constexpr int MK_CHUNK = 4096;

__global__ void mask_count_kernel(const uint8_t* __restrict__ mask, int64_t n, int64_t* __restrict__ counts) {
    __shared__ int sm[SG_BLOCK / 32];
    const int64_t base = (int64_t)blockIdx.x * MK_CHUNK;
    int c = 0;
    for (int e = threadIdx.x; e < MK_CHUNK; e += SG_BLOCK) {
        const int64_t i = base + e;
        if (i < n && mask[i]) ++c;
    }
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int w = 0; w < SG_BLOCK / 32; ++w) t += sm[w];
        counts[blockIdx.x] = t;
    }
}

__global__ void mask_scan_kernel(int64_t* counts, int64_t nchunks, int64_t* total) {
    // one thread: nchunks = n / 4096 is small (<= 262144 for a 2^30 mask); construction-time only
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        int64_t run = 0;
        for (int64_t i = 0; i < nchunks; ++i) {
            int64_t c = counts[i];
            counts[i] = run;
            run += c;
        }
        *total = run;
    }
}

__global__ void mask_write_kernel(const uint8_t* __restrict__ mask, int64_t n, const int64_t* __restrict__ offsets,
                                  int64_t* __restrict__ idx) {
    // one warp per chunk keeps the output sorted: ballot + popc prefix
    const int lane = threadIdx.x & 31;
    const int64_t chunk = (int64_t)blockIdx.x * (SG_BLOCK / 32) + (threadIdx.x >> 5);
    const int64_t base = chunk * MK_CHUNK;
    if (base >= n) return;
    int64_t pos = offsets[chunk];
    for (int e = 0; e < MK_CHUNK; e += 32) {
        const int64_t i = base + e + lane;
        const bool set = (i < n) && mask[i];
        const unsigned b = __ballot_sync(0xffffffffu, set);
        if (set) idx[pos + __popc(b & ((1u << lane) - 1u))] = i;
        pos += __popc(b);
    }
}

extern "C" int64_t pcs_mask_workspace_bytes(int64_t n) { return ((n + MK_CHUNK - 1) / MK_CHUNK + 2) * 8; }

// Phase 1: count.  total_out (device int64) receives popcount(mask); workspace keeps the chunk offsets for phase 2.
extern "C" int pcs_mask_count(const void* mask_u8, int64_t n, void* workspace, void* total_out_i64, void* stream) {
    PCS_REQUIRE(n >= 0 && workspace && total_out_i64 && (n == 0 || mask_u8), "pcs_mask_count: bad arguments");
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t nchunks = (n + MK_CHUNK - 1) / MK_CHUNK;
    if (nchunks > 0) mask_count_kernel<<<(unsigned)nchunks, SG_BLOCK, 0, s>>>((const uint8_t*)mask_u8, n, (int64_t*)workspace);
    mask_scan_kernel<<<1, 32, 0, s>>>((int64_t*)workspace, nchunks, (int64_t*)total_out_i64);
    return check_launch("pcs_mask_count");
}

// Phase 2: write the sorted positions (idx_out must hold popcount(mask) int64).
extern "C" int pcs_mask_indices(const void* mask_u8, int64_t n, const void* workspace, void* idx_out_i64, void* stream) {
    PCS_REQUIRE(n >= 0 && workspace && (n == 0 || (mask_u8 && idx_out_i64)), "pcs_mask_indices: bad arguments");
    const int64_t nchunks = (n + MK_CHUNK - 1) / MK_CHUNK;
    if (nchunks == 0) return PCS_OK;
    const int warps = SG_BLOCK / 32;
    mask_write_kernel<<<(unsigned)((nchunks + warps - 1) / warps), SG_BLOCK, 0, (cudaStream_t)stream>>>(
        (const uint8_t*)mask_u8, n, (const int64_t*)workspace, (int64_t*)idx_out_i64);
    return check_launch("pcs_mask_indices");
}

// ---- allocation-free variant ------------------------------------------------------------------------------------------
// pcs_mask_to_indices_static: mask -> count + sorted positions in ONE call that allocates nothing (no cudaMalloc / cudaFree,
// no allocating library routine): for callers under CUDA graph capture, tight serving loops and harnesses that forbid
// allocation.  The per-chunk counts / offsets that pcs_mask_count keeps in a caller-provided workspace live in a
// file-scope device array instead (one copy per GPU that loads the library, held until exit), which bounds the mask size:
// n <= PCS_MASK_STATIC_MAX_N = MK_STATIC_MAX_CHUNKS * MK_CHUNK = 64 Mi entries (an 8192^2 mask); 128 KiB of scratch.
// No counterpart in the reference.
constexpr int64_t MK_STATIC_MAX_CHUNKS = 16384;
static __device__ __align__(256) unsigned char g_mask_static_scratch[(MK_STATIC_MAX_CHUNKS + 2) * 8];

// as mask_write_kernel, but never writes beyond `capacity` entries (the caller sizes idx_out without a host read-back)
__global__ void mask_write_capped_kernel(const uint8_t* __restrict__ mask, int64_t n, const int64_t* __restrict__ offsets,
                                         int64_t* __restrict__ idx, int64_t capacity) {
    const int lane = threadIdx.x & 31;
    const int64_t chunk = (int64_t)blockIdx.x * (SG_BLOCK / 32) + (threadIdx.x >> 5);
    const int64_t base = chunk * MK_CHUNK;
    if (base >= n) return;
    int64_t pos = offsets[chunk];
    for (int e = 0; e < MK_CHUNK; e += 32) {
        const int64_t i = base + e + lane;
        const bool set = (i < n) && mask[i];
        const unsigned b = __ballot_sync(0xffffffffu, set);
        const int64_t o = pos + __popc(b & ((1u << lane) - 1u));
        if (set && o < capacity) idx[o] = i;
        pos += __popc(b);
    }
}

// Device address of the scratch array ON THE CURRENT DEVICE.  In host code the array's name is only a host-side stand-in
// (not a device address): the address comes from cudaGetSymbolAddress, looked up once per device with that device current.
static int mask_static_scratch_ptr(int64_t** out) {
    constexpr int MAXDEV = 64;
    static void* cache[MAXDEV] = {nullptr};
    static std::mutex mu;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAXDEV) {
        set_error("pcs_mask_to_indices_static: cannot identify the current device");
        return PCS_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (cache[dev] == nullptr) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_mask_static_scratch);
        if (e != cudaSuccess || p == nullptr) {
            set_error("pcs_mask_to_indices_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
            return PCS_ERR_CUDA;
        }
        cache[dev] = p;
    }
    *out = (int64_t*)cache[dev];
    return PCS_OK;
}

extern "C" int64_t pcs_mask_static_max_n(void) { return MK_STATIC_MAX_CHUNKS * (int64_t)MK_CHUNK; }

extern "C" int pcs_mask_to_indices_static(const void* mask_u8, int64_t n, void* total_out_i64, void* idx_out_i64,
                                          int64_t idx_capacity, void* stream) {
    PCS_REQUIRE(n >= 0 && idx_capacity >= 0 && total_out_i64 && (n == 0 || mask_u8) && (idx_capacity == 0 || idx_out_i64),
                "pcs_mask_to_indices_static: bad arguments");
    PCS_REQUIRE(n <= pcs_mask_static_max_n(),
                "pcs_mask_to_indices_static: mask of %lld entries exceeds the static scratch limit of %lld entries "
                "(use pcs_mask_count / pcs_mask_indices with a caller-provided workspace)",
                (long long)n, (long long)pcs_mask_static_max_n());
    int64_t* scratch = nullptr;
    const int rc = mask_static_scratch_ptr(&scratch);
    if (rc != PCS_OK) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t nchunks = (n + MK_CHUNK - 1) / MK_CHUNK;
    // every word of the scratch this call reads is written by this call first (counts[0 .. nchunks) by the count kernel,
    // turned into offsets by the scan): nothing has to start at zero, stale contents of earlier calls are never read
    if (nchunks > 0) mask_count_kernel<<<(unsigned)nchunks, SG_BLOCK, 0, s>>>((const uint8_t*)mask_u8, n, scratch);
    mask_scan_kernel<<<1, 32, 0, s>>>(scratch, nchunks, (int64_t*)total_out_i64);
    if (nchunks > 0 && idx_capacity > 0) {
        const int warps = SG_BLOCK / 32;
        mask_write_capped_kernel<<<(unsigned)((nchunks + warps - 1) / warps), SG_BLOCK, 0, s>>>(
            (const uint8_t*)mask_u8, n, scratch, (int64_t*)idx_out_i64, idx_capacity);
    }
    return check_launch("pcs_mask_to_indices_static");
}

// ------------------------------------------------------------------------------------------------ block pooling
// Pooling (reference linop/sampling.py:394-536 -> skimage.measure.block_reduce with cval = 0): the array (up to three
// axes, C order) is tiled by blocks of b0 x b1 x b2 samples, ragged border blocks being completed with zeros; every
// block is summed (scale = 1) or averaged over the FULL block size (scale = 1 / (b0 b1 b2)).  One thread per pooled
// sample, fp64 accumulation; the threads of a warp walk neighbouring blocks of one row, so their reads cover one
// contiguous span of 32 * b2 samples per block row.
struct PoolGeo { int64_t d[3], b[3], o[3]; };

template <typename T>
__global__ void __launch_bounds__(SG_BLOCK) pool_fwd_kernel(T* __restrict__ out, const T* __restrict__ x, PoolGeo g,
                                                            double scale) {
    const int64_t total = g.o[0] * g.o[1] * g.o[2];
    for (int64_t t = (int64_t)blockIdx.x * SG_BLOCK + threadIdx.x; t < total; t += (int64_t)gridDim.x * SG_BLOCK) {
        const int64_t k = t % g.o[2], j = (t / g.o[2]) % g.o[1], i = t / (g.o[2] * g.o[1]);
        const int64_t i1 = min((i + 1) * g.b[0], g.d[0]), j1 = min((j + 1) * g.b[1], g.d[1]),
                      k1 = min((k + 1) * g.b[2], g.d[2]);
        double acc = 0.0;
        for (int64_t a = i * g.b[0]; a < i1; ++a)
            for (int64_t b = j * g.b[1]; b < j1; ++b) {
                const T* row = x + (a * g.d[1] + b) * g.d[2];
                for (int64_t c = k * g.b[2]; c < k1; ++c) acc += (double)row[c];
            }
        out[t] = (T)(acc * scale);
    }
}

// "Unpooling" (linop/sampling.py:529-536): every sample of a block receives the block's pooled value (np.repeat along
// each axis, cropped to the input shape) -- for 'mean' this is NOT rescaled, exactly as the reference does.
template <typename T>
__global__ void __launch_bounds__(SG_BLOCK) pool_adj_kernel(T* __restrict__ out, const T* __restrict__ y, PoolGeo g) {
    const int64_t total = g.d[0] * g.d[1] * g.d[2];
    for (int64_t t = (int64_t)blockIdx.x * SG_BLOCK + threadIdx.x; t < total; t += (int64_t)gridDim.x * SG_BLOCK) {
        const int64_t c = t % g.d[2], b = (t / g.d[2]) % g.d[1], a = t / (g.d[2] * g.d[1]);
        out[t] = y[((a / g.b[0]) * g.o[1] + b / g.b[1]) * g.o[2] + c / g.b[2]];
    }
}

static bool pool_geo(PoolGeo& g, int64_t d0, int64_t d1, int64_t d2, int64_t b0, int64_t b1, int64_t b2) {
    const int64_t d[3] = {d0, d1, d2}, b[3] = {b0, b1, b2};
    for (int a = 0; a < 3; ++a) {
        if (d[a] < 1 || b[a] < 1) return false;
        g.d[a] = d[a]; g.b[a] = b[a]; g.o[a] = (d[a] + b[a] - 1) / b[a];
    }
    return true;
}

template <typename T>
static int pcs_pool_fwd_impl(T* out, const T* x, int64_t d0, int64_t d1, int64_t d2, int64_t b0, int64_t b1,
                             int64_t b2, double scale, cudaStream_t s) {
    PoolGeo g;
    PCS_REQUIRE(pool_geo(g, d0, d1, d2, b0, b1, b2) && out && x, "pcs_pool_fwd: bad arguments");
    pool_fwd_kernel<T><<<grid_for(g.o[0] * g.o[1] * g.o[2], SG_BLOCK, 8), SG_BLOCK, 0, s>>>(out, x, g, scale);
    return check_launch("pcs_pool_fwd");
}
template <typename T>
static int pcs_pool_adj_impl(T* out, const T* y, int64_t d0, int64_t d1, int64_t d2, int64_t b0, int64_t b1,
                             int64_t b2, cudaStream_t s) {
    PoolGeo g;
    PCS_REQUIRE(pool_geo(g, d0, d1, d2, b0, b1, b2) && out && y, "pcs_pool_adj: bad arguments");
    pool_adj_kernel<T><<<grid_for(d0 * d1 * d2, SG_BLOCK, 8), SG_BLOCK, 0, s>>>(out, y, g);
    return check_launch("pcs_pool_adj");
}
PCS_EXPORT_REAL(pcs_pool_fwd,
                (void* out, const void* x, int64_t d0, int64_t d1, int64_t d2, int64_t b0, int64_t b1, int64_t b2,
                 double scale, void* stream),
                ((T*)out, (const T*)x, d0, d1, d2, b0, b1, b2, scale, (cudaStream_t)stream))
PCS_EXPORT_REAL(pcs_pool_adj,
                (void* out, const void* y, int64_t d0, int64_t d1, int64_t d2, int64_t b0, int64_t b1, int64_t b2,
                 void* stream),
                ((T*)out, (const T*)y, d0, d1, d2, b0, b1, b2, (cudaStream_t)stream))
