// radix_sort.cu — hand-written LSD radix sort of (64-bit key, 32-bit value)
// pairs, used to order primitives by Morton code for the LBVH. It replaces the
// per-node sequential thrust::sort the reference runs from a single device
// thread (hitable/bvh.h:84-93).
//
// 8-bit digits. Per pass three kernels over a fixed grid of at most
// RS_MAX_BLOCKS contiguous chunks:
//   rs_hist    : per-chunk digit histogram (shared-memory atomics)
//   rs_scan    : exclusive scan of the [digit][chunk] table (one block)
//   rs_scatter : stable scatter; each warp ranks its 32x16 slice with
//                __match_any_sync + warp-private counters, a 256-thread step
//                turns the counters into warp offsets, and a running per-digit
//                base carries the chunk from tile to tile.
// Traffic per pass: 8 B/elem (hist) + 12 B read + 12 B write (scatter).
#include "crt_internal.h"

namespace crt {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;   // 4096 elements
constexpr int RS_BINS = 256;
constexpr int RS_MAX_BLOCKS = 148 * 4;

__global__ void __launch_bounds__(RS_THREADS)
rs_hist(const uint64_t* __restrict__ keys, uint32_t n, int shift, uint32_t tiles_per_block,
        uint32_t* __restrict__ hist)
{
    __shared__ uint32_t sh[RS_BINS];
    sh[threadIdx.x] = 0;
    __syncthreads();
    const uint64_t begin = (uint64_t)blockIdx.x * tiles_per_block * RS_TILE;
    uint64_t end = begin + (uint64_t)tiles_per_block * RS_TILE;
    if (end > n) end = n;
    for (uint64_t i = begin + threadIdx.x; i < end; i += RS_THREADS) {
        const uint32_t d = (uint32_t)(keys[i] >> shift) & 255u;
        // aggregate equal digits inside the warp before touching shared memory
        const uint32_t act = __activemask();
        const uint32_t peers = __match_any_sync(act, d);
        if ((peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0u) atomicAdd(&sh[d], __popc(peers));
    }
    __syncthreads();
    hist[threadIdx.x * gridDim.x + blockIdx.x] = sh[threadIdx.x];
}

// In-place exclusive scan of `total` counters by one block of 1024 threads.
__global__ void __launch_bounds__(1024) rs_scan(uint32_t* __restrict__ hist, uint32_t total)
{
    __shared__ uint32_t warp_sums[32];
    const uint32_t per = (total + 1023u) / 1024u;
    const uint32_t b = threadIdx.x * per;
    const uint32_t e = min(b + per, total);
    uint32_t sum = 0;
    for (uint32_t i = b; i < e; i++) sum += hist[i];
    // block exclusive scan of `sum`
    const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = sum;
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    if (lane == 31) warp_sums[w] = inc;
    __syncthreads();
    if (w == 0) {
        uint32_t ws = warp_sums[lane];
        uint32_t winc = ws;
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t v = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += v;
        }
        warp_sums[lane] = winc - ws;
    }
    __syncthreads();
    uint32_t run = warp_sums[w] + inc - sum;
    for (uint32_t i = b; i < e; i++) {
        const uint32_t c = hist[i];
        hist[i] = run;
        run += c;
    }
}

__global__ void __launch_bounds__(RS_THREADS)
rs_scatter(const uint64_t* __restrict__ kin, const uint32_t* __restrict__ vin, uint64_t* __restrict__ kout,
           uint32_t* __restrict__ vout, uint32_t n, int shift, uint32_t tiles_per_block,
           const uint32_t* __restrict__ hist)
{
    __shared__ uint32_t s_base[RS_BINS];
    __shared__ uint32_t s_wcount[RS_WARPS][RS_BINS];
    const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    s_base[tid] = hist[tid * gridDim.x + blockIdx.x];

    for (uint32_t t = 0; t < tiles_per_block; t++) {
        const uint64_t tile_begin = ((uint64_t)blockIdx.x * tiles_per_block + t) * RS_TILE;
        if (tile_begin >= n) break;
#pragma unroll
        for (int i = 0; i < RS_WARPS; i++) s_wcount[i][tid] = 0;
        __syncthreads();

        uint64_t key[RS_ITEMS];
        uint32_t val[RS_ITEMS];
        uint32_t rank[RS_ITEMS];
        const uint64_t sub = tile_begin + (uint64_t)w * (32 * RS_ITEMS);
#pragma unroll
        for (int j = 0; j < RS_ITEMS; j++) {
            const uint64_t i = sub + (uint64_t)j * 32 + lane;
            const bool valid = i < n;
            key[j] = valid ? kin[i] : ~0ull;
            val[j] = valid ? vin[i] : 0u;
        }
#pragma unroll
        for (int j = 0; j < RS_ITEMS; j++) {
            const uint64_t i = sub + (uint64_t)j * 32 + lane;
            const bool valid = i < n;
            const uint32_t d = valid ? ((uint32_t)(key[j] >> shift) & 255u) : 0xFFFFFFFFu;
            const uint32_t peers = __match_any_sync(0xffffffffu, d);
            const uint32_t before = __popc(peers & ((1u << lane) - 1u));
            uint32_t pre = 0;
            if (valid) pre = s_wcount[w][d];
            __syncwarp();
            if (valid && before == 0u) s_wcount[w][d] = pre + __popc(peers);
            __syncwarp();
            rank[j] = pre + before;
        }
        __syncthreads();

        // thread `tid` owns digit `tid`: counters -> exclusive offsets across warps
        uint32_t tile_total = 0;
#pragma unroll
        for (int i = 0; i < RS_WARPS; i++) {
            const uint32_t c = s_wcount[i][tid];
            s_wcount[i][tid] = tile_total;
            tile_total += c;
        }
        __syncthreads();

#pragma unroll
        for (int j = 0; j < RS_ITEMS; j++) {
            const uint64_t i = sub + (uint64_t)j * 32 + lane;
            if (i < n) {
                const uint32_t d = (uint32_t)(key[j] >> shift) & 255u;
                const uint32_t pos = s_base[d] + s_wcount[w][d] + rank[j];
                kout[pos] = key[j];
                vout[pos] = val[j];
            }
        }
        __syncthreads();
        s_base[tid] += tile_total;
    }
}

// Sorts n pairs by the key bits [begin_bit, end_bit). keys[0]/vals[0] hold the
// input; returns the index (0 or 1) of the buffers holding the result.
int radix_sort_pairs(cudaStream_t stream, uint64_t* keys[2], uint32_t* vals[2], uint32_t n, int begin_bit,
                     int end_bit, uint32_t* hist /* RS_BINS * RS_MAX_BLOCKS */)
{
    if (n == 0) return 0;
    const uint32_t tiles = (n + RS_TILE - 1) / RS_TILE;
    uint32_t nblocks = tiles < (uint32_t)RS_MAX_BLOCKS ? tiles : (uint32_t)RS_MAX_BLOCKS;
    const uint32_t tiles_per_block = (tiles + nblocks - 1) / nblocks;
    nblocks = (tiles + tiles_per_block - 1) / tiles_per_block;
    int cur = 0;
    for (int shift = begin_bit; shift < end_bit; shift += 8) {
        rs_hist<<<nblocks, RS_THREADS, 0, stream>>>(keys[cur], n, shift, tiles_per_block, hist);
        rs_scan<<<1, 1024, 0, stream>>>(hist, RS_BINS * nblocks);
        rs_scatter<<<nblocks, RS_THREADS, 0, stream>>>(keys[cur], vals[cur], keys[cur ^ 1], vals[cur ^ 1], n, shift,
                                                       tiles_per_block, hist);
        cur ^= 1;
    }
    return cur;
}

size_t radix_sort_hist_bytes() { return sizeof(uint32_t) * RS_BINS * RS_MAX_BLOCKS; }

}  // namespace crt
