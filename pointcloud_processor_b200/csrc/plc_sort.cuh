// plc_sort.cuh — stable LSD radix sort of (uint32 key, uint32 value) pairs, 8 bits per pass.
//
// Used by K7 (voxel downsample): pcl::VoxelGrid orders its output by ascending voxel index and sums
// each voxel's members in sorted order.  A *stable* sort with value = point
// index reproduces the order of sorting (voxel index, point index) pairs, which makes the float32
// centroid sums reproducible and identical to the CPU restatement.
//
// Per pass: k_radix_hist (digit histogram per 2048-element tile) -> exclusive scan of the
// [digit][tile] table -> k_radix_scatter (warp-synchronous stable ranking with __match_any_sync).
// HBM traffic per pass: N*(8 rd + 8 rd + 8 wr) bytes.
#pragma once
#include "plc_scan.cuh"

namespace plc {

constexpr int kSortWarps = 8;
constexpr int kSortRounds = 8;
constexpr int kSortTile = kSortWarps * 32 * kSortRounds;   // 2048

static __global__ void __launch_bounds__(kSortWarps * 32)
k_radix_hist(const uint32_t* __restrict__ keys, unsigned n, const unsigned* __restrict__ n_dev, int shift,
             uint32_t* __restrict__ table, unsigned n_tiles) {
    __shared__ uint32_t hist[256];
    if (n_dev) n = min(n, *n_dev);          // the count lives on the device: `n` sized the grid, tiles past it count zero
    hist[threadIdx.x] = 0;
    __syncthreads();
    unsigned base = blockIdx.x * kSortTile;
#pragma unroll
    for (int r = 0; r < kSortRounds; ++r) {
        unsigned i = base + r * (kSortWarps * 32) + threadIdx.x;
        if (i < n) atomicAdd(&hist[(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    table[(size_t)threadIdx.x * n_tiles + blockIdx.x] = hist[threadIdx.x];
}

static __global__ void __launch_bounds__(kSortWarps * 32)
k_radix_scatter(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals, unsigned n,
                const unsigned* __restrict__ n_dev, int shift, const uint32_t* __restrict__ table, unsigned n_tiles,
                uint32_t* __restrict__ out_keys, uint32_t* __restrict__ out_vals) {
    __shared__ uint32_t wbase[kSortWarps][256];
    if (n_dev) n = min(n, *n_dev);
    if (blockIdx.x * kSortTile >= n) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int k = threadIdx.x; k < kSortWarps * 256; k += blockDim.x) (&wbase[0][0])[k] = 0;
    __syncthreads();
    // warp `warp` owns elements [tile + warp*256, tile + warp*256 + 256), 32 per round, in order
    const unsigned first = blockIdx.x * kSortTile + warp * (32 * kSortRounds);
    uint32_t key[kSortRounds], val[kSortRounds];
#pragma unroll
    for (int r = 0; r < kSortRounds; ++r) {
        unsigned i = first + r * 32 + lane;
        bool live = i < n;
        key[r] = live ? keys[i] : 0u;
        val[r] = live ? vals[i] : 0u;
        unsigned d = live ? ((key[r] >> shift) & 255u) : 256u;
        unsigned peers = __match_any_sync(0xffffffffu, d);
        if (live && lane == __ffs(peers) - 1) wbase[warp][d] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    // per digit: global base of this tile, then earlier warps of the tile
    {
        unsigned d = threadIdx.x;   // blockDim.x == 256
        uint32_t run = table[(size_t)d * n_tiles + blockIdx.x];
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w) {
            uint32_t t = wbase[w][d];
            wbase[w][d] = run;
            run += t;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kSortRounds; ++r) {
        unsigned i = first + r * 32 + lane;
        bool live = i < n;
        unsigned d = live ? ((key[r] >> shift) & 255u) : 256u;
        unsigned peers = __match_any_sync(0xffffffffu, d);
        uint32_t dst = 0;
        if (live) dst = wbase[warp][d] + __popc(peers & ((1u << lane) - 1u));
        __syncwarp();
        if (live && lane == __ffs(peers) - 1) wbase[warp][d] += __popc(peers);
        __syncwarp();
        if (live) {
            out_keys[dst] = key[r];
            out_vals[dst] = val[r];
        }
    }
}

inline size_t sort_table_entries(size_t n) { return 256 * ((n + kSortTile - 1) / kSortTile); }

// Sorts n pairs by the low `key_bits` bits of the key.  With `n_dev` the number of pairs is min(n, *n_dev), read on
// the device (n then only sizes the launches), so a pipeline need not come back to the host for it.  (keys_a, vals_a) hold the input; the result
// lands in whichever buffer pair the returned index names (0 = a, 1 = b).  `table` needs
// sort_table_entries(n) uint32; `lb` is the context\'s look-back workspace (plc_scan.cuh).
// Returns the number of kernel launches through *launches.
inline int radix_sort_pairs(uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b, unsigned n,
                            int key_bits, uint32_t* table, LookbackWs& lb, cudaStream_t s,
                            uint64_t* launches, const unsigned* n_dev = nullptr) {
    if (n == 0) return 0;
    unsigned n_tiles = (n + kSortTile - 1) / kSortTile;
    int cur = 0;
    for (int shift = 0; shift < key_bits; shift += 8) {
        const uint32_t* kin = cur ? keys_b : keys_a;
        const uint32_t* vin = cur ? vals_b : vals_a;
        uint32_t* kout = cur ? keys_a : keys_b;
        uint32_t* vout = cur ? vals_a : vals_b;
        k_radix_hist<<<n_tiles, kSortWarps * 32, 0, s>>>(kin, n, n_dev, shift, table, n_tiles);
        const int scan_launches = scan_exclusive_u32(table, (size_t)256 * n_tiles, lb, s);
        k_radix_scatter<<<n_tiles, kSortWarps * 32, 0, s>>>(kin, vin, n, n_dev, shift, table, n_tiles, kout, vout);
        if (launches) *launches += 2 + scan_launches;
        cur ^= 1;
    }
    return cur;
}

}  // namespace plc
