// plc_scan.cuh — exclusive prefix sum over uint32 on the device (3 passes, 2048 items per block; one block for short arrays).
// Used by the voxeliser (voxel start offsets), the crop compaction and the radix sort.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace plc {

// ---- exclusive scan over n uint32 (in place), 3 passes, 2048 items per block -----------------
constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;

static __device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total) {
    __shared__ uint32_t warp_sums[kScanThreads / 32];
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane < kScanThreads / 32 ? warp_sums[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += t;
        }
        if (lane < kScanThreads / 32) warp_sums[lane] = w;
    }
    __syncthreads();
    uint32_t before = warp > 0 ? warp_sums[warp - 1] : 0;
    *total = warp_sums[kScanThreads / 32 - 1];
    uint32_t r = before + inc - v;
    __syncthreads();
    return r;
}

static __global__ void __launch_bounds__(kScanThreads)
k_scan_tiles(uint32_t* __restrict__ data, size_t n, uint32_t* __restrict__ tile_sums) {
    size_t base = (size_t)blockIdx.x * kScanTile + (size_t)threadIdx.x * kScanItems;
    uint32_t v[kScanItems], sum = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        v[k] = (base + k < n) ? data[base + k] : 0;
        sum += v[k];
    }
    uint32_t total;
    uint32_t ex = block_exclusive_scan(sum, &total);
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        if (base + k < n) data[base + k] = ex;
        ex += v[k];
    }
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// single block: exclusive scan of the tile sums (any count)
static __global__ void __launch_bounds__(kScanThreads)
k_scan_sums(uint32_t* __restrict__ sums, size_t n) {
    uint32_t carry = 0;
    for (size_t base = 0; base < n; base += kScanThreads) {
        size_t i = base + threadIdx.x;
        uint32_t v = i < n ? sums[i] : 0, total;
        uint32_t ex = block_exclusive_scan(v, &total);
        if (i < n) sums[i] = carry + ex;
        carry += total;
    }
}

static __global__ void __launch_bounds__(kScanThreads)
k_scan_add(uint32_t* __restrict__ data, size_t n, const uint32_t* __restrict__ tile_sums) {
    size_t base = (size_t)blockIdx.x * kScanTile + (size_t)threadIdx.x * kScanItems;
    uint32_t add = tile_sums[blockIdx.x];
#pragma unroll
    for (int k = 0; k < kScanItems; ++k)
        if (base + k < n) data[base + k] += add;
}


// One block walks the whole array, 2048 items per round with a running carry: for the short arrays that dominate
// the node-side stages (sort tables, lattice plans) one launch instead of three.
static __global__ void __launch_bounds__(kScanThreads)
k_scan_small(uint32_t* __restrict__ data, size_t n) {
    uint32_t carry = 0;
    for (size_t round = 0; round < n; round += kScanTile) {
        const size_t base = round + (size_t)threadIdx.x * kScanItems;
        uint32_t v[kScanItems], sum = 0;
#pragma unroll
        for (int k = 0; k < kScanItems; ++k) {
            v[k] = (base + k < n) ? data[base + k] : 0;
            sum += v[k];
        }
        uint32_t total;
        uint32_t ex = carry + block_exclusive_scan(sum, &total);
#pragma unroll
        for (int k = 0; k < kScanItems; ++k) {
            if (base + k < n) data[base + k] = ex;
            ex += v[k];
        }
        carry += total;
    }
}
constexpr size_t kScanSmallLimit = 4 * kScanTile;      // 8192 items: four rounds of one block

inline size_t scan_tmp_bytes(size_t n) { return ((n + kScanTile - 1) / kScanTile + 1) * sizeof(uint32_t); }

// In-place exclusive scan of data[0..n); tmp needs scan_tmp_bytes(n).  Returns the number of kernels launched
// (1 for short arrays, else 3).
inline int scan_exclusive_u32(uint32_t* data, size_t n, uint32_t* tmp, cudaStream_t s) {
    if (n == 0) return 0;
    if (n <= kScanSmallLimit) {
        k_scan_small<<<1, kScanThreads, 0, s>>>(data, n);
        return 1;
    }
    size_t tiles = (n + kScanTile - 1) / kScanTile;
    k_scan_tiles<<<(unsigned)tiles, kScanThreads, 0, s>>>(data, n, tmp);
    k_scan_sums<<<1, kScanThreads, 0, s>>>(tmp, tiles);
    k_scan_add<<<(unsigned)tiles, kScanThreads, 0, s>>>(data, n, tmp);
    return 3;
}

}  // namespace plc
