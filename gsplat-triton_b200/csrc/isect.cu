// isect_tiles (tile-bbox count, key emit, sort) and isect_offset_encode.
//
// Replaces (behaviour, not code): gsplat/triton_impl/isect_tiles.py:14-282, isect_offset.py:9-63,
// gsplat/triton_impl/radix_sort/radix_sort.cu, gsplat/cuda/csrc/IntersectTile.cu:24-344.
// Integer outputs are bit-exact against oracle/gs_oracle.py::isect_tiles / isect_offset_encode.
//
// Key layout (int64): camera << (32 + tile_bits) | tile << 32 | bitcast<int32>(depth).
#include "common.cuh"
#include "gsplat_b200.h"

#include <cub/device/device_radix_sort.cuh>

#include <cstdint>

namespace {

constexpr int kThreads = 256;

struct TileBox {
    int x0, x1, y0, y1;
};

// floor((p - r)/ts) .. ceil((p + r)/ts), clamped to the tile grid; max is exclusive.
// Written as p/ts -/+ r/ts like _torch_impl._isect_tiles:351-358 (identical to the Triton form for
// power-of-two tile sizes, where the division is exact).
__device__ __forceinline__ TileBox tile_box(float mx, float my, int radius, int tile_size, int tw, int th) {
    float ts = (float)tile_size;
    float tr = __fdiv_rn((float)radius, ts);
    float tx = __fdiv_rn(mx, ts), ty = __fdiv_rn(my, ts);
    TileBox b;
    b.x0 = min(max((int)floorf(tx - tr), 0), tw);
    b.x1 = min(max((int)ceilf(tx + tr), 0), tw);
    b.y0 = min(max((int)floorf(ty - tr), 0), th);
    b.y1 = min(max((int)ceilf(ty + tr), 0), th);
    return b;
}

__device__ __forceinline__ int block_exclusive_scan(int v, int *s_warp, int &block_total) {
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    int base = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) {
        int s = s_warp[w];
        base += (w < warp) ? s : 0;
        total += s;
    }
    block_total = total;
    return base + inc - v;
}

__global__ void __launch_bounds__(kThreads)
isect_count_kernel(int64_t P, const float *__restrict__ means2d, const int32_t *__restrict__ radii, int tile_size,
                   int tw, int th, int32_t *__restrict__ tiles_per_gauss, int32_t *__restrict__ block_sums) {
    __shared__ int s_warp[kThreads / 32];
    int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    int cnt = 0;
    if (i < P) {
        int r = radii[i];
        if (r > 0) {
            float2 m = *reinterpret_cast<const float2 *>(means2d + 2 * i);
            TileBox b = tile_box(m.x, m.y, r, tile_size, tw, th);
            cnt = (b.x1 - b.x0) * (b.y1 - b.y0);
        }
        tiles_per_gauss[i] = cnt;
    }
    int s = cnt;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < kThreads / 32; ++w) t += s_warp[w];
        block_sums[blockIdx.x] = t;
    }
}

// single-block exclusive scan int32 -> int64 (n is the number of 256-row blocks: at most a few 100k). Every thread
// owns ceil(n / 1024) consecutive entries: one local sum, one warp scan, one hand-over between the warps, no loop of
// block-wide rounds (3907 entries at cfg2: one round trip instead of four)
__global__ void __launch_bounds__(1024)
scan_block_sums_kernel(int64_t n, const int32_t *__restrict__ in, int64_t *__restrict__ out,
                       int64_t *__restrict__ total_out) {
    gsb_chain_enter();
    __shared__ long long s_warp[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t per = (n + 1023) / 1024;
    const int64_t begin = min(n, (int64_t)threadIdx.x * per), end = min(n, begin + per);
    long long sum = 0;
    for (int64_t i = begin; i < end; ++i) sum += in[i];
    long long inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        long long t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    long long run = inc - sum, total = 0;
#pragma unroll
    for (int w = 0; w < 32; ++w) {
        const long long v = s_warp[w];
        run += (w < warp) ? v : 0;
        total += v;
    }
    for (int64_t i = begin; i < end; ++i) {
        out[i] = run;
        run += in[i];
    }
    if (threadIdx.x == 0) *total_out = total;
}

template <bool PACKED>
__global__ void __launch_bounds__(kThreads)
isect_emit_kernel(int64_t P, int N, const float *__restrict__ means2d, const int32_t *__restrict__ radii,
                  const float *__restrict__ depths, const int64_t *__restrict__ camera_ids, int tile_size, int tw,
                  int th, int tile_n_bits, const int32_t *__restrict__ tiles_per_gauss,
                  const int64_t *__restrict__ block_offsets, int64_t *__restrict__ isect_ids,
                  int32_t *__restrict__ flatten_ids) {
    __shared__ int s_warp[kThreads / 32];
    int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    int cnt = i < P ? tiles_per_gauss[i] : 0;
    int block_total;
    int local = block_exclusive_scan(cnt, s_warp, block_total);
    if (cnt == 0) return;
    int64_t out = block_offsets[blockIdx.x] + local;
    float2 m = *reinterpret_cast<const float2 *>(means2d + 2 * i);
    TileBox b = tile_box(m.x, m.y, radii[i], tile_size, tw, th);
    int64_t cam = PACKED ? camera_ids[i] : (i / N);
    int64_t hi = cam << (32 + tile_n_bits);
    int64_t depth_bits = (int64_t)__float_as_int(depths[i]);  // sign-extending like the reference casts
    for (int y = b.y0; y < b.y1; ++y) {
        for (int x = b.x0; x < b.x1; ++x) {
            int64_t tile = (int64_t)(y * tw + x);
            isect_ids[out] = hi | (tile << 32) | depth_bits;
            flatten_ids[out] = (int32_t)i;
            ++out;
        }
    }
}

__global__ void __launch_bounds__(kThreads)
isect_offsets_kernel(int64_t n_isects, const int64_t *__restrict__ isect_ids, int n_tiles, int tile_n_bits,
                     int64_t total_tiles, int32_t *__restrict__ offsets) {
    int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (i >= n_isects) return;
    const int64_t tile_mask = ((int64_t)1 << tile_n_bits) - 1;
    int64_t id = isect_ids[i] >> 32;
    int64_t cur = (id >> tile_n_bits) * n_tiles + (id & tile_mask);
    if (i == 0) {
        for (int64_t t = 0; t <= cur; ++t) offsets[t] = 0;
    } else {
        int64_t pid = isect_ids[i - 1] >> 32;
        if (pid != id) {
            int64_t prev = (pid >> tile_n_bits) * n_tiles + (pid & tile_mask);
            for (int64_t t = prev + 1; t <= cur; ++t) offsets[t] = (int32_t)i;
        }
    }
    if (i == n_isects - 1) {
        for (int64_t t = cur + 1; t < total_tiles; ++t) offsets[t] = (int32_t)n_isects;
    }
}

// ---- depth-presorted path -----------------------------------------------------------------------------------
// A stable sort of the (camera | tile | depth) keys in emission order equals: order the ROWS by (camera, depth
// bits) with a stable sort (ties keep ascending row index), emit every row's tiles in that order, then stably
// sort the intersections by their (camera | tile) bits alone. The wide pass runs over the P rows (1 M at cfg2)
// instead of the I intersections (6.3 M), and the intersection sort shrinks from ceil((32+tb+cb)/8) = 6 radix
// passes over 12-byte pairs to ceil((tb+cb)/8) = 2 passes over 8-byte pairs.
template <bool PACKED>
__global__ void __launch_bounds__(kThreads)
row_keys_kernel(int64_t P, int N, const float *__restrict__ depths, const int64_t *__restrict__ camera_ids,
                int64_t *__restrict__ keys, int32_t *__restrict__ rows) {
    int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (i >= P) return;
    if (keys != nullptr) {
        int64_t cam = PACKED ? camera_ids[i] : (i / N);
        keys[i] = (cam << 32) | (int64_t)(unsigned)__float_as_int(depths[i]);
    }
    rows[i] = (int32_t)i;
}

// natural order (coalesced): rows[i] = i, the tile count of row i and, for the depth-ordered emission, ONE 16-byte
// record per row (x0 | y0 << 16, width, camera << tile_bits, count): the emit pass then gathers a single sector per
// row instead of four (means2d, radius, count, camera id)
// shared-memory digit histograms of 32-bit keys for the onesweep sort that follows (gsb_radix_prepare): the producer
// of the keys counts the digits itself, the sort then needs no histogram pass. `passes` x 256 counters.
__device__ __forceinline__ void hist_clear(uint32_t *s_hist, int passes) {
    for (int i = threadIdx.x; i < passes * 256; i += blockDim.x) s_hist[i] = 0u;
}
__device__ __forceinline__ void hist_flush(const uint32_t *s_hist, uint32_t *hist, int passes) {
    for (int i = threadIdx.x; i < passes * 256; i += blockDim.x) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&hist[i], c);
    }
}

template <bool PACKED>
__global__ void __launch_bounds__(kThreads)
row_tiles_kernel(int64_t P, int N, const float *__restrict__ means2d, const int32_t *__restrict__ radii,
                 const int64_t *__restrict__ camera_ids, int tile_size, int tw, int th, int tile_n_bits,
                 int32_t *__restrict__ rows, int32_t *__restrict__ tiles_per_gauss, int4 *__restrict__ records,
                 const float *__restrict__ depths, uint32_t *__restrict__ row_keys, uint32_t *__restrict__ key_hist) {
    gsb_chain_enter();
    __shared__ uint32_t s_hist[4 * 256];
    if (key_hist != nullptr) {
        hist_clear(s_hist, 4);
        __syncthreads();
    }
    for (int64_t i0 = (int64_t)blockIdx.x * kThreads; i0 < P; i0 += (int64_t)gridDim.x * kThreads) {
        const int64_t i = i0 + threadIdx.x;
        const bool live = i < P;
        int cnt = 0;
        int4 rec = make_int4(0, 0, 0, 0);
        const int r = live ? radii[i] : 0;
        if (r > 0) {
            float2 m = *reinterpret_cast<const float2 *>(means2d + 2 * i);
            TileBox b = tile_box(m.x, m.y, r, tile_size, tw, th);
            cnt = (b.x1 - b.x0) * (b.y1 - b.y0);
            const uint32_t cam = (uint32_t)(PACKED ? camera_ids[i] : (i / N));
            rec = make_int4(b.x0 | (b.y0 << 16), b.x1 - b.x0, (int)(cam << tile_n_bits), cnt);
        }
        if (live) {
            rows[i] = (int32_t)i;
            tiles_per_gauss[i] = cnt;
            if (records != nullptr) records[i] = rec;
        }
        // the depth-sort key of the row: its depth bits, or all ones for a row without tiles, so that the rows that
        // emit intersections form one contiguous run of the sorted order (the binning kernels of isect_bin.cu rely on it)
        if (row_keys != nullptr) {
            uint32_t key = 0xffffffffu;
            if (live) {
                if (cnt > 0) key = (uint32_t)__float_as_int(depths[i]);
                row_keys[i] = key;
            }
            if (key_hist != nullptr) gsb_hist_add(s_hist, key, live, 4, 32);
        }
    }
    if (key_hist != nullptr) {
        __syncthreads();
        hist_flush(s_hist, key_hist, 4);
    }
}

// k-th thread handles row perm[k]: gathers its 16-byte record into depth order (the binning kernels then read the
// records coalesced) + block sums of the tile counts in that order
__global__ void __launch_bounds__(kThreads)
count_perm_records_kernel(int64_t P, const int32_t *__restrict__ perm, const int4 *__restrict__ records,
                          int4 *__restrict__ sorted_records, int32_t *__restrict__ block_sums) {
    __shared__ int s_warp[kThreads / 32];
    int64_t k = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    int cnt = 0;
    if (k < P) {
        const int4 rec = __ldg(records + perm[k]);
        sorted_records[k] = rec;
        cnt = rec.w;
    }
    int sum = cnt;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < kThreads / 32; ++w) t += s_warp[w];
        block_sums[blockIdx.x] = t;
    }
}

// k-th thread handles row perm[k]: tile count (also stored in natural order) + block sums in permuted order
__global__ void __launch_bounds__(kThreads)
count_perm_kernel(int64_t P, const int32_t *__restrict__ perm, const float *__restrict__ means2d,
                  const int32_t *__restrict__ radii, int tile_size, int tw, int th,
                  int32_t *__restrict__ tiles_per_gauss, int32_t *__restrict__ block_sums) {
    gsb_chain_enter();
    __shared__ int s_warp[kThreads / 32];
    int64_t k = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    int cnt = 0;
    if (k < P) {
        int64_t i = perm[k];
        if (means2d == nullptr) {
            cnt = tiles_per_gauss[i];  // counted in natural order by row_tiles_kernel: one 4-byte gather per row
        } else {
            int r = radii[i];
            if (r > 0) {
                float2 m = *reinterpret_cast<const float2 *>(means2d + 2 * i);
                TileBox b = tile_box(m.x, m.y, r, tile_size, tw, th);
                cnt = (b.x1 - b.x0) * (b.y1 - b.y0);
            }
            tiles_per_gauss[i] = cnt;
        }
    }
    int sum = cnt;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < kThreads / 32; ++w) t += s_warp[w];
        block_sums[blockIdx.x] = t;
    }
}

// Block-cooperative emission: the CTA's 256 rows (in permuted = depth order) describe their tile rectangles in shared
// memory, then the threads walk the CTA's OUTPUT range slot by slot (binary search for the owning row), so every
// store instruction writes 32 consecutive words, whatever the spread of tiles per row.
template <bool PACKED>
__global__ void __launch_bounds__(kThreads)
emit_perm_kernel(int64_t P, int N, const int32_t *__restrict__ perm, const float *__restrict__ means2d,
                 const int32_t *__restrict__ radii, const int64_t *__restrict__ camera_ids, int tile_size, int tw,
                 int th, int tile_n_bits, const int32_t *__restrict__ tiles_per_gauss,
                 const int64_t *__restrict__ block_offsets, uint32_t *__restrict__ tile_keys,
                 int32_t *__restrict__ flatten_ids, int64_t capacity, const int4 *__restrict__ records, int64_t nblk,
                 uint32_t *__restrict__ key_hist, int key_bits) {
    gsb_chain_enter();
    __shared__ int s_warp[kThreads / 32];
    __shared__ int s_start[kThreads];  // first output slot of each row inside the CTA's range
    __shared__ int s_xy[kThreads];     // x0 | y0 << 16
    __shared__ int s_w[kThreads];
    __shared__ int s_row[kThreads];
    __shared__ uint32_t s_hi[kThreads];
    __shared__ uint32_t s_hist[4 * 256];  // digit histograms of the emitted words for the tile sort (key_hist)
    const int tid = threadIdx.x;
    const int passes = (key_bits + 7) / 8;
    if (key_hist != nullptr) hist_clear(s_hist, passes);
    // a CTA takes every gridDim.x-th block of 256 rows (in permuted = depth order)
    for (int64_t blk = blockIdx.x; blk < nblk; blk += gridDim.x) {
        int64_t k = blk * kThreads + tid;
        int64_t i = k < P ? perm[k] : 0;
        int cnt = 0;
        if (records != nullptr) {
            int4 rec = make_int4(0, 0, 0, 0);
            if (k < P) rec = __ldg(records + i);
            cnt = rec.w;
            s_xy[tid] = rec.x;
            s_w[tid] = rec.y;
            s_hi[tid] = (uint32_t)rec.z;
        } else {
            cnt = k < P ? tiles_per_gauss[i] : 0;
            if (cnt > 0) {
                float2 m = *reinterpret_cast<const float2 *>(means2d + 2 * i);
                TileBox b = tile_box(m.x, m.y, radii[i], tile_size, tw, th);
                s_xy[tid] = b.x0 | (b.y0 << 16);
                s_w[tid] = b.x1 - b.x0;
                s_hi[tid] = (uint32_t)(PACKED ? camera_ids[i] : (i / N)) << tile_n_bits;
            }
        }
        s_row[tid] = (int)i;
        int block_total;
        s_start[tid] = block_exclusive_scan(cnt, s_warp, block_total);
        __syncthreads();
        const int64_t base = block_offsets[blk];
        for (int j0 = 0; j0 < block_total; j0 += kThreads) {
            const int j = j0 + tid;
            bool live = j < block_total;
            uint32_t key = 0u;
            if (live) {
                // last row r with s_start[r] <= j (rows without tiles share their successor's start and are skipped)
                int lo = 0, hi_ = kThreads;  // invariant: s_start[lo] <= j < s_start[hi_] (s_start[kThreads] = block_total)
#pragma unroll
                for (int step = 0; step < 8; ++step) {
                    const int mid = (lo + hi_) >> 1;
                    if (s_start[mid] <= j) lo = mid; else hi_ = mid;
                }
                const int t = j - s_start[lo];
                const int w = s_w[lo], xy = s_xy[lo];
                const int dy = t / w, dx = t - dy * w;
                const int64_t out = base + j;
                live = out < capacity;  // speculative capacity (gsb_isect_emit_perm_cap): never write past the buffers
                if (live) {
                    key = s_hi[lo] | (uint32_t)(((xy >> 16) + dy) * tw + (xy & 0xffff) + dx);
                    tile_keys[out] = key;
                    flatten_ids[out] = s_row[lo];
                }
            }
            if (key_hist != nullptr) gsb_hist_add(s_hist, key, live, passes, key_bits);
        }
        __syncthreads();  // the row tables are rewritten by the next block
    }
    if (key_hist != nullptr) {
        __syncthreads();
        hist_flush(s_hist, key_hist, passes);
    }
}

// slots [n_isects, capacity) of a speculatively sized buffer: keys that sort behind every real (camera | tile) word
__global__ void __launch_bounds__(kThreads)
pad_tail_kernel(const int64_t *__restrict__ n_isects_dev, int64_t capacity, uint32_t *__restrict__ tile_keys,
                int32_t *__restrict__ flatten_ids) {
    gsb_chain_enter();
    const int64_t n = *n_isects_dev;
    for (int64_t i = n + (int64_t)blockIdx.x * kThreads + threadIdx.x; i < capacity; i += (int64_t)gridDim.x * kThreads) {
        tile_keys[i] = 0xffffffffu;
        flatten_ids[i] = 0;
    }
}

// sorted (camera | tile) words + row ids -> the reference's 64-bit keys and the per-tile offsets
__global__ void __launch_bounds__(kThreads)
finalize_kernel(int64_t n_isects, const int64_t *__restrict__ n_isects_dev, const uint32_t *__restrict__ tile_keys,
                const int32_t *__restrict__ flatten_ids, const float *__restrict__ depths, int n_tiles,
                int tile_n_bits, int64_t total_tiles, int64_t *__restrict__ isect_ids, int32_t *__restrict__ offsets) {
    gsb_chain_enter();
    int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (n_isects_dev != nullptr) {  // n_isects is the buffer capacity, the real count lives on the device
        const int64_t n = *n_isects_dev;
        n_isects = n < n_isects ? n : n_isects;
        if (n_isects == 0) {
            if (offsets != nullptr)
                for (int64_t t = i; t < total_tiles; t += (int64_t)gridDim.x * kThreads) offsets[t] = 0;
            return;
        }
    }
    if (i >= n_isects) return;
    const uint32_t key = tile_keys[i];
    // either output may be NULL: the ids are not needed by the compositor and can be written on another stream
    if (isect_ids != nullptr)
        isect_ids[i] = ((int64_t)key << 32) | (int64_t)(unsigned)__float_as_int(depths[flatten_ids[i]]);
    if (offsets == nullptr) return;
    const uint32_t tile_mask = (1u << tile_n_bits) - 1u;
    const int64_t cur = (int64_t)(key >> tile_n_bits) * n_tiles + (key & tile_mask);
    if (i == 0) {
        for (int64_t t = 0; t <= cur; ++t) offsets[t] = 0;
    } else {
        const uint32_t pkey = tile_keys[i - 1];
        if (pkey != key) {
            const int64_t prev = (int64_t)(pkey >> tile_n_bits) * n_tiles + (pkey & tile_mask);
            for (int64_t t = prev + 1; t <= cur; ++t) offsets[t] = (int32_t)i;
        }
    }
    if (i == n_isects - 1) {
        for (int64_t t = cur + 1; t < total_tiles; ++t) offsets[t] = (int32_t)n_isects;
    }
}

// The same pass, four consecutive intersections per thread: one 16-byte load of the (camera | tile) words and one of
// the row ids, the four depth gathers in flight together, the 64-bit ids written as two 16-byte stores. (One
// intersection per thread leaves a dependent id -> depth load chain per thread and 4-byte requests: 2.8 TB/s at cfg2.)
// tile_keys, flatten_ids and isect_ids must be 16-byte aligned.
__global__ void __launch_bounds__(kThreads)
finalize4_kernel(int64_t n_isects, const int64_t *__restrict__ n_isects_dev, const uint32_t *__restrict__ tile_keys,
                 const int32_t *__restrict__ flatten_ids, const float *__restrict__ depths, int n_tiles,
                 int tile_n_bits, int64_t total_tiles, int64_t *__restrict__ isect_ids, int32_t *__restrict__ offsets) {
    const int64_t q = (int64_t)blockIdx.x * kThreads + threadIdx.x;
    if (n_isects_dev != nullptr) {
        const int64_t n = *n_isects_dev;
        n_isects = n < n_isects ? n : n_isects;
        if (n_isects == 0) {
            if (offsets != nullptr)
                for (int64_t t = q; t < total_tiles; t += (int64_t)gridDim.x * kThreads) offsets[t] = 0;
            return;
        }
    }
    const int64_t i0 = 4 * q;
    if (i0 >= n_isects) return;
    const int cnt = (int)(n_isects - i0 < 4 ? n_isects - i0 : 4);
    uint32_t key[4] = {0u, 0u, 0u, 0u};
    int32_t row[4] = {0, 0, 0, 0};
    if (cnt == 4) {
        const uint4 k4 = *reinterpret_cast<const uint4 *>(tile_keys + i0);
        const int4 r4 = *reinterpret_cast<const int4 *>(flatten_ids + i0);
        key[0] = k4.x; key[1] = k4.y; key[2] = k4.z; key[3] = k4.w;
        row[0] = r4.x; row[1] = r4.y; row[2] = r4.z; row[3] = r4.w;
    } else {
        for (int v = 0; v < cnt; ++v) {
            key[v] = tile_keys[i0 + v];
            row[v] = flatten_ids[i0 + v];
        }
    }
    const uint32_t before = i0 > 0 ? tile_keys[i0 - 1] : 0u;
    if (isect_ids != nullptr) {
        int64_t id[4];
#pragma unroll
        for (int v = 0; v < 4; ++v)
            id[v] = ((int64_t)key[v] << 32) | (int64_t)(unsigned)__float_as_int(v < cnt ? depths[row[v]] : 0.f);
        if (cnt == 4) {
            longlong2 *dst = reinterpret_cast<longlong2 *>(isect_ids + i0);
            dst[0] = make_longlong2(id[0], id[1]);
            dst[1] = make_longlong2(id[2], id[3]);
        } else {
            for (int v = 0; v < cnt; ++v) isect_ids[i0 + v] = id[v];
        }
    }
    if (offsets == nullptr) return;
    const uint32_t tile_mask = (1u << tile_n_bits) - 1u;
#pragma unroll
    for (int v = 0; v < 4; ++v) {
        if (v >= cnt) break;
        const int64_t i = i0 + v;
        const int64_t cur = (int64_t)(key[v] >> tile_n_bits) * n_tiles + (key[v] & tile_mask);
        if (i == 0) {
            for (int64_t t = 0; t <= cur; ++t) offsets[t] = 0;
        } else {
            const uint32_t pkey = v ? key[v - 1] : before;
            if (pkey != key[v]) {
                const int64_t prev = (int64_t)(pkey >> tile_n_bits) * n_tiles + (pkey & tile_mask);
                for (int64_t t = prev + 1; t <= cur; ++t) offsets[t] = (int32_t)i;
            }
        }
        if (i == n_isects - 1) {
            for (int64_t t = cur + 1; t < total_tiles; ++t) offsets[t] = (int32_t)n_isects;
        }
    }
}

}  // namespace

extern "C" int64_t gsb_isect_num_blocks(int64_t P) { return (P + kThreads - 1) / kThreads; }

extern "C" int gsb_isect_count(int64_t P, const float *means2d, const int32_t *radii, int tile_size, int tile_width,
                               int tile_height, int32_t *tiles_per_gauss, int32_t *block_sums,
                               int64_t *block_offsets, int64_t *n_isects_out, cudaStream_t stream) {
    GSB_REQUIRE(P >= 0 && tile_size > 0 && tile_width >= 0 && tile_height >= 0, "bad sizes");
    int64_t nblk = gsb_isect_num_blocks(P);
    if (P > 0) {
        isect_count_kernel<<<(unsigned)nblk, kThreads, 0, stream>>>(P, means2d, radii, tile_size, tile_width,
                                                                   tile_height, tiles_per_gauss, block_sums);
        if (int e = gsb_check_launch("isect_count")) return e;
    }
    gsb_launch_chained(GSB_CHAIN_ISECT, scan_block_sums_kernel, 1, 1024, 0, stream, nblk, block_sums, block_offsets, n_isects_out);
    return gsb_check_launch("scan_block_sums");
}

extern "C" int gsb_scan_i32_to_i64(int64_t n, const int32_t *in, int64_t *out, int64_t *total_out,
                                   cudaStream_t stream) {
    GSB_REQUIRE(n >= 0, "negative size");
    gsb_launch_chained(GSB_CHAIN_ISECT, scan_block_sums_kernel, 1, 1024, 0, stream, n, in, out, total_out);
    return gsb_check_launch("scan_i32_to_i64");
}

extern "C" int gsb_isect_emit(int64_t P, int N, int packed, const float *means2d, const int32_t *radii,
                              const float *depths, const int64_t *camera_ids, int tile_size, int tile_width,
                              int tile_height, int tile_n_bits, const int32_t *tiles_per_gauss,
                              const int64_t *block_offsets, int64_t *isect_ids, int32_t *flatten_ids,
                              cudaStream_t stream) {
    GSB_REQUIRE(P >= 0 && tile_size > 0, "bad sizes");
    GSB_REQUIRE(!packed || camera_ids != nullptr, "camera_ids required when packed");
    GSB_REQUIRE(packed || N > 0 || P == 0, "N required when not packed");
    if (P == 0) return GSB_OK;
    unsigned nblk = (unsigned)gsb_isect_num_blocks(P);
    if (packed)
        isect_emit_kernel<true><<<nblk, kThreads, 0, stream>>>(P, N, means2d, radii, depths, camera_ids, tile_size,
                                                               tile_width, tile_height, tile_n_bits, tiles_per_gauss,
                                                               block_offsets, isect_ids, flatten_ids);
    else
        isect_emit_kernel<false><<<nblk, kThreads, 0, stream>>>(P, N, means2d, radii, depths, camera_ids, tile_size,
                                                                tile_width, tile_height, tile_n_bits, tiles_per_gauss,
                                                                block_offsets, isect_ids, flatten_ids);
    return gsb_check_launch("isect_emit");
}

extern "C" int64_t gsb_sort_pairs_temp_bytes(int64_t n) {
    size_t bytes = 0;
    cub::DoubleBuffer<int64_t> k(nullptr, nullptr);
    cub::DoubleBuffer<int32_t> v(nullptr, nullptr);
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, k, v, (int64_t)n, 0, 64, (cudaStream_t)0);
    return (int64_t)bytes;
}

// Stable LSD radix sort of (int64 key, int32 value) pairs over key bits [0, end_bit).
// keys_alt / vals_alt are caller-allocated ping-pong buffers; *result_in_alt tells which one holds the result.
extern "C" int gsb_sort_pairs(int64_t n, int end_bit, int64_t *keys, int64_t *keys_alt, int32_t *vals,
                              int32_t *vals_alt, void *temp, int64_t temp_bytes, int *result_in_alt,
                              cudaStream_t stream) {
    GSB_REQUIRE(n >= 0 && end_bit >= 0 && end_bit <= 64, "bad sizes");
    *result_in_alt = 0;
    if (n == 0) return GSB_OK;
    cub::DoubleBuffer<int64_t> k(keys, keys_alt);
    cub::DoubleBuffer<int32_t> v(vals, vals_alt);
    size_t bytes = (size_t)temp_bytes;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(temp, bytes, k, v, n, 0, end_bit, stream);
    if (e != cudaSuccess) {
        gsb_set_error("radix sort: %s", cudaGetErrorString(e));
        return GSB_ERR_CUDA;
    }
    *result_in_alt = (k.Current() == keys_alt) ? 1 : 0;
    return GSB_OK;
}

extern "C" int gsb_isect_offsets(int64_t n_isects, const int64_t *isect_ids, int C, int tile_width, int tile_height,
                                 int tile_n_bits, int32_t *offsets, cudaStream_t stream) {
    GSB_REQUIRE(n_isects >= 0 && C >= 0, "bad sizes");
    int64_t n_tiles = (int64_t)tile_width * tile_height;
    int64_t total = n_tiles * C;
    if (total == 0) return GSB_OK;
    if (n_isects == 0) {
        cudaError_t e = cudaMemsetAsync(offsets, 0, sizeof(int32_t) * total, stream);
        if (e != cudaSuccess) { gsb_set_error("memset offsets: %s", cudaGetErrorString(e)); return GSB_ERR_CUDA; }
        return GSB_OK;
    }
    isect_offsets_kernel<<<gsb_cdiv(n_isects, kThreads), kThreads, 0, stream>>>(n_isects, isect_ids, (int)n_tiles,
                                                                               tile_n_bits, total, offsets);
    return gsb_check_launch("isect_offsets");
}

// ---- depth-presorted isect_tiles (see the kernels above) ------------------------------------------------------
extern "C" int gsb_isect_row_keys(int64_t P, int N, int packed, const float *depths, const int64_t *camera_ids,
                                  int64_t *keys, int32_t *rows, cudaStream_t stream) {
    GSB_REQUIRE(P >= 0, "negative size");
    GSB_REQUIRE(!packed || camera_ids != nullptr || P == 0 || keys == nullptr, "camera_ids required when packed");
    GSB_REQUIRE(packed || N > 0 || P == 0, "N required when not packed");
    GSB_REQUIRE(P < ((int64_t)1 << 31), "more than 2^31-1 rows");
    if (P == 0) return GSB_OK;
    unsigned nblk = (unsigned)gsb_isect_num_blocks(P);
    if (packed)
        row_keys_kernel<true><<<nblk, kThreads, 0, stream>>>(P, N, depths, camera_ids, keys, rows);
    else
        row_keys_kernel<false><<<nblk, kThreads, 0, stream>>>(P, N, depths, camera_ids, keys, rows);
    return gsb_check_launch("isect_row_keys");
}

extern "C" int gsb_isect_row_tiles(int64_t P, int N, int packed, const float *means2d, const int32_t *radii,
                                   const int64_t *camera_ids, int tile_size, int tile_width, int tile_height,
                                   int tile_n_bits, int32_t *rows, int32_t *tiles_per_gauss, int32_t *row_records,
                                   cudaStream_t stream) {
    return gsb_isect_row_tiles_keys(P, N, packed, means2d, radii, nullptr, camera_ids, tile_size, tile_width,
                                    tile_height, tile_n_bits, rows, tiles_per_gauss, row_records, nullptr, nullptr,
                                    stream);
}

extern "C" int gsb_isect_row_tiles_keys(int64_t P, int N, int packed, const float *means2d, const int32_t *radii,
                                        const float *depths, const int64_t *camera_ids, int tile_size, int tile_width,
                                        int tile_height, int tile_n_bits, int32_t *rows, int32_t *tiles_per_gauss,
                                        int32_t *row_records, int32_t *row_keys, void *key_hist, cudaStream_t stream) {
    GSB_REQUIRE(row_keys == nullptr || depths != nullptr, "depths required for row_keys");
    GSB_REQUIRE(key_hist == nullptr || row_keys != nullptr, "key_hist counts the digits of row_keys");
    GSB_REQUIRE(P >= 0 && tile_size > 0 && tile_width >= 0 && tile_height >= 0, "bad sizes");
    GSB_REQUIRE(P < ((int64_t)1 << 31), "more than 2^31-1 rows");
    GSB_REQUIRE(!packed || camera_ids != nullptr || P == 0 || row_records == nullptr, "camera_ids required when packed");
    GSB_REQUIRE(packed || N > 0 || P == 0, "N required when not packed");
    GSB_REQUIRE(tile_width < 65536 && tile_height < 32768, "tile grid too large");
    if (P == 0) return GSB_OK;
    // with a histogram to feed, few CTAs walk the rows (one flush of 1024 counters per CTA); else one CTA per 256 rows
    unsigned nblk = (unsigned)gsb_isect_num_blocks(P);
    if (key_hist != nullptr && nblk > 148 * 6) nblk = 148 * 6;
    auto recs = reinterpret_cast<int4 *>(row_records);
    auto keys = reinterpret_cast<uint32_t *>(row_keys);
    auto hist = reinterpret_cast<uint32_t *>(key_hist);
    if (packed)
        gsb_launch_chained(GSB_CHAIN_ISECT, row_tiles_kernel<true>, nblk, kThreads, 0, stream, P, N, means2d, radii, camera_ids, tile_size,
                           tile_width, tile_height, tile_n_bits, rows, tiles_per_gauss, recs, depths, keys, hist);
    else
        gsb_launch_chained(GSB_CHAIN_ISECT, row_tiles_kernel<false>, nblk, kThreads, 0, stream, P, N, means2d, radii, camera_ids, tile_size,
                           tile_width, tile_height, tile_n_bits, rows, tiles_per_gauss, recs, depths, keys, hist);
    return gsb_check_launch("isect_row_tiles");
}

extern "C" int gsb_isect_count_perm_records(int64_t P, const int32_t *perm, const int32_t *row_records,
                                            int32_t *sorted_records, int32_t *block_sums, int64_t *block_offsets,
                                            int64_t *n_isects_out, cudaStream_t stream) {
    GSB_REQUIRE(P >= 0, "negative size");
    int64_t nblk = gsb_isect_num_blocks(P);
    if (P > 0) {
        count_perm_records_kernel<<<(unsigned)nblk, kThreads, 0, stream>>>(
            P, perm, reinterpret_cast<const int4 *>(row_records), reinterpret_cast<int4 *>(sorted_records), block_sums);
        if (int e = gsb_check_launch("isect_count_perm_records")) return e;
    }
    gsb_launch_chained(GSB_CHAIN_ISECT, scan_block_sums_kernel, 1, 1024, 0, stream, nblk, block_sums, block_offsets, n_isects_out);
    return gsb_check_launch("scan_block_sums");
}

extern "C" int gsb_isect_count_perm(int64_t P, const int32_t *perm, const float *means2d, const int32_t *radii,
                                    int tile_size, int tile_width, int tile_height, int32_t *tiles_per_gauss,
                                    int32_t *block_sums, int64_t *block_offsets, int64_t *n_isects_out,
                                    cudaStream_t stream) {
    GSB_REQUIRE(P >= 0 && tile_size > 0 && tile_width >= 0 && tile_height >= 0, "bad sizes");
    int64_t nblk = gsb_isect_num_blocks(P);
    if (P > 0) {
        gsb_launch_chained(GSB_CHAIN_ISECT, count_perm_kernel, (unsigned)nblk, kThreads, 0, stream, P, perm, means2d, radii, tile_size,
                           tile_width, tile_height, tiles_per_gauss, block_sums);
        if (int e = gsb_check_launch("isect_count_perm")) return e;
    }
    gsb_launch_chained(GSB_CHAIN_ISECT, scan_block_sums_kernel, 1, 1024, 0, stream, nblk, block_sums, block_offsets, n_isects_out);
    return gsb_check_launch("scan_block_sums");
}

extern "C" int gsb_isect_emit_perm(int64_t P, int N, int packed, const int32_t *perm, const float *means2d,
                                   const int32_t *radii, const int64_t *camera_ids, int tile_size, int tile_width,
                                   int tile_height, int tile_n_bits, const int32_t *tiles_per_gauss,
                                   const int64_t *block_offsets, int32_t *tile_keys, int32_t *flatten_ids,
                                   cudaStream_t stream) {
    return gsb_isect_emit_perm_hist(P, N, packed, perm, means2d, radii, camera_ids, tile_size, tile_width, tile_height,
                                    tile_n_bits, tiles_per_gauss, block_offsets, tile_keys, flatten_ids,
                                    INT64_MAX, nullptr, nullptr, nullptr, 0, stream);
}

extern "C" int gsb_isect_emit_perm_cap(int64_t P, int N, int packed, const int32_t *perm, const float *means2d,
                                       const int32_t *radii, const int64_t *camera_ids, int tile_size, int tile_width,
                                       int tile_height, int tile_n_bits, const int32_t *tiles_per_gauss,
                                       const int64_t *block_offsets, int32_t *tile_keys, int32_t *flatten_ids,
                                       int64_t capacity, const int64_t *n_isects_dev, const int32_t *row_records,
                                       cudaStream_t stream) {
    return gsb_isect_emit_perm_hist(P, N, packed, perm, means2d, radii, camera_ids, tile_size, tile_width, tile_height,
                                    tile_n_bits, tiles_per_gauss, block_offsets, tile_keys, flatten_ids, capacity,
                                    n_isects_dev, row_records, nullptr, 0, stream);
}

extern "C" int gsb_isect_emit_perm_hist(int64_t P, int N, int packed, const int32_t *perm, const float *means2d,
                                        const int32_t *radii, const int64_t *camera_ids, int tile_size, int tile_width,
                                        int tile_height, int tile_n_bits, const int32_t *tiles_per_gauss,
                                        const int64_t *block_offsets, int32_t *tile_keys, int32_t *flatten_ids,
                                        int64_t capacity, const int64_t *n_isects_dev, const int32_t *row_records,
                                        void *key_hist, int key_bits, cudaStream_t stream) {
    GSB_REQUIRE(P >= 0 && tile_size > 0 && capacity >= 0, "bad sizes");
    GSB_REQUIRE(!packed || camera_ids != nullptr || P == 0, "camera_ids required when packed");
    GSB_REQUIRE(key_hist == nullptr || (key_bits > 0 && key_bits <= 32), "key_bits: 1..32 with a histogram");
    auto tk = reinterpret_cast<uint32_t *>(tile_keys);
    auto recs = reinterpret_cast<const int4 *>(row_records);
    auto hist = reinterpret_cast<uint32_t *>(key_hist);
    GSB_REQUIRE(tile_width < 65536 && tile_height < 32768, "tile grid too large");
    if (P > 0) {
        const int64_t nblk = gsb_isect_num_blocks(P);
        // with a histogram to feed, few CTAs walk the row blocks (one flush of the counters per CTA)
        unsigned grid = (unsigned)nblk;
        if (hist != nullptr && grid > 148 * 6) grid = 148 * 6;
        if (packed)
            gsb_launch_chained(GSB_CHAIN_ISECT, emit_perm_kernel<true>, grid, kThreads, 0, stream, P, N, perm, means2d, radii, camera_ids,
                               tile_size, tile_width, tile_height, tile_n_bits, tiles_per_gauss, block_offsets, tk,
                               flatten_ids, capacity, recs, nblk, hist, key_bits);
        else
            gsb_launch_chained(GSB_CHAIN_ISECT, emit_perm_kernel<false>, grid, kThreads, 0, stream, P, N, perm, means2d, radii, camera_ids,
                               tile_size, tile_width, tile_height, tile_n_bits, tiles_per_gauss, block_offsets, tk,
                               flatten_ids, capacity, recs, nblk, hist, key_bits);
        if (int e = gsb_check_launch("isect_emit_perm")) return e;
    }
    if (n_isects_dev != nullptr && capacity > 0) {
        gsb_launch_chained(GSB_CHAIN_ISECT, pad_tail_kernel, 148, kThreads, 0, stream, n_isects_dev, capacity, tk, flatten_ids);
        return gsb_check_launch("isect_pad_tail");
    }
    return GSB_OK;
}

extern "C" int64_t gsb_sort_pairs_u32_temp_bytes(int64_t n) {
    size_t bytes = 0;
    cub::DoubleBuffer<uint32_t> k(nullptr, nullptr);
    cub::DoubleBuffer<int32_t> v(nullptr, nullptr);
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, k, v, (int64_t)n, 0, 32, (cudaStream_t)0);
    return (int64_t)bytes;
}

extern "C" int gsb_sort_pairs_u32(int64_t n, int end_bit, int32_t *keys, int32_t *keys_alt, int32_t *vals,
                                  int32_t *vals_alt, void *temp, int64_t temp_bytes, int *result_in_alt,
                                  cudaStream_t stream) {
    GSB_REQUIRE(n >= 0 && end_bit >= 0 && end_bit <= 32, "bad sizes");
    *result_in_alt = 0;
    if (n == 0 || end_bit == 0) return GSB_OK;
    cub::DoubleBuffer<uint32_t> k(reinterpret_cast<uint32_t *>(keys), reinterpret_cast<uint32_t *>(keys_alt));
    cub::DoubleBuffer<int32_t> v(vals, vals_alt);
    size_t bytes = (size_t)temp_bytes;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(temp, bytes, k, v, n, 0, end_bit, stream);
    if (e != cudaSuccess) {
        gsb_set_error("radix sort (u32): %s", cudaGetErrorString(e));
        return GSB_ERR_CUDA;
    }
    *result_in_alt = (reinterpret_cast<int32_t *>(k.Current()) == keys_alt) ? 1 : 0;
    return GSB_OK;
}

extern "C" int gsb_isect_finalize(int64_t n_isects, const int32_t *tile_keys, const int32_t *flatten_ids,
                                  const float *depths, int C, int tile_width, int tile_height, int tile_n_bits,
                                  int64_t *isect_ids, int32_t *offsets, cudaStream_t stream) {
    return gsb_isect_finalize_cap(n_isects, nullptr, tile_keys, flatten_ids, depths, C, tile_width, tile_height,
                                  tile_n_bits, isect_ids, offsets, stream);
}

extern "C" int gsb_isect_finalize_cap(int64_t n_isects, const int64_t *n_isects_dev, const int32_t *tile_keys,
                                      const int32_t *flatten_ids, const float *depths, int C, int tile_width,
                                      int tile_height, int tile_n_bits, int64_t *isect_ids, int32_t *offsets,
                                      cudaStream_t stream) {
    GSB_REQUIRE(n_isects >= 0 && C >= 0, "bad sizes");
    GSB_REQUIRE(n_isects_dev == nullptr || n_isects > 0, "a speculative capacity must be positive");
    int64_t n_tiles = (int64_t)tile_width * tile_height;
    int64_t total = n_tiles * C;
    if (total == 0) return GSB_OK;
    if (n_isects == 0) {
        if (offsets == nullptr) return GSB_OK;
        cudaError_t e = cudaMemsetAsync(offsets, 0, sizeof(int32_t) * total, stream);
        if (e != cudaSuccess) { gsb_set_error("memset offsets: %s", cudaGetErrorString(e)); return GSB_ERR_CUDA; }
        return GSB_OK;
    }
    const bool aligned = (((uintptr_t)tile_keys | (uintptr_t)flatten_ids | (uintptr_t)isect_ids) & 15u) == 0;
    if (aligned && gsb_option(GSB_OPT_FINALIZE_VEC) != 0) {
        finalize4_kernel<<<gsb_cdiv(gsb_cdiv(n_isects, 4), kThreads), kThreads, 0, stream>>>(
            n_isects, n_isects_dev, reinterpret_cast<const uint32_t *>(tile_keys), flatten_ids, depths, (int)n_tiles,
            tile_n_bits, total, isect_ids, offsets);
        return gsb_check_launch("isect_finalize4");
    }
    gsb_launch_chained(GSB_CHAIN_ISECT, finalize_kernel, gsb_cdiv(n_isects, kThreads), kThreads, 0, stream, n_isects, n_isects_dev,
                       reinterpret_cast<const uint32_t *>(tile_keys), flatten_ids, depths, (int)n_tiles, tile_n_bits, total,
                       isect_ids, offsets);
    return gsb_check_launch("isect_finalize");
}
