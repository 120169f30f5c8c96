// Tile binning of depth-ordered rows: the second half of isect_tiles + isect_offset_encode without a global sort.
//
// Replaces (behaviour, not code): the key emit + radix sort + offset encode of gsplat/triton_impl/isect_tiles.py:14-133,
// isect_offset.py:9-63 and gsplat/cuda/csrc/IntersectTile.cu:24-344 — same isect_ids / flatten_ids / offsets, bit for bit.
//
// Input: the rows (projected Gaussians) already sorted by depth (radix_sort.cu over P rows), their 16-byte tile
// rectangles gathered into that order, and the exclusive scan of their tile counts per 256 rows. The intersections in
// "emission order" (row after row in depth order, a row's tiles in row-major order) then only need a STABLE partition
// by (camera | tile): what the reference's 64-bit sort produces. Instead of two more radix passes over all
// intersections (emit 8 B + 2 x (read 8 B + write 8 B) + finalize) this file does a counting sort with one bin per tile:
//
//   bin_count_kernel    chunk c = intersections [c*8192, (c+1)*8192) of the emission order: per-bin counts -> M[c][bin]
//   bin_scan_*          exclusive scan of M down the chunks (two levels) and of the bin totals: `offsets` falls out
//   bin_scatter_kernel  the chunk again: every intersection goes to offsets[bin] + (intersections of that bin in
//                       earlier chunks) + (its rank among the chunk's intersections of that bin, in emission order):
//                       written once, at its final place, as the 64-bit key + the row id.
//
// The rank inside the chunk: shared-memory atomics drop the chunk's slot numbers into per-bin lists (any order), then
// every intersection counts the smaller slot numbers of its list: with ~1 intersection per (chunk, bin) that is a
// handful of instructions. A chunk that piles more than kFastMax intersections on one bin (a far-away object inside
// one tile) takes a stable 2-pass radix sort inside shared memory instead.
//
// No kernel waits on another (no look-back chain), nothing is written twice. The price is the dense chunks x bins count
// matrix (so: C * tiles <= 8192 bins) and, decisively, the final stores: a chunk holds ~1.6 intersections per bin, so
// every 8-byte key and 4-byte id is its own 32-byte sector write (12.7 M sectors per launch at cfg2, 8 x what a radix
// pass writes). Measured at cfg2 (6.35 M intersections, B200): count 46 us + scans 52 us + scatter 178 us under ncu,
// 0.245 ms by CUDA events against 0.238 ms for emit + two radix passes + finalize; with the stores redirected to
// consecutive addresses (experiment) 0.188 ms. Correct and selectable (GSPLAT_B200_ISECT=bin), NOT the default.
#include "common.cuh"

#include <atomic>

#include <cstdint>

namespace {

constexpr int kT = 8192;        // intersections per chunk
constexpr int kThreads = 512;   // CTA of the chunk kernels
constexpr int kWarps = kThreads / 32;
constexpr int kItems = kT / kThreads;
constexpr int kRowBlock = 256;  // rows per block_offsets entry (isect.cu)
constexpr int kMaxBins = kT;    // per-bin tables of the scatter kernel are sized like a chunk
constexpr int kGroups = 32;     // second level of the scan down the chunks
constexpr int kFastMax = 128;   // longest per-bin list the counting rank handles; beyond: radix passes

__device__ __forceinline__ int64_t clamp_count(int64_t capacity, const int64_t *n_dev) {
    if (n_dev == nullptr) return capacity;
    const int64_t n = *n_dev;
    return n < capacity ? (n < 0 ? 0 : n) : capacity;
}

struct ChunkSmem {
    int warp[kWarps];
    uint32_t wmax[kWarps];
    int b0;         // first 256-row block overlapping the chunk
    int first_off;  // tiles of the chunk's first row that belong to the previous chunk
    int maxcnt;
};

// Slot -> owning row for the chunk's slots [a, b_end) of the emission order. Every row starting inside the chunk drops
// (slot << 16 | row) at its first slot (rows are numbered from row 256*b0; the row straddling `a` owns slot 0); an
// inclusive max-scan then carries each marker over the row's slots. Cost per slot: a few shared-memory words, no
// search. On return mark[j] = (first slot of the owner << 16 | owner) for every slot.
__device__ __forceinline__ void build_marks(int64_t a, int64_t b_end, int64_t P, const int4 *__restrict__ srec,
                                            const int64_t *__restrict__ block_offsets, int64_t nblk,
                                            uint32_t *__restrict__ mark, ChunkSmem &s) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint4 *mark4 = reinterpret_cast<uint4 *>(mark);
    for (int i = tid; i < kT / 4; i += kThreads) mark4[i] = make_uint4(0u, 0u, 0u, 0u);
    if (tid == 0) {
        int64_t lo = 0, hi = nblk;  // last block with block_offsets[lo] <= a (block_offsets[0] = 0 <= a)
        while (hi - lo > 1) {
            const int64_t mid = (lo + hi) >> 1;
            if (block_offsets[mid] <= a) lo = mid; else hi = mid;
        }
        s.b0 = (int)lo;
        s.first_off = 0;
    }
    __syncthreads();
    const int64_t b0 = s.b0;
    const int *counts = reinterpret_cast<const int *>(srec) + 3;  // .w of every record
    for (int it = 0;; ++it) {
        const int64_t blk = b0 + 2 * (int64_t)it;
        if (blk >= nblk) break;
        const int64_t base = block_offsets[blk];
        if (base >= b_end) break;
        const int64_t k = blk * kRowBlock + tid;
        const int cnt = k < P ? counts[4 * k] : 0;
        int inc = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s.warp[warp] = inc;
        __syncthreads();
        int wbase = 0;
#pragma unroll
        for (int wi = 0; wi < kWarps; ++wi) wbase += (wi < warp) ? s.warp[wi] : 0;
        const int64_t start = base + wbase + inc - cnt;
        if (cnt > 0 && start < b_end && start + cnt > a) {
            const int krel = (int)(blk - b0) * kRowBlock + tid;
            if (krel > 0xffff) __trap();  // rows without tiles inside the run of emitting rows: contract broken
            if (start < a) {
                s.first_off = (int)(a - start);
                mark[0] = (uint32_t)krel;
            } else {
                const int pos = (int)(start - a);
                mark[pos] = ((uint32_t)pos << 16) | (uint32_t)krel;
            }
        }
        __syncthreads();  // s.warp is rewritten by the next round; also publishes the markers
    }
    // inclusive max-scan, 16 consecutive slots per thread
    uint4 q[kItems / 4];
#pragma unroll
    for (int u = 0; u < kItems / 4; ++u) q[u] = mark4[tid * (kItems / 4) + u];
    uint32_t run = 0u;
#pragma unroll
    for (int u = 0; u < kItems / 4; ++u) {
        run = max(run, q[u].x); q[u].x = run;
        run = max(run, q[u].y); q[u].y = run;
        run = max(run, q[u].z); q[u].z = run;
        run = max(run, q[u].w); q[u].w = run;
    }
    uint32_t inc = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc = max(inc, t);
    }
    if (lane == 31) s.wmax[warp] = inc;
    uint32_t carry = __shfl_up_sync(0xffffffffu, inc, 1);
    if (lane == 0) carry = 0u;
    __syncthreads();
#pragma unroll
    for (int wi = 0; wi < kWarps; ++wi) carry = max(carry, (wi < warp) ? s.wmax[wi] : 0u);
#pragma unroll
    for (int u = 0; u < kItems / 4; ++u) {
        q[u].x = max(q[u].x, carry); q[u].y = max(q[u].y, carry);
        q[u].z = max(q[u].z, carry); q[u].w = max(q[u].w, carry);
        mark4[tid * (kItems / 4) + u] = q[u];
    }
    __syncthreads();
}

// tile (inside its camera), camera word (camera << tile_n_bits) and owning row of slot j (relative to the chunk) from
// its scanned marker
__device__ __forceinline__ int tile_of_slot(uint32_t m, int j, int first_off, const int4 *__restrict__ rows0, int tw,
                                            uint32_t &cam_word, int &krel) {
    krel = (int)(m & 0xffffu);
    const int p = (int)(m >> 16);
    const int t = j - p + (p == 0 ? first_off : 0);
    const int4 rec = __ldg(rows0 + krel);
    const int w = rec.y;
    int dy = (int)__fdividef((float)t + 0.5f, (float)w);  // t / w, fixed up below: no integer division
    int dx = t - dy * w;
    if (dx < 0) { dy -= 1; dx += w; }
    if (dx >= w) { dy += 1; dx -= w; }
    cam_word = (uint32_t)rec.z;
    return ((rec.x >> 16) + dy) * tw + (rec.x & 0xffff) + dx;
}

// (camera, tile) bin of slot j
__device__ __forceinline__ int bin_of_slot(uint32_t m, int j, int first_off, const int4 *__restrict__ rows0, int tw,
                                           int n_tiles, int tile_n_bits, int &krel) {
    uint32_t cam_word;
    const int tile = tile_of_slot(m, j, first_off, rows0, tw, cam_word, krel);
    return (int)(cam_word >> tile_n_bits) * n_tiles + tile;
}

// ---------------------------------------------------------------- emission for the radix path
// The (camera | tile) words + row ids of the chunk's slots in emission order, stored coalesced, and (optionally) the
// digit histograms of the words for the tile sort that follows. Same output as emit_perm_kernel (isect.cu), whose
// per-slot binary search made it issue-bound (ncu: 71 % issue-active, 37 M warp instructions at cfg2).
__global__ void __launch_bounds__(kThreads)
emit_chunks_kernel(int64_t capacity, const int64_t *__restrict__ n_dev, int64_t P, const int4 *__restrict__ srec,
                   const int64_t *__restrict__ block_offsets, int64_t nblk, const int32_t *__restrict__ perm, int tw,
                   uint32_t *__restrict__ tile_keys, int32_t *__restrict__ flatten_ids,
                   uint32_t *__restrict__ key_hist, int key_bits) {
    __shared__ __align__(16) uint32_t mark[kT];
    __shared__ uint32_t s_hist[4 * 256];
    __shared__ ChunkSmem s;
    const int tid = threadIdx.x;
    const int64_t n = clamp_count(capacity, n_dev);
    const int64_t a = (int64_t)blockIdx.x * kT;
    if (a >= n) return;
    const int64_t b_end = a + kT < n ? a + kT : n;
    const int count = (int)(b_end - a);
    const int passes = (key_bits + 7) / 8;
    if (key_hist != nullptr)
        for (int i = tid; i < passes * 256; i += kThreads) s_hist[i] = 0u;
    build_marks(a, b_end, P, srec, block_offsets, nblk, mark, s);
    const int64_t row0 = (int64_t)s.b0 * kRowBlock;
    const int4 *rows0 = srec + row0;
    const int first_off = s.first_off;
#pragma unroll 4
    for (int i = 0; i < kItems; ++i) {
        const int j = i * kThreads + tid;
        const bool live = j < count;
        uint32_t key = 0u;
        if (live) {
            int krel;
            uint32_t cam_word;
            const int tile = tile_of_slot(mark[j], j, first_off, rows0, tw, cam_word, krel);
            key = cam_word | (uint32_t)tile;
            tile_keys[a + j] = key;
            flatten_ids[a + j] = perm[row0 + krel];
        }
        if (key_hist != nullptr) gsb_hist_add(s_hist, key, live, passes, key_bits);
    }
    if (key_hist != nullptr) {
        __syncthreads();
        for (int i = tid; i < passes * 256; i += kThreads) {
            const uint32_t c = s_hist[i];
            if (c) atomicAdd(&key_hist[i], c);
        }
    }
}

// ---------------------------------------------------------------- counts per (chunk, bin)
__global__ void __launch_bounds__(kThreads)
bin_count_kernel(int64_t capacity, const int64_t *__restrict__ n_dev, int64_t P, const int4 *__restrict__ srec,
                 const int64_t *__restrict__ block_offsets, int64_t nblk, int tw, int n_tiles, int tile_n_bits,
                 int bins, uint32_t *__restrict__ M) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t *mark = reinterpret_cast<uint32_t *>(smem_raw);
    uint32_t *hist = mark + kT;
    __shared__ ChunkSmem s;
    const int64_t n = clamp_count(capacity, n_dev);
    const int64_t a = (int64_t)blockIdx.x * kT;
    for (int i = threadIdx.x; i < bins; i += kThreads) hist[i] = 0u;
    if (a < n) {
        const int64_t b_end = a + kT < n ? a + kT : n;
        const int count = (int)(b_end - a);
        build_marks(a, b_end, P, srec, block_offsets, nblk, mark, s);
        const int4 *rows0 = srec + (int64_t)s.b0 * kRowBlock;
        const int first_off = s.first_off;
#pragma unroll 4
        for (int i = 0; i < kItems; ++i) {
            const int j = i * kThreads + threadIdx.x;
            if (j < count) {
                int krel;
                atomicAdd(&hist[bin_of_slot(mark[j], j, first_off, rows0, tw, n_tiles, tile_n_bits, krel)], 1u);
            }
        }
    }
    __syncthreads();
    uint32_t *row = M + (int64_t)blockIdx.x * bins;
    for (int i = threadIdx.x; i < bins; i += kThreads) row[i] = hist[i];
}

// exclusive scan of M down the chunks of one group (in place) + the group's total per bin
__global__ void __launch_bounds__(256)
bin_scan_chunks_kernel(int chunks, int per_group, int bins, uint32_t *__restrict__ M, uint32_t *__restrict__ G) {
    const int bin = blockIdx.x * 256 + threadIdx.x;
    if (bin >= bins) return;
    const int g = blockIdx.y;
    const int c0 = g * per_group, c1 = min(chunks, c0 + per_group);
    uint32_t run = 0u;
    int c = c0;
    for (; c + 4 <= c1; c += 4) {
        uint32_t v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) v[u] = M[(int64_t)(c + u) * bins + bin];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            M[(int64_t)(c + u) * bins + bin] = run;
            run += v[u];
        }
    }
    for (; c < c1; ++c) {
        const uint32_t v = M[(int64_t)c * bins + bin];
        M[(int64_t)c * bins + bin] = run;
        run += v;
    }
    G[(int64_t)g * bins + bin] = run;
}

// exclusive scan over the groups (in place) + the bin's total
__global__ void __launch_bounds__(256)
bin_scan_groups_kernel(int bins, uint32_t *__restrict__ G, uint32_t *__restrict__ totals) {
    const int bin = blockIdx.x * 256 + threadIdx.x;
    if (bin >= bins) return;
    uint32_t v[kGroups];
#pragma unroll
    for (int g = 0; g < kGroups; ++g) v[g] = G[(int64_t)g * bins + bin];
    uint32_t run = 0u;
#pragma unroll
    for (int g = 0; g < kGroups; ++g) {
        G[(int64_t)g * bins + bin] = run;
        run += v[g];
    }
    totals[bin] = run;
}

// exclusive scan of the bin totals = isect_offsets (one CTA; bins <= 8192)
__global__ void __launch_bounds__(1024)
bin_offsets_kernel(int bins, const uint32_t *__restrict__ totals, int32_t *__restrict__ offsets) {
    __shared__ uint32_t s_warp[32];
    constexpr int kPer = kMaxBins / 1024;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t v[kPer], sum = 0u;
#pragma unroll
    for (int u = 0; u < kPer; ++u) {
        const int i = tid * kPer + u;
        v[u] = i < bins ? totals[i] : 0u;
        sum += v[u];
    }
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t run = inc - sum;
    for (int wi = 0; wi < warp; ++wi) run += s_warp[wi];
#pragma unroll
    for (int u = 0; u < kPer; ++u) {
        const int i = tid * kPer + u;
        if (i < bins) offsets[i] = (int32_t)run;
        run += v[u];
    }
}

// ---------------------------------------------------------------- chunk sort + scatter to the final place
// One stable radix pass over the chunk inside shared memory (structure of radix_sort.cu's tile: warp w ranks its
// 512 consecutive items with one ballot per digit bit, per-warp digit counters, digit threads turn the counters into
// first slots).
__device__ __forceinline__ void smem_radix_pass(const uint32_t *__restrict__ src, uint32_t *__restrict__ dst, int count,
                                                int shift, int bits, uint32_t (*s_cnt)[256], uint32_t *s_scan) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nd = 1 << bits, mask = nd - 1;
    for (int i = tid; i < kWarps * 256; i += kThreads) (&s_cnt[0][0])[i] = 0u;
    __syncthreads();
    uint32_t rank[kItems];  // the only per-item state kept across the barriers (keys are re-read from src)
    const int chunk = warp * 32 * kItems;
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const int pos = chunk + i * 32 + lane;
        const bool live = pos < count;
        const int digit = live ? (int)(src[pos] >> shift) & mask : 0;
        uint32_t p = __ballot_sync(0xffffffffu, live);
        if (!live) p = ~p;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            if (b < bits) {  // warp-uniform
                const bool one = (digit >> b) & 1;
                const uint32_t with_one = __ballot_sync(0xffffffffu, one);
                p &= one ? with_one : ~with_one;
            }
        }
        // one warp's shared-memory atomics execute in program order: round i's counter already holds rounds < i
        const int leader = __ffs(p) - 1;
        uint32_t r = 0u;
        if (live && lane == leader) r = atomicAdd(&s_cnt[warp][digit], (uint32_t)__popc(p));
        rank[i] = __shfl_sync(0xffffffffu, r, leader) + __popc(p & lt);
    }
    __syncthreads();
    // digit threads: total of digit d, exclusive scan over the digits, first slot of every (warp, digit)
    uint32_t total = 0u, inc = 0u;
    if (tid < 256) {
        if (tid < nd) {
#pragma unroll
            for (int wi = 0; wi < kWarps; ++wi) total += s_cnt[wi][tid];
        }
        inc = total;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += up;
        }
        if (lane == 31) s_scan[warp] = inc;
    }
    __syncthreads();
    if (tid < nd) {
        uint32_t run = inc - total;
        for (int wi = 0; wi < warp; ++wi) run += s_scan[wi];
#pragma unroll
        for (int wi = 0; wi < kWarps; ++wi) {
            const uint32_t c = s_cnt[wi][tid];
            s_cnt[wi][tid] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const int pos = chunk + i * 32 + lane;
        if (pos < count) {
            const uint32_t key = src[pos];
            dst[s_cnt[warp][(int)(key >> shift) & mask] + rank[i]] = key;
        }
    }
    __syncthreads();
}

// shared memory of the scatter kernel: three chunk-sized tables
//   mark  [kT]            slot markers, later the per-bin slot lists (uint16)     | radix fallback: buf0
//   cnt   [kMaxBins + 16] per-bin counts -> list starts -> list ends               | radix fallback: buf1
//   base  [kMaxBins]      final position of the bin's first intersection of this chunk | radix fallback: warp counters
constexpr size_t kScatterSmem = sizeof(uint32_t) * (kT + (kMaxBins + 16) + kMaxBins);
static_assert(kMaxBins + 16 >= kT && kMaxBins >= kWarps * 256, "fallback buffers alias the tables");

__global__ void __launch_bounds__(kThreads, 2)
bin_scatter_kernel(int64_t capacity, const int64_t *__restrict__ n_dev, int64_t P, const int4 *__restrict__ srec,
                   const int64_t *__restrict__ block_offsets, int64_t nblk, const int32_t *__restrict__ perm,
                   const uint32_t *__restrict__ sorted_keys, int C, int tw, int n_tiles, int tile_n_bits, int bins,
                   int bin_bits, int per_group, const uint32_t *__restrict__ M, const uint32_t *__restrict__ G,
                   const int32_t *__restrict__ offsets, int64_t *__restrict__ isect_ids,
                   int32_t *__restrict__ flatten_ids, int force_radix) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t *mark = reinterpret_cast<uint32_t *>(smem_raw);
    uint32_t *cnt = mark + kT;
    int32_t *base = reinterpret_cast<int32_t *>(cnt + kMaxBins + 16);
    __shared__ ChunkSmem s;
    __shared__ uint32_t s_scan[kWarps];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t n = clamp_count(capacity, n_dev);
    const int64_t a = (int64_t)blockIdx.x * kT;
    if (a >= n) return;
    const int64_t b_end = a + kT < n ? a + kT : n;
    const int count = (int)(b_end - a);
    const int chunk = blockIdx.x, g = chunk / per_group;
    for (int i = tid; i < kMaxBins + 16; i += kThreads) cnt[i] = 0u;
    if (tid == 0) s.maxcnt = 0;
    build_marks(a, b_end, P, srec, block_offsets, nblk, mark, s);
    const int64_t row0 = (int64_t)s.b0 * kRowBlock;
    const int4 *rows0 = srec + row0;
    const int first_off = s.first_off;

    // ---- the chunk's intersections: (bin << 16 | row) in registers, counted per bin
    uint32_t item[kItems];
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const int j = i * kThreads + tid;
        item[i] = 0xffffffffu;
        if (j < count) {
            int krel;
            const int bin = bin_of_slot(mark[j], j, first_off, rows0, tw, n_tiles, tile_n_bits, krel);
            item[i] = ((uint32_t)bin << 16) | (uint32_t)krel;
            atomicAdd(&cnt[bin], 1u);
        }
    }
    // final position of every bin's first intersection of this chunk (dense, coalesced; in flight during the scan)
    for (int b = tid; b < bins; b += kThreads)
        base[b] = offsets[b] + (int32_t)(G[(int64_t)g * bins + b] + M[(int64_t)chunk * bins + b]);
    __syncthreads();

    // ---- exclusive scan of the counts (16 consecutive bins per thread) -> list starts; longest list
    {
        uint4 *cnt4 = reinterpret_cast<uint4 *>(cnt);
        uint4 q[kItems / 4];
        uint32_t sum = 0u, longest = 0u;
#pragma unroll
        for (int u = 0; u < kItems / 4; ++u) {
            q[u] = cnt4[tid * (kItems / 4) + u];
            longest = max(max(longest, q[u].x), max(q[u].y, max(q[u].z, q[u].w)));
            sum += q[u].x + q[u].y + q[u].z + q[u].w;
        }
        uint32_t inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_scan[warp] = inc;
        longest = (uint32_t)warp_max_i((int)longest);
        if (lane == 0 && longest > 0u) atomicMax(&s.maxcnt, (int)longest);
        __syncthreads();
        uint32_t run = inc - sum;
#pragma unroll
        for (int wi = 0; wi < kWarps; ++wi) run += (wi < warp) ? s_scan[wi] : 0u;
#pragma unroll
        for (int u = 0; u < kItems / 4; ++u) {
            uint4 o4;
            o4.x = run; run += q[u].x;
            o4.y = run; run += q[u].y;
            o4.z = run; run += q[u].z;
            o4.w = run; run += q[u].w;
            cnt4[tid * (kItems / 4) + u] = o4;
        }
        __syncthreads();
    }

    if (s.maxcnt <= kFastMax && !force_radix) {
        // ---- drop the slot numbers into the per-bin lists (any order inside a list)
        uint16_t *list = reinterpret_cast<uint16_t *>(mark);
#pragma unroll
        for (int i = 0; i < kItems; ++i) {
            const int j = i * kThreads + tid;
            if (j < count) list[atomicAdd(&cnt[item[i] >> 16], 1u)] = (uint16_t)j;
        }
        __syncthreads();
        // ---- rank inside the list = smaller slot numbers in it (emission order); write at the final place
#pragma unroll
        for (int i = 0; i < kItems; ++i) {
            const int j = i * kThreads + tid;
            if (j < count) {
                const int bin = (int)(item[i] >> 16);
                const int lo = bin ? (int)cnt[bin - 1] : 0, hi = (int)cnt[bin];  // list ends = next list's start
                int rank = 0;
                for (int qi = lo; qi < hi; ++qi) rank += (int)list[qi] < j;
                const int64_t pos = (int64_t)base[bin] + rank;
                if (pos < capacity) {
                    const int64_t k = row0 + (item[i] & 0xffffu);
                    uint32_t hi_word = (uint32_t)bin;
                    if (C > 1) {
                        const int cam = bin / n_tiles;
                        hi_word = ((uint32_t)cam << tile_n_bits) | (uint32_t)(bin - cam * n_tiles);
                    }
                    isect_ids[pos] = ((int64_t)hi_word << 32) | (int64_t)sorted_keys[k];
                    flatten_ids[pos] = perm[k];
                }
            }
        }
        return;
    }

    // ---- a bin with a long list: stable 2-pass radix sort of the chunk by bin inside shared memory
    __syncthreads();  // every thread has read s.maxcnt and the tables are dead
    uint32_t *buf0 = mark, *buf1 = cnt;
    uint32_t(*s_cnt)[256] = reinterpret_cast<uint32_t(*)[256]>(base);
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const int j = i * kThreads + tid;
        if (j < count) buf0[j] = item[i];
    }
    __syncthreads();
    const int lo_bits = (bin_bits + 1) >> 1, hi_bits = bin_bits - lo_bits;
    uint32_t *sorted = buf0, *spare = buf1;
    if (lo_bits > 0) {
        smem_radix_pass(buf0, buf1, count, 16, lo_bits, s_cnt, s_scan);
        sorted = buf1;
        spare = buf0;
    }
    if (hi_bits > 0) {
        smem_radix_pass(sorted, spare, count, 16 + lo_bits, hi_bits, s_cnt, s_scan);
        uint32_t *t = sorted; sorted = spare; spare = t;
    }
    // first slot of every bin present in the chunk -> (final position - slot) of the bin
    int32_t *delta = reinterpret_cast<int32_t *>(spare);
    for (int i = tid; i < count; i += kThreads) {
        const int bin = (int)(sorted[i] >> 16);
        if (i == 0 || (int)(sorted[i - 1] >> 16) != bin)
            delta[bin] = offsets[bin] + (int32_t)(G[(int64_t)g * bins + bin] + M[(int64_t)chunk * bins + bin]) - i;
    }
    __syncthreads();
    for (int i = tid; i < count; i += kThreads) {
        const uint32_t it = sorted[i];
        const int bin = (int)(it >> 16);
        const int64_t k = row0 + (it & 0xffffu);
        const int64_t pos = (int64_t)delta[bin] + i;
        if (pos < capacity) {
            const int cam = bin / n_tiles;
            const uint32_t hi_word = ((uint32_t)cam << tile_n_bits) | (uint32_t)(bin - cam * n_tiles);
            isect_ids[pos] = ((int64_t)hi_word << 32) | (int64_t)sorted_keys[k];
            flatten_ids[pos] = perm[k];
        }
    }
}

}  // namespace

extern "C" int gsb_isect_emit_chunks(int64_t P, int64_t capacity, const int64_t *n_isects_dev,
                                     const int32_t *sorted_records, const int32_t *perm, const int64_t *block_offsets,
                                     int tile_width, int32_t *tile_keys, int32_t *flatten_ids, void *key_hist,
                                     int key_bits, cudaStream_t stream) {
    GSB_REQUIRE(P >= 0 && capacity >= 0, "bad sizes");
    GSB_REQUIRE(P < ((int64_t)1 << 31) && capacity < ((int64_t)1 << 31), "more than 2^31-1 rows or intersections");
    GSB_REQUIRE(key_hist == nullptr || (key_bits > 0 && key_bits <= 32), "key_bits: 1..32 with a histogram");
    if (P == 0 || capacity == 0) return GSB_OK;
    const int chunks = (int)((capacity + kT - 1) / kT);
    const int64_t nblk = (P + kRowBlock - 1) / kRowBlock;
    emit_chunks_kernel<<<chunks, kThreads, 0, stream>>>(
        capacity, n_isects_dev, P, reinterpret_cast<const int4 *>(sorted_records), block_offsets, nblk, perm,
        tile_width, reinterpret_cast<uint32_t *>(tile_keys), flatten_ids, reinterpret_cast<uint32_t *>(key_hist),
        key_bits);
    return gsb_check_launch("isect_emit_chunks");
}

extern "C" int gsb_isect_bin_max_bins(void) { return kMaxBins; }

// M [chunks][bins] | G [kGroups][bins] | totals [bins], uint32
extern "C" int64_t gsb_isect_bin_temp_bytes(int64_t capacity, int bins) {
    const int64_t chunks = (capacity + kT - 1) / kT;
    return 4 * ((chunks > 0 ? chunks : 1) + kGroups + 1) * (int64_t)bins;
}

extern "C" int gsb_isect_bin(int64_t P, int64_t capacity, const int64_t *n_isects_dev, const int32_t *sorted_records,
                             const int32_t *perm, const int32_t *sorted_keys, const int64_t *block_offsets, int C,
                             int tile_width, int tile_height, int tile_n_bits, void *temp, int64_t temp_bytes,
                             int64_t *isect_ids, int32_t *flatten_ids, int32_t *offsets, cudaStream_t stream) {
    GSB_REQUIRE(P >= 0 && capacity >= 0 && C >= 0, "bad sizes");
    GSB_REQUIRE(P < ((int64_t)1 << 31) && capacity < ((int64_t)1 << 31), "more than 2^31-1 rows or intersections");
    const int64_t n_tiles = (int64_t)tile_width * tile_height;
    const int64_t bins64 = n_tiles * C;
    GSB_REQUIRE(bins64 <= kMaxBins, "too many (camera, tile) bins for the binning path: use the radix path");
    if (bins64 == 0) return GSB_OK;
    const int bins = (int)bins64;
    if (capacity == 0 || P == 0) {
        cudaError_t e = cudaMemsetAsync(offsets, 0, sizeof(int32_t) * bins, stream);
        if (e != cudaSuccess) { gsb_set_error("memset offsets: %s", cudaGetErrorString(e)); return GSB_ERR_CUDA; }
        return GSB_OK;
    }
    GSB_REQUIRE(temp != nullptr && temp_bytes >= gsb_isect_bin_temp_bytes(capacity, bins), "temp buffer too small");
    const int chunks = (int)((capacity + kT - 1) / kT);
    const int per_group = (chunks + kGroups - 1) / kGroups;
    const int64_t nblk = (P + kRowBlock - 1) / kRowBlock;
    int bin_bits = 0;
    while ((1 << bin_bits) < bins) ++bin_bits;
    uint32_t *M = reinterpret_cast<uint32_t *>(temp);
    uint32_t *G = M + (int64_t)chunks * bins;
    uint32_t *totals = G + (int64_t)kGroups * bins;
    auto srec = reinterpret_cast<const int4 *>(sorted_records);

    static std::atomic<unsigned long long> configured{0ull};  // per device, like radix_sort.cu
    int device = 0;
    cudaGetDevice(&device);
    if (!((configured.load(std::memory_order_acquire) >> (device & 63)) & 1ull)) {
        cudaError_t e = cudaFuncSetAttribute(bin_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)kScatterSmem);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(bin_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)(sizeof(uint32_t) * (kT + kMaxBins)));
        if (e != cudaSuccess) { gsb_set_error("isect_bin: shared memory: %s", cudaGetErrorString(e)); return GSB_ERR_CUDA; }
        configured.fetch_or(1ull << (device & 63), std::memory_order_release);
    }
    bin_count_kernel<<<chunks, kThreads, sizeof(uint32_t) * (kT + bins), stream>>>(
        capacity, n_isects_dev, P, srec, block_offsets, nblk, tile_width, (int)n_tiles, tile_n_bits, bins, M);
    if (int rc = gsb_check_launch("isect_bin_count")) return rc;
    const unsigned bgrid = gsb_cdiv(bins, 256);
    bin_scan_chunks_kernel<<<dim3(bgrid, kGroups), 256, 0, stream>>>(chunks, per_group, bins, M, G);
    if (int rc = gsb_check_launch("isect_bin_scan_chunks")) return rc;
    bin_scan_groups_kernel<<<bgrid, 256, 0, stream>>>(bins, G, totals);
    if (int rc = gsb_check_launch("isect_bin_scan_groups")) return rc;
    bin_offsets_kernel<<<1, 1024, 0, stream>>>(bins, totals, offsets);
    if (int rc = gsb_check_launch("isect_bin_offsets")) return rc;
    bin_scatter_kernel<<<chunks, kThreads, kScatterSmem, stream>>>(
        capacity, n_isects_dev, P, srec, block_offsets, nblk, perm, reinterpret_cast<const uint32_t *>(sorted_keys), C,
        tile_width, (int)n_tiles, tile_n_bits, bins, bin_bits, per_group, M, G, offsets, isect_ids, flatten_ids,
        gsb_option(GSB_OPT_ISECT_BIN_RADIX));
    return gsb_check_launch("isect_bin_scatter");
}
