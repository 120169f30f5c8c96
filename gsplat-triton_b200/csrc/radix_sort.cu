// Stable LSD radix sort of (key, int32 value) pairs, 8 bits per pass, one kernel per pass ("onesweep":
// chained scan with decoupled look-back), hand-written for sm_100a.
//
// Replaces (behaviour, not code): gsplat/triton_impl/radix_sort/radix_sort.cu:10-62 and the
// cub::DeviceRadixSort::SortPairs call of gsplat/cuda/csrc/IntersectTile.cu (radix_sort_double_buffer).
// Differences that matter on this path:
//   * the item count may live on the DEVICE (`n_dev`): the speculatively sized isect_tiles path sorts exactly the
//     intersections that exist instead of a padded capacity, still without a host read-back;
//   * all scratch (digit histograms, per-tile look-back states, tile tickets) is one caller-provided buffer that a
//     single memset clears; 2 + passes launches per sort.
//
// Pass structure. A tile is kTile = 256 threads x kItems consecutive pairs. Warp w owns a contiguous chunk of
// 32 x kItems pairs, loaded warp-striped (coalesced). Ranks are stable by construction: round i, lane l holds
// position i*32 + l of the chunk; `match.any` groups the lanes with equal digits, the lowest lane of each group
// bumps the warp's running digit counter, so rank = (pairs with this digit earlier in the chunk). Warp counters
// are summed to tile counts; thread d publishes the tile's count of digit d and walks back over the preceding
// tiles' states (LOCAL count or INCLUSIVE prefix) to get its global offset; pairs are staged through shared memory
// in digit order and written out as coalesced runs.
#include "common.cuh"

#include <atomic>

#include <cstdint>

namespace {

// Latency, not instruction count, bounds a pass (ncu: 19 % issue slots at 16 resident warps/SM with 256 x 16-pair
// tiles and 128 registers): 512 threads x 8 pairs keeps the per-thread state at key[8] + rank[8], the value is
// fetched only when the pair is staged, and three tiles (48 warps) fit an SM.
#ifndef GSB_SORT_MIN_BLOCKS
#define GSB_SORT_MIN_BLOCKS 3
#endif
#ifndef GSB_SORT_ITEMS
#define GSB_SORT_ITEMS 8
#endif
constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;
template <typename KeyT>
struct TileOf {
    static constexpr int kItems = GSB_SORT_ITEMS;
    static constexpr int kTile = kThreads * kItems;  // 4096 pairs
    static constexpr size_t kSmem = sizeof(uint32_t) * kWarps * 256 + sizeof(int64_t) * 256 +
                                    sizeof(KeyT) * kTile + sizeof(int32_t) * kTile;
};
constexpr int kBins = 256;
constexpr uint32_t kFlagLocal = 1u << 30, kFlagPrefix = 2u << 30, kValueMask = (1u << 30) - 1u;

template <typename KeyT>
__device__ __forceinline__ int digit_of(KeyT key, int shift, int mask) {
    return (int)((key >> shift) & (KeyT)mask);
}

__device__ __forceinline__ int64_t effective_count(int64_t n, const int64_t *n_dev) {
    if (n_dev == nullptr) return n;
    const int64_t m = *n_dev;
    return m < n ? (m < 0 ? 0 : m) : n;
}

// scratch layout (uint32 words): [passes][256] digit histograms | [passes] tile tickets (padded to 64) |
// [passes][tiles][256] look-back states
__host__ __device__ __forceinline__ int64_t scratch_words(int passes, int64_t tiles) {
    return (int64_t)passes * kBins + 64 + (int64_t)passes * tiles * kBins;
}

template <typename KeyT>
__global__ void __launch_bounds__(kThreads)
histogram_kernel(int64_t n, const int64_t *__restrict__ n_dev, const KeyT *__restrict__ keys, int passes, int end_bit,
                 uint32_t *__restrict__ hist) {
    __shared__ uint32_t s_hist[8 * kBins];
    n = effective_count(n, n_dev);
    for (int i = threadIdx.x; i < passes * kBins; i += kThreads) s_hist[i] = 0u;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kThreads) {
        const KeyT key = keys[i];
        for (int p = 0; p < passes; ++p) {
            const int bits = min(8, end_bit - 8 * p);
            atomicAdd(&s_hist[p * kBins + digit_of(key, 8 * p, (1 << bits) - 1)], 1u);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * kBins; i += kThreads) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&hist[i], c);
    }
}

// Lanes of the warp holding the same digit (dead lanes, digit kBins, match only each other).
// GSB_SORT_MATCH=1: one vote per digit bit, intersected (ALU/vote pipe only); 0: one match.any per item.
#ifndef GSB_SORT_MATCH
#define GSB_SORT_MATCH 1
#endif
__device__ __forceinline__ uint32_t match_digit(int digit, bool live, int bits) {
#if GSB_SORT_MATCH
    uint32_t peers = __ballot_sync(0xffffffffu, live);
    if (!live) peers = ~peers;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        if (b < bits) {  // warp-uniform
            const bool one = (digit >> b) & 1;
            const uint32_t with_one = __ballot_sync(0xffffffffu, one);
            peers &= one ? with_one : ~with_one;
        }
    }
    return peers;
#else
    (void)live;
    (void)bits;
    return __match_any_sync(0xffffffffu, digit);
#endif
}

template <typename KeyT>
__global__ void __launch_bounds__(kThreads, sizeof(KeyT) == 4 ? GSB_SORT_MIN_BLOCKS : 2)
onesweep_kernel(int64_t n, const int64_t *__restrict__ n_dev, const KeyT *__restrict__ keys_in,
                const int32_t *__restrict__ vals_in, KeyT *__restrict__ keys_out, int32_t *__restrict__ vals_out,
                int shift, int bits, const uint32_t *__restrict__ hist, uint32_t *__restrict__ ticket,
                uint32_t *__restrict__ states) {
    gsb_chain_enter();
    constexpr int kItems = TileOf<KeyT>::kItems, kTile = TileOf<KeyT>::kTile;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int32_t *s_global = reinterpret_cast<int32_t *>(smem_raw);     // (global position - staged slot) of each digit
    KeyT *s_keys = reinterpret_cast<KeyT *>(smem_raw + sizeof(int64_t) * kBins);  // staged tile, digit order
    int32_t *s_vals = reinterpret_cast<int32_t *>(s_keys + kTile);
    uint32_t(*s_warp_count)[kBins] = reinterpret_cast<uint32_t(*)[kBins]>(s_vals + kTile);  // per-warp digit counters
    __shared__ uint32_t s_scan[kWarps];
    __shared__ uint32_t s_tile;

    n = effective_count(n, n_dev);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);  // tiles start in ticket order: look-back always finds a live tile
    for (int i = tid; i < kWarps * kBins; i += kThreads) (&s_warp_count[0][0])[i] = 0u;
    __syncthreads();
    const int64_t tile = s_tile;
    const int64_t base = tile * kTile;
    if (base >= n) return;
    const int mask = (1 << bits) - 1;
    const int count = (int)min((int64_t)kTile, n - base);

    // ---- load (warp-striped) and rank inside the warp's chunk
    KeyT key[kItems];
    uint32_t rank[kItems];
    const int chunk = warp * 32 * kItems;
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const int pos = chunk + i * 32 + lane;
        const bool live = pos < count;
        key[i] = live ? keys_in[base + pos] : (KeyT)0;
    }
    // Three sweeps so nothing waits on a round trip: (1) all match.any groupings, (2) the group leaders' shared-memory
    // atomics, issued back to back (one warp's shared-memory operations execute in program order, so round i's
    // counter value already includes rounds < i: the stable rank), (3) the broadcasts of the returned bases.
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t peers[kItems];
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const bool live = chunk + i * 32 + lane < count;
        // dead slots take a digit of their own (kBins: never equal to a live digit) so they match only each other
        peers[i] = match_digit(live ? digit_of(key[i], shift, mask) : kBins, live, bits);
    }
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const bool live = chunk + i * 32 + lane < count;
        rank[i] = 0u;
        if (live && lane == __ffs(peers[i]) - 1)
            rank[i] = atomicAdd(&s_warp_count[warp][digit_of(key[i], shift, mask)], (uint32_t)__popc(peers[i]));
    }
#pragma unroll
    for (int i = 0; i < kItems; ++i)
        rank[i] = __shfl_sync(0xffffffffu, rank[i], __ffs(peers[i]) - 1) + __popc(peers[i] & lt);
    __syncthreads();

    // ---- thread d < 256 (the "digit threads"): tile count of digit d, publish it, start the look-back
    uint32_t total = 0u, inc = 0u, hinc = 0u, h = 0u;
    constexpr int kLook = 4;
    uint32_t look[kLook];  // the first kLook predecessor states: requested now, consumed after staging
    const int d = tid;
    const bool digit_thread = tid < kBins;
    if (digit_thread) {
#pragma unroll
        for (int w = 0; w < kWarps; ++w) total += s_warp_count[w][d];
        __stcg(states + tile * kBins + d, (tile == 0 ? kFlagPrefix : kFlagLocal) | total);
#pragma unroll
        for (int j = 0; j < kLook; ++j)
            look[j] = (tile - 1 - j >= 0) ? __ldcv(states + (tile - 1 - j) * kBins + d) : kFlagPrefix;
        // warp-level inclusive scans: tile counts over digits, and the pass histogram (global digit bases)
        inc = total;
        h = hist[d];
        hinc = h;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, inc, o);
            const uint32_t hup = __shfl_up_sync(0xffffffffu, hinc, o);
            if (lane >= o) {
                inc += up;
                hinc += hup;
            }
        }
        if (lane == 31) {
            s_scan[warp] = inc;
            s_scan[8 + warp] = hinc;
        }
    }
    __syncthreads();
    uint32_t bucket = 0u, hbase = 0u;
    if (digit_thread) {
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            bucket += (w < warp) ? s_scan[w] : 0u;
            hbase += (w < warp) ? s_scan[8 + w] : 0u;
        }
        bucket += inc - total;  // first slot of digit d inside the staged tile
        hbase += hinc - h;      // first global position of digit d in this pass
        // per-warp counters -> first staged slot of (warp, digit): bucket start + exclusive sum over earlier warps
        uint32_t run = bucket;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
            const uint32_t c = s_warp_count[w][d];
            s_warp_count[w][d] = run;
            run += c;
        }
    }
    __syncthreads();

    // ---- stage in digit order (the look-back loads of the digit threads are in flight meanwhile)
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const int pos = chunk + i * 32 + lane;
        if (pos < count) {
            const uint32_t slot = s_warp_count[warp][digit_of(key[i], shift, mask)] + rank[i];
            s_keys[slot] = key[i];
            s_vals[slot] = vals_in[base + pos];
        }
    }
    if (digit_thread) {
        // finish the look-back: kLook states per round trip (independent loads in flight together; with a whole
        // wave of tiles starting at once a one-at-a-time walk costs ~tile/2 serial L2 latencies)
        uint32_t prefix = 0u;
        bool done = (tile == 0);
        int64_t t = tile - 1;
        auto consume = [&](uint32_t v, int64_t at) {
            if ((v >> 30) == 0u) {  // not published yet: poll this one
                const volatile uint32_t *st = states + at * kBins + d;
                unsigned spins = 0;
                while (((v = *st) >> 30) == 0u) {
                    if (++spins > (1u << 26)) __trap();  // a preceding tile never published: fail loudly, never hang
                }
            }
            prefix += v & kValueMask;
            done = (v & kFlagPrefix) != 0u;
        };
#pragma unroll
        for (int j = 0; j < kLook; ++j)
            if (!done) consume(look[j], t - j);
        t -= kLook;
        // further back: 16 states per round trip (key[] / rank[] are dead by now, the registers are free). With a whole
        // wave of tiles starting at once the frontier of finished prefixes grows quadratically in round trips.
        constexpr int kWide = 16;
        while (!done) {
            uint32_t far[kWide];
#pragma unroll
            for (int j = 0; j < kWide; ++j) far[j] = (t - j >= 0) ? __ldcv(states + (t - j) * kBins + d) : kFlagPrefix;
#pragma unroll
            for (int j = 0; j < kWide; ++j)
                if (!done) consume(far[j], t - j);
            t -= kWide;
        }
        if (tile != 0) __stcg(states + tile * kBins + d, kFlagPrefix | ((prefix + total) & kValueMask));
        s_global[d] = (int32_t)(hbase + prefix) - (int32_t)bucket;  // n < 2^30: positions fit 32 bits
    }
    __syncthreads();

    // ---- write coalesced runs
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const int slot = i * kThreads + tid;
        if (slot < count) {
            const KeyT k = s_keys[slot];
            const int64_t dst = (int64_t)(s_global[digit_of(k, shift, mask)] + slot);
            keys_out[dst] = k;
            vals_out[dst] = s_vals[slot];
        }
    }
}

template <typename KeyT>
int sort_pairs(int64_t n, const int64_t *n_dev, int end_bit, KeyT *keys, KeyT *keys_alt, int32_t *vals,
               int32_t *vals_alt, void *temp, int64_t temp_bytes, int *result_in_alt, cudaStream_t stream,
               bool prepared = false) {
    *result_in_alt = 0;
    if (n == 0 || end_bit == 0) return GSB_OK;
    constexpr int kTile = TileOf<KeyT>::kTile;
    const int passes = (end_bit + 7) / 8;
    const int64_t tiles = (n + kTile - 1) / kTile;
    const int64_t words = scratch_words(passes, tiles);
    if (temp == nullptr || temp_bytes < words * 4) {
        gsb_set_error("radix sort: temp buffer too small (%lld < %lld bytes)", (long long)temp_bytes,
                      (long long)(words * 4));
        return GSB_ERR_INVALID;
    }
    cudaError_t e = cudaSuccess;
    if (!prepared) e = cudaMemsetAsync(temp, 0, (size_t)words * 4, stream);
    if (e != cudaSuccess) {
        gsb_set_error("radix sort: memset: %s", cudaGetErrorString(e));
        return GSB_ERR_CUDA;
    }
    // one instantiation per key type; the attribute is per DEVICE: bit d = configured on device d
    static std::atomic<unsigned long long> configured{0ull};
    int device = 0;
    cudaGetDevice(&device);
    if (!((configured.load(std::memory_order_acquire) >> (device & 63)) & 1ull)) {
        e = cudaFuncSetAttribute(onesweep_kernel<KeyT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)TileOf<KeyT>::kSmem);
        if (e != cudaSuccess) {
            gsb_set_error("radix sort: shared memory: %s", cudaGetErrorString(e));
            return GSB_ERR_CUDA;
        }
        configured.fetch_or(1ull << (device & 63), std::memory_order_release);
    }
    uint32_t *hist = reinterpret_cast<uint32_t *>(temp);
    uint32_t *tickets = hist + (int64_t)passes * kBins;
    uint32_t *states = tickets + 64;
    if (!prepared) {  // else: the caller cleared `temp` (gsb_radix_prepare) and its own kernels filled the histograms
        const unsigned hgrid = (unsigned)(tiles < 148 * 8 ? tiles : 148 * 8);
        histogram_kernel<KeyT><<<hgrid, kThreads, 0, stream>>>(n, n_dev, keys, passes, end_bit, hist);
        if (int rc = gsb_check_launch("radix_histogram")) return rc;
    }
    KeyT *kin = keys, *kout = keys_alt;
    int32_t *vin = vals, *vout = vals_alt;
    for (int p = 0; p < passes; ++p) {
        const int bits = end_bit - 8 * p < 8 ? end_bit - 8 * p : 8;
        gsb_launch_chained(GSB_CHAIN_ISECT, onesweep_kernel<KeyT>, (unsigned)tiles, kThreads, TileOf<KeyT>::kSmem, stream, n, n_dev, kin, vin,
                           kout, vout, 8 * p, bits, hist + p * kBins, tickets + p, states + (int64_t)p * tiles * kBins);
        if (int rc = gsb_check_launch("radix_onesweep")) return rc;
        KeyT *tk = kin; kin = kout; kout = tk;
        int32_t *tv = vin; vin = vout; vout = tv;
    }
    *result_in_alt = passes & 1;
    return GSB_OK;
}

}  // namespace

extern "C" int64_t gsb_radix_temp_bytes(int64_t n, int end_bit) {
    const int passes = (end_bit + 7) / 8;
    const int64_t tiles = (n + TileOf<unsigned long long>::kTile - 1) / TileOf<unsigned long long>::kTile;
    return scratch_words(passes > 0 ? passes : 1, tiles > 0 ? tiles : 1) * 4;
}

extern "C" int gsb_radix_sort_pairs_u32(int64_t n, const int64_t *n_dev, int end_bit, int32_t *keys,
                                        int32_t *keys_alt, int32_t *vals, int32_t *vals_alt, void *temp,
                                        int64_t temp_bytes, int *result_in_alt_host, cudaStream_t stream) {
    GSB_REQUIRE(n >= 0 && n < ((int64_t)1 << 30) && end_bit >= 0 && end_bit <= 32, "bad sizes (n < 2^30)");
    GSB_REQUIRE(result_in_alt_host != nullptr, "result_in_alt_host is required");
    return sort_pairs<uint32_t>(n, n_dev, end_bit, reinterpret_cast<uint32_t *>(keys),
                                reinterpret_cast<uint32_t *>(keys_alt), vals, vals_alt, temp, temp_bytes,
                                result_in_alt_host, stream);
}

// The producer of the keys can count the digits itself (it holds every key in a register anyway) and spare the sort
// its histogram pass: gsb_radix_prepare clears `temp` (gsb_radix_temp_bytes(n, end_bit) bytes); the producer adds, for
// every key and every pass p < ceil(end_bit / 8), one count to ((uint32_t *)temp)[p * 256 + ((key >> 8 p) & mask_p)]
// (mask_p = 255, or the remaining bits in the last pass); gsb_radix_sort_pairs_u32_prepared then runs the passes only.
extern "C" int gsb_radix_prepare(int64_t n, int end_bit, void *temp, int64_t temp_bytes, cudaStream_t stream) {
    GSB_REQUIRE(n >= 0 && end_bit >= 0 && end_bit <= 64, "bad sizes");
    const int64_t need = gsb_radix_temp_bytes(n, end_bit);
    GSB_REQUIRE(temp != nullptr && temp_bytes >= need, "temp buffer too small");
    cudaError_t e = cudaMemsetAsync(temp, 0, (size_t)need, stream);
    if (e != cudaSuccess) {
        gsb_set_error("radix prepare: memset: %s", cudaGetErrorString(e));
        return GSB_ERR_CUDA;
    }
    return GSB_OK;
}

extern "C" int gsb_radix_sort_pairs_u32_prepared(int64_t n, const int64_t *n_dev, int end_bit, int32_t *keys,
                                                 int32_t *keys_alt, int32_t *vals, int32_t *vals_alt, void *temp,
                                                 int64_t temp_bytes, int *result_in_alt_host, cudaStream_t stream) {
    GSB_REQUIRE(n >= 0 && n < ((int64_t)1 << 30) && end_bit >= 0 && end_bit <= 32, "bad sizes (n < 2^30)");
    GSB_REQUIRE(result_in_alt_host != nullptr, "result_in_alt_host is required");
    return sort_pairs<uint32_t>(n, n_dev, end_bit, reinterpret_cast<uint32_t *>(keys),
                                reinterpret_cast<uint32_t *>(keys_alt), vals, vals_alt, temp, temp_bytes,
                                result_in_alt_host, stream, true);
}

extern "C" int gsb_radix_sort_pairs_u64(int64_t n, const int64_t *n_dev, int end_bit, int64_t *keys,
                                        int64_t *keys_alt, int32_t *vals, int32_t *vals_alt, void *temp,
                                        int64_t temp_bytes, int *result_in_alt_host, cudaStream_t stream) {
    GSB_REQUIRE(n >= 0 && n < ((int64_t)1 << 30) && end_bit >= 0 && end_bit <= 64, "bad sizes (n < 2^30)");
    GSB_REQUIRE(result_in_alt_host != nullptr, "result_in_alt_host is required");
    return sort_pairs<unsigned long long>(n, n_dev, end_bit, reinterpret_cast<unsigned long long *>(keys),
                                          reinterpret_cast<unsigned long long *>(keys_alt), vals, vals_alt, temp,
                                          temp_bytes, result_in_alt_host, stream);
}

// ---- the same sort for callers that cannot allocate device memory (CUDA graph capture, serving loops, harnesses that
// forbid cudaMalloc). No counterpart in the reference.
// The scratch (digit histograms, tile tickets, look-back states) lives in a file-scope device array instead of a
// caller-provided `temp` buffer: 2.1 MB per GPU that loads the library, held until exit, sized for
// GSB_RADIX_NOALLOC_MAX_ITEMS = 2^21 pairs of 32-bit keys. Every call on a device shares this array: the entry point
// clears it on the caller's stream before the passes, and two calls whose kernels overlap on the GPU (different
// streams, no ordering between them) race on it.
namespace {
constexpr int64_t kNoAllocMaxItems = GSB_RADIX_NOALLOC_MAX_ITEMS;
constexpr int64_t kNoAllocTiles = (kNoAllocMaxItems + TileOf<uint32_t>::kTile - 1) / TileOf<uint32_t>::kTile;
constexpr int64_t kNoAllocWords = 4 * kBins + 64 + 4 * kNoAllocTiles * kBins;  // scratch_words(4 passes, tiles)
constexpr int kNoAllocMaxDevices = 64;
}  // namespace

static __device__ __align__(256) uint32_t g_noalloc_scratch[kNoAllocWords];

namespace {
// Device address of g_noalloc_scratch on the CURRENT device. In host code the array's name is only a handle for the
// runtime, not an address a kernel can use: it is resolved once per device with cudaGetSymbolAddress.
int noalloc_scratch_ptr(uint32_t **out) {
    static uint32_t *cached[kNoAllocMaxDevices] = {nullptr};
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) {
        gsb_set_error("radix sort (noalloc): cudaGetDevice: %s", cudaGetErrorString(e));
        return GSB_ERR_CUDA;
    }
    const bool cacheable = device >= 0 && device < kNoAllocMaxDevices;
    uint32_t *p = cacheable ? __atomic_load_n(&cached[device], __ATOMIC_ACQUIRE) : nullptr;
    if (p == nullptr) {
        void *addr = nullptr;
        e = cudaGetSymbolAddress(&addr, g_noalloc_scratch);
        if (e != cudaSuccess || addr == nullptr) {
            gsb_set_error("radix sort (noalloc): cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return GSB_ERR_CUDA;
        }
        p = static_cast<uint32_t *>(addr);
        if (cacheable) __atomic_store_n(&cached[device], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return GSB_OK;
}
}  // namespace

extern "C" int gsb_radix_sort_pairs_u32_noalloc(int64_t n, const int64_t *n_dev, int end_bit, int32_t *keys,
                                                int32_t *keys_alt, int32_t *vals, int32_t *vals_alt,
                                                int *result_in_alt_host, cudaStream_t stream) {
    GSB_REQUIRE(n >= 0 && end_bit >= 0 && end_bit <= 32, "bad sizes");
    GSB_REQUIRE(result_in_alt_host != nullptr, "result_in_alt_host is required");
    if (n > kNoAllocMaxItems) {
        gsb_set_error("gsb_radix_sort_pairs_u32_noalloc: %lld items exceed the built-in scratch "
                      "(GSB_RADIX_NOALLOC_MAX_ITEMS = %lld); use gsb_radix_sort_pairs_u32 with a temp buffer",
                      (long long)n, (long long)kNoAllocMaxItems);
        return GSB_ERR_INVALID;
    }
    uint32_t *scratch = nullptr;
    if (int rc = noalloc_scratch_ptr(&scratch)) return rc;
    // sort_pairs clears the words this n needs (cudaMemsetAsync on `stream`: no allocation) and runs the passes
    return sort_pairs<uint32_t>(n, n_dev, end_bit, reinterpret_cast<uint32_t *>(keys),
                                reinterpret_cast<uint32_t *>(keys_alt), vals, vals_alt, scratch,
                                kNoAllocWords * (int64_t)sizeof(uint32_t), result_in_alt_host, stream);
}
