// Fused pack + peer-store exchange of projected Gaussians over NVLink (one process per GPU).
//
// Replaces (behaviour, not code) the per-camera all-to-all of the reference's distributed path,
// gsplat/rendering.py:453-494 + gsplat/distributed.py:170-257: cat -> split -> NCCL all_to_all ->
// cat -> reshape_view, twice per direction (floats, then the int32 radii).
//
// Here every rank owns a "mailbox" allocation that its peers map through CUDA IPC. The put kernel
// reads the projected SoA tensors once and stores each (8+D)-word row straight into the mailbox of
// the rank that owns the row's camera, already in that rank's final camera-major order
// [C_local, N_total]; the last CTA to finish publishes an epoch number into every peer's flag slot
// (release, system scope). The receiving side runs ONE kernel that waits for the W flags (acquire)
// and unpacks its mailbox into the SoA tensors the compositor reads. No NCCL call, no staging copy,
// no host synchronisation; the transfer overlaps the packing store by store.
//
// Mailbox layout (identical on every rank):
//   [128 ch, +128)  uint64 flags[ch][16]  slot s = epoch of the last put completed by rank s on channel ch;
//                   channels: 0 forward (geometry, or all fields), 1 backward (geometry, or all fields),
//                   2 forward colours, 3 backward colours (the split exchange: colours travel under the receiver's
//                   intersection pipeline)
//   [768, 784)      uint32 done[4]        CTA completion counters of the local put kernels, per channel
//   [1024, ...)     4 regions of `cap` bytes: (direction = ch & 1, epoch parity) -> 1024 + (2*dir + parity) * cap;
//                   both channels of a direction write disjoint fields of the same region
// Two regions per direction: a peer may start step t+1's put while this rank still unpacks step t,
// but cannot reach step t+2's put before this rank has published step t+1 (stream order), see DESIGN.md 6.
#include "common.cuh"
#include "peer_layout.cuh"

#include <cstring>

namespace {

constexpr int kThreads = 256;
constexpr int kMaxPeers = GSB_MAX_PEERS;
constexpr int64_t kHeaderBytes = 1024;
constexpr unsigned long long kWaitTimeoutNs = 20ull * 1000ull * 1000ull * 1000ull;

struct PeerTable {
    char *base[kMaxPeers];
};
struct RankBases {
    int64_t base[kMaxPeers + 1];
};

__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

constexpr int kChannels = 4;
constexpr int64_t kDoneOffset = 768;

// called by every thread of the CTA after its peer stores: the last CTA of the grid publishes `epoch` on channel `dir`
__device__ __forceinline__ void publish_when_grid_done(const PeerTable &peers, int world, int rank, int dir,
                                                       unsigned long long epoch) {
    __shared__ bool s_last;
    __syncthreads();  // every thread's peer stores happen-before thread 0's fence (cumulative)
    unsigned *done = reinterpret_cast<unsigned *>(peers.base[rank] + kDoneOffset) + dir;
    if (threadIdx.x == 0) {
        __threadfence_system();
        unsigned prev = atomicAdd(done, 1u);
        s_last = (prev == gridDim.x - 1);
        if (s_last) *done = 0u;  // ready for the next put in this direction (stream ordered)
    }
    __syncthreads();
    if (s_last && threadIdx.x < world) {
        __threadfence_system();
        unsigned long long *flag =
            reinterpret_cast<unsigned long long *>(peers.base[threadIdx.x] + 128 * dir) + rank;
        st_release_sys(flag, epoch);
    }
}

// ---- mailbox region layout: structure of arrays ---------------------------------------------------
// A region holding `rows` rows stores field after field (means2d 2, depths 1, conics 3, opacities 1,
// colours D, [radii 1]), each field starting on a 16-byte boundary: the receiving rank's six output tensors
// are views of ONE buffer with the same layout, so unpacking is a straight vectorised copy, and every
// (camera, field) run a sender owns is contiguous on both sides.
constexpr int kFields = 6;
constexpr int kPackedFields = 8;  // + camera ids, Gaussian ids (packed rows only)
struct Fields {
    const float *src[kFields];
    int width[kFields];
    int count;            // number of entries of `active`
    int active[kFields];  // the fields this launch sends (a split exchange sends geometry and colours separately)
};
using peer::field_offset;
using peer::round_up4;

// copies `len` floats src -> dst with this CTA column (blockIdx.x of gridDim.x); src == NULL stores zeros.
// The stores are what crosses NVLink: they are always 16-byte vectors on the 16-byte aligned body of dst (a scalar
// head / tail of < 4 floats around it); the loads are vectors when src has the same alignment phase (unpacked rows
// always; packed runs start at arbitrary rows) and four scalar loads otherwise.
__device__ __forceinline__ void copy_run(const float *__restrict__ src, float *__restrict__ dst, int64_t len) {
    const int64_t tid = (int64_t)blockIdx.x * kThreads + threadIdx.x, step = (int64_t)gridDim.x * kThreads;
    int64_t head = (4 - (int64_t)(((uintptr_t)dst >> 2) & 3)) & 3;  // floats before dst is 16-byte aligned
    if (head > len) head = len;
    const int64_t n4 = (len - head) >> 2;
    float4 *d4 = reinterpret_cast<float4 *>(dst + head);
    if (src == nullptr) {
        for (int64_t i = tid; i < n4; i += step) d4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    } else if ((((uintptr_t)(src + head)) & 15) == 0) {
        const float4 *s4 = reinterpret_cast<const float4 *>(src + head);
        for (int64_t i = tid; i < n4; i += step) d4[i] = __ldg(s4 + i);
    } else {
        const float *s = src + head;
        for (int64_t i = tid; i < n4; i += step)
            d4[i] = make_float4(__ldg(s + 4 * i), __ldg(s + 4 * i + 1), __ldg(s + 4 * i + 2), __ldg(s + 4 * i + 3));
    }
    // head and tail: fewer than 8 floats in total
    if (tid < head) dst[tid] = src ? src[tid] : 0.f;
    const int64_t tail0 = head + (n4 << 2);
    if (tid < len - tail0) dst[tail0 + tid] = src ? src[tail0 + tid] : 0.f;
}

// forward: segment (camera c of all world*C_local cameras, field f): N_local rows -> the rank owning camera c,
// at rows [c_local, n_base .. n_base + N_local) of its [C_local, N_total] table
__global__ void __launch_bounds__(kThreads)
put_fwd_kernel(PeerTable peers, Fields fields, int world, int rank, int C_local, int64_t N_local, int64_t N_total,
               int64_t n_base, int64_t region_offset, unsigned long long epoch, int channel) {
    // segment order: camera fastest and rotated by rank, so the CTAs in flight at any moment cover every
    // destination and no two senders start on the same one (an 8-to-1 incast would share one ingress port)
    const int n_cam = world * C_local;
    const int f = fields.active[blockIdx.y / n_cam];
    const int c = (blockIdx.y % n_cam + (rank + 1) * C_local) % n_cam;
    const int dest = c / C_local, c_local = c - dest * C_local;
    const int w = fields.width[f];
    const float *src = fields.src[f] + (int64_t)c * N_local * w;
    float *dst = reinterpret_cast<float *>(peers.base[dest] + region_offset) +
                 field_offset(fields.width, f, (int64_t)C_local * N_total) + ((int64_t)c_local * N_total + n_base) * w;
    copy_run(src, dst, N_local * w);
    publish_when_grid_done(peers, world, rank, channel, epoch);
}

// backward: segment (local camera c_local, destination rank r, field f): the N_r gradient rows of rank r's
// Gaussians -> rank r, at rows [rank*C_local + c_local, 0 .. N_r) of its [world*C_local, N_r] table
__global__ void __launch_bounds__(kThreads)
put_bwd_kernel(PeerTable peers, Fields fields, RankBases bases, int world, int rank, int C_local, int64_t N_total,
               int64_t region_offset, unsigned long long epoch, int channel) {
    // destination fastest and rotated by rank (see put_fwd_kernel)
    const int n_seg = world * C_local;
    const int f = fields.active[blockIdx.y / n_seg];
    const int q = blockIdx.y % n_seg;
    const int dest = (q + rank + 1) % world, c_local = q / world;
    const int w = fields.width[f];
    const int64_t b0 = bases.base[dest], N_dest = bases.base[dest + 1] - b0;
    const float *src = fields.src[f] ? fields.src[f] + ((int64_t)c_local * N_total + b0) * w : nullptr;
    float *dst = reinterpret_cast<float *>(peers.base[dest] + region_offset) +
                 field_offset(fields.width, f, (int64_t)world * C_local * N_dest) +
                 ((int64_t)rank * C_local + c_local) * N_dest * w;
    copy_run(src, dst, N_dest * w);
    publish_when_grid_done(peers, world, rank, channel, epoch);
}

// ---- packed rows (COO): variable row counts per destination --------------------------------------------
// The sender's packed rows are camera-major, so the rows owned by destination rank d are one contiguous run
// [src_start[d], src_start[d] + count[d]); they land in d's table (dst_rows[d] rows, same SoA layout, two extra
// int32 fields) at row dst_base[d] = rows of lower-ranked senders. The id fields are rebased on the fly: camera ids
// global -> local to the destination, Gaussian ids local -> global (rendering.py:436-451 of the reference).
struct PackedPlan {
    int64_t src_start[kMaxPeers], count[kMaxPeers], dst_base[kMaxPeers], dst_rows[kMaxPeers], cam_base[kMaxPeers];
    int64_t gauss_base;
};
struct PackedFields {
    const void *src[kPackedFields];
    int width[kPackedFields];
};

__global__ void __launch_bounds__(kThreads)
put_packed_fwd_kernel(PeerTable peers, PackedFields fields, PackedPlan plan, int world, int rank,
                      int64_t region_offset, unsigned long long epoch) {
    const int f = blockIdx.y / world;
    const int dest = (blockIdx.y % world + rank + 1) % world;  // rotated by rank: no incast at start
    const int w = fields.width[f];
    const int64_t n = plan.count[dest];
    float *dst = reinterpret_cast<float *>(peers.base[dest] + region_offset) +
                 field_offset(fields.width, f, plan.dst_rows[dest]) + plan.dst_base[dest] * w;
    if (f < kFields) {
        copy_run(reinterpret_cast<const float *>(fields.src[f]) + plan.src_start[dest] * w, dst, n * w);
    } else {
        const int64_t *ids = reinterpret_cast<const int64_t *>(fields.src[f]) + plan.src_start[dest];
        const int64_t shift = (f == kFields) ? -plan.cam_base[dest] : plan.gauss_base;
        int32_t *d32 = reinterpret_cast<int32_t *>(dst);
        const int64_t tid = (int64_t)blockIdx.x * kThreads + threadIdx.x, step = (int64_t)gridDim.x * kThreads;
        int64_t head = (4 - (int64_t)(((uintptr_t)d32 >> 2) & 3)) & 3;
        if (head > n) head = n;
        const int64_t n4 = (n - head) >> 2;
        int4 *d4 = reinterpret_cast<int4 *>(d32 + head);
        const int64_t *s = ids + head;
        for (int64_t i = tid; i < n4; i += step)
            d4[i] = make_int4((int32_t)(s[4 * i] + shift), (int32_t)(s[4 * i + 1] + shift),
                              (int32_t)(s[4 * i + 2] + shift), (int32_t)(s[4 * i + 3] + shift));
        if (tid < head) d32[tid] = (int32_t)(ids[tid] + shift);
        const int64_t tail0 = head + (n4 << 2);
        if (tid < n - tail0) d32[tail0 + tid] = (int32_t)(ids[tail0 + tid] + shift);
    }
    publish_when_grid_done(peers, world, rank, 0, epoch);
}

// backward: the gradient rows this rank holds for sender r's Gaussians ([src_start[r], +count[r]) of its table)
// go back to r, at rows [dst_base[r], +count[r]) of r's packed order (dst_rows[r] = nnz of r)
__global__ void __launch_bounds__(kThreads)
put_packed_bwd_kernel(PeerTable peers, PackedFields fields, PackedPlan plan, int world, int rank,
                      int64_t region_offset, unsigned long long epoch) {
    const int f = blockIdx.y / world;
    const int dest = (blockIdx.y % world + rank + 1) % world;
    const int w = fields.width[f];
    const float *src = reinterpret_cast<const float *>(fields.src[f]);
    float *dst = reinterpret_cast<float *>(peers.base[dest] + region_offset) +
                 field_offset(fields.width, f, plan.dst_rows[dest]) + plan.dst_base[dest] * w;
    copy_run(src ? src + plan.src_start[dest] * w : nullptr, dst, plan.count[dest] * w);
    publish_when_grid_done(peers, world, rank, 1, epoch);
}

// wait until every rank has published `epoch` in direction `dir`, then copy the local region out
__global__ void __launch_bounds__(kThreads)
wait_copy_kernel(const char *__restrict__ local_base, int world, int dir, unsigned long long epoch,
                 int64_t region_offset, int64_t begin0, int64_t end0, int64_t begin1, int64_t end1,
                 float *__restrict__ out, int *timeout_flag) {
    if (threadIdx.x < world) {
        const unsigned long long *flag =
            reinterpret_cast<const unsigned long long *>(local_base + 128 * dir) + threadIdx.x;
        const unsigned long long t0 = global_timer_ns();
        unsigned spins = 0;
        while (ld_acquire_sys(flag) < epoch) {
            if (++spins > 64u) {
                __nanosleep(100);
                if (global_timer_ns() - t0 > kWaitTimeoutNs) {  // a peer never arrived: report, do not hang the GPU
                    if (timeout_flag) *reinterpret_cast<volatile int *>(timeout_flag) = 1 + (int)threadIdx.x;
                    break;
                }
            }
        }
    }
    __syncthreads();
    // region and `out` are 16-byte aligned (region offsets are multiples of 1 KB); peers wrote through L2,
    // so read with .cg: never served from a stale L1 line
    // two float ranges of the region (field offsets are multiples of 4 floats): [begin0, end0) and [begin1, end1)
    const float4 *box = reinterpret_cast<const float4 *>(local_base + region_offset);
    float4 *o4 = reinterpret_cast<float4 *>(out);
    const int64_t b0 = begin0 >> 2, n0 = (end0 - begin0) >> 2, b1 = begin1 >> 2, n1 = (end1 - begin1) >> 2;
    for (int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x; i < n0 + n1; i += (int64_t)gridDim.x * kThreads) {
        const int64_t k = i < n0 ? b0 + i : b1 + (i - n0);
        o4[k] = __ldcg(box + k);
    }
}

int fill_peers(PeerTable &t, const void *const *peer_bases_host, int world) {
    for (int i = 0; i < kMaxPeers; ++i) t.base[i] = nullptr;
    for (int i = 0; i < world; ++i) {
        if (peer_bases_host[i] == nullptr) return -1;
        t.base[i] = reinterpret_cast<char *>(const_cast<void *>(peer_bases_host[i]));
    }
    return 0;
}

#define GSB_CUDA_CALL(expr)                                               \
    do {                                                                  \
        cudaError_t e_ = (expr);                                          \
        if (e_ != cudaSuccess) {                                          \
            gsb_set_error("%s: %s", #expr, cudaGetErrorString(e_));       \
            (void)cudaGetLastError();                                     \
            return GSB_ERR_CUDA;                                          \
        }                                                                 \
    } while (0)

}  // namespace

extern "C" int64_t gsb_peer_header_bytes(void) { return kHeaderBytes; }
extern "C" int gsb_peer_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }

extern "C" int gsb_peer_alloc(int64_t bytes, void **dptr_out_host) {
    GSB_REQUIRE(bytes >= kHeaderBytes && dptr_out_host != nullptr, "bad arguments");
    void *p = nullptr;
    GSB_CUDA_CALL(cudaMalloc(&p, (size_t)bytes));
    GSB_CUDA_CALL(cudaMemset(p, 0, (size_t)kHeaderBytes));
    GSB_CUDA_CALL(cudaDeviceSynchronize());
    *dptr_out_host = p;
    return GSB_OK;
}

extern "C" int gsb_peer_free(void *dptr) {
    if (dptr != nullptr) GSB_CUDA_CALL(cudaFree(dptr));
    return GSB_OK;
}

extern "C" int gsb_peer_get_handle(void *dptr, void *handle_out_host) {
    GSB_REQUIRE(dptr != nullptr && handle_out_host != nullptr, "bad arguments");
    GSB_CUDA_CALL(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t *>(handle_out_host), dptr));
    return GSB_OK;
}

extern "C" int gsb_peer_open_handle(const void *handle_host, void **dptr_out_host) {
    GSB_REQUIRE(handle_host != nullptr && dptr_out_host != nullptr, "bad arguments");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle_host, sizeof(h));
    void *p = nullptr;
    GSB_CUDA_CALL(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *dptr_out_host = p;
    return GSB_OK;
}

extern "C" int gsb_peer_close_handle(void *dptr) {
    if (dptr != nullptr) GSB_CUDA_CALL(cudaIpcCloseMemHandle(dptr));
    return GSB_OK;
}

extern "C" int gsb_peer_flag_alloc(void **flag_out_host) {
    GSB_REQUIRE(flag_out_host != nullptr, "bad arguments");
    int *p = nullptr;
    GSB_CUDA_CALL(cudaHostAlloc(reinterpret_cast<void **>(&p), sizeof(int), cudaHostAllocMapped | cudaHostAllocPortable));
    *p = 0;
    *flag_out_host = p;
    return GSB_OK;
}

extern "C" int gsb_peer_flag_free(void *flag_host) {
    if (flag_host != nullptr) GSB_CUDA_CALL(cudaFreeHost(flag_host));
    return GSB_OK;
}

extern "C" int64_t gsb_a2a_field_offset(int64_t rows, int D, int field) {
    const int width[kPackedFields] = {2, 1, 3, 1, D, 1, 1, 1};
    if (field < 0 || field > kPackedFields) return -1;
    return field_offset(width, field, rows);
}

namespace {
unsigned blocks_per_run(int64_t len) {
    int64_t b = (len / 4 + kThreads - 1) / kThreads;  // one float4 per thread
    return (unsigned)(b < 1 ? 1 : (b > 64 ? 64 : b));
}
}  // namespace

namespace {
// fields selected by `mask` (bit f = field f), in field order
void select_fields(Fields &fields, int n_fields, int mask) {
    fields.count = 0;
    for (int f = 0; f < n_fields; ++f)
        if ((mask >> f) & 1) fields.active[fields.count++] = f;
}
}  // namespace

extern "C" int gsb_a2a_put_fwd_part(int world, int rank, int C_local, int64_t N_local, int64_t N_total, int64_t n_base,
                                    int D, const float *means2d, const float *depths, const float *conics,
                                    const float *opacities, const float *colors, const int32_t *radii,
                                    int fields_mask, int channel, const void *const *peer_bases_host,
                                    int64_t region_offset, int64_t epoch, cudaStream_t stream) {
    GSB_REQUIRE(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world, "bad world/rank");
    GSB_REQUIRE(C_local >= 1 && N_local >= 0 && N_total >= N_local && n_base >= 0 && D >= 1, "bad sizes");
    GSB_REQUIRE((int64_t)world * C_local * kFields <= 65535, "too many cameras for one launch");
    GSB_REQUIRE(region_offset >= kHeaderBytes && epoch > 0, "bad region/epoch");
    GSB_REQUIRE((channel == 0 || channel == 2) && fields_mask > 0 && fields_mask < (1 << kFields), "bad channel/fields");
    PeerTable peers;
    GSB_REQUIRE(fill_peers(peers, peer_bases_host, world) == 0, "null peer mailbox");
    Fields fields = {{means2d, depths, conics, opacities, colors, reinterpret_cast<const float *>(radii)},
                     {2, 1, 3, 1, D, 1}, 0, {0, 0, 0, 0, 0, 0}};
    select_fields(fields, kFields, fields_mask);
    for (int i = 0; i < fields.count; ++i) GSB_REQUIRE(fields.src[fields.active[i]] != nullptr, "null field");
    dim3 grid(blocks_per_run(N_local * (D > 3 ? D : 3)), (unsigned)(world * C_local * fields.count));
    put_fwd_kernel<<<grid, kThreads, 0, stream>>>(peers, fields, world, rank, C_local, N_local, N_total, n_base,
                                                 region_offset, (unsigned long long)epoch, channel);
    return gsb_check_launch("a2a_put_fwd");
}

extern "C" int gsb_a2a_put_fwd(int world, int rank, int C_local, int64_t N_local, int64_t N_total, int64_t n_base,
                               int D, const float *means2d, const float *depths, const float *conics,
                               const float *opacities, const float *colors, const int32_t *radii,
                               const void *const *peer_bases_host, int64_t region_offset, int64_t epoch,
                               cudaStream_t stream) {
    return gsb_a2a_put_fwd_part(world, rank, C_local, N_local, N_total, n_base, D, means2d, depths, conics, opacities,
                                colors, radii, (1 << kFields) - 1, 0, peer_bases_host, region_offset, epoch, stream);
}

extern "C" int gsb_a2a_put_bwd_part(int world, int rank, int C_local, int64_t N_total, const int64_t *n_bases_host,
                                    int D, const float *v_means2d, const float *v_depths, const float *v_conics,
                                    const float *v_opacities, const float *v_colors, int fields_mask, int channel,
                                    const void *const *peer_bases_host, int64_t region_offset, int64_t epoch,
                                    cudaStream_t stream) {
    GSB_REQUIRE(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world, "bad world/rank");
    GSB_REQUIRE(C_local >= 1 && N_total >= 0 && D >= 1 && n_bases_host != nullptr, "bad sizes");
    GSB_REQUIRE((int64_t)world * C_local * kFields <= 65535, "too many cameras for one launch");
    GSB_REQUIRE(region_offset >= kHeaderBytes && epoch > 0, "bad region/epoch");
    GSB_REQUIRE((channel == 1 || channel == 3) && fields_mask > 0 && fields_mask < (1 << (kFields - 1)),
                "bad channel/fields");
    PeerTable peers;
    GSB_REQUIRE(fill_peers(peers, peer_bases_host, world) == 0, "null peer mailbox");
    RankBases bases;
    int64_t widest = 0;
    for (int i = 0; i <= kMaxPeers; ++i) bases.base[i] = (i <= world) ? n_bases_host[i] : n_bases_host[world];
    for (int i = 0; i < world; ++i) {
        GSB_REQUIRE(bases.base[i + 1] >= bases.base[i], "n_bases must be non-decreasing");
        if (bases.base[i + 1] - bases.base[i] > widest) widest = bases.base[i + 1] - bases.base[i];
    }
    GSB_REQUIRE(bases.base[0] == 0 && bases.base[world] == N_total, "n_bases must run from 0 to N_total");
    Fields fields = {{v_means2d, v_depths, v_conics, v_opacities, v_colors, nullptr}, {2, 1, 3, 1, D, 1}, 0,
                     {0, 0, 0, 0, 0, 0}};
    select_fields(fields, kFields - 1, fields_mask);
    dim3 grid(blocks_per_run(widest * (D > 3 ? D : 3)), (unsigned)(world * C_local * fields.count));
    put_bwd_kernel<<<grid, kThreads, 0, stream>>>(peers, fields, bases, world, rank, C_local, N_total, region_offset,
                                                 (unsigned long long)epoch, channel);
    return gsb_check_launch("a2a_put_bwd");
}

extern "C" int gsb_a2a_put_bwd(int world, int rank, int C_local, int64_t N_total, const int64_t *n_bases_host, int D,
                               const float *v_means2d, const float *v_depths, const float *v_conics,
                               const float *v_opacities, const float *v_colors,
                               const void *const *peer_bases_host, int64_t region_offset, int64_t epoch,
                               cudaStream_t stream) {
    return gsb_a2a_put_bwd_part(world, rank, C_local, N_total, n_bases_host, D, v_means2d, v_depths, v_conics,
                                v_opacities, v_colors, (1 << (kFields - 1)) - 1, 1, peer_bases_host, region_offset,
                                epoch, stream);
}

extern "C" int gsb_a2a_wait_copy_part(int world, int channel, int64_t epoch, const void *local_base,
                                      int64_t region_offset, int64_t begin0, int64_t end0, int64_t begin1,
                                      int64_t end1, float *out, int *timeout_flag, cudaStream_t stream) {
    GSB_REQUIRE(world >= 1 && world <= kMaxPeers && channel >= 0 && channel < kChannels, "bad world/channel");
    GSB_REQUIRE(begin0 >= 0 && end0 >= begin0 && begin1 >= 0 && end1 >= begin1 && local_base != nullptr, "bad ranges");
    GSB_REQUIRE(((begin0 | end0 | begin1 | end1) & 3) == 0, "ranges must be multiples of 4 floats");
    GSB_REQUIRE(region_offset >= kHeaderBytes && (region_offset & 15) == 0 && epoch > 0, "bad region/epoch");
    GSB_REQUIRE(((uintptr_t)out & 15) == 0, "output buffer must be 16-byte aligned");
    const int64_t n_floats = (end0 - begin0) + (end1 - begin1);
    int64_t b = (n_floats / 4 + kThreads - 1) / kThreads;
    unsigned grid = (unsigned)(b < 1 ? 1 : (b > 148 * 8 ? 148 * 8 : b));
    wait_copy_kernel<<<grid, kThreads, 0, stream>>>(reinterpret_cast<const char *>(local_base), world, channel,
                                                   (unsigned long long)epoch, region_offset, begin0, end0, begin1, end1,
                                                   out, timeout_flag);
    return gsb_check_launch("a2a_wait_copy");
}

extern "C" int gsb_a2a_wait_copy(int world, int dir, int64_t epoch, const void *local_base, int64_t region_offset,
                                 int64_t n_floats, float *out, int *timeout_flag, cudaStream_t stream) {
    GSB_REQUIRE((dir == 0 || dir == 1) && n_floats >= 0 && (n_floats & 3) == 0, "bad dir/size");
    return gsb_a2a_wait_copy_part(world, dir, epoch, local_base, region_offset, 0, n_floats, 0, 0, out, timeout_flag,
                                  stream);
}

namespace {
int fill_plan(PackedPlan &plan, int world, const int64_t *counts, const int64_t *dst_base, const int64_t *dst_rows,
              const int64_t *cam_base, int64_t gauss_base, int64_t &widest) {
    int64_t at = 0;
    widest = 0;
    for (int i = 0; i < kMaxPeers; ++i) {
        const bool in = i < world;
        plan.src_start[i] = at;
        plan.count[i] = in ? counts[i] : 0;
        plan.dst_base[i] = in ? dst_base[i] : 0;
        plan.dst_rows[i] = in ? dst_rows[i] : 0;
        plan.cam_base[i] = (in && cam_base) ? cam_base[i] : 0;
        if (in && (counts[i] < 0 || dst_base[i] < 0 || dst_base[i] + counts[i] > dst_rows[i])) return -1;
        at += plan.count[i];
        if (plan.count[i] > widest) widest = plan.count[i];
    }
    plan.gauss_base = gauss_base;
    return 0;
}
}  // namespace

extern "C" int gsb_a2a_put_packed_fwd(int world, int rank, int D, const int64_t *send_counts_host,
                                      const int64_t *dst_row_base_host, const int64_t *dst_rows_host,
                                      const int64_t *cam_base_host, int64_t gauss_base, const float *means2d,
                                      const float *depths, const float *conics, const float *opacities,
                                      const float *colors, const int32_t *radii, const int64_t *camera_ids,
                                      const int64_t *gaussian_ids, const void *const *peer_bases_host,
                                      int64_t region_offset, int64_t epoch, cudaStream_t stream) {
    GSB_REQUIRE(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world && D >= 1, "bad world/rank");
    GSB_REQUIRE(send_counts_host && dst_row_base_host && dst_rows_host && cam_base_host, "null plan");
    GSB_REQUIRE(region_offset >= kHeaderBytes && epoch > 0, "bad region/epoch");
    PeerTable peers;
    GSB_REQUIRE(fill_peers(peers, peer_bases_host, world) == 0, "null peer mailbox");
    PackedPlan plan;
    int64_t widest = 0;
    GSB_REQUIRE(fill_plan(plan, world, send_counts_host, dst_row_base_host, dst_rows_host, cam_base_host, gauss_base,
                          widest) == 0, "inconsistent row plan");
    PackedFields fields = {{means2d, depths, conics, opacities, colors, radii, camera_ids, gaussian_ids},
                           {2, 1, 3, 1, D, 1, 1, 1}};
    dim3 grid(blocks_per_run(widest * (D > 3 ? D : 3)), (unsigned)(world * kPackedFields));
    put_packed_fwd_kernel<<<grid, kThreads, 0, stream>>>(peers, fields, plan, world, rank, region_offset,
                                                        (unsigned long long)epoch);
    return gsb_check_launch("a2a_put_packed_fwd");
}

extern "C" int gsb_a2a_put_packed_bwd(int world, int rank, int D, const int64_t *recv_counts_host,
                                      const int64_t *dst_row_base_host, const int64_t *dst_rows_host,
                                      const float *v_means2d, const float *v_depths, const float *v_conics,
                                      const float *v_opacities, const float *v_colors,
                                      const void *const *peer_bases_host, int64_t region_offset, int64_t epoch,
                                      cudaStream_t stream) {
    GSB_REQUIRE(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world && D >= 1, "bad world/rank");
    GSB_REQUIRE(recv_counts_host && dst_row_base_host && dst_rows_host, "null plan");
    GSB_REQUIRE(region_offset >= kHeaderBytes && epoch > 0, "bad region/epoch");
    PeerTable peers;
    GSB_REQUIRE(fill_peers(peers, peer_bases_host, world) == 0, "null peer mailbox");
    PackedPlan plan;
    int64_t widest = 0;
    GSB_REQUIRE(fill_plan(plan, world, recv_counts_host, dst_row_base_host, dst_rows_host, nullptr, 0, widest) == 0,
                "inconsistent row plan");
    PackedFields fields = {{v_means2d, v_depths, v_conics, v_opacities, v_colors, nullptr, nullptr, nullptr},
                           {2, 1, 3, 1, D, 1, 1, 1}};
    dim3 grid(blocks_per_run(widest * (D > 3 ? D : 3)), (unsigned)(world * (kFields - 1)));
    put_packed_bwd_kernel<<<grid, kThreads, 0, stream>>>(peers, fields, plan, world, rank, region_offset,
                                                        (unsigned long long)epoch);
    return gsb_check_launch("a2a_put_packed_bwd");
}

// ---- push exchange: the projection / colour kernels store their rows into the peers' mailboxes themselves ---------
int gsb_fill_push_rows(peer::PushRows &push, int world, int C_local, int64_t N_total, int64_t n_base, int D,
                       const void *const *peer_bases_host, int64_t region_offset, const float *opacities) {
    GSB_REQUIRE(world >= 1 && world <= kMaxPeers && peer_bases_host != nullptr, "bad world / peers");
    GSB_REQUIRE(region_offset >= kHeaderBytes && (region_offset & 15) == 0, "bad region");
    for (int i = 0; i < peer::kMaxPeers; ++i) push.region[i] = nullptr;
    for (int i = 0; i < world; ++i) {
        GSB_REQUIRE(peer_bases_host[i] != nullptr, "null peer mailbox");
        push.region[i] = reinterpret_cast<float *>(reinterpret_cast<char *>(const_cast<void *>(peer_bases_host[i])) +
                                                   region_offset);
    }
    const int width[kFields] = {2, 1, 3, 1, D, 1};
    const int64_t rows = (int64_t)C_local * N_total;
    push.world = world;
    push.C_local = C_local;
    push.N_total = N_total;
    push.n_base = n_base;
    push.off_means2d = field_offset(width, 0, rows);
    push.off_depths = field_offset(width, 1, rows);
    push.off_conics = field_offset(width, 2, rows);
    push.off_opacities = field_offset(width, 3, rows);
    push.off_colors = field_offset(width, 4, rows);
    push.off_radii = field_offset(width, 5, rows);
    push.opacities = opacities;
    return GSB_OK;
}

namespace {
// after the pushing kernels (stream order): tell every peer that this rank's rows of `epoch` have landed
__global__ void publish_kernel(PeerTable peers, int world, int rank, int channel, unsigned long long epoch) {
    if (threadIdx.x < world) {
        __threadfence_system();
        unsigned long long *flag =
            reinterpret_cast<unsigned long long *>(peers.base[threadIdx.x] + 128 * channel) + rank;
        st_release_sys(flag, epoch);
    }
}
}  // namespace

extern "C" int gsb_a2a_publish(int world, int rank, int channel, int64_t epoch, const void *const *peer_bases_host,
                               cudaStream_t stream) {
    GSB_REQUIRE(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world, "bad world/rank");
    GSB_REQUIRE(channel >= 0 && channel < kChannels && epoch > 0, "bad channel/epoch");
    PeerTable peers;
    GSB_REQUIRE(fill_peers(peers, peer_bases_host, world) == 0, "null peer mailbox");
    publish_kernel<<<1, 32, 0, stream>>>(peers, world, rank, channel, (unsigned long long)epoch);
    return gsb_check_launch("a2a_publish");
}
