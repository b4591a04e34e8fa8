// p2pt.cu -- the C ABI of libp2ptunnel_b200.so (include/p2pt.h): context, staging, launches.
//
// No CPU fallback lives here: every entry point either launches the sm_100a kernels or fails with
// an error code.  Host-pointer entry points stage through a device arena and a small ring of streams
// so that host->device copies, kernels and device->host copies of successive chunks overlap (PCIe
// Gen5 is what bounds these calls).  A context owns kLanes such staging sets ("call lanes"): a call
// takes a free lane for its duration, so several threads can be inside host-pointer calls on ONE
// context at the same time - the reference encodes and sends from many tokio tasks over one
// connection (/root/reference/tunnel/src/main.rs:18, serve.rs:131).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "../../include/p2pt.h"
#include "frame_kernels.cuh"
#include "demux_direct_kernels.cuh"
#include "gcm_kernels.cuh"
#include "sctp_kernels.cuh"
#include "dtls_state_kernels.cuh"
#include "sse_kernels.cuh"

using namespace p2pt;

static_assert(sizeof(p2pt_desc) == 24, "p2pt_desc must be 24 bytes");

namespace {

constexpr int kStreams = 4;

struct Arena {
    uint8_t* base = nullptr;
    size_t cap = 0;
    size_t used = 0;
};

}  // namespace

// One staging set of a context: streams, events, a device arena that grows to the largest call it has seen,
// and a mapped pinned buffer for small calls.  Owned by exactly one host-pointer call at a time.
struct Lane {
    cudaStream_t streams[kStreams] = {};
    cudaEvent_t events[kStreams] = {};
    cudaEvent_t sync_events[kStreams] = {};   // cudaEventBlockingSync: the host thread sleeps instead of spinning
    Arena arena;
    bool spin_this_call = false;              // set per call: small calls spin instead of sleeping
    // Small host-pointer calls (the reference's per-frame encode / decode / seal / open call sites, batch of
    // one or a few): inputs are staged into a mapped pinned buffer the kernel reads over PCIe and results
    // come back the same way, so the call is one launch + one stream sync instead of up to eight copies.
    uint8_t* zc_host = nullptr;    // host address of the mapped buffer
    uint8_t* zc_dev = nullptr;     // device alias
    size_t zc_cap = 0;
    struct PendingOut { void* dst; const void* src; size_t bytes; };
    std::vector<PendingOut> zc_pending;
    bool busy = false;
};

constexpr int kLanes = 8;

// A stream beside the caller's, with the two events that fork it off and join it back: the direct demux puts its
// fallback (the sorting path's kernels, which only look at a flag when the direct path succeeded) there, so that
// their launches overlap the direct path's gather instead of following it (seven empty launches: about 15 us).
struct SideStream {
    cudaStream_t st = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
    std::mutex mu;      // record + wait of one event as a pair: two calls that were handed the same side stream must not
                        // wait on each other's record
};
constexpr int kSides = 8;

struct p2pt_ctx {
    int device = 0;
    int sm_count = 148;
    Lane lanes[kLanes];
    SideStream sides[kSides];
    std::atomic<uint32_t> side_next{0};
    std::mutex pool_mu;            // guards Lane::busy
    std::condition_variable pool_cv;
    bool blocking_sync = true;                // big host-pointer calls wait on sync_events (see wait_stream)
    bool small_demux = true;                  // batches of up to 1024 frames take the single-block demux kernel
    bool coop_sort = true;                    // demux sorting passes in one cooperative launch
    bool direct_demux = true;                 // demux calls with max_streams <= 8192 try the sort-free path first
    std::mutex err_mu;
    std::string err;
    SessionKey* keys_local = nullptr;
    SessionKey* keys_remote = nullptr;
    uint32_t n_sessions = 0;
    std::atomic<uint64_t> launches{0};
    uint64_t chunk_bytes = 16u << 20;  // host-pointer calls move records in chunks of about this size
    uint32_t lut_slack = 0;        // bytes needed in front of the T-tables to reach 64 KiB alignment
    int gcm_g = 0;                 // forced group width (env P2PT_GCM_G), 0 = heuristic
    int occ_encode = 4, occ_gather = 4, occ_dgather = 4;   // resident blocks per SM of the copy kernels (set at create)
    size_t zc_cap = 0;             // size of every lane's mapped buffer (0: no mapped memory here, small-call path off)
    uint64_t small_call_bytes = 256u << 10;  // staging footprint up to which a call takes the mapped-buffer path (0 = never)
};

// Takes a free lane of the context (lowest index first, so a single caller keeps reusing lane 0 and its
// arena); blocks while all kLanes are in use.  If the call leaves early with an error, copies and kernels
// it enqueued may still be running into the caller's buffers and the lane's arena: the guard waits for the
// lane's streams before it gives the lane back, unless the call finished normally (dismiss()).
struct LaneGuard {
    p2pt_ctx* ctx;
    Lane* lane = nullptr;
    bool clean = false;
    explicit LaneGuard(p2pt_ctx* c) : ctx(c) {
        std::unique_lock<std::mutex> lk(c->pool_mu);
        for (;;) {
            for (int i = 0; i < kLanes; i++)
                if (!c->lanes[i].busy) { lane = &c->lanes[i]; break; }
            if (lane) break;
            c->pool_cv.wait(lk);
        }
        lane->busy = true;
    }
    void dismiss() { clean = true; }
    ~LaneGuard() {
        if (!clean)
            for (int i = 0; i < kStreams; i++) cudaStreamSynchronize(lane->streams[i]);
        {
            std::lock_guard<std::mutex> lk(ctx->pool_mu);
            lane->busy = false;
        }
        ctx->pool_cv.notify_one();
    }
};

static std::string g_create_err;
static std::mutex g_create_mu;

static int32_t set_err(p2pt_ctx* ctx, int32_t code, const std::string& msg) {
    if (ctx) {
        std::lock_guard<std::mutex> g(ctx->err_mu);
        ctx->err = msg;
    } else {
        std::lock_guard<std::mutex> g(g_create_mu);
        g_create_err = msg;
    }
    return code;
}

#define CUDA_TRY(ctx, expr)                                                                     \
    do {                                                                                        \
        cudaError_t e__ = (expr);                                                               \
        if (e__ != cudaSuccess)                                                                 \
            return set_err(ctx, P2PT_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__)); \
    } while (0)

// One NVTX range per host-pointer call, named after the entry point: what the reference's `tracing` spans are to
// its tokio tasks (/root/reference/tunnel/src/main.rs:21-25) these are to a timeline profiler.  nvtx3 is
// header-only and costs a few nanoseconds when no tool is attached.
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// ---- Te0 table, computed from the S-box definition (FIPS-197 5.1.1) on the host ---------------------
static uint8_t gf8_mul_h(uint8_t a, uint8_t b) {
    uint8_t p = 0;
    while (b) {
        if (b & 1) p ^= a;
        a = (uint8_t)((a << 1) ^ ((a & 0x80) ? 0x1b : 0));
        b >>= 1;
    }
    return p;
}
static void build_te0(uint32_t te0[256]) {
    // inverse via a^254, then the affine map
    for (int x = 0; x < 256; x++) {
        uint8_t inv = 0;
        if (x) {
            uint8_t r = 1, b = (uint8_t)x;
            int e = 254;
            while (e) { if (e & 1) r = gf8_mul_h(r, b); b = gf8_mul_h(b, b); e >>= 1; }
            inv = r;
        }
        uint8_t s = inv ^ (uint8_t)((inv << 1) | (inv >> 7)) ^ (uint8_t)((inv << 2) | (inv >> 6)) ^
                    (uint8_t)((inv << 3) | (inv >> 5)) ^ (uint8_t)((inv << 4) | (inv >> 4)) ^ 0x63;
        uint8_t s2 = gf8_mul_h(s, 2), s3 = (uint8_t)(s2 ^ s);
        te0[x] = (uint32_t)s2 | ((uint32_t)s << 8) | ((uint32_t)s << 16) | ((uint32_t)s3 << 24);
    }
}

// CRC32c slice-by-4 tables and x^(8k) mod P, from the polynomial (RFC 3720 12.1), reflected form
static void build_crc_tables(uint32_t tab[4][256], uint32_t xpow[1280]) {
    for (uint32_t i = 0; i < 256; i++) {
        uint32_t c = i;
        for (int b = 0; b < 8; b++) c = (c >> 1) ^ (0x82F63B78u & (0u - (c & 1u)));
        tab[0][i] = c;
    }
    for (int k = 1; k < 4; k++)
        for (uint32_t i = 0; i < 256; i++) tab[k][i] = (tab[k - 1][i] >> 8) ^ tab[0][tab[k - 1][i] & 0xff];
    xpow[0] = 0x80000000u;                                   // the polynomial "1"
    for (int k = 1; k < 1280; k++) xpow[k] = (xpow[k - 1] >> 8) ^ tab[0][xpow[k - 1] & 0xff];
}

__global__ void smem_base_probe_kernel(uint32_t* out) {
    extern __shared__ __align__(16) uint8_t dyn_smem[];
    *out = shared_addr(dyn_smem);
}

// Wait for one of the ctx's streams.  cudaStreamSynchronize spins on a CPU core; with several worker
// threads per GPU and several GPUs per host that starves the threads that feed the copy engines, so by
// default the wait is a blocking-sync event (the thread sleeps until the interrupt).
static cudaError_t wait_stream(p2pt_ctx* ctx, Lane* L, cudaStream_t st) {
    // a sleeping wait costs a wake-up of 50-100 us; a call that stages a few MB has up to seven of them
    if (!ctx->blocking_sync || L->spin_this_call) return cudaStreamSynchronize(st);
    int idx = 0;
    for (int i = 0; i < kStreams; i++) if (L->streams[i] == st) idx = i;
    cudaError_t e = cudaEventRecord(L->sync_events[idx], st);
    if (e != cudaSuccess) return e;
    return cudaEventSynchronize(L->sync_events[idx]);
}

// host-pointer calls that stage less than this wait by spinning: they finish in well under a millisecond
constexpr size_t kSpinWaitBytes = 8u << 20;

// ---- arena -----------------------------------------------------------------------------------------
static int32_t arena_reserve(p2pt_ctx* ctx, Lane* L, size_t bytes) {
    L->arena.used = 0;
    L->spin_this_call = bytes < kSpinWaitBytes;
    if (bytes <= L->arena.cap) return P2PT_OK;
    if (L->arena.base) cudaFree(L->arena.base);
    L->arena.base = nullptr;
    L->arena.cap = 0;
    size_t want = bytes + bytes / 8 + (1u << 20);
    CUDA_TRY(ctx, cudaMalloc(&L->arena.base, want));
    L->arena.cap = want;
    return P2PT_OK;
}
template <class T>
static T* arena_take(Lane* L, size_t count) {
    size_t bytes = (count * sizeof(T) + 255) & ~(size_t)255;
    T* p = reinterpret_cast<T*>(L->arena.base + L->arena.used);
    L->arena.used += bytes;
    return p;
}
static size_t pad256(size_t b) { return (b + 255) & ~(size_t)255; }

// Carves device-visible pieces out of the ctx's mapped pinned buffer and maps them back to host addresses
// (the small-batch paths of the demux calls; the caller owns the lane and has checked the footprint).
struct ZcCarver {
    Lane* L;
    size_t used = 0;
    uint8_t* take(size_t bytes) { uint8_t* p = L->zc_dev + used; used += pad256(bytes); return p; }
    uint8_t* host_of(const void* d) const { return L->zc_host + ((const uint8_t*)d - L->zc_dev); }
};

// Staging for the inputs and outputs of one host-pointer call: the device arena + async copies, or,
// when everything fits small_call_bytes, the mapped pinned buffer + plain memcpy (see p2pt_ctx).
// The caller owns the lane.
struct IoStage {
    p2pt_ctx* ctx;
    Lane* L;
    bool zc = false;
    size_t zc_used = 0;
    int32_t reserve(size_t bytes) {
        zc = L->zc_host && ctx->small_call_bytes && bytes <= ctx->small_call_bytes && bytes <= L->zc_cap;
        L->zc_pending.clear();
        return zc ? P2PT_OK : arena_reserve(ctx, L, bytes);
    }
    template <class T> T* take(size_t count) {
        if (!zc) return arena_take<T>(L, count);
        T* p = reinterpret_cast<T*>(L->zc_dev + zc_used);
        zc_used += pad256(count * sizeof(T));
        return p;
    }
    cudaError_t in(void* d, const void* h, size_t bytes, cudaStream_t st) {
        if (!zc) return cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, st);
        memcpy(L->zc_host + ((uint8_t*)d - L->zc_dev), h, bytes);
        return cudaSuccess;
    }
    cudaError_t out(void* h, const void* d, size_t bytes, cudaStream_t st) {
        if (!zc) return cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, st);
        L->zc_pending.push_back({h, L->zc_host + ((const uint8_t*)d - L->zc_dev), bytes});
        return cudaSuccess;
    }
    // wait for the call's work on st and deliver the staged results
    cudaError_t finish(cudaStream_t st) {
        if (!zc) return wait_stream(ctx, L, st);
        cudaError_t e = cudaStreamSynchronize(st);        // a spin wait: these calls last microseconds
        if (e != cudaSuccess) return e;
        for (const auto& o : L->zc_pending) memcpy(o.dst, o.src, o.bytes);
        L->zc_pending.clear();
        return cudaSuccess;
    }
};

// ---- launch helpers --------------------------------------------------------------------------------
static inline void count_launch(p2pt_ctx* ctx, int k = 1) { ctx->launches.fetch_add((uint64_t)k, std::memory_order_relaxed); }

static int pick_group(p2pt_ctx* ctx, uint64_t n, uint64_t body_bytes) {
    (void)body_bytes;
    if (ctx->gcm_g) return ctx->gcm_g;
    // A full machine wants G = 4 (measured on B200 at 1200-byte records: G=4 323 GB/s, 8: 313, 16: 290, 32: 213).
    // A batch too small to fill the 148 x 1024 lanes is latency-bound on one record's serial chain of slots
    // (a 1200-byte record on 4 lanes: 20 rounds, ~40 us), so it gets the widest group that still fits.
    const uint64_t lanes = (uint64_t)ctx->sm_count * kGcmThreads;
    if (n * 32 <= lanes) return 32;
    if (n * 16 <= lanes) return 16;
    if (n * 8 <= lanes) return 8;
    return 4;
}

template <int MODE>
static int32_t launch_gcm(p2pt_ctx* ctx, const GcmArgs& a_in, int G, cudaStream_t st) {
    if (a_in.n == 0) return P2PT_OK;
    GcmArgs a = a_in;
    a.m8 = 1u << 8; a.m16 = 1u << 16; a.m24 = 1u << 24;
    const uint32_t smem = gcm_smem_bytes(G, ctx->lut_slack);
    const uint64_t groups_per_cta = kGcmThreads / G;
    uint64_t grid = (a.n + groups_per_cta - 1) / groups_per_cta;
    if (grid > (uint64_t)ctx->sm_count) grid = ctx->sm_count;      // persistent: one CTA per SM
#define P2PT_LAUNCH_G(GG)                                                                              \
    case GG:                                                                                           \
        gcm_record_kernel<GG, MODE><<<(unsigned)grid, kGcmThreads, smem, st>>>(a);                     \
        break;
    switch (G) {
        P2PT_LAUNCH_G(4)
        P2PT_LAUNCH_G(8)
        P2PT_LAUNCH_G(16)
        P2PT_LAUNCH_G(32)
        default: return set_err(ctx, P2PT_ERR_ARG, "group width must be 4, 8, 16 or 32");
    }
#undef P2PT_LAUNCH_G
    count_launch(ctx);
    CUDA_TRY(ctx, cudaGetLastError());
    return P2PT_OK;
}

static unsigned grid_for(p2pt_ctx* ctx, uint64_t items, unsigned per_block, unsigned waves) {
    uint64_t g = (items + per_block - 1) / per_block;
    uint64_t cap = (uint64_t)ctx->sm_count * waves;
    if (g > cap) g = cap;
    if (g == 0) g = 1;
    return (unsigned)g;
}

// Grid for a warp-per-tile grid-stride kernel: at most one resident wave (blocks_per_sm from the
// occupancy query) and as many warps as give every warp the same number of tiles (within one), so that
// no SM idles while the last blocks finish.
static unsigned grid_balanced(p2pt_ctx* ctx, uint64_t tiles, unsigned warps_per_block, unsigned blocks_per_sm) {
    if (tiles == 0) return 1;
    const uint64_t cap_warps = (uint64_t)ctx->sm_count * blocks_per_sm * warps_per_block;
    const uint64_t iters = (tiles + cap_warps - 1) / cap_warps;
    const uint64_t warps = (tiles + iters - 1) / iters;
    return (unsigned)((warps + warps_per_block - 1) / warps_per_block);
}

// Tile size of the two copy kernels (frame_encode_kernel, demux_gather_kernel) in rounds of 8 frames, and
// their grid: 32-frame tiles on one balanced resident wave when every warp still gets >= 16 of them,
// else 8-frame tiles on a grid the hardware scheduler balances.
static int copy_tile_rounds(p2pt_ctx* ctx, uint64_t n, int blocks_per_sm) {
    const uint64_t wave_warps = (uint64_t)ctx->sm_count * blocks_per_sm * 8;
    return (n + 31) / 32 >= 16 * wave_warps ? 4 : 1;
}
static unsigned copy_grid(p2pt_ctx* ctx, uint64_t n, int rounds, int blocks_per_sm) {
    if (rounds == 4) return grid_balanced(ctx, (n + 31) / 32, 8, blocks_per_sm);
    return grid_for(ctx, (n + 7) / 8, 8, 16);
}

// ===================================================================================================
extern "C" {

int32_t p2pt_abi_version(void) { return P2PT_ABI_VERSION; }

const char* p2pt_last_error(p2pt_ctx* ctx) {
    if (ctx) {
        std::lock_guard<std::mutex> g(ctx->err_mu);
        static thread_local std::string copy;
        copy = ctx->err;
        return copy.c_str();
    }
    std::lock_guard<std::mutex> g(g_create_mu);
    static thread_local std::string copy2;
    copy2 = g_create_err;
    return copy2.c_str();
}

}  // extern "C" (reopened below)

// Device-wide tables (g_te0, the CRC32c tables) are uploaded once per device, not once per context: kernels of
// other contexts on the device may be reading them while a new context is being created.
namespace {
constexpr int kMaxDevices = 64;
std::mutex g_tables_mu;
bool g_tables_ready[kMaxDevices] = {};
}  // namespace

static int32_t upload_device_tables(int device) {
    std::lock_guard<std::mutex> lk(g_tables_mu);
    if (g_tables_ready[device]) return P2PT_OK;
    static uint32_t te0[256], ctab[4][256], cxp[1280], cinit[kCrcSeg * 32 + 1];
    static bool built = false;
    if (!built) {
        build_te0(te0);
        build_crc_tables(ctab, cxp);
        // contribution of the 0xFFFFFFFF init value to a packet of w words: the register 0xFFFFFFFF run
        // through 4 w zero bytes (CRC is linear in its initial register)
        uint32_t reg = 0xFFFFFFFFu;
        for (int w = 0; w <= kCrcSeg * 32; w++) {
            cinit[w] = reg;
            for (int b = 0; b < 4; b++) reg = (reg >> 8) ^ ctab[0][reg & 0xff];
        }
        built = true;
    }
    CUDA_TRY(nullptr, cudaMemcpyToSymbol(g_te0, te0, sizeof te0));
    CUDA_TRY(nullptr, cudaMemcpyToSymbol(g_crc_tab, ctab, sizeof ctab));
    CUDA_TRY(nullptr, cudaMemcpyToSymbol(g_crc_xpow, cxp, sizeof cxp));
    CUDA_TRY(nullptr, cudaMemcpyToSymbol(g_crc_init, cinit, sizeof cinit));
    CUDA_TRY(nullptr, cudaDeviceSynchronize());
    g_tables_ready[device] = true;
    return P2PT_OK;
}

// the record and SCTP kernels need more dynamic shared memory than the default limit; set once per context
// (the attribute is per device and function, setting it again is harmless)
template <int MODE>
static cudaError_t configure_gcm_mode(uint32_t slack) {
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(gcm_record_kernel<4, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gcm_smem_bytes(4, slack))) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(gcm_record_kernel<8, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gcm_smem_bytes(8, slack))) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(gcm_record_kernel<16, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gcm_smem_bytes(16, slack))) != cudaSuccess) return e;
    return cudaFuncSetAttribute(gcm_record_kernel<32, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gcm_smem_bytes(32, slack));
}

static int32_t ctx_init(p2pt_ctx* ctx, int device, const cudaDeviceProp& prop) {
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    for (int l = 0; l < kLanes; l++) {
        Lane& L = ctx->lanes[l];
        for (int i = 0; i < kStreams; i++) {
            CUDA_TRY(nullptr, cudaStreamCreateWithFlags(&L.streams[i], cudaStreamNonBlocking));
            CUDA_TRY(nullptr, cudaEventCreateWithFlags(&L.events[i], cudaEventDisableTiming));
            CUDA_TRY(nullptr, cudaEventCreateWithFlags(&L.sync_events[i], cudaEventDisableTiming | cudaEventBlockingSync));
        }
    }
    for (int i = 0; i < kSides; i++) {
        CUDA_TRY(nullptr, cudaStreamCreateWithFlags(&ctx->sides[i].st, cudaStreamNonBlocking));
        CUDA_TRY(nullptr, cudaEventCreateWithFlags(&ctx->sides[i].fork, cudaEventDisableTiming));
        CUDA_TRY(nullptr, cudaEventCreateWithFlags(&ctx->sides[i].join, cudaEventDisableTiming));
    }
    if (int32_t r = upload_device_tables(device)) return r;
    // where does dynamic shared memory start?  The T-tables must sit on a 64 KiB boundary.
    uint32_t* d_probe = nullptr;
    CUDA_TRY(nullptr, cudaMalloc(&d_probe, 4));
    cudaStream_t s0 = ctx->lanes[0].streams[0];
    smem_base_probe_kernel<<<1, 32, 16, s0>>>(d_probe);
    uint32_t sb = 0;
    cudaError_t e = cudaMemcpyAsync(&sb, d_probe, 4, cudaMemcpyDeviceToHost, s0);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s0);
    cudaFree(d_probe);
    CUDA_TRY(nullptr, e);
    ctx->lut_slack = (0x10000u - (sb & 0xffffu)) & 0xffffu;
    CUDA_TRY(nullptr, configure_gcm_mode<kSeal>(ctx->lut_slack));
    CUDA_TRY(nullptr, configure_gcm_mode<kOpen>(ctx->lut_slack));
    CUDA_TRY(nullptr, configure_gcm_mode<kPipeSeal>(ctx->lut_slack));
    CUDA_TRY(nullptr, configure_gcm_mode<kPipeOpen>(ctx->lut_slack));
    CUDA_TRY(nullptr, cudaFuncSetAttribute(sctp_packetize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sctp_packetize_smem_bytes(ctx->lut_slack)));
    CUDA_TRY(nullptr, cudaFuncSetAttribute(sctp_unpack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sctp_unpack_smem_bytes(ctx->lut_slack)));
    CUDA_TRY(nullptr, cudaFuncSetAttribute(demux_direct_rank_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * kDirectIds));
    CUDA_TRY(nullptr, cudaFuncSetAttribute(demux_direct_rank_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 8 * kDirectIds));
    {
        const size_t cap = 1u << 20;
        bool all = true;
        for (int l = 0; l < kLanes && all; l++) {
            void* h = nullptr; void* d = nullptr;
            if (cudaHostAlloc(&h, cap, cudaHostAllocMapped) != cudaSuccess) { all = false; break; }
            if (cudaHostGetDevicePointer(&d, h, 0) != cudaSuccess) { cudaFreeHost(h); all = false; break; }
            ctx->lanes[l].zc_host = (uint8_t*)h; ctx->lanes[l].zc_dev = (uint8_t*)d; ctx->lanes[l].zc_cap = cap;
        }
        (void)cudaGetLastError();      // without mapped memory the small-call path is simply off
        if (all) ctx->zc_cap = cap;
        else
            for (int l = 0; l < kLanes; l++) {
                if (ctx->lanes[l].zc_host) cudaFreeHost(ctx->lanes[l].zc_host);
                ctx->lanes[l].zc_host = ctx->lanes[l].zc_dev = nullptr; ctx->lanes[l].zc_cap = 0;
            }
    }
    CUDA_TRY(nullptr, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx->occ_encode, frame_encode_kernel, 256, 0));
    CUDA_TRY(nullptr, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx->occ_gather, demux_gather_kernel, 256, 0));
    if (ctx->occ_encode < 1) ctx->occ_encode = 1;
    if (ctx->occ_gather < 1) ctx->occ_gather = 1;
    CUDA_TRY(nullptr, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctx->occ_dgather, demux_direct_gather_kernel<true>, 256, 0));
    if (ctx->occ_dgather < 1) ctx->occ_dgather = 1;
    if (const char* g = getenv("P2PT_GCM_G")) {
        int v = atoi(g);
        if (v == 4 || v == 8 || v == 16 || v == 32) ctx->gcm_g = v;
    }
    return P2PT_OK;
}

extern "C" {

int32_t p2pt_ctx_create(int device, p2pt_ctx** out) {
    if (!out) return set_err(nullptr, P2PT_ERR_ARG, "out is null");
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return set_err(nullptr, P2PT_ERR_CUDA, std::string("no CUDA device: ") + cudaGetErrorString(e) +
                                                   " (libp2ptunnel_b200 has no CPU fallback)");
    if (device < 0 || device >= ndev || device >= kMaxDevices) return set_err(nullptr, P2PT_ERR_ARG, "device index out of range");
    DeviceGuard dg(device);
    cudaDeviceProp prop;
    CUDA_TRY(nullptr, cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return set_err(nullptr, P2PT_ERR_CUDA, "libp2ptunnel_b200 is built for sm_100a only");
    p2pt_ctx* ctx = new p2pt_ctx();
    const int32_t r = ctx_init(ctx, device, prop);
    if (r != P2PT_OK) {                // whatever was created so far goes away with the context
        ctx->device = device;
        p2pt_ctx_destroy(ctx);
        return r;
    }
    *out = ctx;
    return P2PT_OK;
}

int32_t p2pt_ctx_destroy(p2pt_ctx* ctx) {
    if (!ctx) return P2PT_OK;
    {
        DeviceGuard dg(ctx->device);
        cudaDeviceSynchronize();
        for (int l = 0; l < kLanes; l++) {
            Lane& L = ctx->lanes[l];
            for (int i = 0; i < kStreams; i++) {
                if (L.streams[i]) cudaStreamDestroy(L.streams[i]);
                if (L.events[i]) cudaEventDestroy(L.events[i]);
                if (L.sync_events[i]) cudaEventDestroy(L.sync_events[i]);
            }
            if (L.arena.base) cudaFree(L.arena.base);
            if (L.zc_host) cudaFreeHost(L.zc_host);
        }
        for (int i = 0; i < kSides; i++) {
            if (ctx->sides[i].st) cudaStreamDestroy(ctx->sides[i].st);
            if (ctx->sides[i].fork) cudaEventDestroy(ctx->sides[i].fork);
            if (ctx->sides[i].join) cudaEventDestroy(ctx->sides[i].join);
        }
        if (ctx->keys_local) cudaFree(ctx->keys_local);
        if (ctx->keys_remote) cudaFree(ctx->keys_remote);
        (void)cudaGetLastError();
    }
    delete ctx;
    return P2PT_OK;
}

void* p2pt_host_alloc(size_t bytes) {
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) return nullptr;
    return p;
}
void p2pt_host_free(void* p) { if (p) cudaFreeHost(p); }

uint64_t p2pt_launch_count(p2pt_ctx* ctx) { return ctx ? ctx->launches.load() : 0; }

int64_t p2pt_ctx_get_info(p2pt_ctx* ctx, const char* name) {
    if (!ctx || !name) return -1;
    if (strcmp(name, "sm_count") == 0) return ctx->sm_count;
    if (strcmp(name, "lut_slack") == 0) return ctx->lut_slack;
    if (strcmp(name, "gcm_group_width") == 0) return ctx->gcm_g;
    if (strcmp(name, "n_sessions") == 0) return ctx->n_sessions;
    if (strcmp(name, "gcm_threads") == 0) return kGcmThreads;
    return -1;
}

int32_t p2pt_ctx_set_option(p2pt_ctx* ctx, const char* name, int64_t value) {
    if (!ctx || !name) return P2PT_ERR_ARG;
    if (strcmp(name, "gcm_group_width") == 0) {
        if (!(value == 0 || value == 4 || value == 8 || value == 16 || value == 32))
            return set_err(ctx, P2PT_ERR_ARG, "gcm_group_width must be 0, 4, 8, 16 or 32");
        ctx->gcm_g = (int)value;
        return P2PT_OK;
    }
    if (strcmp(name, "coop_sort") == 0) {
        ctx->coop_sort = value != 0;
        return P2PT_OK;
    }
    if (strcmp(name, "small_demux") == 0) {
        ctx->small_demux = value != 0;
        return P2PT_OK;
    }
    if (strcmp(name, "direct_demux") == 0) {
        ctx->direct_demux = value != 0;
        return P2PT_OK;
    }
    if (strcmp(name, "blocking_sync") == 0) {
        ctx->blocking_sync = value != 0;
        return P2PT_OK;
    }
    if (strcmp(name, "small_call_bytes") == 0) {
        if (value < 0 || (uint64_t)value > ctx->zc_cap) return set_err(ctx, P2PT_ERR_ARG, "small_call_bytes must be 0 .. 1048576");
        ctx->small_call_bytes = (uint64_t)value;
        return P2PT_OK;
    }
    if (strcmp(name, "chunk_bytes") == 0) {
        if (value < (1 << 16)) return set_err(ctx, P2PT_ERR_ARG, "chunk_bytes must be at least 65536");
        ctx->chunk_bytes = (uint64_t)value;
        return P2PT_OK;
    }
    return set_err(ctx, P2PT_ERR_ARG, std::string("unknown option: ") + name);
}

// ---------------------------------------------------------------------------------------------------
int32_t p2pt_sessions_load(p2pt_ctx* ctx, uint32_t n, const uint8_t* local_keys16, const uint8_t* local_ivs4,
                           const uint8_t* remote_keys16, const uint8_t* remote_ivs4) {
    if (!ctx) return P2PT_ERR_ARG;
    if (!n || !local_keys16 || !local_ivs4 || !remote_keys16 || !remote_ivs4)
        return set_err(ctx, P2PT_ERR_ARG, "sessions_load: null key material or n == 0");
    DeviceGuard dg(ctx->device);
    LaneGuard lg(ctx);
    NvtxRange nvtx_("p2pt_sessions_load");
    Lane* L = lg.lane;
    cudaStream_t st = L->streams[0];
    if (ctx->keys_local) { cudaFree(ctx->keys_local); ctx->keys_local = nullptr; }
    if (ctx->keys_remote) { cudaFree(ctx->keys_remote); ctx->keys_remote = nullptr; }
    ctx->n_sessions = 0;
    CUDA_TRY(ctx, cudaMalloc(&ctx->keys_local, (size_t)n * sizeof(SessionKey)));
    CUDA_TRY(ctx, cudaMalloc(&ctx->keys_remote, (size_t)n * sizeof(SessionKey)));
    uint8_t *dk = nullptr, *di = nullptr;
    CUDA_TRY(ctx, cudaMalloc(&dk, (size_t)n * 16));
    CUDA_TRY(ctx, cudaMalloc(&di, (size_t)n * 4));
    for (int dir = 0; dir < 2; dir++) {
        CUDA_TRY(ctx, cudaMemcpyAsync(dk, dir ? remote_keys16 : local_keys16, (size_t)n * 16, cudaMemcpyHostToDevice, st));
        CUDA_TRY(ctx, cudaMemcpyAsync(di, dir ? remote_ivs4 : local_ivs4, (size_t)n * 4, cudaMemcpyHostToDevice, st));
        session_expand_kernel<<<(n + 127) / 128, 128, 0, st>>>(n, dk, di, dir ? ctx->keys_remote : ctx->keys_local);
        count_launch(ctx);
        CUDA_TRY(ctx, cudaGetLastError());
        CUDA_TRY(ctx, wait_stream(ctx, L, st));
    }
    cudaFree(dk);
    cudaFree(di);
    ctx->n_sessions = n;
    lg.dismiss();
    return P2PT_OK;
}

// ---------------------------------------------------------------------------------------------------
// device-pointer entry points
// ---------------------------------------------------------------------------------------------------
int32_t p2pt_frame_encode_batch_device(p2pt_ctx* ctx, uint64_t n, const uint8_t* types, const uint32_t* sids,
                                       const uint8_t* payload, const uint64_t* pay_off,
                                       uint8_t* out, const uint64_t* out_off, void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!types || !sids || !pay_off || !out || !out_off) return set_err(ctx, P2PT_ERR_ARG, "frame_encode: null argument");
    DeviceGuard dg(ctx->device);
    const int rounds = copy_tile_rounds(ctx, n, ctx->occ_encode);
    frame_encode_kernel<<<copy_grid(ctx, n, rounds, ctx->occ_encode), 256, 0, (cudaStream_t)stream>>>(
        n, types, sids, payload, (const u64*)pay_off, out, (const u64*)out_off, rounds);
    count_launch(ctx);
    CUDA_TRY(ctx, cudaGetLastError());
    return P2PT_OK;
}

int32_t p2pt_frame_decode_batch_device(p2pt_ctx* ctx, uint64_t n, const uint8_t* in, const uint64_t* in_off,
                                       p2pt_desc* desc, void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!in_off || !desc) return set_err(ctx, P2PT_ERR_ARG, "frame_decode: null argument");
    DeviceGuard dg(ctx->device);
    frame_decode_kernel<<<grid_for(ctx, n, 256, 8), 256, 0, (cudaStream_t)stream>>>(n, in, (const u64*)in_off, desc, nullptr);
    count_launch(ctx);
    CUDA_TRY(ctx, cudaGetLastError());
    return P2PT_OK;
}

uint64_t p2pt_frame_demux_scratch_bytes(uint64_t n) { return demux_scratch_bytes(n ? n : 1); }

}  // extern "C"

// The demux behind every entry point.  first_only: see demux_load_quad (the SCTP reassembly drops retransmitted
// fragments this way).  Fused form (in_off != NULL): the frames themselves are the input, `desc` is then where
// descriptors go if anyone wants them (the caller's array, or scratch for the sorting path only).
namespace {
struct DemuxFused {
    const uint64_t* in_off = nullptr;   // frame offsets: decode and demux in one call
    p2pt_desc* desc_user = nullptr;     // descriptors the caller asked for (nullable)
    p2pt_desc* desc_work = nullptr;     // where the sorting path decodes to (desc_user, or scratch)
    uint32_t* perm_work = nullptr;      // the sorting path always writes perm: the caller's, or scratch
};
}  // namespace

static int32_t demux_device_impl(p2pt_ctx* ctx, uint64_t n, const p2pt_desc* desc, const uint32_t* session,
                                 const uint8_t* in,
                                 uint32_t max_streams, uint32_t* uniq_sid, uint32_t* uniq_session,
                                 uint64_t* csr, uint32_t* perm,
                                 uint32_t* k_out, uint8_t* gathered, uint64_t gathered_cap,
                                 uint64_t* gathered_off, void* scratch, uint64_t scratch_bytes, void* stream,
                                 bool first_only, const DemuxFused* fused = nullptr) {
    if (!ctx) return P2PT_ERR_ARG;
    if (!uniq_sid || !csr || (!perm && !fused) || !k_out) return set_err(ctx, P2PT_ERR_ARG, "frame_demux: null output");
    if (n > 0xffffffffull) return set_err(ctx, P2PT_ERR_ARG, "frame_demux: n must fit in 32 bits");
    if (gathered && (!gathered_off || !in)) return set_err(ctx, P2PT_ERR_ARG, "frame_demux: gather needs in and gathered_off");
    DeviceGuard dg(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {
        CUDA_TRY(ctx, cudaMemsetAsync(k_out, 0, 4, st));
        CUDA_TRY(ctx, cudaMemsetAsync(csr, 0, 8, st));
        if (gathered_off) CUDA_TRY(ctx, cudaMemsetAsync(gathered_off, 0, 8, st));
        return P2PT_OK;
    }
    if ((!desc && !fused) || !scratch) return set_err(ctx, P2PT_ERR_ARG, "frame_demux: null desc or scratch");
    if (scratch_bytes < demux_scratch_bytes(n)) return set_err(ctx, P2PT_ERR_CAPACITY, "frame_demux: scratch too small");
    if (((uintptr_t)scratch & 255) != 0) return set_err(ctx, P2PT_ERR_ARG, "frame_demux: scratch must be 256-byte aligned");
    DemuxScratch s = demux_carve(scratch, n);
    DemuxBufs bufs;
    bufs.keys[0] = s.keys[0]; bufs.keys[1] = s.keys[1]; bufs.idx[0] = s.idx[0]; bufs.idx[1] = s.idx[1];
    const bool small = n <= (uint64_t)kSmallDemux && ctx->small_demux;
    const int dst_mode = !gathered ? 0 : gathered_cap <= 0xffffffffull ? 1 : 2;
    uint32_t* perm_sort = perm ? perm : fused->perm_work;       // the sorting path's perm (never NULL)

    // ---- the direct path first (demux_direct_kernels.cuh): no sort when the caller expects few streams.  If the
    //      batch turns out not to fit, it raises run_flag and the sorting path below - enqueued either way - runs.
    const uint32_t* run_flag = nullptr;
    SideStream* side = nullptr;
    if (ctx->direct_demux && !first_only && !small && max_streams <= (uint32_t)kDirectIds) {
        const DirectScratch ds = direct_carve((uint8_t*)scratch + align256(demux_sort_scratch_bytes(n)), n, s);
        CUDA_TRY(ctx, cudaMemsetAsync(ds.ctl, 0, kDirectZeroBytes, st));
        DirectArgs a;
        a.n = n; a.per = direct_per_block(n);
        a.in = in; a.in_off = fused ? (const u64*)fused->in_off : nullptr;
        a.desc_in = desc; a.desc_out = fused ? fused->desc_user : nullptr;
        a.session = session; a.s = ds; a.want_rank = perm != nullptr;
        const unsigned nb = (unsigned)direct_blocks(n);
        if (fused) demux_direct_rank_kernel<true><<<nb, kDirectThreads, 8 * kDirectIds, st>>>(a);
        else demux_direct_rank_kernel<false><<<nb, kDirectThreads, 8 * kDirectIds, st>>>(a);
        demux_direct_streams_kernel<<<dim3(kDirectIds / kOrderT, 64), kOrderT, 0, st>>>(ds, (u64*)csr, k_out, (u64*)gathered_off);
        // the flag is final now: from here the sorting path (as fallback) runs on a side stream, beside prefix + gather
        side = &ctx->sides[ctx->side_next.fetch_add(1, std::memory_order_relaxed) % kSides];
        {
            std::lock_guard<std::mutex> lk(side->mu);
            CUDA_TRY(ctx, cudaEventRecord(side->fork, st));
            CUDA_TRY(ctx, cudaStreamWaitEvent(side->st, side->fork, 0));
        }
        demux_direct_prefix_kernel<<<kDirectIds / 32, 1024, 0, st>>>(ds, nb, max_streams, uniq_sid, uniq_session, (u64*)csr, (u64*)gathered_off);
        count_launch(ctx, 3);
        if (gathered || perm) {
            const int rounds = copy_tile_rounds(ctx, n, ctx->occ_dgather);
            const unsigned grid = copy_grid(ctx, n, rounds, ctx->occ_dgather);
            if (fused) demux_direct_gather_kernel<true><<<grid, 256, 0, st>>>(a, in, gathered, gathered_cap, perm, k_out, max_streams, rounds);
            else demux_direct_gather_kernel<false><<<grid, 256, 0, st>>>(a, in, gathered, gathered_cap, perm, k_out, max_streams, rounds);
            count_launch(ctx);
        }
        CUDA_TRY(ctx, cudaGetLastError());
        run_flag = ds.ctl;
    }
    cudaStream_t user_st = st;
    if (side) st = side->st;                               // everything below is the fallback then
    CUDA_TRY(ctx, cudaMemsetAsync(s.misc, 0, 32, st));     // the sorting path's control words
    if (fused) {
        // the sorting path works on descriptors: decode first (after a direct attempt: only if that gave up)
        frame_decode_kernel<<<grid_for(ctx, n, 256, 8), 256, 0, st>>>(n, in, (const u64*)fused->in_off, fused->desc_work, run_flag);
        count_launch(ctx);
        desc = fused->desc_work;
    }
    if (small) {
        // one block does init + sort + scan + emit; misc stays zero, so the gather finds the per-frame
        // destinations in keys[1] exactly as after a sort without passes
        demux_small_kernel<<<1, kSmallDemux, 0, st>>>((uint32_t)n, desc, session, max_streams, uniq_sid, uniq_session,
                                                      (u64*)csr, perm_sort, k_out, (u64*)gathered_off, s.keys[1], dst_mode, first_only);
        count_launch(ctx);
        if (gathered) {
            const int rounds = copy_tile_rounds(ctx, n, ctx->occ_gather);
            demux_gather_kernel<<<copy_grid(ctx, n, rounds, ctx->occ_gather), 256, 0, st>>>(n, s.misc, bufs, desc, in, gathered,
                                                                        gathered_cap, k_out, max_streams, dst_mode, rounds, nullptr);
            count_launch(ctx);
        }
        CUDA_TRY(ctx, cudaGetLastError());
        return P2PT_OK;
    }
    const bool packed = (session == nullptr);          // key and frame index in one 64-bit word (frame_kernels.cuh)
    demux_init_kernel<<<grid_for(ctx, n, 256, 8), 256, 0, st>>>(n, desc, session, s.keys[0], s.idx[0], s.misc, s.blen, run_flag);
    bool coop_done = false;
    if (ctx->coop_sort) {
        // all nine passes in one cooperative launch (grid-wide syncs between the phases)
        u64 n_arg = n; const uint32_t* desc_arg = s.blen; uint32_t* hist_arg = s.hist; const u64* misc_arg = s.misc;
        uint32_t* rt_arg = s.row_total;
        bool packed_arg = packed;
        const uint32_t* flag_arg = run_flag;
        void* args[] = {&n_arg, &bufs, &desc_arg, &hist_arg, &misc_arg, &rt_arg, &packed_arg, &flag_arg};
        cudaError_t ce = cudaLaunchCooperativeKernel((void*)demux_sort_coop_kernel, dim3(kSortBlocks), dim3(kSortThreads), args, 0, st);
        if (ce == cudaSuccess) { coop_done = true; count_launch(ctx); }
        else { (void)cudaGetLastError(); ctx->coop_sort = false; }      // not supported here: fall back for good
    }
    if (!coop_done) {
        for (int pass = 0; pass < kSortPasses; pass++) {
            demux_hist_kernel<<<kSortBlocks, kSortThreads, 0, st>>>(n, bufs, s.blen, pass, s.hist, s.misc, packed, run_flag);
            demux_rowscan_kernel<<<kRadix / kRowScanWarps, 32 * kRowScanWarps, 0, st>>>(s.hist, pass, s.misc, s.row_total, run_flag);
            demux_scatter_kernel<<<kSortBlocks, kSortThreads, 0, st>>>(n, bufs, s.blen, pass, s.hist, s.misc, s.row_total, packed, run_flag);
        }
        count_launch(ctx, 3 * kSortPasses);
    }
    const u64 nblk = demux_emit_blocks(n);
    const unsigned scan_grid = run_flag && nblk > (u64)ctx->sm_count ? (unsigned)ctx->sm_count : (unsigned)nblk;
    demux_blocksum_kernel<<<scan_grid, kScanThreads, 0, st>>>(n, bufs, s.blen, s.misc, s.blk_heads, s.blk_bytes, nblk, first_only, packed, run_flag);
    demux_scan_blocksums_kernel<<<1, kScanThreads, 0, st>>>(nblk, s.blk_heads, s.blk_bytes, run_flag);
    demux_emit_kernel<<<scan_grid, kScanThreads, 0, st>>>(n, bufs, s.blen, s.misc, s.blk_heads, s.blk_bytes, nblk,
                                                              max_streams, uniq_sid, uniq_session, (u64*)csr, perm_sort, k_out,
                                                              (u64*)gathered_off, dst_mode, first_only, packed, run_flag);
    count_launch(ctx, 1 + 3);
    if (gathered) {
        const int rounds = copy_tile_rounds(ctx, n, ctx->occ_gather);
        demux_gather_kernel<<<copy_grid(ctx, n, rounds, ctx->occ_gather), 256, 0, st>>>(n, s.misc, bufs, desc, in, gathered,
                                                                    gathered_cap, k_out, max_streams, dst_mode, rounds, run_flag);
        count_launch(ctx);
    }
    CUDA_TRY(ctx, cudaGetLastError());
    if (side) {                                            // join: the caller's stream continues when both paths are through
        std::lock_guard<std::mutex> lk(side->mu);
        CUDA_TRY(ctx, cudaEventRecord(side->join, side->st));
        CUDA_TRY(ctx, cudaStreamWaitEvent(user_st, side->join, 0));
    }
    return P2PT_OK;
}

extern "C" {

int32_t p2pt_frame_demux_batch_device(p2pt_ctx* ctx, uint64_t n, const p2pt_desc* desc, const uint32_t* session,
                                      const uint8_t* in,
                                      uint32_t max_streams, uint32_t* uniq_sid, uint32_t* uniq_session,
                                      uint64_t* csr, uint32_t* perm,
                                      uint32_t* k_out, uint8_t* gathered, uint64_t gathered_cap,
                                      uint64_t* gathered_off, void* scratch, uint64_t scratch_bytes, void* stream) {
    return demux_device_impl(ctx, n, desc, session, in, max_streams, uniq_sid, uniq_session, csr, perm, k_out, gathered,
                             gathered_cap, gathered_off, scratch, scratch_bytes, stream, false);
}

uint64_t p2pt_frame_decode_demux_scratch_bytes(uint64_t n) {
    const uint64_t nn = n ? n : 1;
    return align256(demux_scratch_bytes(nn)) + align256(24 * nn) + align256(4 * nn);
}

int32_t p2pt_frame_decode_demux_batch_device(p2pt_ctx* ctx, uint64_t n, const uint8_t* in, const uint64_t* in_off,
                                             const uint32_t* session, p2pt_desc* desc,
                                             uint32_t max_streams, uint32_t* uniq_sid, uint32_t* uniq_session,
                                             uint64_t* csr, uint32_t* perm, uint32_t* k_out,
                                             uint8_t* gathered, uint64_t gathered_cap, uint64_t* gathered_off,
                                             void* scratch, uint64_t scratch_bytes, void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n && (!in_off || !scratch)) return set_err(ctx, P2PT_ERR_ARG, "frame_decode_demux: null in_off or scratch");
    if (n && scratch_bytes < p2pt_frame_decode_demux_scratch_bytes(n))
        return set_err(ctx, P2PT_ERR_CAPACITY, "frame_decode_demux: scratch too small");
    DemuxFused f;
    const uint64_t nn = n ? n : 1;
    f.in_off = in_off;
    f.desc_user = desc;
    uint8_t* p = (uint8_t*)scratch + align256(demux_scratch_bytes(nn));
    f.desc_work = desc ? desc : (p2pt_desc*)p;
    p += align256(24 * nn);
    f.perm_work = perm ? perm : (uint32_t*)p;
    return demux_device_impl(ctx, n, nullptr, session, in, max_streams, uniq_sid, uniq_session, csr, perm, k_out, gathered,
                             gathered_cap, gathered_off, scratch, scratch_bytes, stream, false, &f);
}

// ---------------------------------------------------------------------------------------------------
// SCTP DATA chunks + CRC32c
// ---------------------------------------------------------------------------------------------------
uint64_t p2pt_sctp_fragments(uint64_t msg_len) {
    const uint64_t k = (msg_len + kSctpMaxPayload - 1) / kSctpMaxPayload;
    return k ? k : 1;
}
uint64_t p2pt_sctp_packets_bytes(uint64_t msg_len) {
    const uint64_t k = p2pt_sctp_fragments(msg_len), last = msg_len - (k - 1) * kSctpMaxPayload;
    return (k - 1) * (uint64_t)kSctpMaxPacket + 28 + ((last + 3) & ~(uint64_t)3);
}

int32_t p2pt_sctp_packetize_batch_device(p2pt_ctx* ctx, uint64_t n, const uint8_t* msg, const uint64_t* msg_off,
                                         const uint16_t* src_port, const uint16_t* dst_port, const uint32_t* vtag,
                                         const uint16_t* sid, const uint16_t* ssn, const uint32_t* ppid, const uint32_t* tsn0,
                                         const uint16_t* epoch, const uint64_t* seq48_first,
                                         uint8_t* out, const uint64_t* out_msg_off, const uint64_t* pkt_first,
                                         uint64_t* pkt_off, void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!msg_off || !src_port || !dst_port || !vtag || !sid || !ssn || !ppid || !tsn0 || !out || !out_msg_off || !pkt_first || !pkt_off)
        return set_err(ctx, P2PT_ERR_ARG, "sctp_packetize: null argument");
    if ((epoch == nullptr) != (seq48_first == nullptr))
        return set_err(ctx, P2PT_ERR_ARG, "sctp_packetize: epoch and seq48_first go together");
    DeviceGuard dg(ctx->device);
    SctpPacketizeArgs a;
    a.n = n; a.msg = msg; a.msg_off = (const u64*)msg_off; a.src_port = src_port; a.dst_port = dst_port; a.vtag = vtag;
    a.sid = sid; a.ssn = ssn; a.ppid = ppid; a.tsn0 = tsn0; a.epoch = epoch; a.seq48_first = (const u64*)seq48_first;
    a.out = out; a.out_msg_off = (const u64*)out_msg_off; a.pkt_first = (const u64*)pkt_first; a.pkt_off = (u64*)pkt_off;
    const uint32_t smem = sctp_packetize_smem_bytes(ctx->lut_slack);
    sctp_packetize_kernel<<<grid_for(ctx, (n + 32 / kPkLanes - 1) / (32 / kPkLanes), kSctpWarps, 1), 32 * kSctpWarps, smem, (cudaStream_t)stream>>>(a);
    count_launch(ctx);
    CUDA_TRY(ctx, cudaGetLastError());
    return P2PT_OK;
}

int32_t p2pt_sctp_unpack_batch_device(p2pt_ctx* ctx, uint64_t n_pkts, const uint8_t* in, const uint64_t* in_off,
                                      uint32_t headroom, const uint32_t* tsn_base, p2pt_sctp_chunk* info, p2pt_desc* desc,
                                      void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n_pkts == 0) return P2PT_OK;
    if (!in || !in_off || !info || !desc) return set_err(ctx, P2PT_ERR_ARG, "sctp_unpack: null argument");
    DeviceGuard dg(ctx->device);
    SctpUnpackArgs a;
    a.n = n_pkts; a.in = in; a.in_off = (const u64*)in_off; a.headroom = headroom; a.info = info; a.desc = desc;
    a.tsn_base = tsn_base;
    const uint32_t smem = sctp_unpack_smem_bytes(ctx->lut_slack);
    sctp_unpack_kernel<<<grid_for(ctx, (n_pkts + 1) / 2, kUnpackWarps, 1), 32 * kUnpackWarps, smem, (cudaStream_t)stream>>>(a);
    count_launch(ctx);
    CUDA_TRY(ctx, cudaGetLastError());
    return P2PT_OK;
}

namespace {
struct ReasmScratch {
    uint8_t* demux; uint64_t demux_bytes;
    uint32_t* uniq_tsn; uint32_t* uniq_sess; uint64_t* csr; uint64_t* goff; uint32_t* perm; uint32_t* k; uint32_t* blk;
    uint64_t total;
};
ReasmScratch reasm_carve(void* base, uint64_t n) {
    ReasmScratch r;
    uint8_t* p = (uint8_t*)base;
    const uint64_t nn = n ? n : 1;
    r.demux = p; r.demux_bytes = demux_scratch_bytes(nn); p += align256(r.demux_bytes);
    r.uniq_tsn = (uint32_t*)p; p += align256(4 * (nn + 1));
    r.uniq_sess = (uint32_t*)p; p += align256(4 * (nn + 1));
    r.csr = (uint64_t*)p; p += align256(8 * (nn + 1));
    r.goff = (uint64_t*)p; p += align256(8 * (nn + 1));
    r.perm = (uint32_t*)p; p += align256(4 * nn);
    r.k = (uint32_t*)p; p += 256;
    r.blk = (uint32_t*)p; p += align256(4 * (demux_scan_blocks(nn) + 2));
    r.total = (uint64_t)(p - (uint8_t*)base);
    return r;
}
}  // namespace

uint64_t p2pt_sctp_reassemble_scratch_bytes(uint64_t n_pkts) { return reasm_carve(nullptr, n_pkts).total; }

int32_t p2pt_sctp_reassemble_batch_device(p2pt_ctx* ctx, uint64_t n_pkts, const p2pt_sctp_chunk* info,
                                          const p2pt_desc* desc, const uint32_t* session, const uint8_t* in,
                                          uint32_t max_msgs, uint8_t* msg, uint64_t msg_cap, uint64_t* msg_off,
                                          uint32_t* msg_first, uint8_t* msg_status, uint32_t* n_msgs_out,
                                          void* scratch, uint64_t scratch_bytes, void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (!msg_off || !msg_first || !msg_status || !n_msgs_out) return set_err(ctx, P2PT_ERR_ARG, "sctp_reassemble: null output");
    DeviceGuard dg(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    if (n_pkts == 0) {
        CUDA_TRY(ctx, cudaMemsetAsync(n_msgs_out, 0, 4, st));
        CUDA_TRY(ctx, cudaMemsetAsync(msg_off, 0, 8, st));
        return P2PT_OK;
    }
    if (!info || !desc || !in || !msg || !scratch) return set_err(ctx, P2PT_ERR_ARG, "sctp_reassemble: null argument");
    if (n_pkts > 0xffffffffull) return set_err(ctx, P2PT_ERR_ARG, "sctp_reassemble: n must fit in 32 bits");
    if (((uintptr_t)scratch & 255) != 0) return set_err(ctx, P2PT_ERR_ARG, "sctp_reassemble: scratch must be 256-byte aligned");
    ReasmScratch r = reasm_carve(scratch, n_pkts);
    if (scratch_bytes < r.total) return set_err(ctx, P2PT_ERR_CAPACITY, "sctp_reassemble: scratch too small");
    // order the fragments by (session, TSN) and lay their payloads out back to back: the demux does
    // exactly that with key = desc.stream_id (= TSN, set by the unpack kernel).  first_only: a second packet
    // with the same (session, TSN) is a retransmission; only the first to arrive contributes its payload
    // (webrtc-sctp's reassembly queue ignores a chunk whose TSN it already holds)
    if (int32_t rc = demux_device_impl(ctx, n_pkts, desc, session, in, (uint32_t)n_pkts, r.uniq_tsn,
                                       session ? r.uniq_sess : nullptr, r.csr, r.perm, r.k, msg, msg_cap,
                                       r.goff, r.demux, r.demux_bytes, st, true)) return rc;
    CUDA_TRY(ctx, cudaMemsetAsync(msg_status, 0, max_msgs, st));
    const u64 nblk = demux_scan_blocks(n_pkts);
    sctp_bounds_blocksum_kernel<<<(unsigned)nblk, kScanThreads, 0, st>>>(r.k, r.perm, (const u64*)r.csr, info, r.blk);
    sctp_bounds_scan_kernel<<<1, kScanThreads, 0, st>>>(nblk, r.blk);
    sctp_bounds_emit_kernel<<<(unsigned)nblk, kScanThreads, 0, st>>>(nblk, r.k, r.perm, (const u64*)r.csr, (const u64*)r.goff,
                                                                   session ? r.uniq_sess : nullptr, info, r.blk, max_msgs,
                                                                   (u64*)msg_off, msg_first, msg_status, n_msgs_out);
    count_launch(ctx, 3);
    CUDA_TRY(ctx, cudaGetLastError());
    return P2PT_OK;
}

}  // extern "C" (reopened below)

// ---- demux without any allocation ------------------------------------------------------------------
// Scratch for up to P2PT_NOALLOC_MAX_FRAMES frames, laid out exactly as demux_carve() expects.  It is a
// file-scope device array: in host code its name is only a handle, so the device address is looked up
// once per device with cudaGetSymbolAddress (below) and that pointer is what the kernels receive.
namespace {
constexpr uint64_t kNoAllocFrames = P2PT_NOALLOC_MAX_FRAMES;
constexpr uint64_t kNoAllocBytes = demux_scratch_bytes(kNoAllocFrames);      // same layout function the kernels use
constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_addr[kNoAllocMaxDevices] = {};
}  // namespace

static __device__ __align__(256) unsigned char g_demux_static_scratch[kNoAllocBytes];

extern "C" {

int32_t p2pt_frame_demux_batch_device_noalloc(p2pt_ctx* ctx, uint64_t n, const p2pt_desc* desc, const uint32_t* session,
                                              const uint8_t* in,
                                              uint32_t max_streams, uint32_t* uniq_sid, uint32_t* uniq_session,
                                              uint64_t* csr, uint32_t* perm,
                                              uint32_t* k_out, uint8_t* gathered, uint64_t gathered_cap,
                                              uint64_t* gathered_off, void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n > kNoAllocFrames)
        return set_err(ctx, P2PT_ERR_CAPACITY,
                       "frame_demux (no-alloc): n = " + std::to_string(n) + " exceeds P2PT_NOALLOC_MAX_FRAMES (" +
                           std::to_string(kNoAllocFrames) + "); split the batch or pass your own scratch to "
                           "p2pt_frame_demux_batch_device");
    if (ctx->device < 0 || ctx->device >= kNoAllocMaxDevices)
        return set_err(ctx, P2PT_ERR_ARG, "frame_demux (no-alloc): device index beyond the per-device address table");
    if (demux_scratch_bytes(n ? n : 1) > kNoAllocBytes)
        return set_err(ctx, P2PT_ERR_STATE, "frame_demux (no-alloc): static scratch smaller than the layout needs");
    DeviceGuard dg(ctx->device);                       // the symbol address is per device: look it up with ours current
    void* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!g_noalloc_addr[ctx->device]) {            // first call on this device: one-time set-up
            void* p = nullptr;
            cudaError_t e = cudaGetSymbolAddress(&p, g_demux_static_scratch);
            if (e != cudaSuccess || !p)
                return set_err(ctx, P2PT_ERR_CUDA, std::string("cudaGetSymbolAddress(static demux scratch): ") + cudaGetErrorString(e));
            g_noalloc_addr[ctx->device] = p;
        }
        scratch = g_noalloc_addr[ctx->device];
    }
    // same kernels, same order as the caller-scratch entry point; it zeroes its control words itself
    return p2pt_frame_demux_batch_device(ctx, n, desc, session, in, max_streams, uniq_sid, uniq_session, csr, perm, k_out,
                                         gathered, gathered_cap, gathered_off, scratch, kNoAllocBytes, stream);
}

static int32_t check_sessions(p2pt_ctx* ctx) {
    if (!ctx->n_sessions) return set_err(ctx, P2PT_ERR_STATE, "no sessions loaded: call p2pt_sessions_load first");
    return P2PT_OK;
}

int32_t p2pt_dtls_gcm_seal_batch_device(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                                        const uint8_t* nonce8, const uint8_t* raw, const uint64_t* raw_off,
                                        uint8_t* rec, const uint64_t* rec_off, uint8_t* status, void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!session || !nonce8 || !raw || !raw_off || !rec || !rec_off || !status)
        return set_err(ctx, P2PT_ERR_ARG, "dtls_gcm_seal: null argument");
    if (int32_t r = check_sessions(ctx)) return r;
    DeviceGuard dg(ctx->device);
    GcmArgs a = {};
    a.n = n; a.session = session; a.keys = ctx->keys_local; a.n_sessions = ctx->n_sessions;
    a.nonce8 = nonce8; a.in = raw; a.in_off = (const u64*)raw_off; a.out = rec; a.out_off = (const u64*)rec_off;
    a.status = status;
    return launch_gcm<kSeal>(ctx, a, ctx->gcm_g ? ctx->gcm_g : 4, (cudaStream_t)stream);
}

int32_t p2pt_dtls_gcm_open_batch_device(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                                        const uint8_t* rec, const uint64_t* rec_off,
                                        uint8_t* out, const uint64_t* out_off,
                                        uint32_t* out_len, uint8_t* status, void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!session || !rec || !rec_off || !out || !out_off || !out_len || !status)
        return set_err(ctx, P2PT_ERR_ARG, "dtls_gcm_open: null argument");
    if (int32_t r = check_sessions(ctx)) return r;
    DeviceGuard dg(ctx->device);
    GcmArgs a = {};
    a.n = n; a.session = session; a.keys = ctx->keys_remote; a.n_sessions = ctx->n_sessions;
    a.in = rec; a.in_off = (const u64*)rec_off; a.out = out; a.out_off = (const u64*)out_off;
    a.status = status; a.out_len = out_len;
    return launch_gcm<kOpen>(ctx, a, ctx->gcm_g ? ctx->gcm_g : 4, (cudaStream_t)stream);
}

int32_t p2pt_pipeline_seal_device(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                                  const uint16_t* epoch, const uint64_t* seq48, const uint8_t* nonce8,
                                  const uint8_t* types, const uint32_t* sids,
                                  const uint8_t* payload, const uint64_t* pay_off,
                                  uint8_t* rec, const uint64_t* rec_off, uint8_t* status, void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!session || !epoch || !seq48 || !nonce8 || !types || !sids || !pay_off || !rec || !rec_off || !status)
        return set_err(ctx, P2PT_ERR_ARG, "pipeline_seal: null argument");
    if (int32_t r = check_sessions(ctx)) return r;
    DeviceGuard dg(ctx->device);
    GcmArgs a = {};
    a.n = n; a.session = session; a.keys = ctx->keys_local; a.n_sessions = ctx->n_sessions;
    a.nonce8 = nonce8; a.in = payload; a.in_off = (const u64*)pay_off; a.out = rec; a.out_off = (const u64*)rec_off;
    a.status = status; a.epoch = epoch; a.seq48 = (const u64*)seq48; a.types = types; a.sids = sids;
    return launch_gcm<kPipeSeal>(ctx, a, ctx->gcm_g ? ctx->gcm_g : 4, (cudaStream_t)stream);
}

int32_t p2pt_pipeline_open_device(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                                  const uint8_t* rec, const uint64_t* rec_off,
                                  uint8_t* out, const uint64_t* out_off, p2pt_desc* desc, void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!session || !rec || !rec_off || !out || !out_off || !desc)
        return set_err(ctx, P2PT_ERR_ARG, "pipeline_open: null argument");
    if (int32_t r = check_sessions(ctx)) return r;
    DeviceGuard dg(ctx->device);
    GcmArgs a = {};
    a.n = n; a.session = session; a.keys = ctx->keys_remote; a.n_sessions = ctx->n_sessions;
    a.in = rec; a.in_off = (const u64*)rec_off; a.out = out; a.out_off = (const u64*)out_off;
    a.desc = desc;
    return launch_gcm<kPipeOpen>(ctx, a, ctx->gcm_g ? ctx->gcm_g : 4, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------------------
// host-pointer entry points: stage, run, fetch
// ---------------------------------------------------------------------------------------------------
static bool offsets_monotonic(const uint64_t* off, uint64_t n) {
    for (uint64_t i = 0; i < n; i++)
        if (off[i + 1] < off[i]) return false;
    return true;
}

int32_t p2pt_frame_encode_batch(p2pt_ctx* ctx, uint64_t n, const uint8_t* types, const uint32_t* sids,
                                const uint8_t* payload, const uint64_t* pay_off,
                                uint8_t* out, const uint64_t* out_off) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!types || !sids || !pay_off || !out || !out_off) return set_err(ctx, P2PT_ERR_ARG, "frame_encode: null argument");
    if (!offsets_monotonic(pay_off, n) || !offsets_monotonic(out_off, n))
        return set_err(ctx, P2PT_ERR_ARG, "frame_encode: offsets must be non-decreasing");
    for (uint64_t i = 0; i < n; i++)
        if (out_off[i + 1] - out_off[i] < 5 + (pay_off[i + 1] - pay_off[i]))
            return set_err(ctx, P2PT_ERR_CAPACITY, "frame_encode: output slot smaller than 5 + payload");
    DeviceGuard dg(ctx->device);
    LaneGuard lg(ctx);
    NvtxRange nvtx_("p2pt_frame_encode_batch");
    Lane* L = lg.lane;
    cudaStream_t st = L->streams[0];
    const size_t pbytes = pay_off[n] - pay_off[0], obytes = out_off[n] - out_off[0];
    if (pbytes && !payload) return set_err(ctx, P2PT_ERR_ARG, "frame_encode: null payload");
    IoStage io{ctx, L};
    if (int32_t r = io.reserve(pad256(n) + pad256(4 * n) + 2 * pad256(8 * (n + 1)) + pad256(pbytes + 1) + pad256(obytes + 1) + 4096)) return r;
    uint8_t* d_types = io.take<uint8_t>(n);
    uint32_t* d_sids = io.take<uint32_t>(n);
    uint64_t* d_po = io.take<uint64_t>(n + 1);
    uint64_t* d_oo = io.take<uint64_t>(n + 1);
    uint8_t* d_pay = io.take<uint8_t>(pbytes + 1);
    uint8_t* d_out = io.take<uint8_t>(obytes + 1);
    CUDA_TRY(ctx, io.in(d_types, types, n, st));
    CUDA_TRY(ctx, io.in(d_sids, sids, 4 * n, st));
    CUDA_TRY(ctx, io.in(d_po, pay_off, 8 * (n + 1), st));
    CUDA_TRY(ctx, io.in(d_oo, out_off, 8 * (n + 1), st));
    if (pbytes) CUDA_TRY(ctx, io.in(d_pay, payload + pay_off[0], pbytes, st));
    if (int32_t r = p2pt_frame_encode_batch_device(ctx, n, d_types, d_sids, d_pay - pay_off[0], d_po,
                                                   d_out - out_off[0], d_oo, st)) return r;
    // only the bytes the frames cover are defined; gaps between slots keep the caller's contents
    bool dense = true;
    for (uint64_t i = 0; i < n && dense; i++) dense = (out_off[i + 1] - out_off[i] == 5 + (pay_off[i + 1] - pay_off[i]));
    if (dense) {
        CUDA_TRY(ctx, io.out(out + out_off[0], d_out, obytes, st));
    } else {
        for (uint64_t i = 0; i < n; i++) {
            const size_t len = 5 + (pay_off[i + 1] - pay_off[i]);
            CUDA_TRY(ctx, io.out(out + out_off[i], d_out + (out_off[i] - out_off[0]), len, st));
        }
    }
    CUDA_TRY(ctx, io.finish(st));
    lg.dismiss();
    return P2PT_OK;
}

int32_t p2pt_frame_decode_batch(p2pt_ctx* ctx, uint64_t n, const uint8_t* in, const uint64_t* in_off, p2pt_desc* desc) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!in_off || !desc) return set_err(ctx, P2PT_ERR_ARG, "frame_decode: null argument");
    if (!offsets_monotonic(in_off, n)) return set_err(ctx, P2PT_ERR_ARG, "frame_decode: offsets must be non-decreasing");
    DeviceGuard dg(ctx->device);
    LaneGuard lg(ctx);
    NvtxRange nvtx_("p2pt_frame_decode_batch");
    Lane* L = lg.lane;
    cudaStream_t st = L->streams[0];
    const size_t ibytes = in_off[n] - in_off[0];
    if (ibytes && !in) return set_err(ctx, P2PT_ERR_ARG, "frame_decode: null input");
    IoStage io{ctx, L};
    if (int32_t r = io.reserve(pad256(8 * (n + 1)) + pad256(ibytes + 1) + pad256(24 * n) + 4096)) return r;
    uint64_t* d_io = io.take<uint64_t>(n + 1);
    uint8_t* d_in = io.take<uint8_t>(ibytes + 1);
    p2pt_desc* d_desc = io.take<p2pt_desc>(n);
    CUDA_TRY(ctx, io.in(d_io, in_off, 8 * (n + 1), st));
    if (ibytes) CUDA_TRY(ctx, io.in(d_in, in + in_off[0], ibytes, st));
    if (int32_t r = p2pt_frame_decode_batch_device(ctx, n, d_in - in_off[0], d_io, d_desc, st)) return r;
    CUDA_TRY(ctx, io.out(desc, d_desc, 24 * n, st));
    CUDA_TRY(ctx, io.finish(st));
    lg.dismiss();
    return P2PT_OK;
}

int32_t p2pt_frame_demux_batch(p2pt_ctx* ctx, uint64_t n, const p2pt_desc* desc, const uint32_t* session,
                               const uint8_t* in, uint64_t in_bytes,
                               uint32_t max_streams, uint32_t* uniq_sid, uint32_t* uniq_session,
                               uint64_t* csr, uint32_t* perm,
                               uint32_t* k_out, uint8_t* gathered, uint64_t gathered_cap, uint64_t* gathered_off) {
    if (!ctx) return P2PT_ERR_ARG;
    if (!uniq_sid || !csr || !perm || !k_out) return set_err(ctx, P2PT_ERR_ARG, "frame_demux: null output");
    if (n == 0) { *k_out = 0; csr[0] = 0; if (gathered_off) gathered_off[0] = 0; return P2PT_OK; }
    if (!desc) return set_err(ctx, P2PT_ERR_ARG, "frame_demux: null desc");
    if (gathered && (!in || !gathered_off)) return set_err(ctx, P2PT_ERR_ARG, "frame_demux: gather needs in and gathered_off");
    DeviceGuard dg(ctx->device);
    LaneGuard lg(ctx);
    NvtxRange nvtx_("p2pt_frame_demux_batch");
    Lane* L = lg.lane;
    cudaStream_t st = L->streams[0];
    const uint64_t sbytes = demux_scratch_bytes(n);
    const size_t gcap = gathered ? (size_t)gathered_cap : 0;
    const size_t ib = gathered ? (size_t)in_bytes : 0;
    if (ctx->small_demux && L->zc_host && ctx->small_call_bytes && n <= (uint64_t)kSmallDemux) {
        // small batch: descriptors, frames and every result live in the mapped pinned buffer; two launches, one wait
        const size_t ms1 = (size_t)(max_streams < n ? max_streams : n) + 1;
        const size_t g2 = gcap < ib ? gcap : ib;                  // the bodies cannot be longer than the input
        const size_t foot = pad256(24 * n) + pad256(4 * n) + pad256(ib + 1) + 2 * pad256(4 * ms1) + 2 * pad256(8 * ms1) +
                            pad256(4 * n) + pad256(g2 + 1) + 256;
        if (foot <= ctx->small_call_bytes && foot <= L->zc_cap) {
            if (int32_t r = arena_reserve(ctx, L, pad256(sbytes) + 4096)) return r;
            uint8_t* d_scratch = arena_take<uint8_t>(L, sbytes);
            ZcCarver zcv{L};
            auto zc = [&](size_t bytes) { return zcv.take(bytes); };
            auto host_of = [&](const void* d) { return zcv.host_of(d); };
            p2pt_desc* z_desc = (p2pt_desc*)zc(24 * n);
            uint32_t* z_sess = (uint32_t*)zc(4 * n);
            uint8_t* z_in = zc(ib + 1);
            uint32_t* z_uniq = (uint32_t*)zc(4 * ms1);
            uint32_t* z_usess = (uint32_t*)zc(4 * ms1);
            uint64_t* z_csr = (uint64_t*)zc(8 * ms1);
            uint64_t* z_goff = (uint64_t*)zc(8 * ms1);
            uint32_t* z_perm = (uint32_t*)zc(4 * n);
            uint8_t* z_g = zc(g2 + 1);
            uint32_t* z_k = (uint32_t*)zc(4);
            memcpy(host_of(z_desc), desc, 24 * n);
            if (session) memcpy(host_of(z_sess), session, 4 * n);
            if (ib) memcpy(host_of(z_in), in, ib);
            if (int32_t r = p2pt_frame_demux_batch_device(ctx, n, z_desc, session ? z_sess : nullptr, z_in, max_streams, z_uniq,
                                                          uniq_session ? z_usess : nullptr, z_csr, z_perm, z_k,
                                                          gathered ? z_g : nullptr, g2, gathered_off ? z_goff : nullptr,
                                                          d_scratch, sbytes, st)) return r;
            CUDA_TRY(ctx, cudaStreamSynchronize(st));
            const uint32_t k = *(const uint32_t*)host_of(z_k);
            *k_out = k;
            if (k > max_streams) return set_err(ctx, P2PT_ERR_CAPACITY, "frame_demux: more streams than max_streams");
            memcpy(uniq_sid, host_of(z_uniq), 4 * (size_t)k);
            if (uniq_session) memcpy(uniq_session, host_of(z_usess), 4 * (size_t)k);
            memcpy(csr, host_of(z_csr), 8 * ((size_t)k + 1));
            const uint64_t m = csr[k];
            if (m) memcpy(perm, host_of(z_perm), 4 * m);
            if (gathered_off) {
                memcpy(gathered_off, host_of(z_goff), 8 * ((size_t)k + 1));
                if (gathered) {
                    if (gathered_off[k] > gathered_cap) return set_err(ctx, P2PT_ERR_CAPACITY, "frame_demux: gathered_cap too small");
                    if (gathered_off[k]) memcpy(gathered, host_of(z_g), gathered_off[k]);
                }
            }
            lg.dismiss();
            return P2PT_OK;
        }
    }
    if (int32_t r = arena_reserve(ctx, L, pad256(24 * n) + pad256(4 * n) + pad256(ib) + 2 * pad256(4 * ((size_t)max_streams + 1)) +
                                           2 * pad256(8 * ((size_t)max_streams + 1)) + pad256(4 * n) + 256 +
                                           pad256(gcap) + pad256(sbytes) + 8192)) return r;
    p2pt_desc* d_desc = arena_take<p2pt_desc>(L, n);
    uint32_t* d_sess = arena_take<uint32_t>(L, n);
    uint8_t* d_in = arena_take<uint8_t>(L, ib + 1);
    uint32_t* d_uniq = arena_take<uint32_t>(L, max_streams + 1);
    uint32_t* d_usess = arena_take<uint32_t>(L, max_streams + 1);
    uint64_t* d_csr = arena_take<uint64_t>(L, (size_t)max_streams + 1);
    uint64_t* d_goff = arena_take<uint64_t>(L, (size_t)max_streams + 1);
    uint32_t* d_perm = arena_take<uint32_t>(L, n);
    uint32_t* d_k = arena_take<uint32_t>(L, 1);
    uint8_t* d_g = arena_take<uint8_t>(L, gcap + 1);
    uint8_t* d_scratch = arena_take<uint8_t>(L, sbytes);
    CUDA_TRY(ctx, cudaMemcpyAsync(d_desc, desc, 24 * n, cudaMemcpyHostToDevice, st));
    if (session) CUDA_TRY(ctx, cudaMemcpyAsync(d_sess, session, 4 * n, cudaMemcpyHostToDevice, st));
    if (ib) CUDA_TRY(ctx, cudaMemcpyAsync(d_in, in, ib, cudaMemcpyHostToDevice, st));
    if (int32_t r = p2pt_frame_demux_batch_device(ctx, n, d_desc, session ? d_sess : nullptr, d_in, max_streams, d_uniq,
                                                  uniq_session ? d_usess : nullptr, d_csr, d_perm, d_k,
                                                  gathered ? d_g : nullptr, gathered_cap,
                                                  gathered_off ? d_goff : nullptr, d_scratch, sbytes, st)) return r;
    uint32_t k = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(&k, d_k, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    *k_out = k;
    if (k > max_streams) return set_err(ctx, P2PT_ERR_CAPACITY, "frame_demux: more streams than max_streams");
    CUDA_TRY(ctx, cudaMemcpyAsync(uniq_sid, d_uniq, 4 * (size_t)k, cudaMemcpyDeviceToHost, st));
    if (uniq_session) CUDA_TRY(ctx, cudaMemcpyAsync(uniq_session, d_usess, 4 * (size_t)k, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(csr, d_csr, 8 * ((size_t)k + 1), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    const uint64_t m = csr[k];
    if (m) CUDA_TRY(ctx, cudaMemcpyAsync(perm, d_perm, 4 * m, cudaMemcpyDeviceToHost, st));
    if (gathered_off) {
        CUDA_TRY(ctx, cudaMemcpyAsync(gathered_off, d_goff, 8 * ((size_t)k + 1), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(ctx, wait_stream(ctx, L, st));
        if (gathered) {
            if (gathered_off[k] > gathered_cap) return set_err(ctx, P2PT_ERR_CAPACITY, "frame_demux: gathered_cap too small");
            if (gathered_off[k]) CUDA_TRY(ctx, cudaMemcpyAsync(gathered, d_g, gathered_off[k], cudaMemcpyDeviceToHost, st));
        }
    }
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    lg.dismiss();
    return P2PT_OK;
}

// The reader loop of proxy.rs:109-170 / serve.rs:82-156 on frames that are already plaintext: decode, demux and
// body gather in one call.  Frames and offsets go up once; descriptors and perm come back only if asked for.
int32_t p2pt_frame_decode_demux_batch(p2pt_ctx* ctx, uint64_t n, const uint8_t* in, const uint64_t* in_off,
                                      const uint32_t* session, p2pt_desc* desc,
                                      uint32_t max_streams, uint32_t* uniq_sid, uint32_t* uniq_session,
                                      uint64_t* csr, uint32_t* perm, uint32_t* k_out,
                                      uint8_t* gathered, uint64_t gathered_cap, uint64_t* gathered_off) {
    if (!ctx) return P2PT_ERR_ARG;
    if (!uniq_sid || !csr || !k_out) return set_err(ctx, P2PT_ERR_ARG, "frame_decode_demux: null output");
    if (n == 0) { *k_out = 0; csr[0] = 0; if (gathered_off) gathered_off[0] = 0; return P2PT_OK; }
    if (!in_off) return set_err(ctx, P2PT_ERR_ARG, "frame_decode_demux: null in_off");
    if (gathered && !gathered_off) return set_err(ctx, P2PT_ERR_ARG, "frame_decode_demux: gather needs gathered_off");
    if (n > 0xffffffffull) return set_err(ctx, P2PT_ERR_ARG, "frame_decode_demux: n must fit in 32 bits");
    if (!offsets_monotonic(in_off, n)) return set_err(ctx, P2PT_ERR_ARG, "frame_decode_demux: offsets must be non-decreasing");
    const size_t ibytes = in_off[n] - in_off[0];
    if (ibytes && !in) return set_err(ctx, P2PT_ERR_ARG, "frame_decode_demux: null input");
    DeviceGuard dg(ctx->device);
    LaneGuard lg(ctx);
    NvtxRange nvtx_("p2pt_frame_decode_demux_batch");
    Lane* L = lg.lane;
    cudaStream_t st = L->streams[0];
    const uint64_t sbytes = p2pt_frame_decode_demux_scratch_bytes(n);
    const size_t gcap = gathered ? (size_t)gathered_cap : 0;
    if (int32_t r = arena_reserve(ctx, L, pad256(8 * (n + 1)) + pad256(4 * n) + pad256(ibytes + 1) + pad256(24 * n) +
                                           2 * pad256(4 * ((size_t)max_streams + 1)) + 2 * pad256(8 * ((size_t)max_streams + 1)) +
                                           pad256(4 * n) + 256 + pad256(gcap + 1) + pad256(sbytes) + 8192)) return r;
    uint64_t* d_io = arena_take<uint64_t>(L, n + 1);
    uint32_t* d_sess = arena_take<uint32_t>(L, n);
    uint8_t* d_in = arena_take<uint8_t>(L, ibytes + 1);
    p2pt_desc* d_desc = arena_take<p2pt_desc>(L, n);
    uint32_t* d_uniq = arena_take<uint32_t>(L, (size_t)max_streams + 1);
    uint32_t* d_usess = arena_take<uint32_t>(L, (size_t)max_streams + 1);
    uint64_t* d_csr = arena_take<uint64_t>(L, (size_t)max_streams + 1);
    uint64_t* d_goff = arena_take<uint64_t>(L, (size_t)max_streams + 1);
    uint32_t* d_perm = arena_take<uint32_t>(L, n);
    uint32_t* d_k = arena_take<uint32_t>(L, 1);
    uint8_t* d_g = arena_take<uint8_t>(L, gcap + 1);
    uint8_t* d_scratch = arena_take<uint8_t>(L, sbytes);
    CUDA_TRY(ctx, cudaMemcpyAsync(d_io, in_off, 8 * (n + 1), cudaMemcpyHostToDevice, st));
    if (session) CUDA_TRY(ctx, cudaMemcpyAsync(d_sess, session, 4 * n, cudaMemcpyHostToDevice, st));
    if (ibytes) CUDA_TRY(ctx, cudaMemcpyAsync(d_in, in + in_off[0], ibytes, cudaMemcpyHostToDevice, st));
    if (int32_t r = p2pt_frame_decode_demux_batch_device(ctx, n, d_in - in_off[0], d_io, session ? d_sess : nullptr,
                                                         desc ? d_desc : nullptr, max_streams, d_uniq,
                                                         uniq_session ? d_usess : nullptr, d_csr, perm ? d_perm : nullptr, d_k,
                                                         gathered ? d_g : nullptr, gathered_cap,
                                                         gathered_off ? d_goff : nullptr, d_scratch, sbytes, st)) return r;
    if (desc) CUDA_TRY(ctx, cudaMemcpyAsync(desc, d_desc, 24 * n, cudaMemcpyDeviceToHost, st));
    uint32_t k = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(&k, d_k, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    *k_out = k;
    if (k > max_streams) return set_err(ctx, P2PT_ERR_CAPACITY, "frame_decode_demux: more streams than max_streams");
    CUDA_TRY(ctx, cudaMemcpyAsync(uniq_sid, d_uniq, 4 * (size_t)k, cudaMemcpyDeviceToHost, st));
    if (uniq_session) CUDA_TRY(ctx, cudaMemcpyAsync(uniq_session, d_usess, 4 * (size_t)k, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(csr, d_csr, 8 * ((size_t)k + 1), cudaMemcpyDeviceToHost, st));
    if (gathered_off) CUDA_TRY(ctx, cudaMemcpyAsync(gathered_off, d_goff, 8 * ((size_t)k + 1), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    const uint64_t m = csr[k];
    if (m && perm) CUDA_TRY(ctx, cudaMemcpyAsync(perm, d_perm, 4 * m, cudaMemcpyDeviceToHost, st));
    if (gathered) {
        if (gathered_off[k] > gathered_cap) return set_err(ctx, P2PT_ERR_CAPACITY, "frame_decode_demux: gathered_cap too small");
        if (gathered_off[k]) CUDA_TRY(ctx, cudaMemcpyAsync(gathered, d_g, gathered_off[k], cudaMemcpyDeviceToHost, st));
    }
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    lg.dismiss();
    return P2PT_OK;
}

// Shared staging for the four record entry points.  Records are processed in chunks round-robin
// over kStreams streams so copies and kernels of neighbouring chunks overlap.
namespace {
struct RecordCall {
    int mode;
    uint64_t n;
    const uint32_t* session;
    const uint8_t* nonce8;
    const uint8_t* in; const uint64_t* in_off;
    uint8_t* out; const uint64_t* out_off;
    uint8_t* status; uint32_t* out_len;
    const uint16_t* epoch; const uint64_t* seq48; const uint8_t* types; const uint32_t* sids;
    p2pt_desc* desc;
};
}  // namespace

static int32_t run_record_call(p2pt_ctx* ctx, const RecordCall& c, const char* what) {
    const uint64_t n = c.n;
    if (!offsets_monotonic(c.in_off, n) || !offsets_monotonic(c.out_off, n))
        return set_err(ctx, P2PT_ERR_ARG, std::string(what) + ": offsets must be non-decreasing");
    const bool seal = (c.mode == kSeal || c.mode == kPipeSeal);
    for (uint64_t i = 0; i < n; i++) {
        const uint64_t il = c.in_off[i + 1] - c.in_off[i], ol = c.out_off[i + 1] - c.out_off[i];
        uint64_t need = c.mode == kSeal ? il + 24 : c.mode == kPipeSeal ? il + 5 + 13 + 24 : il;
        if (c.mode == kSeal && il < 13) need = 0;
        if (ol < need) return set_err(ctx, P2PT_ERR_CAPACITY, std::string(what) + ": output slot too small");
    }
    if (int32_t r = check_sessions(ctx)) return r;
    DeviceGuard dg(ctx->device);
    LaneGuard lg(ctx);
    NvtxRange nvtx_(what);
    Lane* L = lg.lane;
    const size_t ibytes = c.in_off[n] - c.in_off[0], obytes = c.out_off[n] - c.out_off[0];
    IoStage io{ctx, L};
    if (int32_t r = io.reserve(pad256(4 * n) + pad256(8 * n) + 2 * pad256(8 * (n + 1)) + pad256(ibytes + 1) +
                                           pad256(obytes + 1) + pad256(n) + pad256(4 * n) + pad256(2 * n) + pad256(8 * n) +
                                           pad256(n) + pad256(4 * n) + pad256(24 * n) + 16384)) return r;
    uint32_t* d_sess = io.take<uint32_t>(n);
    uint8_t* d_nonce = io.take<uint8_t>(8 * n);
    uint64_t* d_io = io.take<uint64_t>(n + 1);
    uint64_t* d_oo = io.take<uint64_t>(n + 1);
    uint8_t* d_in = io.take<uint8_t>(ibytes + 1);
    uint8_t* d_out = io.take<uint8_t>(obytes + 1);
    uint8_t* d_status = io.take<uint8_t>(n);
    uint32_t* d_olen = io.take<uint32_t>(n);
    uint16_t* d_epoch = io.take<uint16_t>(n);
    uint64_t* d_seq = io.take<uint64_t>(n);
    uint8_t* d_types = io.take<uint8_t>(n);
    uint32_t* d_sids = io.take<uint32_t>(n);
    p2pt_desc* d_desc = io.take<p2pt_desc>(n);

    cudaStream_t s0 = L->streams[0];
    CUDA_TRY(ctx, io.in(d_io, c.in_off, 8 * (n + 1), s0));
    CUDA_TRY(ctx, io.in(d_oo, c.out_off, 8 * (n + 1), s0));
    CUDA_TRY(ctx, io.in(d_sess, c.session, 4 * n, s0));
    if (seal) CUDA_TRY(ctx, io.in(d_nonce, c.nonce8, 8 * n, s0));
    if (c.mode == kPipeSeal) {
        CUDA_TRY(ctx, io.in(d_epoch, c.epoch, 2 * n, s0));
        CUDA_TRY(ctx, io.in(d_seq, c.seq48, 8 * n, s0));
        CUDA_TRY(ctx, io.in(d_types, c.types, n, s0));
        CUDA_TRY(ctx, io.in(d_sids, c.sids, 4 * n, s0));
    }
    if (!io.zc) CUDA_TRY(ctx, cudaEventRecord(L->events[0], s0));     // the other streams wait for the metadata

    const int G = pick_group(ctx, n, ibytes);
    // chunk so that each chunk moves ~16 MiB of input
    uint64_t chunk = n;
    if (ibytes > 2 * ctx->chunk_bytes) {
        const uint64_t avg = ibytes / n + 1;
        chunk = ctx->chunk_bytes / avg;
        if (chunk < 1024) chunk = 1024;
    }
    int si = 0;
    for (uint64_t r0 = 0; r0 < n; r0 += chunk, si = (si + 1) % kStreams) {
        const uint64_t r1 = r0 + chunk < n ? r0 + chunk : n, m = r1 - r0;
        cudaStream_t st = L->streams[si];
        if (!io.zc && (si != 0 || r0 == 0)) CUDA_TRY(ctx, cudaStreamWaitEvent(st, L->events[0], 0));
        const size_t ib0 = c.in_off[r0] - c.in_off[0], ib1 = c.in_off[r1] - c.in_off[0];
        const size_t ob0 = c.out_off[r0] - c.out_off[0], ob1 = c.out_off[r1] - c.out_off[0];
        if (ib1 > ib0) CUDA_TRY(ctx, io.in(d_in + ib0, c.in + c.in_off[r0], ib1 - ib0, st));
        GcmArgs a = {};
        a.n = m; a.session = d_sess + r0; a.n_sessions = ctx->n_sessions;
        a.keys = seal ? ctx->keys_local : ctx->keys_remote;
        a.nonce8 = d_nonce + 8 * r0;
        a.in = d_in - c.in_off[0]; a.in_off = (const u64*)(d_io + r0);
        a.out = d_out - c.out_off[0]; a.out_off = (const u64*)(d_oo + r0);
        a.status = d_status + r0; a.out_len = d_olen + r0;
        a.epoch = d_epoch + r0; a.seq48 = (const u64*)(d_seq + r0); a.types = d_types + r0; a.sids = d_sids + r0;
        a.desc = d_desc + r0;
        int32_t r;
        switch (c.mode) {
            case kSeal: r = launch_gcm<kSeal>(ctx, a, G, st); break;
            case kOpen: r = launch_gcm<kOpen>(ctx, a, G, st); break;
            case kPipeSeal: r = launch_gcm<kPipeSeal>(ctx, a, G, st); break;
            default: r = launch_gcm<kPipeOpen>(ctx, a, G, st); break;
        }
        if (r) return r;
        if (ob1 > ob0) CUDA_TRY(ctx, io.out(c.out + c.out_off[r0], d_out + ob0, ob1 - ob0, st));
        if (c.status) CUDA_TRY(ctx, io.out(c.status + r0, d_status + r0, m, st));
        if (c.mode == kOpen) CUDA_TRY(ctx, io.out(c.out_len + r0, d_olen + r0, 4 * m, st));
        if (c.mode == kPipeOpen) CUDA_TRY(ctx, io.out(c.desc + r0, d_desc + r0, 24 * m, st));
    }
    if (io.zc) {
        CUDA_TRY(ctx, io.finish(s0));                                    // a small call is one chunk on stream 0
    } else {
        for (int i = 0; i < kStreams; i++) CUDA_TRY(ctx, wait_stream(ctx, L, L->streams[i]));
    }
    // pipe-open descriptors are relative to the caller's `out`, which is how the kernel wrote them
    lg.dismiss();
    return P2PT_OK;
}

int32_t p2pt_dtls_gcm_seal_batch(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                                 const uint8_t* nonce8, const uint8_t* raw, const uint64_t* raw_off,
                                 uint8_t* rec, const uint64_t* rec_off, uint8_t* status) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!session || !nonce8 || !raw || !raw_off || !rec || !rec_off || !status)
        return set_err(ctx, P2PT_ERR_ARG, "dtls_gcm_seal: null argument");
    RecordCall c = {};
    c.mode = kSeal; c.n = n; c.session = session; c.nonce8 = nonce8; c.in = raw; c.in_off = raw_off;
    c.out = rec; c.out_off = rec_off; c.status = status;
    return run_record_call(ctx, c, "dtls_gcm_seal");
}

int32_t p2pt_dtls_gcm_open_batch(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                                 const uint8_t* rec, const uint64_t* rec_off,
                                 uint8_t* out, const uint64_t* out_off, uint32_t* out_len, uint8_t* status) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!session || !rec || !rec_off || !out || !out_off || !out_len || !status)
        return set_err(ctx, P2PT_ERR_ARG, "dtls_gcm_open: null argument");
    RecordCall c = {};
    c.mode = kOpen; c.n = n; c.session = session; c.in = rec; c.in_off = rec_off;
    c.out = out; c.out_off = out_off; c.status = status; c.out_len = out_len;
    return run_record_call(ctx, c, "dtls_gcm_open");
}

int32_t p2pt_pipeline_seal(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                           const uint16_t* epoch, const uint64_t* seq48, const uint8_t* nonce8,
                           const uint8_t* types, const uint32_t* sids,
                           const uint8_t* payload, const uint64_t* pay_off,
                           uint8_t* rec, const uint64_t* rec_off, uint8_t* status) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!session || !epoch || !seq48 || !nonce8 || !types || !sids || !pay_off || !rec || !rec_off || !status)
        return set_err(ctx, P2PT_ERR_ARG, "pipeline_seal: null argument");
    if (pay_off[n] > pay_off[0] && !payload) return set_err(ctx, P2PT_ERR_ARG, "pipeline_seal: null payload");
    RecordCall c = {};
    c.mode = kPipeSeal; c.n = n; c.session = session; c.nonce8 = nonce8; c.in = payload; c.in_off = pay_off;
    c.out = rec; c.out_off = rec_off; c.status = status;
    c.epoch = epoch; c.seq48 = seq48; c.types = types; c.sids = sids;
    return run_record_call(ctx, c, "pipeline_seal");
}

int32_t p2pt_pipeline_open(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                           const uint8_t* rec, const uint64_t* rec_off,
                           uint8_t* out, const uint64_t* out_off, p2pt_desc* desc) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!session || !rec || !rec_off || !out || !out_off || !desc)
        return set_err(ctx, P2PT_ERR_ARG, "pipeline_open: null argument");
    RecordCall c = {};
    c.mode = kPipeOpen; c.n = n; c.session = session; c.in = rec; c.in_off = rec_off;
    c.out = out; c.out_off = out_off; c.desc = desc;
    return run_record_call(ctx, c, "pipeline_open");
}

int32_t p2pt_sctp_packetize_batch(p2pt_ctx* ctx, uint64_t n, const uint8_t* msg, const uint64_t* msg_off,
                                  const uint16_t* src_port, const uint16_t* dst_port, const uint32_t* vtag,
                                  const uint16_t* sid, const uint16_t* ssn, const uint32_t* ppid, const uint32_t* tsn0,
                                  const uint16_t* epoch, const uint64_t* seq48_first,
                                  uint8_t* out, uint64_t out_cap, uint64_t* pkt_off, uint64_t pkt_cap,
                                  uint64_t* n_pkts_out) {
    if (!ctx) return P2PT_ERR_ARG;
    if (!n_pkts_out || !pkt_off) return set_err(ctx, P2PT_ERR_ARG, "sctp_packetize: null output");
    if (n == 0) { *n_pkts_out = 0; if (pkt_cap) pkt_off[0] = 0; return P2PT_OK; }
    if (!msg_off || !src_port || !dst_port || !vtag || !sid || !ssn || !ppid || !tsn0 || !out)
        return set_err(ctx, P2PT_ERR_ARG, "sctp_packetize: null argument");
    if ((epoch == nullptr) != (seq48_first == nullptr)) return set_err(ctx, P2PT_ERR_ARG, "sctp_packetize: epoch and seq48_first go together");
    if (!offsets_monotonic(msg_off, n)) return set_err(ctx, P2PT_ERR_ARG, "sctp_packetize: offsets must be non-decreasing");
    const uint64_t hdr = epoch ? 13 : 0;
    std::vector<uint64_t> first(n + 1), moff(n);
    uint64_t pk = 0, bytes = 0;
    for (uint64_t i = 0; i < n; i++) {
        const uint64_t len = msg_off[i + 1] - msg_off[i], k = p2pt_sctp_fragments(len);
        first[i] = pk; moff[i] = bytes;
        pk += k; bytes += p2pt_sctp_packets_bytes(len) + hdr * k;
    }
    first[n] = pk;
    *n_pkts_out = pk;
    if (bytes > out_cap || pk + 1 > pkt_cap) return set_err(ctx, P2PT_ERR_CAPACITY, "sctp_packetize: out_cap or pkt_cap too small");
    const size_t mbytes = msg_off[n] - msg_off[0];
    if (mbytes && !msg) return set_err(ctx, P2PT_ERR_ARG, "sctp_packetize: null msg");
    DeviceGuard dg(ctx->device);
    LaneGuard lg(ctx);
    NvtxRange nvtx_("p2pt_sctp_packetize_batch");
    Lane* L = lg.lane;
    cudaStream_t st = L->streams[0];
    if (int32_t r = arena_reserve(ctx, L, pad256(mbytes) + 3 * pad256(8 * (n + 1)) + 4 * pad256(4 * n) + 4 * pad256(2 * n) +
                                           pad256(8 * n) + pad256(bytes) + pad256(8 * (pk + 1)) + 16384)) return r;
    uint8_t* d_msg = arena_take<uint8_t>(L, mbytes + 1);
    uint64_t* d_moff = arena_take<uint64_t>(L, n + 1);
    uint64_t* d_first = arena_take<uint64_t>(L, n + 1);
    uint64_t* d_omoff = arena_take<uint64_t>(L, n + 1);
    uint16_t* d_sp = arena_take<uint16_t>(L, n); uint16_t* d_dp = arena_take<uint16_t>(L, n);
    uint16_t* d_sid = arena_take<uint16_t>(L, n); uint16_t* d_ssn = arena_take<uint16_t>(L, n);
    uint32_t* d_vt = arena_take<uint32_t>(L, n); uint32_t* d_pp = arena_take<uint32_t>(L, n); uint32_t* d_tsn = arena_take<uint32_t>(L, n);
    uint16_t* d_ep = arena_take<uint16_t>(L, n); uint64_t* d_sq = arena_take<uint64_t>(L, n);
    uint8_t* d_out = arena_take<uint8_t>(L, bytes + 1);
    uint64_t* d_poff = arena_take<uint64_t>(L, pk + 1);
#define P2PT_UP(dst, src, cnt) CUDA_TRY(ctx, cudaMemcpyAsync(dst, src, (cnt), cudaMemcpyHostToDevice, st))
    if (mbytes) P2PT_UP(d_msg, msg + msg_off[0], mbytes);
    P2PT_UP(d_moff, msg_off, 8 * (n + 1)); P2PT_UP(d_first, first.data(), 8 * (n + 1)); P2PT_UP(d_omoff, moff.data(), 8 * n);
    P2PT_UP(d_sp, src_port, 2 * n); P2PT_UP(d_dp, dst_port, 2 * n); P2PT_UP(d_sid, sid, 2 * n); P2PT_UP(d_ssn, ssn, 2 * n);
    P2PT_UP(d_vt, vtag, 4 * n); P2PT_UP(d_pp, ppid, 4 * n); P2PT_UP(d_tsn, tsn0, 4 * n);
    if (epoch) { P2PT_UP(d_ep, epoch, 2 * n); P2PT_UP(d_sq, seq48_first, 8 * n); }
#undef P2PT_UP
    if (int32_t r = p2pt_sctp_packetize_batch_device(ctx, n, d_msg - msg_off[0], d_moff, d_sp, d_dp, d_vt, d_sid, d_ssn, d_pp, d_tsn,
                                                     epoch ? d_ep : nullptr, epoch ? d_sq : nullptr, d_out, d_omoff, d_first, d_poff, st)) return r;
    CUDA_TRY(ctx, cudaMemcpyAsync(out, d_out, bytes, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(pkt_off, d_poff, 8 * (pk + 1), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    lg.dismiss();
    return P2PT_OK;
}

int32_t p2pt_sctp_unpack_reassemble_batch(p2pt_ctx* ctx, uint64_t n_pkts, const uint8_t* in, const uint64_t* in_off,
                                          uint32_t headroom, const uint32_t* session, const uint32_t* tsn_base, p2pt_sctp_chunk* info,
                                          uint32_t max_msgs, uint8_t* msg, uint64_t msg_cap, uint64_t* msg_off,
                                          uint32_t* msg_first, uint8_t* msg_status, uint32_t* n_msgs_out) {
    if (!ctx) return P2PT_ERR_ARG;
    if (!msg_off || !msg_first || !msg_status || !n_msgs_out || !info) return set_err(ctx, P2PT_ERR_ARG, "sctp_unpack_reassemble: null output");
    if (n_pkts == 0) { *n_msgs_out = 0; msg_off[0] = 0; return P2PT_OK; }
    if (!in || !in_off || !msg) return set_err(ctx, P2PT_ERR_ARG, "sctp_unpack_reassemble: null argument");
    if (!offsets_monotonic(in_off, n_pkts)) return set_err(ctx, P2PT_ERR_ARG, "sctp_unpack_reassemble: offsets must be non-decreasing");
    DeviceGuard dg(ctx->device);
    LaneGuard lg(ctx);
    NvtxRange nvtx_("p2pt_sctp_unpack_reassemble_batch");
    Lane* L = lg.lane;
    cudaStream_t st = L->streams[0];
    const uint64_t n = n_pkts;
    const size_t ibytes = in_off[n] - in_off[0];
    const uint64_t sbytes = p2pt_sctp_reassemble_scratch_bytes(n);
    if (int32_t r = arena_reserve(ctx, L, pad256(ibytes) + pad256(8 * (n + 1)) + 2 * pad256(4 * n) + pad256(16 * n) + pad256(24 * n) +
                                           pad256(msg_cap) + pad256(8 * ((size_t)max_msgs + 1)) + pad256(4 * (size_t)max_msgs) +
                                           pad256(max_msgs) + 256 + pad256(sbytes) + 16384)) return r;
    uint8_t* d_in = arena_take<uint8_t>(L, ibytes + 1);
    uint64_t* d_ioff = arena_take<uint64_t>(L, n + 1);
    uint32_t* d_sess = arena_take<uint32_t>(L, n);
    uint32_t* d_base = arena_take<uint32_t>(L, n);
    p2pt_sctp_chunk* d_info = arena_take<p2pt_sctp_chunk>(L, n);
    p2pt_desc* d_desc = arena_take<p2pt_desc>(L, n);
    uint8_t* d_msg = arena_take<uint8_t>(L, msg_cap + 1);
    uint64_t* d_moff = arena_take<uint64_t>(L, (size_t)max_msgs + 1);
    uint32_t* d_mfirst = arena_take<uint32_t>(L, (size_t)max_msgs + 1);
    uint8_t* d_mstat = arena_take<uint8_t>(L, (size_t)max_msgs + 1);
    uint32_t* d_nm = arena_take<uint32_t>(L, 1);
    uint8_t* d_scratch = arena_take<uint8_t>(L, sbytes);
    CUDA_TRY(ctx, cudaMemcpyAsync(d_in, in + in_off[0], ibytes, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_ioff, in_off, 8 * (n + 1), cudaMemcpyHostToDevice, st));
    if (session) CUDA_TRY(ctx, cudaMemcpyAsync(d_sess, session, 4 * n, cudaMemcpyHostToDevice, st));
    if (tsn_base) CUDA_TRY(ctx, cudaMemcpyAsync(d_base, tsn_base, 4 * n, cudaMemcpyHostToDevice, st));
    if (int32_t r = p2pt_sctp_unpack_batch_device(ctx, n, d_in - in_off[0], d_ioff, headroom, tsn_base ? d_base : nullptr, d_info,
                                                  d_desc, st)) return r;
    if (int32_t r = p2pt_sctp_reassemble_batch_device(ctx, n, d_info, d_desc, session ? d_sess : nullptr, d_in - in_off[0],
                                                      max_msgs, d_msg, msg_cap, d_moff, d_mfirst, d_mstat, d_nm,
                                                      d_scratch, sbytes, st)) return r;
    uint32_t nm = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(&nm, d_nm, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(info, d_info, 16 * n, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    *n_msgs_out = nm;
    if (nm > max_msgs) return set_err(ctx, P2PT_ERR_CAPACITY, "sctp_unpack_reassemble: more messages than max_msgs");
    CUDA_TRY(ctx, cudaMemcpyAsync(msg_off, d_moff, 8 * ((size_t)nm + 1), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(msg_first, d_mfirst, 4 * (size_t)nm, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(msg_status, d_mstat, nm, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    if (msg_off[nm] > msg_cap) return set_err(ctx, P2PT_ERR_CAPACITY, "sctp_unpack_reassemble: msg_cap too small");
    if (msg_off[nm]) CUDA_TRY(ctx, cudaMemcpyAsync(msg, d_msg, msg_off[nm], cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    lg.dismiss();
    return P2PT_OK;
}

// ---------------------------------------------------------------------------------------------------
// SSE event scan
// ---------------------------------------------------------------------------------------------------
uint64_t p2pt_sse_scan_scratch_bytes(uint64_t k) { return align256(8 * (demux_scan_blocks(k ? k : 1) + 2)) + 256; }

int32_t p2pt_sse_scan_batch_device(p2pt_ctx* ctx, uint64_t k, const uint8_t* body, const uint64_t* body_off,
                                   uint64_t max_events, uint64_t* ev_csr, uint64_t* ev_end, uint8_t* done,
                                   uint64_t* n_events_out, void* scratch, uint64_t scratch_bytes, void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (!ev_csr || !n_events_out) return set_err(ctx, P2PT_ERR_ARG, "sse_scan: null output");
    DeviceGuard dg(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    if (k == 0) {
        CUDA_TRY(ctx, cudaMemsetAsync(n_events_out, 0, 8, st));
        CUDA_TRY(ctx, cudaMemsetAsync(ev_csr, 0, 8, st));
        return P2PT_OK;
    }
    if (!body_off || !ev_end || !done || !scratch) return set_err(ctx, P2PT_ERR_ARG, "sse_scan: null argument");
    if (scratch_bytes < p2pt_sse_scan_scratch_bytes(k)) return set_err(ctx, P2PT_ERR_CAPACITY, "sse_scan: scratch too small");
    u64* blk = (u64*)scratch;
    const u64 nblk = demux_scan_blocks(k);
    const unsigned grid = grid_for(ctx, k, 8, 16);
    sse_scan_kernel<false><<<grid, 256, 0, st>>>(k, body, (const u64*)body_off, (u64*)ev_csr, nullptr, 0, nullptr);
    u64_scan_blocksum_kernel<<<(unsigned)nblk, kScanThreads, 0, st>>>(k, (const u64*)ev_csr, blk);
    u64_scan_blocksums_kernel<<<1, kScanThreads, 0, st>>>(nblk, blk);
    u64_scan_apply_kernel<<<(unsigned)nblk, kScanThreads, 0, st>>>(k, nblk, (u64*)ev_csr, blk, (u64*)n_events_out);
    sse_scan_kernel<true><<<grid, 256, 0, st>>>(k, body, (const u64*)body_off, (u64*)ev_csr, (u64*)ev_end, max_events, done);
    count_launch(ctx, 5);
    CUDA_TRY(ctx, cudaGetLastError());
    return P2PT_OK;
}

namespace {
struct SseOnepassScratch { u64* state; u64* gstate; u64* gfinal; unsigned long long* ticket; uint32_t* tile_first; uint64_t zero_bytes; uint64_t total; };
SseOnepassScratch sse_onepass_carve(void* base, uint64_t ntiles) {
    SseOnepassScratch r;
    uint8_t* p = (uint8_t*)base;
    r.ticket = (unsigned long long*)p; p += 256;
    r.state = (u64*)p; p += align256(8 * ntiles);
    r.gstate = (u64*)p; p += align256(8 * (ntiles / 32 + 1));
    r.gfinal = (u64*)p; p += align256(8 * (ntiles / 32 + 1));
    r.zero_bytes = (uint64_t)(p - (uint8_t*)base);          // ticket + tile states: cleared per call
    r.tile_first = (uint32_t*)p; p += align256(4 * ntiles);
    r.total = (uint64_t)(p - (uint8_t*)base);
    return r;
}
}  // namespace

uint64_t p2pt_sse_scan_onepass_scratch_bytes(uint64_t k, uint64_t body_bytes) {
    (void)k;
    return sse_onepass_carve(nullptr, sse_onepass_tiles(body_bytes)).total;
}

int32_t p2pt_sse_scan_onepass_batch_device(p2pt_ctx* ctx, uint64_t k, const uint8_t* body, const uint64_t* body_off,
                                           uint64_t body_bytes, uint64_t max_events, uint64_t* ev_csr, uint64_t* ev_end,
                                           uint8_t* done, uint64_t* n_events_out, void* scratch, uint64_t scratch_bytes,
                                           void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (!ev_csr || !n_events_out) return set_err(ctx, P2PT_ERR_ARG, "sse_scan: null output");
    DeviceGuard dg(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    if (k == 0) {
        CUDA_TRY(ctx, cudaMemsetAsync(n_events_out, 0, 8, st));
        CUDA_TRY(ctx, cudaMemsetAsync(ev_csr, 0, 8, st));
        return P2PT_OK;
    }
    if (!body_off || !ev_end || !done || !scratch) return set_err(ctx, P2PT_ERR_ARG, "sse_scan: null argument");
    if (k > 0xfffffffeull) return set_err(ctx, P2PT_ERR_ARG, "sse_scan: k must fit in 32 bits");
    if (((uintptr_t)scratch & 255) != 0) return set_err(ctx, P2PT_ERR_ARG, "sse_scan: scratch must be 256-byte aligned");
    const u64 ntiles = sse_onepass_tiles(body_bytes);
    SseOnepassScratch r = sse_onepass_carve(scratch, ntiles);
    if (scratch_bytes < r.total) return set_err(ctx, P2PT_ERR_CAPACITY, "sse_scan: scratch too small");
    SseOnepassArgs a;
    a.k = k; a.body = body; a.body_off = (const u64*)body_off; a.ntiles = ntiles;
    a.ev_csr = (u64*)ev_csr; a.ev_end = (u64*)ev_end; a.cap = max_events; a.done = done; a.n_events_out = (u64*)n_events_out;
    a.state = r.state; a.gstate = r.gstate; a.gfinal = r.gfinal; a.ticket = r.ticket; a.tile_first = r.tile_first;
    CUDA_TRY(ctx, cudaMemsetAsync(scratch, 0, r.zero_bytes, st));
    CUDA_TRY(ctx, cudaMemsetAsync(r.tile_first, 0xff, 4 * ntiles, st));
    CUDA_TRY(ctx, cudaMemsetAsync(done, 0, k, st));
    sse_tile_first_kernel<<<grid_for(ctx, k + 1, 256, 8), 256, 0, st>>>(a);
    static std::atomic<int> occ{0};                      // resident blocks per SM (the same on every B200)
    int per_sm = occ.load();
    if (!per_sm) {
        CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sse_onepass_kernel, 256, 0));
        if (per_sm < 1) per_sm = 1;
        occ.store(per_sm);
    }
    const u64 resident = (u64)ctx->sm_count * (u64)per_sm;   // persistent: tiles are handed out by ticket
    sse_onepass_kernel<<<(unsigned)(ntiles < resident ? ntiles : resident), 256, 0, st>>>(a);
    count_launch(ctx, 2);
    CUDA_TRY(ctx, cudaGetLastError());
    return P2PT_OK;
}

int32_t p2pt_sse_scan_batch(p2pt_ctx* ctx, uint64_t k, const uint8_t* body, const uint64_t* body_off,
                            uint64_t max_events, uint64_t* ev_csr, uint64_t* ev_end, uint8_t* done,
                            uint64_t* n_events_out) {
    if (!ctx) return P2PT_ERR_ARG;
    if (!ev_csr || !n_events_out) return set_err(ctx, P2PT_ERR_ARG, "sse_scan: null output");
    if (k == 0) { *n_events_out = 0; ev_csr[0] = 0; return P2PT_OK; }
    if (!body_off || !ev_end || !done) return set_err(ctx, P2PT_ERR_ARG, "sse_scan: null argument");
    if (!offsets_monotonic(body_off, k)) return set_err(ctx, P2PT_ERR_ARG, "sse_scan: offsets must be non-decreasing");
    const size_t bytes = body_off[k] - body_off[0];
    if (bytes && !body) return set_err(ctx, P2PT_ERR_ARG, "sse_scan: null body");
    DeviceGuard dg(ctx->device);
    LaneGuard lg(ctx);
    NvtxRange nvtx_("p2pt_sse_scan_batch");
    Lane* L = lg.lane;
    cudaStream_t st = L->streams[0];
    const uint64_t sb = p2pt_sse_scan_onepass_scratch_bytes(k, bytes);
    if (int32_t r = arena_reserve(ctx, L, pad256(bytes) + 2 * pad256(8 * (k + 1)) + pad256(8 * (max_events + 1)) + pad256(k) + 256 + pad256(sb) + 8192)) return r;
    uint8_t* d_body = arena_take<uint8_t>(L, bytes + 1);
    uint64_t* d_off = arena_take<uint64_t>(L, k + 1);
    uint64_t* d_csr = arena_take<uint64_t>(L, k + 1);
    uint64_t* d_end = arena_take<uint64_t>(L, max_events + 1);
    uint8_t* d_done = arena_take<uint8_t>(L, k);
    uint64_t* d_n = arena_take<uint64_t>(L, 1);
    uint8_t* d_scr = arena_take<uint8_t>(L, sb);
    if (bytes) CUDA_TRY(ctx, cudaMemcpyAsync(d_body, body + body_off[0], bytes, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_off, body_off, 8 * (k + 1), cudaMemcpyHostToDevice, st));
    if (int32_t r = p2pt_sse_scan_onepass_batch_device(ctx, k, d_body - body_off[0], d_off, bytes, max_events, d_csr, d_end, d_done, d_n, d_scr, sb, st)) return r;
    uint64_t total = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(&total, d_n, 8, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(ev_csr, d_csr, 8 * (k + 1), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(done, d_done, k, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    *n_events_out = total;
    if (total > max_events) return set_err(ctx, P2PT_ERR_CAPACITY, "sse_scan: more events than max_events");
    if (total) CUDA_TRY(ctx, cudaMemcpyAsync(ev_end, d_end, 8 * total, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    lg.dismiss();
    return P2PT_OK;
}

// ---------------------------------------------------------------------------------------------------
// DTLS record-layer bookkeeping
// ---------------------------------------------------------------------------------------------------
uint64_t p2pt_dtls_datagram_split_scratch_bytes(uint64_t n) {
    const uint64_t nn = n ? n : 1;
    return align256(4 * nn) + align256(4 * (demux_scan_blocks(nn) + 2)) + 256;
}

int32_t p2pt_dtls_datagram_split_batch_device(p2pt_ctx* ctx, uint64_t n, const uint8_t* buf, const uint64_t* dgram_off,
                                              uint64_t max_records, uint64_t* rec_off, uint32_t* rec_dgram,
                                              uint8_t* dgram_status, uint32_t* n_records_out,
                                              void* scratch, uint64_t scratch_bytes, void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (!rec_off || !rec_dgram || !n_records_out) return set_err(ctx, P2PT_ERR_ARG, "datagram_split: null output");
    DeviceGuard dg(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) { CUDA_TRY(ctx, cudaMemsetAsync(n_records_out, 0, 4, st)); return P2PT_OK; }
    if (!buf || !dgram_off || !dgram_status || !scratch) return set_err(ctx, P2PT_ERR_ARG, "datagram_split: null argument");
    if (scratch_bytes < p2pt_dtls_datagram_split_scratch_bytes(n)) return set_err(ctx, P2PT_ERR_CAPACITY, "datagram_split: scratch too small");
    uint32_t* count = (uint32_t*)scratch;
    uint32_t* blk = (uint32_t*)((uint8_t*)scratch + align256(4 * n));
    const u64 nblk = demux_scan_blocks(n);
    dgram_count_kernel<<<(unsigned)nblk, kScanThreads, 0, st>>>(n, buf, (const u64*)dgram_off, count, dgram_status, blk);
    u32_scan_blocksums_kernel<<<1, kScanThreads, 0, st>>>(nblk, blk);
    dgram_emit_kernel<<<(unsigned)nblk, kScanThreads, 0, st>>>(n, nblk, buf, (const u64*)dgram_off, count, blk, max_records,
                                                              (u64*)rec_off, rec_dgram, n_records_out);
    count_launch(ctx, 3);
    CUDA_TRY(ctx, cudaGetLastError());
    return P2PT_OK;
}

namespace {
struct ReplayScratch { p2pt_desc* desc; uint32_t* uniq; uint64_t* csr; uint32_t* perm; uint32_t* k; uint8_t* demux; uint64_t demux_bytes; uint64_t total; };
ReplayScratch replay_carve(void* base, uint64_t n) {
    ReplayScratch r;
    uint8_t* p = (uint8_t*)base;
    const uint64_t nn = n ? n : 1;
    r.desc = (p2pt_desc*)p; p += align256(24 * nn);
    r.uniq = (uint32_t*)p; p += align256(4 * (nn + 1));
    r.csr = (uint64_t*)p; p += align256(8 * (nn + 1));
    r.perm = (uint32_t*)p; p += align256(4 * nn);
    r.k = (uint32_t*)p; p += 256;
    r.demux = p; r.demux_bytes = demux_scratch_bytes(nn); p += align256(r.demux_bytes);
    r.total = (uint64_t)(p - (uint8_t*)base);
    return r;
}
}  // namespace

uint64_t p2pt_dtls_replay_filter_scratch_bytes(uint64_t n) { return replay_carve(nullptr, n).total; }

int32_t p2pt_dtls_replay_filter_batch_device(p2pt_ctx* ctx, uint64_t n, const uint32_t* session, const uint64_t* seq48,
                                             const uint8_t* authentic, uint32_t n_sessions,
                                             uint64_t* latest, uint64_t* mask, uint8_t* started, uint8_t* accept,
                                             void* scratch, uint64_t scratch_bytes, void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!session || !seq48 || !latest || !mask || !started || !accept || !scratch)
        return set_err(ctx, P2PT_ERR_ARG, "replay_filter: null argument");
    if (n > 0xffffffffull) return set_err(ctx, P2PT_ERR_ARG, "replay_filter: n must fit in 32 bits");
    if (((uintptr_t)scratch & 255) != 0) return set_err(ctx, P2PT_ERR_ARG, "replay_filter: scratch must be 256-byte aligned");
    ReplayScratch r = replay_carve(scratch, n);
    if (scratch_bytes < r.total) return set_err(ctx, P2PT_ERR_CAPACITY, "replay_filter: scratch too small");
    DeviceGuard dg(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    replay_keys_kernel<<<grid_for(ctx, n, 256, 8), 256, 0, st>>>(n, session, r.desc);
    // group the records by session, arrival order kept inside a session (the stable demux, key = session)
    if (int32_t rc = p2pt_frame_demux_batch_device(ctx, n, r.desc, nullptr, nullptr, (uint32_t)n, r.uniq, nullptr, r.csr, r.perm,
                                                   r.k, nullptr, 0, nullptr, r.demux, r.demux_bytes, st)) return rc;
    replay_walk_kernel<<<grid_for(ctx, n, 256, 8), 256, 0, st>>>(r.k, r.uniq, (const u64*)r.csr, r.perm, (const u64*)seq48,
                                                                authentic, n_sessions, (u64*)latest, (u64*)mask, started, accept);
    count_launch(ctx, 2);
    CUDA_TRY(ctx, cudaGetLastError());
    return P2PT_OK;
}

uint64_t p2pt_dtls_seq_assign_scratch_bytes(uint64_t n) { return replay_carve(nullptr, n).total; }

int32_t p2pt_dtls_seq_assign_batch_device(p2pt_ctx* ctx, uint64_t n, const uint32_t* session, uint32_t n_sessions,
                                          uint64_t* next_seq, uint64_t* seq48, uint8_t* ok,
                                          void* scratch, uint64_t scratch_bytes, void* stream) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!session || !next_seq || !seq48 || !ok || !scratch) return set_err(ctx, P2PT_ERR_ARG, "seq_assign: null argument");
    if (n > 0xffffffffull) return set_err(ctx, P2PT_ERR_ARG, "seq_assign: n must fit in 32 bits");
    if (((uintptr_t)scratch & 255) != 0) return set_err(ctx, P2PT_ERR_ARG, "seq_assign: scratch must be 256-byte aligned");
    ReplayScratch r = replay_carve(scratch, n);
    if (scratch_bytes < r.total) return set_err(ctx, P2PT_ERR_CAPACITY, "seq_assign: scratch too small");
    DeviceGuard dg(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    replay_keys_kernel<<<grid_for(ctx, n, 256, 8), 256, 0, st>>>(n, session, r.desc);
    // the records of a session in arrival order (the stable demux, key = session)
    if (int32_t rc = p2pt_frame_demux_batch_device(ctx, n, r.desc, nullptr, nullptr, (uint32_t)n, r.uniq, nullptr, r.csr, r.perm,
                                                   r.k, nullptr, 0, nullptr, r.demux, r.demux_bytes, st)) return rc;
    seq_assign_kernel<<<grid_for(ctx, n, 256, 8), 256, 0, st>>>(n, r.k, r.uniq, (const u64*)r.csr, r.perm, n_sessions,
                                                                (const u64*)next_seq, (u64*)seq48, ok);
    seq_advance_kernel<<<grid_for(ctx, n, 256, 8), 256, 0, st>>>(r.k, r.uniq, (const u64*)r.csr, n_sessions, (u64*)next_seq);
    count_launch(ctx, 3);
    CUDA_TRY(ctx, cudaGetLastError());
    return P2PT_OK;
}

int32_t p2pt_dtls_datagram_split_batch(p2pt_ctx* ctx, uint64_t n, const uint8_t* buf, const uint64_t* dgram_off,
                                       uint64_t max_records, uint64_t* rec_off, uint32_t* rec_dgram,
                                       uint8_t* dgram_status, uint32_t* n_records_out) {
    if (!ctx) return P2PT_ERR_ARG;
    if (!rec_off || !rec_dgram || !n_records_out) return set_err(ctx, P2PT_ERR_ARG, "datagram_split: null output");
    if (n == 0) { *n_records_out = 0; return P2PT_OK; }
    if (!buf || !dgram_off || !dgram_status) return set_err(ctx, P2PT_ERR_ARG, "datagram_split: null argument");
    if (!offsets_monotonic(dgram_off, n)) return set_err(ctx, P2PT_ERR_ARG, "datagram_split: offsets must be non-decreasing");
    DeviceGuard dg(ctx->device);
    LaneGuard lg(ctx);
    NvtxRange nvtx_("p2pt_dtls_datagram_split_batch");
    Lane* L = lg.lane;
    cudaStream_t st = L->streams[0];
    const size_t bytes = dgram_off[n] - dgram_off[0];
    const uint64_t sb = p2pt_dtls_datagram_split_scratch_bytes(n);
    if (int32_t r = arena_reserve(ctx, L, pad256(bytes) + pad256(8 * (n + 1)) + pad256(8 * (max_records + 1)) + pad256(4 * max_records) +
                                           pad256(n) + 256 + pad256(sb) + 8192)) return r;
    uint8_t* d_buf = arena_take<uint8_t>(L, bytes + 1);
    uint64_t* d_off = arena_take<uint64_t>(L, n + 1);
    uint64_t* d_roff = arena_take<uint64_t>(L, max_records + 1);
    uint32_t* d_rd = arena_take<uint32_t>(L, max_records + 1);
    uint8_t* d_st = arena_take<uint8_t>(L, n);
    uint32_t* d_n = arena_take<uint32_t>(L, 1);
    uint8_t* d_scr = arena_take<uint8_t>(L, sb);
    CUDA_TRY(ctx, cudaMemcpyAsync(d_buf, buf + dgram_off[0], bytes, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_off, dgram_off, 8 * (n + 1), cudaMemcpyHostToDevice, st));
    if (int32_t r = p2pt_dtls_datagram_split_batch_device(ctx, n, d_buf - dgram_off[0], d_off, max_records, d_roff, d_rd, d_st, d_n,
                                                          d_scr, sb, st)) return r;
    uint32_t total = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(&total, d_n, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(dgram_status, d_st, n, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    *n_records_out = total;
    if (total > max_records) return set_err(ctx, P2PT_ERR_CAPACITY, "datagram_split: more records than max_records");
    CUDA_TRY(ctx, cudaMemcpyAsync(rec_off, d_roff, 8 * ((size_t)total + 1), cudaMemcpyDeviceToHost, st));
    if (total) CUDA_TRY(ctx, cudaMemcpyAsync(rec_dgram, d_rd, 4 * (size_t)total, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    lg.dismiss();
    return P2PT_OK;
}

int32_t p2pt_dtls_replay_filter_batch(p2pt_ctx* ctx, uint64_t n, const uint32_t* session, const uint64_t* seq48,
                                      const uint8_t* authentic, uint32_t n_sessions,
                                      uint64_t* latest, uint64_t* mask, uint8_t* started, uint8_t* accept) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!session || !seq48 || !latest || !mask || !started || !accept || !n_sessions)
        return set_err(ctx, P2PT_ERR_ARG, "replay_filter: null argument");
    DeviceGuard dg(ctx->device);
    LaneGuard lg(ctx);
    NvtxRange nvtx_("p2pt_dtls_replay_filter_batch");
    Lane* L = lg.lane;
    cudaStream_t st = L->streams[0];
    const uint64_t sb = p2pt_dtls_replay_filter_scratch_bytes(n);
    const size_t S = n_sessions;
    if (int32_t r = arena_reserve(ctx, L, pad256(4 * n) + pad256(8 * n) + 2 * pad256(n) + 2 * pad256(8 * S) + pad256(S) + pad256(sb) + 8192)) return r;
    uint32_t* d_sess = arena_take<uint32_t>(L, n);
    uint64_t* d_seq = arena_take<uint64_t>(L, n);
    uint8_t* d_auth = arena_take<uint8_t>(L, n);
    uint8_t* d_acc = arena_take<uint8_t>(L, n);
    uint64_t* d_lat = arena_take<uint64_t>(L, S);
    uint64_t* d_mask = arena_take<uint64_t>(L, S);
    uint8_t* d_started = arena_take<uint8_t>(L, S);
    uint8_t* d_scr = arena_take<uint8_t>(L, sb);
    CUDA_TRY(ctx, cudaMemcpyAsync(d_sess, session, 4 * n, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_seq, seq48, 8 * n, cudaMemcpyHostToDevice, st));
    if (authentic) CUDA_TRY(ctx, cudaMemcpyAsync(d_auth, authentic, n, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_lat, latest, 8 * S, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_mask, mask, 8 * S, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_started, started, S, cudaMemcpyHostToDevice, st));
    if (int32_t r = p2pt_dtls_replay_filter_batch_device(ctx, n, d_sess, d_seq, authentic ? d_auth : nullptr, n_sessions,
                                                         d_lat, d_mask, d_started, d_acc, d_scr, sb, st)) return r;
    CUDA_TRY(ctx, cudaMemcpyAsync(accept, d_acc, n, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(latest, d_lat, 8 * S, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(mask, d_mask, 8 * S, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(started, d_started, S, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    lg.dismiss();
    return P2PT_OK;
}

int32_t p2pt_dtls_seq_assign_batch(p2pt_ctx* ctx, uint64_t n, const uint32_t* session, uint32_t n_sessions,
                                   uint64_t* next_seq, uint64_t* seq48, uint8_t* ok) {
    if (!ctx) return P2PT_ERR_ARG;
    if (n == 0) return P2PT_OK;
    if (!session || !next_seq || !seq48 || !ok || !n_sessions) return set_err(ctx, P2PT_ERR_ARG, "seq_assign: null argument");
    DeviceGuard dg(ctx->device);
    LaneGuard lg(ctx);
    NvtxRange nvtx_("p2pt_dtls_seq_assign_batch");
    Lane* L = lg.lane;
    cudaStream_t st = L->streams[0];
    const uint64_t sb = p2pt_dtls_seq_assign_scratch_bytes(n);
    const size_t S = n_sessions;
    if (int32_t r = arena_reserve(ctx, L, pad256(4 * n) + pad256(8 * n) + pad256(n) + pad256(8 * S) + pad256(sb) + 8192)) return r;
    uint32_t* d_sess = arena_take<uint32_t>(L, n);
    uint64_t* d_seq = arena_take<uint64_t>(L, n);
    uint8_t* d_ok = arena_take<uint8_t>(L, n);
    uint64_t* d_next = arena_take<uint64_t>(L, S);
    uint8_t* d_scr = arena_take<uint8_t>(L, sb);
    CUDA_TRY(ctx, cudaMemcpyAsync(d_sess, session, 4 * n, cudaMemcpyHostToDevice, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_next, next_seq, 8 * S, cudaMemcpyHostToDevice, st));
    if (int32_t r = p2pt_dtls_seq_assign_batch_device(ctx, n, d_sess, n_sessions, d_next, d_seq, d_ok, d_scr, sb, st)) return r;
    CUDA_TRY(ctx, cudaMemcpyAsync(seq48, d_seq, 8 * n, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(ok, d_ok, n, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, cudaMemcpyAsync(next_seq, d_next, 8 * S, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(ctx, wait_stream(ctx, L, st));
    lg.dismiss();
    return P2PT_OK;
}

}  // extern "C"

// p2pt_pipeline_open_demux for a reader loop that drains a few frames per wake-up (n <= kSmallDemux and a
// staging footprint within small_call_bytes): inputs go through the mapped pinned buffer, the three kernels
// (open, single-block demux, gather) write the demux results straight back into it, only `desc` (and `out`
// if asked for) take an asynchronous copy, and there is ONE stream wait.  Entries beyond k are never
// copied to the caller, exactly like the general path.  *taken stays false if the call does not fit.
static int32_t open_demux_small(p2pt_ctx* ctx, Lane* L, uint64_t n, const uint32_t* session, const uint8_t* rec,
                                const uint64_t* rec_off, uint8_t* out, p2pt_desc* desc, uint32_t max_streams,
                                uint32_t* uniq_sid, uint32_t* uniq_session, uint64_t* csr, uint32_t* perm,
                                uint32_t* k_out, uint8_t* gathered, uint64_t gathered_cap, uint64_t* gathered_off,
                                bool* taken) {
    *taken = false;
    if (!ctx->small_demux || !L->zc_host || !ctx->small_call_bytes || n > (uint64_t)kSmallDemux) return P2PT_OK;
    const size_t ibytes = rec_off[n] - rec_off[0];
    const size_t ms1 = (size_t)(max_streams < n ? max_streams : n) + 1;
    const size_t gcap = gathered ? (size_t)(gathered_cap < ibytes ? gathered_cap : ibytes) : 0;
    const size_t foot = pad256(4 * n) + pad256(8 * (n + 1)) + pad256(ibytes + 1) + 2 * pad256(4 * ms1) + 2 * pad256(8 * ms1) +
                        pad256(4 * n) + pad256(gcap + 1) + 256;
    if (foot > ctx->small_call_bytes || foot > L->zc_cap) return P2PT_OK;
    *taken = true;
    const uint64_t sbytes = demux_scratch_bytes(n);
    if (int32_t r = arena_reserve(ctx, L, pad256(ibytes + 1) + pad256(24 * n) + pad256(sbytes) + 4096)) return r;
    uint8_t* d_out = arena_take<uint8_t>(L, ibytes + 1);
    p2pt_desc* d_desc = arena_take<p2pt_desc>(L, n);
    uint8_t* d_scratch = arena_take<uint8_t>(L, sbytes);
    ZcCarver zcv{L};
    auto zc = [&](size_t bytes) { return zcv.take(bytes); };
    auto host_of = [&](const void* d) { return zcv.host_of(d); };
    uint32_t* z_sess = (uint32_t*)zc(4 * n);
    uint64_t* z_io = (uint64_t*)zc(8 * (n + 1));
    uint8_t* z_in = zc(ibytes + 1);
    uint32_t* z_uniq = (uint32_t*)zc(4 * ms1);
    uint32_t* z_usess = (uint32_t*)zc(4 * ms1);
    uint64_t* z_csr = (uint64_t*)zc(8 * ms1);
    uint64_t* z_goff = (uint64_t*)zc(8 * ms1);
    uint32_t* z_perm = (uint32_t*)zc(4 * n);
    uint8_t* z_g = zc(gcap + 1);
    uint32_t* z_k = (uint32_t*)zc(4);
    memcpy(host_of(z_sess), session, 4 * n);
    memcpy(host_of(z_io), rec_off, 8 * (n + 1));
    if (ibytes) memcpy(host_of(z_in), rec + rec_off[0], ibytes);
    cudaStream_t st = L->streams[0];
    GcmArgs a = {};
    a.n = n; a.session = z_sess; a.n_sessions = ctx->n_sessions; a.keys = ctx->keys_remote;
    a.in = z_in - rec_off[0]; a.in_off = (const u64*)z_io;
    a.out = d_out - rec_off[0]; a.out_off = (const u64*)z_io;
    a.desc = d_desc;
    if (int32_t r = launch_gcm<kPipeOpen>(ctx, a, pick_group(ctx, n, ibytes), st)) return r;
    if (desc) CUDA_TRY(ctx, cudaMemcpyAsync(desc, d_desc, 24 * n, cudaMemcpyDeviceToHost, st));
    if (out && ibytes) CUDA_TRY(ctx, cudaMemcpyAsync(out + rec_off[0], d_out, ibytes, cudaMemcpyDeviceToHost, st));
    if (int32_t r = p2pt_frame_demux_batch_device(ctx, n, d_desc, z_sess, d_out - rec_off[0], max_streams, z_uniq,
                                                  uniq_session ? z_usess : nullptr, z_csr, z_perm, z_k,
                                                  gathered ? z_g : nullptr, gcap, gathered_off ? z_goff : nullptr,
                                                  d_scratch, sbytes, st)) return r;
    CUDA_TRY(ctx, cudaStreamSynchronize(st));
    const uint32_t k = *(const uint32_t*)host_of(z_k);
    *k_out = k;
    if (k > max_streams) return set_err(ctx, P2PT_ERR_CAPACITY, "pipeline_open_demux: more streams than max_streams");
    memcpy(uniq_sid, host_of(z_uniq), 4 * (size_t)k);
    if (uniq_session) memcpy(uniq_session, host_of(z_usess), 4 * (size_t)k);
    memcpy(csr, host_of(z_csr), 8 * ((size_t)k + 1));
    if (gathered_off) memcpy(gathered_off, host_of(z_goff), 8 * ((size_t)k + 1));
    const uint64_t m = csr[k];
    if (m && perm) memcpy(perm, host_of(z_perm), 4 * m);
    if (gathered) {
        if (gathered_off[k] > gathered_cap) return set_err(ctx, P2PT_ERR_CAPACITY, "pipeline_open_demux: gathered_cap too small");
        memcpy(gathered, host_of(z_g), gathered_off[k]);
    }
    return P2PT_OK;
}

extern "C" {

int32_t p2pt_pipeline_open_demux(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                                 const uint8_t* rec, const uint64_t* rec_off,
                                 uint8_t* out, p2pt_desc* desc,
                                 uint32_t max_streams, uint32_t* uniq_sid, uint32_t* uniq_session,
                                 uint64_t* csr, uint32_t* perm, uint32_t* k_out,
                                 uint8_t* gathered, uint64_t gathered_cap, uint64_t* gathered_off) {
    if (!ctx) return P2PT_ERR_ARG;
    if (!uniq_sid || !csr || !k_out) return set_err(ctx, P2PT_ERR_ARG, "pipeline_open_demux: null output");
    if (n == 0) { *k_out = 0; csr[0] = 0; if (gathered_off) gathered_off[0] = 0; return P2PT_OK; }
    if (!session || !rec || !rec_off) return set_err(ctx, P2PT_ERR_ARG, "pipeline_open_demux: null argument");
    if (gathered && !gathered_off) return set_err(ctx, P2PT_ERR_ARG, "pipeline_open_demux: gather needs gathered_off");
    if (n > 0xffffffffull) return set_err(ctx, P2PT_ERR_ARG, "pipeline_open_demux: n must fit in 32 bits");
    if (!offsets_monotonic(rec_off, n)) return set_err(ctx, P2PT_ERR_ARG, "pipeline_open_demux: offsets must be non-decreasing");
    if (int32_t r = check_sessions(ctx)) return r;
    DeviceGuard dg(ctx->device);
    LaneGuard lg(ctx);
    NvtxRange nvtx_("p2pt_pipeline_open_demux");
    Lane* L = lg.lane;
    {
        bool taken = false;
        const int32_t r = open_demux_small(ctx, L, n, session, rec, rec_off, out, desc, max_streams, uniq_sid, uniq_session, csr, perm,
                                           k_out, gathered, gathered_cap, gathered_off, &taken);
        if (taken) { if (r == P2PT_OK) lg.dismiss(); return r; }
    }
    const size_t ibytes = rec_off[n] - rec_off[0];
    const uint64_t sbytes = demux_scratch_bytes(n);
    const size_t gcap = gathered ? (size_t)gathered_cap : 0;
    if (int32_t r = arena_reserve(ctx, L, pad256(4 * n) + pad256(8 * (n + 1)) + 2 * pad256(ibytes) + pad256(24 * n) +
                                           2 * pad256(4 * ((size_t)max_streams + 1)) + 2 * pad256(8 * ((size_t)max_streams + 1)) +
                                           pad256(4 * n) + 256 + pad256(gcap) + pad256(sbytes) + 16384)) return r;
    uint32_t* d_sess = arena_take<uint32_t>(L, n);
    uint64_t* d_io = arena_take<uint64_t>(L, n + 1);
    uint8_t* d_in = arena_take<uint8_t>(L, ibytes + 1);
    uint8_t* d_out = arena_take<uint8_t>(L, ibytes + 1);
    p2pt_desc* d_desc = arena_take<p2pt_desc>(L, n);
    uint32_t* d_uniq = arena_take<uint32_t>(L, (size_t)max_streams + 1);
    uint32_t* d_usess = arena_take<uint32_t>(L, (size_t)max_streams + 1);
    uint64_t* d_csr = arena_take<uint64_t>(L, (size_t)max_streams + 1);
    uint64_t* d_goff = arena_take<uint64_t>(L, (size_t)max_streams + 1);
    uint32_t* d_perm = arena_take<uint32_t>(L, n);
    uint32_t* d_k = arena_take<uint32_t>(L, 1);
    uint8_t* d_g = arena_take<uint8_t>(L, gcap + 1);
    uint8_t* d_scratch = arena_take<uint8_t>(L, sbytes);

    cudaStream_t s0 = L->streams[0];
    CUDA_TRY(ctx, cudaMemcpyAsync(d_io, rec_off, 8 * (n + 1), cudaMemcpyHostToDevice, s0));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_sess, session, 4 * n, cudaMemcpyHostToDevice, s0));
    CUDA_TRY(ctx, cudaEventRecord(L->events[0], s0));
    const int G = pick_group(ctx, n, ibytes);
    uint64_t chunk = n;
    if (ibytes > 2 * ctx->chunk_bytes) {
        chunk = ctx->chunk_bytes / (ibytes / n + 1);
        if (chunk < 1024) chunk = 1024;
    }
    int si = 0;
    for (uint64_t r0 = 0; r0 < n; r0 += chunk, si = (si + 1) % kStreams) {
        const uint64_t r1 = r0 + chunk < n ? r0 + chunk : n;
        cudaStream_t st = L->streams[si];
        CUDA_TRY(ctx, cudaStreamWaitEvent(st, L->events[0], 0));
        const size_t b0 = rec_off[r0] - rec_off[0], b1 = rec_off[r1] - rec_off[0];
        if (b1 > b0) CUDA_TRY(ctx, cudaMemcpyAsync(d_in + b0, rec + rec_off[r0], b1 - b0, cudaMemcpyHostToDevice, st));
        GcmArgs a = {};
        a.n = r1 - r0; a.session = d_sess + r0; a.n_sessions = ctx->n_sessions; a.keys = ctx->keys_remote;
        a.in = d_in - rec_off[0]; a.in_off = (const u64*)(d_io + r0);
        a.out = d_out - rec_off[0]; a.out_off = (const u64*)(d_io + r0);
        a.desc = d_desc + r0;
        if (int32_t r = launch_gcm<kPipeOpen>(ctx, a, G, st)) return r;
        if (out && b1 > b0) CUDA_TRY(ctx, cudaMemcpyAsync(out + rec_off[r0], d_out + b0, b1 - b0, cudaMemcpyDeviceToHost, st));
    }
    for (int i = 1; i < kStreams; i++) {          // join the chunk streams into stream 0
        CUDA_TRY(ctx, cudaEventRecord(L->events[i], L->streams[i]));
        CUDA_TRY(ctx, cudaStreamWaitEvent(s0, L->events[i], 0));
    }
    if (desc) CUDA_TRY(ctx, cudaMemcpyAsync(desc, d_desc, 24 * n, cudaMemcpyDeviceToHost, s0));       // desc / perm only when asked for
    if (int32_t r = p2pt_frame_demux_batch_device(ctx, n, d_desc, d_sess, d_out - rec_off[0], max_streams, d_uniq,
                                                  uniq_session ? d_usess : nullptr, d_csr, d_perm, d_k,
                                                  gathered ? d_g : nullptr, gathered_cap,
                                                  gathered_off ? d_goff : nullptr, d_scratch, sbytes, s0)) return r;
    uint32_t k = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(&k, d_k, 4, cudaMemcpyDeviceToHost, s0));
    CUDA_TRY(ctx, wait_stream(ctx, L, s0));
    *k_out = k;
    if (k > max_streams) return set_err(ctx, P2PT_ERR_CAPACITY, "pipeline_open_demux: more streams than max_streams");
    CUDA_TRY(ctx, cudaMemcpyAsync(uniq_sid, d_uniq, 4 * (size_t)k, cudaMemcpyDeviceToHost, s0));
    if (uniq_session) CUDA_TRY(ctx, cudaMemcpyAsync(uniq_session, d_usess, 4 * (size_t)k, cudaMemcpyDeviceToHost, s0));
    CUDA_TRY(ctx, cudaMemcpyAsync(csr, d_csr, 8 * ((size_t)k + 1), cudaMemcpyDeviceToHost, s0));
    if (gathered_off) CUDA_TRY(ctx, cudaMemcpyAsync(gathered_off, d_goff, 8 * ((size_t)k + 1), cudaMemcpyDeviceToHost, s0));
    CUDA_TRY(ctx, wait_stream(ctx, L, s0));
    const uint64_t m = csr[k];
    if (m && perm) CUDA_TRY(ctx, cudaMemcpyAsync(perm, d_perm, 4 * m, cudaMemcpyDeviceToHost, s0));
    if (gathered) {
        if (gathered_off[k] > gathered_cap) return set_err(ctx, P2PT_ERR_CAPACITY, "pipeline_open_demux: gathered_cap too small");
        // split the big copy over the streams so both DMA engines stay busy
        const uint64_t total = gathered_off[k];
        const uint64_t piece = (total + kStreams - 1) / kStreams;
        CUDA_TRY(ctx, cudaEventRecord(L->events[0], s0));
        for (int i = 0; i < kStreams && piece; i++) {
            const uint64_t o = (uint64_t)i * piece;
            if (o >= total) break;
            const uint64_t len = o + piece <= total ? piece : total - o;
            CUDA_TRY(ctx, cudaStreamWaitEvent(L->streams[i], L->events[0], 0));
            CUDA_TRY(ctx, cudaMemcpyAsync(gathered + o, d_g + o, len, cudaMemcpyDeviceToHost, L->streams[i]));
        }
    }
    for (int i = 0; i < kStreams; i++) CUDA_TRY(ctx, wait_stream(ctx, L, L->streams[i]));
    lg.dismiss();
    return P2PT_OK;
}

}  // extern "C"
