// demux_direct_kernels.cuh -- demux without a sort, for batches that hold at most kDirectIds distinct streams.
//
// The reference's reader loops (/root/reference/tunnel/src/proxy.rs:109-170, serve.rs:82-156) decode a frame,
// probe a HashMap with its stream id and append the body to that stream's Vec<u8>.  The sorting demux
// (frame_kernels.cuh) pays nine passes' worth of machinery for an order the map never had; when the streams of a
// batch are few (an SSE token stream: 10 M frames over 4096 streams) the same result - streams ascending by
// (session, stream_id), arrival order inside a stream - comes out of ONE ranking pass over the frames:
//
//   rank     every (session, stream_id) gets a dense id from a hash table in global memory (L2-resident, 16-byte
//            slots, insert by compare-and-swap).  A block walks its contiguous share of the frames in arrival order
//            and keeps, per id, the running (frame count, body bytes) of ITS share in shared memory; a frame's
//            position inside (block, id) is that running value when its turn comes.  The order inside a block is kept
//            without block-wide barriers: warps take tiles of 128 consecutive frames, resolve ids and the prefix
//            inside each group of 32 on their own (ballots on the id bits), and only the read-modify-write of the
//            shared counters goes round the warps in turn, handed on through named barriers.
//   streams  a key's stream number = the number of keys smaller than it, its base = the sum of their totals: all
//            pairs of the <= 8192 distinct keys, spread over the machine (no sort of the keys either).
//   prefix   per id an exclusive scan of the per-block counters over the blocks, started at the stream's base;
//            writes uniq_sid / uniq_session / csr / gathered_off.
//   gather   frames in arrival order (each DRAM sector read once): destination = matrix[block][id] + position.
//
// A batch that does not fit (more distinct keys than kDirectIds, 4 GiB or more of bodies, the one key the table
// cannot hold) raises a flag in device memory; the sorting path's kernels, enqueued behind these, run only then.
#pragma once
#include "frame_kernels.cuh"

namespace p2pt {

constexpr int kDirectIds = 8192;                    // distinct (session, stream_id) keys the direct path can hold
constexpr int kDirectSlots = 4 * kDirectIds;        // hash slots (load factor <= 0.25: most keys are found by the first probe)
#ifndef P2PT_DIRECT_WARPS
#define P2PT_DIRECT_WARPS 16
#endif
constexpr int kDirectWarps = P2PT_DIRECT_WARPS;     // 8 (three blocks per SM: 24 warps, 72 registers) or 16 (two blocks: 32 warps at
                                                    // 60 registers, and 128 KiB of L1 left beside the counters): C4 rank 0.173 -> 0.153 ms
constexpr int kDirectBlocksPerSm = kDirectWarps == 8 ? 3 : 2;
constexpr int kDirectRounds = 4;                    // groups of 32 consecutive frames per warp and tile
constexpr int kDirectThreads = 32 * kDirectWarps;
constexpr int kDirectTile = kDirectThreads * kDirectRounds;      // frames per block iteration (2048 with 16 warps)
constexpr int kDirectMaxBlocks = 148 * kDirectBlocksPerSm;      // every block resident
constexpr uint32_t kNoLoc = 0xffffffffu;            // per-frame position: contributes nothing to the gather
constexpr uint16_t kNoId = 0xffff;                  // per-frame id: not decodable
constexpr uint32_t kDirectNone = 0xffffffffu;

__host__ __device__ constexpr u64 direct_blocks(u64 n) {
    const u64 g = (n + 4 * kDirectTile - 1) / (4 * kDirectTile);
    return g < 1 ? 1 : g > (u64)kDirectMaxBlocks ? (u64)kDirectMaxBlocks : g;
}
// frames per block: a multiple of the tile, so that a gather warp's 32 frames never straddle two blocks
__host__ __device__ constexpr u64 direct_per_block(u64 n) {
    const u64 g = direct_blocks(n);
    return ((n + g - 1) / g + kDirectTile - 1) / kDirectTile * kDirectTile;
}

constexpr u64 kDirectZeroBytes = 256 + (8ull + 4 + 8) * kDirectIds + 16ull * kDirectSlots;   // control words, totals, stream numbers, bases, hash slots
constexpr u64 direct_scratch_bytes(u64 n) {
    return kDirectZeroBytes + 8ull * kDirectIds + align256(4ull * kDirectMaxBlocks) + direct_blocks(n) * 8ull * kDirectIds;
}

// everything a demux call may need: the sorting path's arrays, then the direct path's
constexpr u64 demux_scratch_bytes(u64 n) { return align256(demux_sort_scratch_bytes(n)) + direct_scratch_bytes(n); }

struct DirectScratch {
    uint32_t* ctl;      // [0] fail flag (the sorting path runs iff set), [1] ids handed out, [2..3] body bytes of the batch (u64)
    u64* tot;           // per id: (frames << 32 | body bytes) over the whole batch
    uint32_t* ord;      // per id: how many keys are smaller than its key = its stream number
    uint4* slots;       // hash table: x,y = ~key (0 = empty), z = id + 1 (0 = not published yet)
    u64* idkey;         // per id: its key
    u64* base;          // per id: (csr << 32 | gathered_off) of its stream = the totals of all smaller keys
    uint32_t* rowmax;   // per block: ids [0, rowmax) of its matrix row are written
    u64* matrix;        // [block][id]: the block's own (frames, bytes) first, then (after prefix) where its share starts
    uint32_t* loc;      // per frame: body position inside (block, id), kNoLoc = nothing to copy
    uint32_t* rank;     // per frame: frame position inside (block, id)
    uint16_t* fid;      // per frame: dense id, kNoId = not decodable
};

// region = the bytes behind the sorting path's scratch; the per-frame arrays alias the sort's ping-pong buffers
// (only one of the two paths ever uses them in a call)
inline DirectScratch direct_carve(void* region, u64 n, const DemuxScratch& sort) {
    uint8_t* p = (uint8_t*)region;
    DirectScratch s;
    s.ctl = (uint32_t*)p; p += 256;
    s.tot = (u64*)p; p += 8ull * kDirectIds;
    s.ord = (uint32_t*)p; p += 4ull * kDirectIds;
    s.base = (u64*)p; p += 8ull * kDirectIds;
    s.slots = (uint4*)p; p += 16ull * kDirectSlots;
    s.idkey = (u64*)p; p += 8ull * kDirectIds;
    s.rowmax = (uint32_t*)p; p += align256(4ull * kDirectMaxBlocks);
    s.matrix = (u64*)p;
    s.loc = reinterpret_cast<uint32_t*>(sort.keys[0]);
    s.rank = reinterpret_cast<uint32_t*>(sort.keys[1]);
    s.fid = reinterpret_cast<uint16_t*>(sort.idx[0]);
    (void)n;
    return s;
}

struct DirectArgs {
    u64 n, per;
    const uint8_t* in; const u64* in_off;      // frames (fused decode) ...
    const p2pt_desc* desc_in;                  // ... or descriptors
    p2pt_desc* desc_out;                       // fused decode: descriptors wanted by the caller (nullable)
    const uint32_t* session;
    DirectScratch s;
    int want_rank;
};

__device__ __forceinline__ uint4 ld_slot(const uint4* p) {
    uint4 r;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ uint32_t ld_u32_volatile(const uint32_t* p) {
    uint32_t r;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(r) : "l"(p) : "memory");
    return r;
}
__device__ __forceinline__ void st_u32_volatile(uint32_t* p, uint32_t v) {
    asm volatile("st.volatile.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
// two 32-bit multiplicative hashes (stream id, session), top bits: a handful of instructions instead of a 64-bit mixer
__device__ __forceinline__ uint32_t direct_hash(u64 k) {
    const uint32_t h = (uint32_t)k * 0x9E3779B1u ^ (uint32_t)(k >> 32) * 0x85EBCA77u;
    return h >> 17;                                        // the top bits of a multiplicative hash are the good ones
}
static_assert(kDirectSlots == (1 << 15), "direct_hash returns 15 bits");
// named-barrier hand-over of the counters from one warp to the next: one barrier per warp.  With 8 warps ids 1 .. 8
// (0 is __syncthreads); with 16 warps all sixteen, id 0 being free between the kernel's first and last __syncthreads
constexpr int kTurnBase = kDirectWarps == 16 ? 0 : 1;
__device__ __forceinline__ void turn_wait(int warp) { asm volatile("bar.sync %0, 64;" :: "r"(kTurnBase + warp) : "memory"); }
__device__ __forceinline__ void turn_pass(int next_warp) { asm volatile("bar.arrive %0, 64;" :: "r"(kTurnBase + next_warp) : "memory"); }

// One look at `slot` (its contents v just loaded) for `key`: 0 = resolved (id set), 1 = look at the same slot again
// (its id is not published yet, or somebody took the slot in between), 2 = another key lives there, try the next slot.
__device__ __forceinline__ int direct_probe(const DirectScratch& s, u64 key, uint32_t slot, const uint4& v, uint32_t& id) {
    const u64 nk = ~key;
    const u64 sk = ((u64)v.y << 32) | v.x;
    if (sk == nk) {
        if (v.z) { id = v.z - 1; return 0; }
        return 1;
    }
    if (sk != 0) return 2;
    const u64 prev = atomicCAS(reinterpret_cast<u64*>(s.slots + slot), 0ull, nk);
    if (prev != 0) return 1;
    const uint32_t nid = atomicAdd(s.ctl + 1, 1u);
    if (nid < (uint32_t)kDirectIds) s.idkey[nid] = key;
    else st_u32_volatile(s.ctl, 1u);                                  // too many streams: the sorting path takes over
    st_u32_volatile(reinterpret_cast<uint32_t*>(s.slots + slot) + 2, nid + 1);
    id = nid;
    return 0;
}

// Dense ids for up to kDirectRounds keys per lane.  Every lane of the warp calls; need[r] marks the keys to resolve.
// The first probe of every key is straight-line code with all loads in flight together and finds nearly all of
// them (a key is inserted once and looked up thousands of times); what is left - collisions, inserts - goes
// through a loop that takes ONE pending key per lane and step.  No lane ever waits for another inside a step (a key
// whose id is not published yet is simply looked at again in the next step), so a lane that depends on an insert
// made by another lane of the same warp cannot starve it.
__device__ __forceinline__ void direct_lookup(const bool (&need)[kDirectRounds], const u64 (&key)[kDirectRounds],
                                              uint32_t (&id)[kDirectRounds], const DirectScratch& s) {
    constexpr int R = kDirectRounds;
    uint32_t slot[R];
    uint4 v[R];
    uint32_t pend = 0;
#pragma unroll
    for (int r = 0; r < R; r++) {
        id[r] = kDirectNone;
        slot[r] = direct_hash(key[r]);
        v[r] = make_uint4(0, 0, 0, 0);
        if (need[r]) v[r] = ld_slot(s.slots + slot[r]);
    }
#pragma unroll
    for (int r = 0; r < R; r++) {
        const u64 nk = ~key[r];
        const u64 sk = ((u64)v[r].y << 32) | v[r].x;
        if (need[r]) {
            if (nk == 0) st_u32_volatile(s.ctl, 1u);                   // the one key a slot cannot hold
            else if (sk == nk && v[r].z) id[r] = v[r].z - 1;
            else {
                pend |= 1u << r;
                if (sk != nk && sk != 0) slot[r] = (slot[r] + 1) & (kDirectSlots - 1);
            }
        }
    }
    uint32_t probes = 0;
    while (__any_sync(0xffffffffu, pend != 0)) {
        if (pend) {
            const int r = __ffs(pend) - 1;
            u64 k = key[0];
            uint32_t sl = slot[0];
#pragma unroll
            for (int q = 1; q < R; q++) if (r == q) { k = key[q]; sl = slot[q]; }
            const uint4 vv = ld_slot(s.slots + sl);
            uint32_t got = kDirectNone;
            int st = direct_probe(s, k, sl, vv, got);
            if (st == 2) {
                sl = (sl + 1) & (kDirectSlots - 1);
                if ((++probes & 15u) == 0 && (probes > (uint32_t)kDirectSlots || ld_u32_volatile(s.ctl))) {
                    st_u32_volatile(s.ctl, 1u); st = 0;                // table full, or the batch already failed: give up
                }
            }
#pragma unroll
            for (int q = 0; q < R; q++) if (r == q) { slot[q] = sl; if (st == 0) id[q] = got; }
            if (st == 0) pend &= pend - 1;
        }
    }
}

template <bool FROM_FRAMES>
__global__ void __launch_bounds__(kDirectThreads, kDirectBlocksPerSm) demux_direct_rank_kernel(const DirectArgs a) {
    extern __shared__ __align__(16) u64 bins[];         // per id: (frames << 32 | body bytes) of this block so far
    __shared__ uint32_t s_maxid;
    for (int j = threadIdx.x; j < kDirectIds; j += kDirectThreads) bins[j] = 0;
    if (threadIdx.x == 0) s_maxid = 0;
    __syncthreads();
    constexpr int R = kDirectRounds;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t lt = (1u << lane) - 1u;
    const u64 lo = (u64)blockIdx.x * a.per;
    const u64 hi = lo + a.per < a.n ? lo + a.per : a.n;
    const uint32_t ntiles = lo < hi ? (uint32_t)((hi - lo + kDirectTile - 1) / kDirectTile) : 0u;
    uint32_t maxid1 = 0;
    u64 mybytes = 0;
    for (uint32_t t = 0; t < ntiles; t++) {
        const u64 w0 = lo + (u64)t * kDirectTile + (u64)warp * (32 * R);
        // once the batch is known not to fit, the warps only keep the hand-over chain going
        const bool failed = __shfl_sync(0xffffffffu, ld_u32_volatile(a.s.ctl), 0) != 0;
        u64 key[R];
        uint32_t body[R];
        bool valid[R], live[R];
        // ---- this lane's frame of every round, in stages without branches in between: the loads of a stage are
        //      all in flight together (offsets -> header words -> decode; or descriptor -> decode)
        u64 ii[R];
        uint32_t sess[R];
#pragma unroll
        for (int r = 0; r < R; r++) {
            const u64 i = w0 + 32 * r + lane;
            live[r] = !failed && i < hi;
            ii[r] = live[r] ? i : lo;                             // a frame that exists, so that nothing below needs a branch
            sess[r] = a.session ? __ldg(a.session + ii[r]) : 0u;
        }
        if (FROM_FRAMES) {
            u64 o[R], len[R];
#pragma unroll
            for (int r = 0; r < R; r++) { o[r] = __ldg(a.in_off + ii[r]); len[r] = __ldg(a.in_off + ii[r] + 1) - o[r]; }
            uint32_t x0[R], x1[R], tb[R];
#pragma unroll
            for (int r = 0; r < R; r++) {
                // the five header bytes lie inside the two aligned words that start at the frame's word
                const uint8_t* f = a.in + o[r];
                const uint32_t* w = reinterpret_cast<const uint32_t*>(f - ((uintptr_t)f & 3));
                x0[r] = x1[r] = tb[r] = 0;
                if (len[r] >= 5) { x0[r] = __ldg(w); x1[r] = __ldg(w + 1); }
                else if (len[r]) tb[r] = __ldg(f);
            }
#pragma unroll
            for (int r = 0; r < R; r++) {
                const unsigned sh = (unsigned)((uintptr_t)(a.in + o[r]) & 3);
                const u64 x = (((u64)x1[r] << 32) | x0[r]) >> (8 * sh);
                const bool full = len[r] >= 5;                                   // protocol.rs:158-160
                const uint32_t type = full ? ((uint32_t)x & 0xffu) : tb[r];
                const bool okh = full && msg_type_valid(type);                  // protocol.rs:161-163
                const uint32_t sid = okh ? __byte_perm((uint32_t)(x >> 8), 0, 0x0123) : 0u;     // :165 get_u32 (BE)
                const uint32_t plen = okh ? (uint32_t)(len[r] - 5) : 0u;
                if (a.desc_out && live[r]) {
                    const u64 poff = okh ? o[r] + 5 : o[r];
                    const uint32_t status = !full ? P2PT_ST_TOO_SHORT : okh ? P2PT_ST_OK : P2PT_ST_UNKNOWN_TYPE;
                    uint8_t* d = reinterpret_cast<uint8_t*>(a.desc_out + ii[r]);
                    *reinterpret_cast<uint2*>(d) = make_uint2((uint32_t)poff, (uint32_t)(poff >> 32));
                    *reinterpret_cast<uint2*>(d + 8) = make_uint2(plen, sid);
                    *reinterpret_cast<uint2*>(d + 16) = make_uint2(type | (status << 8), 0u);
                }
                valid[r] = live[r] && okh;
                key[r] = ((u64)sess[r] << 32) | sid;
                body[r] = (valid[r] && (type == 11 || type == 21)) ? plen : 0u;  // ReqBody / ResBody (serve.rs:120-124)
            }
        } else {
#pragma unroll
            for (int r = 0; r < R; r++) {
                const uint2* d = reinterpret_cast<const uint2*>(a.desc_in + ii[r]);
                const uint2 d1 = __ldg(d + 1), d2 = __ldg(d + 2);
                const uint32_t type = d2.x & 0xffu;
                valid[r] = live[r] && ((d2.x >> 8) & 0xffu) == P2PT_ST_OK;
                key[r] = ((u64)sess[r] << 32) | d1.y;
                body[r] = (valid[r] && (type == 11 || type == 21)) ? d1.x : 0u;
            }
        }
        // ---- dense ids first (every decodable frame looks its key up), then the groups of equal ids inside each
        //      round of 32 from thirteen ballots.  MATCH.ANY on the 64-bit keys was the bottleneck of this kernel:
        //      its unit takes about 8 cycles per distinct value and warp, and SSE frames of 4096 interleaved
        //      streams are all distinct (measured 260 us of MATCH unit time for 10 M frames); ballots cost the same
        //      whatever the data.
        uint32_t id[R];
#pragma unroll
        for (int r = 0; r < R; r++) mybytes += body[r];
        direct_lookup(valid, key, id, a.s);
        uint32_t peers[R];
        u64 pre[R];
        bool ok[R];
#pragma unroll
        for (int r = 0; r < R; r++) {
            ok[r] = valid[r] && id[r] < (uint32_t)kDirectIds;
            // (A cheaper test for "no two lanes share an id" - every lane writes its number to a per-id tag in shared
            // memory and reads it back - measured slower: the 8 KiB of tags push the block over the 196 KiB shared-memory
            // carve-out and L1 shrinks from 60 to 28 KiB: +73 us.  So did L2 prefetches of the next tile: no change.)
            uint32_t pm = __ballot_sync(0xffffffffu, ok[r]), pb = 0;
#pragma unroll
            for (int b = 0; b < 13; b++) {
                static_assert(kDirectIds <= (1 << 13), "thirteen id bits");
                const uint32_t bm = __ballot_sync(0xffffffffu, (id[r] >> b) & 1u);
                pm &= bm ^ (((id[r] >> b) & 1u) - 1u);            // lanes whose bit b equals mine
            }
            // body bytes of the group in front of this lane: only lanes of groups with several members matter
            for (uint32_t m = __ballot_sync(0xffffffffu, ok[r] && (pm & (pm - 1)) != 0); m; m &= m - 1) {
                const int src = __ffs(m) - 1;
                const uint32_t v = __shfl_sync(0xffffffffu, body[r], src);
                if (((pm >> src) & 1u) && src < lane) pb += v;
            }
            peers[r] = ok[r] ? pm : 0u;
            pre[r] = ((u64)__popc(peers[r] & lt) << 32) | pb;
            if (ok[r] && id[r] + 1 > maxid1) maxid1 = id[r] + 1;
        }
        // ---- my turn at the block's counters: warps in order, tiles in order
        __syncwarp();
        if (t | (uint32_t)warp) turn_wait(warp);
        u64 res[R];
#pragma unroll
        for (int r = 0; r < R; r++) {
            res[r] = 0;
            if (ok[r]) res[r] = bins[id[r]] + pre[r];
            __syncwarp();
            if (ok[r] && (peers[r] >> lane) == 1u) bins[id[r]] = res[r] + ((1ull << 32) | body[r]);    // highest lane of the group
            __syncwarp();
        }
        __syncwarp();                                     // bar.arrive orders the counter stores before the next warp's loads
        if (!(t == ntiles - 1 && warp == kDirectWarps - 1)) turn_pass(warp + 1 == kDirectWarps ? 0 : warp + 1);
        // ---- per-frame results
#pragma unroll
        for (int r = 0; r < R; r++) {
            if (!live[r]) continue;
            a.s.fid[ii[r]] = ok[r] ? (uint16_t)id[r] : kNoId;
            a.s.loc[ii[r]] = (ok[r] && body[r]) ? (uint32_t)res[r] : kNoLoc;
            if (a.want_rank) a.s.rank[ii[r]] = (uint32_t)(res[r] >> 32);
        }
    }
    // ---- this block's row of the matrix, its share of the totals
    maxid1 = __reduce_max_sync(0xffffffffu, maxid1);
    if (lane == 0 && maxid1) atomicMax(&s_maxid, maxid1);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mybytes += __shfl_xor_sync(0xffffffffu, mybytes, o);
    if (lane == 0 && mybytes) atomicAdd(reinterpret_cast<u64*>(a.s.ctl + 2), mybytes);
    __syncthreads();
    const uint32_t mx = s_maxid;
    u64* row = a.s.matrix + (u64)blockIdx.x * kDirectIds;
    for (uint32_t j = threadIdx.x; j < mx; j += kDirectThreads) {
        const u64 v = bins[j];
        row[j] = v;
        if (v) atomicAdd(a.s.tot + j, v);
    }
    if (threadIdx.x == 0) a.s.rowmax[blockIdx.x] = mx;
}

// Stream numbers and bases without a sort: a key's stream number is the number of keys smaller than it, its base
// (csr << 32 | gathered_off) the sum of their totals.  Blocks of 256 keys compare against chunks of 128 keys - all
// pairs, spread over the machine: 4096 keys are 512 blocks of 128 comparisons per thread - and add what they find
// with atomics.  (A bitonic sort of the keys in one block took 63 us for 4096 keys; this takes 9.5.)
constexpr int kOrderT = 256;
constexpr int kOrderChunk = 128;           // keys a block compares its 256 keys against per step
__global__ void __launch_bounds__(kOrderT) demux_direct_streams_kernel(
    const DirectScratch s, u64* __restrict__ csr, uint32_t* __restrict__ k_out, u64* __restrict__ gathered_off) {
    __shared__ u64 ck[kOrderChunk], ct[kOrderChunk];
    const uint32_t tid = threadIdx.x;
    const bool first_block = blockIdx.x == 0 && blockIdx.y == 0;
    const uint32_t nid = s.ctl[1];
    const u64 total = *reinterpret_cast<const u64*>(s.ctl + 2);
    if (s.ctl[0] || nid > (uint32_t)kDirectIds || total > 0xffffffffull) {      // the same for every block, whoever looks first
        if (first_block && tid == 0) s.ctl[0] = 1;
        return;
    }
    if (first_block && tid == 0) {
        *k_out = nid;
        if (nid == 0) { csr[0] = 0; if (gathered_off) gathered_off[0] = 0; }
    }
    const uint32_t nb = (nid + kOrderT - 1) / kOrderT;
    if (blockIdx.x >= nb) return;
    const uint32_t id = blockIdx.x * kOrderT + tid;
    const u64 key = id < nid ? s.idkey[id] : 0ull;
    uint32_t cnt = 0;
    u64 sum = 0;                                                 // (frames << 32 | bytes) sums cannot carry: both totals fit 32 bits
    const uint32_t nchunks = (nid + kOrderChunk - 1) / kOrderChunk;
    for (uint32_t c = blockIdx.y; c < nchunks; c += gridDim.y) {
        const uint32_t j0 = c * kOrderChunk;
        __syncthreads();
        if (tid < kOrderChunk) {
            ck[tid] = j0 + tid < nid ? s.idkey[j0 + tid] : ~0ull;    // no real key is all ones (direct_lookup refuses it)
            ct[tid] = j0 + tid < nid ? s.tot[j0 + tid] : 0ull;
        }
        __syncthreads();
#pragma unroll 16
        for (int j = 0; j < kOrderChunk; j++) {
            if (ck[j] < key) { cnt++; sum += ct[j]; }              // two compares and three predicated adds per pair
        }
    }
    if (id < nid && cnt) { atomicAdd(s.ord + id, cnt); atomicAdd(s.base + id, sum); }
}

// Per id: exclusive scan of the blocks' counters, started at the stream's base.  32 ids x 32 chunks of blocks per
// thread block; a thread holds its chunk (<= 14 blocks) in registers.
constexpr int kDirectChunk = (kDirectMaxBlocks + 31) / 32;
__global__ void __launch_bounds__(1024) demux_direct_prefix_kernel(
    const DirectScratch s, uint32_t nblocks, uint32_t max_streams, uint32_t* __restrict__ uniq_sid,
    uint32_t* __restrict__ uniq_session, u64* __restrict__ csr, u64* __restrict__ gathered_off) {
    if (s.ctl[0]) return;
    const uint32_t nid = s.ctl[1];
    if (blockIdx.x * 32u >= nid) return;
    __shared__ uint32_t rm[kDirectMaxBlocks];
    __shared__ u64 tots[32][33];
    for (uint32_t g = threadIdx.x; g < nblocks; g += 1024) rm[g] = s.rowmax[g];
    __syncthreads();
    const uint32_t x = threadIdx.x & 31, y = threadIdx.x >> 5;
    const uint32_t id = blockIdx.x * 32u + x;
    const uint32_t c = (nblocks + 31) / 32, g0 = y * c;
    u64 v[kDirectChunk];
#pragma unroll
    for (int q = 0; q < kDirectChunk; q++) {
        const uint32_t g = g0 + q;
        v[q] = ((uint32_t)q < c && g < nblocks && id < rm[g]) ? s.matrix[(u64)g * kDirectIds + id] : 0ull;
    }
    u64 run = 0;
#pragma unroll
    for (int q = 0; q < kDirectChunk; q++) { const u64 t = v[q]; v[q] = run; run += t; }
    tots[y][x] = run;
    __syncthreads();
    u64 off = id < nid ? s.base[id] : 0ull;
    if (y == 0 && id < nid) {                                    // the stream table entry of this id
        const uint32_t p = s.ord[id];
        const u64 k = s.idkey[id];
        if (p < max_streams) {
            uniq_sid[p] = (uint32_t)k;
            if (uniq_session) uniq_session[p] = (uint32_t)(k >> 32);
            csr[p] = off >> 32;
            if (gathered_off) gathered_off[p] = off & 0xffffffffull;
        }
        if (p == nid - 1 && nid <= max_streams) {                // the largest key closes both offset arrays
            const u64 all = off + s.tot[id];
            csr[nid] = all >> 32;
            if (gathered_off) gathered_off[nid] = all & 0xffffffffull;
        }
    }
    for (uint32_t yy = 0; yy < y; yy++) off += tots[yy][x];
#pragma unroll
    for (int q = 0; q < kDirectChunk; q++) {
        const uint32_t g = g0 + q;
        if ((uint32_t)q < c && g < nblocks && id < rm[g]) s.matrix[(u64)g * kDirectIds + id] = off + v[q];
    }
}

// Frames in arrival order -> perm and the per-stream bodies.  Same tiles and copies as demux_gather_kernel: lane l
// holds the metadata of frame l of the tile; four lanes per short frame in rounds of eight frames, the whole warp
// per long one.  When every frame of the tile is at most 64 bytes and lands on whole aligned vectors (bodies that
// are multiples of 16 bytes - the fixed-size SSE chunk of BASELINE.json config #4) the four rounds issue all their
// loads before the first store: taken one after the other each round waited for its own loads, and the kernel sat
// at 81 % long-scoreboard stalls.  (A frame per LANE - all 32 in flight - measured ten times slower: 3.8 ms.)
template <bool FROM_FRAMES>
#ifndef P2PT_DIRECT_GATHER_BLOCKS
#define P2PT_DIRECT_GATHER_BLOCKS P2PT_GATHER_MIN_BLOCKS
#endif
__global__ void __launch_bounds__(256, P2PT_DIRECT_GATHER_BLOCKS) demux_direct_gather_kernel(
    const DirectArgs a, const uint8_t* __restrict__ in, uint8_t* __restrict__ gathered, u64 gathered_cap,
    uint32_t* __restrict__ perm, const uint32_t* __restrict__ k_out, uint32_t max_streams, int rounds) {
    if (a.s.ctl[0]) return;
    if (*k_out > max_streams) return;
    const int lane = threadIdx.x & 31, quad = lane >> 2, q = lane & 3;
    const u64 warp = ((u64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u64 nwarps = ((u64)gridDim.x * blockDim.x) >> 5;
    const int tile = 8 * rounds;
    // (Keeping the metadata of the next two tiles in flight while a tile is copied - a two-stage software pipeline -
    // measured slower: 293 .. 507 us instead of 313 on 10 M x 64 B, 736 instead of 437 us on 1 M x 1200 B.)
    // (Issuing the body loads of a short-frame tile BEFORE the destination look-up is used - they need only the frame's own
    // offsets: two memory round trips per tile instead of three - keeps eight vectors live across that look-up: 160 bytes
    // of spills at 64 registers and 0.656 instead of 0.512 ms per 10 M x 64 B in one call; at 80 registers, three blocks per
    // SM, 0.500 ms there but 0.61 instead of 0.53 ms on 1 M x 1200 B.)
    for (u64 i0 = warp * tile; i0 < a.n; i0 += nwarps * tile) {
        const u64 i = i0 + lane;
        const u64* row = a.s.matrix + (i0 / a.per) * kDirectIds;      // the whole tile lies in one rank block's share
        u64 so = 0, o = 0;
        uint32_t len = 0;
        if (lane < tile && i < a.n) {
            const uint32_t f = a.s.fid[i];
            const uint32_t l = a.s.loc[i];
            u64 po; uint32_t pl;
            if (FROM_FRAMES) {
                const u64 fo = __ldg(a.in_off + i);
                po = fo + 5; pl = (uint32_t)(__ldg(a.in_off + i + 1) - fo - 5);
            } else {
                const uint2* d = reinterpret_cast<const uint2*>(a.desc_in + i);
                const uint2 d0 = __ldg(d), d1 = __ldg(d + 1);
                po = ((u64)d0.y << 32) | d0.x; pl = d1.x;
            }
            if (f != kNoId) {
                const u64 b = __ldg(row + f);
                if (perm) perm[(b >> 32) + a.s.rank[i]] = (uint32_t)i;
                if (gathered && l != kNoLoc) {
                    o = (b & 0xffffffffull) + l;
                    if (o <= gathered_cap && pl <= gathered_cap - o) { len = pl; so = po; }   // overflow: caller sees it in gathered_off[k]
                }
            }
        }
        if (rounds == 4 && __all_sync(0xffffffffu, len <= 64 && (((uintptr_t)(gathered + o) | len) & 15) == 0)) {
            uint4 va[4], vb[4];
            uint8_t* dv[4];
            unsigned sh[4];
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int f = 8 * r + quad;
                const u64 fso = __shfl_sync(0xffffffffu, so, f), fo = __shfl_sync(0xffffffffu, o, f);
                const uint32_t flen = __shfl_sync(0xffffffffu, len, f);
                const uint8_t* src = in + fso;
                sh[r] = (unsigned)((uintptr_t)src & 15);
                const uint8_t* sb = src - sh[r] + 16 * q;
                dv[r] = nullptr;
                va[r] = vb[r] = make_uint4(0, 0, 0, 0);
                if (16u * q < flen) {                          // this lane's vector of the frame: bytes [16 q, 16 q + 16)
                    dv[r] = gathered + fo + 16 * q;
                    va[r] = ldg128(sb);
                    if (sh[r]) vb[r] = ldg128(sb + 16);        // holds item bytes whenever the source is not aligned
                }
            }
#pragma unroll
            for (int r = 0; r < 4; r++)
                if (dv[r]) stg128(dv[r], sh[r] ? funnel16(va[r], vb[r], sh[r]) : va[r]);
        } else if (__all_sync(0xffffffffu, len <= kQuadCopyMax)) {
#pragma unroll
            for (int r = 0; r < 4; r++) {
                if (r >= rounds) break;
                const int f = 8 * r + quad;
                const u64 fso = __shfl_sync(0xffffffffu, so, f), fo = __shfl_sync(0xffffffffu, o, f);
                const uint32_t flen = __shfl_sync(0xffffffffu, len, f);
                if (flen) quad_copy_bytes(gathered + fo, in + fso, flen, q);
            }
        } else {
#pragma unroll 1
            for (int f = 0; f < tile; f++) {
                const u64 fso = __shfl_sync(0xffffffffu, so, f), fo = __shfl_sync(0xffffffffu, o, f);
                const uint32_t flen = __shfl_sync(0xffffffffu, len, f);
                if (flen) warp_copy_bytes(gathered + fo, in + fso, flen, lane);
            }
        }
    }
}

}  // namespace p2pt
