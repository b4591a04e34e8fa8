// sse_kernels.cuh -- SSE event scan over the per-stream bodies the demux gathered.
// The tunnel forwards server-sent events untouched ("sse" feature, /root/reference/tunnel/src/protocol.rs:36,68);
// the framing is the upstream's: `data: {...}\n\n` ... `data: [DONE]\n\n` (/root/reference/tmp/mock_llm.py:54,68-69).
// One warp per stream walks the body in 512-byte rows (aligned 16-byte loads, bytes outside the stream read
// as zero), marks every position that closes a "\n\n" pair, and either counts them (pass 1) or writes the
// event end offsets in order (pass 2).  HBM-bound: every body byte is read once per pass.
#pragma once
#include "frame_kernels.cuh"

namespace p2pt {

// 0x80 in every byte b of the result whose byte in w is '\n' and whose predecessor (byte b - 1, the top byte of
// prev_word for b = 0) is '\n' too - the byte that closes an event.  One funnel shift lines the predecessors up, one
// LOP3 ORs the two differences from 0x0a, and an exact SWAR zero-byte test (no borrow between bytes) flags the bytes
// where both vanish: five instructions per word (testing for '\n' first and pairing the flags afterwards took eight,
// and the scan kernels were issue-bound at 68 / 135 instructions per 512-byte row).
__device__ __forceinline__ uint32_t lf_pair_flags(uint32_t w, uint32_t prev_word) {
    const uint32_t s = __funnelshift_l(prev_word, w, 8);                // bytes [prev.3, w.0, w.1, w.2]
    const uint32_t z = (w ^ 0x0a0a0a0au) | (s ^ 0x0a0a0a0au);           // zero byte <=> both are '\n'
    return ~(((z & 0x7f7f7f7fu) + 0x7f7f7f7fu) | z) & 0x80808080u;
}
struct PairFlags { uint32_t f[4]; };
__device__ __forceinline__ PairFlags lf_pairs16(const uint4& v, uint32_t prev_word) {
    PairFlags r;
    r.f[0] = lf_pair_flags(v.x, prev_word); r.f[1] = lf_pair_flags(v.y, v.x);
    r.f[2] = lf_pair_flags(v.z, v.y);       r.f[3] = lf_pair_flags(v.w, v.z);
    return r;
}
__device__ __forceinline__ uint32_t pair_count(const PairFlags& p) { return __popc(p.f[0]) + __popc(p.f[1]) + __popc(p.f[2]) + __popc(p.f[3]); }
// bit b = byte b of the vector closes an event (only lanes that hold an event pay for this: the multiply gathers the
// four flag bits of a word into one nibble)
__device__ __forceinline__ uint32_t pair_mask16(const PairFlags& p) {
    uint32_t m = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) m |= (((p.f[k] >> 7) * 0x01020408u) >> 24) << (4 * k);
    return m;
}

// keep bytes [lo, hi) of v, zero the others (0 <= lo <= hi <= 16); zero is not '\n'
__device__ __forceinline__ uint4 keep_bytes16(const uint4& v, int lo, int hi) {
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int l = lo - 4 * k, h = hi - 4 * k;                        // range inside this word
        const uint32_t ml = l <= 0 ? 0xffffffffu : l >= 4 ? 0u : (0xffffffffu << (8 * l));
        const uint32_t mh = h >= 4 ? 0xffffffffu : h <= 0 ? 0u : (0xffffffffu >> (32 - 8 * h));
        w[k] &= ml & mh;
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

// does an event that starts at body + pos read `data: [DONE]\n\n`?  Byte 6 tells `data: [` from `data: {`,
// so ordinary events cost one byte load.
__device__ __forceinline__ bool sse_is_done_event(const uint8_t* __restrict__ body, u64 pos, u64 stream_end) {
    if (stream_end - pos < 14 || __ldg(body + pos + 6) != (uint8_t)'[') return false;
    const char* D = "data: [DONE]\n\n";
    bool eq = true;
    for (int q = 0; q < 14; q++) eq &= (__ldg(body + pos + q) == (uint8_t)D[q]);
    return eq;
}

// (Splitting every stream into four parts, one warp each, to raise occupancy on a few thousand long streams
// measured the same 0.78-0.79 ms per 1.2 GB: the passes are issue-bound on the per-byte tests, not short of warps.)
// One warp per stream; the stream is walked in groups of kSseRows rows of 512 bytes whose ALIGNED 16-byte
// loads are all issued before the first is used (one vector per lane and row; a vector that sticks out of
// the stream is masked, never byte-loaded).  The count pass only sums; the emit pass needs each lane's
// exclusive event index, two rows share one 5-step shuffle scan (16-bit fields).
#ifndef P2PT_SSE_ROWS
#define P2PT_SSE_ROWS 4          // measured per 1.2 GB of bodies: 2 rows 0.84 ms, 4 rows 0.79 ms, 8 rows 0.91 ms (80 registers)
#endif
constexpr int kSseRows = P2PT_SSE_ROWS;

template <bool EMIT>
__global__ void __launch_bounds__(256) sse_scan_kernel(u64 k, const uint8_t* __restrict__ body, const u64* __restrict__ body_off,
                                                      u64* __restrict__ ev_csr, u64* __restrict__ ev_end, u64 cap,
                                                      uint8_t* __restrict__ done) {
    const int lane = threadIdx.x & 31;
    const u64 warp = ((u64)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((u64)gridDim.x * blockDim.x) >> 5;
    for (u64 s = warp; s < k; s += nwarps) {
        const u64 a = __ldg(body_off + s), b = __ldg(body_off + s + 1);
        const uint8_t* lo = body + a; const uint8_t* hi = body + b;
        const uint8_t* row0 = lo - ((uintptr_t)lo & 15);                  // aligned start of the first row
        u64 count = 0;
        uint32_t lane_count = 0;
        const u64 out0 = EMIT ? ev_csr[s] : 0;
        uint32_t carry = 0;                                   // last word of the previous row (0 at the stream start)
        bool is_done = false;
        if (EMIT && lane == 0) is_done = sse_is_done_event(body, a, b);   // an event starts at the stream start
        for (const uint8_t* rp = row0; rp < hi; rp += 512 * kSseRows) {
            uint4 v[kSseRows];
#pragma unroll
            for (int r = 0; r < kSseRows; r++) {
                const uint8_t* p = rp + 512 * r + 16 * lane;
                v[r] = (p < hi) ? ldg128(p) : make_uint4(0, 0, 0, 0);
            }
            PairFlags m[kSseRows];
            uint32_t c[kSseRows];
#pragma unroll
            for (int r = 0; r < kSseRows; r++) {
                const uint8_t* p = rp + 512 * r + 16 * lane;
                if (p < hi && (p < lo || p + 16 > hi))
                    v[r] = keep_bytes16(v[r], p < lo ? (int)(lo - p) : 0, p + 16 > hi ? (int)(hi - p) : 16);
                uint32_t prev = __shfl_up_sync(0xffffffffu, v[r].w, 1);
                if (lane == 0) prev = carry;
                carry = __shfl_sync(0xffffffffu, v[r].w, 31);
                m[r] = lf_pairs16(v[r], prev);
                c[r] = pair_count(m[r]);
            }
            if (!EMIT) {
#pragma unroll
                for (int r = 0; r < kSseRows; r++) lane_count += c[r];
            } else if (!__any_sync(0xffffffffu, (c[0] | c[1] | c[2] | c[3]) > 1u)) {
                // the usual case - no lane holds two events of a row (events are longer than 16 bytes): a lane's event
                // index is a ballot and a popcount, no shuffle scan, and there is no loop over the lane's events
                static_assert(kSseRows == 4, "four rows per group");
#pragma unroll
                for (int r = 0; r < kSseRows; r++) {
                    const uint32_t bal = __ballot_sync(0xffffffffu, c[r] != 0);
                    if (c[r]) {
                        const u64 slot = out0 + count + __popc(bal & ((1u << lane) - 1u));
                        const u64 end = (u64)(rp + 512 * r + 16 * lane - body) + __ffs(pair_mask16(m[r]));   // just past the second '\n'
                        if (slot < cap) ev_end[slot] = end;
                        is_done |= sse_is_done_event(body, end, b);            // does the next event read `data: [DONE]`?
                    }
                    count += __popc(bal);
                }
            } else {
#pragma unroll
                for (int r2 = 0; r2 < kSseRows; r2 += 2) {
                    // a row holds at most 512 events, so two rows' inclusive scans fit 16-bit fields
                    const uint32_t both = c[r2] | (c[r2 + 1] << 16);
                    uint32_t incl = both;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
                    const uint32_t tot = __shfl_sync(0xffffffffu, incl, 31);
#pragma unroll
                    for (int e = 0; e < 2; e++) {
                        const int r = r2 + e;
                        const uint32_t inc_r = (incl >> (16 * e)) & 0xffffu;
                        u64 slot = out0 + count + (inc_r - c[r]);
                        uint32_t mm = c[r] ? pair_mask16(m[r]) : 0u;
                        const u64 pos0 = (u64)(rp + 512 * r + 16 * lane - body);
                        while (mm) {
                            const int bit = __ffs(mm) - 1; mm &= mm - 1;
                            const u64 end = pos0 + bit + 1;                    // offset just past the second '\n'
                            if (slot < cap) ev_end[slot] = end;
                            slot++;
                            is_done |= sse_is_done_event(body, end, b);        // does the next event read `data: [DONE]`?
                        }
                        count += (tot >> (16 * e)) & 0xffffu;
                    }
                }
            }
        }
        if (!EMIT) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) lane_count += __shfl_xor_sync(0xffffffffu, lane_count, o);
            if (lane == 0) ev_csr[s] = lane_count;             // per-stream counts, scanned afterwards
        } else {
            const bool any = __any_sync(0xffffffffu, is_done);
            if (lane == 0) done[s] = any ? 1 : 0;
        }
    }
}

// ---- one pass ---------------------------------------------------------------------------------------------------
// The two-pass scan above reads every body byte twice (count per stream, scan the counts, emit).  The order of ev_end
// is simply the byte order of the bodies (streams lie back to back, body_off is monotone), so the whole thing is ONE
// stream compaction over the byte range with a decoupled look-back: blocks take 16 KiB tiles in ticket order, flag the
// bytes that close an event, publish the tile's count, sum the counts of the tiles in front (waiting only for tiles
// whose ticket is lower - those are held by blocks that are already running, and every block publishes its count
// before it waits), and write the tile's event ends at the running index.  A '\n' pair must not reach across a stream
// boundary: the stream starts that lie in a tile are marked in a bitmap and their flags cleared; the same streams get
// their ev_csr (events in front of their first byte) from the tile.  Which streams start in a tile comes from a small
// pre-pass (sse_tile_first_kernel).  The body is read once.
#ifndef P2PT_SSE_UNROLL
#define P2PT_SSE_UNROLL 4      // vectors in flight per thread in the flag phase
#endif
constexpr int kSseUnroll = P2PT_SSE_UNROLL;
constexpr uint32_t kSseTileBytes = 65536;                                // per block and ticket: what is in flight per SM has to cover
constexpr int kSseTileVecs = kSseTileBytes / 16;                         // the latencies of a tile (look-back, offsets): 16 KiB tiles
constexpr int kSseVecsPerThread = kSseTileVecs / 256;                    // measured 1.26 ms per 1.2 GB, twice the two-pass scan
__host__ __device__ inline u64 sse_onepass_tiles(u64 body_bytes) { return (body_bytes + 15) / kSseTileBytes + 1; }

struct SseOnepassArgs {
    u64 k; const uint8_t* body; const u64* body_off; u64 ntiles;
    u64* ev_csr; u64* ev_end; u64 cap; uint8_t* done; u64* n_events_out;
    u64* state;                 // [ntiles] count << 2 | (1 = the tile's own count, 2 = count of all tiles up to it); zeroed
    u64* gstate;                // [ntiles / 32 + 1] per group of 32 tiles: sum of the counts published so far << 8 | how many; zeroed
    u64* gfinal;                // [ntiles / 32 + 1] count of all tiles up to the group's last one << 1 | 1 once known; zeroed
    unsigned long long* ticket; // zeroed
    uint32_t* tile_first;       // [ntiles] lowest stream index (0 .. k) whose first byte lies in the tile; 0xffffffff = none
};

__global__ void __launch_bounds__(256) sse_tile_first_kernel(const SseOnepassArgs a) {
    const u64 b0 = __ldg(a.body_off);
    const uint32_t lead = (uint32_t)((uintptr_t)(a.body + b0) & 15);
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 s = (u64)blockIdx.x * blockDim.x + threadIdx.x; s <= a.k; s += stride) {
        u64 t = (__ldg(a.body_off + s) - b0 + lead) / kSseTileBytes;
        if (t >= a.ntiles) t = a.ntiles - 1;
        atomicMin(a.tile_first + t, (uint32_t)s);
    }
}

__device__ __forceinline__ u64 ld_relaxed_u64(const u64* p) {
    u64 v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u64(u64* p, u64 v) {
    asm volatile("st.release.gpu.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}

// first index u in [0, k] with body_off[u] > pos (k + 1 if none)
__device__ __forceinline__ u64 sse_upper_bound(const u64* __restrict__ body_off, u64 k, u64 pos) {
    u64 lo = 0, hi = k + 1;
    while (lo < hi) { const u64 mid = (lo + hi) >> 1; if (__ldg(body_off + mid) <= pos) lo = mid + 1; else hi = mid; }
    return lo;
}

__global__ void __launch_bounds__(256) sse_onepass_kernel(const SseOnepassArgs a) {
    __shared__ __align__(16) uint16_t s_m16[kSseTileVecs];       // per vector: bit b = byte b closes an event
    __shared__ __align__(16) uint16_t s_pre[kSseTileVecs];       // events of the tile in front of the vector (<= 65535 per tile)
    __shared__ __align__(16) uint32_t s_start[kSseTileVecs / 2]; // bit per byte of the tile: a stream starts here
    __shared__ uint32_t s_wsum[8];
    __shared__ u64 s_base;
    __shared__ unsigned long long s_tile;
    const int t = threadIdx.x, lane = t & 31, wib = t >> 5;
    const u64 b0 = __ldg(a.body_off), b1 = __ldg(a.body_off + a.k);
    const uint8_t* lo = a.body + b0; const uint8_t* hi = a.body + b1;
    const uint32_t lead = (uint32_t)((uintptr_t)lo & 15);
    const uint8_t* base0 = lo - lead;                  // aligned start of tile 0
    const uint16_t* s_start16 = reinterpret_cast<const uint16_t*>(s_start);
    for (;;) {
        if (t == 0) s_tile = atomicAdd(a.ticket, 1ull);
#pragma unroll
        for (int q = 0; q < kSseTileVecs / 2 / 256; q++) s_start[t + 256 * q] = 0;
        __syncthreads();
        const u64 tile = s_tile;
        if (tile >= a.ntiles) return;
        const uint8_t* tp = base0 + tile * kSseTileBytes;
        const u64 rel0 = tile * kSseTileBytes;         // distance of the tile's first byte from base0
        const bool last_tile = tile == a.ntiles - 1;
        const uint32_t fs = a.tile_first[tile];        // needed further down: in flight meanwhile
        // ---- flags of every vector (bytes outside the bodies count as zero); a vector's predecessor byte is its left
        //      neighbour's last one: the lane to the left holds it, lane 0 reads it
        if (tp > lo && tp + kSseTileBytes <= hi) {
            // a tile inside the bodies (all but the first and the last): no range tests
#pragma unroll 1
            for (int j0 = 0; j0 < kSseVecsPerThread; j0 += kSseUnroll) {
                uint4 v[kSseUnroll];
                const uint8_t* p0 = tp + 16 * (t + 256 * j0);
                uint32_t pb[kSseUnroll];                                   // lane 0: the byte in front of its vector
#pragma unroll
                for (int u = 0; u < kSseUnroll; u++) {                     // all loads of the group before the first use
                    v[u] = ldg128(p0 + 4096 * u);
                    pb[u] = lane == 0 ? (uint32_t)__ldg(p0 + 4096 * u - 1) : 0u;
                }
#pragma unroll
                for (int u = 0; u < kSseUnroll; u++) {
                    uint32_t prev = __shfl_up_sync(0xffffffffu, v[u].w, 1);
                    if (lane == 0) prev = pb[u] << 24;
                    const PairFlags f = lf_pairs16(v[u], prev);
                    s_m16[t + 256 * (j0 + u)] = (f.f[0] | f.f[1] | f.f[2] | f.f[3]) ? (uint16_t)pair_mask16(f) : (uint16_t)0;
                }
            }
        } else {
#pragma unroll 1
            for (int j = 0; j < kSseVecsPerThread; j++) {
                const int vi = t + 256 * j;
                const uint8_t* p = tp + 16 * vi;
                uint4 v = make_uint4(0, 0, 0, 0);
                if (p < hi && p + 16 > lo) {
                    v = ldg128(p);
                    if (p < lo || p + 16 > hi) v = keep_bytes16(v, p < lo ? (int)(lo - p) : 0, p + 16 > hi ? (int)(hi - p) : 16);
                }
                uint32_t prev = __shfl_up_sync(0xffffffffu, v.w, 1);
                if (lane == 0) prev = (p > lo && p <= hi) ? ((uint32_t)__ldg(p - 1) << 24) : 0u;
                const PairFlags f = lf_pairs16(v, prev);
                s_m16[vi] = (f.f[0] | f.f[1] | f.f[2] | f.f[3]) ? (uint16_t)pair_mask16(f) : (uint16_t)0;
            }
        }
        // ---- the streams whose first byte lies in this tile: a pair must not reach across it
        if (fs != 0xffffffffu) {
            for (u64 s = (u64)fs + t; s <= a.k; s += 256) {
                const u64 r = __ldg(a.body_off + s) - b0 + lead - rel0;
                if (r >= kSseTileBytes) break;          // later tiles (offsets are monotone); behind the last tile there is no byte
                atomicOr(&s_start[r >> 5], 1u << (r & 31));
            }
        }
        __syncthreads();
        // ---- thread t owns vectors 16 t .. 16 t + 15 from here on: clear the stream starts, count, prefix
        uint32_t mine = 0;
        uint32_t cw[kSseVecsPerThread / 2];                                // bit b of word i: byte 32 i + b of this thread's 256 closes an event
        {
            uint4* mq = reinterpret_cast<uint4*>(s_m16 + kSseVecsPerThread * t);
            const uint4* sq = reinterpret_cast<const uint4*>(s_start16 + kSseVecsPerThread * t);
#pragma unroll
            for (int h = 0; h < kSseVecsPerThread / 8; h++) {
                uint4 mm = mq[h];
                const uint4 ss = sq[h];
                mm.x &= ~ss.x; mm.y &= ~ss.y; mm.z &= ~ss.z; mm.w &= ~ss.w;
                if (fs != 0xffffffffu) mq[h] = mm;                         // the streams that start here look their counts up below
                cw[4 * h] = mm.x; cw[4 * h + 1] = mm.y; cw[4 * h + 2] = mm.z; cw[4 * h + 3] = mm.w;
                mine += __popc(mm.x) + __popc(mm.y) + __popc(mm.z) + __popc(mm.w);
            }
        }
        uint32_t incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t x = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += x; }
        if (lane == 31) s_wsum[wib] = incl;
        __syncthreads();
        uint32_t wbase = 0, agg = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) { const uint32_t x = s_wsum[w]; if (w < wib) wbase += x; agg += x; }
        const uint32_t tpre = wbase + incl - mine;     // events of the tile in front of this thread's vectors
        if (fs != 0xffffffffu) {                       // the streams that start here look their counts up below
            uint32_t run = tpre;
#pragma unroll
            for (int q = 0; q < kSseVecsPerThread; q++) {
                s_pre[kSseVecsPerThread * t + q] = (uint16_t)run;
                run += __popc((uint32_t)s_m16[kSseVecsPerThread * t + q]);
            }
        }
        // ---- look-back by warp 0: the count of all tiles in front.  All resident blocks (up to 1184) reach this point at
        //      about the same time, so the nearest tile with a FINAL count lies about that far back: walking the tile
        //      states 32 at a time took ~30 us per tile (most of the kernel), 256 at a time flooded L2 with polling.  So
        //      tiles also add their count to their GROUP of 32 (sum and number of tiles in one 64-bit word, one atomic),
        //      and the last tile of a group publishes the final count at the group's end: a walk is the tiles of the own
        //      group in front (one step) and then whole groups, 32 of them = 1024 tiles per step.
        if (wib == 0) {
            u64 base = 0;
            const u64 g = tile >> 5;
            const int r = (int)(tile & 31);
            if (lane == 0) {
                st_release_u64(a.state + tile, ((u64)agg << 2) | 1u);
                atomicAdd(reinterpret_cast<unsigned long long*>(a.gstate + g), ((unsigned long long)agg << 8) | 1ull);
            }
            bool found = false;
            if (r > 0) {                                                    // tiles tile - 1 .. tile - r of this group
                for (;;) {
                    const u64 st = lane < r ? ld_relaxed_u64(a.state + tile - 1 - lane) : 0ull;
                    const uint32_t flag = (uint32_t)(st & 3u);
                    const uint32_t fin = __ballot_sync(0xffffffffu, lane < r && flag == 2u);
                    const int first_fin = fin ? __ffs(fin) - 1 : r;        // nearest tile with a final count (r: none)
                    if (__ballot_sync(0xffffffffu, lane < r && lane <= first_fin && flag == 0u)) { __nanosleep(64); continue; }
                    u64 part = (lane < r && lane <= first_fin) ? (st >> 2) : 0ull;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
                    base += part;
                    found = first_fin < r;
                    break;
                }
            }
            if (!found && g > 0) {                                          // whole groups g - 1, g - 2, ...
                long long jg = (long long)g - 1;
                for (;;) {
                    const long long idx = jg - lane;
                    const u64 gf = idx >= 0 ? ld_relaxed_u64(a.gfinal + idx) : 1ull;   // in front of group 0: nothing, final
                    const u64 gs = idx >= 0 ? ld_relaxed_u64(a.gstate + idx) : 32ull;
                    const uint32_t fin = __ballot_sync(0xffffffffu, (gf & 1ull) != 0);
                    const int first_fin = fin ? __ffs(fin) - 1 : 32;
                    const bool sums = lane < first_fin;                     // groups in front of the nearest final one: whole sums
                    if (__ballot_sync(0xffffffffu, sums && (gs & 0xffull) != 32ull)) { __nanosleep(64); continue; }
                    u64 part = sums ? (gs >> 8) : (lane == first_fin ? (gf >> 1) : 0ull);
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
                    base += part;
                    if (first_fin < 32) break;
                    jg -= 32;
                }
            }
            if (lane == 0) {
                st_release_u64(a.state + tile, ((base + agg) << 2) | 2u);
                if (r == 31) st_release_u64(a.gfinal + g, ((base + agg) << 1) | 1u);
                s_base = base;
            }
        }
        __syncthreads();
        const u64 base = s_base;
        // ---- emit the event ends of this thread's vectors, in order
        if (mine) {
            u64 slot = base + tpre;
            const u64 pos0 = (u64)(tp - a.body) + 16u * (uint32_t)(kSseVecsPerThread * t);
            // Does the event behind an end read `data: [DONE]`?  Byte 6 tells `data: [` from `data: {`, and only a full match
            // looks for the stream it lies in (it has to end inside that stream).  That byte is a load from L2 per event;
            // asked for one event at a time in the loop below the loads ran one after the other (the emit phase was 26 % of the
            // instructions and 43 % of the stall samples): four ends are collected, then their bytes are fetched together.
            uint32_t pe0 = 0, pe1 = 0, pe2 = 0, pe3 = 0;                       // ends, relative to pos0 (1 .. 256)
            int np = 0;
            auto check = [&](uint32_t rel) {
                const u64 end = pos0 + rel;
                if (sse_is_done_event(a.body, end, b1)) {
                    const u64 s = sse_upper_bound(a.body_off, a.k, end) - 1;
                    if (s < a.k && __ldg(a.body_off + s + 1) - end >= 14) a.done[s] = 1;
                }
            };
            auto flush = [&]() {
                const uint8_t* q = a.body + pos0 + 6;
                const bool h0 = np > 0 && b1 - (pos0 + pe0) >= 14, h1 = np > 1 && b1 - (pos0 + pe1) >= 14;
                const bool h2 = np > 2 && b1 - (pos0 + pe2) >= 14, h3 = np > 3 && b1 - (pos0 + pe3) >= 14;
                const uint32_t c0 = h0 ? __ldg(q + pe0) : 0u, c1 = h1 ? __ldg(q + pe1) : 0u;
                const uint32_t c2 = h2 ? __ldg(q + pe2) : 0u, c3 = h3 ? __ldg(q + pe3) : 0u;
                if (c0 == (uint32_t)'[') check(pe0);
                if (c1 == (uint32_t)'[') check(pe1);
                if (c2 == (uint32_t)'[') check(pe2);
                if (c3 == (uint32_t)'[') check(pe3);
                np = 0;
            };
#pragma unroll
            for (int i = 0; i < kSseVecsPerThread / 2; i++) {
                uint32_t mm = cw[i];
                while (mm) {
                    const int bit = __ffs(mm) - 1; mm &= mm - 1;
                    const uint32_t rel = 32 * i + bit + 1;                     // just past the second '\n'
                    if (slot < a.cap) a.ev_end[slot] = pos0 + rel;
                    slot++;
                    pe3 = pe2; pe2 = pe1; pe1 = pe0; pe0 = rel;
                    if (++np == 4) flush();
                }
            }
            if (np) flush();
        }
        // ---- the streams that start in this tile: events in front of them, and an event starts at their first byte
        if (fs != 0xffffffffu) {
            for (u64 s = (u64)fs + t; s <= a.k; s += 256) {
                const u64 so = __ldg(a.body_off + s);
                const u64 r = so - b0 + lead - rel0;
                if (r >= kSseTileBytes && !last_tile) break;               // the last tile also takes what lies behind it (nothing there)
                const uint32_t vi = (uint32_t)(r >> 4);
                const u64 before = r >= kSseTileBytes ? (u64)agg
                                                      : (u64)s_pre[vi] + __popc((uint32_t)s_m16[vi] & ((1u << (r & 15)) - 1u));
                a.ev_csr[s] = base + before;
                if (s < a.k && sse_is_done_event(a.body, so, __ldg(a.body_off + s + 1))) a.done[s] = 1;
            }
        }
        if (last_tile && t == 0) *a.n_events_out = base + agg;
        __syncthreads();                                                      // shared arrays are rewritten for the next tile
    }
}

// exclusive scan of k u64 counts in place (k + 1 entries out): block sums -> single block scan -> add
__global__ void __launch_bounds__(kScanThreads) u64_scan_blocksum_kernel(u64 k, const u64* __restrict__ v, u64* __restrict__ blk) {
    const u64 i = (u64)blockIdx.x * kScanThreads + threadIdx.x;
    HeadBytes x; x.heads = 0; x.bytes = i < k ? v[i] : 0;
    HeadBytes tot;
    (void)block_scan_excl(x, tot);
    if (threadIdx.x == 0) blk[blockIdx.x] = tot.bytes;
}
__global__ void __launch_bounds__(kScanThreads) u64_scan_blocksums_kernel(u64 nblk, u64* __restrict__ blk) {
    __shared__ u64 carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (u64 b0 = 0; b0 < nblk; b0 += kScanThreads) {
        const u64 b = b0 + threadIdx.x;
        HeadBytes x; x.heads = 0; x.bytes = b < nblk ? blk[b] : 0;
        HeadBytes tot;
        HeadBytes ex = block_scan_excl(x, tot);
        const u64 c = carry;
        if (b < nblk) blk[b] = c + ex.bytes;
        __syncthreads();
        if (threadIdx.x == 0) carry = c + tot.bytes;
        __syncthreads();
    }
    if (threadIdx.x == 0) blk[nblk] = carry;
}
__global__ void __launch_bounds__(kScanThreads) u64_scan_apply_kernel(u64 k, u64 nblk, u64* __restrict__ v, const u64* __restrict__ blk,
                                                                     u64* __restrict__ total_out) {
    const u64 i = (u64)blockIdx.x * kScanThreads + threadIdx.x;
    HeadBytes x; x.heads = 0; x.bytes = i < k ? v[i] : 0;
    HeadBytes tot;
    HeadBytes ex = block_scan_excl(x, tot);
    if (i < k) v[i] = blk[blockIdx.x] + ex.bytes;
    if (i == 0) { v[k] = blk[nblk]; *total_out = blk[nblk]; }
}

}  // namespace p2pt
