// frame_kernels.cuh -- batched TunnelMessage encode / decode / demux for sm_100a.
//
// What the reference does one frame at a time:
//   encode  /root/reference/tunnel/src/protocol.rs:148-154   [type u8][stream_id u32 BE][payload]
//   decode  /root/reference/tunnel/src/protocol.rs:157-172   len<5 -> too short; tag set :103-118
//   demux   /root/reference/tunnel/src/proxy.rs:118-153 (HashMap<u32,_> probe per frame) and
//           /root/reference/tunnel/src/serve.rs:112-139 (ReqBody appended to a per-stream Vec<u8>)
// All of it is HBM-bound byte work: the kernels move aligned 16-byte vectors (common.cuh), size
// their grids in multiples of the SM count, and never touch a tensor core.
#pragma once
#include <cooperative_groups.h>
#include "common.cuh"
#include "../../include/p2pt.h"

namespace p2pt {

// 256-bit membership mask of the 11 valid tags (protocol.rs:88-100): 1,2,3,4,10,11,12,20,21,22,99
__device__ __forceinline__ bool msg_type_valid(uint32_t b) {
    // bits 0..31: 1,2,3,4,10,11,12,20,21,22 ; bits 96..127: 99 -> bit 3 of word 3
    const uint32_t w0 = (1u << 1) | (1u << 2) | (1u << 3) | (1u << 4) | (1u << 10) | (1u << 11) |
                        (1u << 12) | (1u << 20) | (1u << 21) | (1u << 22);
    if (b < 32) return (w0 >> b) & 1u;
    return b == 99;
}

// Resident 256-thread blocks per SM the two copy kernels are compiled for (register budget 64 K / (256 x blocks)).
// Measured on 1 M x 1200 B with the two-loads-per-vector warp copy, three vectors in flight per lane:
// encode 0.447 / 0.463 / 0.435 ms at 4 / 5 / 6 blocks; demux + gather 0.573 / 0.588 / 0.588 ms (the gather
// spills at 40 registers).
#ifndef P2PT_ENCODE_MIN_BLOCKS
#define P2PT_ENCODE_MIN_BLOCKS 4
#endif
#ifndef P2PT_GATHER_MIN_BLOCKS
#define P2PT_GATHER_MIN_BLOCKS 4
#endif

// ---- encode ----------------------------------------------------------------------------------------
// Two regimes, chosen once per launch from the average output slot (two loads every thread makes):
//
//  * general (slots above kSmallFrameAvg bytes on average): a warp takes a tile of 8 * rounds consecutive frames
//    (rounds = 1 or 4).  Lane l loads the metadata of frame l (coalesced) and writes its 5-byte header; the payload
//    copies then run in rounds of eight frames, four lanes per frame, when all of the tile are short (<= 256 B),
//    else frame after frame with all 32 lanes (aligned 16-byte vectors either way).  Tiles of 32 amortise the
//    metadata latency over four times the frames; the host picks them when the batch is large enough that every
//    warp still gets many tiles (copy_tile_rounds), else the last warps would run alone.
//  * small frames (the SSE token regime, 10 M x 64 B): a 69-byte stride makes two of every five aligned
//    destination vectors partial, and with frames written straight to global memory 13 of 32 lanes were live
//    on average (0.30 of the HBM roofline).  Here a block takes kSmallTile consecutive frames at a time: their
//    payload span is contiguous (offsets are back to back by construction) and comes into shared memory as whole
//    aligned vectors; two lanes per frame then lay header + payload out in a second shared buffer at the
//    frame's output position (32-bit shared-memory words, one funnel shift each, any alignment on both sides);
//    and the output span - contiguous when the slots have no gaps - leaves as whole aligned 16-byte vectors.
//    Only the two ends of a span are partial.  A tile whose slots have gaps or whose span does not fit the
//    buffers takes the general path's four-lanes-per-frame copy.
constexpr int kSmallTile = 128;                  // frames per block iteration in the small-frame regime
constexpr uint32_t kSmallTileIn = 8192;          // payload span held in shared memory
constexpr uint32_t kSmallTileOut = kSmallTileIn + 5 * kSmallTile;
constexpr u64 kSmallFrameAvg = 128;              // average slot (bytes) up to which a launch runs in the small-frame regime

// The two shared buffers of the small-frame regime are PADDED by one word after every 16: logical word w lives at
// physical word w + (w >> 4).  Frames of 64 bytes (the SSE chunk of BASELINE.json config #4) start 16 words apart,
// so without the padding the lanes of a warp - each on its own frame - hit two banks; measured 67 % of the
// shared-memory wavefronts were bank conflicts and the L1 pipe sat at 95 % of its peak.  With the padding
// consecutive frames start 17 words apart.
__device__ __forceinline__ uint32_t pad_word(uint32_t w) { return w + (w >> 4); }
struct PaddedBuf {
    uint32_t* w;                                       // physical words
    __device__ __forceinline__ uint32_t& word(uint32_t i) const { return w[pad_word(i)]; }
    __device__ __forceinline__ uint8_t& byte(uint32_t b) const { return reinterpret_cast<uint8_t*>(w)[4 * pad_word(b >> 2) + (b & 3)]; }
};

// len bytes from logical byte offset so of src to logical byte offset d of dst, any alignment on either side, by
// `parts` cooperating lanes (part = 0 .. parts - 1).  Whole destination words come from two source words and a
// funnel shift; each lane walks a contiguous run of words so that consecutive words share a source word.
__device__ __forceinline__ void smem_copy_unaligned(const PaddedBuf& dst, uint32_t d, const PaddedBuf& src, uint32_t so,
                                                    uint32_t len, int part, int parts) {
    const uint32_t d0 = (4 - (d & 3)) & 3;                                       // bytes up to the first aligned dst word
    if (len < d0 + 4) {                                                          // no whole word: bytes
        for (uint32_t b = part; b < len; b += parts) dst.byte(d + b) = src.byte(so + b);
        return;
    }
    const uint32_t nw = (len - d0) >> 2;                                         // whole destination words
    if (part == 0) for (uint32_t b = 0; b < d0; b++) dst.byte(d + b) = src.byte(so + b);
    if (part == parts - 1) for (uint32_t b = d0 + 4 * nw; b < len; b++) dst.byte(d + b) = src.byte(so + b);
    const uint32_t per = (nw + parts - 1) / parts, w0 = per * part;
    uint32_t w1 = w0 + per; if (w1 > nw) w1 = nw;
    if (w0 >= w1) return;
    const uint32_t dw = (d + d0) >> 2;                                           // first whole destination word
    const uint32_t sb = so + d0 + 4 * w0;                                        // source byte of destination word w0
    const unsigned sh = (sb & 3) * 8;
    const uint32_t sw = sb >> 2;
    uint32_t lo = src.word(sw);
    for (uint32_t k = 0; k < w1 - w0; k++) {
        const uint32_t hi = src.word(sw + k + 1);   // for an aligned source the funnel ignores it; the buffers carry spare words
        dst.word(dw + w0 + k) = __funnelshift_r(lo, hi, sh);
        lo = hi;
    }
}

__global__ void __launch_bounds__(256, P2PT_ENCODE_MIN_BLOCKS) frame_encode_kernel(
    u64 n, const uint8_t* __restrict__ types, const uint32_t* __restrict__ sids,
    const uint8_t* __restrict__ payload, const u64* __restrict__ pay_off,
    uint8_t* __restrict__ out, const u64* __restrict__ out_off, int rounds) {
    const int lane = threadIdx.x & 31, quad = lane >> 2, q = lane & 3;
    if (__ldg(out_off + n) - __ldg(out_off) > n * kSmallFrameAvg) {
        // ------------------------------------------------------------------ general regime
        const u64 warp = ((u64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
        const u64 nwarps = ((u64)gridDim.x * blockDim.x) >> 5;
        const int tile = 8 * rounds;
        for (u64 i0 = warp * tile; i0 < n; i0 += nwarps * tile) {
            const u64 i = i0 + lane;
            u64 po = 0, oo = 0;
            uint32_t len = 0;
            bool live = lane < tile && i < n;
            if (live) {
                po = __ldg(pay_off + i); oo = __ldg(out_off + i);
                const u64 l64 = __ldg(pay_off + i + 1) - po;
                len = (uint32_t)l64;                       // payload lengths are 32-bit throughout (p2pt_desc.pay_len)
                const uint32_t sid = __ldg(sids + i);      // header: [type][stream_id BE]
                uint8_t* o = out + oo;
                o[0] = __ldg(types + i);
                o[1] = (uint8_t)(sid >> 24); o[2] = (uint8_t)(sid >> 16); o[3] = (uint8_t)(sid >> 8); o[4] = (uint8_t)sid;
            }
            if (__all_sync(0xffffffffu, len <= kQuadCopyMax)) {
#pragma unroll
                for (int r = 0; r < 4; r++) {
                    if (r >= rounds) break;
                    const int f = 8 * r + quad;
                    const u64 fpo = __shfl_sync(0xffffffffu, po, f), foo = __shfl_sync(0xffffffffu, oo, f);
                    const uint32_t flen = __shfl_sync(0xffffffffu, len, f);
                    if (flen) quad_copy_bytes(out + foo + 5, payload + fpo, flen, q);
                }
            } else {
#pragma unroll 1
                for (int f = 0; f < tile; f++) {
                    const u64 fpo = __shfl_sync(0xffffffffu, po, f), foo = __shfl_sync(0xffffffffu, oo, f);
                    const uint32_t flen = __shfl_sync(0xffffffffu, len, f);
                    if (flen) warp_copy_bytes(out + foo + 5, payload + fpo, flen, lane);
                }
            }
        }
        return;
    }
    // ---------------------------------------------------------------------- small-frame regime
    // sin: the payload span behind kSinFront spare bytes, padded (pad_word); sout: the output span, plain words, all
    // zero between tiles (boundary words are OR-ed together from the frames that share them)
    constexpr uint32_t kSinFront = 16;
    __shared__ __align__(16) uint32_t sin_w[(kSmallTileIn + 64) / 4 + (kSmallTileIn + 64) / 64 + 2];
    __shared__ __align__(16) uint32_t sout_w[(kSmallTileOut + 32) / 4];
    const PaddedBuf sin{sin_w};
    __shared__ u64 s_po[kSmallTile + 1], s_oo[kSmallTile + 1];
    __shared__ uint32_t s_sid[kSmallTile];
    __shared__ uint8_t s_type[kSmallTile];
    const int t = threadIdx.x;
    const u64 ntile = (n + kSmallTile - 1) / kSmallTile;
    // Software pipeline over the block's tiles: the global loads of tile k + 1 (offsets, stream ids, types and, when
    // the tile is dense, its payload span) are issued into registers before tile k is shuffled and flushed, so a
    // tile costs one memory round trip and two barriers.
    struct TileRegs {
        u64 pb, pe, ob, oe;        // ends of the payload span and of the slot span (every thread loads them: broadcast)
        u64 po, oo;                // offsets of frame t (t <= T)
        uint32_t sid; uint8_t type;
        uint4 v[3];                // this thread's aligned vectors of the payload span (<= 513 over 256 threads)
        int T; bool dense;
        bool tiny;                 // frame t has fewer than 3 payload bytes (see the whole-word shuffle below)
    };
    auto load_tile = [&](u64 tile, TileRegs& r) {
        const u64 i0 = tile * kSmallTile;
        r.T = n - i0 < (u64)kSmallTile ? (int)(n - i0) : kSmallTile;
        r.pb = __ldg(pay_off + i0); r.pe = __ldg(pay_off + i0 + r.T);
        r.ob = __ldg(out_off + i0); r.oe = __ldg(out_off + i0 + r.T);
        r.po = r.oo = 0; r.sid = 0; r.type = 0;
        r.tiny = false;
        if (t <= r.T) { r.po = __ldg(pay_off + i0 + t); r.oo = __ldg(out_off + i0 + t); }
        if (t < r.T) { r.sid = __ldg(sids + i0 + t); r.type = __ldg(types + i0 + t); r.tiny = __ldg(pay_off + i0 + t + 1) - r.po < 3; }
        r.dense = (r.oe - r.ob == (r.pe - r.pb) + 5ull * (u64)r.T) && (r.pe - r.pb <= kSmallTileIn);
        if (r.dense) {
            // whole aligned vectors (an aligned vector that holds one byte of the span lies in the same page as that
            // byte, so the few bytes read beyond either end of the span are always mapped)
            const uint8_t* gsrc = payload + r.pb;
            const uint32_t ia = (uint32_t)((uintptr_t)gsrc & 15);
            const uint32_t nvin = (ia + (uint32_t)(r.pe - r.pb) + 15) >> 4;
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const uint32_t vi = t + 256 * k;
                if (vi < nvin) r.v[k] = ldg128_stream(gsrc - ia + 16 * (size_t)vi);
            }
        }
    };
    TileRegs cur;
    cur.tiny = false;
    bool sout_dirty = false;                           // block-uniform: sout holds words of a whole-word tile
    if (blockIdx.x < ntile) load_tile(blockIdx.x, cur);
    for (uint32_t w = t; w < sizeof(sout_w) / 4; w += 256) sout_w[w] = 0;     // ordered before the first shuffle by the tile's barrier
    for (u64 tile = blockIdx.x; tile < ntile; tile += gridDim.x) {
        const int T = cur.T;
        const u64 pb = cur.pb, ob = cur.ob;
        const uint32_t ispan = (uint32_t)(cur.pe - cur.pb);
        const u64 ospan = cur.oe - cur.ob;
        const bool dense = cur.dense;
        const uint32_t ia = (uint32_t)((uintptr_t)(payload + pb) & 15);
        // registers -> shared memory (every thread is past the previous tile's shuffle: its second barrier)
        if (t <= T) { s_po[t] = cur.po; s_oo[t] = cur.oo; }
        if (t < T) { s_sid[t] = cur.sid; s_type[t] = cur.type; }
        if (dense) {
            const uint32_t nvin = (ia + ispan + 15) >> 4;
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const uint32_t vi = t + 256 * k;
                if (vi < nvin) {                                       // four words that share one padding offset
                    uint32_t* q4 = sin_w + pad_word(kSinFront / 4 + 4 * vi);
                    q4[0] = cur.v[k].x; q4[1] = cur.v[k].y; q4[2] = cur.v[k].z; q4[3] = cur.v[k].w;
                }
            }
        }
        const bool whole = !__syncthreads_or(cur.tiny);                // tile in shared memory; previous flush behind everyone
        if (tile + gridDim.x < ntile) load_tile(tile + gridDim.x, cur);   // next tile's loads fly during the rest
        if (dense) {
            uint8_t* gdst = out + ob;
            const uint32_t oa = (uint32_t)((uintptr_t)gdst & 15);      // sout keeps the span's 16-byte phase
            if (whole) {
                // Every payload of the tile has at least 3 bytes (an SSE chunk, any body frame): then every aligned output
                // WORD has one owner and is written whole - no byte masks, no OR-ing, no zeroed buffer.  Frame f owns the
                // words from the one its first byte lies in up to, not including, the one the next frame starts in.  The
                // bytes in front of the frame in its first word are the last payload bytes of frame f - 1, and those lie
                // right in front of this frame's payload in the input span (payloads are back to back): the output is
                // the input with five bytes inserted per frame.  So the first two words are funnel shifts of
                // (4 bytes before the payload | header | 4 bytes of payload), and every word after them lies wholly
                // inside the payload.  Two lanes per frame: one head word and half of the payload words each.  (The
                // bytes in front of the tile's first frame and behind its last are not stored by the flush.)
                sout_dirty = true;
                const int f = t >> 1, h = t & 1;
                if (f < T) {
                    const uint32_t po = (uint32_t)(s_po[f] - pb);
                    const uint32_t S = oa + (uint32_t)(s_oo[f] - ob), E = oa + (uint32_t)(s_oo[f + 1] - ob);
                    const uint32_t P0 = kSinFront + ia + po;                                   // payload start in sin
                    const uint32_t wS = S >> 2, wE = f == T - 1 ? (E + 3) >> 2 : E >> 2;
                    {
                        const uint32_t base = P0 - 4, bw = (base >> 2) + h;
                        const uint32_t v = __funnelshift_r(sin.word(bw), sin.word(bw + 1), (base & 3) * 8);   // h = 0: the 4 bytes before the payload, 1: its first 4
                        const uint32_t sidbe = __byte_perm(s_sid[f], 0, 0x0123);
                        const uint32_t hlo = (uint32_t)s_type[f] | (sidbe << 8);               // [type][stream_id BE bytes 0..2]
                        const uint32_t lo = h ? hlo : v, hi = h ? ((sidbe >> 24) | (v << 8)) : hlo;
                        sout_w[wS + h] = __funnelshift_rc(lo, hi, 32 - 8 * (S & 3));
                    }
                    const uint32_t w0 = wS + 2;
                    if (wE > w0) {
                        const int32_t delta = (int32_t)P0 - (int32_t)(S + 5);                  // source byte = output byte + delta
                        const int32_t dq = delta >> 2;
                        const unsigned sh = (unsigned)(delta & 3) * 8;
                        const uint32_t per = (wE - w0 + 1) >> 1, a = w0 + h * per;
                        const uint32_t b = a + per < wE ? a + per : wE;
                        if (a < b) {
                            uint32_t lo = sin.word((uint32_t)((int32_t)a + dq));
                            for (uint32_t w = a; w < b; w++) {
                                const uint32_t hi = sin.word((uint32_t)((int32_t)w + dq + 1));
                                sout_w[w] = __funnelshift_r(lo, hi, sh);
                                lo = hi;
                            }
                        }
                    }
                }
            } else {
                if (sout_dirty) {                                      // the whole-word shuffle leaves its words behind
                    for (uint32_t w = t; w < sizeof(sout_w) / 4; w += 256) sout_w[w] = 0;
                    __syncthreads();
                    sout_dirty = false;
                }
                // Two lanes per frame, one aligned output WORD at a time.  Words that lie wholly inside the payload are
                // two source words and a funnel shift (the byte shift is the same for the whole frame); the at most three
                // words at the head of the frame (header bytes, start of the payload) and the partial word at its end are
                // built with byte masks and OR-ed in, because the neighbouring frame owns the other bytes of such a word.
                const int f = t >> 1, h = t & 1;
                if (f < T) {
                    const uint32_t po = (uint32_t)(s_po[f] - pb), len = (uint32_t)(s_po[f + 1] - s_po[f]);
                    const uint32_t S = oa + (uint32_t)(s_oo[f] - ob), E = S + 5 + len;
                    const int32_t delta = (int32_t)(kSinFront + ia + po) - (int32_t)(S + 5);   // source byte = output byte + delta
                    const int32_t dq = delta >> 2;                                             // (4 w + delta) >> 2 = w + dq
                    const unsigned sh = (unsigned)(delta & 3) * 8;
                    const uint32_t wS = S >> 2, w_lo = (S + 8) >> 2, w_hi = E >> 2, wE = (E + 3) >> 2;
                    if (w_hi > w_lo) {                                 // words wholly inside the payload: a contiguous run per lane
                        const uint32_t per = (w_hi - w_lo + 1) >> 1, a = w_lo + h * per;
                        const uint32_t b = a + per < w_hi ? a + per : w_hi;
                        if (a < b) {
                            uint32_t lo = sin.word((uint32_t)((int32_t)a + dq));
                            for (uint32_t w = a; w < b; w++) {
                                const uint32_t hi = sin.word((uint32_t)((int32_t)w + dq + 1));
                                sout_w[w] = __funnelshift_r(lo, hi, sh);
                                lo = hi;
                            }
                        }
                    }
                    // edge words: lane 0 the head of the frame, lane 1 the partial word at its end
                    const uint32_t he = w_lo < wE ? w_lo : wE;
                    const uint32_t e0 = h == 0 ? wS : w_hi, e1 = h == 0 ? he : ((w_hi >= w_lo && 4 * w_hi < E) ? w_hi + 1 : w_hi);
                    if (e0 < e1) {
                        const uint32_t sid = s_sid[f];
                        const u64 H = (u64)s_type[f] | ((u64)__byte_perm(sid, 0, 0x0123) << 8);     // [type][stream_id BE]
                        for (uint32_t w = e0; w < e1; w++) {
                            const int32_t rel = (int32_t)(4 * w) - (int32_t)S;             // position of the word's byte 0 in header || payload
                            const uint32_t sa = (uint32_t)((int32_t)(4 * w) + delta);
                            const uint32_t pw = __funnelshift_r(sin.word(sa >> 2), sin.word((sa >> 2) + 1), (sa & 3) * 8);
                            const uint32_t hw = rel >= 5 ? 0u : rel >= 0 ? (uint32_t)(H >> (8 * rel)) : (uint32_t)(H << (8 * -rel));
                            // bytes j of the word with lo <= rel + j < hi, as a mask
                            auto bytes = [rel](int32_t lo, int32_t hi) -> uint32_t {
                                int32_t j0 = lo - rel, j1 = hi - rel;
                                j0 = j0 < 0 ? 0 : j0; j1 = j1 > 4 ? 4 : j1;
                                if (j1 <= j0) return 0u;
                                const uint32_t upto = j1 >= 4 ? 0xffffffffu : ((1u << (8 * j1)) - 1u);
                                return upto & ~((1u << (8 * j0)) - 1u);
                            };
                            const uint32_t v = (hw & bytes(0, 5)) | (pw & bytes(5, 5 + (int32_t)len));
                            atomicOr(&sout_w[w], v);
                        }
                    }
                }
            }
            __syncthreads();
            const uint32_t total = oa + (uint32_t)ospan;               // [oa, total) of sout is the span
            uint8_t* g0 = gdst - oa;
            for (uint32_t lo = 16u * t; lo < total; lo += 16u * 256u) {
                uint4* q4 = reinterpret_cast<uint4*>(sout_w + (lo >> 2));
                const uint4 x = *q4;
                if (!whole) *q4 = make_uint4(0, 0, 0, 0);              // zero again for the next tile's OR-ed words
                if (lo >= oa && lo + 16 <= total) stg128(g0 + lo, x);
                else store_bytes(g0 + lo, x, lo < oa ? (int)(oa - lo) : 0, total - lo < 16u ? (int)(total - lo) : 16);
            }
        } else {
            // slack between the slots, or a span beyond the buffers: four lanes per frame, straight to global memory
            for (int f = t >> 2; f < T; f += 64) {
                const u64 po = s_po[f];
                const uint32_t len = (uint32_t)(s_po[f + 1] - po);
                uint8_t* o = out + s_oo[f];
                if ((t & 3) == 0) {
                    const uint32_t sid = s_sid[f];
                    o[0] = s_type[f];
                    o[1] = (uint8_t)(sid >> 24); o[2] = (uint8_t)(sid >> 16); o[3] = (uint8_t)(sid >> 8); o[4] = (uint8_t)sid;
                }
                if (len) quad_copy_bytes(o + 5, payload + po, len, t & 3);
            }
            __syncthreads();                                           // s_po / s_oo are rewritten for the next tile
        }
    }
}

// ---- decode: one thread per frame (offsets coalesced; the header itself is two aligned word loads) ------------
__global__ void __launch_bounds__(256) frame_decode_kernel(
    u64 n, const uint8_t* __restrict__ in, const u64* __restrict__ in_off, p2pt_desc* __restrict__ desc,
    const uint32_t* run_flag) {
    if (run_flag && *run_flag == 0) return;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const u64 o = __ldg(in_off + i), e = __ldg(in_off + i + 1);
        const u64 len = e - o;
        const uint8_t* f = in + o;
        uint32_t type = 0, sid = 0, status, plen = 0;
        u64 poff = o;
        if (len < 5) {                                  // protocol.rs:158-160
            status = P2PT_ST_TOO_SHORT;
            if (len) type = __ldg(f);
        } else {
            // the five header bytes lie inside the two aligned words that start at the frame's word: two loads, one
            // funnel and one byte permute instead of five byte loads and four shifts
            const unsigned sh = (unsigned)((uintptr_t)f & 3);
            const uint32_t* w = reinterpret_cast<const uint32_t*>(f - sh);
            const u64 x = (((u64)__ldg(w + 1) << 32) | __ldg(w)) >> (8 * sh);
            type = (uint32_t)x & 0xffu;
            if (!msg_type_valid(type)) {                // protocol.rs:161-163
                status = P2PT_ST_UNKNOWN_TYPE;
            } else {
                status = P2PT_ST_OK;
                sid = __byte_perm((uint32_t)(x >> 8), 0, 0x0123);                // :165 get_u32 (BE)
                poff = o + 5;
                plen = (uint32_t)(len - 5);
            }
        }
        // one 16-byte + one 8-byte store per descriptor
        uint4 lo;
        lo.x = (uint32_t)poff; lo.y = (uint32_t)(poff >> 32); lo.z = plen; lo.w = sid;
        uint2 hi;
        hi.x = type | (status << 8); hi.y = 0;
        uint8_t* d = reinterpret_cast<uint8_t*>(desc + i);
        *reinterpret_cast<uint2*>(d) = make_uint2(lo.x, lo.y);
        *reinterpret_cast<uint2*>(d + 8) = make_uint2(lo.z, lo.w);
        *reinterpret_cast<uint2*>(d + 16) = hi;
    }
}

// ---- demux ----------------------------------------------------------------------------------------
// Stable LSD radix sort of (key, frame index), key = stream_id or (session << 32 | stream_id) when the
// batch spans several tunnel sessions (each session has its own stream_id namespace,
// /root/reference/tunnel/src/proxy.rs:52), with a last pass on the "not decodable" flag, so the
// status-OK frames end up first, grouped by key ascending, arrival order kept inside a stream.
// Each pass = per-block digit histogram -> scan -> stable scatter with warp match/ballot ranking.
// A pass whose digit is the same for every key degenerates to a copy.
//
// Scratch layout (p2pt_frame_demux_scratch_bytes): two key and two index arrays for ping-pong, the
// [digit][block] histogram matrix, and the block sums of the final (stream count, body bytes) scan.
//
// Packed mode: without a session array the key is the 32-bit stream id, so key and frame index travel as
// ONE 64-bit word (stream_id << 32 | index) and the index arrays are not touched at all: a sorting pass
// moves 8 bytes per frame instead of 12 (the digit of key byte b is then byte 4 + b of the word).

constexpr int kSortBlocks = 592;        // 4 x 148 SMs
constexpr int kSortThreads = 256;
#ifndef P2PT_SORT_ITEMS
#define P2PT_SORT_ITEMS 4
#endif
constexpr int kSortItems = P2PT_SORT_ITEMS;   // keys per thread per chunk
constexpr int kRadix = 256;
constexpr int kScanThreads = 1024;
constexpr int kKeyPasses = 8;           // 8 key bytes, then the validity pass
constexpr int kSortPasses = kKeyPasses + 1;

struct DemuxScratch {
    u64* keys[2];
    uint32_t* idx[2];
    uint32_t* hist;        // kRadix * kSortBlocks
    u64* blk_bytes;        // per scan block: body bytes
    uint32_t* blk_heads;   // per scan block: stream heads
    u64* misc;             // [1] = OR of keys, [2] = OR of ~keys, [3] = number of undecodable frames
    uint32_t* row_total;   // kRadix digit totals of the current pass
    uint32_t* blen;        // per frame, arrival order: body bytes it contributes to the gather, kNotDecodable if status != ok
};
constexpr uint32_t kNotDecodable = 0xffffffffu;

__host__ __device__ constexpr u64 align256(u64 x) { return (x + 255) & ~(u64)255; }

constexpr u64 demux_scan_blocks(u64 n) { return (n + kScanThreads - 1) / kScanThreads; }

constexpr u64 demux_sort_scratch_bytes(u64 n) {
    u64 b = 0;
    b += 2 * align256(n * sizeof(u64));
    b += 2 * align256(n * sizeof(uint32_t));
    b += align256((u64)kRadix * kSortBlocks * sizeof(uint32_t));
    b += align256((demux_scan_blocks(n) + 1) * sizeof(u64));
    b += align256((demux_scan_blocks(n) + 1) * sizeof(uint32_t));
    b += align256(256 + kRadix * sizeof(uint32_t));      // control words + row totals
    b += align256(n * sizeof(uint32_t));                 // blen
    return b;
}

inline DemuxScratch demux_carve(void* scratch, u64 n) {
    uint8_t* p = (uint8_t*)scratch;
    DemuxScratch s;
    s.keys[0] = (u64*)p; p += align256(n * sizeof(u64));
    s.keys[1] = (u64*)p; p += align256(n * sizeof(u64));
    s.idx[0] = (uint32_t*)p; p += align256(n * sizeof(uint32_t));
    s.idx[1] = (uint32_t*)p; p += align256(n * sizeof(uint32_t));
    s.hist = (uint32_t*)p; p += align256((u64)kRadix * kSortBlocks * sizeof(uint32_t));
    s.blk_bytes = (u64*)p; p += align256((demux_scan_blocks(n) + 1) * sizeof(u64));
    s.blk_heads = (uint32_t*)p; p += align256((demux_scan_blocks(n) + 1) * sizeof(uint32_t));
    s.misc = (u64*)p; p += 256;
    s.row_total = (uint32_t*)p; p += align256(256 + kRadix * sizeof(uint32_t)) - 256;
    s.blen = (uint32_t*)p;
    return s;
}

// digit of element i of the pass's source buffers; the index array is only touched by the validity pass
__device__ __forceinline__ uint32_t demux_digit(const u64* keys, const uint32_t* idx, u64 i, int pass, const uint32_t* blen, bool packed) {
    if (pass < kKeyPasses) return (uint32_t)(__ldcg(keys + i) >> (8 * pass + (packed ? 32 : 0))) & 0xffu;
    const uint32_t ix = packed ? (uint32_t)__ldcg(keys + i) : __ldcg(idx + i);
    return __ldg(blen + ix) == kNotDecodable ? 1u : 0u;
}

__global__ void __launch_bounds__(256) demux_init_kernel(u64 n, const p2pt_desc* __restrict__ desc,
                                                         const uint32_t* __restrict__ session,
                                                         u64* __restrict__ keys, uint32_t* __restrict__ idx,
                                                         u64* __restrict__ misc, uint32_t* __restrict__ blen,
                                                         const uint32_t* run_flag) {
    if (run_flag && *run_flag == 0) return;
    u64 vor = 0, vnor = 0;
    uint32_t bad = 0;
    const u64 stride = (u64)gridDim.x * blockDim.x;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const p2pt_desc d = desc[i];
        u64 k = d.stream_id;
        if (session) {
            k |= (u64)session[i] << 32;
            keys[i] = k;
            idx[i] = (uint32_t)i;
        } else {
            keys[i] = (k << 32) | (u64)(uint32_t)i;      // packed mode
        }
        vor |= k; vnor |= ~k;
        const bool ok = d.status == P2PT_ST_OK;
        bad += !ok;
        // ReqBody / ResBody payloads are what the gather lays out (serve.rs:120-124)
        blen[i] = !ok ? kNotDecodable : (d.msg_type == 11 || d.msg_type == 21) ? d.pay_len : 0u;
    }
    bad = __reduce_add_sync(0xffffffffu, bad);
    if ((threadIdx.x & 31) == 0 && bad) atomicAdd(misc + 3, (u64)bad);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        vor |= __shfl_xor_sync(0xffffffffu, vor, o);
        vnor |= __shfl_xor_sync(0xffffffffu, vnor, o);
    }
    if ((threadIdx.x & 31) == 0) { atomicOr(misc + 1, vor); atomicOr(misc + 2, vnor); }
}

// a key byte that is identical in every key needs no sorting pass; neither does the validity pass
// when every frame decoded
__device__ __forceinline__ bool demux_pass_trivial(const u64* misc, int pass) {
    if (pass >= kKeyPasses) return misc[3] == 0;
    return (((misc[1] & misc[2]) >> (8 * pass)) & 0xffu) == 0;
}
__device__ __forceinline__ u64 demux_valid_count(u64 n, const u64* misc) { return n - misc[3]; }
// A trivial pass moves nothing: it leaves the data in the buffer it is in.  Which of the two
// ping-pong buffers holds the input of `pass` (pass == kSortPasses: the final result) is the parity
// of the number of real passes before it.
__device__ __forceinline__ int demux_src_buffer(const u64* misc, int pass) {
    int b = 0;
    for (int p = 0; p < pass; p++) b ^= demux_pass_trivial(misc, p) ? 0 : 1;
    return b;
}
struct DemuxBufs { u64* keys[2]; uint32_t* idx[2]; };

// run_flag (nullable): the kernels of the sorting path are also enqueued behind the direct path
// (demux_direct_kernels.cuh) as its fallback; they then run only if that path raised its flag.
__device__ __forceinline__ bool demux_skip(const uint32_t* run_flag) { return run_flag && *run_flag == 0; }

__device__ __forceinline__ void demux_tile(u64 n, u64& lo, u64& hi) {
    const u64 per = (n + kSortBlocks - 1) / kSortBlocks;
    const u64 per_r = (per + kSortThreads - 1) / kSortThreads * kSortThreads;
    lo = (u64)blockIdx.x * per_r;
    hi = lo + per_r;
    if (lo > n) lo = n;
    if (hi > n) hi = n;
}

// Digit ranking uses MATCH.ANY.  On B200 it costs about 8 cycles per DISTINCT value in the warp (a hist
// pass over 10 M keys: 104 us with 256 evenly used digits, 66 us with 16); eight ballots + LOP3 cost the
// same whatever the data and made that pass 62 us in isolation, but their longer dependent chain made the
// latency-bound cooperative sort slower where it matters (1 M frames: demux + gather 0.65 -> 0.70 ms;
// 10 M: 1.11 -> 1.13 ms), so MATCH.ANY stays.
// The three phases of a sorting pass, as device functions so that they can run either as three
// launches or inside one cooperative kernel (grid-wide syncs in between).  Everything another block
// wrote (keys, indices, the histogram matrix, row totals) is read with __ldcg, i.e. from L2: inside a
// single kernel an L1 line cached in an earlier pass would be stale.
__device__ __forceinline__ void demux_hist_phase(u64 n, const DemuxBufs& bufs, const uint32_t* blen, int pass,
                                                 uint32_t* hist, const u64* misc, bool packed) {
    const int sb = demux_src_buffer(misc, pass);
    const u64* keys = bufs.keys[sb];
    const uint32_t* idx = bufs.idx[sb];
    // per-warp private counters: the leaders of a warp's match groups hold distinct digits, so a plain
    // load + store per group is race-free inside the warp and no shared-memory atomics are needed
    // (measured the same speed as warp-aggregated shared atomics; kept because it has no worst case)
    constexpr int W = kSortThreads / 32;
    __shared__ uint32_t wh[W][kRadix];
    const int warp = threadIdx.x >> 5;
#pragma unroll
    for (int w = 0; w < W; w++) wh[w][threadIdx.x] = 0;
    __syncthreads();
    u64 lo, hi;
    demux_tile(n, lo, hi);
    // the key words of the next chunk are in flight while this one is ranked (the loop is latency-bound otherwise)
    u64 nkey[kSortItems];
#pragma unroll
    for (int r = 0; r < kSortItems; r++) {
        const u64 i = lo + r * kSortThreads + threadIdx.x;
        nkey[r] = i < hi ? __ldcg(keys + i) : 0;
    }
    for (u64 c0 = lo; c0 < hi; c0 += kSortThreads * kSortItems) {
        uint32_t d[kSortItems];
        bool live[kSortItems];
        u64 key[kSortItems];
#pragma unroll
        for (int r = 0; r < kSortItems; r++) {
            key[r] = nkey[r];
            const u64 i = c0 + kSortThreads * kSortItems + r * kSortThreads + threadIdx.x;
            nkey[r] = i < hi ? __ldcg(keys + i) : 0;
        }
#pragma unroll
        for (int r = 0; r < kSortItems; r++) {
            const u64 i = c0 + r * kSortThreads + threadIdx.x;
            live[r] = i < hi;
            d[r] = 0;
            if (live[r]) {
                if (pass < kKeyPasses) d[r] = (uint32_t)(key[r] >> (8 * pass + (packed ? 32 : 0))) & 0xffu;
                else d[r] = __ldg(blen + (packed ? (uint32_t)key[r] : __ldcg(idx + i))) == kNotDecodable ? 1u : 0u;
            }
        }
#pragma unroll
        for (int r = 0; r < kSortItems; r++) {
            const uint32_t act = __ballot_sync(0xffffffffu, live[r]);
            if (live[r]) {
                const uint32_t peers = __match_any_sync(act, d[r]);
                if ((__ffs(peers) - 1) == (int)(threadIdx.x & 31)) wh[warp][d[r]] += __popc(peers);
            }
            __syncwarp();
        }
    }
    __syncthreads();
    uint32_t tot = 0;
#pragma unroll
    for (int w = 0; w < W; w++) tot += wh[w][threadIdx.x];
    hist[(u64)threadIdx.x * kSortBlocks + blockIdx.x] = tot;
}

// Per digit row: exclusive scan over the kSortBlocks block counts (one warp per row, all loads of
// the row issued up front) and the row total.  The scan ACROSS rows is redone by every scatter
// block in shared memory (256 values), which is cheaper than a serial single-block pass.
constexpr int kRowScanWarps = 4;
__device__ __forceinline__ void demux_rowscan_row(uint32_t* hist, int row, uint32_t* row_total) {
    constexpr int CH = (kSortBlocks + 31) / 32;
    const int lane = threadIdx.x & 31;
    uint32_t* r = hist + (u64)row * kSortBlocks;
    uint32_t v[CH];
#pragma unroll
    for (int i = 0; i < CH; i++) { const int b = 32 * i + lane; v[i] = b < kSortBlocks ? __ldcg(r + b) : 0u; }
    uint32_t carry = 0;
#pragma unroll
    for (int i = 0; i < CH; i++) {
        uint32_t incl = v[i];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        const int b = 32 * i + lane;
        if (b < kSortBlocks) r[b] = carry + incl - v[i];
        carry += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 0) row_total[row] = carry;
}

__device__ __forceinline__ void demux_scatter_phase(u64 n, const DemuxBufs& bufs, const uint32_t* blen, int pass,
                                                    const uint32_t* hist, const u64* misc, const uint32_t* row_total, bool packed) {
    const int sb = demux_src_buffer(misc, pass);
    const u64* keys = bufs.keys[sb];
    const uint32_t* idx = bufs.idx[sb];
    u64* keys_out = bufs.keys[sb ^ 1];
    uint32_t* idx_out = bufs.idx[sb ^ 1];
    u64 lo, hi;
    demux_tile(n, lo, hi);
    static_assert(kSortThreads == kRadix, "one thread per digit");
    constexpr int W = kSortThreads / 32;
    constexpr int kChunk = kSortThreads * kSortItems;
    // Chunks of W x 32 x kSortItems consecutive keys; warp w owns 32 x kSortItems consecutive ones and
    // takes them in rounds of 32 (lane order = arrival order).  The chunk is first sorted by digit into
    // shared memory (position = start of the digit inside the chunk + keys of that digit in earlier
    // warps + rank inside the warp), then written out in that order: neighbouring threads then store
    // to neighbouring addresses inside a digit's run, whole sectors instead of one partial sector per key
    // (measured on 10 M keys with 256 evenly used digits: 253 -> 159 us per pass).
    __shared__ uint32_t wcount[W][kRadix];     // per-warp digit counts of the current chunk, then their prefix over warps
    __shared__ uint32_t dstart[kRadix];        // first chunk-local position of each digit
    __shared__ uint32_t gofs[kRadix];          // global position of a key = gofs[digit] + its chunk-local position
    __shared__ uint32_t wtot[W];
    __shared__ u64 skey[kChunk];
    __shared__ uint32_t sidx[kChunk];
    __shared__ uint8_t sdig[kChunk];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int col = threadIdx.x;
    uint32_t base;                             // running output position of digit `col` for this block
    {
        // exclusive scan of the 256 row totals -> where each digit's output range starts
        const uint32_t v = __ldcg(row_total + threadIdx.x);
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) wtot[warp] = incl;
        __syncthreads();
        uint32_t wbase = 0;
#pragma unroll
        for (int w = 0; w < W; w++) wbase += (w < warp) ? wtot[w] : 0u;
        base = wbase + incl - v + __ldcg(hist + (u64)threadIdx.x * kSortBlocks + blockIdx.x);
    }
    u64 nkey[kSortItems];                      // next chunk's key words, in flight while this chunk is ranked and written
#pragma unroll
    for (int r = 0; r < kSortItems; r++) {
        const u64 i = lo + (u64)warp * (32 * kSortItems) + r * 32 + lane;
        nkey[r] = i < hi ? __ldcg(keys + i) : 0;
    }
    for (u64 c0 = lo; c0 < hi; c0 += kChunk) {
#pragma unroll
        for (int w = 0; w < W; w++) wcount[w][col] = 0;
        u64 key[kSortItems];
        uint32_t ix[kSortItems], d[kSortItems], rank[kSortItems];
        bool live[kSortItems];
#pragma unroll
        for (int r = 0; r < kSortItems; r++) {
            const u64 i = c0 + (u64)warp * (32 * kSortItems) + r * 32 + lane;
            live[r] = i < hi;
            key[r] = nkey[r];
            ix[r] = 0;
            if (live[r]) ix[r] = packed ? (uint32_t)key[r] : __ldcg(idx + i);
            const u64 inext = i + kChunk;
            nkey[r] = inext < hi ? __ldcg(keys + inext) : 0;
        }
#pragma unroll
        for (int r = 0; r < kSortItems; r++) {
            d[r] = 0;
            if (live[r])
                d[r] = pass < kKeyPasses ? (uint32_t)(key[r] >> (8 * pass + (packed ? 32 : 0))) & 0xffu
                                         : (__ldg(blen + ix[r]) == kNotDecodable ? 1u : 0u);
        }
        __syncthreads();                                 // counters zeroed; the previous chunk's write-out is done
        // rank among equal digits inside the warp's 32 x kSortItems keys, in arrival order
#pragma unroll
        for (int r = 0; r < kSortItems; r++) {
            const uint32_t act = __ballot_sync(0xffffffffu, live[r]);
            uint32_t peers = 0, before = 0;
            if (live[r]) {
                peers = __match_any_sync(act, d[r]);
                before = wcount[warp][d[r]];
                rank[r] = before + __popc(peers & ((1u << lane) - 1));
            }
            __syncwarp();
            if (live[r] && (peers >> lane) == 1u) wcount[warp][d[r]] = before + __popc(peers);   // highest peer lane
            __syncwarp();
        }
        __syncthreads();
        // exclusive prefix over warps for this thread's digit column, then over the digits
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < W; w++) { const uint32_t t = wcount[w][col]; wcount[w][col] = run; run += t; }
        uint32_t incl = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) wtot[warp] = incl;
        __syncthreads();
        {
            uint32_t wbase = 0;
#pragma unroll
            for (int w = 0; w < W; w++) wbase += (w < warp) ? wtot[w] : 0u;
            const uint32_t ds = wbase + incl - run;
            dstart[col] = ds;
            gofs[col] = base - ds;
            base += run;
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < kSortItems; r++) {
            if (live[r]) {
                const uint32_t lp = dstart[d[r]] + wcount[warp][d[r]] + rank[r];
                skey[lp] = key[r]; sdig[lp] = (uint8_t)d[r];
                if (!packed) sidx[lp] = ix[r];
            }
        }
        __syncthreads();
        const uint32_t cnt = hi - c0 < (u64)kChunk ? (uint32_t)(hi - c0) : (uint32_t)kChunk;
#pragma unroll
        for (int r = 0; r < kSortItems; r++) {
            const uint32_t t = r * kSortThreads + threadIdx.x;
            if (t < cnt) {
                const uint32_t o = gofs[sdig[t]] + t;
                keys_out[o] = skey[t];
                if (!packed) idx_out[o] = sidx[t];
            }
        }
    }
}

// three launches per pass (fallback path) ...
__global__ void __launch_bounds__(kSortThreads) demux_hist_kernel(
    u64 n, const DemuxBufs bufs, const uint32_t* __restrict__ blen,
    int pass, uint32_t* __restrict__ hist, const u64* __restrict__ misc, bool packed, const uint32_t* run_flag) {
    if (demux_skip(run_flag) || demux_pass_trivial(misc, pass)) return;
    demux_hist_phase(n, bufs, blen, pass, hist, misc, packed);
}
__global__ void __launch_bounds__(32 * kRowScanWarps) demux_rowscan_kernel(uint32_t* __restrict__ hist, int pass,
                                                                          u64* __restrict__ misc,
                                                                          uint32_t* __restrict__ row_total, const uint32_t* run_flag) {
    if (demux_skip(run_flag) || demux_pass_trivial(misc, pass)) return;
    const int row = blockIdx.x * kRowScanWarps + (threadIdx.x >> 5);
    if (row < kRadix) demux_rowscan_row(hist, row, row_total);
}
__global__ void __launch_bounds__(kSortThreads) demux_scatter_kernel(
    u64 n, const DemuxBufs bufs, const uint32_t* __restrict__ blen,
    int pass, const uint32_t* __restrict__ hist, const u64* __restrict__ misc,
    const uint32_t* __restrict__ row_total, bool packed, const uint32_t* run_flag) {
    if (demux_skip(run_flag) || demux_pass_trivial(misc, pass)) return;          // nothing to move: the data stays where it is
    demux_scatter_phase(n, bufs, blen, pass, hist, misc, row_total, packed);
}

// ... or all passes in ONE cooperative launch: the nine passes are launch-latency bound at 1 M items
// (27 launches of a few microseconds each), so the phases are separated by grid-wide syncs instead.
__global__ void __launch_bounds__(kSortThreads, 4) demux_sort_coop_kernel(
    u64 n, const DemuxBufs bufs, const uint32_t* blen, uint32_t* hist, const u64* misc, uint32_t* row_total, bool packed,
    const uint32_t* run_flag) {
    if (demux_skip(run_flag)) return;                     // uniform over the whole grid
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    for (int pass = 0; pass < kSortPasses; pass++) {
        if (demux_pass_trivial(misc, pass)) continue;     // uniform over the whole grid
        demux_hist_phase(n, bufs, blen, pass, hist, misc, packed);
        grid.sync();
        const int row = blockIdx.x * (kSortThreads / 32) + (threadIdx.x >> 5);
        if (row < kRadix) demux_rowscan_row(hist, row, row_total);
        grid.sync();
        demux_scatter_phase(n, bufs, blen, pass, hist, misc, row_total, packed);
        grid.sync();
    }
}

// After the sort: flag stream heads, and scan (heads, body bytes) over the m decodable frames.
struct HeadBytes { uint32_t heads; u64 bytes; };

__device__ __forceinline__ HeadBytes demux_item(u64 j, u64 m, const u64* keys, const uint32_t* idx, const uint32_t* blen) {
    HeadBytes r; r.heads = 0; r.bytes = 0;
    if (j < m) {
        r.heads = (j == 0 || keys[j - 1] != keys[j]) ? 1u : 0u;
        r.bytes = blen[idx[j]];          // the 4 n-byte array stays in L2; the 24-byte descriptors would not
    }
    return r;
}

__device__ __forceinline__ HeadBytes block_scan_excl(HeadBytes v, HeadBytes& total) {
    __shared__ uint32_t sh[32];
    __shared__ u64 sb[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t ih = v.heads; u64 ib = v.bytes;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t th = __shfl_up_sync(0xffffffffu, ih, o);
        u64 tb = __shfl_up_sync(0xffffffffu, ib, o);
        if (lane >= o) { ih += th; ib += tb; }
    }
    if (lane == 31) { sh[warp] = ih; sb[warp] = ib; }
    __syncthreads();
    if (warp == 0) {
        uint32_t wh = sh[lane]; u64 wb = sb[lane];
        uint32_t xh = wh; u64 xb = wb;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t th = __shfl_up_sync(0xffffffffu, xh, o);
            u64 tb = __shfl_up_sync(0xffffffffu, xb, o);
            if (lane >= o) { xh += th; xb += tb; }
        }
        sh[lane] = xh - wh; sb[lane] = xb - wb;
        if (lane == 31) { total.heads = xh; total.bytes = xb; }
    }
    __syncthreads();
    HeadBytes r;
    r.heads = sh[warp] + ih - v.heads;
    r.bytes = sb[warp] + ib - v.bytes;
    // broadcast total from warp 0 lane 31 through shared memory
    __shared__ uint32_t th_; __shared__ u64 tb_;
    if (threadIdx.x == 31) { th_ = total.heads; tb_ = total.bytes; }
    __syncthreads();
    total.heads = th_; total.bytes = tb_;
    return r;
}

// Four consecutive sorted positions per thread (vector loads, four independent blen look-ups in
// flight, a quarter of the block-wide scans per item).
constexpr int kDemuxItems = 4;
constexpr u64 demux_emit_blocks(u64 n) { return (n + (u64)kScanThreads * kDemuxItems - 1) / ((u64)kScanThreads * kDemuxItems); }

struct DemuxQuad { u64 key[kDemuxItems]; uint32_t ix[kDemuxItems]; HeadBytes it[kDemuxItems]; HeadBytes sum; };

// first_only: only the first frame of every key contributes body bytes (the SCTP reassembly, where a second
// packet with the same (session, TSN) is a retransmission to be dropped)
__device__ __forceinline__ DemuxQuad demux_load_quad(u64 j, u64 m, u64 n, const u64* __restrict__ keys,
                                                     const uint32_t* __restrict__ idx, const uint32_t* __restrict__ blen,
                                                     bool first_only, bool packed) {
    DemuxQuad r;
    if (j + kDemuxItems <= n) {                          // arrays are 256-byte aligned, j is a multiple of 4
        const ulonglong2 k01 = *reinterpret_cast<const ulonglong2*>(keys + j);
        const ulonglong2 k23 = *reinterpret_cast<const ulonglong2*>(keys + j + 2);
        r.key[0] = k01.x; r.key[1] = k01.y; r.key[2] = k23.x; r.key[3] = k23.y;
        if (!packed) {
            const uint4 x = *reinterpret_cast<const uint4*>(idx + j);
            r.ix[0] = x.x; r.ix[1] = x.y; r.ix[2] = x.z; r.ix[3] = x.w;
        }
    } else {
#pragma unroll
        for (int e = 0; e < kDemuxItems; e++) {
            r.key[e] = j + e < n ? keys[j + e] : 0;
            r.ix[e] = (!packed && j + e < n) ? idx[j + e] : 0;
        }
    }
    if (packed) {                                        // word = stream_id << 32 | frame index
#pragma unroll
        for (int e = 0; e < kDemuxItems; e++) { r.ix[e] = (uint32_t)r.key[e]; r.key[e] >>= 32; }
    }
    // the key before this thread's first one: the neighbour lane holds it, lane 0 reads it
    u64 prev = __shfl_up_sync(0xffffffffu, r.key[kDemuxItems - 1], 1);
    if ((threadIdx.x & 31) == 0 && j > 0 && j < m) prev = packed ? keys[j - 1] >> 32 : keys[j - 1];
    r.sum.heads = 0; r.sum.bytes = 0;
#pragma unroll
    for (int e = 0; e < kDemuxItems; e++) {
        HeadBytes h; h.heads = 0; h.bytes = 0;
        if (j + e < m) {
            h.heads = (j + e == 0 || prev != r.key[e]) ? 1u : 0u;
            h.bytes = (first_only && !h.heads) ? 0u : blen[r.ix[e]];   // the 4 n-byte array stays in L2; the 24-byte descriptors would not
        }
        prev = r.key[e];
        r.it[e] = h;
        r.sum.heads += h.heads; r.sum.bytes += h.bytes;
    }
    return r;
}

__global__ void __launch_bounds__(kScanThreads) demux_blocksum_kernel(
    u64 n, const DemuxBufs bufs, const uint32_t* __restrict__ blen,
    const u64* __restrict__ misc, uint32_t* __restrict__ blk_heads, u64* __restrict__ blk_bytes, u64 nblk, bool first_only, bool packed,
    const uint32_t* run_flag) {
    if (demux_skip(run_flag)) return;
    const int fb = demux_src_buffer(misc, kSortPasses);
    const u64 m = demux_valid_count(n, misc);
    // one step per scan block; as the direct path's fallback the grid is small (a launch of thousands of blocks
    // that only look at the flag costs 10+ us) and every block takes several steps
    for (u64 vb = blockIdx.x; vb < nblk; vb += gridDim.x) {
        const u64 j = (vb * kScanThreads + threadIdx.x) * kDemuxItems;
        const DemuxQuad qd = demux_load_quad(j, m, n, bufs.keys[fb], bufs.idx[fb], blen, first_only, packed);
        HeadBytes tot;
        (void)block_scan_excl(qd.sum, tot);
        if (threadIdx.x == 0) { blk_heads[vb] = tot.heads; blk_bytes[vb] = tot.bytes; }
        __syncthreads();
    }
}

// single block: exclusive scan of the per-block sums, in place; element [nblk] receives the total
__global__ void __launch_bounds__(kScanThreads) demux_scan_blocksums_kernel(
    u64 nblk, uint32_t* __restrict__ blk_heads, u64* __restrict__ blk_bytes, const uint32_t* run_flag) {
    if (demux_skip(run_flag)) return;
    __shared__ uint32_t carry_h; __shared__ u64 carry_b;
    if (threadIdx.x == 0) { carry_h = 0; carry_b = 0; }
    __syncthreads();
    for (u64 b0 = 0; b0 < nblk; b0 += kScanThreads) {
        const u64 b = b0 + threadIdx.x;
        HeadBytes v; v.heads = b < nblk ? blk_heads[b] : 0; v.bytes = b < nblk ? blk_bytes[b] : 0;
        HeadBytes tot;
        HeadBytes ex = block_scan_excl(v, tot);
        const uint32_t ch = carry_h; const u64 cb = carry_b;
        if (b < nblk) { blk_heads[b] = ch + ex.heads; blk_bytes[b] = cb + ex.bytes; }
        __syncthreads();
        if (threadIdx.x == 0) { carry_h = ch + tot.heads; carry_b = cb + tot.bytes; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { blk_heads[nblk] = carry_h; blk_bytes[nblk] = carry_b; }
}

// Emit perm / uniq_sid / csr / gathered_off and per-frame gather destinations.
// dst[i] (in scratch, reusing the idle key buffer) = byte offset of frame i (ARRIVAL index) in gathered:
// the gather walks the input in arrival order, see demux_gather_kernel.  The offsets are stored as 32-bit
// words (saturated) when the caller's gathered_cap fits 32 bits - the scattered writes then land in an
// array half the size, which L2 absorbs - and as 64-bit words otherwise; not at all without a gather.
__global__ void __launch_bounds__(kScanThreads) demux_emit_kernel(
    u64 n, const DemuxBufs bufs, const uint32_t* __restrict__ blen,
    const u64* __restrict__ misc, const uint32_t* __restrict__ blk_heads, const u64* __restrict__ blk_bytes,
    u64 nblk, uint32_t max_streams, uint32_t* __restrict__ uniq_sid, uint32_t* __restrict__ uniq_session,
    u64* __restrict__ csr, uint32_t* __restrict__ perm, uint32_t* __restrict__ k_out,
    u64* __restrict__ gathered_off, int dst_mode, bool first_only, bool packed,    // dst_mode: 0 none, 1 = 32-bit, 2 = 64-bit offsets
    const uint32_t* run_flag) {
    if (demux_skip(run_flag)) return;
    const int fb = demux_src_buffer(misc, kSortPasses);
    u64* __restrict__ dst64 = bufs.keys[fb ^ 1];
    uint32_t* __restrict__ dst32 = reinterpret_cast<uint32_t*>(bufs.keys[fb ^ 1]);
    const u64 m = demux_valid_count(n, misc);
    for (u64 vb = blockIdx.x; vb < nblk; vb += gridDim.x) {       // see demux_blocksum_kernel
    const u64 j = (vb * kScanThreads + threadIdx.x) * kDemuxItems;
    const DemuxQuad qd = demux_load_quad(j, m, n, bufs.keys[fb], bufs.idx[fb], blen, first_only, packed);
    HeadBytes tot;
    HeadBytes ex = block_scan_excl(qd.sum, tot);
    uint32_t heads = blk_heads[vb] + ex.heads;
    u64 bytes = blk_bytes[vb] + ex.bytes;
    if (j + kDemuxItems <= m) *reinterpret_cast<uint4*>(perm + j) = make_uint4(qd.ix[0], qd.ix[1], qd.ix[2], qd.ix[3]);
#pragma unroll
    for (int e = 0; e < kDemuxItems; e++) {
        if (j + e < m) {
            if (j + kDemuxItems > m) perm[j + e] = qd.ix[e];
            const bool skip = first_only && !qd.it[e].heads;      // all-ones = beyond any capacity: the gather leaves it out
            if (dst_mode == 1) dst32[qd.ix[e]] = (skip || bytes > 0xffffffffull) ? 0xffffffffu : (uint32_t)bytes;
            else if (dst_mode == 2) dst64[qd.ix[e]] = skip ? ~0ull : bytes;
            if (qd.it[e].heads) {
                if (heads < max_streams) {
                    uniq_sid[heads] = (uint32_t)qd.key[e];
                    if (uniq_session) uniq_session[heads] = (uint32_t)(qd.key[e] >> 32);
                    csr[heads] = j + e;
                    if (gathered_off) gathered_off[heads] = bytes;
                }
            }
        }
        heads += qd.it[e].heads; bytes += qd.it[e].bytes;
    }
    if (j == 0) {
        const uint32_t k = blk_heads[nblk];
        *k_out = k;
        if (k <= max_streams) {          // csr / gathered_off hold max_streams + 1 entries
            csr[k] = m;
            if (gathered_off) gathered_off[k] = blk_bytes[nblk];
        }
    }
    __syncthreads();
    }
}

// ---- small batches: the whole demux (all but the gather) in ONE block -------------------------------
// Up to kSmallDemux frames: every thread ranks its frame by counting the decodable frames that sort in
// front of it (key order, arrival order among equal keys) - n^2 / 1024 broadcast shared-memory reads per
// thread - and the scan / emit step follows in the same kernel.  A reader loop that drains a few dozen
// frames per wake-up pays two launches instead of six and no cooperative launch.
constexpr int kSmallDemux = 1024;

__global__ void __launch_bounds__(kSmallDemux) demux_small_kernel(
    uint32_t n, const p2pt_desc* __restrict__ desc, const uint32_t* __restrict__ session,
    uint32_t max_streams, uint32_t* __restrict__ uniq_sid, uint32_t* __restrict__ uniq_session,
    u64* __restrict__ csr, uint32_t* __restrict__ perm, uint32_t* __restrict__ k_out,
    u64* __restrict__ gathered_off, u64* __restrict__ dst_buf, int dst_mode, bool first_only) {
    __shared__ u64 skey[kSmallDemux];
    __shared__ uint8_t sok[kSmallDemux];
    __shared__ u64 okey[kSmallDemux];
    __shared__ uint32_t oidx[kSmallDemux], olen[kSmallDemux];
    __shared__ uint32_t m_sh;
    const uint32_t i = threadIdx.x;
    u64 key = 0; bool ok = false; uint32_t bl = 0;
    if (i < n) {
        const p2pt_desc d = desc[i];
        key = d.stream_id;
        if (session) key |= (u64)session[i] << 32;
        ok = d.status == P2PT_ST_OK;
        bl = (d.msg_type == 11 || d.msg_type == 21) ? d.pay_len : 0u;
    }
    skey[i] = key; sok[i] = ok ? 1 : 0;
    if (i == 0) m_sh = 0;
    __syncthreads();
    if (ok) {
        uint32_t rank = 0;
        for (uint32_t j = 0; j < n; j++) {
            const u64 kj = skey[j];
            rank += (sok[j] && (kj < key || (kj == key && j < i))) ? 1u : 0u;
        }
        okey[rank] = key; oidx[rank] = i; olen[rank] = bl;
        atomicAdd(&m_sh, 1u);
    }
    __syncthreads();
    const uint32_t m = m_sh, s = i;
    HeadBytes it; it.heads = 0; it.bytes = 0;
    if (s < m) {
        it.heads = (s == 0 || okey[s - 1] != okey[s]) ? 1u : 0u;
        it.bytes = (first_only && !it.heads) ? 0u : olen[s];
    }
    HeadBytes tot;
    const HeadBytes ex = block_scan_excl(it, tot);
    if (s < m) {
        const uint32_t ix = oidx[s];
        perm[s] = ix;
        const bool skip = first_only && !it.heads;
        if (dst_mode == 1) reinterpret_cast<uint32_t*>(dst_buf)[ix] = (skip || ex.bytes > 0xffffffffull) ? 0xffffffffu : (uint32_t)ex.bytes;
        else if (dst_mode == 2) dst_buf[ix] = skip ? ~0ull : ex.bytes;
        if (it.heads && ex.heads < max_streams) {
            uniq_sid[ex.heads] = (uint32_t)okey[s];
            if (uniq_session) uniq_session[ex.heads] = (uint32_t)(okey[s] >> 32);
            csr[ex.heads] = s;
            if (gathered_off) gathered_off[ex.heads] = ex.bytes;
        }
    }
    if (i == 0) {
        *k_out = tot.heads;
        if (tot.heads <= max_streams) {
            csr[tot.heads] = m;
            if (gathered_off) gathered_off[tot.heads] = tot.bytes;
        }
    }
}

// Gather: body frames only (serve.rs:120-124 extend_from_slice).  Same 32-frames-per-warp scheme as
// the encoder: four lanes per short frame, the whole warp per long one.
// The frames are walked in ARRIVAL order and written to their place in the stream-grouped output
// (dst_off from the emit kernel): descriptors and frame bytes are then read front to back, each DRAM
// sector once, and the writes advance one front per stream, which L2 merges into whole lines.  Walking
// the sorted order instead reads 69-byte frames and 24-byte descriptors at random (measured 4.7x read
// amplification on 10 M x 64 B).
__global__ void __launch_bounds__(256, P2PT_GATHER_MIN_BLOCKS) demux_gather_kernel(
    u64 n, const u64* __restrict__ misc, const DemuxBufs bufs,
    const p2pt_desc* __restrict__ desc, const uint8_t* __restrict__ in,
    uint8_t* __restrict__ gathered, u64 gathered_cap, const uint32_t* __restrict__ k_out, uint32_t max_streams,
    int dst_mode, int rounds, const uint32_t* run_flag) {
    if (demux_skip(run_flag)) return;
    if (*k_out > max_streams) return;
    const u64* __restrict__ dst64 = bufs.keys[demux_src_buffer(misc, kSortPasses) ^ 1];
    const uint32_t* __restrict__ dst32 = reinterpret_cast<const uint32_t*>(dst64);
    const int lane = threadIdx.x & 31, quad = lane >> 2, q = lane & 3;
    const u64 warp = ((u64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u64 nwarps = ((u64)gridDim.x * blockDim.x) >> 5;
    const int tile = 8 * rounds;                       // see frame_encode_kernel
    for (u64 i0 = warp * tile; i0 < n; i0 += nwarps * tile) {
        const u64 i = i0 + lane;                       // lane l holds the metadata of frame i0 + l
        u64 so = 0, o = 0;
        uint32_t len = 0;
        if (lane < tile && i < n) {
            const p2pt_desc d = desc[i];
            if (d.status == P2PT_ST_OK && (d.msg_type == 11 || d.msg_type == 21) && d.pay_len) {
                o = dst_mode == 1 ? (u64)dst32[i] : dst64[i];      // a saturated 32-bit offset is beyond any 32-bit cap
                if (o <= gathered_cap && d.pay_len <= gathered_cap - o) { len = d.pay_len; so = d.pay_off; }   // overflow: caller sees it in gathered_off[k]
            }
        }
        if (__all_sync(0xffffffffu, len <= kQuadCopyMax)) {
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
