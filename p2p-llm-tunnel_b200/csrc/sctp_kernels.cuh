// sctp_kernels.cuh -- SCTP DATA-chunk packetise / unpack with CRC32c, and message reassembly, for
// sm_100a.  This is the step that sits between the tunnel frame and the DTLS record in the
// reference's stack: webrtc-data hands the frame to webrtc-sctp 0.10.0
// (UN-VENDORED, /root/reference/tunnel/Cargo.lock:2910-2911), which cuts it into DATA chunks of at
// most 1200 bytes, one per packet here, and checksums every packet with CRC32c (crc 3.4.0,
// Cargo.lock:420-421).  Layout: RFC 4960 3.1 / 3.3.1 / appendix B.
//
// HBM-bound byte work again, and instruction-bound in practice.  HALF a warp per packet (unpack) or per message
// (packetise): what is not the CRC - offsets, geometry, staging, header work, stores - is warp-wide instruction
// stream whatever the number of lanes that need it, so two items per warp halve it per packet.  Unpack: the packet
// comes into shared memory as whole ALIGNED 16-byte vectors by asynchronous copies (cp.async), double buffered - the
// next packet's bytes are in flight while this one is checksummed - and is read where it lies (a packet that is not
// word-aligned in memory goes through a funnel shift).  Packetise: the fragment is staged through registers into a
// 16-byte aligned position behind the header words.  Either way every lane runs a table CRC over its own contiguous
// 20-word segment, the 16 partial remainders are shifted to their place with a carry-less multiply by
// x^(8 * bytes after the segment) mod P and XOR-reduced by shuffles.
#pragma once
#include "common.cuh"
#include "gcm_math.cuh"
#include "frame_kernels.cuh"
#include "../../include/p2pt.h"

namespace p2pt {

constexpr uint32_t kSctpMaxPayload = P2PT_SCTP_MAX_PAYLOAD;          // 1200
constexpr uint32_t kSctpMaxPacket = 28 + kSctpMaxPayload;            // 1228
constexpr int kSctpWarps = 32;                                       // warps (packets in flight) per block, 1 block per SM
// One staging buffer: kSctpFront bytes in front (zeros for the checksum's leading words; the packetiser builds its
// headers there), then the raw aligned vectors (15 + 1228 + 15 bytes at most), then one spare vector (an unaligned
// word read looks one word ahead).  The unpack kernel's buffers (two per half warp: double buffering).
constexpr uint32_t kSctpFront = 48;
constexpr uint32_t kSctpRawBytes = 1248;
constexpr uint32_t kSctpBufBytes = kSctpFront + kSctpRawBytes + 16;        // (padding it to 16 mod 32 words like kPkBufBytes: no gain here)
constexpr uint32_t kCrcLutBytes = 65536;                             // 256 rows x (T0 | T1) x 32 replicas x 4 B

// slice-by-4 tables and x^(8k) mod P for k = 0..1279, reflected representation; filled by the host
__device__ uint32_t g_crc_tab[4][256];
__device__ uint32_t g_crc_xpow[1280];

// Shared-memory map, like the record kernel's: the CRC tables sit on a 64 KiB boundary of the shared
// window so that one PRMT builds a lookup address (byte << 8 | lane << 2 | base); every table entry is
// replicated 32x, lane l only touches bank l, so data-dependent lookups never conflict.  The staging
// buffers fill the gap below the tables first and continue behind them (sctp_*_smem_bytes below).
struct CrcLut {
    uint32_t lb;     // 64 KiB-aligned shared address of the tables + 4 * lane
    template <int K, int T> __device__ __forceinline__ uint32_t look(uint32_t w) const {
        const uint32_t idx = __byte_perm(w, lb, 0x7604 | (K << 4));
        uint32_t v;
        if (T == 0) asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(idx));
        else        asm("ld.shared.u32 %0, [%1+128];" : "=r"(v) : "r"(idx));
        return v;
    }
};

// one 32-bit word through the CRC register: two slice-by-2 steps (T1 = T0 advanced by one byte)
__device__ __forceinline__ uint32_t crc_word(const CrcLut& t, uint32_t crc, uint32_t w) {
    const uint32_t x = crc ^ w;
    const uint32_t c1 = t.look<0, 1>(x) ^ t.look<1, 0>(x) ^ (x >> 16);
    return t.look<0, 1>(c1) ^ t.look<1, 0>(c1) ^ (c1 >> 16);
}

// (a * b) mod P for reflected 32-bit residues, b given as its pre-masked classes: carry-less product
// (integer-multiply-with-holes, gcm_math.cuh), realigned by one bit, then the upper 32 coefficients are
// folded with one word step.
__device__ __forceinline__ uint32_t crc_mulmod(const CrcLut& t, uint32_t a, const uint32_t (&bm)[kGhashClasses]) {
    const uint64_t p = clmul32(a, bm) << 1;
    return crc_word(t, 0u, (uint32_t)p) ^ (uint32_t)(p >> 32);
}

constexpr int kCrcSeg = 10;                              // g_crc_init covers packets of up to 32 x 10 words = 1280 bytes
__device__ uint32_t g_crc_init[kCrcSeg * 32 + 1];        // init-value contribution for a packet of w words

// packet word i (32-bit, little-endian load) of a staged packet
__device__ __forceinline__ uint32_t staged_word(const uint8_t* buf, uint32_t pkt0, int i) {
    const uint32_t* q = reinterpret_cast<const uint32_t*>(buf) + (pkt0 >> 2) + i;
    return __funnelshift_r(q[0], q[1], (pkt0 & 3) * 8);
}

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all_but_one() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

// ---- the packetiser stages through registers (the asynchronous-copy form of it measured slower: 1.10 instead of
//      0.93 ms per 1 M packets - byte-wise header and padding stores, and its message metadata still waits) ----
constexpr uint32_t kSctpPktAt = 20;      // packet byte 0 sits at buffer byte 20 -> payload (packet byte 28) at 48, 16-aligned;
                                         // an optional DTLS header occupies 7..19
// 16 payload bytes at src + 16 v for staging, built from the two ALIGNED source vectors that cover them
// (one funnel when the source is not 16-byte aligned).  An aligned vector is read only if it holds at
// least one byte of the item, so nothing outside the item's own 16-byte blocks is touched; with ZERO_TAIL
// the bytes past the item's end read as zero (the packetiser's padding), else they are unspecified.
// No lane ever falls back to byte loads: the two lanes at the item's end used to spend ~200
// single-lane instructions per packet there.
// (Word-granular staging - ten 32-bit loads per lane instead of three vectors - measured 30-60 % slower.)
template <bool ZERO_TAIL>
__device__ __forceinline__ uint4 stage_load16(const uint8_t* src, uint32_t v, uint32_t len) {
    const unsigned sa = (unsigned)((uintptr_t)src & 15);
    const uint8_t* base = src - sa + 16 * (size_t)v;
    const uint32_t end = len + sa;                         // item end relative to the aligned base of vector 0
    uint4 r = make_uint4(0, 0, 0, 0);
    if (16 * v < end) r = ldg128(base);
    if (sa) {
        uint4 b = make_uint4(0, 0, 0, 0);
        if (16 * v + 16 < end) b = ldg128(base + 16);
        r = funnel16(r, b, sa);
    }
    if (ZERO_TAIL && 16 * v + 16 > len) {
        const int valid = (int)len - (int)(16 * v);        // < 16 bytes of this vector belong to the item
        uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int keep = valid - 4 * k;
            w[k] = keep >= 4 ? w[k] : keep <= 0 ? 0u : (w[k] & (0xffffffffu >> (32 - 8 * keep)));
        }
        r = make_uint4(w[0], w[1], w[2], w[3]);
    }
    return r;
}

struct SctpPacketizeArgs {
    u64 n;                                  // messages
    const uint8_t* msg; const u64* msg_off;
    const uint16_t* src_port; const uint16_t* dst_port; const uint32_t* vtag;
    const uint16_t* sid; const uint16_t* ssn; const uint32_t* ppid; const uint32_t* tsn0;
    const uint16_t* epoch; const u64* seq48_first;        // both null: bare packets; else DTLS header in front
    uint8_t* out; const u64* out_msg_off;                 // byte offset of each message's first packet
    const u64* pkt_first;                                 // index of each message's first packet (n + 1)
    u64* pkt_off;                                         // out: byte offset of every packet (+ 1 terminator)
};

// A message per GROUP of kPkLanes lanes (half or quarter of a warp), like the unpack kernel's packet per half warp (see
// there): the per-message and per-packet bookkeeping is warp-wide instruction stream, several messages per warp divide
// it per packet; a lane's CRC segment grows to 320 / kPkLanes words at the same cost per packet.  One staging buffer per
// group; they fill the gap below the tables first.  Measured per 1 M packets: a warp per message 0.915 ms, half a warp
// 0.767 ms, a quarter 0.804 ms (40-word CRC chains per lane, ten staging vectors per lane: the serial part grows).
#ifndef P2PT_SCTP_PK_LANES
#define P2PT_SCTP_PK_LANES 16
#endif
constexpr int kPkLanes = P2PT_SCTP_PK_LANES;
static_assert(kPkLanes == 8 || kPkLanes == 16, "a message per quarter or half warp");
constexpr int kPkSegWords = 320 / kPkLanes;                            // kPkLanes x kPkSegWords words = 1280 bytes
constexpr uint32_t kPkBufBytes = 1344;                                // packet at byte 20 + 1228 + slack; 336 words = 16 mod 32: the two
                                                                      // halves of a warp, each in its own buffer, use disjoint banks
constexpr uint32_t kPkBufs = kSctpWarps * (32 / kPkLanes);
__host__ __device__ constexpr uint32_t sctp_packetize_smem_bytes(uint32_t slack) {
    const uint32_t low = slack / kPkBufBytes;
    return slack + kCrcLutBytes + (kPkBufs > low ? (kPkBufs - low) * kPkBufBytes : 0u);
}
// CRC32c of a packet staged word-aligned at `words` (nwords <= 320; checksum field zeroed by the caller), by the kPkLanes
// lanes of a group; every lane of the group gets the result.  The segments are aligned to the END of the packet: lane l
// owns the kPkSegWords words that end (kPkLanes - 1 - l) segments before the end, so the shift x^(8 * bytes after) is a
// per-lane constant (its classes `xpm`, computed once per kernel) and every lane runs the same number of steps; words
// before the packet start count as zero (they leave a zero register unchanged), and the 0xFFFFFFFF init value is added
// by linearity from a table indexed by the packet length.
__device__ __forceinline__ uint32_t group_crc32c_words(const CrcLut& t, const uint32_t* words, uint32_t nwords, int gl, uint32_t gmask,
                                                       const uint32_t (&xpm)[kGhashClasses]) {
    const int first = (int)nwords - (kPkLanes - gl) * kPkSegWords;   // may be negative: leading zero words
    uint32_t c = 0;
#pragma unroll 5
    for (int i = 0; i < kPkSegWords; i++) {
        const int wi = first + i;
        const uint32_t v = wi >= 0 ? words[wi] : 0u;
        c = crc_word(t, c, v);
    }
    uint32_t part = crc_mulmod(t, c, xpm);
#pragma unroll
    for (int o = kPkLanes / 2; o > 0; o >>= 1) part ^= __shfl_xor_sync(gmask, part, o, kPkLanes);
    return ~(part ^ g_crc_init[nwords]);
}
// smem byte stream [start, start + nbytes) of a 4-byte aligned buffer -> dst (any alignment) by the kPkLanes lanes of a
// group: aligned 32-bit stores built from two adjacent shared words, byte stores only at the two ends
__device__ __forceinline__ void group_stream_to_global(const uint8_t* sbuf, uint32_t start, uint32_t nbytes, uint8_t* dst, int gl) {
    uint32_t head = (uint32_t)((4 - ((uintptr_t)dst & 3)) & 3);
    if (head > nbytes) head = nbytes;
    if ((uint32_t)gl < head) dst[gl] = sbuf[start + gl];
    const uint32_t body = nbytes - head, nfull = body >> 2;
    const uint32_t so = start + head;                    // stream offset of the first aligned destination word
    const uint32_t* sw = reinterpret_cast<const uint32_t*>(sbuf) + (so >> 2);
    const unsigned sh = (so & 3) * 8;
    uint32_t* dw = reinterpret_cast<uint32_t*>(dst + head);
#pragma unroll 5
    for (int it = 0; it < kPkSegWords; it++) {
        const uint32_t j = gl + kPkLanes * it;
        if (j < nfull) dw[j] = __funnelshift_r(sw[j], sw[j + 1], sh);
    }
    for (uint32_t j = gl + 320; j < nfull; j += kPkLanes) dw[j] = __funnelshift_r(sw[j], sw[j + 1], sh);
    for (uint32_t b = (nfull << 2) + gl; b < body; b += kPkLanes) dst[head + b] = sbuf[so + b];
}

__global__ void __launch_bounds__(32 * kSctpWarps, 1) sctp_packetize_kernel(const SctpPacketizeArgs a) {
    extern __shared__ __align__(16) uint8_t sctp_dyn_smem[];
    constexpr int GPW = 32 / kPkLanes;                                   // groups (messages in flight) per warp
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, grp = lane / kPkLanes, gl = lane % kPkLanes;
    const uint32_t gmask = (kPkLanes == 16 ? 0xffffu : 0xffu) << (kPkLanes * grp);
    // ---- tables on their 64 KiB boundary, staging buffers below them first
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(sctp_dyn_smem);
    const uint32_t tbase = (sbase + 0xffffu) & ~0xffffu;
    {
        uint32_t* tb = reinterpret_cast<uint32_t*>(sctp_dyn_smem + (tbase - sbase));
        for (int idx = threadIdx.x; idx < 256 * 32; idx += blockDim.x) {
            const int x = idx >> 5, l = idx & 31;
            tb[x * 64 + l] = g_crc_tab[0][x];
            tb[x * 64 + 32 + l] = g_crc_tab[1][x];
        }
    }
    __syncthreads();
    CrcLut tabs; tabs.lb = tbase + 4u * (uint32_t)lane;
    uint32_t xpm[kGhashClasses];
    ghash_classes(g_crc_xpow[(kPkLanes - 1 - gl) * kPkSegWords * 4], xpm);   // the shift of this lane's segment
    const uint32_t n_low = (tbase - sbase) / kPkBufBytes, q = (uint32_t)(wib * GPW + grp);
    uint8_t* buf = q < n_low ? sctp_dyn_smem + q * kPkBufBytes : sctp_dyn_smem + (tbase - sbase) + kCrcLutBytes + (q - n_low) * kPkBufBytes;
    uint32_t* bw = reinterpret_cast<uint32_t*>(buf);
    const bool dtls = a.epoch != nullptr;
    const uint32_t hdr = dtls ? 13u : 0u;
    const u64 warp = (u64)blockIdx.x * kSctpWarps + wib, nwarps = (u64)gridDim.x * kSctpWarps;
    const u64 bundles = (a.n + GPW - 1) / GPW;                           // GPW consecutive messages per warp and step
    // this group's next message is pulled towards L2 while the current one is packetised (its offset was fetched a round
    // earlier, so nothing waits for it): the payload loads below then find their lines a cache level closer
    u64 mo_next = GPW * (warp + nwarps) + grp < a.n ? __ldg(a.msg_off + GPW * (warp + nwarps) + grp) : 0;
    for (u64 pp = warp; pp < bundles; pp += nwarps) {
        const u64 i = GPW * pp + grp;
        const bool have = i < a.n;
        u64 mo = 0, mlen = 0, p0 = 0, oo = 0;
        uint32_t sp = 0, dp = 0, vt = 0, sid = 0, ssn = 0, ppid = 0, t0 = 0, k = 0;
        if (have) {
            mo = __ldg(a.msg_off + i); mlen = __ldg(a.msg_off + i + 1) - mo;
            k = mlen == 0 ? 1u : mlen < 0xfffff000ull ? ((uint32_t)mlen + kSctpMaxPayload - 1) / kSctpMaxPayload
                                                      : (uint32_t)((mlen + kSctpMaxPayload - 1) / kSctpMaxPayload);
            p0 = __ldg(a.pkt_first + i); oo = __ldg(a.out_msg_off + i);
            sp = __ldg(a.src_port + i); dp = __ldg(a.dst_port + i); vt = __ldg(a.vtag + i);
            sid = __ldg(a.sid + i); ssn = __ldg(a.ssn + i); ppid = __ldg(a.ppid + i); t0 = __ldg(a.tsn0 + i);
        }
        {
            const u64 inx = GPW * (pp + nwarps) + grp;
            if (inx < a.n) for (int l = gl; l < 10; l += kPkLanes) asm volatile("prefetch.global.L2 [%0];" :: "l"(a.msg + mo_next + 128 * l));
            const u64 inn = GPW * (pp + 2 * nwarps) + grp;
            mo_next = inn < a.n ? __ldg(a.msg_off + inn) : 0;
        }
        const uint8_t* m = a.msg + mo;
        uint32_t kmax = k;
#pragma unroll
        for (int o = kPkLanes; o < 32; o <<= 1) { const uint32_t ko = __shfl_xor_sync(0xffffffffu, kmax, o); kmax = ko > kmax ? ko : kmax; }
        for (uint32_t f = 0; f < kmax; f++) {
            if (f < k) {                                                         // the same for all lanes of the group
                const u64 off = (u64)f * kSctpMaxPayload;
                const uint32_t fl = (uint32_t)(mlen - off < kSctpMaxPayload ? mlen - off : kSctpMaxPayload);
                const uint32_t padded = (fl + 3) & ~3u, plen = 28 + padded;
                __syncwarp(gmask);
                // ---- header words (packet bytes 0..27 -> buffer words 5..11), big-endian fields
                if (gl < 7) {
                    uint32_t be;
                    const uint32_t flags = (f == 0 ? 2u : 0u) | (f == k - 1 ? 1u : 0u);
                    switch (gl) {
                        case 0: be = (sp << 16) | dp; break;
                        case 1: be = vt; break;
                        case 2: be = 0; break;                                   // checksum, patched below
                        case 3: be = (0u << 24) | (flags << 16) | (16 + fl); break;
                        case 4: be = t0 + f; break;
                        case 5: be = (sid << 16) | ssn; break;
                        default: be = ppid; break;
                    }
                    bw[kSctpPktAt / 4 + gl] = bswap32(be);
                }
                if (dtls) {                                                      // DTLS record header in front
                    const uint32_t ep = __ldg(a.epoch + i);
                    const u64 sq = __ldg(a.seq48_first + i) + f;
                    for (int h = gl; h < 13; h += kPkLanes) {
                        uint32_t b;
                        switch (h) {
                            case 0: b = 23; break; case 1: b = 0xfe; break; case 2: b = 0xfd; break;
                            case 3: b = ep >> 8; break; case 4: b = ep; break;
                            case 11: b = plen >> 8; break; case 12: b = plen; break;
                            default: b = (uint32_t)(sq >> (8 * (10 - h))); break;  // bytes 5..10: seq bytes 5..0
                        }
                        buf[kSctpPktAt - 13 + h] = (uint8_t)b;
                    }
                }
                // ---- payload, zero padded, into the 16-byte aligned part of the buffer
#pragma unroll 5
                for (int it = 0; it < 80 / kPkLanes; it++) {                      // 75 vectors for a full fragment
                    const uint32_t v = gl + kPkLanes * it;
                    if (v * 16 < padded) *reinterpret_cast<uint4*>(buf + kSctpPktAt + 28 + 16 * v) = stage_load16<true>(m + off, v, fl);
                }
                __syncwarp(gmask);
                const uint32_t crc = group_crc32c_words(tabs, bw + kSctpPktAt / 4, plen >> 2, gl, gmask, xpm);
                if (gl == 0) bw[kSctpPktAt / 4 + 2] = crc;                       // stored little-endian (webrtc-sctp put_u32_le)
                __syncwarp(gmask);
                group_stream_to_global(buf, kSctpPktAt - hdr, plen + hdr, a.out + oo, gl);
                if (gl == 0) a.pkt_off[p0 + f] = oo;
                oo += plen + hdr;
            }
        }
        if (have && gl == 0 && i == a.n - 1) a.pkt_off[p0 + k] = oo;             // terminator after the last packet
    }
}

struct SctpUnpackArgs {
    u64 n;                                   // packets
    const uint8_t* in; const u64* in_off; uint32_t headroom;
    p2pt_sctp_chunk* info; p2pt_desc* desc;
    const uint32_t* tsn_base;                // nullable: per packet, the cumulative TSN ack point of its association
};

// The unpack kernel gives a packet to HALF a warp: everything that is not the CRC itself - offsets, geometry, staging,
// fix-ups, header parse, the stores of info / desc - is warp-wide instruction stream whatever the number of lanes that
// need it (it was 290 of the 492 warp instructions per packet with a warp per packet), so two packets per warp halve it
// per packet; a lane's CRC segment doubles to 20 words, which costs the same per packet.  30 warps per block: their
// 120 staging buffers (two packets x double buffering, 1312 bytes each) fill the gap below the tables first (24 warps
// with 1376-byte buffers: 0.507 ms per 1 M packets, 30 warps: 0.478 ms; 32 would need 16 bytes more than an SM has).
constexpr int kUnpackWarps = 30;
constexpr int kCrcSegHalf = 20;                                       // 16 lanes x 20 words = 1280 bytes
constexpr uint32_t kUnpackBufs = kUnpackWarps * 4;
__host__ __device__ constexpr uint32_t sctp_unpack_smem_bytes(uint32_t slack) {
    const uint32_t low = slack / kSctpBufBytes;
    return slack + kCrcLutBytes + (kUnpackBufs > low ? (kUnpackBufs - low) * kSctpBufBytes : 0u);
}

// The same for a packet that lies anywhere in its staging buffer (byte 0 at offset pkt0; ALIGNED = a multiple of 4), held by
// 16 lanes (hl = lane & 15), 20 words each: the bytes in front of the packet are zero, so there is no bounds test per word
template <bool ALIGNED>
__device__ __forceinline__ uint32_t half_crc32c(const CrcLut& t, const uint8_t* buf, uint32_t pkt0, uint32_t nwords, int hl,
                                                uint32_t hmask, const uint32_t (&xpm)[kGhashClasses]) {
    const int first = (int)nwords - (16 - hl) * kCrcSegHalf;
    uint32_t c = 0;
    if (first > -kCrcSegHalf) {
        // words before the packet start: zeros lie in front of it, but only kSctpFront bytes of them - clamp
        const uint32_t* q0 = reinterpret_cast<const uint32_t*>(buf) + (int)(pkt0 >> 2);
        const int lo = first < -(int)(kSctpFront / 4 - 4) ? -(int)(kSctpFront / 4 - 4) : first;      // still all zeros in front
        const uint32_t* q = q0 + lo;
        const int steps = first + kCrcSegHalf - lo;                   // <= 20; the skipped leading words are zero
        if (ALIGNED) {
#pragma unroll 5
            for (int i = 0; i < steps; i++) c = crc_word(t, c, q[i]);
        } else {
            const unsigned sh = (pkt0 & 3) * 8;
            uint32_t lo_w = q[0];
#pragma unroll 5
            for (int i = 0; i < steps; i++) {
                const uint32_t hi_w = q[i + 1];
                c = crc_word(t, c, __funnelshift_r(lo_w, hi_w, sh));
                lo_w = hi_w;
            }
        }
    }
    uint32_t part = crc_mulmod(t, c, xpm);
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) part ^= __shfl_xor_sync(hmask, part, o, 16);     // the other half may be elsewhere
    return ~(part ^ g_crc_init[nwords]);
}

__global__ void __launch_bounds__(32 * kUnpackWarps, 1) sctp_unpack_kernel(const SctpUnpackArgs a) {
    extern __shared__ __align__(16) uint8_t sctp_dyn_smem[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, half = lane >> 4, hl = lane & 15;
    // ---- tables on their 64 KiB boundary, staging buffers below them first
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(sctp_dyn_smem);
    const uint32_t tbase = (sbase + 0xffffu) & ~0xffffu;
    {
        uint32_t* tabs = reinterpret_cast<uint32_t*>(sctp_dyn_smem + (tbase - sbase));
        for (int idx = threadIdx.x; idx < 256 * 32; idx += blockDim.x) {
            const int x = idx >> 5, l = idx & 31;
            tabs[x * 64 + l] = g_crc_tab[0][x];
            tabs[x * 64 + 32 + l] = g_crc_tab[1][x];
        }
    }
    CrcLut tabs; tabs.lb = tbase + 4u * (uint32_t)lane;
    uint32_t xpm[kGhashClasses];
    ghash_classes(g_crc_xpow[(15 - hl) * kCrcSegHalf * 4], xpm);
    const uint32_t n_low = (tbase - sbase) / kSctpBufBytes;
    auto buffer = [&](uint32_t q) -> uint8_t* {
        return q < n_low ? sctp_dyn_smem + q * kSctpBufBytes : sctp_dyn_smem + (tbase - sbase) + kCrcLutBytes + (q - n_low) * kSctpBufBytes;
    };
    uint8_t* const buf_a = buffer(wib * 4 + half * 2);
    uint8_t* const buf_b = buffer(wib * 4 + half * 2 + 1);
    for (uint32_t w = hl; w < kSctpBufBytes / 4; w += 16) { reinterpret_cast<uint32_t*>(buf_a)[w] = 0; reinterpret_cast<uint32_t*>(buf_b)[w] = 0; }
    const uint32_t hmask = 0xffffu << (16 * half);
    __syncthreads();

    const u64 pairs = (a.n + 1) / 2;
    const u64 warp = (u64)blockIdx.x * kUnpackWarps + wib, nwarps = (u64)gridDim.x * kUnpackWarps;
    // a packet we handle: headroom + 12 .. 1228 bytes, a multiple of 4 (else it is reported malformed and not staged)
    auto sized = [&](u64 ilen) { return !(ilen < (u64)a.headroom + 12 || ilen - a.headroom > kSctpMaxPacket || ((ilen - a.headroom) & 3)); };
    auto stage_half = [&](uint8_t* buf, u64 i, u64& io, u64& ilen) {      // start the copy of packet i (this half's)
        io = 0; ilen = 0;
        if (i < a.n) { io = __ldg(a.in_off + i); ilen = __ldg(a.in_off + i + 1) - io; }
        const uint8_t* src = a.in + io + a.headroom;
        const uint32_t len = (i < a.n && sized(ilen)) ? (uint32_t)(ilen - a.headroom) : 0u;
        const uint32_t sa = (uint32_t)((uintptr_t)src & 15);
        const uint32_t nvec = len ? (sa + len + 15) >> 4 : 0;
#pragma unroll
        for (int it = 0; it < 5; it++) {                                  // <= 79 vectors over 16 lanes
            const uint32_t v = hl + 16 * it;
            if (v < nvec) cp_async16(buf + kSctpFront + 16 * v, src - sa + 16 * (size_t)v);
        }
        cp_async_commit();
    };
    int cur = 0;
    u64 io, ilen;
    stage_half(buf_a, 2 * warp + half, io, ilen);
    for (u64 pp = warp; pp < pairs; pp += nwarps) {
        const u64 i = 2 * pp + half;
        uint8_t* buf = cur ? buf_b : buf_a;
        u64 io_n, ilen_n;
        stage_half(cur ? buf_a : buf_b, 2 * (pp + nwarps) + half, io_n, ilen_n);   // the next pair flies while this one is checked
        cp_async_wait_all_but_one();
        __syncwarp();
        uint32_t st = P2PT_ST_OK;
        uint32_t flags = 0, tsn = 0, sid = 0, ssn = 0, pp_id = 0, plen_pay = 0;
        if (i < a.n) {
            if (!sized(ilen)) {
                st = P2PT_ST_SCTP_MALFORMED;                                 // not a single-DATA-chunk packet we handle
            } else {
                const uint32_t n = (uint32_t)(ilen - a.headroom);
                const uint32_t sa = (uint32_t)((uintptr_t)(a.in + io + a.headroom) & 15);
                const uint32_t pkt0 = kSctpFront + sa;
                if ((uint32_t)hl < sa) buf[kSctpFront + hl] = 0;             // bytes of the first vector in front of the packet
                __syncwarp(hmask);
                const uint32_t theirs = staged_word(buf, pkt0, 2);
                __syncwarp(hmask);
                if (hl < 4) buf[pkt0 + 8 + hl] = 0;                          // checksum field counts as zero
                __syncwarp(hmask);
                const uint32_t crc = (pkt0 & 3) ? half_crc32c<false>(tabs, buf, pkt0, n >> 2, hl, hmask, xpm)
                                                : half_crc32c<true>(tabs, buf, pkt0, n >> 2, hl, hmask, xpm);
                if (crc != theirs) st = P2PT_ST_SCTP_BAD_CHECKSUM;           // checked first, like Packet::unmarshal
                else if (n < 28 || buf[pkt0 + 12] != 0) st = P2PT_ST_SCTP_MALFORMED;
                else {
                    const uint32_t w3 = bswap32(staged_word(buf, pkt0, 3));
                    const uint32_t clen = w3 & 0xffff;
                    if (clen < 16 || 12 + ((clen + 3) & ~3u) != n) st = P2PT_ST_SCTP_MALFORMED;
                    else {
                        flags = (w3 >> 16) & 0xff; tsn = bswap32(staged_word(buf, pkt0, 4));
                        const uint32_t w5 = bswap32(staged_word(buf, pkt0, 5));
                        sid = w5 >> 16; ssn = w5 & 0xffff; pp_id = bswap32(staged_word(buf, pkt0, 6)); plen_pay = clen - 16;
                    }
                }
            }
            // every lane of the half holds the results: lanes 0..3 store the four words of the chunk info, lanes 4..6 the
            // three 8-byte words of a descriptor p2pt_frame_demux_batch can order: key = TSN, "body" type so the gather
            // takes it.  TSNs are serial numbers (RFC 4960 1.6: arithmetic modulo 2^32) and the peer picks where they
            // start, so a message may straddle 0xFFFFFFFF -> 0: with tsn_base the key is the distance from the
            // association's cumulative ack point, which orders any window of fewer than 2^32 TSNs correctly (webrtc-sctp
            // compares with sna32)
            const uint32_t key = a.tsn_base ? tsn - __ldg(a.tsn_base + i) : tsn;
            const u64 poff = st == P2PT_ST_OK ? io + a.headroom + 28 : io;
            const uint32_t plen = st == P2PT_ST_OK ? plen_pay : 0u;
            if (hl < 4) {
                const uint32_t w = hl == 0 ? tsn : hl == 1 ? (sid & 0xffffu) | (ssn << 16) : hl == 2 ? pp_id : (flags & 0xffu) | (st << 8);
                reinterpret_cast<uint32_t*>(a.info + i)[hl] = w;
            } else if (hl < 7) {
                const uint2 w = hl == 4 ? make_uint2((uint32_t)poff, (uint32_t)(poff >> 32)) : hl == 5 ? make_uint2(plen, key)
                                        : make_uint2(21u | (st << 8), 0u);
                reinterpret_cast<uint2*>(a.desc + i)[hl - 4] = w;
            }
        }
        __syncwarp();                                                        // the buffers are staged into again next round
        cur ^= 1; io = io_n; ilen = ilen_n;
    }
}

// ---- message bounds after the fragments were ordered by (session, TSN) with the demux ---------------
// Sorted position s starts a message iff its fragment carries the B flag.  Exclusive scan of the B
// flags numbers the messages; a message is complete when its fragments have consecutive TSNs in one
// session and the last one carries E.
__global__ void __launch_bounds__(kScanThreads) sctp_bounds_blocksum_kernel(
    const uint32_t* __restrict__ k_ptr, const uint32_t* __restrict__ perm, const u64* __restrict__ csr,
    const p2pt_sctp_chunk* __restrict__ info, uint32_t* __restrict__ blk) {
    const u64 k = *k_ptr, s = (u64)blockIdx.x * kScanThreads + threadIdx.x;
    HeadBytes v; v.heads = 0; v.bytes = 0;
    if (s < k) v.heads = (info[perm[csr[s]]].flags & 2) ? 1u : 0u;
    HeadBytes tot;
    (void)block_scan_excl(v, tot);
    if (threadIdx.x == 0) blk[blockIdx.x] = tot.heads;
}

__global__ void __launch_bounds__(kScanThreads) sctp_bounds_scan_kernel(u64 nblk, uint32_t* __restrict__ blk) {
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (u64 b0 = 0; b0 < nblk; b0 += kScanThreads) {
        const u64 b = b0 + threadIdx.x;
        HeadBytes v; v.heads = b < nblk ? blk[b] : 0; v.bytes = 0;
        HeadBytes tot;
        HeadBytes ex = block_scan_excl(v, tot);
        const uint32_t c = carry;
        if (b < nblk) blk[b] = c + ex.heads;
        __syncthreads();
        if (threadIdx.x == 0) carry = c + tot.heads;
        __syncthreads();
    }
    if (threadIdx.x == 0) blk[nblk] = carry;
}

__global__ void __launch_bounds__(kScanThreads) sctp_bounds_emit_kernel(
    u64 nblk, const uint32_t* __restrict__ k_ptr, const uint32_t* __restrict__ perm, const u64* __restrict__ csr,
    const u64* __restrict__ gathered_off, const uint32_t* __restrict__ uniq_session,
    const p2pt_sctp_chunk* __restrict__ info, const uint32_t* __restrict__ blk, uint32_t max_msgs,
    u64* __restrict__ msg_off, uint32_t* __restrict__ msg_first, uint8_t* __restrict__ msg_status,
    uint32_t* __restrict__ n_msgs_out) {
    const u64 k = *k_ptr, s = (u64)blockIdx.x * kScanThreads + threadIdx.x;
    HeadBytes v; v.heads = 0; v.bytes = 0;
    p2pt_sctp_chunk me; me.flags = 0; me.tsn = 0;
    if (s < k) { me = info[perm[csr[s]]]; v.heads = (me.flags & 2) ? 1u : 0u; }
    HeadBytes tot;
    HeadBytes ex = block_scan_excl(v, tot);
    const uint32_t base = blk[blockIdx.x];
    if (s < k) {
        const uint32_t before = base + ex.heads;                 // messages that started before me
        if (v.heads && before < max_msgs) { msg_off[before] = gathered_off[s]; msg_first[before] = perm[csr[s]]; }
        // which message do I belong to, and is the chain to my successor intact?
        const uint32_t mine = v.heads ? before : before - 1;     // before == 0 and no B: orphan fragment
        bool bad = (!v.heads && before == 0);
        if (!(me.flags & 1)) {                                    // not the end: successor must continue the message
            if (s + 1 >= k) bad = true;
            else {
                const p2pt_sctp_chunk nx = info[perm[csr[s + 1]]];
                const bool same_sess = !uniq_session || uniq_session[s + 1] == uniq_session[s];
                if (!same_sess || nx.tsn != me.tsn + 1 || (nx.flags & 2)) bad = true;
            }
        } else if (s + 1 < k) {                                   // the end: successor must begin a new message
            if (!(info[perm[csr[s + 1]]].flags & 2)) { /* successor is an orphan; it flags itself */ }
        }
        if (!v.heads && before > 0) {                             // continuation: predecessor must not have ended
            const p2pt_sctp_chunk pv = info[perm[csr[s - 1]]];
            if (pv.flags & 1) bad = true;
        }
        if (bad && before + (v.heads ? 1 : 0) > 0 && mine < max_msgs) msg_status[mine] = P2PT_ST_SCTP_INCOMPLETE;
    }
    if (s == 0) {
        const uint32_t nm = blk[nblk];
        *n_msgs_out = nm;
        if (nm <= max_msgs) msg_off[nm] = gathered_off[k];
    }
}

}  // namespace p2pt
