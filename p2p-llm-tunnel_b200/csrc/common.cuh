// common.cuh -- byte-stream helpers shared by the frame and record kernels (sm_100a).
//
// Frames and records are packed back to back at arbitrary byte offsets (a 5-byte frame header, a
// 13-byte record header and an 8-byte nonce shift everything off 16-byte alignment), so every
// kernel moves data as ALIGNED 16-byte vectors and re-aligns in registers with a byte funnel.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace p2pt {

typedef unsigned long long u64;

__device__ __forceinline__ uint4 ldg128(const void* p) {
    return __ldg(reinterpret_cast<const uint4*>(p));
}
// streaming variants: data touched once, keep it out of L1
__device__ __forceinline__ uint4 ldg128_stream(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void stg128(void* p, const uint4& v) {
    *reinterpret_cast<uint4*>(p) = v;
}

// bytes [s, s+16) of the 32-byte string a || b, 0 <= s <= 15.
__device__ __forceinline__ uint4 funnel16(const uint4& a, const uint4& b, unsigned s) {
    uint32_t w0 = a.x, w1 = a.y, w2 = a.z, w3 = a.w, w4 = b.x, w5 = b.y, w6 = b.z, w7 = b.w;
    if (s & 4) { w0 = w1; w1 = w2; w2 = w3; w3 = w4; w4 = w5; w5 = w6; w6 = w7; }
    if (s & 8) { w0 = w2; w1 = w3; w2 = w4; w3 = w5; w4 = w6; }
    const unsigned bs = (s & 3) * 8;
    uint4 r;
    r.x = __funnelshift_r(w0, w1, bs);
    r.y = __funnelshift_r(w1, w2, bs);
    r.z = __funnelshift_r(w2, w3, bs);
    r.w = __funnelshift_r(w3, w4, bs);
    return r;
}

// Load the 16 bytes at p (any alignment).  Bytes outside [lo, hi) read as zero and are never
// touched in memory.  Fast path: two aligned 16-byte loads + funnel; edges: byte loads.
__device__ __forceinline__ uint4 load16_bounded(const uint8_t* p, const uint8_t* lo, const uint8_t* hi) {
    const unsigned s = (unsigned)((uintptr_t)p & 15);
    const uint8_t* a = p - s;
    if (a >= lo && a + 32 <= hi) {
        uint4 v0 = ldg128(a);
        uint4 v1 = ldg128(a + 16);
        return funnel16(v0, v1, s);
    }
    if (s == 0 && p >= lo && p + 16 <= hi) return ldg128(p);
    uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const uint8_t* q = p + i;
        uint32_t b = (q >= lo && q < hi) ? (uint32_t)__ldg(q) : 0u;
        w[i >> 2] |= b << (8 * (i & 3));
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

__device__ __forceinline__ uint32_t word_of(const uint4& v, int i) {
    return i == 0 ? v.x : i == 1 ? v.y : i == 2 ? v.z : v.w;
}

// Store bytes [from, to) of the 16-byte block v to dst + from .. (byte stores; edges only).
__device__ __forceinline__ void store_bytes(uint8_t* dst, const uint4& v, int from, int to) {
#pragma unroll
    for (int i = 0; i < 16; i++) {
        if (i >= from && i < to) dst[i] = (uint8_t)(word_of(v, i >> 2) >> (8 * (i & 3)));
    }
}

#ifndef P2PT_COPY_BATCH
#define P2PT_COPY_BATCH 3
#endif
constexpr int kCopyBatch = P2PT_COPY_BATCH;

// funnel16 for a shift that is the same in every lane of the warp: the word part of the shift is a
// (uniform) branch instead of twelve selects
__device__ __forceinline__ uint4 funnel16_uniform(const uint4& a, const uint4& b, unsigned s) {
    uint32_t w0, w1, w2, w3, w4;
    switch (s >> 2) {
        case 0:  w0 = a.x; w1 = a.y; w2 = a.z; w3 = a.w; w4 = b.x; break;
        case 1:  w0 = a.y; w1 = a.z; w2 = a.w; w3 = b.x; w4 = b.y; break;
        case 2:  w0 = a.z; w1 = a.w; w2 = b.x; w3 = b.y; w4 = b.z; break;
        default: w0 = a.w; w1 = b.x; w2 = b.y; w3 = b.z; w4 = b.w; break;
    }
    const unsigned bs = (s & 3) * 8;
    return make_uint4(__funnelshift_r(w0, w1, bs), __funnelshift_r(w1, w2, bs),
                      __funnelshift_r(w2, w3, bs), __funnelshift_r(w3, w4, bs));
}

// Warp-cooperative copy of len bytes, any alignment on either side.  All 32 lanes must call.
// Every lane builds whole ALIGNED destination vectors from the two aligned source vectors that cover
// them (the second one is its neighbour's first and comes out of L1; an earlier version passed it
// between lanes with eight shuffles and cost 365 instructions per 1200-byte frame, 65 % of the issue
// slots).  Both source vectors of a whole destination vector hold bytes of the item, so nothing outside
// the item's own 16-byte blocks is read; the < 16 bytes at either end are copied bytewise.
__device__ __forceinline__ void warp_copy_bytes(uint8_t* dst, const uint8_t* src, uint32_t len, int lane) {
    if (len < 64) {
        for (uint32_t i = lane; i < len; i += 32) dst[i] = __ldg(src + i);
        return;
    }
    const uint32_t head = (uint32_t)((16 - ((uintptr_t)dst & 15)) & 15);
    if ((uint32_t)lane < head) dst[lane] = __ldg(src + lane);
    uint8_t* d2 = dst + head;
    const uint8_t* s2 = src + head;
    const uint32_t rem = len - head, nvec = rem >> 4;
    const unsigned s = (unsigned)((uintptr_t)s2 & 15);
    const uint8_t* sa = s2 - s;
    // kCopyBatch destination vectors per lane and pass, all loads issued before the first store
    const uint4 zero = make_uint4(0, 0, 0, 0);
    for (uint32_t base = 0; base < nvec; base += 32 * kCopyBatch) {
        uint4 a[kCopyBatch], b[kCopyBatch];
#pragma unroll
        for (int g = 0; g < kCopyBatch; g++) {
            const uint32_t j = base + 32 * g + lane;
            a[g] = j < nvec ? ldg128(sa + 16 * j) : zero;
            b[g] = (s && j < nvec) ? ldg128(sa + 16 * j + 16) : zero;
        }
#pragma unroll
        for (int g = 0; g < kCopyBatch; g++) {
            const uint32_t j = base + 32 * g + lane;
            if (j < nvec) stg128(d2 + 16 * j, s ? funnel16_uniform(a[g], b[g], s) : a[g]);
        }
    }
    for (uint32_t i = (nvec << 4) + lane; i < rem; i += 32) d2[i] = __ldg(s2 + i);
}

// Four lanes (q = 0..3) copy one short item (any alignment).  For frames of a few dozen to a few
// hundred bytes (SSE token chunks) a whole warp per item wastes 7/8 of every load; eight items per
// warp keeps the lanes busy.  All four lanes of the quad must call.
__device__ __forceinline__ void quad_copy_bytes(uint8_t* dst, const uint8_t* src, uint32_t len, int q) {
    if ((((uintptr_t)dst | len) & 15) == 0) {
        // Whole aligned destination vectors (the gather of 16-byte-multiple payloads): each one is built from
        // the two aligned source vectors that cover it; one of those is read only if it holds item bytes.
        const unsigned r = (unsigned)((uintptr_t)src & 15);
        const uint8_t* sb = src - r;
        const uint32_t nv = len >> 4;
        for (uint32_t v = q; v < nv; v += 4) {
            const uint4 a = ldg128(sb + 16 * v);
            uint4 o = a;
            if (r) o = funnel16(a, ldg128(sb + 16 * v + 16), r);      // r != 0: vector v + 1 holds item bytes [16 (v + 1) - r, ..)
            stg128(dst + 16 * v, o);
        }
        return;
    }
    // General case: head bytes up to the first aligned destination vector, whole vectors, then words and
    // bytes.  Two leaner-looking alternatives measured SLOWER on 10 M x 64 B frames at a 69-byte stride
    // (encode 0.80 ms): aligned destination vectors with piecewise stores at both ends (0.89 ms), and
    // aligned 32-bit destination words funnelled from two source words (0.88 ms, LSU-bound).
    uint32_t head = (uint32_t)((16 - ((uintptr_t)dst & 15)) & 15);
    if (head > len) head = len;
    for (uint32_t b = q; b < head; b += 4) dst[b] = __ldg(src + b);
    dst += head; src += head; len -= head;
    uint32_t nvec = len >> 4;
    const unsigned s = (unsigned)((uintptr_t)src & 15);
    const uint8_t* sa = src - s;
    if (s) {   // the second aligned source vector must not run past the end of the item
        const uint8_t* send = src + len;
        const uint32_t safe = (send >= sa + 32) ? (uint32_t)((send - (sa + 32)) >> 4) + 1 : 0;
        if (nvec > safe) nvec = safe;
    }
    for (uint32_t v = q; v < nvec; v += 4) {
        const uint4 a = ldg128(sa + 16 * v);
        uint4 o = a;
        if (s) o = funnel16(a, ldg128(sa + 16 * v + 16), s);
        stg128(dst + 16 * v, o);
    }
    // tail (< 32 bytes): aligned 32-bit destination words, each from the two aligned source words
    // that cover it (an aligned word that holds at least one byte of the item is always readable),
    // then at most three single bytes
    const uint32_t done = nvec << 4, rem = len - done;
    const uint32_t nword = rem >> 2;
    {
        const uint8_t* sp = src + done;
        const unsigned ws = (unsigned)((uintptr_t)sp & 3);
        const uint32_t* sw = reinterpret_cast<const uint32_t*>(sp - ws);
        uint32_t* dw = reinterpret_cast<uint32_t*>(dst + done);
        for (uint32_t w = q; w < nword; w += 4) {
            const uint32_t lo = __ldg(sw + w);
            const uint32_t hi = ws ? __ldg(sw + w + 1) : 0u;
            dw[w] = __funnelshift_r(lo, hi, ws * 8);
        }
    }
    for (uint32_t b = done + (nword << 2) + q; b < len; b += 4) dst[b] = __ldg(src + b);
}

constexpr uint32_t kQuadCopyMax = 256;     // items up to this size take the 4-lane path


}  // namespace p2pt
