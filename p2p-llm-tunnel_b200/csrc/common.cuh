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

// Warp-cooperative copy of len bytes, any alignment on either side.  All 32 lanes must call.
__device__ __forceinline__ void warp_copy_bytes(uint8_t* dst, const uint8_t* src, u64 len, int lane) {
    if (len < 64) {
        for (u64 i = lane; i < len; i += 32) dst[i] = __ldg(src + i);
        return;
    }
    const unsigned head = (unsigned)((16 - ((uintptr_t)dst & 15)) & 15);
    if ((unsigned)lane < head) dst[lane] = __ldg(src + lane);
    uint8_t* d2 = dst + head;
    const uint8_t* s2 = src + head;
    const u64 rem = len - head;
    u64 nvec = rem >> 4;
    const unsigned s = (unsigned)((uintptr_t)s2 & 15);
    const uint8_t* sa = s2 - s;
    if (s) {
        // the second aligned source vector of the last few destination vectors may run past the
        // end of the source: leave those to the byte tail
        const uint8_t* send = src + len;
        u64 safe = (send >= sa + 32) ? (u64)((send - (sa + 32)) >> 4) + 1 : 0;
        if (nvec > safe) nvec = safe;
    }
    // Each aligned source vector is loaded exactly once.  A lane keeps three vectors in flight
    // (96 per warp pass); its right neighbour's vector comes by shuffle, the neighbour of lane 31
    // is lane 0's vector of the next group.
    const u64 nsrc = nvec + (s ? 1 : 0);
    const uint4 zero = make_uint4(0, 0, 0, 0);
    for (u64 base = 0; base < nvec; base += 96) {
        uint4 v[4];
#pragma unroll
        for (int g = 0; g < 3; g++) {
            const u64 idx = base + 32 * g + lane;
            v[g] = idx < nsrc ? ldg128_stream(sa + 16 * idx) : zero;
        }
        {   // only lane 0's copy of the 4th group is ever used (as lane 31's neighbour in group 2)
            const u64 idx = base + 96 + lane;
            v[3] = (s && lane == 0 && idx < nsrc) ? ldg128_stream(sa + 16 * idx) : zero;
        }
#pragma unroll
        for (int g = 0; g < 3; g++) {
            const u64 idx = base + 32 * g + lane;
            uint4 out = v[g];
            if (s) {
                uint4 nb, n0;
                nb.x = __shfl_down_sync(0xffffffffu, v[g].x, 1); n0.x = __shfl_sync(0xffffffffu, v[g + 1].x, 0);
                nb.y = __shfl_down_sync(0xffffffffu, v[g].y, 1); n0.y = __shfl_sync(0xffffffffu, v[g + 1].y, 0);
                nb.z = __shfl_down_sync(0xffffffffu, v[g].z, 1); n0.z = __shfl_sync(0xffffffffu, v[g + 1].z, 0);
                nb.w = __shfl_down_sync(0xffffffffu, v[g].w, 1); n0.w = __shfl_sync(0xffffffffu, v[g + 1].w, 0);
                if (lane == 31) nb = n0;
                out = funnel16(v[g], nb, s);
            }
            if (idx < nvec) stg128(d2 + 16 * idx, out);
        }
    }
    for (u64 i = (nvec << 4) + lane; i < rem; i += 32) d2[i] = __ldg(s2 + i);
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
