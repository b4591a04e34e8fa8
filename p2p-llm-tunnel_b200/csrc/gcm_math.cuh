// gcm_math.cuh -- GF(2^128) arithmetic for GHASH without a carry-less multiply instruction.
//
// Everything here is __host__ __device__ so that the exact device arithmetic can be unit-tested on
// the CPU (tests/test_host_logic.py builds a tiny host harness around this header and compares
// it with the oracle's SP 800-38D bitwise multiply).  It is NOT a CPU fallback: the library only
// calls it from kernels.
//
// Representation ("reflected integer"): a 16-byte GHASH block is read as a big-endian 128-bit
// integer X; w[3] holds bytes 0..3, w[0] bytes 12..15.  GHASH's coefficient of alpha^i is integer
// bit 127-i.  In that domain the field product is  reduce((X (x) Y) << 1)  with (x) a plain
// carry-less integer multiply (the layout used by every CLMUL implementation of GHASH, e.g.
// ghash 0.5.1 / polyval 0.6.2 behind aes-gcm 0.10.3, /root/reference/tunnel/Cargo.lock:27-29,810,1542).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define P2PT_HD __host__ __device__ __forceinline__
#else
#define P2PT_HD inline
#endif

namespace p2pt {

struct U128 { uint32_t w[4]; };   // w[0] = least-significant 32 bits of the BE integer

// The carry-less 32x32 product is built from integer multiplies of operands that keep only every S-th bit
// ("holes"): a product column then holds a COUNT of partial bits whose lowest bit is the XOR wanted, and the
// count must not reach the next kept column S bits up.
//   S = 4: 4 x 4 = 16 multiplies per product, a column of ONE multiply counts at most 8 (< 16), so the four
//          multiplies that feed an output class have to be combined with XORs (LOP3, ALU pipe).
//   S = 6: 6 x 6 = 36 multiplies, the six that feed an output class count at most 6+6+5+5+5+5 = 32 (< 64)
//          TOGETHER, so they are simply accumulated by the multiply-add itself (IMAD.WIDE with a 64-bit addend, FMA
//          pipe) and no XOR is needed: per product 16 ALU + 36 FMA instructions instead of 24 ALU + 16 FMA.  The
//          record kernel is bound by the ALU pipe while the FMA pipe idles at 21 %, which is what this trades.
#ifndef P2PT_GHASH_SPACING
#define P2PT_GHASH_SPACING 4
#endif
constexpr int kGhashClasses = P2PT_GHASH_SPACING;
static_assert(kGhashClasses == 4 || kGhashClasses == 6, "hole spacing 4 or 6");

// the classes of a 32-bit operand: out[k] = bits of v at positions k, k + S, k + 2 S, ...
P2PT_HD void ghash_classes(uint32_t v, uint32_t out[kGhashClasses]) {
    if (kGhashClasses == 4) {
        out[0] = v & 0x11111111u; out[1] = v & 0x22222222u; out[2] = v & 0x44444444u; out[3] = v & 0x88888888u;
    } else {
        out[0] = v & 0x41041041u; out[1] = v & 0x82082082u; out[2] = v & 0x04104104u;
        out[3 % kGhashClasses] = v & 0x08208208u; out[4 % kGhashClasses] = v & 0x10410410u; out[5 % kGhashClasses] = v & 0x20820820u;
    }
}

// Multiplier prepared once per (session, power): the 9 Karatsuba operands of Y, each split into its classes.
struct GhashKey { uint32_t m[9][kGhashClasses]; };

P2PT_HD void ghash_key_prepare(const U128& y, GhashKey& k) {
    uint32_t o[9];
    o[0] = y.w[0]; o[1] = y.w[1]; o[2] = y.w[0] ^ y.w[1];
    o[3] = y.w[2]; o[4] = y.w[3]; o[5] = y.w[2] ^ y.w[3];
    o[6] = y.w[0] ^ y.w[2]; o[7] = y.w[1] ^ y.w[3]; o[8] = o[6] ^ o[7];
#pragma unroll
    for (int i = 0; i < 9; i++) ghash_classes(o[i], k.m[i]);
}

// 32x32 -> 64 carry-less multiply, b given as its classes.
P2PT_HD uint64_t clmul32(uint32_t a, const uint32_t b[kGhashClasses]) {
    uint32_t ac[kGhashClasses];
    ghash_classes(a, ac);
    if (kGhashClasses == 4) {
        // every operand class keeps every 4th bit: a column of one product sums at most 8 partial bits and never
        // carries into the next kept column; the four products of an output class are XORed
        const uint32_t a0 = ac[0], a1 = ac[1], a2 = ac[2], a3 = ac[3 % kGhashClasses];
        uint64_t z0 = ((uint64_t)a0 * b[0]) ^ ((uint64_t)a1 * b[3 % kGhashClasses]) ^ ((uint64_t)a2 * b[2]) ^ ((uint64_t)a3 * b[1]);
        uint64_t z1 = ((uint64_t)a0 * b[1]) ^ ((uint64_t)a1 * b[0]) ^ ((uint64_t)a2 * b[3 % kGhashClasses]) ^ ((uint64_t)a3 * b[2]);
        uint64_t z2 = ((uint64_t)a0 * b[2]) ^ ((uint64_t)a1 * b[1]) ^ ((uint64_t)a2 * b[0]) ^ ((uint64_t)a3 * b[3 % kGhashClasses]);
        uint64_t z3 = ((uint64_t)a0 * b[3 % kGhashClasses]) ^ ((uint64_t)a1 * b[2]) ^ ((uint64_t)a2 * b[1]) ^ ((uint64_t)a3 * b[0]);
        return (z0 & 0x1111111111111111ull) | (z1 & 0x2222222222222222ull) |
               (z2 & 0x4444444444444444ull) | (z3 & 0x8888888888888888ull);
    } else {
        // every 6th bit: the six products of output class k (a_i x b_(k-i)) are ADDED - a column counts at most 32
        // partial bits, six bits wide - and the class's bit of the result is the lowest bit of its column
        uint64_t r = 0;
#pragma unroll
        for (int k = kGhashClasses - 1; k >= 0; k--) {
            uint64_t z = 0;
#pragma unroll
            for (int i = 0; i < kGhashClasses; i++) z += (uint64_t)ac[i] * b[(k - i + kGhashClasses) % kGhashClasses];
            const uint64_t mk = 0x1041041041041041ull << k;        // bits k, k + 6, ... (< 64)
            r = (z & mk) | (r & ~mk);                               // one bit-select per 32-bit half
        }
        return r;
    }
}

// 64x64 -> 128 Karatsuba from three clmul32; operands (lo, hi, lo^hi) and key rows (r, r+1, r+2).
P2PT_HD void clmul64(uint32_t lo, uint32_t hi, uint32_t mid, const GhashKey& k, int r,
                     uint64_t& out_lo, uint64_t& out_hi) {
    uint64_t p = clmul32(lo, k.m[r]);
    uint64_t q = clmul32(hi, k.m[r + 1]);
    uint64_t m = clmul32(mid, k.m[r + 2]) ^ p ^ q;
    out_lo = p ^ (m << 32);
    out_hi = q ^ (m >> 32);
}

// Z = X * Y in GF(2^128) (GHASH convention), Y given prepared.
// PRE = false: Y is the field element itself and the 256-bit integer product is shifted left by one bit (reflected
// domain) before it is folded.  PRE = true: Y was multiplied by alpha^-1 beforehand (gf128_premul, once per key power),
// which is that shift done on the multiplier: the product is folded as it is - eight funnel shifts fewer per
// multiply, on the pipe that bounds the record kernel.  (The "H << 1 mod P" form of the CLMUL implementations.)
template <bool PRE>
P2PT_HD U128 gf128_fold(uint64_t v0, uint64_t v1, uint64_t v2, uint64_t v3) {
    if (!PRE) {
        v3 = (v3 << 1) | (v2 >> 63);
        v2 = (v2 << 1) | (v1 >> 63);
        v1 = (v1 << 1) | (v0 >> 63);
        v0 = (v0 << 1);
    }
    // fold the low 128 bits with x^128 = x^7 + x^2 + x + 1 (reflected: right shifts by 1, 2, 7)
    v2 ^= v0 ^ (v0 >> 1) ^ (v0 >> 2) ^ (v0 >> 7);
    v1 ^= (v0 << 63) ^ (v0 << 62) ^ (v0 << 57);
    v3 ^= v1 ^ (v1 >> 1) ^ (v1 >> 2) ^ (v1 >> 7);
    v2 ^= (v1 << 63) ^ (v1 << 62) ^ (v1 << 57);
    U128 z;
    z.w[0] = (uint32_t)v2; z.w[1] = (uint32_t)(v2 >> 32);
    z.w[2] = (uint32_t)v3; z.w[3] = (uint32_t)(v3 >> 32);
    return z;
}

template <bool PRE>
P2PT_HD U128 gf128_mul_t(const U128& x, const GhashKey& k) {
    const uint32_t x0 = x.w[0], x1 = x.w[1], x2 = x.w[2], x3 = x.w[3];
    uint64_t pl, ph, ql, qh, rl, rh;
    clmul64(x0, x1, x0 ^ x1, k, 0, pl, ph);                           // low  half x low  half
    clmul64(x2, x3, x2 ^ x3, k, 3, ql, qh);                           // high half x high half
    clmul64(x0 ^ x2, x1 ^ x3, x0 ^ x1 ^ x2 ^ x3, k, 6, rl, rh);       // (lo^hi) x (lo^hi)
    rl ^= pl ^ ql;
    rh ^= ph ^ qh;
    return gf128_fold<PRE>(pl, ph ^ rl, ql ^ rh, qh);                 // 256-bit carry-less product v3:v2:v1:v0
}
P2PT_HD U128 gf128_mul(const U128& x, const GhashKey& k) { return gf128_mul_t<false>(x, k); }
P2PT_HD U128 gf128_mul_pre(const U128& x, const GhashKey& k) { return gf128_mul_t<true>(x, k); }

// Y * alpha^-1: one bit up in the reflected integer; the alpha^0 coefficient that falls out of the top comes back as
// alpha^-1 = alpha^127 + alpha^6 + alpha + 1 (bits 0, 121, 126, 127).
P2PT_HD U128 gf128_premul(const U128& y) {
    const uint32_t carry = y.w[3] >> 31;
    U128 r;
    r.w[3] = (y.w[3] << 1) | (y.w[2] >> 31);
    r.w[2] = (y.w[2] << 1) | (y.w[1] >> 31);
    r.w[1] = (y.w[1] << 1) | (y.w[0] >> 31);
    r.w[0] = y.w[0] << 1;
    if (carry) { r.w[0] ^= 1u; r.w[3] ^= 0xC2000000u; }
    return r;
}

P2PT_HD uint32_t bswap32(uint32_t v) {
#if defined(__CUDA_ARCH__)
    return __byte_perm(v, 0, 0x0123);
#else
    return (v >> 24) | ((v >> 8) & 0xff00u) | ((v << 8) & 0xff0000u) | (v << 24);
#endif
}

// 16 bytes held as four little-endian words (memory order) -> reflected integer, and back.
P2PT_HD U128 block_to_u128(uint32_t b0, uint32_t b1, uint32_t b2, uint32_t b3) {
    U128 r;
    r.w[3] = bswap32(b0); r.w[2] = bswap32(b1); r.w[1] = bswap32(b2); r.w[0] = bswap32(b3);
    return r;
}

}  // namespace p2pt
