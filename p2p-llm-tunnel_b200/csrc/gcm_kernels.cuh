// gcm_kernels.cuh -- DTLS 1.2 AES-128-GCM record seal / open for sm_100a, plus the fused
// frame-encode->seal and open->frame-decode passes.
//
// Reference behaviour restated: webrtc-dtls 0.10.0 CryptoGcm::encrypt / decrypt and
// generate_aead_additional_data (UN-VENDORED, /root/reference/tunnel/Cargo.lock:2822-2824; reached
// from /root/reference/tunnel/src/serve.rs:274 and tunnel/src/rtc.rs:88-92), over aes-gcm 0.10.3
// (Cargo.lock:27-29).  Layout: SURVEY.md App. A.2.
//
// Design (integer / LSU pipes, no tensor cores -- nothing here is a contraction):
//   * G lanes (4/8/16/32) cooperate on one record.  Slot u = 0 is the J0 block (tag mask) paired
//     with the AAD GHASH block; slots 1..nc are CTR blocks paired with their ciphertext GHASH
//     block; the last slot is the GHASH length block.  Lane j takes slots j, j+G, ...
//   * AES: four T-tables in shared memory, each entry replicated 32x so lane l only ever touches
//     bank l (zero bank conflicts for data-dependent indices).  The byte extract, the <<8 and the
//     lane offset are ONE PRMT: idx = prmt(state, lanebase, 0x76k4).  Round keys live in registers.
//   * GHASH: per lane Horner with H^G (integer-multiply carry-less product, gcm_math.cuh), one final
//     multiply by H^(slots left), XOR-reduced over the G lanes with shuffles.
//   * I/O: aligned 16-byte vectors; the output stream is re-aligned with a byte funnel against the
//     neighbouring lane's block (shuffle), so interior stores are full STG.128 at any byte offset.
#pragma once
#include "common.cuh"
#include "gcm_math.cuh"
#include "../../include/p2pt.h"

namespace p2pt {

// One direction of one session.  704 bytes, 16-byte aligned.
struct SessionKey {
    uint32_t rk[44];        // FIPS-197 key schedule, little-endian words
    uint32_t iv;            // 4-byte implicit nonce (write IV), little-endian word
    uint32_t pad[3];
    uint32_t hpow[32][4];   // H^1 .. H^32, each times alpha^-1 (gf128_premul: the form gf128_mul_pre takes), U128 word order
};
static_assert(sizeof(SessionKey) == 704, "SessionKey layout");

// Te0[x] = {2S, S, S, 3S} as a little-endian word; filled by the host at ctx creation.
__device__ uint32_t g_te0[256];

__device__ __forceinline__ uint32_t rotl32(uint32_t v, int s) { return __funnelshift_l(v, v, s); }

// ---- AES-128 with a lookup policy -------------------------------------------------------------
struct GlobalLut {   // setup kernels only (key expansion): plain global table
    template <int K, int T> __device__ __forceinline__ uint32_t look(uint32_t w) const {
        uint32_t v = g_te0[(w >> (8 * K)) & 0xffu];
        return T == 0 ? v : rotl32(v, 8 * T);
    }
};

// Measured on B200 (round 1): moving the byte-0 / byte-3 index arithmetic to the FMA pipe
// (IMAD / IMAD.WIDE with run-time multipliers) took the ALU pipe from 630 to 550 instructions per
// block but made the kernel 9 % SLOWER (3.89 -> 4.25 ms per 1 M records): the extra instructions
// cost more issue slots than the ALU relief buys.  Moving only the byte-3 extraction (value 2: one PRMT ->
// IMAD.WIDE + IMAD, a quarter of the look-ups) is still 3 % slower (3.74 -> 3.85 ms).  Kept behind a switch, off.
#ifndef P2PT_FMA_EXTRACT
#define P2PT_FMA_EXTRACT 0
#endif
constexpr uint32_t kLutBytes = 131072;   // 4 tables x 256 entries x 32 replicas x 4 B

struct SmemLut {     // hot path: bank-replicated shared-memory tables
    uint32_t lb;     // 64 KiB-aligned shared address of the tables + 4 * lane
    uint32_t m8, m16, m24;   // 1<<8, 1<<16, 1<<24 as RUN-TIME values (kernel arguments): multiplying by
                             // them keeps the byte-0 / byte-3 index arithmetic on the FMA pipe (IMAD,
                             // IMAD.HI); as literals ptxas would turn them back into ALU-pipe shifts
    template <int K, int T> __device__ __forceinline__ uint32_t look(uint32_t w) const {
        // address = base | byte_K(w) << 8 | lane << 2
        uint32_t idx;
#if P2PT_FMA_EXTRACT
        if (K == 0 && P2PT_FMA_EXTRACT == 1) {            // ((w << 24) * 2^16 + (lb << 32)) >> 32 = (byte0 << 8) + lb, via IMAD + IMAD.WIDE
            const uint32_t t = w * m24;
            const unsigned long long r = (unsigned long long)t * m16 + ((unsigned long long)lb << 32);
            idx = (uint32_t)(r >> 32);
        } else if (K == 3) {     // (w * 2^8) >> 32 = byte3, via IMAD.WIDE + IMAD
            const unsigned long long r = (unsigned long long)w * m8;
            idx = (uint32_t)(r >> 32) * m8 + lb;
        } else
#endif
        {
            idx = __byte_perm(w, lb, 0x7604 | (K << 4));     // a single PRMT (ALU pipe)
        }
        uint32_t v;
        if (T == 0)      asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(idx));
        else if (T == 1) asm("ld.shared.u32 %0, [%1+128];" : "=r"(v) : "r"(idx));
        else if (T == 2) asm("ld.shared.u32 %0, [%1+65536];" : "=r"(v) : "r"(idx));
        else             asm("ld.shared.u32 %0, [%1+65664];" : "=r"(v) : "r"(idx));
        return v;
    }
};

// Round keys either in registers (set-up kernel) or in the group's shared-memory slot (hot path:
// one LDS.128 per round, every lane of the group reads the same address -> broadcast).
struct RegKeys {
    const uint32_t (&rk)[44];
    __device__ __forceinline__ uint4 get(int r) const { return make_uint4(rk[4 * r], rk[4 * r + 1], rk[4 * r + 2], rk[4 * r + 3]); }
};
struct SmemKeys {
    uint32_t addr;     // shared address of 11 x 16 bytes
    __device__ __forceinline__ uint4 get(int r) const {
        uint4 v;
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr + 16u * (uint32_t)r));
        return v;
    }
};

// rounds FIRST..9 and the final round; the state enters as the input of round FIRST
template <int FIRST, class Lut, class Keys>
__device__ __forceinline__ void aes128_rounds(const Lut& lut, const Keys& keys,
                                              uint32_t& s0, uint32_t& s1, uint32_t& s2, uint32_t& s3) {
#pragma unroll
    for (int r = FIRST; r < 10; r++) {
        const uint4 k = keys.get(r);
        uint32_t t0 = lut.template look<0, 0>(s0) ^ lut.template look<1, 1>(s1) ^ lut.template look<2, 2>(s2) ^ lut.template look<3, 3>(s3) ^ k.x;
        uint32_t t1 = lut.template look<0, 0>(s1) ^ lut.template look<1, 1>(s2) ^ lut.template look<2, 2>(s3) ^ lut.template look<3, 3>(s0) ^ k.y;
        uint32_t t2 = lut.template look<0, 0>(s2) ^ lut.template look<1, 1>(s3) ^ lut.template look<2, 2>(s0) ^ lut.template look<3, 3>(s1) ^ k.z;
        uint32_t t3 = lut.template look<0, 0>(s3) ^ lut.template look<1, 1>(s0) ^ lut.template look<2, 2>(s1) ^ lut.template look<3, 3>(s2) ^ k.w;
        s0 = t0; s1 = t1; s2 = t2; s3 = t3;
    }
    // last round: SubBytes + ShiftRows only.  S[x] sits in byte 0 of Te2, byte 1 of Te3, byte 2 of
    // Te0 and byte 3 of Te1, i.e. already in the byte lane it is needed in.
    const uint4 k = keys.get(10);
#define P2PT_LAST(a, b, c, d, kw)                                                            \
    (__byte_perm(__byte_perm(lut.template look<0, 2>(a), lut.template look<1, 3>(b), 0x0050), \
                 __byte_perm(lut.template look<2, 0>(c), lut.template look<3, 1>(d), 0x7200), 0x7610) ^ kw)
    uint32_t o0 = P2PT_LAST(s0, s1, s2, s3, k.x);
    uint32_t o1 = P2PT_LAST(s1, s2, s3, s0, k.y);
    uint32_t o2 = P2PT_LAST(s2, s3, s0, s1, k.z);
    uint32_t o3 = P2PT_LAST(s3, s0, s1, s2, k.w);
#undef P2PT_LAST
    s0 = o0; s1 = o1; s2 = o2; s3 = o3;
}

template <class Lut, class Keys>
__device__ __forceinline__ void aes128_encrypt(const Lut& lut, const Keys& keys,
                                               uint32_t& s0, uint32_t& s1, uint32_t& s2, uint32_t& s3) {
    const uint4 k = keys.get(0);
    s0 ^= k.x; s1 ^= k.y; s2 ^= k.z; s3 ^= k.w;
    aes128_rounds<1>(lut, keys, s0, s1, s2, s3);
}

// CTR blocks of one record differ only in the 32-bit counter, and for counters below 2^16 only in
// two state bytes after the first AddRoundKey: of the 16 table lookups of round 1, 14 are the same
// for every block of the record.  They are folded into four words once per record.
struct CtrRound1 {
    uint32_t c0, c1, c2, c3;     // round-1 outputs without the two counter-dependent lookups
    uint32_t r3;                 // round key 0, word 3 (the counter word is XORed into it)
    template <class Lut, class Keys>
    __device__ __forceinline__ void prepare(const Lut& lut, const Keys& keys, uint32_t iv, uint32_t n0, uint32_t n1) {
        const uint4 k0 = keys.get(0), k1 = keys.get(1);
        const uint32_t s0 = iv ^ k0.x, s1 = n0 ^ k0.y, s2 = n1 ^ k0.z, s3 = k0.w;   // counter bytes 0, 1 are zero
        r3 = k0.w;
        c0 = lut.template look<0, 0>(s0) ^ lut.template look<1, 1>(s1) ^ lut.template look<2, 2>(s2) ^ k1.x;
        c1 = lut.template look<0, 0>(s1) ^ lut.template look<1, 1>(s2) ^ lut.template look<3, 3>(s0) ^ k1.y;
        c2 = lut.template look<0, 0>(s2) ^ lut.template look<1, 1>(s3) ^ lut.template look<2, 2>(s0) ^ lut.template look<3, 3>(s1) ^ k1.z;
        c3 = lut.template look<0, 0>(s3) ^ lut.template look<1, 1>(s0) ^ lut.template look<2, 2>(s1) ^ lut.template look<3, 3>(s2) ^ k1.w;
    }
    // the same with the four words and round key 0 word 3 read back from the group's slot (P2PT_GCM_PARK)
    template <class Lut, class Keys>
    static __device__ __forceinline__ void block_parked(const Lut& lut, const Keys& keys, uint32_t slot, uint32_t r1_off, uint32_t ctr,
                                                        uint32_t& o0, uint32_t& o1, uint32_t& o2, uint32_t& o3) {
        uint32_t r3w, q0, q1, q2, q3;
        asm volatile("ld.shared.u32 %0, [%1+12];" : "=r"(r3w) : "r"(slot));
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(q0), "=r"(q1), "=r"(q2), "=r"(q3) : "r"(slot + r1_off));
        const uint32_t s3 = bswap32(ctr) ^ r3w;
        o0 = q0 ^ lut.template look<3, 3>(s3);
        o1 = q1 ^ lut.template look<2, 2>(s3);
        o2 = q2; o3 = q3;
        aes128_rounds<2>(lut, keys, o0, o1, o2, o3);
    }
    // Records of at most 254 blocks (4064 bytes: every DTLS record that fits a path MTU) only ever count in the LAST
    // counter byte, so after round 1 a single state word depends on the counter and of the 16 look-ups of round 2 twelve
    // are the same for every block of the record: they, round key 2 and the constant look-up of round 1 are folded into
    // four words once per record (kept where round key 1 lay, which nobody reads after prepare()).  A block then costs
    // 1 + 4 look-ups in rounds 1 and 2 instead of 2 + 16.
    template <class Lut, class Keys>
    __device__ __forceinline__ void fold_round2(const Lut& lut, const Keys& keys, uint32_t& q0, uint32_t& q1, uint32_t& q2, uint32_t& q3) const {
        const uint4 k2 = keys.get(2);
        const uint32_t v1 = c1 ^ lut.template look<2, 2>(r3), v2 = c2, v3 = c3;      // counter bytes 0..2 are zero
        q0 = lut.template look<1, 1>(v1) ^ lut.template look<2, 2>(v2) ^ lut.template look<3, 3>(v3) ^ k2.x;
        q1 = lut.template look<0, 0>(v1) ^ lut.template look<1, 1>(v2) ^ lut.template look<2, 2>(v3) ^ k2.y;
        q2 = lut.template look<0, 0>(v2) ^ lut.template look<1, 1>(v3) ^ lut.template look<3, 3>(v1) ^ k2.z;
        q3 = lut.template look<0, 0>(v3) ^ lut.template look<2, 2>(v1) ^ lut.template look<3, 3>(v2) ^ k2.w;
    }
    template <class Lut, class Keys>
    static __device__ __forceinline__ void block_small(const Lut& lut, const Keys& keys, uint32_t slot, uint32_t r1_off, uint32_t ctr,
                                                       uint32_t& o0, uint32_t& o1, uint32_t& o2, uint32_t& o3) {
        uint32_t r3w, c0w, q0, q1, q2, q3;
        asm volatile("ld.shared.u32 %0, [%1+12];" : "=r"(r3w) : "r"(slot));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(c0w) : "r"(slot + r1_off));
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4+16];" : "=r"(q0), "=r"(q1), "=r"(q2), "=r"(q3) : "r"(slot));
        const uint32_t s3 = (ctr << 24) ^ r3w;                                        // bswap32(ctr) for ctr < 256
        const uint32_t v0 = c0w ^ lut.template look<3, 3>(s3);
        o0 = q0 ^ lut.template look<0, 0>(v0);
        o1 = q1 ^ lut.template look<3, 3>(v0);
        o2 = q2 ^ lut.template look<2, 2>(v0);
        o3 = q3 ^ lut.template look<1, 1>(v0);
        aes128_rounds<3>(lut, keys, o0, o1, o2, o3);
    }
    // keystream block for counter ctr < 65536
    template <class Lut, class Keys>
    __device__ __forceinline__ void block(const Lut& lut, const Keys& keys, uint32_t ctr,
                                          uint32_t& o0, uint32_t& o1, uint32_t& o2, uint32_t& o3) const {
        const uint32_t s3 = bswap32(ctr) ^ r3;
        o0 = c0 ^ lut.template look<3, 3>(s3);
        o1 = c1 ^ lut.template look<2, 2>(s3);
        o2 = c2; o3 = c3;
        aes128_rounds<2>(lut, keys, o0, o1, o2, o3);
    }
};

// The keystream block of a record longer than 4064 bytes (its counters need two bytes), out of line: such records are
// rare, and inlined beside block_small the second path cost the hot loop registers (spills at the 64-register cap).
__device__ __noinline__ uint4 ctr_block_wide(uint32_t lut_lb, uint32_t slot, uint32_t r1_off, uint32_t ctr);

// GF(2^128) multiply with the prepared multiplier in shared memory (9 rows of 4 class masks):
// same arithmetic as gf128_mul in gcm_math.cuh, one LDS.128 per 32x32 product instead of 36 registers.
// A prepared multiplier in the group's slot: nine rows of classes 0..3 (16 bytes each) and, with hole spacing 6,
// nine rows of classes 4..5 (8 bytes each) behind them: one LDS.128 (+ one LDS.64) per 32x32 product.
constexpr uint32_t kGkeyHiOff = 9 * 16;
constexpr uint32_t kGkeyBytes = kGhashClasses == 4 ? 9 * 16 : 9 * 16 + 80;
__device__ __forceinline__ uint64_t clmul32_s(uint32_t a, uint32_t key_addr, int row) {
    uint32_t b[8];
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(b[0]), "=r"(b[1]), "=r"(b[2]), "=r"(b[3]) : "r"(key_addr + 16u * (uint32_t)row));
    if (kGhashClasses == 6)
        asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(b[4]), "=r"(b[5]) : "r"(key_addr + kGkeyHiOff + 8u * (uint32_t)row));
    return clmul32(a, b);
}
__device__ __forceinline__ void clmul64_s(uint32_t lo, uint32_t hi, uint32_t mid, uint32_t key_addr, int r,
                                          uint64_t& out_lo, uint64_t& out_hi) {
    uint64_t p = clmul32_s(lo, key_addr, r);
    uint64_t q = clmul32_s(hi, key_addr, r + 1);
    uint64_t m = clmul32_s(mid, key_addr, r + 2) ^ p ^ q;
    out_lo = p ^ (m << 32);
    out_hi = q ^ (m >> 32);
}
// the multiplier rows in the slot come from a key power that was pre-multiplied by alpha^-1 (gf128_premul): no shift
__device__ __forceinline__ U128 gf128_reduce(uint64_t pl, uint64_t ph, uint64_t ql, uint64_t qh, uint64_t rl, uint64_t rh) {
    rl ^= pl ^ ql;
    rh ^= ph ^ qh;
    return gf128_fold<true>(pl, ph ^ rl, ql ^ rh, qh);
}
__device__ __forceinline__ U128 gf128_mul_s(const U128& x, uint32_t key_addr) {
    const uint32_t x0 = x.w[0], x1 = x.w[1], x2 = x.w[2], x3 = x.w[3];
    uint64_t pl, ph, ql, qh, rl, rh;
    clmul64_s(x0, x1, x0 ^ x1, key_addr, 0, pl, ph);
    clmul64_s(x2, x3, x2 ^ x3, key_addr, 3, ql, qh);
    clmul64_s(x0 ^ x2, x1 ^ x3, x0 ^ x1 ^ x2 ^ x3, key_addr, 6, rl, rh);
    return gf128_reduce(pl, ph, ql, qh, rl, rh);
}

__device__ __noinline__ uint4 ctr_block_wide(uint32_t lut_lb, uint32_t slot, uint32_t r1_off, uint32_t ctr) {
    SmemLut lut; lut.lb = lut_lb; lut.m8 = 1u << 8; lut.m16 = 1u << 16; lut.m24 = 1u << 24;
    SmemKeys keys; keys.addr = slot;
    uint4 o;
    CtrRound1::block_parked(lut, keys, slot, r1_off, ctr, o.x, o.y, o.z, o.w);
    return o;
}

// ---- session set-up: key schedule, H, H^1..H^32 (one thread per key) ------------------------------
__global__ void __launch_bounds__(128) session_expand_kernel(uint32_t n, const uint8_t* __restrict__ keys16,
                                                             const uint8_t* __restrict__ ivs4,
                                                             SessionKey* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t rk[44];
#pragma unroll
    for (int w = 0; w < 4; w++) {
        const uint8_t* k = keys16 + 16 * (u64)i + 4 * w;
        rk[w] = (uint32_t)k[0] | ((uint32_t)k[1] << 8) | ((uint32_t)k[2] << 16) | ((uint32_t)k[3] << 24);
    }
    uint32_t rcon = 1;
#pragma unroll
    for (int w = 4; w < 44; w++) {
        uint32_t t = rk[w - 1];
        if ((w & 3) == 0) {
            t = (t >> 8) | (t << 24);                       // RotWord on a little-endian word
            t = ((g_te0[t & 0xff] >> 8) & 0xffu) | (((g_te0[(t >> 8) & 0xff] >> 8) & 0xffu) << 8) |
                (((g_te0[(t >> 16) & 0xff] >> 8) & 0xffu) << 16) | (((g_te0[t >> 24] >> 8) & 0xffu) << 24);
            t ^= rcon;
            rcon = (rcon << 1) ^ ((rcon & 0x80) ? 0x11b : 0);
        }
        rk[w] = rk[w - 4] ^ t;
    }
    SessionKey* sk = out + i;
#pragma unroll
    for (int w = 0; w < 44; w++) sk->rk[w] = rk[w];
    const uint8_t* iv = ivs4 + 4 * (u64)i;
    sk->iv = (uint32_t)iv[0] | ((uint32_t)iv[1] << 8) | ((uint32_t)iv[2] << 16) | ((uint32_t)iv[3] << 24);
    sk->pad[0] = sk->pad[1] = sk->pad[2] = 0;
    // H = AES_K(0^128)
    uint32_t h0 = 0, h1 = 0, h2 = 0, h3 = 0;
    GlobalLut lut;
    RegKeys rkeys{rk};
    aes128_encrypt(lut, rkeys, h0, h1, h2, h3);
    const U128 H = block_to_u128(h0, h1, h2, h3);
    GhashKey hk;
    ghash_key_prepare(H, hk);
    U128 p = H;
    for (int e = 0; e < 32; e++) {
        const U128 pm = gf128_premul(p);
        sk->hpow[e][0] = pm.w[0]; sk->hpow[e][1] = pm.w[1]; sk->hpow[e][2] = pm.w[2]; sk->hpow[e][3] = pm.w[3];
        p = gf128_mul(p, hk);
    }
}

// ---- the record kernel -----------------------------------------------------------------------------
enum GcmMode { kSeal = 0, kOpen = 1, kPipeSeal = 2, kPipeOpen = 3 };

struct GcmArgs {
    u64 n;
    const uint32_t* session;
    const SessionKey* keys;
    uint32_t n_sessions;
    const uint8_t* nonce8;        // seal modes: n x 8 explicit nonces (4-byte aligned)
    const uint8_t* in;            // seal: raw records; open: sealed records; pipe-seal: payload bytes
    const u64* in_off;
    uint8_t* out;
    const u64* out_off;
    uint8_t* status;              // may be null in pipe-open (status lands in desc)
    uint32_t* out_len;            // open only
    const uint16_t* epoch;        // pipe-seal
    const u64* seq48;             // pipe-seal
    const uint8_t* types;         // pipe-seal
    const uint32_t* sids;         // pipe-seal
    p2pt_desc* desc;              // pipe-open
    uint32_t m8, m16, m24;        // 1<<8, 1<<16, 1<<24 (see SmemLut)
};

#ifndef P2PT_GCM_THREADS
#define P2PT_GCM_THREADS (P2PT_GHASH_SPACING == 6 ? 768 : 1024)     // the wider multiplier rows of spacing 6 need the room
#endif
#ifndef P2PT_GCM_UNROLL
#define P2PT_GCM_UNROLL 1
#endif
// P2PT_GCM_PIPELINE 1: the keystream block of round t + 1 is computed in round t, ahead of the GHASH multiply
// of round t, so that the two chains (table look-ups + PRMT on one side, IMAD.WIDE + LOP3 on the other) sit
// in one basic block and can be interleaved; costs four live registers across the loop.
#ifndef P2PT_GCM_PIPELINE
#define P2PT_GCM_PIPELINE 0
#endif
// P2PT_GCM_INTERIOR 1: rounds in which every lane of the group holds a complete interior block (five readable
// source words, a complete successor block, a multiply to follow) are recognised by ONE group-uniform test on the
// round index and skip the per-slot edge bookkeeping (fast-path range tests, active / last-slot tracking, tail
// stores); the AES and GHASH code itself exists once.
#ifndef P2PT_GCM_INTERIOR
#define P2PT_GCM_INTERIOR 0
#endif
// P2PT_GCM_PARK 1: values that live for the whole record but are used once per block or once per record are
// parked in the group's shared-memory slot instead of registers - the four round-1 words of the counter block
// (one broadcast LDS.128 per block), round key 0 word 3 (already in the slot) and E_K(J0) (read once for the
// tag) - and the three body pointers of the edge path are rebuilt there.  At the 64-register cap the compiler
// otherwise spills and re-derives loop invariants inside the loop (64-bit range tests, the shared window
// address from S2R).
#ifndef P2PT_GCM_PARK
#define P2PT_GCM_PARK 1
#endif
constexpr int kGcmThreads = P2PT_GCM_THREADS;
constexpr int kGcmUnroll = P2PT_GCM_UNROLL;
#if P2PT_GCM_PARK
// 11 x 16 B round keys, the GHASH multiplier rows, E_K(J0), the four round-1 words; an odd multiple of 16 bytes keeps
// neighbouring groups' rows in different banks (368 = 23 x 16 with hole spacing 4, 432 = 27 x 16 with spacing 6)
constexpr uint32_t kSlotEj0 = 176 + kGkeyBytes, kSlotR1 = kSlotEj0 + 16;
constexpr uint32_t kGroupSlotBytes = ((kSlotR1 + 16) / 16) % 2 ? kSlotR1 + 16 : kSlotR1 + 32;
static_assert((kGroupSlotBytes / 16) % 2 == 1, "slot stride must be an odd multiple of 16 bytes");
#else
constexpr uint32_t kGroupSlotBytes = 336;   // 11 x 16 B round keys + 9 x 16 B GHASH multiplier rows, padded
                                            // so that neighbouring groups' rows fall in different banks
#endif

// Shared-memory map of the record kernel.  The T-tables must start on a 64 KiB boundary of the
// shared window (PRMT addressing) and the window itself starts a little above 0 (the first KiB is
// reserved by the system), so there is a gap of `slack` bytes below the tables: the per-group key
// slots fill that gap first and continue after the tables.
__host__ __device__ constexpr uint32_t gcm_smem_bytes(int G, uint32_t slack) {
    const uint32_t groups = kGcmThreads / G, low = slack / kGroupSlotBytes;
    return slack + kLutBytes + (groups > low ? (groups - low) * kGroupSlotBytes : 0u);
}

__device__ __forceinline__ uint32_t shared_addr(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" :: "r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// bytes beyond `nb` of a 16-byte block (four LE words) cleared
__device__ __forceinline__ void mask_block(uint4& y, uint32_t nb) {
    if (nb >= 16) return;
    const uint32_t m0 = nb >= 4 ? 0xffffffffu : (nb ? (0xffffffffu >> (32 - 8 * nb)) : 0u);
    const uint32_t m1 = nb >= 8 ? 0xffffffffu : (nb > 4 ? (0xffffffffu >> (32 - 8 * (nb - 4))) : 0u);
    const uint32_t m2 = nb >= 12 ? 0xffffffffu : (nb > 8 ? (0xffffffffu >> (32 - 8 * (nb - 8))) : 0u);
    const uint32_t m3 = nb > 12 ? (0xffffffffu >> (32 - 8 * (nb - 12))) : 0u;
    y.x &= m0; y.y &= m1; y.z &= m2; y.w &= m3;
}

template <int G, int MODE>
__global__ void __launch_bounds__(kGcmThreads, 1) gcm_record_kernel(const GcmArgs a) {
    extern __shared__ __align__(16) uint8_t dyn_smem[];
    constexpr bool SEAL = (MODE == kSeal || MODE == kPipeSeal);
    constexpr uint32_t GMASK_BASE = (G == 32) ? 0xffffffffu : ((1u << G) - 1u);
    constexpr uint32_t PREFIX = SEAL ? 21u : 13u;

    const int lane = threadIdx.x & 31;
    const int j = lane & (G - 1);                        // lane inside the record group
    const uint32_t gmask = GMASK_BASE << (lane & ~(G - 1));

    // ---- fill the bank-replicated T-tables (once per persistent CTA) ----
    const uint32_t sbase = shared_addr(dyn_smem);
    const uint32_t tbase = (sbase + 0xffffu) & ~0xffffu;          // 64 KiB aligned (PRMT addressing)
    {
        uint32_t* tabs = reinterpret_cast<uint32_t*>(dyn_smem + (tbase - sbase));
        for (int idx = threadIdx.x; idx < 256 * 32; idx += blockDim.x) {
            const int x = idx >> 5, l = idx & 31;
            const uint32_t t0 = g_te0[x];
            tabs[x * 64 + l] = t0;
            tabs[x * 64 + 32 + l] = rotl32(t0, 8);
            tabs[16384 + x * 64 + l] = rotl32(t0, 16);
            tabs[16384 + x * 64 + 32 + l] = rotl32(t0, 24);
        }
    }
    __syncthreads();
    SmemLut lut;
    lut.lb = tbase + 4u * (uint32_t)lane;
    lut.m8 = a.m8; lut.m16 = a.m16; lut.m24 = a.m24;
    // this group's key slot: round keys at +0, GHASH multiplier rows at +176
    const uint32_t n_low = (tbase - sbase) / kGroupSlotBytes, gi = (uint32_t)(threadIdx.x / G);
    const uint32_t slot = gi < n_low ? sbase + gi * kGroupSlotBytes : tbase + kLutBytes + (gi - n_low) * kGroupSlotBytes;
    SmemKeys rkeys; rkeys.addr = slot;
    const uint32_t gkey = slot + 176;

    const u64 groups_per_cta = kGcmThreads / G;
    const u64 group0 = (u64)blockIdx.x * groups_per_cta + (threadIdx.x / G);
    const u64 group_stride = (u64)gridDim.x * groups_per_cta;

    for (u64 i = group0; i < a.n; i += group_stride) {
        // ------------------------------------------------------------------ record geometry
        const u64 ioff = __ldg(a.in_off + i);
        const u64 ilen = __ldg(a.in_off + i + 1) - ioff;
        const uint8_t* ibase = a.in + ioff;
        const u64 ooff = __ldg(a.out_off + i);
        uint8_t* obase = a.out + ooff;
        const u64 oslot = __ldg(a.out_off + i + 1) - ooff;      // bytes of the slot beyond the item are cleared below:
                                                                // nothing stale from an earlier batch may reach the caller
        const uint32_t sidx = __ldg(a.session + i);
        const SessionKey* sk = a.keys + (sidx < a.n_sessions ? sidx : 0);

        uint32_t L;                     // plaintext / ciphertext length
        uint32_t st = P2PT_ST_OK;
        uint32_t hdr_type = 23, hdr_v0 = 0xfe, hdr_v1 = 0xfd;
        // input body bytes live in [body_lo, body_hi); body0 = address of input body byte 0 (may precede body_lo:
        // pipe-seal).  All three follow from ibase, ilen and L (P2PT_BODY_* below), so the edge path can rebuild them.
        const uint8_t* body_lo;
        const uint8_t* body_hi;
        const uint8_t* body0;
        if (MODE == kSeal) {
            if (ilen < 13) st = P2PT_ST_BAD_HEADER;
            L = (uint32_t)(ilen - 13);
            body0 = body_lo = ibase + 13; body_hi = ibase + ilen;
        } else if (MODE == kPipeSeal) {
            L = (uint32_t)ilen + 5;
            body_lo = ibase; body_hi = ibase + ilen; body0 = ibase - 5;
        } else {
            if (ilen < 13) st = P2PT_ST_BAD_HEADER;
            else {
                hdr_type = __ldg(ibase); hdr_v0 = __ldg(ibase + 1); hdr_v1 = __ldg(ibase + 2);
                if (!(hdr_v0 == 0xfe && (hdr_v1 == 0xfd || hdr_v1 == 0xff))) st = P2PT_ST_BAD_HEADER;
                else if (hdr_type == 20) st = P2PT_ST_PASS_THROUGH;
                else if (ilen <= 21) st = P2PT_ST_NO_ROOM_FOR_NONCE;
                else if (ilen < 37) st = P2PT_ST_AUTH_FAIL;
            }
            L = (uint32_t)(ilen - 37);
            body0 = body_lo = ibase + 21; body_hi = ibase + 21 + L;
        }
        if (sidx >= a.n_sessions && st == P2PT_ST_OK) st = P2PT_ST_BAD_HEADER;
        if (ilen > 1000000ull && st == P2PT_ST_OK) st = P2PT_ST_BAD_HEADER;       // far beyond any DTLS record (<= 2^16); keeps every
                                                                                  // CTR counter below 2^16 (CtrRound1) and block math in 32 bits

        if (st != P2PT_ST_OK) {         // group-uniform early outs
            if (MODE == kOpen || MODE == kPipeOpen) {
                uint32_t olen = 0;
                if (st == P2PT_ST_PASS_THROUGH) {           // ChangeCipherSpec: copy unmodified
                    for (u64 b = j; b < ilen; b += G) obase[b] = __ldg(ibase + b);
                    olen = (uint32_t)ilen;
                }
                for (u64 b = olen + j; b < oslot; b += G) obase[b] = 0;
                if (j == 0) {
                    if (MODE == kOpen) a.out_len[i] = olen;
                    if (MODE == kPipeOpen) {
                        p2pt_desc d; d.pay_off = __ldg(a.out_off + i); d.pay_len = 0; d.stream_id = 0;
                        d.msg_type = 0; d.status = (uint8_t)st;
                        for (int q = 0; q < 6; q++) d.pad[q] = 0;
                        a.desc[i] = d;
                    }
                }
            }
            else {
                for (u64 b = j; b < oslot; b += G) obase[b] = 0;
            }
            if (j == 0 && a.status) a.status[i] = (uint8_t)st;
            continue;
        }

        const uint32_t nc = (L + 15) >> 4;          // CTR / ciphertext blocks
        const uint32_t nfull = L >> 4;              // of which complete
        const uint32_t slots = nc + 2;              // + J0/AAD slot + length slot
        uint8_t* ob0 = obase + PREFIX;              // output body byte 0

        // Input side: blocks whose five covering aligned words lie inside the body are loaded as
        // 32-bit words + funnel shift; the others (record edges) go through the bounded byte path.
        const unsigned ish = (unsigned)((uintptr_t)body0 & 3);
        const uint8_t* ib4 = body0 - ish;           // word-aligned address of block 0's first covering word
        int c_fast_lo = 0, c_fast_hi = -1;          // inclusive range of blocks on the fast input path
        {
            // the covering words must stay inside this item's own input bytes [ibase, ibase + ilen):
            // ib4 + 16c >= ibase  and  ib4 + 16c + (ish ? 20 : 16) <= ibase + ilen
            const long long lo_gap = (long long)(ibase - ib4);
            c_fast_lo = lo_gap > 0 ? (int)((lo_gap + 15) >> 4) : 0;
            const long long room = (long long)((ibase + ilen) - ib4) - (ish ? 20 : 16);
            c_fast_hi = room >= 0 ? (int)(room >> 4) : -1;
            if (c_fast_hi > (int)nfull - 1) c_fast_hi = (int)nfull - 1;
        }
        // Output side: every complete block is stored as four aligned 32-bit words, each word
        // funnel-shifted against the previous block's last word (from the neighbouring lane).
        const unsigned osh = (unsigned)((uintptr_t)ob0 & 3);
        uint8_t* ob4 = ob0 - osh;

        // ------------------------------------------------------------------ per-record key material -> shared
        __syncwarp(gmask);                                   // previous record's readers are done
        {
            const uint4* rp = reinterpret_cast<const uint4*>(sk->rk);
            for (int q = j; q < 11; q += G) {
                const uint4 v = __ldg(rp + q);
                sts128(slot + 16u * q, v.x, v.y, v.z, v.w);
            }
            const uint4 hv = __ldg(reinterpret_cast<const uint4*>(sk->hpow[G - 1]));     // H^G
            uint32_t o[9];
            o[0] = hv.x; o[1] = hv.y; o[2] = hv.x ^ hv.y; o[3] = hv.z; o[4] = hv.w; o[5] = hv.z ^ hv.w;
            o[6] = hv.x ^ hv.z; o[7] = hv.y ^ hv.w; o[8] = o[6] ^ o[7];
#pragma unroll
            for (int q = 0; q < 9; q++) {
                if ((q % G) == j) {
                    uint32_t cl[8];
                    ghash_classes(o[q], cl);
                    sts128(gkey + 16u * q, cl[0], cl[1], cl[2], cl[3]);
                    if (kGhashClasses == 6)
                        asm volatile("st.shared.v2.u32 [%0], {%1,%2};" :: "r"(gkey + kGkeyHiOff + 8u * q), "r"(cl[4]), "r"(cl[5]) : "memory");
                }
            }
        }
        const uint32_t ivw = __ldg(&sk->iv);
        uint32_t n0, n1;                                     // explicit nonce, LE words
        if (SEAL) {
            const uint32_t* np = reinterpret_cast<const uint32_t*>(a.nonce8 + 8 * i);
            n0 = __ldg(np); n1 = __ldg(np + 1);
        } else {
            n0 = (uint32_t)__ldg(ibase + 13) | ((uint32_t)__ldg(ibase + 14) << 8) |
                 ((uint32_t)__ldg(ibase + 15) << 16) | ((uint32_t)__ldg(ibase + 16) << 24);
            n1 = (uint32_t)__ldg(ibase + 17) | ((uint32_t)__ldg(ibase + 18) << 8) |
                 ((uint32_t)__ldg(ibase + 19) << 16) | ((uint32_t)__ldg(ibase + 20) << 24);
        }
        __syncwarp(gmask);                                   // key slot visible to the whole group
        CtrRound1 r1;                                        // counters stay below 2^16: items are capped at 1 MB above
        r1.prepare(lut, rkeys, ivw, n0, n1);
#if P2PT_GCM_PARK
        const bool ctr_small = nc + 1 < 256;                 // every counter of the record (1 .. nc + 1) fits one byte
        __syncwarp(gmask);                                   // everyone has read round keys 0 and 1
        if (j == 0) {
            sts128(slot + kSlotR1, r1.c0, r1.c1, r1.c2, r1.c3);
            if (ctr_small) {
                uint32_t q0, q1, q2, q3;
                r1.fold_round2(lut, rkeys, q0, q1, q2, q3);
                sts128(slot + 16, q0, q1, q2, q3);           // over round key 1
            }
        }
        __syncwarp(gmask);
#define P2PT_KS(ctr, a0, a1, a2, a3)                                                                   \
    do {                                                                                               \
        if (ctr_small) CtrRound1::block_small(lut, rkeys, slot, kSlotR1, (ctr), a0, a1, a2, a3);      \
        else { const uint4 ks_ = ctr_block_wide(lut.lb, slot, kSlotR1, (ctr)); a0 = ks_.x; a1 = ks_.y; a2 = ks_.z; a3 = ks_.w; } \
    } while (0)
#else
#define P2PT_KS(ctr, a0, a1, a2, a3) r1.block(lut, rkeys, (ctr), a0, a1, a2, a3)
#endif

        U128 acc; acc.w[0] = acc.w[1] = acc.w[2] = acc.w[3] = 0;
#if !P2PT_GCM_PARK
        uint4 ej0 = make_uint4(0, 0, 0, 0);                  // E_K(J0), valid in group lane 0
#endif
        uint32_t carry3 = 0;                                 // last word of lane G-1's block, previous round
        uint32_t fh0 = 0, fh1 = 0;                           // pipe-open: first 8 plaintext bytes (lane 1)
        const uint32_t rounds = (slots + G - 1) / G;
#if P2PT_GCM_INTERIOR
        // rounds t_in_lo..t_in_hi: for every lane j, block c = tG + j - 1 lies in [max(c_fast_lo, 0), min(c_fast_hi,
        // nfull - 2)] and slot u + G exists
        int t_in_lo, t_in_hi;
        {
            const int lo = c_fast_lo > 0 ? c_fast_lo : 0;
            int hi = c_fast_hi < (int)nfull - 2 ? c_fast_hi : (int)nfull - 2;
            t_in_lo = (lo + 1 + G - 1) / G;
            const int a_ = hi + 2 - G, b_ = (int)slots - 2 * G;
            const int m_ = a_ < b_ ? a_ : b_;
            t_in_hi = m_ >= 0 ? m_ / G : -1;
        }
#endif
#if P2PT_GCM_PIPELINE
        uint32_t nk0, nk1, nk2, nk3;                         // keystream of the coming round
        P2PT_KS((uint32_t)(j + 1), nk0, nk1, nk2, nk3);
#endif
#pragma unroll kGcmUnroll
        for (uint32_t t = 0; t < rounds; t++) {
            const uint32_t u = t * G + j;
            const int c = (int)u - 1;                        // body block index; -1 = J0/AAD slot
            const bool active = u < slots;

            // ---- AES-CTR keystream for counter c + 2 (J0 = counter 1)
            uint32_t k0, k1, k2, k3;
#if P2PT_GCM_PIPELINE
            k0 = nk0; k1 = nk1; k2 = nk2; k3 = nk3;
#else
            P2PT_KS((uint32_t)(c + 2), k0, k1, k2, k3);
#endif

            uint4 oblk = make_uint4(0, 0, 0, 0);             // bytes this lane contributes to the output stream
            U128 gin; gin.w[0] = gin.w[1] = gin.w[2] = gin.w[3] = 0;
#if P2PT_GCM_INTERIOR
            const bool inner = ((int)t >= t_in_lo && (int)t <= t_in_hi);
            const bool fast_in = inner || (c >= c_fast_lo && c <= c_fast_hi);
            const bool full_blk = inner || (c >= 0 && (uint32_t)c < nfull);
#else
            const bool inner = false;
            const bool fast_in = (c >= c_fast_lo && c <= c_fast_hi);
            const bool full_blk = (c >= 0 && (uint32_t)c < nfull);
#endif
            int ovalid_lo = 0, ovalid_hi = 0;                // valid byte range of oblk (edge slots)

            if (fast_in) {
                // -------- interior block: five aligned words, one funnel shift each
                const uint32_t* wp = reinterpret_cast<const uint32_t*>(ib4 + 16 * (size_t)c);
                const uint32_t w0 = __ldg(wp), w1 = __ldg(wp + 1), w2 = __ldg(wp + 2), w3 = __ldg(wp + 3);
                const uint32_t w4 = ish ? __ldg(wp + 4) : 0u;
                const unsigned sh = ish * 8;
                uint4 x;
                x.x = __funnelshift_r(w0, w1, sh); x.y = __funnelshift_r(w1, w2, sh);
                x.z = __funnelshift_r(w2, w3, sh); x.w = __funnelshift_r(w3, w4, sh);
                const uint4 y = make_uint4(x.x ^ k0, x.y ^ k1, x.z ^ k2, x.w ^ k3);
                const uint4 ct = SEAL ? y : x;
                gin = block_to_u128(ct.x, ct.y, ct.z, ct.w);
                oblk = y; ovalid_hi = 16;
                if (MODE == kPipeOpen && c == 0) { fh0 = y.x; fh1 = y.y; }
            } else if (active) {
                const bool is_j0 = (u == 0);
                const bool is_len = (u == slots - 1);
                if (!is_j0 && !is_len) {
                    // -------- edge block: bounded byte loads, zero padding
                    const uint32_t pos = 16u * (uint32_t)c;
                    const uint32_t nb = (L - pos) < 16u ? (L - pos) : 16u;
#if P2PT_GCM_PARK
                    // rebuilt here (three adds on the rare path) rather than kept live across the loop
                    const uint8_t* e_b0 = MODE == kSeal ? ibase + 13 : MODE == kPipeSeal ? ibase - 5 : ibase + 21;
                    const uint8_t* e_lo = MODE == kPipeSeal ? ibase : e_b0;
                    const uint8_t* e_hi = (MODE == kSeal || MODE == kPipeSeal) ? ibase + ilen : ibase + 21 + L;
                    uint4 x = load16_bounded(e_b0 + pos, e_lo, e_hi);
#else
                    uint4 x = load16_bounded(body0 + pos, body_lo, body_hi);
#endif
                    if (MODE == kPipeSeal && c == 0) {       // synthesise [type][sid BE] in front of the payload
                        const uint32_t sid = __ldg(a.sids + i);
                        x.x = (uint32_t)__ldg(a.types + i) | ((sid >> 24) << 8) | (((sid >> 16) & 0xff) << 16) | (((sid >> 8) & 0xff) << 24);
                        x.y = (x.y & 0xffffff00u) | (sid & 0xff);
                    }
                    uint4 y = make_uint4(x.x ^ k0, x.y ^ k1, x.z ^ k2, x.w ^ k3);
                    mask_block(y, nb);
                    const uint4 ct = SEAL ? y : x;           // GHASH always runs over the ciphertext
                    gin = block_to_u128(ct.x, ct.y, ct.z, ct.w);
                    oblk = y; ovalid_hi = (int)nb;
                    if (MODE == kPipeOpen && c == 0) { fh0 = y.x; fh1 = y.y; }
                } else if (is_j0) {
#if P2PT_GCM_PARK
                    sts128(slot + kSlotEj0, k0, k1, k2, k3);      // E_K(J0): read back by every lane for the tag
#else
                    ej0 = make_uint4(k0, k1, k2, k3);
#endif
                    // AAD = epoch(2) seq(6) type ver(2) len(2); record header = type ver(2) epoch(2) seq(6) len(2)
                    uint32_t e_hi, s_w0, s_w1;               // header bytes 3..4, 5..8, 9..10 as BE numbers
                    if (MODE == kPipeSeal) {
                        const uint32_t ep = __ldg(a.epoch + i);
                        const u64 sq = __ldg(a.seq48 + i);
                        e_hi = ep & 0xffff; s_w0 = (uint32_t)(sq >> 16); s_w1 = (uint32_t)(sq & 0xffff);
                    } else {
                        if (MODE == kSeal) { hdr_type = __ldg(ibase); hdr_v0 = __ldg(ibase + 1); hdr_v1 = __ldg(ibase + 2); }
                        e_hi = ((uint32_t)__ldg(ibase + 3) << 8) | __ldg(ibase + 4);
                        s_w0 = ((uint32_t)__ldg(ibase + 5) << 24) | ((uint32_t)__ldg(ibase + 6) << 16) |
                               ((uint32_t)__ldg(ibase + 7) << 8) | __ldg(ibase + 8);
                        s_w1 = ((uint32_t)__ldg(ibase + 9) << 8) | __ldg(ibase + 10);
                    }
                    const uint32_t plen = L & 0xffff;        // `as u16` in the reference
                    gin.w[3] = (e_hi << 16) | (s_w0 >> 16);
                    gin.w[2] = (s_w0 << 16) | s_w1;
                    gin.w[1] = (hdr_type << 24) | (hdr_v0 << 16) | (hdr_v1 << 8) | (plen >> 8);
                    gin.w[0] = (plen & 0xff) << 24;
                    // the 16 output bytes that precede the body
                    if (SEAL) {
                        const uint32_t rlen = (8 + L + 16) & 0xffff;
                        obase[0] = (uint8_t)hdr_type; obase[1] = (uint8_t)hdr_v0; obase[2] = (uint8_t)hdr_v1;
                        obase[3] = (uint8_t)(e_hi >> 8); obase[4] = (uint8_t)e_hi;
                        oblk.x = bswap32(s_w0);
                        oblk.y = ((s_w1 >> 8) & 0xff) | ((s_w1 & 0xff) << 8) | ((rlen >> 8) << 16) | ((rlen & 0xff) << 24);
                        oblk.z = n0; oblk.w = n1;
                        ovalid_lo = 0; ovalid_hi = 16;
                    } else {
                        const uint32_t l_w = ((uint32_t)__ldg(ibase + 11) << 8) | __ldg(ibase + 12);
                        oblk.x = hdr_type << 24;
                        oblk.y = hdr_v0 | (hdr_v1 << 8) | ((e_hi >> 8) << 16) | ((e_hi & 0xff) << 24);
                        oblk.z = bswap32(s_w0);
                        oblk.w = ((s_w1 >> 8) & 0xff) | ((s_w1 & 0xff) << 8) | ((l_w >> 8) << 16) | ((l_w & 0xff) << 24);
                        ovalid_lo = 3; ovalid_hi = 16;
                    }
                } else {
                    gin.w[2] = 13 * 8;                       // len(AAD) in bits
                    gin.w[0] = L << 3;                       // len(C) in bits
                }
            }

#if P2PT_GCM_PIPELINE
            P2PT_KS((uint32_t)(c + 2 + G), nk0, nk1, nk2, nk3);      // next round's counter (unused after the last round)
#endif
            // ---- GHASH: Horner in H^G; the lane's last block is multiplied after the loop
            if (inner || active) {
                acc.w[0] ^= gin.w[0]; acc.w[1] ^= gin.w[1]; acc.w[2] ^= gin.w[2]; acc.w[3] ^= gin.w[3];
            }
            if (inner || u + G < slots) acc = gf128_mul_s(acc, gkey);

            // ---- output: the word that straddles into my block needs the previous block's last word
            uint32_t prev3 = __shfl_up_sync(gmask, oblk.w, 1, G);
            if (j == 0) prev3 = carry3;
            carry3 = __shfl_sync(gmask, oblk.w, G - 1, G);
            if (full_blk) {
                uint32_t* op = reinterpret_cast<uint32_t*>(ob4 + 16 * (size_t)c);
                const unsigned sh = osh * 8;
                op[0] = __funnelshift_l(prev3, oblk.x, sh);
                op[1] = __funnelshift_l(oblk.x, oblk.y, sh);
                op[2] = __funnelshift_l(oblk.y, oblk.z, sh);
                op[3] = __funnelshift_l(oblk.z, oblk.w, sh);
                // my last `osh` bytes ride in the next block's first word; if there is no complete
                // next block, store them here
                if (!inner && osh && (uint32_t)(c + 1) >= nfull) store_bytes(ob0 + 16 * (size_t)c, oblk, 16 - (int)osh, 16);
            } else if (active && ovalid_hi > ovalid_lo) {
                // edge slots (prefix block, partial last block): byte stores; the prefix's last `osh`
                // bytes are covered by block 0's first word when block 0 is complete
                int hi = ovalid_hi;
                if (c == -1 && nfull > 0 && osh) hi = 16 - (int)osh;
                store_bytes(ob0 + 16 * (ptrdiff_t)c, oblk, ovalid_lo, hi);
            }
        }

        // ------------------------------------------------------------------ finish GHASH
        {
            // the lane's last slot is the largest u = tG + j below `slots`; a lane without any slot holds acc = 0
            const uint32_t last_u = (uint32_t)j < slots ? ((slots - 1 - (uint32_t)j) / G) * G + (uint32_t)j : 0xffffffffu;
            const uint32_t fp = (last_u == 0xffffffffu) ? 1u : (slots - last_u);     // 1..G
            uint4 v = __ldg(reinterpret_cast<const uint4*>(sk->hpow[fp - 1]));
            U128 hf; hf.w[0] = v.x; hf.w[1] = v.y; hf.w[2] = v.z; hf.w[3] = v.w;
            GhashKey kF;
            ghash_key_prepare(hf, kF);
            acc = gf128_mul_pre(acc, kF);                        // lanes that never ran hold 0
        }
#pragma unroll
        for (int o = G >> 1; o > 0; o >>= 1) {
            acc.w[0] ^= __shfl_xor_sync(gmask, acc.w[0], o, G);
            acc.w[1] ^= __shfl_xor_sync(gmask, acc.w[1], o, G);
            acc.w[2] ^= __shfl_xor_sync(gmask, acc.w[2], o, G);
            acc.w[3] ^= __shfl_xor_sync(gmask, acc.w[3], o, G);
        }
        uint4 tag;                                               // little-endian words, memory order
#if P2PT_GCM_PARK
        {
            __syncwarp(gmask);
            uint4 e;                                             // lane 0 of the group stored it in round 0
            asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(e.x), "=r"(e.y), "=r"(e.z), "=r"(e.w) : "r"(slot + kSlotEj0) : "memory");
            tag.x = bswap32(acc.w[3]) ^ e.x; tag.y = bswap32(acc.w[2]) ^ e.y;
            tag.z = bswap32(acc.w[1]) ^ e.z; tag.w = bswap32(acc.w[0]) ^ e.w;
        }
#else
        tag.x = bswap32(acc.w[3]) ^ __shfl_sync(gmask, ej0.x, 0, G);
        tag.y = bswap32(acc.w[2]) ^ __shfl_sync(gmask, ej0.y, 0, G);
        tag.z = bswap32(acc.w[1]) ^ __shfl_sync(gmask, ej0.z, 0, G);
        tag.w = bswap32(acc.w[0]) ^ __shfl_sync(gmask, ej0.w, 0, G);
#endif

        if (SEAL) {
            uint8_t* tp = ob0 + L;
            for (int b = j; b < 16; b += G) tp[b] = (uint8_t)(word_of(tag, b >> 2) >> (8 * (b & 3)));
            for (u64 b = (u64)PREFIX + L + 16 + j; b < oslot; b += G) obase[b] = 0;
            if (j == 0 && a.status) a.status[i] = P2PT_ST_OK;
        } else {
            // constant-time style compare: OR of all byte differences, no early exit
            const uint8_t* tp = ibase + 21 + L;
            uint32_t diff = 0;
            for (int b = j; b < 16; b += G)
                diff |= ((word_of(tag, b >> 2) >> (8 * (b & 3))) & 0xffu) ^ (uint32_t)__ldg(tp + b);
#pragma unroll
            for (int o = G >> 1; o > 0; o >>= 1) diff |= __shfl_xor_sync(gmask, diff, o, G);
            const bool ok = (diff == 0);
            for (u64 b = (ok ? 13 + (u64)L : 0) + j; b < oslot; b += G) obase[b] = 0;     // a record that fails is cleared whole
            if (MODE == kOpen) {
                if (j == 0) { a.out_len[i] = ok ? 13 + L : 0; a.status[i] = ok ? P2PT_ST_OK : P2PT_ST_AUTH_FAIL; }
            } else {
                // frame decode on the opened plaintext (protocol.rs:157-172), done by the lane that
                // decrypted body block 0
                const uint32_t f0 = __shfl_sync(gmask, fh0, 1 & (G - 1), G);
                const uint32_t f1 = __shfl_sync(gmask, fh1, 1 & (G - 1), G);
                if (j == 0) {
                    p2pt_desc d;
                    const u64 frame_off = __ldg(a.out_off + i) + 13;
                    d.pay_off = frame_off; d.pay_len = 0; d.stream_id = 0; d.msg_type = 0;
                    for (int q = 0; q < 6; q++) d.pad[q] = 0;
                    if (!ok) d.status = P2PT_ST_AUTH_FAIL;
                    else if (L < 5) { d.status = P2PT_ST_TOO_SHORT; d.msg_type = L ? (uint8_t)(f0 & 0xff) : 0; }
                    else {
                        const uint32_t ty = f0 & 0xff;
                        d.msg_type = (uint8_t)ty;
                        if (!msg_type_valid(ty)) d.status = P2PT_ST_UNKNOWN_TYPE;
                        else {
                            d.status = P2PT_ST_OK;
                            d.stream_id = (((f0 >> 8) & 0xff) << 24) | (((f0 >> 16) & 0xff) << 16) | ((f0 >> 24) << 8) | (f1 & 0xff);
                            d.pay_off = frame_off + 5;
                            d.pay_len = L - 5;
                        }
                    }
                    a.desc[i] = d;
                    if (a.status) a.status[i] = d.status;
                }
            }
        }
    }
}

}  // namespace p2pt
