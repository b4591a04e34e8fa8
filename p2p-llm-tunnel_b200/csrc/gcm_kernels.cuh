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
    uint32_t hpow[32][4];   // H^1 .. H^32, U128 word order
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

constexpr uint32_t kLutBytes = 131072;   // 4 tables x 256 entries x 32 replicas x 4 B

struct SmemLut {     // hot path: bank-replicated shared-memory tables
    uint32_t lb;     // 64 KiB-aligned shared address of the tables + 4 * lane
    template <int K, int T> __device__ __forceinline__ uint32_t look(uint32_t w) const {
        // address = base | byte_K(w) << 8 | lane << 2, built by a single PRMT
        const uint32_t idx = __byte_perm(w, lb, 0x7604 | (K << 4));
        uint32_t v;
        if (T == 0)      asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(idx));
        else if (T == 1) asm("ld.shared.u32 %0, [%1+128];" : "=r"(v) : "r"(idx));
        else if (T == 2) asm("ld.shared.u32 %0, [%1+65536];" : "=r"(v) : "r"(idx));
        else             asm("ld.shared.u32 %0, [%1+65664];" : "=r"(v) : "r"(idx));
        return v;
    }
};

template <class Lut>
__device__ __forceinline__ void aes128_encrypt(const Lut& lut, const uint32_t (&rk)[44],
                                               uint32_t& s0, uint32_t& s1, uint32_t& s2, uint32_t& s3) {
    s0 ^= rk[0]; s1 ^= rk[1]; s2 ^= rk[2]; s3 ^= rk[3];
#pragma unroll
    for (int r = 1; r < 10; r++) {
        uint32_t t0 = lut.template look<0, 0>(s0) ^ lut.template look<1, 1>(s1) ^ lut.template look<2, 2>(s2) ^ lut.template look<3, 3>(s3) ^ rk[4 * r + 0];
        uint32_t t1 = lut.template look<0, 0>(s1) ^ lut.template look<1, 1>(s2) ^ lut.template look<2, 2>(s3) ^ lut.template look<3, 3>(s0) ^ rk[4 * r + 1];
        uint32_t t2 = lut.template look<0, 0>(s2) ^ lut.template look<1, 1>(s3) ^ lut.template look<2, 2>(s0) ^ lut.template look<3, 3>(s1) ^ rk[4 * r + 2];
        uint32_t t3 = lut.template look<0, 0>(s3) ^ lut.template look<1, 1>(s0) ^ lut.template look<2, 2>(s1) ^ lut.template look<3, 3>(s2) ^ rk[4 * r + 3];
        s0 = t0; s1 = t1; s2 = t2; s3 = t3;
    }
    // last round: SubBytes + ShiftRows only.  S[x] sits in byte 0 of Te2, byte 1 of Te3, byte 2 of
    // Te0 and byte 3 of Te1, i.e. already in the byte lane it is needed in.
#define P2PT_LAST(a, b, c, d, k)                                                             \
    (__byte_perm(__byte_perm(lut.template look<0, 2>(a), lut.template look<1, 3>(b), 0x0050), \
                 __byte_perm(lut.template look<2, 0>(c), lut.template look<3, 1>(d), 0x7200), 0x7610) ^ rk[40 + k])
    uint32_t o0 = P2PT_LAST(s0, s1, s2, s3, 0);
    uint32_t o1 = P2PT_LAST(s1, s2, s3, s0, 1);
    uint32_t o2 = P2PT_LAST(s2, s3, s0, s1, 2);
    uint32_t o3 = P2PT_LAST(s3, s0, s1, s2, 3);
#undef P2PT_LAST
    s0 = o0; s1 = o1; s2 = o2; s3 = o3;
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
    aes128_encrypt(lut, rk, h0, h1, h2, h3);
    const U128 H = block_to_u128(h0, h1, h2, h3);
    GhashKey hk;
    ghash_key_prepare(H, hk);
    U128 p = H;
    for (int e = 0; e < 32; e++) {
        sk->hpow[e][0] = p.w[0]; sk->hpow[e][1] = p.w[1]; sk->hpow[e][2] = p.w[2]; sk->hpow[e][3] = p.w[3];
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
};

constexpr int kGcmThreads = 512;

__device__ __forceinline__ uint32_t shared_addr(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

template <int G, int MODE>
__global__ void __launch_bounds__(kGcmThreads, 1) gcm_record_kernel(const GcmArgs a) {
    extern __shared__ __align__(16) uint8_t dyn_smem[];
    constexpr bool SEAL = (MODE == kSeal || MODE == kPipeSeal);
    constexpr uint32_t GMASK_BASE = (G == 32) ? 0xffffffffu : ((1u << G) - 1u);

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

    const u64 groups_per_cta = kGcmThreads / G;
    const u64 group0 = (u64)blockIdx.x * groups_per_cta + (threadIdx.x / G);
    const u64 group_stride = (u64)gridDim.x * groups_per_cta;

    for (u64 i = group0; i < a.n; i += group_stride) {
        // ------------------------------------------------------------------ record geometry
        const u64 ioff = __ldg(a.in_off + i);
        const u64 ilen = __ldg(a.in_off + i + 1) - ioff;
        const uint8_t* ibase = a.in + ioff;
        uint8_t* obase = a.out + __ldg(a.out_off + i);
        const uint32_t sidx = __ldg(a.session + i);
        const SessionKey* sk = a.keys + (sidx < a.n_sessions ? sidx : 0);

        uint32_t L;                     // plaintext / ciphertext length
        uint32_t st = P2PT_ST_OK;
        uint32_t hdr_type = 23, hdr_v0 = 0xfe, hdr_v1 = 0xfd;
        const uint8_t* body_lo;         // input body bytes live in [body_lo, body_hi)
        const uint8_t* body_hi;
        const uint8_t* body0;           // address of input body byte 0 (may precede body_lo: pipe-seal)
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

        if (st != P2PT_ST_OK) {         // group-uniform early outs
            if (MODE == kOpen || MODE == kPipeOpen) {
                uint32_t olen = 0;
                if (st == P2PT_ST_PASS_THROUGH) {           // ChangeCipherSpec: copy unmodified
                    for (u64 b = j; b < ilen; b += G) obase[b] = __ldg(ibase + b);
                    olen = (uint32_t)ilen;
                }
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
            if (j == 0 && a.status) a.status[i] = (uint8_t)st;
            continue;
        }

        const uint32_t nc = (L + 15) >> 4;          // CTR / ciphertext blocks
        const uint32_t slots = nc + 2;              // + J0/AAD slot + length slot
        constexpr uint32_t PREFIX = SEAL ? 21u : 13u;
        uint8_t* ob0 = obase + PREFIX;              // output body byte 0
        const unsigned so = (unsigned)((uintptr_t)ob0 & 15);

        // ------------------------------------------------------------------ per-record key material
        uint32_t rk[44];
        {
            const uint4* p = reinterpret_cast<const uint4*>(sk->rk);
#pragma unroll
            for (int q = 0; q < 11; q++) {
                uint4 v = __ldg(p + q);
                rk[4 * q] = v.x; rk[4 * q + 1] = v.y; rk[4 * q + 2] = v.z; rk[4 * q + 3] = v.w;
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
        GhashKey kG;
        {
            uint4 v = __ldg(reinterpret_cast<const uint4*>(sk->hpow[G - 1]));
            U128 hg; hg.w[0] = v.x; hg.w[1] = v.y; hg.w[2] = v.z; hg.w[3] = v.w;
            ghash_key_prepare(hg, kG);
        }

        U128 acc; acc.w[0] = acc.w[1] = acc.w[2] = acc.w[3] = 0;
        uint4 ej0 = make_uint4(0, 0, 0, 0);                  // E_K(J0), valid in group lane 0
        uint4 carry = make_uint4(0, 0, 0, 0);                // output block of lane G-1, previous round
        uint32_t fh0 = 0, fh1 = 0;                           // pipe-open: first 8 plaintext bytes (lane 1)
        uint32_t last_u = 0xffffffffu;

        const uint32_t rounds = (slots + G - 1) / G;
        for (uint32_t t = 0; t < rounds; t++) {
            const uint32_t u = t * G + j;
            const int c = (int)u - 1;                        // body block index; -1 = J0/AAD slot
            const bool active = u < slots;
            const bool is_j0 = (u == 0);
            const bool is_len = (u == slots - 1);
            const bool is_data = active && !is_j0 && !is_len;

            // ---- AES-CTR keystream for counter c + 2 (J0 = counter 1)
            uint32_t k0 = ivw, k1 = n0, k2 = n1, k3 = bswap32((uint32_t)(c + 2));
            aes128_encrypt(lut, rk, k0, k1, k2, k3);

            // ---- the slot's block
            uint4 oblk = make_uint4(0, 0, 0, 0);             // bytes this lane contributes to the output stream
            int ovalid_lo = 0, ovalid_hi = 0;                // valid byte range of oblk
            U128 gin; gin.w[0] = gin.w[1] = gin.w[2] = gin.w[3] = 0;
            if (is_data) {
                const uint32_t pos = 16u * (uint32_t)c;
                const uint32_t nb = (L - pos) < 16u ? (L - pos) : 16u;
                uint4 x = load16_bounded(body0 + pos, body_lo, body_hi);
                if (MODE == kPipeSeal && c == 0) {           // synthesise [type][sid BE] in front of the payload
                    const uint32_t sid = __ldg(a.sids + i);
                    x.x = (uint32_t)__ldg(a.types + i) | ((sid >> 24) << 8) | (((sid >> 16) & 0xff) << 16) | (((sid >> 8) & 0xff) << 24);
                    x.y = (x.y & 0xffffff00u) | (sid & 0xff);
                }
                uint4 y = make_uint4(x.x ^ k0, x.y ^ k1, x.z ^ k2, x.w ^ k3);
                if (nb < 16) {                               // zero the bytes past the end of the body
                    const uint32_t m0 = nb >= 4 ? 0xffffffffu : (nb ? (0xffffffffu >> (32 - 8 * nb)) : 0u);
                    const uint32_t m1 = nb >= 8 ? 0xffffffffu : (nb > 4 ? (0xffffffffu >> (32 - 8 * (nb - 4))) : 0u);
                    const uint32_t m2 = nb >= 12 ? 0xffffffffu : (nb > 8 ? (0xffffffffu >> (32 - 8 * (nb - 8))) : 0u);
                    const uint32_t m3 = nb > 12 ? (0xffffffffu >> (32 - 8 * (nb - 12))) : 0u;
                    y.x &= m0; y.y &= m1; y.z &= m2; y.w &= m3;
                }
                const uint4 ct = SEAL ? y : x;               // GHASH always runs over the ciphertext
                gin = block_to_u128(ct.x, ct.y, ct.z, ct.w);
                oblk = y; ovalid_lo = 0; ovalid_hi = (int)nb;
                if (MODE == kPipeOpen && c == 0) { fh0 = y.x; fh1 = y.y; }
            } else if (is_j0) {
                ej0 = make_uint4(k0, k1, k2, k3);
                // AAD = epoch(2) seq(6) type ver(2) len(2); record header = type ver(2) epoch(2) seq(6) len(2)
                uint32_t e_hi, s_w0, s_w1;                   // header bytes 3..4, 5..8, 9..10 as BE numbers
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
                const uint32_t plen = L & 0xffff;            // `as u16` in the reference
                gin.w[3] = (e_hi << 16) | (s_w0 >> 16);
                gin.w[2] = (s_w0 << 16) | s_w1;
                gin.w[1] = (hdr_type << 24) | (hdr_v0 << 16) | (hdr_v1 << 8) | (plen >> 8);
                gin.w[0] = (plen & 0xff) << 24;
                // the 16 output bytes that precede the body
                if (SEAL) {
                    const uint32_t rlen = (8 + L + 16) & 0xffff;
                    // record bytes 5..20: seq(6) len(2) nonce(8); bytes 0..4 are stored right here
                    obase[0] = (uint8_t)hdr_type; obase[1] = (uint8_t)hdr_v0; obase[2] = (uint8_t)hdr_v1;
                    obase[3] = (uint8_t)(e_hi >> 8); obase[4] = (uint8_t)e_hi;
                    oblk.x = bswap32(s_w0);
                    oblk.y = ((s_w1 >> 8) & 0xff) | ((s_w1 & 0xff) << 8) | ((rlen >> 8) << 16) | ((rlen & 0xff) << 24);
                    oblk.z = n0; oblk.w = n1;
                    ovalid_lo = 0; ovalid_hi = 16;
                } else {
                    // output bytes -3..12: (3 bytes not ours) then the 13 header bytes unchanged
                    const uint32_t l_w = ((uint32_t)__ldg(ibase + 11) << 8) | __ldg(ibase + 12);
                    oblk.x = hdr_type << 24;
                    oblk.y = hdr_v0 | (hdr_v1 << 8) | ((e_hi >> 8) << 16) | ((e_hi & 0xff) << 24);
                    oblk.z = bswap32(s_w0);
                    oblk.w = ((s_w1 >> 8) & 0xff) | ((s_w1 & 0xff) << 8) | ((l_w >> 8) << 16) | ((l_w & 0xff) << 24);
                    ovalid_lo = 3; ovalid_hi = 16;
                }
            } else if (is_len) {
                gin.w[2] = 13 * 8;                           // len(AAD) in bits
                gin.w[0] = L << 3;                           // len(C) in bits (L < 2^16)
            }

            // ---- GHASH: Horner in H^G; the lane's last block is multiplied after the loop
            if (active) {
                acc.w[0] ^= gin.w[0]; acc.w[1] ^= gin.w[1]; acc.w[2] ^= gin.w[2]; acc.w[3] ^= gin.w[3];
                last_u = u;
            }
            if (u + G < slots) acc = gf128_mul(acc, kG);

            // ---- output stream: aligned vector V_c = tail of block c-1 || head of block c
            uint4 prev;
            prev.x = __shfl_up_sync(gmask, oblk.x, 1, G); prev.y = __shfl_up_sync(gmask, oblk.y, 1, G);
            prev.z = __shfl_up_sync(gmask, oblk.z, 1, G); prev.w = __shfl_up_sync(gmask, oblk.w, 1, G);
            if (j == 0) prev = carry;
            carry.x = __shfl_sync(gmask, oblk.x, G - 1, G); carry.y = __shfl_sync(gmask, oblk.y, G - 1, G);
            carry.z = __shfl_sync(gmask, oblk.z, G - 1, G); carry.w = __shfl_sync(gmask, oblk.w, G - 1, G);

            // W(c): vector c is a full interior vector
            auto vec_written = [&](int cc) -> bool {
                if (cc < 0 || (uint32_t)(16 * cc + 16) > L) return false;
                if (!SEAL && cc == 0 && so > 13) return false;   // would start before the output record
                return true;
            };
            if (active && !is_len) {
                const bool wc = vec_written(c), wn = vec_written(c + 1);
                uint8_t* bp = ob0 + 16 * (int64_t)c;             // where this lane's block starts
                if (wc) stg128(bp - so, so ? funnel16(prev, oblk, 16 - so) : oblk);
                // bytes of my block no full vector covers
                const int split = 16 - (int)so;                  // head [0,split) belongs to V_c, tail to V_c+1
                if (!wc) store_bytes(bp, oblk, ovalid_lo, ovalid_hi < split ? ovalid_hi : split);
                if (!wn && so) store_bytes(bp, oblk, ovalid_lo > split ? ovalid_lo : split, ovalid_hi);
            }
        }

        // ------------------------------------------------------------------ finish GHASH
        {
            const uint32_t fp = (last_u == 0xffffffffu) ? 1u : (slots - last_u);     // 1..G
            uint4 v = __ldg(reinterpret_cast<const uint4*>(sk->hpow[fp - 1]));
            U128 hf; hf.w[0] = v.x; hf.w[1] = v.y; hf.w[2] = v.z; hf.w[3] = v.w;
            ghash_key_prepare(hf, kG);
            acc = gf128_mul(acc, kG);                            // lanes that never ran hold 0
        }
#pragma unroll
        for (int o = G >> 1; o > 0; o >>= 1) {
            acc.w[0] ^= __shfl_xor_sync(gmask, acc.w[0], o, G);
            acc.w[1] ^= __shfl_xor_sync(gmask, acc.w[1], o, G);
            acc.w[2] ^= __shfl_xor_sync(gmask, acc.w[2], o, G);
            acc.w[3] ^= __shfl_xor_sync(gmask, acc.w[3], o, G);
        }
        uint4 tag;                                               // little-endian words, memory order
        tag.x = bswap32(acc.w[3]) ^ __shfl_sync(gmask, ej0.x, 0, G);
        tag.y = bswap32(acc.w[2]) ^ __shfl_sync(gmask, ej0.y, 0, G);
        tag.z = bswap32(acc.w[1]) ^ __shfl_sync(gmask, ej0.z, 0, G);
        tag.w = bswap32(acc.w[0]) ^ __shfl_sync(gmask, ej0.w, 0, G);

        if (SEAL) {
            uint8_t* tp = ob0 + L;
            for (int b = j; b < 16; b += G) tp[b] = (uint8_t)(word_of(tag, b >> 2) >> (8 * (b & 3)));
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
            if (!ok) {
                for (u64 b = j; b < 13 + (u64)L; b += G) obase[b] = 0;
            }
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
