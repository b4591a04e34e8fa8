/*
 * p2pt_oracle.c -- CPU oracle (plain C, scalar, no intrinsics).  TEST INFRASTRUCTURE ONLY:
 * see the header comment in p2pt_oracle.h for who may use it and for the parity-pinning status
 * (codec, AES-GCM arithmetic, DTLS record layout and replay window are pinned; the SCTP packet layout is "parity unpinned").
 *
 * Every function cites the reference file:line (or the standard, where the arithmetic lives in an
 * un-vendored crate) it restates.  Nothing here is copied from the reference: the codec is 20
 * lines of Rust restated in C, AES/GHASH are written from FIPS-197 / SP 800-38D.
 */
#include "p2pt_oracle.h"
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------ */
/* Frame codec                                                                                */
/* ------------------------------------------------------------------------------------------ */

/* MessageType::from_u8 -- /root/reference/tunnel/src/protocol.rs:103-118 (tag set :88-100). */
int orc_msg_type_valid(uint8_t b) {
    switch (b) {
    case 1: case 2: case 3: case 4:
    case 10: case 11: case 12:
    case 20: case 21: case 22:
    case 99:
        return 1;
    default:
        return 0;
    }
}

/* TunnelMessage::encode -- protocol.rs:148-154: [type u8][stream_id u32 BE][payload]. */
void orc_frame_encode_batch(uint64_t n, const uint8_t* types, const uint32_t* sids,
                            const uint8_t* payload, const uint64_t* pay_off,
                            uint8_t* out, const uint64_t* out_off) {
    for (uint64_t i = 0; i < n; i++) {
        uint8_t* o = out + out_off[i];
        uint64_t len = pay_off[i + 1] - pay_off[i];
        o[0] = types[i];
        o[1] = (uint8_t)(sids[i] >> 24);
        o[2] = (uint8_t)(sids[i] >> 16);
        o[3] = (uint8_t)(sids[i] >> 8);
        o[4] = (uint8_t)(sids[i]);
        if (len) memcpy(o + 5, payload + pay_off[i], len);
    }
}

/* TunnelMessage::decode -- protocol.rs:157-172.  Check order is the reference's:
 * (1) len < 5 -> too short (:158-160); (2) tag via from_u8 -> unknown type (:161-163);
 * (3) BE stream_id from bytes 1..5 (:165); payload = the rest (may be empty, :416-421). */
void orc_frame_decode_batch(uint64_t n, const uint8_t* in, const uint64_t* in_off, orc_desc* desc) {
    for (uint64_t i = 0; i < n; i++) {
        const uint8_t* f = in + in_off[i];
        uint64_t len = in_off[i + 1] - in_off[i];
        orc_desc d;
        memset(&d, 0, sizeof d);
        if (len < 5) {
            d.status = ORC_TOO_SHORT;
            d.pay_off = in_off[i];
            d.pay_len = 0;
            d.msg_type = len ? f[0] : 0;
        } else if (!orc_msg_type_valid(f[0])) {
            d.status = ORC_UNKNOWN_TYPE;
            d.msg_type = f[0];
            d.pay_off = in_off[i];
            d.pay_len = 0;
        } else {
            d.status = ORC_OK;
            d.msg_type = f[0];
            d.stream_id = ((uint32_t)f[1] << 24) | ((uint32_t)f[2] << 16) | ((uint32_t)f[3] << 8) | f[4];
            d.pay_off = in_off[i] + 5;
            d.pay_len = (uint32_t)(len - 5);
        }
        desc[i] = d;
    }
}

/* Demux by stream_id -- the batched form of the per-frame routing in
 * /root/reference/tunnel/src/proxy.rs:118-153 (HashMap<u32, sender> lookup per frame, per-stream
 * order preserved) and the body reassembly in /root/reference/tunnel/src/serve.rs:112-139
 * (ReqBody payloads appended to a contiguous per-stream Vec<u8>).
 * Undecodable frames are skipped (proxy.rs:110-116 / serve.rs:103-109 warn + continue).
 * Output: streams ascending by id; perm = frame indices grouped by stream, arrival order kept;
 * csr[k+1] offsets into perm; gathered = per-stream concatenation of the payloads of body frames
 * (ReqBody = 11, ResBody = 21); gathered_off[k+1]. */
typedef struct { uint64_t sid; uint32_t idx; } orc_pair;   /* sid = session << 32 | stream_id */

static void orc_stable_sort_pairs(orc_pair* a, orc_pair* tmp, uint64_t n) {
    /* bottom-up merge sort: stable and deterministic; result always left in a */
    for (uint64_t w = 1; w < n; w *= 2) {
        for (uint64_t lo = 0; lo < n; lo += 2 * w) {
            uint64_t mid = lo + w < n ? lo + w : n;
            uint64_t hi = lo + 2 * w < n ? lo + 2 * w : n;
            uint64_t i = lo, j = mid, k = lo;
            while (i < mid && j < hi) tmp[k++] = (a[j].sid < a[i].sid) ? a[j++] : a[i++];
            while (i < mid) tmp[k++] = a[i++];
            while (j < hi) tmp[k++] = a[j++];
        }
        memcpy(a, tmp, n * sizeof(orc_pair));
    }
}

uint32_t orc_frame_demux_batch(uint64_t n, const orc_desc* desc, const uint32_t* session, const uint8_t* in,
                               uint32_t* uniq_sid, uint32_t* uniq_session, uint64_t* csr, uint32_t* perm,
                               uint8_t* gathered, uint64_t* gathered_off) {
    orc_pair* a = (orc_pair*)malloc((n ? n : 1) * sizeof(orc_pair));
    orc_pair* b = (orc_pair*)malloc((n ? n : 1) * sizeof(orc_pair));
    uint64_t m = 0;
    for (uint64_t i = 0; i < n; i++)
        if (desc[i].status == ORC_OK) {
            /* one stream_id namespace per tunnel session (proxy.rs:52) */
            a[m].sid = desc[i].stream_id | (session ? (uint64_t)session[i] << 32 : 0);
            a[m].idx = (uint32_t)i;
            m++;
        }
    orc_stable_sort_pairs(a, b, m);
    orc_pair* s = a;

    uint32_t k = 0;
    uint64_t goff = 0;
    for (uint64_t j = 0; j < m; j++) {
        if (j == 0 || s[j].sid != s[j - 1].sid) {
            uniq_sid[k] = (uint32_t)s[j].sid;
            if (uniq_session) uniq_session[k] = (uint32_t)(s[j].sid >> 32);
            csr[k] = j;
            if (gathered_off) gathered_off[k] = goff;
            k++;
        }
        perm[j] = s[j].idx;
        const orc_desc* d = &desc[s[j].idx];
        if (d->msg_type == 11 || d->msg_type == 21) {
            if (gathered && d->pay_len) memcpy(gathered + goff, in + d->pay_off, d->pay_len);
            goff += d->pay_len;
        }
    }
    csr[k] = m;
    if (gathered_off) gathered_off[k] = goff;
    free(a);
    free(b);
    return k;
}

/* ------------------------------------------------------------------------------------------ */
/* AES-128 (FIPS-197).  The reference reaches this through aes-gcm 0.10.3 -> aes 0.8.4          */
/* (/root/reference/tunnel/Cargo.lock:27-29, :16-17); un-vendored, so restated from the standard. */
/* ------------------------------------------------------------------------------------------ */

static uint8_t g_sbox[256], g_mul2[256], g_mul3[256];
static volatile int g_sbox_ready = 0;

static uint8_t gf8_mul(uint8_t a, uint8_t b) {
    uint8_t p = 0;
    for (int i = 0; i < 8; i++) {
        if (b & 1) p ^= a;
        uint8_t hi = a & 0x80;
        a <<= 1;
        if (hi) a ^= 0x1b; /* x^8 + x^4 + x^3 + x + 1 (FIPS-197 4.2) */
        b >>= 1;
    }
    return p;
}

/* S-box computed from its definition (FIPS-197 5.1.1): multiplicative inverse then affine map. */
static void sbox_init(void) {
    if (g_sbox_ready) return;
    for (int x = 0; x < 256; x++) {
        uint8_t inv = 0;
        if (x) {
            for (int y = 1; y < 256; y++)
                if (gf8_mul((uint8_t)x, (uint8_t)y) == 1) { inv = (uint8_t)y; break; }
        }
        uint8_t s = inv;
        uint8_t r = inv;
        for (int i = 0; i < 4; i++) {
            r = (uint8_t)((r << 1) | (r >> 7));
            s ^= r;
        }
        g_sbox[x] = s ^ 0x63;
        g_mul2[x] = gf8_mul((uint8_t)x, 2);
        g_mul3[x] = gf8_mul((uint8_t)x, 3);
    }
    g_sbox_ready = 1;   /* idempotent: concurrent first calls write identical bytes */
}

/* KeyExpansion, FIPS-197 5.2, Nk = 4, Nr = 10: 44 words = 176 bytes. */
void orc_aes128_expand(const uint8_t key[16], uint8_t rk[176]) {
    sbox_init();
    memcpy(rk, key, 16);
    uint8_t rcon = 1;
    for (int i = 4; i < 44; i++) {
        uint8_t t[4];
        memcpy(t, rk + 4 * (i - 1), 4);
        if (i % 4 == 0) {
            uint8_t t0 = t[0];
            t[0] = g_sbox[t[1]] ^ rcon;
            t[1] = g_sbox[t[2]];
            t[2] = g_sbox[t[3]];
            t[3] = g_sbox[t0];
            rcon = gf8_mul(rcon, 2);
        }
        for (int j = 0; j < 4; j++) rk[4 * i + j] = rk[4 * (i - 4) + j] ^ t[j];
    }
}

/* Cipher, FIPS-197 5.1.  State is column-major: s[4*c + r]. */
void orc_aes128_encrypt_block(const uint8_t rk[176], const uint8_t in[16], uint8_t out[16]) {
    sbox_init();
    uint8_t s[16], t[16];
    for (int i = 0; i < 16; i++) s[i] = in[i] ^ rk[i];
    for (int round = 1; round <= 10; round++) {
        /* SubBytes + ShiftRows: row r rotates left by r columns */
        for (int c = 0; c < 4; c++)
            for (int r = 0; r < 4; r++)
                t[4 * c + r] = g_sbox[s[4 * ((c + r) & 3) + r]];
        if (round < 10) {
            for (int c = 0; c < 4; c++) {
                uint8_t a0 = t[4 * c], a1 = t[4 * c + 1], a2 = t[4 * c + 2], a3 = t[4 * c + 3];
                /* MixColumns, FIPS-197 5.1.3: {02} {03} {01} {01} circulant */
                s[4 * c + 0] = g_mul2[a0] ^ g_mul3[a1] ^ a2 ^ a3;
                s[4 * c + 1] = a0 ^ g_mul2[a1] ^ g_mul3[a2] ^ a3;
                s[4 * c + 2] = a0 ^ a1 ^ g_mul2[a2] ^ g_mul3[a3];
                s[4 * c + 3] = g_mul3[a0] ^ a1 ^ a2 ^ g_mul2[a3];
            }
        } else {
            memcpy(s, t, 16);
        }
        for (int i = 0; i < 16; i++) s[i] ^= rk[16 * round + i];
    }
    memcpy(out, s, 16);
}

/* ------------------------------------------------------------------------------------------ */
/* GHASH / GCM (NIST SP 800-38D).  Reference: ghash 0.5.1 / polyval 0.6.2 (Cargo.lock:810, 1542). */
/* ------------------------------------------------------------------------------------------ */

/* SP 800-38D Algorithm 1: X * Y in GF(2^128), bit 0 = MSB of byte 0, R = 11100001 || 0^120. */
void orc_gf128_mul(const uint8_t x[16], const uint8_t y[16], uint8_t out[16]) {
    uint8_t z[16] = {0}, v[16];
    memcpy(v, y, 16);
    for (int i = 0; i < 128; i++) {
        if ((x[i >> 3] >> (7 - (i & 7))) & 1)
            for (int j = 0; j < 16; j++) z[j] ^= v[j];
        int lsb = v[15] & 1;
        for (int j = 15; j > 0; j--) v[j] = (uint8_t)((v[j] >> 1) | (v[j - 1] << 7));
        v[0] >>= 1;
        if (lsb) v[0] ^= 0xe1;
    }
    memcpy(out, z, 16);
}

static void ghash_absorb(uint8_t y[16], const uint8_t h[16], const uint8_t* p, size_t len) {
    while (len) {
        size_t n = len < 16 ? len : 16;
        for (size_t j = 0; j < n; j++) y[j] ^= p[j];
        orc_gf128_mul(y, h, y);
        p += n;
        len -= n;
    }
}

/* SP 800-38D 6.4 / 7.1 step 5: GHASH_H(A || 0^v || C || 0^u || [len(A)]64 || [len(C)]64). */
void orc_ghash(const uint8_t h[16], const uint8_t* aad, size_t aad_len,
               const uint8_t* c, size_t c_len, uint8_t out[16]) {
    uint8_t y[16] = {0}, lb[16];
    ghash_absorb(y, h, aad, aad_len);
    ghash_absorb(y, h, c, c_len);
    uint64_t abits = (uint64_t)aad_len * 8, cbits = (uint64_t)c_len * 8;
    for (int i = 0; i < 8; i++) {
        lb[i] = (uint8_t)(abits >> (56 - 8 * i));
        lb[8 + i] = (uint8_t)(cbits >> (56 - 8 * i));
    }
    for (int j = 0; j < 16; j++) y[j] ^= lb[j];
    orc_gf128_mul(y, h, y);
    memcpy(out, y, 16);
}

static void inc32(uint8_t cb[16]) {
    for (int i = 15; i >= 12; i--)
        if (++cb[i]) break;
}

static void gctr(const uint8_t rk[176], uint8_t cb[16], const uint8_t* in, size_t len, uint8_t* out) {
    uint8_t ks[16];
    while (len) {
        size_t n = len < 16 ? len : 16;
        orc_aes128_encrypt_block(rk, cb, ks);
        for (size_t j = 0; j < n; j++) out[j] = in[j] ^ ks[j];
        inc32(cb);
        in += n; out += n; len -= n;
    }
}

/* SP 800-38D Algorithm 4 (GCM-AE), 96-bit IV: J0 = IV || 0^31 || 1. */
void orc_gcm_encrypt(const uint8_t key[16], const uint8_t iv12[12],
                     const uint8_t* aad, size_t aad_len,
                     const uint8_t* pt, size_t len, uint8_t* ct, uint8_t tag[16]) {
    uint8_t rk[176], h[16], z[16] = {0}, j0[16], cb[16], s[16], ej0[16];
    orc_aes128_expand(key, rk);
    orc_aes128_encrypt_block(rk, z, h);
    memcpy(j0, iv12, 12);
    j0[12] = 0; j0[13] = 0; j0[14] = 0; j0[15] = 1;
    memcpy(cb, j0, 16);
    inc32(cb);
    gctr(rk, cb, pt, len, ct);
    orc_ghash(h, aad, aad_len, ct, len, s);
    orc_aes128_encrypt_block(rk, j0, ej0);
    for (int i = 0; i < 16; i++) tag[i] = s[i] ^ ej0[i];
}

/* SP 800-38D Algorithm 5 (GCM-AD).  Tag compare accumulates differences (constant time in the
 * reference via subtle 2.6.1, Cargo.lock:2141).  On failure pt is zeroed. */
int orc_gcm_decrypt(const uint8_t key[16], const uint8_t iv12[12],
                    const uint8_t* aad, size_t aad_len,
                    const uint8_t* ct, size_t len, const uint8_t tag[16], uint8_t* pt) {
    uint8_t rk[176], h[16], z[16] = {0}, j0[16], cb[16], s[16], ej0[16];
    orc_aes128_expand(key, rk);
    orc_aes128_encrypt_block(rk, z, h);
    memcpy(j0, iv12, 12);
    j0[12] = 0; j0[13] = 0; j0[14] = 0; j0[15] = 1;
    orc_ghash(h, aad, aad_len, ct, len, s);
    orc_aes128_encrypt_block(rk, j0, ej0);
    uint8_t diff = 0;
    for (int i = 0; i < 16; i++) diff |= (uint8_t)(s[i] ^ ej0[i] ^ tag[i]);
    if (diff) {
        if (len) memset(pt, 0, len);
        return -1;
    }
    memcpy(cb, j0, 16);
    inc32(cb);
    gctr(rk, cb, ct, len, pt);
    return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* DTLS 1.2 record layer, AES-128-GCM suite.  Restates webrtc-dtls 0.10.0 crypto/crypto_gcm.rs  */
/* (CryptoGcm::encrypt / decrypt) and crypto/mod.rs generate_aead_additional_data -- UN-VENDORED */
/* (/root/reference/tunnel/Cargo.lock:2822-2824); reached from the reference at every dc.send   */
/* (/root/reference/tunnel/src/serve.rs:274) and dc.on_message (tunnel/src/rtc.rs:88-92).       */
/* ------------------------------------------------------------------------------------------ */

/* AAD(13) = epoch u16 || seq u48 || content_type || version(2) || payload_len u16, all BE.
 * hdr13 = type(1) ver(2) epoch(2) seq(6) len(2) (RFC 6347 4.1). */
void orc_dtls_aad(const uint8_t hdr13[13], uint16_t payload_len, uint8_t aad[13]) {
    memcpy(aad, hdr13 + 3, 8);   /* epoch || seq48 */
    aad[8] = hdr13[0];           /* content type */
    aad[9] = hdr13[1];           /* version major */
    aad[10] = hdr13[2];          /* version minor */
    aad[11] = (uint8_t)(payload_len >> 8);
    aad[12] = (uint8_t)payload_len;
}

int orc_dtls_seal(const uint8_t key[16], const uint8_t iv4[4], const uint8_t nonce8[8],
                  const uint8_t* raw, uint64_t raw_len, uint8_t* out) {
    if (raw_len < 13) return ORC_BAD_HEADER;
    uint64_t L = raw_len - 13;
    uint8_t iv[12], aad[13];
    memcpy(iv, iv4, 4);
    memcpy(iv + 4, nonce8, 8);   /* nonce = write_iv[0..4] || explicit 8 (RFC 5288 3) */
    orc_dtls_aad(raw, (uint16_t)L, aad);
    memcpy(out, raw, 13);
    memcpy(out + 13, nonce8, 8);
    orc_gcm_encrypt(key, iv, aad, 13, raw + 13, (size_t)L, out + 21, out + 21 + L);
    uint16_t rlen = (uint16_t)(8 + L + 16);   /* header length rewritten to include nonce + tag */
    out[11] = (uint8_t)(rlen >> 8);
    out[12] = (uint8_t)rlen;
    return ORC_OK;
}

int orc_dtls_open(const uint8_t key[16], const uint8_t iv4[4],
                  const uint8_t* rec, uint64_t rec_len, uint8_t* out, uint64_t* out_len) {
    *out_len = 0;
    if (rec_len < 13) return ORC_BAD_HEADER;                       /* unmarshal: unexpected EOF */
    if (!(rec[1] == 0xfe && (rec[2] == 0xfd || rec[2] == 0xff)))   /* DTLS 1.2 or 1.0 only */
        return ORC_BAD_HEADER;
    if (rec[0] == 20) {                                            /* ChangeCipherSpec: unmodified */
        memcpy(out, rec, rec_len);
        *out_len = rec_len;
        return ORC_PASS_THROUGH;
    }
    if (rec_len <= 13 + 8) return ORC_NO_ROOM_FOR_NONCE;
    if (rec_len < 13 + 8 + 16) return ORC_AUTH_FAIL;               /* shorter than a tag */
    uint64_t L = rec_len - 13 - 8 - 16;
    uint8_t iv[12], aad[13];
    memcpy(iv, iv4, 4);
    memcpy(iv + 4, rec + 13, 8);
    orc_dtls_aad(rec, (uint16_t)L, aad);
    memcpy(out, rec, 13);                                          /* header NOT rewritten on open */
    if (orc_gcm_decrypt(key, iv, aad, 13, rec + 21, (size_t)L, rec + 21 + L, out + 13) != 0) {
        memset(out, 0, 13 + L);
        return ORC_AUTH_FAIL;
    }
    *out_len = 13 + L;
    return ORC_OK;
}

void orc_dtls_seal_batch(uint64_t n, const uint32_t* session,
                         const uint8_t* keys16, const uint8_t* ivs4, const uint8_t* nonce8,
                         const uint8_t* raw, const uint64_t* raw_off,
                         uint8_t* rec, const uint64_t* rec_off, uint8_t* status) {
    for (uint64_t i = 0; i < n; i++) {
        uint32_t s = session[i];
        status[i] = (uint8_t)orc_dtls_seal(keys16 + 16 * (uint64_t)s, ivs4 + 4 * (uint64_t)s,
                                           nonce8 + 8 * i, raw + raw_off[i],
                                           raw_off[i + 1] - raw_off[i], rec + rec_off[i]);
    }
}

void orc_dtls_open_batch(uint64_t n, const uint32_t* session,
                         const uint8_t* keys16, const uint8_t* ivs4,
                         const uint8_t* rec, const uint64_t* rec_off,
                         uint8_t* out, const uint64_t* out_off, uint32_t* out_len, uint8_t* status) {
    for (uint64_t i = 0; i < n; i++) {
        uint32_t s = session[i];
        uint64_t ol = 0;
        status[i] = (uint8_t)orc_dtls_open(keys16 + 16 * (uint64_t)s, ivs4 + 4 * (uint64_t)s,
                                           rec + rec_off[i], rec_off[i + 1] - rec_off[i],
                                           out + out_off[i], &ol);
        out_len[i] = (uint32_t)ol;
    }
}

/* ------------------------------------------------------------------------------------------ */
/* SCTP DATA chunks + CRC32c.  webrtc-sctp 0.10.0 (un-vendored): packet.rs marshal / unmarshal,  */
/* chunk_payload_data.rs, util.rs generate_packet_checksum.  Layout per RFC 4960 3.1 / 3.3.1.    */
/* ------------------------------------------------------------------------------------------ */

/* CRC32c (Castagnoli), bit-serial and reflected: polynomial 0x1EDC6F41 -> 0x82F63B78, init and
 * final xor 0xFFFFFFFF (RFC 3720 12.1 / RFC 4960 appendix B). */
uint32_t orc_crc32c(const uint8_t* p, size_t len) {
    uint32_t crc = 0xFFFFFFFFu;
    for (size_t i = 0; i < len; i++) {
        crc ^= p[i];
        for (int b = 0; b < 8; b++) crc = (crc >> 1) ^ (0x82F63B78u & (0u - (crc & 1u)));
    }
    return ~crc;
}

uint64_t orc_sctp_fragments(uint64_t msg_len) {
    uint64_t k = (msg_len + ORC_SCTP_MAX_PAYLOAD - 1) / ORC_SCTP_MAX_PAYLOAD;
    return k ? k : 1;
}

uint64_t orc_sctp_packets_bytes(uint64_t msg_len) {
    uint64_t k = orc_sctp_fragments(msg_len);
    uint64_t last = msg_len - (k - 1) * ORC_SCTP_MAX_PAYLOAD;
    return (k - 1) * (28 + ORC_SCTP_MAX_PAYLOAD) + 28 + ((last + 3) & ~(uint64_t)3);
}

static void put16(uint8_t* p, uint32_t v) { p[0] = (uint8_t)(v >> 8); p[1] = (uint8_t)v; }
static void put32(uint8_t* p, uint32_t v) { p[0] = (uint8_t)(v >> 24); p[1] = (uint8_t)(v >> 16); p[2] = (uint8_t)(v >> 8); p[3] = (uint8_t)v; }

uint64_t orc_sctp_packetize(const uint8_t* msg, uint64_t len, uint16_t src_port, uint16_t dst_port, uint32_t vtag,
                            uint16_t sid, uint16_t ssn, uint32_t ppid, uint32_t tsn0,
                            int dtls_hdr, uint16_t epoch, uint64_t seq48_first, uint8_t* out) {
    uint64_t k = orc_sctp_fragments(len), w = 0;
    for (uint64_t f = 0; f < k; f++) {
        uint64_t off = f * ORC_SCTP_MAX_PAYLOAD;
        uint32_t fl = (uint32_t)(len - off < ORC_SCTP_MAX_PAYLOAD ? len - off : ORC_SCTP_MAX_PAYLOAD);
        uint32_t padded = (fl + 3) & ~3u, plen = 28 + padded;
        if (dtls_hdr) {
            uint8_t* h = out + w;
            uint64_t seq = seq48_first + f;
            h[0] = 23; h[1] = 0xfe; h[2] = 0xfd; put16(h + 3, epoch);
            for (int b = 0; b < 6; b++) h[5 + b] = (uint8_t)(seq >> (8 * (5 - b)));
            put16(h + 11, plen);
            w += 13;
        }
        uint8_t* p = out + w;
        put16(p, src_port); put16(p + 2, dst_port); put32(p + 4, vtag);
        p[8] = p[9] = p[10] = p[11] = 0;                                  /* checksum placeholder */
        p[12] = 0;                                                        /* chunk type DATA */
        p[13] = (uint8_t)((f == 0 ? 2 : 0) | (f == k - 1 ? 1 : 0));       /* B / E flags */
        put16(p + 14, 16 + fl);                                           /* chunk length excludes padding */
        put32(p + 16, tsn0 + (uint32_t)f);
        put16(p + 20, sid); put16(p + 22, ssn); put32(p + 24, ppid);
        if (fl) memcpy(p + 28, msg + off, fl);
        for (uint32_t z = fl; z < padded; z++) p[28 + z] = 0;
        uint32_t c = orc_crc32c(p, plen);
        p[8] = (uint8_t)c; p[9] = (uint8_t)(c >> 8); p[10] = (uint8_t)(c >> 16); p[11] = (uint8_t)(c >> 24);   /* little-endian */
        w += plen;
    }
    return w;
}

int orc_sctp_unpack(const uint8_t* pkt, uint64_t len, uint32_t headroom, orc_sctp_chunk* info,
                    uint64_t* pay_off, uint32_t* pay_len) {
    memset(info, 0, sizeof *info);
    *pay_off = 0; *pay_len = 0;
    if (len < (uint64_t)headroom + 12) { info->status = ORC_SCTP_MALFORMED; return info->status; }
    const uint8_t* p = pkt + headroom;
    uint64_t n = len - headroom;
    /* checksum over the packet with the checksum field zeroed (verified first, like Packet::unmarshal) */
    uint8_t* tmp = (uint8_t*)malloc(n);
    memcpy(tmp, p, n);
    tmp[8] = tmp[9] = tmp[10] = tmp[11] = 0;
    uint32_t c = orc_crc32c(tmp, n);
    free(tmp);
    uint32_t theirs = (uint32_t)p[8] | ((uint32_t)p[9] << 8) | ((uint32_t)p[10] << 16) | ((uint32_t)p[11] << 24);
    if (c != theirs) { info->status = ORC_SCTP_BAD_CHECKSUM; return info->status; }
    if (n < 28 || p[12] != 0) { info->status = ORC_SCTP_MALFORMED; return info->status; }
    uint32_t clen = ((uint32_t)p[14] << 8) | p[15];
    if (clen < 16 || 12 + (uint64_t)((clen + 3) & ~3u) != n) { info->status = ORC_SCTP_MALFORMED; return info->status; }
    info->flags = p[13];
    info->tsn = ((uint32_t)p[16] << 24) | ((uint32_t)p[17] << 16) | ((uint32_t)p[18] << 8) | p[19];
    info->stream_id = (uint16_t)(((uint32_t)p[20] << 8) | p[21]);
    info->ssn = (uint16_t)(((uint32_t)p[22] << 8) | p[23]);
    info->ppid = ((uint32_t)p[24] << 24) | ((uint32_t)p[25] << 16) | ((uint32_t)p[26] << 8) | p[27];
    info->status = ORC_OK;
    *pay_off = (uint64_t)headroom + 28;
    *pay_len = clen - 16;
    return ORC_OK;
}

/* ------------------------------------------------------------------------------------------ */
/* DTLS record-layer bookkeeping: datagram -> records, anti-replay window. */
/* webrtc-dtls 0.10.0 (un-vendored): record_layer/mod.rs unpack_datagram, and the replaydetector */
/* crate's SlidingWindowDetector (window 64) used from conn.rs handle_incoming_packet.           */
/* ------------------------------------------------------------------------------------------ */

/* unpack_datagram: records are 13 + length bytes each; a datagram with a truncated record is
 * rejected as a whole (ErrInvalidPacketLength).  The remaining-bytes test is `<= header size` there, so a
 * trailing header with nothing behind it (a zero-length record at the end) also drops the datagram.
 * Returns the number of records, or -1. */
int64_t orc_dtls_datagram_split(const uint8_t* d, uint64_t len, uint64_t* rec_off /* cap entries */, uint64_t cap) {
    uint64_t off = 0, k = 0;
    while (off < len) {
        if (len - off <= 13) return -1;
        uint64_t plen = 13 + (((uint64_t)d[off + 11] << 8) | d[off + 12]);
        if (off + plen > len) return -1;
        if (k < cap) rec_off[k] = off;
        k++;
        off += plen;
    }
    return (int64_t)k;
}

/* SlidingWindowDetector, one (latest, mask, started) triple per session; records are visited in
 * arrival order.  check() happens before decryption, accept() only for records that authenticated
 * (authentic[i] != 0).  accept_out[i] = 1 if the record is delivered. */
void orc_replay_filter(uint64_t n, const uint32_t* session, const uint64_t* seq48, const uint8_t* authentic,
                       uint64_t* latest, uint64_t* mask, uint8_t* started, uint8_t* accept_out) {
    const uint64_t WINDOW = 64, MAXSEQ = (1ull << 48) - 1;
    for (uint64_t i = 0; i < n; i++) {
        uint32_t s = session[i];
        uint64_t q = seq48[i];
        int ok = 1;
        if (q > MAXSEQ) ok = 0;
        else if (started[s] && q <= latest[s]) {
            if (latest[s] >= WINDOW + q) ok = 0;                         /* too old */
            else if ((mask[s] >> (latest[s] - q)) & 1) ok = 0;           /* duplicate */
        }
        if (ok && authentic[i]) {                                        /* accept() */
            if (!started[s]) { started[s] = 1; latest[s] = q; mask[s] = 1; }
            else if (q > latest[s]) {
                uint64_t sh = q - latest[s];
                mask[s] = sh >= 64 ? 0 : mask[s] << sh;
                latest[s] = q;
                mask[s] |= 1;
            } else mask[s] |= 1ull << (latest[s] - q);
            accept_out[i] = 1;
        } else accept_out[i] = 0;
    }
}

/* Sequence numbers of outgoing records: webrtc-dtls 0.10.0 conn.rs (un-vendored) takes the epoch's counter with
 * fetch_add(1) for every record it marshals and fails with ErrSequenceNumberOverflow beyond 2^48 - 1.  Records are
 * visited in batch order; ok[i] = 0 (seq 0) for a record that cannot be numbered. */
void orc_dtls_seq_assign(uint64_t n, const uint32_t* session, uint32_t n_sessions, uint64_t* next_seq,
                         uint64_t* seq48, uint8_t* ok) {
    const uint64_t MAXSEQ = (1ull << 48) - 1;
    for (uint64_t i = 0; i < n; i++) {
        uint32_t s = session[i];
        seq48[i] = 0; ok[i] = 0;
        if (s >= n_sessions || next_seq[s] > MAXSEQ) continue;
        seq48[i] = next_seq[s]++;
        ok[i] = 1;
    }
}

/* ------------------------------------------------------------------------------------------ */
/* SSE scan over reassembled response bodies.  The reference forwards SSE */
/* bytes untouched (feature flag "sse", /root/reference/tunnel/src/protocol.rs:36,68); the event  */
/* framing is the upstream's: `data: {...}\n\n` ... `data: [DONE]\n\n`                            */
/* (/root/reference/tmp/mock_llm.py:54, :68-69).  An event ends after every pair of consecutive   */
/* line feeds inside one stream; done[s] says whether an event `data: [DONE]` starts an event.   */
/* ------------------------------------------------------------------------------------------ */
uint64_t orc_sse_scan(uint64_t k, const uint8_t* body, const uint64_t* body_off,
                      uint64_t* ev_csr /* k+1 */, uint64_t* ev_end /* cap */, uint64_t cap, uint8_t* done /* k */) {
    static const char DONE[] = "data: [DONE]\n\n";
    uint64_t total = 0;
    for (uint64_t s = 0; s < k; s++) {
        uint64_t a = body_off[s], b = body_off[s + 1];
        ev_csr[s] = total;
        done[s] = 0;
        uint64_t start = a;                       /* start of the current event */
        for (uint64_t i = a; i < b; i++) {
            if (i == start && b - i >= 14 && memcmp(body + i, DONE, 14) == 0) done[s] = 1;
            if (i > a && body[i] == '\n' && body[i - 1] == '\n') {
                if (total < cap) ev_end[total] = i + 1;
                total++;
                start = i + 1;
            }
        }
    }
    ev_csr[k] = total;
    return total;
}
