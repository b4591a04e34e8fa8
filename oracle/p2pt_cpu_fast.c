/*
 * p2pt_cpu_fast.c -- the timed CPU baseline ("port" of the reference's CPU path).  TEST / BENCH
 * INFRASTRUCTURE ONLY, same rules as p2pt_oracle.h: never linked into or called by the product.
 *
 * The reference seals records with RustCrypto aes-gcm 0.10.3, which at run time selects AES-NI
 * (aes 0.8.4, 8 blocks interleaved) and PCLMULQDQ (polyval 0.6.2, one multiply + reduction per
 * block) through cpufeatures 0.2.17 (/root/reference/tunnel/Cargo.lock:16,27,411,1542).  A scalar C
 * port would flatter the GPU, so this file restates that path with the same instruction classes:
 * AES-NI CTR 8 blocks at a time, PCLMULQDQ GHASH with a per-block Karatsuba multiply and
 * reduction, the webrtc-dtls record layout around it, the protocol.rs frame codec, and the
 * per-frame HashMap<u32, Vec<u8>> append of serve.rs:112-139 for the demux.  Work is sharded over
 * host threads by session (contiguous record ranges), the reference's one-task-per-connection
 * concurrency (proxy.rs:108-109).  tests/test_cpu_fast.py checks every output against the scalar
 * oracle.
 */
#include <immintrin.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "p2pt_oracle.h"

int cpufast_supported(void) {
    return __builtin_cpu_supports("aes") && __builtin_cpu_supports("pclmul") && __builtin_cpu_supports("ssse3");
}

/* ---- AES-128 with AES-NI ---- */
typedef struct { __m128i rk[11]; __m128i h; } fast_key;

#define EXPAND_STEP(i, rcon)                                                     \
    do {                                                                         \
        __m128i t = _mm_aeskeygenassist_si128(k->rk[i - 1], rcon);              \
        t = _mm_shuffle_epi32(t, 0xff);                                          \
        __m128i p = k->rk[i - 1];                                                \
        p = _mm_xor_si128(p, _mm_slli_si128(p, 4));                              \
        p = _mm_xor_si128(p, _mm_slli_si128(p, 4));                              \
        p = _mm_xor_si128(p, _mm_slli_si128(p, 4));                              \
        k->rk[i] = _mm_xor_si128(p, t);                                          \
    } while (0)

static inline __m128i aes_block(const fast_key* k, __m128i b) {
    b = _mm_xor_si128(b, k->rk[0]);
    for (int r = 1; r < 10; r++) b = _mm_aesenc_si128(b, k->rk[r]);
    return _mm_aesenclast_si128(b, k->rk[10]);
}

static inline __m128i bswap128(__m128i v) {
    const __m128i m = _mm_set_epi8(0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15);
    return _mm_shuffle_epi8(v, m);
}

static void fast_key_init(fast_key* k, const uint8_t key[16]) {
    k->rk[0] = _mm_loadu_si128((const __m128i*)key);
    EXPAND_STEP(1, 0x01); EXPAND_STEP(2, 0x02); EXPAND_STEP(3, 0x04); EXPAND_STEP(4, 0x08);
    EXPAND_STEP(5, 0x10); EXPAND_STEP(6, 0x20); EXPAND_STEP(7, 0x40); EXPAND_STEP(8, 0x80);
    EXPAND_STEP(9, 0x1b); EXPAND_STEP(10, 0x36);
    k->h = bswap128(aes_block(k, _mm_setzero_si128()));
}

/* GF(2^128) multiply, GHASH bit order, operands byte-reversed (Intel CLMUL white paper alg. 1 + 5):
 * Karatsuba 3 x PCLMULQDQ, shift left by one, reduce. */
static inline __m128i gf_mul(__m128i a, __m128i b) {
    __m128i lo = _mm_clmulepi64_si128(a, b, 0x00);
    __m128i hi = _mm_clmulepi64_si128(a, b, 0x11);
    __m128i am = _mm_xor_si128(a, _mm_srli_si128(a, 8));
    __m128i bm = _mm_xor_si128(b, _mm_srli_si128(b, 8));
    __m128i mid = _mm_clmulepi64_si128(am, bm, 0x00);
    mid = _mm_xor_si128(mid, _mm_xor_si128(lo, hi));
    lo = _mm_xor_si128(lo, _mm_slli_si128(mid, 8));
    hi = _mm_xor_si128(hi, _mm_srli_si128(mid, 8));
    /* shift the 256-bit product left by one bit */
    __m128i lo_c = _mm_srli_epi32(lo, 31), hi_c = _mm_srli_epi32(hi, 31);
    lo = _mm_slli_epi32(lo, 1); hi = _mm_slli_epi32(hi, 1);
    __m128i cross = _mm_srli_si128(lo_c, 12);
    hi_c = _mm_slli_si128(hi_c, 4); lo_c = _mm_slli_si128(lo_c, 4);
    lo = _mm_or_si128(lo, lo_c);
    hi = _mm_or_si128(_mm_or_si128(hi, hi_c), cross);
    /* reduce modulo x^128 + x^7 + x^2 + x + 1 */
    __m128i t7 = _mm_slli_epi32(lo, 31), t8 = _mm_slli_epi32(lo, 30), t9 = _mm_slli_epi32(lo, 25);
    t7 = _mm_xor_si128(_mm_xor_si128(t7, t8), t9);
    t8 = _mm_srli_si128(t7, 4);
    t7 = _mm_slli_si128(t7, 12);
    lo = _mm_xor_si128(lo, t7);
    __m128i t2 = _mm_srli_epi32(lo, 1), t3 = _mm_srli_epi32(lo, 2), t4 = _mm_srli_epi32(lo, 7);
    t2 = _mm_xor_si128(_mm_xor_si128(t2, t3), _mm_xor_si128(t4, t8));
    lo = _mm_xor_si128(lo, t2);
    return _mm_xor_si128(hi, lo);
}

static inline __m128i load_partial(const uint8_t* p, size_t n) {
    uint8_t b[16] = {0};
    memcpy(b, p, n);
    return _mm_loadu_si128((const __m128i*)b);
}

/* GCM over (aad13, data): if enc, out = data ^ keystream and GHASH runs over out; else GHASH over data. */
static void gcm_crypt(const fast_key* k, const uint8_t iv12[12], const uint8_t aad13[13],
                      const uint8_t* in, size_t len, uint8_t* out, int enc, uint8_t tag[16]) {
    uint8_t cb[16];
    memcpy(cb, iv12, 12);
    cb[12] = 0; cb[13] = 0; cb[14] = 0; cb[15] = 1;
    const __m128i j0 = _mm_loadu_si128((const __m128i*)cb);
    const __m128i ej0 = aes_block(k, j0);
    const __m128i be_j0 = bswap128(j0);
    __m128i y = gf_mul(bswap128(load_partial(aad13, 13)), k->h);
    uint32_t ctr = 2;
    size_t off = 0;
    while (len - off >= 128) {                       /* 8 blocks interleaved, like aes 0.8.4 */
        __m128i ks[8];
        for (int b = 0; b < 8; b++) {
            __m128i c = _mm_add_epi32(_mm_and_si128(be_j0, _mm_set_epi32(-1, -1, -1, 0)), _mm_set_epi32(0, 0, 0, (int)(ctr + b)));
            ks[b] = _mm_xor_si128(bswap128(c), k->rk[0]);
        }
        for (int r = 1; r < 10; r++)
            for (int b = 0; b < 8; b++) ks[b] = _mm_aesenc_si128(ks[b], k->rk[r]);
        for (int b = 0; b < 8; b++) {
            ks[b] = _mm_aesenclast_si128(ks[b], k->rk[10]);
            __m128i d = _mm_loadu_si128((const __m128i*)(in + off + 16 * b));
            __m128i o = _mm_xor_si128(d, ks[b]);
            _mm_storeu_si128((__m128i*)(out + off + 16 * b), o);
            y = gf_mul(_mm_xor_si128(y, bswap128(enc ? o : d)), k->h);
        }
        ctr += 8;
        off += 128;
    }
    while (off < len) {
        size_t n = len - off < 16 ? len - off : 16;
        __m128i c = _mm_add_epi32(_mm_and_si128(be_j0, _mm_set_epi32(-1, -1, -1, 0)), _mm_set_epi32(0, 0, 0, (int)ctr));
        __m128i ks = aes_block(k, bswap128(c));
        __m128i d = load_partial(in + off, n);
        __m128i o = _mm_xor_si128(d, ks);
        uint8_t ob[16];
        _mm_storeu_si128((__m128i*)ob, o);
        memcpy(out + off, ob, n);
        if (n < 16) memset(ob + n, 0, 16 - n);
        __m128i g = enc ? _mm_loadu_si128((const __m128i*)ob) : d;
        y = gf_mul(_mm_xor_si128(y, bswap128(g)), k->h);
        ctr++;
        off += n;
    }
    __m128i lb = _mm_set_epi64x((long long)(13 * 8), (long long)((uint64_t)len * 8));
    y = gf_mul(_mm_xor_si128(y, lb), k->h);
    _mm_storeu_si128((__m128i*)tag, _mm_xor_si128(bswap128(y), ej0));
}

/* webrtc-dtls CryptoGcm::encrypt around the fast AEAD (same layout as orc_dtls_seal) */
static int fast_seal(const fast_key* k, const uint8_t iv4[4], const uint8_t nonce8[8],
                     const uint8_t* raw, uint64_t raw_len, uint8_t* out) {
    if (raw_len < 13) return ORC_BAD_HEADER;
    uint64_t L = raw_len - 13;
    uint8_t iv[12], aad[13];
    memcpy(iv, iv4, 4); memcpy(iv + 4, nonce8, 8);
    orc_dtls_aad(raw, (uint16_t)L, aad);
    memcpy(out, raw, 13);
    memcpy(out + 13, nonce8, 8);
    gcm_crypt(k, iv, aad, raw + 13, (size_t)L, out + 21, 1, out + 21 + L);
    uint16_t rlen = (uint16_t)(8 + L + 16);
    out[11] = (uint8_t)(rlen >> 8); out[12] = (uint8_t)rlen;
    return ORC_OK;
}

static int fast_open(const fast_key* k, const uint8_t iv4[4], const uint8_t* rec, uint64_t rec_len,
                     uint8_t* out, uint64_t* out_len) {
    *out_len = 0;
    if (rec_len < 13) return ORC_BAD_HEADER;
    if (!(rec[1] == 0xfe && (rec[2] == 0xfd || rec[2] == 0xff))) return ORC_BAD_HEADER;
    if (rec[0] == 20) { memcpy(out, rec, rec_len); *out_len = rec_len; return ORC_PASS_THROUGH; }
    if (rec_len <= 21) return ORC_NO_ROOM_FOR_NONCE;
    if (rec_len < 37) return ORC_AUTH_FAIL;
    uint64_t L = rec_len - 37;
    uint8_t iv[12], aad[13], tag[16];
    memcpy(iv, iv4, 4); memcpy(iv + 4, rec + 13, 8);
    orc_dtls_aad(rec, (uint16_t)L, aad);
    memcpy(out, rec, 13);
    gcm_crypt(k, iv, aad, rec + 21, (size_t)L, out + 13, 0, tag);
    uint8_t diff = 0;
    for (int i = 0; i < 16; i++) diff |= (uint8_t)(tag[i] ^ rec[21 + L + i]);
    if (diff) { memset(out, 0, 13 + L); return ORC_AUTH_FAIL; }
    *out_len = 13 + L;
    return ORC_OK;
}

/* ---- per-stream reassembly map: HashMap<(session, sid), Vec<u8>> (serve.rs:65, :120-124) ---- */
typedef struct { uint64_t key; uint8_t* buf; uint64_t len, cap; int used; } stream_slot;
typedef struct { stream_slot* slots; uint64_t nslots, count; } stream_map;

static void map_init(stream_map* m, uint64_t cap) {
    m->nslots = 64; while (m->nslots < cap * 2) m->nslots *= 2;
    m->slots = (stream_slot*)calloc(m->nslots, sizeof(stream_slot));
    m->count = 0;
}
static stream_slot* map_probe(stream_slot* slots, uint64_t nslots, uint64_t key) {
    uint64_t h = key * 0x9E3779B97F4A7C15ull;
    for (uint64_t i = h >> 20;; i++) {
        stream_slot* s = &slots[i & (nslots - 1)];
        if (!s->used || s->key == key) return s;
    }
}
static stream_slot* map_get(stream_map* m, uint64_t key) {
    if ((m->count + 1) * 2 > m->nslots) {            /* grow like HashMap does */
        uint64_t nn = m->nslots * 2;
        stream_slot* ns = (stream_slot*)calloc(nn, sizeof(stream_slot));
        for (uint64_t q = 0; q < m->nslots; q++)
            if (m->slots[q].used) *map_probe(ns, nn, m->slots[q].key) = m->slots[q];
        free(m->slots);
        m->slots = ns; m->nslots = nn;
    }
    stream_slot* s = map_probe(m->slots, m->nslots, key);
    if (!s->used) { s->used = 1; s->key = key; m->count++; }
    return s;
}
static void map_append(stream_slot* s, const uint8_t* p, uint64_t n) {
    if (s->len + n > s->cap) {
        uint64_t c = s->cap ? s->cap * 2 : 4096;
        while (c < s->len + n) c *= 2;
        s->buf = (uint8_t*)realloc(s->buf, c);
        s->cap = c;
    }
    memcpy(s->buf + s->len, p, n);
    s->len += n;
}
static void map_free(stream_map* m) {
    for (uint64_t i = 0; i < m->nslots; i++) free(m->slots[i].buf);
    free(m->slots);
}

/* second AEAD backend: OpenSSL EVP (p2pt_cpu_openssl.c) */
int cpuossl_available(void);
void* cpuossl_key_new(void);
void cpuossl_key_free(void* p);
void cpuossl_key_set_enc(void* p, const uint8_t key[16]);
void cpuossl_key_set_dec(void* p, const uint8_t key[16]);
int cpuossl_seal(void* p, const uint8_t iv4[4], const uint8_t nonce8[8], const uint8_t* raw, uint64_t raw_len, uint8_t* out);
int cpuossl_open(void* p, const uint8_t iv4[4], const uint8_t* rec, uint64_t rec_len, uint8_t* out, uint64_t* out_len);

/* ---- the threaded pipeline ---- */
typedef struct {
    /* shared inputs */
    uint64_t n; const uint32_t* session; const uint8_t* keys16; const uint8_t* ivs4; uint32_t n_sessions;
    const uint16_t* epoch; const uint64_t* seq48; const uint8_t* nonce8; const uint8_t* types; const uint32_t* sids;
    const uint8_t* payload; const uint64_t* pay_off; uint8_t* rec; const uint64_t* rec_off; uint8_t* out;
    int do_seal, do_open, no_demux;
    int backend;   /* 0 = the RustCrypto-style port above, 1 = OpenSSL EVP_aes_128_gcm */
    /* per thread */
    uint64_t r0, r1; uint64_t bad; uint64_t streams; uint64_t body_bytes; uint64_t checksum;
} job;

static void* worker(void* arg) {
    job* j = (job*)arg;
    fast_key fk; uint32_t cur = 0xffffffffu;
    uint8_t* scratch = (uint8_t*)malloc(70000);
    void* ok = j->backend ? cpuossl_key_new() : NULL;
    if (j->do_seal) {
        for (uint64_t i = j->r0; i < j->r1; i++) {
            uint32_t s = j->session[i];
            if (s != cur) {
                if (j->backend) cpuossl_key_set_enc(ok, j->keys16 + 16 * (uint64_t)s);
                else fast_key_init(&fk, j->keys16 + 16 * (uint64_t)s);
                cur = s;
            }
            /* TunnelMessage::encode (protocol.rs:148-154) into raw = hdr13 || frame */
            uint64_t plen = j->pay_off[i + 1] - j->pay_off[i], flen = plen + 5;
            uint8_t* raw = scratch;
            raw[0] = 23; raw[1] = 0xfe; raw[2] = 0xfd; raw[3] = (uint8_t)(j->epoch[i] >> 8); raw[4] = (uint8_t)j->epoch[i];
            for (int b = 0; b < 6; b++) raw[5 + b] = (uint8_t)(j->seq48[i] >> (8 * (5 - b)));
            raw[11] = (uint8_t)(flen >> 8); raw[12] = (uint8_t)flen;
            raw[13] = j->types[i];
            raw[14] = (uint8_t)(j->sids[i] >> 24); raw[15] = (uint8_t)(j->sids[i] >> 16);
            raw[16] = (uint8_t)(j->sids[i] >> 8); raw[17] = (uint8_t)j->sids[i];
            memcpy(raw + 18, j->payload + j->pay_off[i], plen);
            const int st = j->backend ? cpuossl_seal(ok, j->ivs4 + 4 * (uint64_t)s, j->nonce8 + 8 * i, raw, 13 + flen, j->rec + j->rec_off[i])
                                      : fast_seal(&fk, j->ivs4 + 4 * (uint64_t)s, j->nonce8 + 8 * i, raw, 13 + flen, j->rec + j->rec_off[i]);
            if (st) j->bad++;
        }
    }
    if (j->do_open) {
        stream_map m; map_init(&m, 1024);
        cur = 0xffffffffu;
        for (uint64_t i = j->r0; i < j->r1; i++) {
            uint32_t s = j->session[i];
            if (s != cur) {
                if (j->backend) cpuossl_key_set_dec(ok, j->keys16 + 16 * (uint64_t)s);
                else fast_key_init(&fk, j->keys16 + 16 * (uint64_t)s);
                cur = s;
            }
            uint64_t olen = 0;
            uint8_t* o = j->out ? j->out + j->rec_off[i] : scratch;
            int st = j->backend ? cpuossl_open(ok, j->ivs4 + 4 * (uint64_t)s, j->rec + j->rec_off[i], j->rec_off[i + 1] - j->rec_off[i], o, &olen)
                                : fast_open(&fk, j->ivs4 + 4 * (uint64_t)s, j->rec + j->rec_off[i], j->rec_off[i + 1] - j->rec_off[i], o, &olen);
            if (st != ORC_OK) { j->bad++; continue; }
            /* TunnelMessage::decode (protocol.rs:157-172) + demux append (serve.rs:120-124) */
            uint64_t flen = olen - 13; const uint8_t* f = o + 13;
            if (flen < 5 || !orc_msg_type_valid(f[0])) { j->bad++; continue; }
            uint32_t sid = ((uint32_t)f[1] << 24) | ((uint32_t)f[2] << 16) | ((uint32_t)f[3] << 8) | f[4];
            if (j->no_demux) { if (f[0] == 11 || f[0] == 21) j->body_bytes += flen - 5; continue; }
            stream_slot* sl = map_get(&m, ((uint64_t)s << 32) | sid);
            if (f[0] == 11 || f[0] == 21) map_append(sl, f + 5, flen - 5);
        }
        j->streams = m.count;
        for (uint64_t q = 0; q < m.nslots; q++)
            if (m.slots[q].used) {
                j->body_bytes += m.slots[q].len;
                if (m.slots[q].len) j->checksum += m.slots[q].buf[0] + m.slots[q].buf[m.slots[q].len - 1];
            }
        map_free(&m);
    }
    free(scratch);
    if (ok) cpuossl_key_free(ok);
    return NULL;
}

static double now_s(void) {
    struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

/* Runs encode->seal over all records, then open->decode->demux, on nthreads threads sharded by
 * contiguous record ranges cut at session boundaries.  times[0] = seal seconds, times[1] = open
 * seconds.  stats[0] = bad records, [1] = streams, [2] = body bytes reassembled, [3] = checksum.
 * phases: bit 0 = seal, bit 1 = open + decode + demux append, bit 2 = open + decode WITHOUT the per-stream map
 * (the AEAD alone; its seconds go to times[2], so only callers that set bit 2 need a third slot; stats then
 * describe that last pass). */
int cpufast_pipeline_ex(uint64_t n, const uint32_t* session, const uint8_t* keys16, const uint8_t* ivs4, uint32_t n_sessions,
                        const uint16_t* epoch, const uint64_t* seq48, const uint8_t* nonce8,
                        const uint8_t* types, const uint32_t* sids, const uint8_t* payload, const uint64_t* pay_off,
                        uint8_t* rec, const uint64_t* rec_off, uint8_t* out, int nthreads, int phases,
                        double* times, uint64_t* stats, int backend) {
    if (backend == 0 && !cpufast_supported()) return -1;
    if (backend == 1 && !cpuossl_available()) return -2;
    if (nthreads < 1) nthreads = 1;
    job* jobs = (job*)calloc((size_t)nthreads, sizeof(job));
    pthread_t* th = (pthread_t*)calloc((size_t)nthreads, sizeof(pthread_t));
    uint64_t prev = 0;
    for (int t = 0; t < nthreads; t++) {
        uint64_t end = (t == nthreads - 1) ? n : n * (uint64_t)(t + 1) / (uint64_t)nthreads;
        while (end < n && end > 0 && session[end] == session[end - 1]) end++;   /* keep a session on one thread */
        if (end < prev) end = prev;
        jobs[t].n = n; jobs[t].session = session; jobs[t].keys16 = keys16; jobs[t].ivs4 = ivs4; jobs[t].n_sessions = n_sessions;
        jobs[t].epoch = epoch; jobs[t].seq48 = seq48; jobs[t].nonce8 = nonce8; jobs[t].types = types; jobs[t].sids = sids;
        jobs[t].payload = payload; jobs[t].pay_off = pay_off; jobs[t].rec = rec; jobs[t].rec_off = rec_off; jobs[t].out = out;
        jobs[t].r0 = prev; jobs[t].r1 = end; jobs[t].backend = backend;
        prev = end;
    }
    memset(stats, 0, 4 * sizeof(uint64_t));
    times[0] = times[1] = 0;
    for (int phase = 0; phase < 3; phase++) {
        if (!(phases & (1 << phase))) continue;
        for (int t = 0; t < nthreads; t++) {
            jobs[t].do_seal = (phase == 0); jobs[t].do_open = (phase >= 1); jobs[t].no_demux = (phase == 2);
            jobs[t].streams = jobs[t].body_bytes = jobs[t].checksum = 0;      /* bad records add up over the phases */
        }
        double t0 = now_s();
        for (int t = 0; t < nthreads; t++) pthread_create(&th[t], NULL, worker, &jobs[t]);
        for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
        times[phase] = now_s() - t0;
    }
    for (int t = 0; t < nthreads; t++) {
        stats[0] += jobs[t].bad; stats[1] += jobs[t].streams; stats[2] += jobs[t].body_bytes; stats[3] += jobs[t].checksum;
    }
    free(jobs); free(th);
    return 0;
}

int cpufast_pipeline(uint64_t n, const uint32_t* session, const uint8_t* keys16, const uint8_t* ivs4, uint32_t n_sessions,
                     const uint16_t* epoch, const uint64_t* seq48, const uint8_t* nonce8,
                     const uint8_t* types, const uint32_t* sids, const uint8_t* payload, const uint64_t* pay_off,
                     uint8_t* rec, const uint64_t* rec_off, uint8_t* out, int nthreads, int phases,
                     double* times, uint64_t* stats) {
    return cpufast_pipeline_ex(n, session, keys16, ivs4, n_sessions, epoch, seq48, nonce8, types, sids, payload, pay_off,
                               rec, rec_off, out, nthreads, phases, times, stats, 0);
}

/* batch seal / open through OpenSSL EVP (parity tests against the scalar oracle) */
void cpuossl_seal_batch(uint64_t n, const uint32_t* session, const uint8_t* keys16, const uint8_t* ivs4,
                        const uint8_t* nonce8, const uint8_t* raw, const uint64_t* raw_off,
                        uint8_t* rec, const uint64_t* rec_off, uint8_t* status) {
    void* k = cpuossl_key_new();
    for (uint64_t i = 0; i < n; i++) {
        cpuossl_key_set_enc(k, keys16 + 16 * (uint64_t)session[i]);
        status[i] = (uint8_t)cpuossl_seal(k, ivs4 + 4 * (uint64_t)session[i], nonce8 + 8 * i, raw + raw_off[i],
                                          raw_off[i + 1] - raw_off[i], rec + rec_off[i]);
    }
    cpuossl_key_free(k);
}
void cpuossl_open_batch(uint64_t n, const uint32_t* session, const uint8_t* keys16, const uint8_t* ivs4,
                        const uint8_t* rec, const uint64_t* rec_off, uint8_t* out, const uint64_t* out_off,
                        uint32_t* out_len, uint8_t* status) {
    void* k = cpuossl_key_new();
    for (uint64_t i = 0; i < n; i++) {
        uint64_t ol = 0;
        cpuossl_key_set_dec(k, keys16 + 16 * (uint64_t)session[i]);
        status[i] = (uint8_t)cpuossl_open(k, ivs4 + 4 * (uint64_t)session[i], rec + rec_off[i],
                                          rec_off[i + 1] - rec_off[i], out + out_off[i], &ol);
        out_len[i] = (uint32_t)ol;
    }
    cpuossl_key_free(k);
}

/* plain batch seal / open with the fast path (for parity tests against the scalar oracle) */
void cpufast_seal_batch(uint64_t n, const uint32_t* session, const uint8_t* keys16, const uint8_t* ivs4,
                        const uint8_t* nonce8, const uint8_t* raw, const uint64_t* raw_off,
                        uint8_t* rec, const uint64_t* rec_off, uint8_t* status) {
    fast_key fk;
    for (uint64_t i = 0; i < n; i++) {
        fast_key_init(&fk, keys16 + 16 * (uint64_t)session[i]);
        status[i] = (uint8_t)fast_seal(&fk, ivs4 + 4 * (uint64_t)session[i], nonce8 + 8 * i, raw + raw_off[i],
                                       raw_off[i + 1] - raw_off[i], rec + rec_off[i]);
    }
}
void cpufast_open_batch(uint64_t n, const uint32_t* session, const uint8_t* keys16, const uint8_t* ivs4,
                        const uint8_t* rec, const uint64_t* rec_off, uint8_t* out, const uint64_t* out_off,
                        uint32_t* out_len, uint8_t* status) {
    fast_key fk;
    for (uint64_t i = 0; i < n; i++) {
        uint64_t ol = 0;
        fast_key_init(&fk, keys16 + 16 * (uint64_t)session[i]);
        status[i] = (uint8_t)fast_open(&fk, ivs4 + 4 * (uint64_t)session[i], rec + rec_off[i],
                                       rec_off[i + 1] - rec_off[i], out + out_off[i], &ol);
        out_len[i] = (uint32_t)ol;
    }
}

/* ---- frame codec baseline (protocol.rs:148-172 + the reader loop's map append, serve.rs:112-139):
 * nthreads threads each take a contiguous range of frames, as if each were one connection's
 * reader task (proxy.rs:108-109).  phase 0 = encode, phase 1 = decode + demux append. ---- */
typedef struct {
    uint64_t r0, r1; const uint8_t* types; const uint32_t* sids; const uint8_t* payload; const uint64_t* pay_off;
    uint8_t* frames; const uint64_t* frame_off; int phase; uint64_t bad, streams, body_bytes;
} codec_job;

static void* codec_worker(void* arg) {
    codec_job* j = (codec_job*)arg;
    if (j->phase == 0) {
        for (uint64_t i = j->r0; i < j->r1; i++) {
            uint8_t* o = j->frames + j->frame_off[i];
            uint64_t len = j->pay_off[i + 1] - j->pay_off[i];
            o[0] = j->types[i];
            o[1] = (uint8_t)(j->sids[i] >> 24); o[2] = (uint8_t)(j->sids[i] >> 16);
            o[3] = (uint8_t)(j->sids[i] >> 8); o[4] = (uint8_t)j->sids[i];
            memcpy(o + 5, j->payload + j->pay_off[i], len);
        }
    } else {
        stream_map m; map_init(&m, 1024);
        for (uint64_t i = j->r0; i < j->r1; i++) {
            const uint8_t* f = j->frames + j->frame_off[i];
            uint64_t len = j->frame_off[i + 1] - j->frame_off[i];
            if (len < 5 || !orc_msg_type_valid(f[0])) { j->bad++; continue; }
            uint32_t sid = ((uint32_t)f[1] << 24) | ((uint32_t)f[2] << 16) | ((uint32_t)f[3] << 8) | f[4];
            stream_slot* sl = map_get(&m, sid);
            if (f[0] == 11 || f[0] == 21) map_append(sl, f + 5, len - 5);
        }
        j->streams = m.count;
        for (uint64_t q = 0; q < m.nslots; q++) if (m.slots[q].used) j->body_bytes += m.slots[q].len;
        map_free(&m);
    }
    return NULL;
}

int cpufast_codec(uint64_t n, const uint8_t* types, const uint32_t* sids, const uint8_t* payload, const uint64_t* pay_off,
                  uint8_t* frames, const uint64_t* frame_off, int nthreads, double* times, uint64_t* stats) {
    if (nthreads < 1) nthreads = 1;
    codec_job* jobs = (codec_job*)calloc((size_t)nthreads, sizeof(codec_job));
    pthread_t* th = (pthread_t*)calloc((size_t)nthreads, sizeof(pthread_t));
    memset(stats, 0, 3 * sizeof(uint64_t));
    for (int phase = 0; phase < 2; phase++) {
        for (int t = 0; t < nthreads; t++) {
            jobs[t].r0 = n * (uint64_t)t / (uint64_t)nthreads; jobs[t].r1 = n * (uint64_t)(t + 1) / (uint64_t)nthreads;
            jobs[t].types = types; jobs[t].sids = sids; jobs[t].payload = payload; jobs[t].pay_off = pay_off;
            jobs[t].frames = frames; jobs[t].frame_off = frame_off; jobs[t].phase = phase;
            jobs[t].bad = jobs[t].streams = jobs[t].body_bytes = 0;
        }
        double t0 = now_s();
        for (int t = 0; t < nthreads; t++) pthread_create(&th[t], NULL, codec_worker, &jobs[t]);
        for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
        times[phase] = now_s() - t0;
        if (phase == 1)
            for (int t = 0; t < nthreads; t++) { stats[0] += jobs[t].bad; stats[1] += jobs[t].streams; stats[2] += jobs[t].body_bytes; }
    }
    free(jobs); free(th);
    return 0;
}

/* ---- SCTP packetise + unpack/verify baseline: table-driven CRC32c, slice-by-8 (the `crc` crate the
 * reference pulls in, crc 3.4.0, is table-driven software too -- no SSE4.2 crc32 instruction). ---- */
static uint32_t g_c8[8][256];
static volatile int g_c8_ready = 0;
static void c8_init(void) {
    if (g_c8_ready) return;
    for (uint32_t i = 0; i < 256; i++) {
        uint32_t c = i;
        for (int b = 0; b < 8; b++) c = (c >> 1) ^ (0x82F63B78u & (0u - (c & 1u)));
        g_c8[0][i] = c;
    }
    for (int k = 1; k < 8; k++)
        for (uint32_t i = 0; i < 256; i++) g_c8[k][i] = (g_c8[k - 1][i] >> 8) ^ g_c8[0][g_c8[k - 1][i] & 0xff];
    g_c8_ready = 1;
}
uint32_t cpufast_crc32c(const uint8_t* p, size_t n) {
    c8_init();
    uint32_t c = 0xFFFFFFFFu;
    while (n >= 8) {
        uint32_t a, b;
        memcpy(&a, p, 4); memcpy(&b, p + 4, 4);
        a ^= c;
        c = g_c8[7][a & 0xff] ^ g_c8[6][(a >> 8) & 0xff] ^ g_c8[5][(a >> 16) & 0xff] ^ g_c8[4][a >> 24] ^
            g_c8[3][b & 0xff] ^ g_c8[2][(b >> 8) & 0xff] ^ g_c8[1][(b >> 16) & 0xff] ^ g_c8[0][b >> 24];
        p += 8; n -= 8;
    }
    while (n--) c = (c >> 8) ^ g_c8[0][(c ^ *p++) & 0xff];
    return ~c;
}

typedef struct { uint64_t r0, r1; const uint8_t* msg; const uint64_t* msg_off; uint8_t* pk; const uint64_t* pk_off; int phase; uint64_t bad, bytes; } sctp_job;

static void* sctp_worker(void* arg) {
    sctp_job* j = (sctp_job*)arg;
    for (uint64_t i = j->r0; i < j->r1; i++) {
        uint8_t* p = j->pk + j->pk_off[i];
        uint64_t plen = j->pk_off[i + 1] - j->pk_off[i];
        if (j->phase == 0) {          /* one fragment per message here (messages <= 1200 bytes) */
            uint64_t fl = j->msg_off[i + 1] - j->msg_off[i];
            p[0] = 0x13; p[1] = 0x88; p[2] = 0x13; p[3] = 0x88; p[4] = 0; p[5] = 0; p[6] = 0; p[7] = 1;
            p[8] = p[9] = p[10] = p[11] = 0;
            p[12] = 0; p[13] = 3; p[14] = (uint8_t)((16 + fl) >> 8); p[15] = (uint8_t)(16 + fl);
            p[16] = (uint8_t)(i >> 24); p[17] = (uint8_t)(i >> 16); p[18] = (uint8_t)(i >> 8); p[19] = (uint8_t)i;
            p[20] = 0; p[21] = 1; p[22] = (uint8_t)(i >> 8); p[23] = (uint8_t)i; p[24] = 0; p[25] = 0; p[26] = 0; p[27] = 53;
            memcpy(p + 28, j->msg + j->msg_off[i], fl);
            for (uint64_t z = 28 + fl; z < plen; z++) p[z] = 0;
            uint32_t c = cpufast_crc32c(p, plen);
            p[8] = (uint8_t)c; p[9] = (uint8_t)(c >> 8); p[10] = (uint8_t)(c >> 16); p[11] = (uint8_t)(c >> 24);
        } else {
            uint32_t theirs = (uint32_t)p[8] | ((uint32_t)p[9] << 8) | ((uint32_t)p[10] << 16) | ((uint32_t)p[11] << 24);
            uint8_t tmp[1300];
            memcpy(tmp, p, plen); tmp[8] = tmp[9] = tmp[10] = tmp[11] = 0;
            if (cpufast_crc32c(tmp, plen) != theirs) j->bad++;
            j->bytes += ((uint32_t)p[14] << 8 | p[15]) - 16;
        }
    }
    return NULL;
}

int cpufast_sctp(uint64_t n, const uint8_t* msg, const uint64_t* msg_off, uint8_t* pk, const uint64_t* pk_off,
                 int nthreads, double* times, uint64_t* stats) {
    c8_init();
    if (nthreads < 1) nthreads = 1;
    sctp_job* jobs = (sctp_job*)calloc((size_t)nthreads, sizeof(sctp_job));
    pthread_t* th = (pthread_t*)calloc((size_t)nthreads, sizeof(pthread_t));
    stats[0] = stats[1] = 0;
    for (int phase = 0; phase < 2; phase++) {
        for (int t = 0; t < nthreads; t++) {
            jobs[t].r0 = n * (uint64_t)t / (uint64_t)nthreads; jobs[t].r1 = n * (uint64_t)(t + 1) / (uint64_t)nthreads;
            jobs[t].msg = msg; jobs[t].msg_off = msg_off; jobs[t].pk = pk; jobs[t].pk_off = pk_off; jobs[t].phase = phase;
            jobs[t].bad = jobs[t].bytes = 0;
        }
        double t0 = now_s();
        for (int t = 0; t < nthreads; t++) pthread_create(&th[t], NULL, sctp_worker, &jobs[t]);
        for (int t = 0; t < nthreads; t++) pthread_join(th[t], NULL);
        times[phase] = now_s() - t0;
        if (phase == 1) for (int t = 0; t < nthreads; t++) { stats[0] += jobs[t].bad; stats[1] += jobs[t].bytes; }
    }
    free(jobs); free(th);
    return 0;
}
