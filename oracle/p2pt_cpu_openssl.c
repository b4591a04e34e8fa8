/*
 * p2pt_cpu_openssl.c -- the second timed CPU baseline: the oracle's DTLS record layout around OpenSSL's
 * EVP_aes_128_gcm (AES-NI / VAES + (V)PCLMULQDQ with aggregated reduction, stitched CTR+GHASH assembly).
 * TEST / BENCH INFRASTRUCTURE ONLY, same rules as p2pt_oracle.h: never linked into or called by the
 * product.
 *
 * Why a second one: the reference seals with RustCrypto aes-gcm 0.10.3 (/root/reference/tunnel/Cargo.lock:27-29),
 * which p2pt_cpu_fast.c restates (one PCLMULQDQ multiply + reduction per block).  OpenSSL's implementation of the
 * same AEAD is several times faster per core; SURVEY.md 8(d) names it as the baseline that is conservative for the
 * GPU side, so bench.py times both and prints both.
 *
 * One EVP_CIPHER_CTX per direction per worker thread; the key is set when the worker moves to another session
 * (16 records per session in BASELINE.json configs[4]) and only the 12-byte IV per record, which is how a DTLS
 * stack holds its write state.  Layout = orc_dtls_seal / orc_dtls_open (oracle/p2pt_oracle.c), checked against
 * them byte for byte in tests/test_cpu_fast.py.
 */
#ifndef P2PT_NO_OPENSSL
#include <openssl/evp.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "p2pt_oracle.h"

typedef struct { EVP_CIPHER_CTX* enc; EVP_CIPHER_CTX* dec; } ossl_key;

int cpuossl_available(void) { return 1; }

void* cpuossl_key_new(void) {
    ossl_key* k = (ossl_key*)calloc(1, sizeof(ossl_key));
    if (!k) return NULL;
    k->enc = EVP_CIPHER_CTX_new();
    k->dec = EVP_CIPHER_CTX_new();
    return k;
}
void cpuossl_key_free(void* p) {
    ossl_key* k = (ossl_key*)p;
    if (!k) return;
    EVP_CIPHER_CTX_free(k->enc);
    EVP_CIPHER_CTX_free(k->dec);
    free(k);
}
/* new session: the seal phase keys only the encrypt context, the open phase only the decrypt context */
void cpuossl_key_set_enc(void* p, const uint8_t key[16]) {
    EVP_EncryptInit_ex(((ossl_key*)p)->enc, EVP_aes_128_gcm(), NULL, key, NULL);
}
void cpuossl_key_set_dec(void* p, const uint8_t key[16]) {
    EVP_DecryptInit_ex(((ossl_key*)p)->dec, EVP_aes_128_gcm(), NULL, key, NULL);
}

/* CryptoGcm::encrypt, layout as orc_dtls_seal */
int cpuossl_seal(void* p, const uint8_t iv4[4], const uint8_t nonce8[8], const uint8_t* raw, uint64_t raw_len, uint8_t* out) {
    ossl_key* k = (ossl_key*)p;
    if (raw_len < 13) return ORC_BAD_HEADER;
    uint64_t L = raw_len - 13;
    uint8_t iv[12], aad[13];
    int l = 0;
    memcpy(iv, iv4, 4); memcpy(iv + 4, nonce8, 8);
    orc_dtls_aad(raw, (uint16_t)L, aad);
    memcpy(out, raw, 13);
    memcpy(out + 13, nonce8, 8);
    EVP_EncryptInit_ex(k->enc, NULL, NULL, NULL, iv);
    EVP_EncryptUpdate(k->enc, NULL, &l, aad, 13);
    if (L) EVP_EncryptUpdate(k->enc, out + 21, &l, raw + 13, (int)L);
    EVP_EncryptFinal_ex(k->enc, out + 21 + L, &l);
    EVP_CIPHER_CTX_ctrl(k->enc, EVP_CTRL_GCM_GET_TAG, 16, out + 21 + L);
    uint16_t rlen = (uint16_t)(8 + L + 16);
    out[11] = (uint8_t)(rlen >> 8); out[12] = (uint8_t)rlen;
    return ORC_OK;
}

/* CryptoGcm::decrypt, layout and status codes as orc_dtls_open */
int cpuossl_open(void* p, const uint8_t iv4[4], const uint8_t* rec, uint64_t rec_len, uint8_t* out, uint64_t* out_len) {
    ossl_key* k = (ossl_key*)p;
    *out_len = 0;
    if (rec_len < 13) return ORC_BAD_HEADER;
    if (!(rec[1] == 0xfe && (rec[2] == 0xfd || rec[2] == 0xff))) return ORC_BAD_HEADER;
    if (rec[0] == 20) { memcpy(out, rec, rec_len); *out_len = rec_len; return ORC_PASS_THROUGH; }
    if (rec_len <= 21) return ORC_NO_ROOM_FOR_NONCE;
    if (rec_len < 37) return ORC_AUTH_FAIL;
    uint64_t L = rec_len - 37;
    uint8_t iv[12], aad[13], tag[16];
    int l = 0;
    memcpy(iv, iv4, 4); memcpy(iv + 4, rec + 13, 8);
    orc_dtls_aad(rec, (uint16_t)L, aad);
    memcpy(out, rec, 13);
    memcpy(tag, rec + 21 + L, 16);
    EVP_DecryptInit_ex(k->dec, NULL, NULL, NULL, iv);
    EVP_DecryptUpdate(k->dec, NULL, &l, aad, 13);
    if (L) EVP_DecryptUpdate(k->dec, out + 13, &l, rec + 21, (int)L);
    EVP_CIPHER_CTX_ctrl(k->dec, EVP_CTRL_GCM_SET_TAG, 16, tag);
    if (EVP_DecryptFinal_ex(k->dec, out + 13 + L, &l) <= 0) { memset(out, 0, 13 + L); return ORC_AUTH_FAIL; }
    *out_len = 13 + L;
    return ORC_OK;
}

#else  /* built without libcrypto: the baseline reports itself absent, nothing else changes */
#include <stdint.h>
#include <stddef.h>
int cpuossl_available(void) { return 0; }
void* cpuossl_key_new(void) { return NULL; }
void cpuossl_key_free(void* p) { (void)p; }
void cpuossl_key_set_enc(void* p, const uint8_t key[16]) { (void)p; (void)key; }
void cpuossl_key_set_dec(void* p, const uint8_t key[16]) { (void)p; (void)key; }
int cpuossl_seal(void* p, const uint8_t iv4[4], const uint8_t nonce8[8], const uint8_t* raw, uint64_t raw_len, uint8_t* out) {
    (void)p; (void)iv4; (void)nonce8; (void)raw; (void)raw_len; (void)out; return -1;
}
int cpuossl_open(void* p, const uint8_t iv4[4], const uint8_t* rec, uint64_t rec_len, uint8_t* out, uint64_t* out_len) {
    (void)p; (void)iv4; (void)rec; (void)rec_len; (void)out; (void)out_len; return -1;
}
#endif
