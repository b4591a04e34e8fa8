/*
 * p2pt_oracle.h -- CPU oracle for the tunnel frame codec + DTLS 1.2 AES-128-GCM record path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it,
 * and only as the checker or as the timed CPU baseline.  The product (libp2ptunnel_b200.so)
 * never links or calls it.
 *
 * Parity pinning:
 *   - frame codec: PINNED by the reference's own literal vectors and round-trip values
 *     (/root/reference/tunnel/src/protocol.rs:392-460, :271-388) -> tests/golden/frames_kat.json.
 *   - AES-128-GCM arithmetic: pinned by FIPS-197 App. B / C.1 and the GCM spec test cases 1-4,
 *     and cross-checked against Python `cryptography` (OpenSSL) on every golden record.
 *   - DTLS record LAYOUT (nonce / AAD / header length fix-up): it lives in the un-vendored crate
 *     webrtc-dtls 0.10.0 (/root/reference/tunnel/Cargo.lock:2822-2824), whose source is not under
 *     /root/reference and which no reference test touches.  Restated from RFC 6347 4.1 / RFC 5288 3 in
 *     orc_dtls_aad()/orc_dtls_seal()/orc_dtls_open() and PINNED to wire records captured from OpenSSL's
 *     DTLS 1.2 stack (tests/golden/gen_dtls_openssl.py -> dtls_openssl_kat.json,
 *     tests/test_dtls_openssl_pin.py): what that crate has to interoperate with.
 *   - anti-replay window: PINNED the same way, to what a live `openssl s_server -dtls1_2` accepted of a
 *     scripted sequence of replayed / reordered / stale / forged records
 *     (tests/golden/gen_replay_openssl.py -> replay_openssl_kat.json).
 *   - SCTP packet layout: "parity unpinned" (un-vendored webrtc-sctp 0.10.0, no SCTP stack in this image);
 *     CRC32c itself is pinned by the RFC 3720 B.4 vectors.
 */
#ifndef P2PT_ORACLE_H
#define P2PT_ORACLE_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* per-record status codes shared with include/p2pt.h */
enum {
    ORC_OK = 0,
    ORC_TOO_SHORT = 1,        /* protocol.rs:158-160 "message too short" */
    ORC_UNKNOWN_TYPE = 2,     /* protocol.rs:161-163 "unknown message type" */
    ORC_AUTH_FAIL = 3,        /* aes-gcm decrypt_in_place Err */
    ORC_NO_ROOM_FOR_NONCE = 4,/* webrtc-dtls ErrNotEnoughRoomForNonce */
    ORC_PASS_THROUGH = 5,     /* ChangeCipherSpec record returned unmodified */
    ORC_BAD_HEADER = 6,       /* RecordLayerHeader::unmarshal failure (short / version) */
    ORC_SCTP_BAD_CHECKSUM = 7,/* webrtc-sctp ErrChecksumMismatch */
    ORC_SCTP_MALFORMED = 8    /* shorter than the 12-byte common header, or not exactly one DATA chunk */
};

/* 24-byte frame descriptor, identical to p2pt_desc in include/p2pt.h */
typedef struct {
    uint64_t pay_off;   /* absolute byte offset of the payload inside the input buffer */
    uint32_t pay_len;
    uint32_t stream_id;
    uint8_t  msg_type;
    uint8_t  status;
    uint8_t  pad[6];
} orc_desc;

/* ---- frame codec (protocol.rs) ---- */
int  orc_msg_type_valid(uint8_t b);
void orc_frame_encode_batch(uint64_t n, const uint8_t* types, const uint32_t* sids,
                            const uint8_t* payload, const uint64_t* pay_off,
                            uint8_t* out, const uint64_t* out_off);
void orc_frame_decode_batch(uint64_t n, const uint8_t* in, const uint64_t* in_off, orc_desc* desc);
/* stable group-by (session, stream_id); returns number of streams k. Streams ascending by key.
 * session and uniq_session may be NULL (single-session batch). */
uint32_t orc_frame_demux_batch(uint64_t n, const orc_desc* desc, const uint32_t* session, const uint8_t* in,
                               uint32_t* uniq_sid, uint32_t* uniq_session, uint64_t* csr, uint32_t* perm,
                               uint8_t* gathered, uint64_t* gathered_off);

/* ---- AES-128 / GHASH / GCM (FIPS-197, SP 800-38D) ---- */
void orc_aes128_expand(const uint8_t key[16], uint8_t rk[176]);
void orc_aes128_encrypt_block(const uint8_t rk[176], const uint8_t in[16], uint8_t out[16]);
void orc_gf128_mul(const uint8_t x[16], const uint8_t y[16], uint8_t out[16]);
void orc_ghash(const uint8_t h[16], const uint8_t* aad, size_t aad_len,
               const uint8_t* c, size_t c_len, uint8_t out[16]);
void orc_gcm_encrypt(const uint8_t key[16], const uint8_t iv12[12],
                     const uint8_t* aad, size_t aad_len,
                     const uint8_t* pt, size_t len, uint8_t* ct, uint8_t tag[16]);
int  orc_gcm_decrypt(const uint8_t key[16], const uint8_t iv12[12],
                     const uint8_t* aad, size_t aad_len,
                     const uint8_t* ct, size_t len, const uint8_t tag[16], uint8_t* pt);

/* ---- DTLS 1.2 record seal / open (webrtc-dtls 0.10.0 CryptoGcm, restated) ---- */
void orc_dtls_aad(const uint8_t hdr13[13], uint16_t payload_len, uint8_t aad[13]);
/* raw = hdr13 || plaintext; out gets hdr13' || nonce8 || ct || tag16 (raw_len + 24 bytes).
 * returns status (ORC_OK or ORC_BAD_HEADER when raw_len < 13). */
int  orc_dtls_seal(const uint8_t key[16], const uint8_t iv4[4], const uint8_t nonce8[8],
                   const uint8_t* raw, uint64_t raw_len, uint8_t* out);
/* rec = sealed record; out gets hdr13 || plaintext. *out_len set on OK / PASS_THROUGH. */
int  orc_dtls_open(const uint8_t key[16], const uint8_t iv4[4],
                   const uint8_t* rec, uint64_t rec_len, uint8_t* out, uint64_t* out_len);

void orc_dtls_seal_batch(uint64_t n, const uint32_t* session,
                         const uint8_t* keys16, const uint8_t* ivs4, const uint8_t* nonce8,
                         const uint8_t* raw, const uint64_t* raw_off,
                         uint8_t* rec, const uint64_t* rec_off, uint8_t* status);
void orc_dtls_open_batch(uint64_t n, const uint32_t* session,
                         const uint8_t* keys16, const uint8_t* ivs4,
                         const uint8_t* rec, const uint64_t* rec_off,
                         uint8_t* out, const uint64_t* out_off, uint32_t* out_len, uint8_t* status);

/* ---- SCTP DATA-chunk packetise / unpack + CRC32c.  Restates webrtc-sctp 0.10.0
 *      (UN-VENDORED, /root/reference/tunnel/Cargo.lock:2910-2911; CRC via crc 3.4.0, :420-421): RFC 4960
 *      3.1 common header, 3.3.1 DATA chunk, appendix B checksum.  CRC32c itself is pinned by the RFC 3720
 *      B.4 vectors; the packet layout is "parity unpinned". ---- */
#define ORC_SCTP_MAX_PAYLOAD 1200u   /* INITIAL_MTU 1228 - 12 common header - 16 DATA chunk header */
typedef struct {
    uint32_t tsn; uint16_t stream_id; uint16_t ssn; uint32_t ppid;
    uint8_t flags;      /* bit0 E(nd), bit1 B(egin), bit2 U(nordered), bit3 I(mmediate) */
    uint8_t status; uint8_t pad[2];
} orc_sctp_chunk;
uint32_t orc_crc32c(const uint8_t* p, size_t len);
uint64_t orc_sctp_fragments(uint64_t msg_len);                 /* max(1, ceil(len / 1200)) */
uint64_t orc_sctp_packets_bytes(uint64_t msg_len);             /* total bytes of the message's packets */
/* message -> packets (one DATA chunk per packet, fragments of <= 1200 bytes, B/E flags, consecutive TSNs).
 * If dtls_hdr != 0 every packet is preceded by a 13-byte DTLS record header (type 23, fe fd, epoch,
 * seq48_first + fragment index, length = packet length).  Returns bytes written. */
uint64_t orc_sctp_packetize(const uint8_t* msg, uint64_t len, uint16_t src_port, uint16_t dst_port, uint32_t vtag,
                            uint16_t sid, uint16_t ssn, uint32_t ppid, uint32_t tsn0,
                            int dtls_hdr, uint16_t epoch, uint64_t seq48_first, uint8_t* out);
/* one packet (after `headroom` bytes) -> chunk info + payload slice; status OK / BAD_CHECKSUM / MALFORMED */
int orc_sctp_unpack(const uint8_t* pkt, uint64_t len, uint32_t headroom, orc_sctp_chunk* info,
                    uint64_t* pay_off, uint32_t* pay_len);

/* ---- DTLS record-layer bookkeeping (8(f) row 3; webrtc-dtls unpack_datagram, replaydetector) ---- */
int64_t orc_dtls_datagram_split(const uint8_t* d, uint64_t len, uint64_t* rec_off, uint64_t cap);
void orc_replay_filter(uint64_t n, const uint32_t* session, const uint64_t* seq48, const uint8_t* authentic,
                       uint64_t* latest, uint64_t* mask, uint8_t* started, uint8_t* accept_out);
void orc_dtls_seq_assign(uint64_t n, const uint32_t* session, uint32_t n_sessions, uint64_t* next_seq,
                         uint64_t* seq48, uint8_t* ok);

/* ---- SSE event scan over per-stream bodies (8(f) row 4) ---- */
uint64_t orc_sse_scan(uint64_t k, const uint8_t* body, const uint64_t* body_off,
                      uint64_t* ev_csr, uint64_t* ev_end, uint64_t cap, uint8_t* done);

#ifdef __cplusplus
}
#endif
#endif
