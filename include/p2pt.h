/*
 * p2pt.h -- C ABI of libp2ptunnel_b200.so: the batched, B200-resident form of the one byte path of
 * michaelneale/p2p-llm-tunnel that every tunnelled response byte crosses:
 *
 *   TunnelMessage::encode / decode      /root/reference/tunnel/src/protocol.rs:148-154, :157-172
 *   MessageType::from_u8                /root/reference/tunnel/src/protocol.rs:103-118
 *   demux by stream_id                  /root/reference/tunnel/src/proxy.rs:105-172, serve.rs:103-139
 *   CryptoGcm::encrypt / decrypt        webrtc-dtls 0.10.0 (un-vendored; Cargo.lock:2822-2824),
 *                                       reached at every dc.send (serve.rs:274) / on_message (rtc.rs:88)
 *
 * Conventions
 *   - Plain pointers and sizes only.  The caller owns every buffer it passes; the library keeps no
 *     pointer past return.  A batch of n variable-length items is (base pointer, u64 off[n+1]):
 *     item i occupies [off[i], off[i+1]) -- frames carry no length field (protocol.rs:157-172).
 *     Offsets must be non-decreasing.
 *   - Return value: 0 = the call ran; negative = the whole call failed (bad argument, CUDA error,
 *     capacity).  p2pt_last_error() gives the text.  One bad item never fails a batch: it gets a
 *     per-item status byte, mirroring the reference's warn!+continue (serve.rs:103-109,
 *     proxy.rs:110-116).
 *   - Entry points without a suffix take HOST pointers and do host->device copy, kernels,
 *     device->host copy, synchronously.  *_device entry points take DEVICE pointers plus a
 *     cudaStream_t (as void*) and only enqueue kernels; they allocate nothing.  Work is ordered on `stream`
 *     as usual; the demux calls may run part of it on a stream of the library's own that is forked off
 *     `stream` and joined back into it before the call returns (two events; no host synchronisation).
 *   - Re-entrant: any thread may call on any ctx, and up to eight host-pointer calls run on one ctx at
 *     the same time (each takes one of the ctx's call lanes - streams, staging arena, mapped buffer -
 *     for its duration; a ninth caller waits for a lane).  This is the reference's calling pattern:
 *     many tokio tasks encode and send over one connection (main.rs:18, serve.rs:131).  One ctx per
 *     GPU.  There is no CPU fallback: without a CUDA device p2pt_ctx_create fails.
 */
#ifndef P2PT_H
#define P2PT_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define P2PT_ABI_VERSION 2

/* whole-call return codes */
#define P2PT_OK            0
#define P2PT_ERR_ARG      -1
#define P2PT_ERR_CUDA     -2
#define P2PT_ERR_CAPACITY -3   /* an output array the caller sized is too small */
#define P2PT_ERR_STATE    -4   /* e.g. seal/open before p2pt_sessions_load */

/* per-item status bytes */
#define P2PT_ST_OK                0
#define P2PT_ST_TOO_SHORT         1  /* protocol.rs:158-160 "message too short: {len} bytes" */
#define P2PT_ST_UNKNOWN_TYPE      2  /* protocol.rs:161-163 "unknown message type: {b}" */
#define P2PT_ST_AUTH_FAIL         3  /* AEAD tag mismatch (aes-gcm decrypt_in_place Err) */
#define P2PT_ST_NO_ROOM_FOR_NONCE 4  /* webrtc-dtls ErrNotEnoughRoomForNonce: len <= 13 + 8 */
#define P2PT_ST_PASS_THROUGH      5  /* ChangeCipherSpec (type 20) record copied unmodified */
#define P2PT_ST_BAD_HEADER        6  /* RecordLayerHeader::unmarshal failure: < 13 bytes or version */
#define P2PT_ST_SCTP_BAD_CHECKSUM 7  /* webrtc-sctp ErrChecksumMismatch (CRC32c) */
#define P2PT_ST_SCTP_MALFORMED    8  /* not a packet with exactly one DATA chunk of <= 1200 bytes */
#define P2PT_ST_SCTP_INCOMPLETE   9  /* message: fragments missing between B and E in this batch */

/* protocol.rs:9-12 */
#define P2PT_MAX_FRAME_SIZE 65536u
#define P2PT_MAX_BODY_CHUNK 65408u
#define P2PT_FRAME_HEADER   5u
/* DTLS record geometry (RFC 6347 4.1, RFC 5288 3) */
#define P2PT_REC_HEADER     13u
#define P2PT_REC_NONCE      8u
#define P2PT_REC_TAG        16u
#define P2PT_REC_OVERHEAD   24u   /* sealed = raw + 8 + 16 */

/* SCTP geometry (webrtc-sctp 0.10.0: INITIAL_MTU 1228 - 12 common header - 16 DATA chunk header) */
#define P2PT_SCTP_MAX_PAYLOAD 1200u
#define P2PT_SCTP_OVERHEAD    28u

typedef struct p2pt_ctx p2pt_ctx;

/* Result of decoding one frame: the batched TunnelMessage (protocol.rs:139-144).  payload is the
 * zero-copy slice [pay_off, pay_off + pay_len) of the input buffer, like Bytes::advance(5). */
typedef struct {
    uint64_t pay_off;
    uint32_t pay_len;
    uint32_t stream_id;
    uint8_t  msg_type;
    uint8_t  status;     /* P2PT_ST_OK / TOO_SHORT / UNKNOWN_TYPE */
    uint8_t  pad[6];
} p2pt_desc;             /* 24 bytes */

/* One parsed SCTP DATA chunk (RFC 4960 3.3.1).  flags: bit0 E(nd), bit1 B(egin), bit2 U, bit3 I. */
typedef struct {
    uint32_t tsn;
    uint16_t stream_id;
    uint16_t ssn;
    uint32_t ppid;
    uint8_t  flags;
    uint8_t  status;     /* P2PT_ST_OK / SCTP_BAD_CHECKSUM / SCTP_MALFORMED */
    uint8_t  pad[2];
} p2pt_sctp_chunk;       /* 16 bytes */

/* ---- context ---------------------------------------------------------------------------- */
int32_t     p2pt_abi_version(void);
int32_t     p2pt_ctx_create(int device, p2pt_ctx** out);
int32_t     p2pt_ctx_destroy(p2pt_ctx* ctx);
const char* p2pt_last_error(p2pt_ctx* ctx);            /* ctx may be NULL: last create error */
/* pinned host memory for callers that want full PCIe speed on the host-pointer entry points */
void*       p2pt_host_alloc(size_t bytes);
void        p2pt_host_free(void* p);
/* number of kernels this ctx has launched so far (bench.py's gpu_launches evidence) */
uint64_t    p2pt_launch_count(p2pt_ctx* ctx);
/* tuning knobs.  "gcm_group_width": lanes that share one record in the seal/open kernels
 * (4, 8, 16 or 32; 0 = library default).  "chunk_bytes": size of the pieces in which host-pointer
 * calls overlap host->device copies, kernels and device->host copies (default 16 MiB).  "blocking_sync":
 * 1 (default) = host-pointer calls sleep while they wait for the GPU, 0 = they spin (lowest latency for
 * single small frames, but a spinning thread per call starves other workers on a busy host).
 * "small_call_bytes" (default 262144, 0 = off, at most 1048576): p2pt_frame_encode_batch,
 * p2pt_frame_decode_batch, the two dtls_gcm calls and p2pt_pipeline_seal / _open whose staged inputs and
 * outputs fit this many bytes - the reference's per-frame call sites, batches of one or a few - run out
 * of a mapped pinned buffer: one kernel launch and one (spinning) stream wait, no copies; about half the
 * latency of the copy path for a single 64-byte frame.  "small_demux" (default 1): demux calls of up to
 * 1024 frames run init + sort + scan + emit in one single-block kernel (two launches with the gather
 * instead of six and a cooperative launch).  "direct_demux" (default 1): demux calls with max_streams <= 8192
 * try the sort-free path first (see p2pt_frame_decode_demux_batch).  Host-pointer calls that stage less than 8 MiB always wait
 * by spinning, whatever "blocking_sync" says: a sleeping wait costs 50-100 us per wake-up. */
int32_t     p2pt_ctx_set_option(p2pt_ctx* ctx, const char* name, int64_t value);
/* read-only facts: "sm_count", "lut_slack", "gcm_group_width", "n_sessions", "gcm_threads"; -1 if unknown */
int64_t     p2pt_ctx_get_info(p2pt_ctx* ctx, const char* name);

/* ---- DTLS session keys: CryptoGcm::new(local_key, local_write_iv, remote_key, remote_write_iv)
 *      (webrtc-dtls 0.10.0 crypto_gcm.rs).  n sessions; keys n x 16 bytes, ivs n x 4 bytes.
 *      Expands AES-128 round keys, H = AES_K(0^128) and H^1..H^32 on the device.  Host pointers. */
int32_t p2pt_sessions_load(p2pt_ctx* ctx, uint32_t n,
                           const uint8_t* local_keys16, const uint8_t* local_ivs4,
                           const uint8_t* remote_keys16, const uint8_t* remote_ivs4);

/* ---- frame codec ------------------------------------------------------------------------- */
/* TunnelMessage::encode (protocol.rs:148-154): out[out_off[i]..] = [types[i]][sids[i] BE][payload_i].
 * Needs out_off[i+1] - out_off[i] >= 5 + payload length. */
int32_t p2pt_frame_encode_batch(p2pt_ctx* ctx, uint64_t n, const uint8_t* types, const uint32_t* sids,
                                const uint8_t* payload, const uint64_t* pay_off,
                                uint8_t* out, const uint64_t* out_off);
/* TunnelMessage::decode (protocol.rs:157-172), check order as the reference. */
int32_t p2pt_frame_decode_batch(p2pt_ctx* ctx, uint64_t n, const uint8_t* in, const uint64_t* in_off,
                                p2pt_desc* desc);
/* Demux (proxy.rs:118-153, serve.rs:112-139): stable group-by stream_id of the status-OK frames.
 * session (nullable, n entries): the tunnel session each frame arrived on; when given, streams are
 * keyed by (session, stream_id) -- every connection has its own stream_id counter (proxy.rs:52) --
 * and uniq_session (nullable) receives the session of each stream.
 * Streams ascending by (session, stream_id); csr[k+1] offsets into perm; perm[csr[s]..csr[s+1]) =
 * frame indices of stream s in arrival order.  If gathered != NULL, body frames (ReqBody 11 /
 * ResBody 21) are concatenated per stream into gathered and gathered_off[k+1] delimits streams
 * (the Vec<u8> of serve.rs:120-124).  max_streams sizes uniq_* / csr / gathered_off (csr and
 * gathered_off need max_streams + 1 entries); gathered_cap sizes gathered. */
int32_t p2pt_frame_demux_batch(p2pt_ctx* ctx, uint64_t n, const p2pt_desc* desc, const uint32_t* session,
                               const uint8_t* in, uint64_t in_bytes,
                               uint32_t max_streams, uint32_t* uniq_sid, uint32_t* uniq_session,
                               uint64_t* csr, uint32_t* perm,
                               uint32_t* k_out, uint8_t* gathered, uint64_t gathered_cap,
                               uint64_t* gathered_off);

/* The reader loop in one call: decode + demux + body gather of frames that are already plaintext
 * (proxy.rs:109-170: decode :109-116, HashMap probe per frame :118-153; serve.rs:82-156: decode :103-109,
 * per-stream Vec<u8> append :112-139).  in / in_off as in p2pt_frame_decode_batch, everything else as in
 * p2pt_frame_demux_batch.  desc (n entries) and perm are nullable: a reader that only wants the per-stream
 * bodies never pays for 24-byte descriptors or the frame permutation.
 * When max_streams <= 8192 the demux (here and in p2pt_frame_demux_batch) first tries a path without any
 * sort - dense stream ids from a hash table, one ordered counting pass, one gather - and falls back to the
 * sorting path on the device, without the host noticing, if the batch holds more distinct streams than that
 * (k_out is exact either way).  Give max_streams honestly: a caller that passes n "to be safe" always sorts. */
int32_t p2pt_frame_decode_demux_batch(p2pt_ctx* ctx, uint64_t n, const uint8_t* in, const uint64_t* in_off,
                                      const uint32_t* session, p2pt_desc* desc,
                                      uint32_t max_streams, uint32_t* uniq_sid, uint32_t* uniq_session,
                                      uint64_t* csr, uint32_t* perm, uint32_t* k_out,
                                      uint8_t* gathered, uint64_t gathered_cap, uint64_t* gathered_off);

/* ---- DTLS 1.2 AES-128-GCM record layer ------------------------------------------------------ */
/* CryptoGcm::encrypt(pkt_rlh, raw): raw_i = hdr13 || plaintext; rec_i = hdr13' || nonce8 ||
 * ciphertext || tag16 with the header length field rewritten to 8 + len + 16.  nonce8 (n x 8) is
 * the explicit nonce the reference draws from rand::thread_rng(); it is an INPUT here so results
 * are reproducible.  session[i] indexes the table given to p2pt_sessions_load (local key / iv).
 * Needs rec_off[i+1] - rec_off[i] >= raw_len_i + 24.  status: OK or BAD_HEADER (raw_len < 13, session index
 * out of range, or an item above 1 000 000 bytes).  Bytes of an output slot beyond the item written into it are
 * ZERO after the call (seal, open and the fused forms alike: the kernel clears them, so nothing an earlier batch
 * left in the staging memory can reach the caller), and so is the whole slot of an item that was refused.  Do
 * not call p2pt_sessions_load while seal / open calls on the same ctx are in flight. */
int32_t p2pt_dtls_gcm_seal_batch(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                                 const uint8_t* nonce8, const uint8_t* raw, const uint64_t* raw_off,
                                 uint8_t* rec, const uint64_t* rec_off, uint8_t* status);
/* CryptoGcm::decrypt(r): out_i = hdr13 (unchanged) || plaintext, out_len[i] its length.
 * Uses the remote key / iv.  status: OK, AUTH_FAIL (output zeroed, out_len 0), NO_ROOM_FOR_NONCE,
 * PASS_THROUGH (record copied), BAD_HEADER.  Needs out_off[i+1] - out_off[i] >= rec_len_i. */
int32_t p2pt_dtls_gcm_open_batch(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                                 const uint8_t* rec, const uint64_t* rec_off,
                                 uint8_t* out, const uint64_t* out_off,
                                 uint32_t* out_len, uint8_t* status);

/* ---- fused frame + record pipeline (one frame = one data-channel message = one record) ------- */
/* encode -> seal in one pass: record i = hdr13 || nonce8 || E([type][sid BE][payload_i]) || tag.
 * hdr13 = content_type 23, version fe fd, epoch[i], seq48[i], length.  (serve.rs:269-276 followed
 * by dc.send.) */
int32_t p2pt_pipeline_seal(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                           const uint16_t* epoch, const uint64_t* seq48, const uint8_t* nonce8,
                           const uint8_t* types, const uint32_t* sids,
                           const uint8_t* payload, const uint64_t* pay_off,
                           uint8_t* rec, const uint64_t* rec_off, uint8_t* status);
/* open -> decode in one pass: out_i = hdr13 || frame_i; desc[i] describes the frame inside out
 * (pay_off relative to out).  desc[i].status carries the record status when the record did not
 * open, else the frame decode status.  Follow with p2pt_frame_demux_batch on desc/out. */
int32_t p2pt_pipeline_open(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                           const uint8_t* rec, const uint64_t* rec_off,
                           uint8_t* out, const uint64_t* out_off, p2pt_desc* desc);

/* The consumer-side reader loop in one call (proxy.rs:109-170 / serve.rs:82-156 batched):
 * open -> decode -> demux -> gather.  The opened frames stay on the device unless `out` (nullable,
 * sized like rec, with out_off = rec_off) is given; desc / uniq_* / csr / perm / gathered* are as in
 * p2pt_pipeline_open and p2pt_frame_demux_batch (desc[i].pay_off is relative to `out`).  desc and perm are
 * nullable: a reader that only wants the per-stream bodies (uniq_*, csr counts, gathered, gathered_off) saves
 * 28 bytes per record of device->host traffic by passing NULL. */
int32_t p2pt_pipeline_open_demux(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                                 const uint8_t* rec, const uint64_t* rec_off,
                                 uint8_t* out, p2pt_desc* desc,
                                 uint32_t max_streams, uint32_t* uniq_sid, uint32_t* uniq_session,
                                 uint64_t* csr, uint32_t* perm, uint32_t* k_out,
                                 uint8_t* gathered, uint64_t gathered_cap, uint64_t* gathered_off);

/* ---- SCTP DATA chunks + CRC32c: the layer between the frame and the record.
 *      webrtc-sctp 0.10.0 (un-vendored, Cargo.lock:2910-2911) cuts every data-channel message into
 *      DATA chunks of <= 1200 bytes and checksums each packet with CRC32c (RFC 4960 3.1, 3.3.1, app. B).
 *      Here: one DATA chunk per packet, fragments of one message get consecutive TSNs, B / E flags. */
uint64_t p2pt_sctp_fragments(uint64_t msg_len);        /* max(1, ceil(len / 1200)) */
uint64_t p2pt_sctp_packets_bytes(uint64_t msg_len);    /* bytes of all packets of one message, no DTLS header */
/* messages -> packets, packed back to back into out.  Per message: ports, verification tag, stream id,
 * stream sequence number, PPID (53 = binary), TSN of its first fragment.  If epoch != NULL every packet
 * is preceded by a 13-byte DTLS record header (type 23, fe fd, epoch[i], seq48_first[i] + fragment,
 * length), i.e. the output is directly the `raw` input of p2pt_dtls_gcm_seal_batch.
 * pkt_off receives n_pkts + 1 byte offsets; pkt_cap = entries available in pkt_off. */
int32_t p2pt_sctp_packetize_batch(p2pt_ctx* ctx, uint64_t n, const uint8_t* msg, const uint64_t* msg_off,
                                  const uint16_t* src_port, const uint16_t* dst_port, const uint32_t* vtag,
                                  const uint16_t* sid, const uint16_t* ssn, const uint32_t* ppid, const uint32_t* tsn0,
                                  const uint16_t* epoch, const uint64_t* seq48_first,
                                  uint8_t* out, uint64_t out_cap, uint64_t* pkt_off, uint64_t pkt_cap,
                                  uint64_t* n_pkts_out);
/* packets (each after `headroom` bytes, 13 when they come out of p2pt_dtls_gcm_open_batch) -> CRC check,
 * chunk parse, and reassembly: fragments are ordered by (session, TSN).  TSNs are serial numbers modulo 2^32
 * whose start the PEER chooses, so a message may straddle 0xFFFFFFFF -> 0: give tsn_base (n_pkts entries, the
 * cumulative TSN ack point of each packet's association - what an SCTP stack always has at hand) and fragments
 * are ordered by their distance from it, like webrtc-sctp's sna32 comparisons; with tsn_base == NULL raw TSNs
 * are compared and must not wrap inside a batch.  Of several packets with the same (session, TSN) - retransmissions - only the first to arrive
 * counts, as in webrtc-sctp's reassembly queue; messages = runs from a B to an E fragment, their payloads
 * concatenated into msg with msg_off[n_msgs + 1] -- exactly the (buffer, offsets) input of
 * p2pt_frame_decode_batch.  msg_first[m] = packet index of message m's first fragment (its info gives
 * stream / ssn / ppid); msg_status[m] = OK or SCTP_INCOMPLETE.  info (n entries) is always filled. */
int32_t p2pt_sctp_unpack_reassemble_batch(p2pt_ctx* ctx, uint64_t n_pkts, const uint8_t* in, const uint64_t* in_off,
                                          uint32_t headroom, const uint32_t* session, const uint32_t* tsn_base,
                                          p2pt_sctp_chunk* info,
                                          uint32_t max_msgs, uint8_t* msg, uint64_t msg_cap, uint64_t* msg_off,
                                          uint32_t* msg_first, uint8_t* msg_status, uint32_t* n_msgs_out);

/* ---- DTLS record-layer bookkeeping (webrtc-dtls 0.10.0, un-vendored) ------------
 * unpack_datagram: a UDP datagram carries one or more records back to back (13-byte header + length).
 * rec_off[j] = byte offset of record j inside buf, rec_dgram[j] = its datagram; a record ends where the next
 * one of the same datagram starts, or at the end of its datagram.  A datagram with a truncated record is
 * dropped whole (dgram_status = BAD_HEADER), like ErrInvalidPacketLength in the reference.
 * max_records sizes rec_off (max_records + 1 entries) and rec_dgram. */
int32_t p2pt_dtls_datagram_split_batch(p2pt_ctx* ctx, uint64_t n_dgrams, const uint8_t* buf, const uint64_t* dgram_off,
                                       uint64_t max_records, uint64_t* rec_off, uint32_t* rec_dgram,
                                       uint8_t* dgram_status, uint32_t* n_records_out);
/* Anti-replay sliding window (replaydetector SlidingWindowDetector, 64 entries), one state per session:
 * latest / mask / started are arrays of n_sessions entries that the caller keeps between batches (host
 * arrays here).  Records are judged in arrival order within their session: accept[i] = 1 if the record
 * is neither older than the window nor a duplicate AND authentic[i] != 0 (authentic may be NULL = all);
 * only accepted records advance the window, exactly as check() / accept() do around the AEAD open. */
int32_t p2pt_dtls_replay_filter_batch(p2pt_ctx* ctx, uint64_t n, const uint32_t* session, const uint64_t* seq48,
                                      const uint8_t* authentic, uint32_t n_sessions,
                                      uint64_t* latest, uint64_t* mask, uint8_t* started, uint8_t* accept);
/* Sequence numbers of outgoing records (webrtc-dtls 0.10.0 conn.rs, the write path, un-vendored: every record takes
 * its epoch's counter with fetch_add(1); more than 2^48 - 1 is ErrSequenceNumberOverflow).  next_seq[s] is the number
 * the next record of session s gets; the caller keeps the array (n_sessions entries, host array here) between batches
 * and restarts a session at 0 when its epoch changes.  seq48[i] = next_seq[session[i]] + the number of records of the
 * same session in front of record i in this batch, ok[i] = 1; afterwards every next_seq[s] has advanced by the
 * session's records.  ok[i] = 0 and seq48[i] = 0 for a record whose number would pass 2^48 - 1 or whose session index is
 * not below n_sessions: it must not be sealed.  The result feeds the seq48 argument of the seal calls. */
int32_t p2pt_dtls_seq_assign_batch(p2pt_ctx* ctx, uint64_t n, const uint32_t* session, uint32_t n_sessions,
                                   uint64_t* next_seq, uint64_t* seq48, uint8_t* ok);

/* ---- SSE event scan over per-stream bodies.  body / body_off[k + 1] are the
 *      gathered / gathered_off outputs of the demux.  An event ends after every "\n\n" inside one stream:
 *      ev_csr[k + 1] = running event count per stream, ev_end[j] = byte offset (into body) just past event j;
 *      done[s] = 1 if some event of stream s begins with `data: [DONE]\n\n`
 *      (/root/reference/tmp/mock_llm.py:68-69).  max_events sizes ev_end. */
int32_t p2pt_sse_scan_batch(p2pt_ctx* ctx, uint64_t k, const uint8_t* body, const uint64_t* body_off,
                            uint64_t max_events, uint64_t* ev_csr, uint64_t* ev_end, uint8_t* done,
                            uint64_t* n_events_out);

/* ---- device-pointer variants: same arguments, all pointers are device pointers, `stream` is a
 *      cudaStream_t.  They enqueue on `stream` and return without synchronising. -------------- */
int32_t p2pt_frame_encode_batch_device(p2pt_ctx* ctx, uint64_t n, const uint8_t* types, const uint32_t* sids,
                                       const uint8_t* payload, const uint64_t* pay_off,
                                       uint8_t* out, const uint64_t* out_off, void* stream);
int32_t p2pt_frame_decode_batch_device(p2pt_ctx* ctx, uint64_t n, const uint8_t* in, const uint64_t* in_off,
                                       p2pt_desc* desc, void* stream);
/* scratch: device memory of at least p2pt_frame_demux_scratch_bytes(n) bytes, 256-byte aligned.
 * k_out is a device pointer to one u32. */
uint64_t p2pt_frame_demux_scratch_bytes(uint64_t n);
int32_t p2pt_frame_demux_batch_device(p2pt_ctx* ctx, uint64_t n, const p2pt_desc* desc, const uint32_t* session,
                                      const uint8_t* in,
                                      uint32_t max_streams, uint32_t* uniq_sid, uint32_t* uniq_session,
                                      uint64_t* csr, uint32_t* perm,
                                      uint32_t* k_out, uint8_t* gathered, uint64_t gathered_cap,
                                      uint64_t* gathered_off, void* scratch, uint64_t scratch_bytes,
                                      void* stream);
/* device form of p2pt_frame_decode_demux_batch; scratch of p2pt_frame_decode_demux_scratch_bytes(n) bytes,
 * 256-byte aligned (it also holds descriptors / perm for the sorting path when the caller passes NULL). */
uint64_t p2pt_frame_decode_demux_scratch_bytes(uint64_t n);
int32_t p2pt_frame_decode_demux_batch_device(p2pt_ctx* ctx, uint64_t n, const uint8_t* in, const uint64_t* in_off,
                                             const uint32_t* session, p2pt_desc* desc,
                                             uint32_t max_streams, uint32_t* uniq_sid, uint32_t* uniq_session,
                                             uint64_t* csr, uint32_t* perm, uint32_t* k_out,
                                             uint8_t* gathered, uint64_t gathered_cap, uint64_t* gathered_off,
                                             void* scratch, uint64_t scratch_bytes, void* stream);
/* No counterpart in the reference (it has no device memory to manage).
 * Same work and same results as p2pt_frame_demux_batch_device, for callers that may not allocate
 * device memory while they run and have nowhere to take `scratch` from: for example a reader task
 * that was handed a fixed device-memory budget at start-up because an inference server on the same
 * GPU owns the rest of HBM, or a per-connection loop that cannot afford cudaMalloc / cudaFree
 * because those take the driver's allocator lock and synchronise the whole device, stalling every
 * other session's stream.  The scratch lives in a static device array inside the library
 * (about 10.8 MB, resident on every GPU that loads the library until the process exits), which
 * bounds the batch: n <= P2PT_NOALLOC_MAX_FRAMES, larger batches are refused with
 * P2PT_ERR_CAPACITY.  The call itself performs no cudaMalloc / cudaFree and uses no allocating
 * library routine.
 * The first call on each device looks up the array's device address (cudaGetSymbolAddress, which
 * may make the driver load the module); make that first call during start-up, before any stream
 * capture begins.  All calls on a device share the one array and it keeps the previous call's
 * contents; the call clears the few words that must start at zero itself.  Two calls whose
 * kernels can overlap on the GPU (different streams, or different contexts/threads on the same
 * device) race on the array: serialise them. */
#define P2PT_NOALLOC_MAX_FRAMES 262144u
int32_t p2pt_frame_demux_batch_device_noalloc(p2pt_ctx* ctx, uint64_t n, const p2pt_desc* desc, const uint32_t* session,
                                              const uint8_t* in,
                                              uint32_t max_streams, uint32_t* uniq_sid, uint32_t* uniq_session,
                                              uint64_t* csr, uint32_t* perm,
                                              uint32_t* k_out, uint8_t* gathered, uint64_t gathered_cap,
                                              uint64_t* gathered_off, void* stream);
/* device forms of the SCTP calls.  packetize: the caller supplies the layout -- out_msg_off[i] = byte offset
 * of message i's first packet (its packets follow back to back), pkt_first[n + 1] = running packet count;
 * pkt_off[pkt_first[n] + 1] is written.  reassemble: scratch of p2pt_sctp_reassemble_scratch_bytes(n). */
int32_t p2pt_sctp_packetize_batch_device(p2pt_ctx* ctx, uint64_t n, const uint8_t* msg, const uint64_t* msg_off,
                                         const uint16_t* src_port, const uint16_t* dst_port, const uint32_t* vtag,
                                         const uint16_t* sid, const uint16_t* ssn, const uint32_t* ppid, const uint32_t* tsn0,
                                         const uint16_t* epoch, const uint64_t* seq48_first,
                                         uint8_t* out, const uint64_t* out_msg_off, const uint64_t* pkt_first,
                                         uint64_t* pkt_off, void* stream);
int32_t p2pt_sctp_unpack_batch_device(p2pt_ctx* ctx, uint64_t n_pkts, const uint8_t* in, const uint64_t* in_off,
                                      uint32_t headroom, const uint32_t* tsn_base, p2pt_sctp_chunk* info, p2pt_desc* desc,
                                      void* stream);
uint64_t p2pt_sctp_reassemble_scratch_bytes(uint64_t n_pkts);
int32_t p2pt_sctp_reassemble_batch_device(p2pt_ctx* ctx, uint64_t n_pkts, const p2pt_sctp_chunk* info,
                                          const p2pt_desc* desc, const uint32_t* session, const uint8_t* in,
                                          uint32_t max_msgs, uint8_t* msg, uint64_t msg_cap, uint64_t* msg_off,
                                          uint32_t* msg_first, uint8_t* msg_status, uint32_t* n_msgs_out,
                                          void* scratch, uint64_t scratch_bytes, void* stream);
uint64_t p2pt_dtls_datagram_split_scratch_bytes(uint64_t n_dgrams);
int32_t p2pt_dtls_datagram_split_batch_device(p2pt_ctx* ctx, uint64_t n_dgrams, const uint8_t* buf, const uint64_t* dgram_off,
                                              uint64_t max_records, uint64_t* rec_off, uint32_t* rec_dgram,
                                              uint8_t* dgram_status, uint32_t* n_records_out,
                                              void* scratch, uint64_t scratch_bytes, void* stream);
uint64_t p2pt_dtls_seq_assign_scratch_bytes(uint64_t n);
int32_t p2pt_dtls_seq_assign_batch_device(p2pt_ctx* ctx, uint64_t n, const uint32_t* session, uint32_t n_sessions,
                                          uint64_t* next_seq, uint64_t* seq48, uint8_t* ok,
                                          void* scratch, uint64_t scratch_bytes, void* stream);
uint64_t p2pt_dtls_replay_filter_scratch_bytes(uint64_t n);
int32_t p2pt_dtls_replay_filter_batch_device(p2pt_ctx* ctx, uint64_t n, const uint32_t* session, const uint64_t* seq48,
                                             const uint8_t* authentic, uint32_t n_sessions,
                                             uint64_t* latest, uint64_t* mask, uint8_t* started, uint8_t* accept,
                                             void* scratch, uint64_t scratch_bytes, void* stream);
/* The same scan in ONE pass over the bodies (the call above reads them twice: count, then emit): the caller says how
 * many body bytes there are at most (body_bytes >= body_off[k] - body_off[0], e.g. the gathered_cap it gave the demux) -
 * the tile states of the scan need scratch in proportion to it, and body_off lives on the device.  Same outputs. */
uint64_t p2pt_sse_scan_onepass_scratch_bytes(uint64_t k, uint64_t body_bytes);
int32_t p2pt_sse_scan_onepass_batch_device(p2pt_ctx* ctx, uint64_t k, const uint8_t* body, const uint64_t* body_off,
                                           uint64_t body_bytes, uint64_t max_events, uint64_t* ev_csr, uint64_t* ev_end,
                                           uint8_t* done, uint64_t* n_events_out, void* scratch, uint64_t scratch_bytes,
                                           void* stream);
uint64_t p2pt_sse_scan_scratch_bytes(uint64_t k);
int32_t p2pt_sse_scan_batch_device(p2pt_ctx* ctx, uint64_t k, const uint8_t* body, const uint64_t* body_off,
                                   uint64_t max_events, uint64_t* ev_csr, uint64_t* ev_end, uint8_t* done,
                                   uint64_t* n_events_out, void* scratch, uint64_t scratch_bytes, void* stream);
int32_t p2pt_dtls_gcm_seal_batch_device(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                                        const uint8_t* nonce8, const uint8_t* raw, const uint64_t* raw_off,
                                        uint8_t* rec, const uint64_t* rec_off, uint8_t* status, void* stream);
int32_t p2pt_dtls_gcm_open_batch_device(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                                        const uint8_t* rec, const uint64_t* rec_off,
                                        uint8_t* out, const uint64_t* out_off,
                                        uint32_t* out_len, uint8_t* status, void* stream);
int32_t p2pt_pipeline_seal_device(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                                  const uint16_t* epoch, const uint64_t* seq48, const uint8_t* nonce8,
                                  const uint8_t* types, const uint32_t* sids,
                                  const uint8_t* payload, const uint64_t* pay_off,
                                  uint8_t* rec, const uint64_t* rec_off, uint8_t* status, void* stream);
int32_t p2pt_pipeline_open_device(p2pt_ctx* ctx, uint64_t n, const uint32_t* session,
                                  const uint8_t* rec, const uint64_t* rec_off,
                                  uint8_t* out, const uint64_t* out_off, p2pt_desc* desc, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* P2PT_H */
