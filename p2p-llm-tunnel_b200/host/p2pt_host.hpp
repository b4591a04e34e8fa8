// p2pt_host.hpp -- C++ host-side mirror of the reference's interface for the hot path, over the C ABI.
//
// The reference is compiled code (Rust); no Rust toolchain exists in the build image, so the host
// layer a maintainer would write in Rust is provided here in C++ with the same
// names, argument meaning and error behaviour:
//   protocol::MessageType / from_u8 / TunnelMessage::{encode, decode} + constructors
//       /root/reference/tunnel/src/protocol.rs:86-118, :139-262, limits :9-12
//   dtls::CryptoGcm::{encrypt, decrypt}
//       webrtc-dtls 0.10.0 crypto_gcm.rs (un-vendored, /root/reference/tunnel/Cargo.lock:2822-2824)
//   reader::ProxyReader / reader::ServeReassembler -- the two per-frame reader loops in batched form
//       /root/reference/tunnel/src/proxy.rs:105-172 (PendingResponses :23-24, StreamEvent :28-37)
//       /root/reference/tunnel/src/serve.rs:82-156 (pending_requests :65)
// Every encode / decode / encrypt / decrypt runs on the GPU (batch of one through the batch entry
// points; encode_batch / decode_batch for real batches).  Nothing here computes on the CPU.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/p2pt.h"

namespace p2pt_host {

using Bytes = std::vector<uint8_t>;

struct Error : std::runtime_error {          // anyhow::Error in the reference
    using std::runtime_error::runtime_error;
};

class Context {                               // one per GPU
public:
    explicit Context(int device = 0) {
        if (p2pt_ctx_create(device, &ctx_) != P2PT_OK) throw Error(std::string("p2pt_ctx_create: ") + p2pt_last_error(nullptr));
    }
    ~Context() { if (ctx_) p2pt_ctx_destroy(ctx_); }
    Context(const Context&) = delete;
    Context& operator=(const Context&) = delete;
    p2pt_ctx* get() const { return ctx_; }
    void check(int32_t rc) const { if (rc != P2PT_OK) throw Error(p2pt_last_error(ctx_)); }
private:
    p2pt_ctx* ctx_ = nullptr;
};

namespace protocol {

constexpr uint32_t PROTOCOL_VERSION = 1;                       // protocol.rs:6
constexpr const char* PROTOCOL_NAME = "httptunnel";            // protocol.rs:7
constexpr size_t MAX_FRAME_SIZE = 64 * 1024;                   // protocol.rs:10
constexpr size_t MAX_BODY_CHUNK = MAX_FRAME_SIZE - 128;        // protocol.rs:12

enum class MessageType : uint8_t {                             // protocol.rs:86-100
    Hello = 1, Agree = 2, Ping = 3, Pong = 4,
    ReqHeaders = 10, ReqBody = 11, ReqEnd = 12,
    ResHeaders = 20, ResBody = 21, ResEnd = 22,
    Error = 99,
};

inline std::optional<MessageType> from_u8(uint8_t v) {         // protocol.rs:103-118
    switch (v) {
        case 1: case 2: case 3: case 4: case 10: case 11: case 12: case 20: case 21: case 22: case 99:
            return static_cast<MessageType>(v);
        default: return std::nullopt;
    }
}

struct TunnelMessage {                                         // protocol.rs:139-144
    MessageType msg_type;
    uint32_t stream_id;
    Bytes payload;

    // Encode: [1 byte type][4 byte stream_id BE][payload...]   protocol.rs:148-154
    Bytes encode(const Context& c) const {
        const uint8_t t = static_cast<uint8_t>(msg_type);
        const uint64_t pay_off[2] = {0, payload.size()}, out_off[2] = {0, 5 + payload.size()};
        Bytes out(5 + payload.size());
        const uint8_t dummy = 0;
        c.check(p2pt_frame_encode_batch(c.get(), 1, &t, &stream_id, payload.empty() ? &dummy : payload.data(), pay_off,
                                        out.data(), out_off));
        return out;
    }

    // Decode from raw bytes                                     protocol.rs:157-172
    static TunnelMessage decode(const Context& c, const Bytes& data) {
        const uint64_t off[2] = {0, data.size()};
        p2pt_desc d{};
        const uint8_t dummy = 0;
        c.check(p2pt_frame_decode_batch(c.get(), 1, data.empty() ? &dummy : data.data(), off, &d));
        if (d.status == P2PT_ST_TOO_SHORT) throw Error("message too short: " + std::to_string(data.size()) + " bytes");
        if (d.status == P2PT_ST_UNKNOWN_TYPE) throw Error("unknown message type: " + std::to_string((int)d.msg_type));
        TunnelMessage m{static_cast<MessageType>(d.msg_type), d.stream_id, {}};
        m.payload.assign(data.begin() + (ptrdiff_t)d.pay_off, data.begin() + (ptrdiff_t)(d.pay_off + d.pay_len));
        return m;
    }

    // Convenience constructors                                   protocol.rs:176-262 (JSON payloads are opaque here)
    static TunnelMessage hello(const std::string& json) { return {MessageType::Hello, 0, Bytes(json.begin(), json.end())}; }
    static TunnelMessage agree(const std::string& json) { return {MessageType::Agree, 0, Bytes(json.begin(), json.end())}; }
    static TunnelMessage req_headers(uint32_t sid, const std::string& json) { return {MessageType::ReqHeaders, sid, Bytes(json.begin(), json.end())}; }
    static TunnelMessage req_body(uint32_t sid, Bytes data) { return {MessageType::ReqBody, sid, std::move(data)}; }
    static TunnelMessage req_end(uint32_t sid) { return {MessageType::ReqEnd, sid, {}}; }
    static TunnelMessage res_headers(uint32_t sid, const std::string& json) { return {MessageType::ResHeaders, sid, Bytes(json.begin(), json.end())}; }
    static TunnelMessage res_body(uint32_t sid, Bytes data) { return {MessageType::ResBody, sid, std::move(data)}; }
    static TunnelMessage res_end(uint32_t sid) { return {MessageType::ResEnd, sid, {}}; }
    static TunnelMessage error(uint32_t sid, const std::string& msg) { return {MessageType::Error, sid, Bytes(msg.begin(), msg.end())}; }
    static TunnelMessage ping() { return {MessageType::Ping, 0, {}}; }
    static TunnelMessage pong() { return {MessageType::Pong, 0, {}}; }
};

// serve.rs:269-276 / proxy.rs:318-330: sub-chunk a body at MAX_BODY_CHUNK
inline std::vector<TunnelMessage> chunk_body(uint32_t sid, const Bytes& body, bool response = true) {
    std::vector<TunnelMessage> out;
    for (size_t o = 0; o < body.size(); o += MAX_BODY_CHUNK) {
        const size_t e = std::min(o + MAX_BODY_CHUNK, body.size());
        Bytes sub(body.begin() + (ptrdiff_t)o, body.begin() + (ptrdiff_t)e);
        out.push_back(response ? TunnelMessage::res_body(sid, std::move(sub)) : TunnelMessage::req_body(sid, std::move(sub)));
    }
    return out;
}

// Batched forms: one GPU call for many messages.
inline Bytes encode_batch(const Context& c, const std::vector<TunnelMessage>& msgs, std::vector<uint64_t>& frame_off) {
    const size_t n = msgs.size();
    std::vector<uint8_t> types(n);
    std::vector<uint32_t> sids(n);
    std::vector<uint64_t> pay_off(n + 1, 0);
    frame_off.assign(n + 1, 0);
    for (size_t i = 0; i < n; i++) {
        types[i] = static_cast<uint8_t>(msgs[i].msg_type); sids[i] = msgs[i].stream_id;
        pay_off[i + 1] = pay_off[i] + msgs[i].payload.size();
        frame_off[i + 1] = frame_off[i] + 5 + msgs[i].payload.size();
    }
    Bytes payload(pay_off[n] + 1), out(frame_off[n] + 1);
    for (size_t i = 0; i < n; i++) std::memcpy(payload.data() + pay_off[i], msgs[i].payload.data(), msgs[i].payload.size());
    if (n) c.check(p2pt_frame_encode_batch(c.get(), n, types.data(), sids.data(), payload.data(), pay_off.data(), out.data(), frame_off.data()));
    out.resize(frame_off[n]);
    return out;
}

}  // namespace protocol

// ---- batched reader / dispatcher ------------------------------------------------
// The reference decodes one frame, takes a mutex, probes a HashMap and pushes one event, per frame.  Here a
// whole batch is decoded and demultiplexed on the GPU (p2pt_frame_decode_demux_batch) and
// the host walks each stream's frames once, in arrival order: the same per-stream event sequences come out.
namespace reader {

using protocol::MessageType;

struct StreamEvent {                                           // proxy.rs:28-37
    enum Kind { Headers, Body, End, Error } kind;
    Bytes value;                                               // headers JSON / body bytes / error text (lossy UTF-8 is the caller's)
    bool operator==(const StreamEvent& o) const { return kind == o.kind && value == o.value; }
};

// The top-level unsigned integer field `key` of a JSON object (RequestHeaders / ResponseHeaders carry the stream
// id there, protocol.rs:122-136; headers frames are routed by it, not by the frame header: proxy.rs:129,
// serve.rs:113-118).  Nothing else of the JSON is interpreted here.
inline std::optional<uint32_t> json_u32_field(const uint8_t* p, size_t n, const char* key) {
    const size_t klen = std::strlen(key);
    int depth = 0;
    for (size_t i = 0; i < n; i++) {
        const uint8_t ch = p[i];
        if (ch == '"') {
            size_t j = i + 1;
            while (j < n && p[j] != '"') j += (p[j] == '\\') ? 2 : 1;
            if (j >= n) return std::nullopt;                    // unterminated string: not JSON
            if (depth == 1 && j - i - 1 == klen && std::memcmp(p + i + 1, key, klen) == 0) {
                size_t q = j + 1;
                while (q < n && (p[q] == ' ' || p[q] == '\t' || p[q] == '\n' || p[q] == '\r')) q++;
                if (q < n && p[q] == ':') {
                    q++;
                    while (q < n && (p[q] == ' ' || p[q] == '\t' || p[q] == '\n' || p[q] == '\r')) q++;
                    uint64_t v = 0; size_t digits = 0;
                    while (q < n && p[q] >= '0' && p[q] <= '9' && digits < 11) { v = v * 10 + (p[q] - '0'); q++; digits++; }
                    if (digits && v <= 0xffffffffull) return (uint32_t)v;
                    return std::nullopt;
                }
            }
            i = j;
        } else if (ch == '{' || ch == '[') depth++;
        else if (ch == '}' || ch == ']') depth--;
    }
    return std::nullopt;
}
inline bool json_looks_complete(const uint8_t* p, size_t n) {   // a whole object: what serde_json::from_slice needs at least
    int depth = 0; bool any = false;
    for (size_t i = 0; i < n; i++) {
        if (p[i] == '"') { size_t j = i + 1; while (j < n && p[j] != '"') j += (p[j] == '\\') ? 2 : 1; if (j >= n) return false; i = j; }
        else if (p[i] == '{') { depth++; any = true; }
        else if (p[i] == '}') depth--;
    }
    return any && depth == 0;
}

// one decoded + demultiplexed batch, walked per target stream in arrival order
struct Routed { uint64_t idx; uint8_t type; const uint8_t* data; size_t len; };

inline std::map<int64_t, std::vector<Routed>> decode_demux(const Context& c, const std::vector<Bytes>& frames, uint8_t header_type,
                                                           Bytes& buf, Bytes& gathered, size_t& decode_failures) {
    const size_t n = frames.size();
    std::vector<uint64_t> off(n + 1, 0);
    for (size_t i = 0; i < n; i++) off[i + 1] = off[i] + frames[i].size();
    buf.assign(off[n] + 1, 0);
    for (size_t i = 0; i < n; i++) std::memcpy(buf.data() + off[i], frames[i].data(), frames[i].size());
    std::vector<p2pt_desc> desc(n);
    std::vector<uint32_t> uniq(n + 1), perm(n);
    std::vector<uint64_t> csr(n + 1), goff(n + 1);
    gathered.assign(off[n] + 1, 0);
    uint32_t k = 0;
    // decode + demux + body gather in ONE call (descriptors and perm wanted here: the host walks every frame)
    c.check(p2pt_frame_decode_demux_batch(c.get(), n, buf.data(), off.data(), nullptr, desc.data(), (uint32_t)n, uniq.data(), nullptr,
                                          csr.data(), perm.data(), &k, gathered.data(), off[n], goff.data()));
    for (const auto& d : desc) decode_failures += d.status != P2PT_ST_OK;          // warn! + continue
    std::map<int64_t, std::vector<Routed>> out;
    for (uint32_t s = 0; s < k; s++) {
        uint64_t cursor = goff[s];
        for (uint64_t q = csr[s]; q < csr[s + 1]; q++) {
            const p2pt_desc& d = desc[perm[q]];
            Routed r{perm[q], d.msg_type, nullptr, d.pay_len};
            if (d.msg_type == 11 || d.msg_type == 21) { r.data = gathered.data() + cursor; cursor += d.pay_len; }   // body: slice of the stream's gathered buffer
            else r.data = buf.data() + d.pay_off;
            int64_t target = uniq[s];
            if (d.msg_type == header_type) {
                auto sid = json_looks_complete(r.data, r.len) ? json_u32_field(r.data, r.len, "stream_id") : std::nullopt;
                target = sid ? (int64_t)*sid : -1;               // unparsable headers
            }
            out[target].push_back(r);
        }
    }
    for (auto& kv : out) std::sort(kv.second.begin(), kv.second.end(), [](const Routed& a, const Routed& b) { return a.idx < b.idx; });
    return out;
}

// Consumer side: routes RES_* frames to per-stream event queues (proxy.rs:105-172).
class ProxyReader {
public:
    explicit ProxyReader(const Context& c) : c_(c) {}
    std::map<uint32_t, std::vector<StreamEvent>> pending, closed;   // pending = the mpsc senders; closed = finished streams' queues
    size_t pongs_to_send = 0, decode_failures = 0;
    void register_stream(uint32_t sid) { pending[sid]; }            // proxy.rs:266-270 pending.insert(stream_id, event_tx)
    void feed(const std::vector<Bytes>& frames) {
        if (frames.empty()) return;
        Bytes buf, gathered;
        auto per = decode_demux(c_, frames, (uint8_t)MessageType::ResHeaders, buf, gathered, decode_failures);
        for (auto& kv : per) {
            for (const Routed& r : kv.second) {
                const auto t = (MessageType)r.type;
                if (t == MessageType::Ping) { pongs_to_send++; continue; }                       // proxy.rs:154-162
                if (kv.first < 0) continue;                                                       // failed to parse response headers
                const uint32_t sid = (uint32_t)kv.first;
                auto it = pending.find(sid);
                if (it == pending.end()) continue;                                                // unknown stream: silently dropped (:135)
                const Bytes v(r.data, r.data + r.len);
                if (t == MessageType::ResHeaders) it->second.push_back({StreamEvent::Headers, v});
                else if (t == MessageType::ResBody) it->second.push_back({StreamEvent::Body, v});
                else if (t == MessageType::ResEnd || t == MessageType::Error) {                   // remove + terminal event (:139-153)
                    it->second.push_back({t == MessageType::ResEnd ? StreamEvent::End : StreamEvent::Error, t == MessageType::ResEnd ? Bytes() : v});
                    closed[sid] = std::move(it->second);
                    pending.erase(it);
                }
                // Pong and everything else: ignored (:163-169)
            }
        }
    }
private:
    const Context& c_;
};

// Provider side: REQ_HEADERS / REQ_BODY / REQ_END -> completed requests (serve.rs:82-156).
class ServeReassembler {
public:
    struct Request { std::string headers_json; Bytes body; };
    explicit ServeReassembler(const Context& c) : c_(c) {}
    std::map<uint32_t, Request> pending_requests;                   // serve.rs:65
    size_t pongs_to_send = 0, decode_failures = 0;
    // the requests completed by this batch, in REQ_END arrival order
    std::vector<Request> feed(const std::vector<Bytes>& frames) {
        std::vector<std::pair<uint64_t, Request>> done;
        if (frames.empty()) return {};
        Bytes buf, gathered;
        auto per = decode_demux(c_, frames, (uint8_t)MessageType::ReqHeaders, buf, gathered, decode_failures);
        for (auto& kv : per) {
            for (const Routed& r : kv.second) {
                const auto t = (MessageType)r.type;
                if (t == MessageType::Ping) { pongs_to_send++; continue; }
                if (t == MessageType::ReqHeaders) {
                    if (kv.first < 0) throw Error("invalid request headers JSON");               // serde `?` ends run_serve (serve.rs:113)
                    pending_requests[(uint32_t)kv.first] = Request{std::string(r.data, r.data + r.len), {}};   // :112-119 insert
                    continue;
                }
                if (kv.first < 0) continue;
                auto it = pending_requests.find((uint32_t)kv.first);
                if (it == pending_requests.end()) continue;                                       // body / end for an unknown request
                if (t == MessageType::ReqBody) it->second.body.insert(it->second.body.end(), r.data, r.data + r.len);   // :120-124 extend_from_slice
                else if (t == MessageType::ReqEnd) { done.push_back({r.idx, std::move(it->second)}); pending_requests.erase(it); }   // :125-139
            }
        }
        std::sort(done.begin(), done.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
        std::vector<Request> out;
        for (auto& d : done) out.push_back(std::move(d.second));
        return out;
    }
private:
    const Context& c_;
};

}  // namespace reader

namespace dtls {

constexpr size_t RECORD_LAYER_HEADER_SIZE = 13;

struct ErrNotEnoughRoomForNonce : Error { ErrNotEnoughRoomForNonce() : Error("buffer not long enough to contain nonce") {} };
struct ErrDecrypt : Error { ErrDecrypt() : Error("aead decrypt failed") {} };
struct ErrBadHeader : Error { ErrBadHeader() : Error("record layer header invalid") {} };

// CryptoGcm::new(local_key, local_write_iv, remote_key, remote_write_iv)
class CryptoGcm {
public:
    CryptoGcm(const Context& c, const uint8_t local_key[16], const uint8_t local_write_iv[4],
              const uint8_t remote_key[16], const uint8_t remote_write_iv[4]) : c_(c) {
        c_.check(p2pt_sessions_load(c_.get(), 1, local_key, local_write_iv, remote_key, remote_write_iv));
    }
    // encrypt(pkt_rlh, raw): raw = marshalled header (13) + payload.  The 8-byte explicit nonce, random in the
    // reference, is passed in so results are reproducible.
    Bytes encrypt(const Bytes& raw, const uint8_t nonce8[8]) const {
        const uint32_t sess = 0;
        const uint64_t in_off[2] = {0, raw.size()}, out_off[2] = {0, raw.size() + 24};
        Bytes out(raw.size() + 24);
        uint8_t st = 0;
        c_.check(p2pt_dtls_gcm_seal_batch(c_.get(), 1, &sess, nonce8, raw.data(), in_off, out.data(), out_off, &st));
        if (st != P2PT_ST_OK) throw ErrBadHeader();
        return out;
    }
    Bytes decrypt(const Bytes& r) const {
        const uint32_t sess = 0;
        const uint64_t off[2] = {0, r.size()};
        Bytes out(r.size() + 1);
        uint32_t olen = 0; uint8_t st = 0;
        const uint8_t dummy = 0;
        c_.check(p2pt_dtls_gcm_open_batch(c_.get(), 1, &sess, r.empty() ? &dummy : r.data(), off, out.data(), off, &olen, &st));
        if (st == P2PT_ST_NO_ROOM_FOR_NONCE) throw ErrNotEnoughRoomForNonce();
        if (st == P2PT_ST_BAD_HEADER) throw ErrBadHeader();
        if (st != P2PT_ST_OK && st != P2PT_ST_PASS_THROUGH) throw ErrDecrypt();
        out.resize(olen);
        return out;
    }
private:
    const Context& c_;
};

// ---- record-layer bookkeeping (webrtc-dtls 0.10.0 conn / record layer, un-vendored) ----

// unpack_datagram for a batch: the records of every datagram, back to back; a datagram with a truncated record (or a
// trailing bare header) is dropped whole.
struct Unpacked { std::vector<uint64_t> rec_off; std::vector<uint32_t> rec_dgram; std::vector<uint8_t> dgram_status; };
inline Unpacked unpack_datagrams(const Context& c, const Bytes& buf, const std::vector<uint64_t>& dgram_off) {
    Unpacked u;
    const uint64_t n = dgram_off.empty() ? 0 : dgram_off.size() - 1;
    const uint64_t cap = buf.size() / RECORD_LAYER_HEADER_SIZE + 1;
    u.rec_off.assign(cap + 1, 0); u.rec_dgram.assign(cap, 0); u.dgram_status.assign(n ? n : 1, 0);
    uint32_t k = 0;
    const uint8_t dummy = 0;
    c.check(p2pt_dtls_datagram_split_batch(c.get(), n, buf.empty() ? &dummy : buf.data(), dgram_off.data(), cap, u.rec_off.data(),
                                           u.rec_dgram.data(), u.dgram_status.data(), &k));
    u.rec_off.resize(k + 1); u.rec_dgram.resize(k); u.dgram_status.resize(n);
    return u;
}

// replaydetector::SlidingWindowDetector (window 64) for every session of a connection table: check() before the record is
// opened, accept() only for records that authenticated - a batch at a time, in arrival order per session.
class ReplayDetector {
public:
    ReplayDetector(const Context& c, size_t n_sessions) : c_(c), latest_(n_sessions, 0), mask_(n_sessions, 0), started_(n_sessions, 0) {}
    std::vector<uint8_t> filter(const std::vector<uint32_t>& session, const std::vector<uint64_t>& seq48,
                                const std::vector<uint8_t>* authentic = nullptr) {
        std::vector<uint8_t> accept(session.size() ? session.size() : 1, 0);
        c_.check(p2pt_dtls_replay_filter_batch(c_.get(), session.size(), session.data(), seq48.data(),
                                               authentic ? authentic->data() : nullptr, (uint32_t)latest_.size(), latest_.data(),
                                               mask_.data(), started_.data(), accept.data()));
        accept.resize(session.size());
        return accept;
    }
    uint64_t latest(size_t s) const { return latest_[s]; }
private:
    const Context& c_;
    std::vector<uint64_t> latest_, mask_;
    std::vector<uint8_t> started_;
};

// the write path's sequence numbers: every record takes its session's (epoch's) counter, fetch_add(1); more than 2^48 - 1
// is ErrSequenceNumberOverflow (ok[i] = 0 here)
class SequenceNumbers {
public:
    SequenceNumbers(const Context& c, size_t n_sessions) : c_(c), next_(n_sessions, 0) {}
    std::vector<uint64_t> assign(const std::vector<uint32_t>& session, std::vector<uint8_t>* ok_out = nullptr) {
        std::vector<uint64_t> seq(session.size() ? session.size() : 1, 0);
        std::vector<uint8_t> ok(seq.size(), 0);
        c_.check(p2pt_dtls_seq_assign_batch(c_.get(), session.size(), session.data(), (uint32_t)next_.size(), next_.data(), seq.data(), ok.data()));
        seq.resize(session.size()); ok.resize(session.size());
        if (ok_out) *ok_out = std::move(ok);
        return seq;
    }
    void restart(size_t s) { next_[s] = 0; }                    // a new epoch
    uint64_t next(size_t s) const { return next_[s]; }
private:
    const Context& c_;
    std::vector<uint64_t> next_;
};

}  // namespace dtls

// ---- SSE event scan over per-stream bodies (upstream framing of tmp/mock_llm.py:54, :68-69) ----
namespace sse {
struct Events { std::vector<uint64_t> ev_csr, ev_end; std::vector<uint8_t> done; };
inline Events scan(const Context& c, const Bytes& body, const std::vector<uint64_t>& body_off) {
    Events e;
    const uint64_t k = body_off.empty() ? 0 : body_off.size() - 1;
    const uint64_t cap = body.size() / 2 + 1;
    e.ev_csr.assign(k + 1, 0); e.ev_end.assign(cap, 0); e.done.assign(k ? k : 1, 0);
    uint64_t total = 0;
    const uint8_t dummy = 0;
    c.check(p2pt_sse_scan_batch(c.get(), k, body.empty() ? &dummy : body.data(), body_off.data(), cap, e.ev_csr.data(), e.ev_end.data(),
                                e.done.data(), &total));
    e.ev_end.resize(total); e.done.resize(k);
    return e;
}
}  // namespace sse
}  // namespace p2pt_host
