// p2pt_host.hpp -- C++ host-side mirror of the reference's interface for the hot path, over the C ABI.
//
// The reference is compiled code (Rust); no Rust toolchain exists in the build image, so the host
// layer a maintainer would write in Rust is provided here in C++ with the same
// names, argument meaning and error behaviour:
//   protocol::MessageType / from_u8 / TunnelMessage::{encode, decode} + constructors
//       /root/reference/tunnel/src/protocol.rs:86-118, :139-262, limits :9-12
//   dtls::CryptoGcm::{encrypt, decrypt}
//       webrtc-dtls 0.10.0 crypto_gcm.rs (un-vendored, /root/reference/tunnel/Cargo.lock:2822-2824)
// Every encode / decode / encrypt / decrypt runs on the GPU (batch of one through the batch entry
// points; encode_batch / decode_batch for real batches).  Nothing here computes on the CPU.
#pragma once
#include <cstdint>
#include <cstring>
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

}  // namespace dtls
}  // namespace p2pt_host
