// test_protocol.cpp -- the reference's own unit tests for the codec
// (/root/reference/tunnel/src/protocol.rs:265-460), restated one for one against the C++ host mirror, plus the
// DTLS record example of SURVEY.md App. A.2.  Needs a B200: every encode / decode runs on the GPU.
// Build + run: tests/test_gpu_host_cpp.py (pytest -m gpu).
#include <cstdio>
#include <cstdlib>
#include <string>

#include "p2pt_host.hpp"

using namespace p2pt_host;
using namespace p2pt_host::protocol;

static int g_failed = 0, g_run = 0;
#define CHECK(cond)                                                                      \
    do { if (!(cond)) { std::printf("  CHECK failed: %s (%s:%d)\n", #cond, __FILE__, __LINE__); g_failed++; } } while (0)
#define TEST(name) static void name(const Context& c); \
    static struct name##_reg { name##_reg() { tests().push_back({#name, name}); } } name##_inst; \
    static void name([[maybe_unused]] const Context& c)
struct T { const char* n; void (*f)(const Context&); };
static std::vector<T>& tests() { static std::vector<T> v; return v; }
static Bytes B(const std::string& s) { return Bytes(s.begin(), s.end()); }
static bool throws_with(const Context& c, const Bytes& b, const char* what) {
    try { TunnelMessage::decode(c, b); } catch (const Error& e) { return std::string(e.what()).find(what) != std::string::npos; }
    return false;
}

// --- Encode/Decode roundtrip tests (protocol.rs:271-388) ---
TEST(test_roundtrip_hello) {
    auto enc = TunnelMessage::hello(R"({"proto":"httptunnel","min_version":1,"max_version":1,"features":["sse"]})").encode(c);
    auto d = TunnelMessage::decode(c, enc);
    CHECK(d.msg_type == MessageType::Hello); CHECK(d.stream_id == 0);
    CHECK(std::string(d.payload.begin(), d.payload.end()).find("\"proto\":\"httptunnel\"") != std::string::npos);
}
TEST(test_roundtrip_agree) {
    auto d = TunnelMessage::decode(c, TunnelMessage::agree(R"({"version":1,"features":["sse"]})").encode(c));
    CHECK(d.msg_type == MessageType::Agree);
}
TEST(test_roundtrip_req_headers) {
    auto d = TunnelMessage::decode(c, TunnelMessage::req_headers(42, R"({"stream_id":42,"method":"POST","path":"/v1/chat/completions","headers":{}})").encode(c));
    CHECK(d.msg_type == MessageType::ReqHeaders); CHECK(d.stream_id == 42);
}
TEST(test_roundtrip_req_body) {
    auto body = B("hello world");
    auto d = TunnelMessage::decode(c, TunnelMessage::req_body(7, body).encode(c));
    CHECK(d.msg_type == MessageType::ReqBody); CHECK(d.stream_id == 7); CHECK(d.payload == body);
}
TEST(test_roundtrip_req_end) {
    auto d = TunnelMessage::decode(c, TunnelMessage::req_end(7).encode(c));
    CHECK(d.msg_type == MessageType::ReqEnd); CHECK(d.stream_id == 7); CHECK(d.payload.empty());
}
TEST(test_roundtrip_res_headers) {
    auto d = TunnelMessage::decode(c, TunnelMessage::res_headers(99, R"({"stream_id":99,"status":200,"headers":{}})").encode(c));
    CHECK(d.msg_type == MessageType::ResHeaders); CHECK(d.stream_id == 99);
}
TEST(test_roundtrip_res_body) {
    auto data = B("data: {\"token\": \"hello\"}\n\n");
    auto d = TunnelMessage::decode(c, TunnelMessage::res_body(99, data).encode(c));
    CHECK(d.msg_type == MessageType::ResBody); CHECK(d.stream_id == 99); CHECK(d.payload == data);
}
TEST(test_roundtrip_res_end) {
    auto d = TunnelMessage::decode(c, TunnelMessage::res_end(99).encode(c));
    CHECK(d.msg_type == MessageType::ResEnd); CHECK(d.stream_id == 99); CHECK(d.payload.empty());
}
TEST(test_roundtrip_error) {
    auto d = TunnelMessage::decode(c, TunnelMessage::error(5, "upstream timeout").encode(c));
    CHECK(d.msg_type == MessageType::Error); CHECK(d.stream_id == 5); CHECK(d.payload == B("upstream timeout"));
}
// --- Negative: corrupt/truncated messages (protocol.rs:392-422) ---
TEST(test_decode_empty) { CHECK(throws_with(c, {}, "too short")); }
TEST(test_decode_too_short) { CHECK(throws_with(c, {1, 0, 0}, "too short")); }
TEST(test_decode_unknown_type) { CHECK(throws_with(c, {255, 0, 0, 0, 0}, "unknown message type")); }
TEST(test_decode_exactly_header_no_payload) {
    auto m = TunnelMessage::decode(c, {12, 0, 0, 0, 0});
    CHECK(m.msg_type == MessageType::ReqEnd); CHECK(m.stream_id == 0); CHECK(m.payload.empty());
}
// --- Boundary: edge cases (protocol.rs:426-460) ---
TEST(test_zero_stream_id) { CHECK(TunnelMessage::decode(c, TunnelMessage::req_end(0).encode(c)).stream_id == 0); }
TEST(test_max_stream_id) { CHECK(TunnelMessage::decode(c, TunnelMessage::req_end(UINT32_MAX).encode(c)).stream_id == UINT32_MAX); }
TEST(test_large_payload) {
    Bytes data(MAX_BODY_CHUNK, 0xAB);
    auto enc = TunnelMessage::res_body(1, data).encode(c);
    CHECK(enc.size() == 5 + MAX_BODY_CHUNK);
    auto d = TunnelMessage::decode(c, enc);
    CHECK(d.payload.size() == MAX_BODY_CHUNK); CHECK(d.payload == data);
}
TEST(test_empty_payload_body) {
    auto d = TunnelMessage::decode(c, TunnelMessage::res_body(1, {}).encode(c));
    CHECK(d.msg_type == MessageType::ResBody); CHECK(d.payload.empty());
}
// --- not in the reference's tests: wire bytes, from_u8, the chunker, the batched encoder ---
TEST(test_wire_bytes_and_from_u8) {
    auto enc = TunnelMessage::res_body(0x01020304, B("ab")).encode(c);
    CHECK((enc == Bytes{21, 1, 2, 3, 4, 'a', 'b'}));
    int valid = 0;
    for (int v = 0; v < 256; v++) valid += from_u8((uint8_t)v).has_value();
    CHECK(valid == 11); CHECK(!from_u8(0).has_value()); CHECK(from_u8(99).value() == MessageType::Error);
}
TEST(test_chunker_and_batch) {
    Bytes body(2 * MAX_BODY_CHUNK + 100);
    for (size_t i = 0; i < body.size(); i++) body[i] = (uint8_t)(i * 7);
    auto msgs = chunk_body(9, body);
    CHECK(msgs.size() == 3); CHECK(msgs[2].payload.size() == 100);
    std::vector<uint64_t> off;
    auto frames = encode_batch(c, msgs, off);
    CHECK(off.back() == body.size() + 15);
    Bytes f1(frames.begin() + (ptrdiff_t)off[1], frames.begin() + (ptrdiff_t)off[2]);
    auto d = TunnelMessage::decode(c, f1);
    CHECK(d.stream_id == 9); CHECK(d.payload == msgs[1].payload);
}
// --- DTLS record: the worked example of SURVEY.md App. A.2 (computed with OpenSSL via `cryptography`) ---
TEST(test_crypto_gcm_known_record) {
    uint8_t key[16], iv[4] = {1, 2, 3, 4}, nonce[8];
    for (int i = 0; i < 16; i++) key[i] = (uint8_t)i;
    for (int i = 0; i < 8; i++) nonce[i] = (uint8_t)i;
    dtls::CryptoGcm gcm(c, key, iv, key, iv);
    auto frame = TunnelMessage::res_body(99, B("data: {\"token\": \"hello\"}\n\n")).encode(c);
    Bytes raw = {23, 0xfe, 0xfd, 0, 1, 0, 0, 0, 0, 0, 7, 0, (uint8_t)frame.size()};
    raw.insert(raw.end(), frame.begin(), frame.end());
    auto rec = gcm.encrypt(raw, nonce);
    const char* expect = "17fefd00010000000000070037000102030405060738c4bc820b39df30be70dd0aa3eff22646ed80547c0e10efc21ea0e482efc15b2c93b3e5915c58f5e5dca8810d2964";
    std::string hex;
    for (uint8_t b : rec) { char t[3]; std::snprintf(t, 3, "%02x", b); hex += t; }
    CHECK(hex == expect);
    auto plain = gcm.decrypt(rec);
    CHECK(plain.size() == raw.size()); CHECK(Bytes(plain.begin() + 13, plain.end()) == frame);
    rec.back() ^= 1;
    bool failed = false;
    try { gcm.decrypt(rec); } catch (const dtls::ErrDecrypt&) { failed = true; }
    CHECK(failed);
    bool noroom = false;
    try { gcm.decrypt(Bytes(rec.begin(), rec.begin() + 20)); } catch (const dtls::ErrNotEnoughRoomForNonce&) { noroom = true; }
    CHECK(noroom);
}

// --- the two reader loops, batched (reader::ProxyReader = proxy.rs:109-170, reader::ServeReassembler = serve.rs:82-156),
//     against the same loops written frame by frame the way the reference runs them ---
static Bytes F(uint8_t t, uint32_t sid, const Bytes& p = {}) {
    Bytes f = {t, (uint8_t)(sid >> 24), (uint8_t)(sid >> 16), (uint8_t)(sid >> 8), (uint8_t)sid};
    f.insert(f.end(), p.begin(), p.end());
    return f;
}
static uint32_t lcg(uint32_t& s) { s = s * 1664525u + 1013904223u; return s >> 8; }
static bool valid_tag(uint8_t b) { return from_u8(b).has_value(); }

TEST(test_proxy_reader_matches_the_per_frame_loop) {
    using namespace p2pt_host::reader;
    uint32_t seed = 5;
    std::vector<std::vector<Bytes>> streams;
    for (uint32_t sid = 1; sid <= 30; sid++) {
        std::vector<Bytes> seq;
        seq.push_back(F(20, sid, B("{\"stream_id\":" + std::to_string(sid) + ",\"status\":200,\"headers\":{\"stream_id\":\"7\"}}")));
        const int nb = (int)(lcg(seed) % 9);
        for (int b = 0; b < nb; b++) {
            Bytes body(lcg(seed) % 300);
            for (auto& x : body) x = (uint8_t)lcg(seed);
            seq.push_back(F(21, sid, body));
        }
        seq.push_back(sid % 5 ? F(22, sid) : F(99, sid, B("upstream timeout")));
        if (sid % 7 == 0) seq.push_back(F(21, sid, B("after the end")));            // dropped: stream already removed
        streams.push_back(seq);
    }
    std::vector<Bytes> inter;                                                       // interleave, each stream's own order kept
    std::vector<size_t> at(streams.size(), 0);
    for (size_t left = 1; left;) {
        left = 0;
        const size_t j = lcg(seed) % streams.size();
        if (at[j] < streams[j].size()) inter.push_back(streams[j][at[j]++]);
        for (size_t q = 0; q < streams.size(); q++) left += at[q] < streams[q].size();
    }
    const std::vector<Bytes> extra = {F(3, 0), Bytes{250, 0, 0, 0, 1}, Bytes{0x15, 0}, F(4, 0), F(21, 9999, B("unknown stream")),
                                      F(20, 3, B("{not json")), F(20, 77, B("{\"stream_id\":2,\"status\":404,\"headers\":{}}"))};
    inter.insert(inter.begin() + 10, extra.begin(), extra.end());
    // per-frame loop (proxy.rs:109-170)
    std::map<uint32_t, std::vector<StreamEvent>> pending, closed;
    size_t pongs = 0, fails = 0;
    for (uint32_t s = 1; s <= 30; s++) if (s != 13) pending[s];
    for (const Bytes& f : inter) {
        if (f.size() < 5 || !valid_tag(f[0])) { fails++; continue; }
        const uint8_t t = f[0];
        uint32_t sid = ((uint32_t)f[1] << 24) | ((uint32_t)f[2] << 16) | ((uint32_t)f[3] << 8) | f[4];
        const Bytes p(f.begin() + 5, f.end());
        if (t == 20) {
            if (!json_looks_complete(p.data(), p.size())) continue;
            auto js = json_u32_field(p.data(), p.size(), "stream_id");
            if (!js) continue;
            auto it = pending.find(*js);
            if (it != pending.end()) it->second.push_back({StreamEvent::Headers, p});
        } else if (t == 21) { auto it = pending.find(sid); if (it != pending.end()) it->second.push_back({StreamEvent::Body, p}); }
        else if (t == 22 || t == 99) {
            auto it = pending.find(sid);
            if (it != pending.end()) { it->second.push_back({t == 22 ? StreamEvent::End : StreamEvent::Error, t == 22 ? Bytes() : p}); closed[sid] = it->second; pending.erase(it); }
        } else if (t == 3) pongs++;
    }
    ProxyReader rd(c);
    for (uint32_t s = 1; s <= 30; s++) if (s != 13) rd.register_stream(s);
    const size_t cut = inter.size() / 2;
    rd.feed(std::vector<Bytes>(inter.begin(), inter.begin() + (ptrdiff_t)cut));     // two batches: state carries over
    rd.feed(std::vector<Bytes>(inter.begin() + (ptrdiff_t)cut, inter.end()));
    CHECK(rd.pongs_to_send == pongs); CHECK(pongs == 1); CHECK(rd.decode_failures == fails); CHECK(fails == 2);
    CHECK(rd.pending == pending); CHECK(rd.closed == closed);
    CHECK(closed.size() >= 28);
}

TEST(test_serve_reassembler_matches_the_per_frame_loop) {
    using namespace p2pt_host::reader;
    uint32_t seed = 6;
    std::vector<std::vector<Bytes>> streams;
    std::map<uint32_t, Bytes> bodies;
    for (uint32_t sid = 1; sid <= 20; sid++) {
        Bytes body(sid == 3 ? 2 * MAX_BODY_CHUNK + 777 : lcg(seed) % 2000);
        for (auto& x : body) x = (uint8_t)lcg(seed);
        bodies[sid] = body;
        std::vector<Bytes> seq;
        seq.push_back(F(10, sid, B("{\"stream_id\":" + std::to_string(sid) + ",\"method\":\"POST\",\"path\":\"/v1/x\",\"headers\":{}}")));
        for (const auto& m : chunk_body(sid, body, false)) seq.push_back(F(11, sid, m.payload));
        if (sid != 9) seq.push_back(F(12, sid));                                    // stream 9 stays pending
        streams.push_back(seq);
    }
    std::vector<Bytes> inter;
    std::vector<size_t> at(streams.size(), 0);
    for (size_t left = 1; left;) {
        left = 0;
        const size_t j = lcg(seed) % streams.size();
        if (at[j] < streams[j].size()) inter.push_back(streams[j][at[j]++]);
        for (size_t q = 0; q < streams.size(); q++) left += at[q] < streams[q].size();
    }
    const std::vector<Bytes> extra = {F(11, 555, B("body for an unknown request")), F(3, 0), F(12, 556), Bytes{0, 1, 2}};
    inter.insert(inter.begin() + 5, extra.begin(), extra.end());
    std::vector<uint32_t> end_order;                                                // REQ_END arrival order
    for (const Bytes& f : inter) if (f.size() >= 5 && f[0] == 12) { uint32_t sid = ((uint32_t)f[3] << 8) | f[4]; if (sid <= 20) end_order.push_back(sid); }
    ServeReassembler rs(c);
    std::vector<ServeReassembler::Request> done;
    const size_t cut = inter.size() / 3;
    for (int part = 0; part < 3; part++) {
        auto a = inter.begin() + (ptrdiff_t)(part * cut), b = part == 2 ? inter.end() : inter.begin() + (ptrdiff_t)((part + 1) * cut);
        for (auto& r : rs.feed(std::vector<Bytes>(a, b))) done.push_back(std::move(r));
    }
    CHECK(rs.pongs_to_send == 1); CHECK(rs.decode_failures == 1);
    CHECK(done.size() == end_order.size()); CHECK(done.size() == 19);
    for (size_t i = 0; i < done.size() && i < end_order.size(); i++) {
        auto sid = json_u32_field((const uint8_t*)done[i].headers_json.data(), done[i].headers_json.size(), "stream_id");
        CHECK(sid && *sid == end_order[i]);
        CHECK(done[i].body == bodies[end_order[i]]);
    }
    CHECK(rs.pending_requests.size() == 1 && rs.pending_requests.count(9) == 1);
}

// --- record-layer bookkeeping and the SSE scan against scalar restatements written here (not the oracle: this file is
//     product-side) ---
TEST(test_replay_detector_matches_the_sliding_window) {
    // the window as replaydetector keeps it: latest, a 64-bit mask of seen numbers behind it
    struct Win { bool started = false; uint64_t latest = 0, mask = 0;
                 bool check(uint64_t q) const { if (!started || q > latest) return true; if (latest - q >= 64) return false; return !((mask >> (latest - q)) & 1); }
                 void accept(uint64_t q) { if (!started) { started = true; latest = q; mask = 1; }
                                           else if (q > latest) { uint64_t sh = q - latest; mask = sh >= 64 ? 0 : mask << sh; mask |= 1; latest = q; }
                                           else mask |= 1ull << (latest - q); } };
    const size_t S = 5;
    std::vector<Win> ref(S);
    dtls::ReplayDetector det(c, S);
    uint64_t x = 88172645463325252ull;
    auto rnd = [&]() { x ^= x << 13; x ^= x >> 7; x ^= x << 17; return x; };
    std::vector<uint64_t> nxt(S, 0);
    for (int batch = 0; batch < 3; batch++) {
        std::vector<uint32_t> sess; std::vector<uint64_t> seq; std::vector<uint8_t> auth, want;
        for (int i = 0; i < 4000; i++) {
            const uint32_t s = (uint32_t)(rnd() % S);
            const uint64_t r = rnd() % 100;
            uint64_t q;
            if (r < 60) q = nxt[s]++;
            else if (r < 85) { const uint64_t back = rnd() % 80; q = nxt[s] > back ? nxt[s] - back : 0; }
            else { nxt[s] += rnd() % 90; q = nxt[s]++; }
            const uint8_t a = (rnd() % 25) != 0;
            sess.push_back(s); seq.push_back(q); auth.push_back(a);
            const bool ok = ref[s].check(q) && a;
            if (ok) ref[s].accept(q);
            want.push_back(ok);
        }
        CHECK(det.filter(sess, seq, &auth) == want);
        for (size_t s = 0; s < S; s++) CHECK(det.latest(s) == ref[s].latest);
    }
}
TEST(test_sequence_numbers_count_per_session) {
    dtls::SequenceNumbers sn(c, 4);
    const std::vector<uint32_t> sess = {2, 0, 2, 2, 1, 0, 7, 2};
    std::vector<uint8_t> ok;
    auto seq = sn.assign(sess, &ok);
    CHECK((seq == std::vector<uint64_t>{0, 0, 1, 2, 0, 1, 0, 3}));
    CHECK((ok == std::vector<uint8_t>{1, 1, 1, 1, 1, 1, 0, 1}));                     // session 7 is not in the table
    seq = sn.assign({2, 2, 3});
    CHECK((seq == std::vector<uint64_t>{4, 5, 0}));
    CHECK(sn.next(2) == 6 && sn.next(0) == 2 && sn.next(3) == 1);
    sn.restart(2);
    CHECK(sn.assign({2})[0] == 0);
}
TEST(test_unpack_datagrams_and_drop_the_truncated_one) {
    auto rec = [](uint8_t type, size_t len) { Bytes r = {type, 0xfe, 0xfd, 0, 1, 0, 0, 0, 0, 0, 9, (uint8_t)(len >> 8), (uint8_t)len}; r.resize(13 + len, 0x5a); return r; };
    Bytes buf; std::vector<uint64_t> off = {0};
    auto dgram = [&](std::initializer_list<Bytes> recs, size_t cut = 0) { for (auto& r : recs) buf.insert(buf.end(), r.begin(), r.end()); buf.resize(buf.size() - cut); off.push_back(buf.size()); };
    dgram({rec(22, 40), rec(20, 1), rec(22, 64)});
    dgram({rec(23, 1200)});
    dgram({rec(23, 100), rec(23, 50)}, 7);                                          // the second record is cut short
    dgram({rec(23, 10), rec(23, 0)});                                               // a trailing bare header: dropped as well
    auto u = dtls::unpack_datagrams(c, buf, off);
    CHECK((u.dgram_status == std::vector<uint8_t>{0, 0, P2PT_ST_BAD_HEADER, P2PT_ST_BAD_HEADER}));
    CHECK((u.rec_dgram == std::vector<uint32_t>{0, 0, 0, 1}));
    CHECK(u.rec_off.size() == 5 && u.rec_off[0] == 0 && u.rec_off[1] == 53 && u.rec_off[2] == 67 && u.rec_off[3] == off[1]);
}
TEST(test_sse_scan_finds_the_events_of_every_stream) {
    const std::vector<std::string> streams = {"data: {\"a\":1}\n\ndata: {\"b\":2}\n\ndata: [DONE]\n\n", "", "\n", "data: {\"c\":3}\n\ntail without an end",
                                              "data: [DONE]\n\n", "x data: [DONE]\n\n"};
    Bytes body; std::vector<uint64_t> off = {0};
    for (auto& s : streams) { body.insert(body.end(), s.begin(), s.end()); off.push_back(body.size()); }
    auto e = sse::scan(c, body, off);
    std::vector<uint64_t> want_csr = {0}, want_end; std::vector<uint8_t> want_done;
    for (size_t s = 0; s < streams.size(); s++) {
        const std::string& t = streams[s];
        bool done = t.compare(0, 14, "data: [DONE]\n\n") == 0;
        for (size_t i = 1; i < t.size(); i++)
            if (t[i] == '\n' && t[i - 1] == '\n') { want_end.push_back(off[s] + i + 1); done |= t.compare(i + 1, 14, "data: [DONE]\n\n") == 0; }
        want_csr.push_back(want_end.size()); want_done.push_back(done);
    }
    CHECK(e.ev_csr == want_csr); CHECK(e.ev_end == want_end); CHECK(e.done == want_done);
    CHECK((want_done == std::vector<uint8_t>{1, 0, 0, 0, 1, 0}));
}

int main() {
    try {
        Context c(0);
        for (auto& t : tests()) { int before = g_failed; t.f(c); g_run++; std::printf("%s %s\n", g_failed == before ? "ok  " : "FAIL", t.n); }
    } catch (const std::exception& e) { std::printf("fatal: %s\n", e.what()); return 2; }
    std::printf("%d tests, %d failed checks\n", g_run, g_failed);
    return g_failed ? 1 : 0;
}
