"""Generates the committed golden fixtures.  Run from the repo root:  python tests/golden/gen_golden.py

Sources of truth:
  * frames_kat.json   -- literal vectors and round-trip values of the reference's own tests
                         (/root/reference/tunnel/src/protocol.rs:271-460), restated as bytes.
  * gcm_nist_kat.json -- FIPS-197 App. B / C.1 and GCM spec (SP 800-38D) test cases 1-4, typed in from
                         the standards and RE-DERIVED here with `cryptography` (OpenSSL); the script
                         aborts if a typed-in value disagrees.
  * dtls_records_kat.json -- DTLS 1.2 AES-128-GCM records computed with `cryptography` AESGCM around
                         the webrtc-dtls 0.10.0 record layout, independent of the
                         oracle.  The layout itself is "parity unpinned" (un-vendored crate).
"""
import json
import os

from cryptography.hazmat.primitives.ciphers import Cipher, algorithms, modes
from cryptography.hazmat.primitives.ciphers.aead import AESGCM

HERE = os.path.dirname(os.path.abspath(__file__))
h = bytes.fromhex


def frames():
    MAX_BODY_CHUNK = 65536 - 128
    enc = lambda t, sid, p: bytes([t]) + sid.to_bytes(4, "big") + p
    ok = []
    def add(name, t, sid, payload, cite):
        ok.append({"name": name, "type": t, "stream_id": sid, "payload_hex": payload.hex() if len(payload) < 256 else None,
                   "payload_fill": None if len(payload) < 256 else [payload[0], len(payload)],
                   "frame_len": 5 + len(payload), "frame_head_hex": enc(t, sid, payload)[:32].hex(), "cite": cite})
    add("req_body", 11, 7, b"hello world", "protocol.rs:321-330")
    add("req_end", 12, 7, b"", "protocol.rs:332-340")
    add("res_body_sse", 21, 99, b'data: {"token": "hello"}\n\n', "protocol.rs:360-369")
    add("res_end", 22, 99, b"", "protocol.rs:371-379")
    add("error", 99, 5, b"upstream timeout", "protocol.rs:381-388")
    add("header_only_req_end", 12, 0, b"", "protocol.rs:414-422 literal [12,0,0,0,0]")
    add("zero_stream_id", 12, 0, b"", "protocol.rs:426-432")
    add("max_stream_id", 12, 0xFFFFFFFF, b"", "protocol.rs:434-440")
    add("large_payload", 21, 1, bytes([0xAB]) * MAX_BODY_CHUNK, "protocol.rs:442-451 (65408 -> 65413 bytes)")
    add("empty_payload_body", 21, 1, b"", "protocol.rs:453-460")
    add("ping", 3, 0, b"", "protocol.rs:248-254")
    add("pong", 4, 0, b"", "protocol.rs:256-262")
    add("hello_json", 1, 0, b'{"proto":"httptunnel","min_version":1,"max_version":1,"features":["sse"]}', "protocol.rs:271-283")
    add("req_headers_sid42", 10, 42, b'{"stream_id":42,"method":"POST","path":"/v1/chat/completions","headers":{"content-type":"application/json"}}', "protocol.rs:300-318")
    bad = [
        {"name": "empty", "frame_hex": "", "status": 1, "error": "message too short: 0 bytes", "cite": "protocol.rs:392-397"},
        {"name": "too_short", "frame_hex": "010000", "status": 1, "error": "message too short: 3 bytes", "cite": "protocol.rs:399-404"},
        {"name": "unknown_type", "frame_hex": "ff00000000", "status": 2, "error": "unknown message type: 255", "cite": "protocol.rs:406-411"},
        {"name": "four_bytes", "frame_hex": "15000000", "status": 1, "error": "message too short: 4 bytes", "cite": "protocol.rs:158-160"},
        {"name": "type_zero", "frame_hex": "0000000001ff", "status": 2, "error": "unknown message type: 0", "cite": "protocol.rs:103-118"},
    ]
    return {"valid_tags": [1, 2, 3, 4, 10, 11, 12, 20, 21, 22, 99], "ok": ok, "bad": bad}


def aes_ecb(key, block):
    e = Cipher(algorithms.AES(key), modes.ECB()).encryptor()
    return e.update(block) + e.finalize()


def nist():
    aes = [
        {"name": "fips197_appB", "key": "2b7e151628aed2a6abf7158809cf4f3c", "pt": "3243f6a8885a308d313198a2e0370734",
         "ct": "3925841d02dc09fbdc118597196a0b32"},
        {"name": "fips197_appC1", "key": "000102030405060708090a0b0c0d0e0f", "pt": "00112233445566778899aabbccddeeff",
         "ct": "69c4e0d86a7b0430d8cdb78070b4c55a"},
    ]
    for v in aes:
        assert aes_ecb(h(v["key"]), h(v["pt"])).hex() == v["ct"], v["name"]
    P = ("d9313225f88406e5a55909c5aff5269a86a7a9531534f7da2e4c303d8a318a72"
         "1c3c0c95956809532fcf0e2449a6b525b16aedf5aa0de657ba637b391aafd255")
    C3 = ("42831ec2217774244b7221b784d0d49ce3aa212f2c02a4e035c17e2329aca12e"
          "21d514b25466931c7d8f6a5aac84aa051ba30b396a0aac973d58e091473f5985")
    gcm = [
        {"name": "gcm_tc1", "key": "00" * 16, "iv": "00" * 12, "aad": "", "pt": "", "ct": "",
         "tag": "58e2fccefa7e3061367f1d57a4e7455a"},
        {"name": "gcm_tc2", "key": "00" * 16, "iv": "00" * 12, "aad": "", "pt": "00" * 16,
         "ct": "0388dace60b6a392f328c2b971b2fe78", "tag": "ab6e47d42cec13bdf53a67b21257bddf"},
        {"name": "gcm_tc3", "key": "feffe9928665731c6d6a8f9467308308", "iv": "cafebabefacedbaddecaf888", "aad": "",
         "pt": P, "ct": C3, "tag": "4d5c2af327cd64a62cf35abd2ba6fab4"},
        {"name": "gcm_tc4", "key": "feffe9928665731c6d6a8f9467308308", "iv": "cafebabefacedbaddecaf888",
         "aad": "feedfacedeadbeeffeedfacedeadbeefabaddad2", "pt": P[:120], "ct": C3[:120],
         "tag": "5bc94fbc3221a5db94fae95ae7121a47"},
    ]
    for v in gcm:
        out = AESGCM(h(v["key"])).encrypt(h(v["iv"]), h(v["pt"]), h(v["aad"]))
        assert out.hex() == v["ct"] + v["tag"], v["name"]
    # GHASH H for tc2 (= AES_0(0)) as an extra fixed point
    assert aes_ecb(bytes(16), bytes(16)).hex() == "66e94bd4ef8a2c3b884cfa59ca342b2e"
    return {"aes128": aes, "gcm": gcm, "h_zero_key": "66e94bd4ef8a2c3b884cfa59ca342b2e"}


def dtls_record(key, iv4, nonce8, ctype, epoch, seq, plaintext, version=(0xFE, 0xFD)):
    L = len(plaintext)
    hdr = bytes([ctype, version[0], version[1]]) + epoch.to_bytes(2, "big") + seq.to_bytes(6, "big")
    aad = epoch.to_bytes(2, "big") + seq.to_bytes(6, "big") + bytes([ctype, version[0], version[1]]) + (L & 0xFFFF).to_bytes(2, "big")
    ct = AESGCM(key).encrypt(iv4 + nonce8, plaintext, aad)
    raw = hdr + (L & 0xFFFF).to_bytes(2, "big") + plaintext
    rec = hdr + ((8 + L + 16) & 0xFFFF).to_bytes(2, "big") + nonce8 + ct
    return raw, rec


def dtls():
    key = bytes(range(16)); iv = h("01020304")
    recs = []
    frame = bytes([0x15, 0, 0, 0, 0x63]) + b'data: {"token": "hello"}\n\n'
    def add(name, plaintext, seq=7, epoch=1, nonce=bytes(range(8)), ctype=23, k=key, s=iv, version=(0xFE, 0xFD)):
        raw, rec = dtls_record(k, s, nonce, ctype, epoch, seq, plaintext, version)
        recs.append({"name": name, "key": k.hex(), "iv": s.hex(), "nonce": nonce.hex(), "raw": raw.hex(), "rec": rec.hex()})
    add("survey_appendix_A2_example", frame)
    assert recs[0]["rec"] == ("17fefd00010000000000070037" "0001020304050607"
                              "38c4bc820b39df30be70dd0aa3eff22646ed80547c0e10efc21ea0e482efc1"
                              "5b2c93b3e5915c58f5e5dca8810d2964")
    add("empty_plaintext", b"")
    add("one_byte", b"\x7f")
    add("fifteen", bytes(range(15)))
    add("sixteen", bytes(range(16)))
    add("seventeen", bytes(range(17)))
    add("len_1200", bytes((i * 7 + 3) & 0xFF for i in range(1200)), seq=0x0000A1B2C3D4, epoch=3)
    add("len_1205_frame", bytes([21, 0, 0, 0x0F, 0xFF]) + bytes((i * 13 + 1) & 0xFF for i in range(1200)), seq=1)
    add("max_seq", b"seq48 max", seq=(1 << 48) - 1, epoch=0xFFFF, nonce=h("ffffffffffffffff"))
    add("handshake_type22_v10", b"\x01\x02\x03\x04", ctype=22, version=(0xFE, 0xFF))
    add("other_key", bytes(100), k=h("feffe9928665731c6d6a8f9467308308"), s=h("cafebabe"), nonce=h("facedbaddecaf888"))
    add("len_4096", bytes((i * 31 + 5) & 0xFF for i in range(4096)), seq=9)
    return {"records": recs}


def crc32c_vectors():
    """RFC 3720 B.4 CRC32c examples (the polynomial SCTP uses, RFC 4960 appendix B)."""
    return {"vectors": [
        {"name": "ascii_123456789", "data": b"123456789".hex(), "crc": "e3069283"},
        {"name": "32_zeros", "data": "00" * 32, "crc": "8a9136aa"},
        {"name": "32_ones", "data": "ff" * 32, "crc": "62a8ab43"},
        {"name": "32_incrementing", "data": bytes(range(32)).hex(), "crc": "46dd794e"},
        {"name": "32_decrementing", "data": bytes(range(31, -1, -1)).hex(), "crc": "113fdb5c"},
        {"name": "iscsi_read10_pdu", "data": "01c00000" + "00" * 12 + "14000000" + "00000400" + "00000014" + "00000018" + "28000000" + "00000000" + "02000000" + "00000000", "crc": "d9963a56"},
    ]}


if __name__ == "__main__":
    for name, fn in (("frames_kat.json", frames), ("gcm_nist_kat.json", nist), ("dtls_records_kat.json", dtls),
                     ("crc32c_kat.json", crc32c_vectors)):
        with open(os.path.join(HERE, name), "w") as f:
            json.dump(fn(), f, indent=1)
        print("wrote", name)
