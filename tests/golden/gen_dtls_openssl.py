#!/usr/bin/env python3
"""Capture DTLS 1.2 wire records from a real, interoperable stack and commit them as a fixture.

Why: the DTLS record layout the reference uses (nonce / AAD / header length fix-up) lives in the un-vendored
crate webrtc-dtls 0.10.0 (/root/reference/tunnel/Cargo.lock:2822-2825), reached at every dc.send
(/root/reference/tunnel/src/serve.rs:274) and on_message (/root/reference/tunnel/src/rtc.rs:88-92).  Neither
the crate nor a Rust toolchain is in this image, so the layout is pinned to what webrtc-rs has to interoperate
with: records of TLS_ECDHE_ECDSA_WITH_AES_128_GCM_SHA256 as OpenSSL's DTLS 1.2 stack puts them on the wire.

How: `openssl s_server -dtls1_2` and `openssl s_client -dtls1_2` talk through a UDP relay in this process that
records every datagram.  client_random / server_random are parsed from the captured hellos, the master secret
comes from the key log, the TLS 1.2 key block (PRF-SHA256, label "key expansion", RFC 5246 6.3) gives the two
write keys (16 B) and write IVs (4 B).  Nothing here calls the repo's oracle or kernels; `cryptography` is used
once per record as a sanity check that the derived keys are the right ones (the generator refuses to write a
fixture otherwise).

Output: tests/golden/dtls_openssl_kat.json
  keys     : client/server write key + iv (hex)
  records  : every epoch >= 1 record of both directions: {dir, type, epoch, seq, rec (wire bytes), plain}
  datagrams: every captured datagram (hex) with its direction, for the datagram splitter
Run:  python tests/golden/gen_dtls_openssl.py     (needs the openssl CLI; the committed JSON does not)
"""
import hashlib
import hmac
import json
import os
import select
import socket
import subprocess
import sys
import tempfile
import threading
import time

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "dtls_openssl_kat.json")
CIPHER = "ECDHE-ECDSA-AES128-GCM-SHA256"


def frame(msg_type, sid, payload):
    """[type u8][stream_id u32 BE][payload] -- /root/reference/tunnel/src/protocol.rs:148-154"""
    return bytes([msg_type]) + sid.to_bytes(4, "big") + payload


def payloads():
    """(direction, bytes) application writes, one record each.  Sizes 1, 15, 16, 17, a real SSE frame, a
    1200-byte body frame (1205 B), 1200 B, and a 1228-byte SCTP-sized packet."""
    pat = lambda n, a, b: bytes((i * a + b) & 0xFF for i in range(n))
    sse = frame(21, 99, b'data: {"token": "hello"}\n\n')            # tmp/mock_llm.py:45-54 chunk, protocol.rs:200
    c2s = [b"\x7f", pat(15, 3, 1), pat(16, 5, 2), pat(17, 7, 3), frame(10, 42, b'{"stream_id":42,"method":"POST"}'),
           frame(11, 42, pat(1200, 13, 1)), frame(12, 42, b"")]
    s2c = [frame(20, 42, b'{"stream_id":42,"status":200}'), sse, frame(21, 42, pat(1200, 7, 3)), pat(1200, 11, 9),
           pat(1228, 17, 4), frame(22, 42, b""), frame(3, 0, b"")]
    return c2s, s2c


class Relay(threading.Thread):
    """UDP relay client <-> server on localhost that keeps every datagram it forwards."""

    def __init__(self, server_port):
        super().__init__(daemon=True)
        self.sock_c = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)   # faces the client
        self.sock_c.bind(("127.0.0.1", 0))
        self.sock_s = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)   # faces the server
        self.sock_s.bind(("127.0.0.1", 0))
        self.port = self.sock_c.getsockname()[1]
        self.server = ("127.0.0.1", server_port)
        self.client = None
        self.log = []            # (direction, bytes)
        self.stop = False

    def run(self):
        while not self.stop:
            r, _, _ = select.select([self.sock_c, self.sock_s], [], [], 0.1)
            for s in r:
                data, addr = s.recvfrom(65536)
                if s is self.sock_c:
                    self.client = addr
                    self.log.append(("c2s", data))
                    self.sock_s.sendto(data, self.server)
                elif self.client is not None:
                    self.log.append(("s2c", data))
                    self.sock_c.sendto(data, self.client)


def free_udp_port():
    s = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def split_records(dgram):
    """RFC 6347 4.1: records back to back, 13-byte header, length in bytes 11..13."""
    out, off = [], 0
    while off < len(dgram):
        assert len(dgram) - off >= 13, "truncated record header in a captured datagram"
        ln = int.from_bytes(dgram[off + 11:off + 13], "big")
        assert off + 13 + ln <= len(dgram)
        out.append(dgram[off:off + 13 + ln])
        off += 13 + ln
    return out


def p_sha256(secret, seed, n):
    """RFC 5246 5: P_hash with HMAC-SHA256."""
    out, a = b"", seed
    while len(out) < n:
        a = hmac.new(secret, a, hashlib.sha256).digest()
        out += hmac.new(secret, a + seed, hashlib.sha256).digest()
    return out[:n]


def capture(workdir):
    key, cert = os.path.join(workdir, "key.pem"), os.path.join(workdir, "cert.pem")
    subprocess.check_call(["openssl", "ecparam", "-genkey", "-name", "prime256v1", "-noout", "-out", key])
    subprocess.check_call(["openssl", "req", "-x509", "-new", "-key", key, "-out", cert, "-days", "2", "-subj",
                           "/CN=p2pt-dtls-fixture"], stderr=subprocess.DEVNULL)
    sport = free_udp_port()
    klog = os.path.join(workdir, "keys.log")
    srv = subprocess.Popen(["openssl", "s_server", "-dtls1_2", "-accept", f"127.0.0.1:{sport}", "-cert", cert, "-key", key,
                            "-cipher", CIPHER, "-keylogfile", klog, "-quiet", "-mtu", "1500"],
                           stdin=subprocess.PIPE, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    time.sleep(0.5)
    relay = Relay(sport)
    relay.start()
    cli = subprocess.Popen(["openssl", "s_client", "-dtls1_2", "-connect", f"127.0.0.1:{relay.port}", "-cipher", CIPHER,
                            "-quiet", "-mtu", "1500"],
                           stdin=subprocess.PIPE, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    # wait for the handshake: both Finished flights carry an epoch-1 record
    t0 = time.time()
    while time.time() - t0 < 10:
        e1 = {d for d, g in relay.log for r in split_records(g) if int.from_bytes(r[3:5], "big") >= 1}
        if e1 == {"c2s", "s2c"}:
            break
        time.sleep(0.05)
    else:
        raise RuntimeError("DTLS handshake did not complete")
    c2s, s2c = payloads()
    # one write() per record; the pause keeps s_client / s_server from coalescing two writes into one read
    for i in range(max(len(c2s), len(s2c))):
        if i < len(c2s):
            cli.stdin.write(c2s[i]); cli.stdin.flush()
        if i < len(s2c):
            srv.stdin.write(s2c[i]); srv.stdin.flush()
        time.sleep(0.25)
    time.sleep(0.5)
    cli.stdin.close()
    time.sleep(0.5)
    relay.stop = True
    for p in (cli, srv):
        p.kill()
        p.wait()
    relay.join()
    with open(klog) as f:
        keylog = f.read()
    return relay.log, keylog, c2s, s2c


def hello_random(log, direction, hs_type):
    """random of the ClientHello (1) / ServerHello (2): record hdr 13 + handshake hdr 12 + version 2."""
    last = None
    for d, g in log:
        if d != direction:
            continue
        for r in split_records(g):
            if r[0] == 22 and int.from_bytes(r[3:5], "big") == 0 and r[13] == hs_type and int.from_bytes(r[19:22], "big") == 0:
                last = r[13 + 12 + 2:13 + 12 + 2 + 32]
    assert last is not None and len(last) == 32
    return last


def main():
    from cryptography.hazmat.primitives.ciphers.aead import AESGCM
    with tempfile.TemporaryDirectory() as wd:
        log, keylog, c2s, s2c = capture(wd)
    crand, srand = hello_random(log, "c2s", 1), hello_random(log, "s2c", 2)
    master = None
    for line in keylog.splitlines():
        p = line.split()
        if len(p) == 3 and p[0] == "CLIENT_RANDOM" and bytes.fromhex(p[1]) == crand:
            master = bytes.fromhex(p[2])
    assert master is not None and len(master) == 48, "master secret not in the key log"
    kb = p_sha256(master, b"key expansion" + srand + crand, 40)       # RFC 5246 6.3; AEAD suites have no MAC keys
    keys = {"client_write_key": kb[0:16], "server_write_key": kb[16:32], "client_write_iv": kb[32:36], "server_write_iv": kb[36:40]}
    records = []
    for d, g in log:
        for r in split_records(g):
            epoch = int.from_bytes(r[3:5], "big")
            if epoch == 0:
                continue
            k = keys["client_write_key" if d == "c2s" else "server_write_key"]
            iv = keys["client_write_iv" if d == "c2s" else "server_write_iv"]
            L = len(r) - 13 - 8 - 16
            aad = r[3:11] + r[0:3] + L.to_bytes(2, "big")
            plain = AESGCM(k).decrypt(iv + r[13:21], r[21:], aad)       # raises if keys / layout are wrong
            records.append({"dir": d, "type": r[0], "epoch": epoch, "seq": int.from_bytes(r[5:11], "big"),
                            "rec": r.hex(), "plain": plain.hex()})
    app = {"c2s": [bytes.fromhex(x["plain"]) for x in records if x["dir"] == "c2s" and x["type"] == 23],
           "s2c": [bytes.fromhex(x["plain"]) for x in records if x["dir"] == "s2c" and x["type"] == 23]}
    assert app["c2s"] == c2s, "client writes did not map one to one onto records"
    assert app["s2c"] == s2c, "server writes did not map one to one onto records"
    multi = sum(1 for _, g in log if len(split_records(g)) > 1)
    assert multi >= 1, "no multi-record datagram captured"
    ver = subprocess.run(["openssl", "version"], capture_output=True, text=True).stdout.strip()
    fixture = {
        "source": f"{ver}: s_server/s_client -dtls1_2 -cipher {CIPHER} through a recording UDP relay (gen_dtls_openssl.py)",
        "suite": "TLS_ECDHE_ECDSA_WITH_AES_128_GCM_SHA256",
        "client_random": crand.hex(), "server_random": srand.hex(),
        "keys": {k: v.hex() for k, v in keys.items()},
        "records": records,
        "datagrams": [{"dir": d, "bytes": g.hex(), "n_records": len(split_records(g))} for d, g in log],
    }
    with open(OUT, "w") as f:
        json.dump(fixture, f, indent=1)
    print(f"wrote {OUT}: {len(records)} protected records ({len(app['c2s'])}+{len(app['s2c'])} application_data), "
          f"{len(log)} datagrams, {multi} with more than one record")


if __name__ == "__main__":
    sys.exit(main())
