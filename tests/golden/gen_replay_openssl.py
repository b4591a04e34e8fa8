#!/usr/bin/env python3
"""What a real DTLS 1.2 receiver does with replayed, reordered and stale records: a fixture for the anti-replay window.

Why: the window the reference relies on is webrtc-dtls 0.10.0's replaydetector (SlidingWindowDetector, 64 wide; the
crate is un-vendored, /root/reference/tunnel/Cargo.lock:2822-2825; every record of rtc.rs:88-92 passes it).  It
implements RFC 6347 4.1.2.6, and so does OpenSSL's DTLS stack (dtls1_record_replay_check, a 64-bit bitmap), which
webrtc-rs interoperates with.  This script lets OpenSSL decide: which of a scripted sequence of records a live
`openssl s_server -dtls1_2` hands to the application.

How: same relay as gen_dtls_openssl.py, but after the handshake the client's datagrams are HELD.  The client writes
one short numbered message per record; the relay then delivers the held records in a scripted order - duplicates,
reordering inside the window, jumps of more than 64, records exactly 63 and 64 behind the newest, and copies with one
ciphertext bit flipped (they fail authentication and must neither be delivered nor move the window).  What the server
prints is what it accepted.

Output: tests/golden/replay_openssl_kat.json
  arrivals: [{seq, authentic, accepted}] in arrival order, one session, epoch 1
Nothing here calls the repo's oracle or kernels.  Run: python tests/golden/gen_replay_openssl.py
"""
import json
import os
import select
import subprocess
import sys
import tempfile
import time

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from gen_dtls_openssl import CIPHER, Relay, free_udp_port, split_records      # noqa: E402

OUT = os.path.join(HERE, "replay_openssl_kat.json")
N_MSG = 200

# (sequence number, tampered?) in arrival order.  Application records start at seq 1 (seq 0 is the client's Finished).
SCRIPT = (
    [(1, 0), (2, 0), (2, 0), (4, 0), (3, 0), (3, 0), (10, 0), (5, 0), (10, 0), (9, 0), (9, 0),
     (20, 1), (20, 0), (20, 0),                      # a forged copy first: it must not mark 20 as seen
     (75, 0),                                        # newest jumps by 55
     (11, 0), (12, 0), (12, 0),                      # 64 and 63 behind 75
     (74, 0), (75, 0), (76, 0), (13, 0), (14, 0),
     (150, 1),                                       # a forged record far ahead must not move the window ...
     (15, 0), (16, 0),                               # ... so these (61 and 60 behind 76) are still inside it
     (150, 0),                                       # now it moves: jump of 74, the whole bitmap is shifted out
     (149, 0), (87, 0), (86, 0), (85, 0), (87, 0), (76, 0), (100, 0), (100, 0),
     (151, 0), (152, 0), (88, 0), (89, 0), (90, 0), (153, 0), (89, 0), (90, 0), (91, 0)]
    + [(s, 0) for s in range(154, 190)]              # a long in-order run
    + [(s, 0) for s in range(189, 120, -3)]          # and a walk backwards through and out of the window
)


class HoldingRelay(Relay):
    """forwards everything until `hold` is set; from then on client datagrams are kept in `held` instead"""

    def __init__(self, server_port):
        super().__init__(server_port)
        self.hold = False
        self.held = []

    def run(self):
        while not self.stop:
            r, _, _ = select.select([self.sock_c, self.sock_s], [], [], 0.05)
            for s in r:
                data, addr = s.recvfrom(65536)
                if s is self.sock_c:
                    self.client = addr
                    if self.hold:
                        self.held.append(data)
                    else:
                        self.log.append(("c2s", data))
                        self.sock_s.sendto(data, self.server)
                elif self.client is not None:
                    self.log.append(("s2c", data))
                    self.sock_c.sendto(data, self.client)


def drain(pipe, quiet_s):
    """everything the process has written to `pipe` until it has been silent for quiet_s"""
    out = b""
    fd = pipe.fileno()
    while True:
        r, _, _ = select.select([fd], [], [], quiet_s)
        if not r:
            return out
        chunk = os.read(fd, 65536)
        if not chunk:
            return out
        out += chunk


def main():
    with tempfile.TemporaryDirectory() as wd:
        key, cert = os.path.join(wd, "key.pem"), os.path.join(wd, "cert.pem")
        subprocess.check_call(["openssl", "ecparam", "-genkey", "-name", "prime256v1", "-noout", "-out", key])
        subprocess.check_call(["openssl", "req", "-x509", "-new", "-key", key, "-out", cert, "-days", "2", "-subj",
                               "/CN=p2pt-replay-fixture"], stderr=subprocess.DEVNULL)
        sport = free_udp_port()
        srv = subprocess.Popen(["openssl", "s_server", "-dtls1_2", "-accept", f"127.0.0.1:{sport}", "-cert", cert, "-key", key,
                                "-cipher", CIPHER, "-quiet", "-mtu", "1500"],
                               stdin=subprocess.PIPE, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        time.sleep(0.5)
        relay = HoldingRelay(sport)
        relay.start()
        cli = subprocess.Popen(["openssl", "s_client", "-dtls1_2", "-connect", f"127.0.0.1:{relay.port}", "-cipher", CIPHER,
                                "-quiet", "-mtu", "1500"],
                               stdin=subprocess.PIPE, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        t0 = time.time()
        while time.time() - t0 < 10:
            e1 = {d for d, g in relay.log for r in split_records(g) if int.from_bytes(r[3:5], "big") >= 1}
            if e1 == {"c2s", "s2c"}:
                break
            time.sleep(0.05)
        else:
            raise RuntimeError("DTLS handshake did not complete")
        time.sleep(0.3)
        relay.hold = True
        for j in range(N_MSG):                                   # message j travels in the record with seq j + 1
            cli.stdin.write(b"m%04d\n" % (j + 1)); cli.stdin.flush()
            time.sleep(0.02)
        time.sleep(0.5)
        by_seq = {}
        for g in relay.held:
            recs = split_records(g)
            assert len(recs) == 1 and recs[0][0] == 23 and int.from_bytes(recs[0][3:5], "big") == 1, "one application record per datagram"
            by_seq[int.from_bytes(recs[0][5:11], "big")] = g
        assert sorted(by_seq) == list(range(1, N_MSG + 1)), "client writes did not map one to one onto records 1..N"
        drain(srv.stdout, 0.2)
        for seq, forged in SCRIPT:
            g = bytearray(by_seq[seq])
            if forged:
                g[30] ^= 0x40                                    # one ciphertext bit (header 13 + nonce 8 + 9)
            relay.sock_s.sendto(bytes(g), relay.server)
            time.sleep(0.01)
        printed = drain(srv.stdout, 1.0)
        relay.stop = True
        for p in (cli, srv):
            p.kill()
            p.wait()
        relay.join()
    delivered = [int(x[1:]) for x in printed.decode().split()]
    assert all(1 <= d <= N_MSG for d in delivered)
    arrivals, ptr = [], 0
    for seq, forged in SCRIPT:
        acc = int(not forged and ptr < len(delivered) and delivered[ptr] == seq)
        ptr += acc
        arrivals.append({"seq": seq, "authentic": int(not forged), "accepted": acc})
    assert ptr == len(delivered), "the server printed something the script cannot account for"
    assert len(set(delivered)) == len(delivered), "a record was delivered twice"
    ver = subprocess.run(["openssl", "version"], capture_output=True, text=True).stdout.strip()
    fixture = {
        "source": f"{ver}: s_server -dtls1_2 -cipher {CIPHER}; records of one s_client held by a UDP relay and delivered "
                  "in a scripted order (gen_replay_openssl.py); accepted = the server printed the record's message",
        "window": 64, "epoch": 1,
        "arrivals": arrivals,
    }
    with open(OUT, "w") as f:
        json.dump(fixture, f, indent=1)
    na = sum(a["accepted"] for a in arrivals)
    print(f"wrote {OUT}: {len(arrivals)} arrivals, {na} accepted, {len(arrivals) - na} dropped "
          f"({sum(1 - a['authentic'] for a in arrivals)} forged)")


if __name__ == "__main__":
    sys.exit(main())
