"""CPU: the AES-NI / PCLMULQDQ baseline (oracle/p2pt_cpu_fast.c, the timed "reference port") must
agree with the scalar oracle byte for byte before bench.py is allowed to time it."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_binding as ob

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def fast(built):
    lib = C.CDLL(os.path.join(ROOT, "oracle", "libcpufast.so"))
    if not lib.cpufast_supported():
        pytest.skip("host CPU lacks AES-NI / PCLMULQDQ")
    return lib


def _p(a):
    return C.c_void_p(a.ctypes.data)


def test_fast_seal_open_match_oracle(fast):
    rng = np.random.default_rng(9)
    ns, n = 5, 700
    keys = rng.integers(0, 256, (ns, 16), dtype=np.uint8); ivs = rng.integers(0, 256, (ns, 4), dtype=np.uint8)
    lens = rng.integers(0, 1500, n).astype(np.uint64)
    lens[:40] = np.arange(40)
    raw_off = np.zeros(n + 1, np.uint64); np.cumsum(lens + np.uint64(13), out=raw_off[1:])
    raw = rng.integers(0, 256, int(raw_off[-1]), dtype=np.uint8)
    for i in range(n):
        o = int(raw_off[i]); raw[o] = 23; raw[o + 1] = 0xFE; raw[o + 2] = 0xFD
    session = rng.integers(0, ns, n).astype(np.uint32)
    nonce = rng.integers(0, 256, 8 * n, dtype=np.uint8)
    erec, rec_off, est = ob.dtls_seal(session, keys, ivs, nonce, raw, raw_off)
    rec = np.zeros_like(erec); st = np.zeros(n, np.uint8)
    fast.cpufast_seal_batch(C.c_uint64(n), _p(session), _p(keys), _p(ivs), _p(nonce), _p(raw), _p(raw_off), _p(rec), _p(rec_off), _p(st))
    assert np.array_equal(st, est) and np.array_equal(rec, erec)
    bad = erec.copy()
    for i in range(0, n, 13):
        bad[int(rec_off[i]) + 21 + int(lens[i]) // 2] ^= 0x10
    eout, _, eolen, est = ob.dtls_open(session, keys, ivs, bad, rec_off)
    out = np.zeros_like(eout); olen = np.zeros(n, np.uint32)
    fast.cpufast_open_batch(C.c_uint64(n), _p(session), _p(keys), _p(ivs), _p(bad), _p(rec_off), _p(out), _p(rec_off), _p(olen), _p(st))
    assert np.array_equal(st, est) and np.array_equal(olen, eolen) and np.array_equal(out, eout)


def test_fast_pipeline_matches_oracle_and_counts_streams(fast):
    from p2p_llm_tunnel_b200 import workload
    keys, ivs = workload.sessions_keys(48)
    w = workload.pipeline_workload(48, 8, 333)
    # many streams per thread: the per-thread map must grow (HashMap semantics)
    big = workload.pipeline_workload(3000, 4, 20)
    bk, bi = workload.sessions_keys(3000)
    brec = np.zeros(int(big["rec_off"][-1]), np.uint8)
    t2 = (C.c_double * 2)(); s2 = (C.c_uint64 * 4)()
    assert fast.cpufast_pipeline(C.c_uint64(big["n"]), _p(big["session"]), _p(bk), _p(bi), 3000, _p(big["epoch"]),
                                 _p(big["seq48"]), _p(big["nonce8"]), _p(big["types"]), _p(big["sids"]),
                                 _p(big["payload"]), _p(big["pay_off"]), _p(brec), _p(big["rec_off"]), None, 2, 3, t2, s2) == 0
    assert s2[0] == 0 and s2[1] == 3000 * 4 and s2[2] == big["n"] * 20
    n = w["n"]
    rec = np.zeros(int(w["rec_off"][-1]), np.uint8); out = np.zeros_like(rec)
    times = (C.c_double * 2)(); stats = (C.c_uint64 * 4)()
    rc = fast.cpufast_pipeline(C.c_uint64(n), _p(w["session"]), _p(keys), _p(ivs), 48, _p(w["epoch"]), _p(w["seq48"]),
                               _p(w["nonce8"]), _p(w["types"]), _p(w["sids"]), _p(w["payload"]), _p(w["pay_off"]),
                               _p(rec), _p(w["rec_off"]), _p(out), 3, 3, times, stats)
    assert rc == 0
    assert np.array_equal(rec, ob.pipeline_seal(w, keys, ivs))
    assert stats[0] == 0 and stats[1] == 48 * 4 and stats[2] == n * 333


def test_openssl_evp_baseline_matches_oracle(fast):
    """the second timed baseline (oracle/p2pt_cpu_openssl.c: the record layout around EVP_aes_128_gcm) against
    the scalar oracle, seal, open with tampering, and the threaded pipeline"""
    if not fast.cpuossl_available():
        pytest.skip("libcpufast.so was built without libcrypto")
    rng = np.random.default_rng(10)
    ns, n = 5, 600
    keys = rng.integers(0, 256, (ns, 16), dtype=np.uint8); ivs = rng.integers(0, 256, (ns, 4), dtype=np.uint8)
    lens = rng.integers(0, 1500, n).astype(np.uint64)
    lens[:40] = np.arange(40)
    raw_off = np.zeros(n + 1, np.uint64); np.cumsum(lens + np.uint64(13), out=raw_off[1:])
    raw = rng.integers(0, 256, int(raw_off[-1]), dtype=np.uint8)
    for i in range(n):
        o = int(raw_off[i]); raw[o] = 23; raw[o + 1] = 0xFE; raw[o + 2] = 0xFD
    session = rng.integers(0, ns, n).astype(np.uint32)
    nonce = rng.integers(0, 256, 8 * n, dtype=np.uint8)
    erec, rec_off, est = ob.dtls_seal(session, keys, ivs, nonce, raw, raw_off)
    rec = np.zeros_like(erec); st = np.zeros(n, np.uint8)
    fast.cpuossl_seal_batch(C.c_uint64(n), _p(session), _p(keys), _p(ivs), _p(nonce), _p(raw), _p(raw_off), _p(rec), _p(rec_off), _p(st))
    assert np.array_equal(st, est) and np.array_equal(rec, erec)
    bad = erec.copy()
    for i in range(0, n, 13):
        bad[int(rec_off[i]) + 21 + int(lens[i]) // 2] ^= 0x10
    eout, _, eolen, est = ob.dtls_open(session, keys, ivs, bad, rec_off)
    out = np.zeros_like(eout); olen = np.zeros(n, np.uint32)
    fast.cpuossl_open_batch(C.c_uint64(n), _p(session), _p(keys), _p(ivs), _p(bad), _p(rec_off), _p(out), _p(rec_off), _p(olen), _p(st))
    assert np.array_equal(st, est) and np.array_equal(olen, eolen) and np.array_equal(out, eout)

    from p2p_llm_tunnel_b200 import workload
    keys, ivs = workload.sessions_keys(48)
    w = workload.pipeline_workload(48, 8, 333)
    rec = np.zeros(int(w["rec_off"][-1]), np.uint8)
    times = (C.c_double * 2)(); stats = (C.c_uint64 * 4)()
    rc = fast.cpufast_pipeline_ex(C.c_uint64(w["n"]), _p(w["session"]), _p(keys), _p(ivs), 48, _p(w["epoch"]), _p(w["seq48"]),
                                  _p(w["nonce8"]), _p(w["types"]), _p(w["sids"]), _p(w["payload"]), _p(w["pay_off"]),
                                  _p(rec), _p(w["rec_off"]), None, 3, 3, times, stats, 1)
    assert rc == 0
    assert np.array_equal(rec, ob.pipeline_seal(w, keys, ivs))
    assert stats[0] == 0 and stats[1] == 48 * 4 and stats[2] == w["n"] * 333
