"""GPU: BASELINE.json's full-size configurations through the C ABI, checked with size-independent
properties (round trips, permutation / stability, checksums) plus sampled records against
Python `cryptography`.  The oracle is too slow for these sizes; it covers the same code paths at
small sizes in test_gpu_parity.py."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _check_demux_properties(r, sids, n_valid, session=None):
    k = r["k"]
    csr, perm = r["csr"].astype(np.int64), r["perm"].astype(np.int64)
    assert csr[0] == 0 and csr[k] == n_valid and np.all(np.diff(csr) > 0)
    assert np.array_equal(np.sort(perm), np.arange(n_valid))                 # a permutation of the valid frames
    keys = sids[perm].astype(np.uint64)
    if session is not None:
        keys |= session[perm].astype(np.uint64) << np.uint64(32)
    starts = np.zeros(n_valid, bool); starts[csr[:k]] = True
    assert np.all(np.diff(keys.astype(np.int64))[~starts[1:]] == 0)          # one key per stream
    assert np.all(np.diff(keys.astype(np.int64))[starts[1:]] > 0)            # streams ascending
    assert np.all(np.diff(perm)[~starts[1:]] > 0)                            # arrival order kept inside a stream
    assert np.array_equal(r["uniq_sid"], (keys[csr[:k]] & np.uint64(0xFFFFFFFF)).astype(np.uint32))


def test_c2_one_million_frames_decode_demux(ctx):
    """configs[1]: 1 M RES_BODY frames x 1200 B across 4096 stream ids, decode + demux + gather."""
    from p2p_llm_tunnel_b200 import workload
    n, P = 1_000_000, 1200
    w = workload.frames_workload(n, P)
    frames, off = ctx.frame_encode(w["types"], w["sids"], w["payload"], w["pay_off"])
    assert np.array_equal(off, w["frame_off"])
    f2 = frames.reshape(n, P + 5)
    assert np.all(f2[:, 0] == 21)
    assert np.array_equal(f2[:, 1:5].copy().view(">u4").reshape(-1), w["sids"])     # big-endian stream id
    assert np.array_equal(f2[:, 5:].reshape(-1), w["payload"])
    desc = ctx.frame_decode(frames, off)
    assert not desc["status"].any() and np.all(desc["msg_type"] == 21) and np.all(desc["pay_len"] == P)
    assert np.array_equal(desc["stream_id"], w["sids"]) and np.array_equal(desc["pay_off"], off[:-1] + np.uint64(5))
    r = ctx.frame_demux(desc, frames, max_streams=4096)
    assert r["k"] == 4096 and r["uniq_sid"][0] == 1 and r["uniq_sid"][-1] == 4096
    _check_demux_properties(r, w["sids"], n)
    assert np.array_equal(r["gathered_off"], r["csr"] * np.uint64(P))
    assert np.array_equal(r["gathered"].reshape(n, P), w["payload"].reshape(n, P)[r["perm"]])


def test_c2x_corrupt_frames_are_skipped(ctx):
    from p2p_llm_tunnel_b200 import workload
    n, P = 200_000, 1200
    w = workload.frames_workload(n, P, seed_shift=9)
    frames, off = ctx.frame_encode(w["types"], w["sids"], w["payload"], w["pay_off"])
    frames = frames.copy()
    bad = (workload.splitmix64(workload.SEED_CORRUPT, n) % np.uint64(1000)) == 0        # 0.1 %
    frames[off[:-1][bad].astype(np.int64)] = 0xEE
    desc = ctx.frame_decode(frames, off)
    assert np.array_equal(desc["status"] != 0, bad) and np.all(desc["status"][bad] == 2)
    r = ctx.frame_demux(desc, frames, max_streams=4096)
    good = np.nonzero(~bad)[0]
    assert np.array_equal(np.sort(r["perm"].astype(np.int64)), good)
    assert np.array_equal(r["gathered"].reshape(-1, P), w["payload"].reshape(n, P)[r["perm"]])


def test_c2p_padded_slots_keep_what_lies_between_the_frames(ctx):
    """SURVEY.md 8(d) variant C2p: 1 M x 1200 B frames into 1216-byte slots (16-byte aligned frame starts); the
    11 bytes of every slot behind its frame keep the caller's contents (frame_encode leaves gaps alone)."""
    from p2p_llm_tunnel_b200 import workload
    n, P, stride = 1_000_000, 1200, 1216
    w = workload.frames_workload(n, P, out_stride=stride)
    import ctypes as C
    out = np.full(n * stride, 0x5A, np.uint8)
    ctx._check(ctx.lib.p2pt_frame_encode_batch(ctx.h, n, w["types"].ctypes.data, w["sids"].ctypes.data, w["payload"].ctypes.data,
                                               w["pay_off"].ctypes.data, out.ctypes.data, w["frame_off"].ctypes.data))
    f2 = out.reshape(n, stride)
    assert np.all(f2[:, 0] == 21) and np.array_equal(f2[:, 1:5].copy().view(">u4").reshape(-1), w["sids"])
    assert np.array_equal(f2[:, 5:5 + P].reshape(-1), w["payload"]) and np.all(f2[:, 5 + P:] == 0x5A)


def test_c2x_stream_lifecycles_with_corrupt_frames(ctx):
    """SURVEY.md 8(d) variant C2x at full size: 4096 streams x (ResHeaders, 242 x ResBody, ResEnd) interleaved,
    0.1 % of the frames damaged.  What proxy.rs:109-153 does frame by frame must come out of decode + demux:
    damaged frames are skipped (warn! + continue), every stream's frames keep their order, the first surviving
    frame of an undamaged stream is its ResHeaders, the last its ResEnd, and the gathered body is the
    concatenation of its surviving ResBody payloads."""
    from p2p_llm_tunnel_b200 import workload
    S, K, P = 4096, 242, 1200
    w = workload.lifecycle_workload(S, K, P)
    n = w["n"]
    frames, off = ctx.frame_encode(w["types"], w["sids"], w["payload"], w["pay_off"])
    assert np.array_equal(off, w["frame_off"])
    frames = frames.copy()
    bad = np.flatnonzero(w["corrupt"])
    assert 500 < len(bad) < 1500
    in_off = off.copy()
    starts = off[:-1].astype(np.int64)
    frames[starts[bad[::2]]] = 0xEE                                   # a tag that is no MessageType
    # truncation below five bytes: the frame keeps its first 3 bytes, the rest of its slot becomes a second,
    # also undecodable "message" - the batch just has more items
    cut = bad[1::2]
    extra = np.sort(starts[cut] + 3).astype(np.uint64)
    in_off = np.sort(np.concatenate([off, extra]))
    desc = ctx.frame_decode(frames, in_off)
    is_extra = np.isin(in_off[:-1], extra)
    is_cut_head = np.isin(in_off[:-1], starts[cut].astype(np.uint64))
    lens = (in_off[1:] - in_off[:-1]).astype(np.int64)
    exp_status = np.zeros(len(desc), np.uint8)
    exp_status[is_cut_head] = 1                                       # 3 bytes: too short
    exp_status[is_extra & (lens < 5)] = 1
    tail = is_extra & (lens >= 5)                                     # the cut-off remainder decodes as whatever its first byte says
    ok_tags = np.isin(frames[in_off[:-1][tail].astype(np.int64)], (1, 2, 3, 4, 10, 11, 12, 20, 21, 22, 99))
    exp_status[np.flatnonzero(tail)[~ok_tags]] = 2
    orig = ~is_extra                                                  # items that are (what is left of) the original frames
    bad_tag = np.zeros(n, bool); bad_tag[bad[::2]] = True
    es = exp_status.copy(); es[np.flatnonzero(orig)[bad_tag]] = 2
    assert np.array_equal(desc["status"], es)
    r = ctx.frame_demux(desc, frames, max_streams=len(desc))
    # streams 1..4096 are all there (the remainders of truncated frames may add a few stray ids)
    sid_of = {int(s): i for i, s in enumerate(r["uniq_sid"])}
    oidx = np.flatnonzero(orig)                                       # item index of original frame j
    item_to_frame = np.full(len(desc), -1, np.int64); item_to_frame[oidx] = np.arange(n)
    good = ~w["corrupt"]
    for s in (1, 2, 77, 1000, 4095, 4096):
        k = sid_of[s]
        items = r["perm"][int(r["csr"][k]):int(r["csr"][k + 1])].astype(np.int64)
        fr = item_to_frame[items]
        fr = fr[fr >= 0]                                              # stray remainders carry random ids; ignore any that landed here
        mine = np.flatnonzero((w["sids"] == s) & good)
        assert np.array_equal(fr[np.isin(fr, mine)], mine)            # arrival order kept, damaged frames gone
        if good[w["sids"] == s].all() and len(fr) == len(mine):
            assert w["types"][fr[0]] == 20 and w["types"][fr[-1]] == 22
            body = np.concatenate([w["payload"][int(w["pay_off"][j]):int(w["pay_off"][j + 1])] for j in mine if w["types"][j] == 21])
            got = r["gathered"][int(r["gathered_off"][k]):int(r["gathered_off"][k + 1])]
            assert np.array_equal(got, body)
    total_body = int(((w["types"] == 21) & good).sum()) * P
    assert int(r["gathered_off"][r["k"]]) >= total_body              # + whatever the stray remainders decoded to


@pytest.mark.parametrize("n_sessions", [1, 4096])
def test_c3_one_million_records_seal_open(ctx, n_sessions):
    """configs[2]: 1 M records x 1200 B, fixed keys / injected nonces; seal -> open round trip, sampled
    records against `cryptography`, 0.1 % tampered tags must fail and only those."""
    from cryptography.hazmat.primitives.ciphers.aead import AESGCM
    from p2p_llm_tunnel_b200 import workload
    n, L = 1_000_000, 1200
    keys, ivs = workload.sessions_keys(n_sessions, seed_shift=n_sessions)
    ctx.sessions_load(keys, ivs, keys, ivs)
    w = workload.records_workload(n, L, n_sessions=n_sessions)
    rec, rec_off, st = ctx.dtls_seal(w["session"], w["nonce8"], w["raw"], w["raw_off"], w["rec_off"])
    assert not st.any()
    raw2, rec2 = w["raw"].reshape(n, 13 + L), rec.reshape(n, 37 + L)
    assert np.array_equal(rec2[:, :11], raw2[:, :11]) and np.all(rec2[:, 11] == (L + 24) >> 8) and np.all(rec2[:, 12] == (L + 24) & 0xFF)
    assert np.array_equal(rec2[:, 13:21].reshape(-1), w["nonce8"])
    for i in [0, 1, 4095, 4096, 65537, 123456, 999_999] + list(range(500_000, 500_020)):
        s = int(w["session"][i])
        hdr = raw2[i, :13].tobytes()
        aad = hdr[3:11] + hdr[0:3] + L.to_bytes(2, "big")
        exp = AESGCM(keys[s].tobytes()).encrypt(ivs[s].tobytes() + w["nonce8"][8 * i:8 * i + 8].tobytes(), raw2[i, 13:].tobytes(), aad)
        assert rec2[i, 21:].tobytes() == exp, i
    bad = (workload.splitmix64(workload.SEED_CORRUPT + 1, n) % np.uint64(1000)) == 0
    rec = rec.copy(); rec.reshape(n, 37 + L)[bad, -1] ^= 0x01
    out, _, olen, st = ctx.dtls_open(w["session"], rec, rec_off)
    assert np.array_equal(st != 0, bad) and np.all(st[bad] == 3) and np.all(olen[~bad] == 13 + L) and np.all(olen[bad] == 0)
    out2 = out.reshape(n, 37 + L)
    assert np.array_equal(out2[~bad, 13:13 + L], raw2[~bad, 13:])           # plaintext restored
    assert np.array_equal(out2[~bad, :13], rec.reshape(n, 37 + L)[~bad, :13])  # header not rewritten on open
    assert not out2[bad, :13 + L].any()                                     # failed records are zeroed


@pytest.mark.parametrize("sse", [True, False])
def test_c4_ten_million_tiny_sse_frames(ctx, sse):
    """configs[3]: 10 M RES_BODY frames x 64 B, 4096 stream ids; payload bytes from the SSE chunk template and the
    uniform-random variant of SURVEY.md 8(d)."""
    from p2p_llm_tunnel_b200 import workload
    n, P = 10_000_000, 64
    w = workload.frames_workload(n, P, sse=sse)
    frames, off = ctx.frame_encode(w["types"], w["sids"], w["payload"], w["pay_off"])
    f2 = frames.reshape(n, P + 5)
    assert np.all(f2[:, 0] == 21) and np.array_equal(f2[:, 5:].reshape(-1), w["payload"])
    desc = ctx.frame_decode(frames, off)
    assert not desc["status"].any() and np.array_equal(desc["stream_id"], w["sids"])
    r = ctx.frame_demux(desc, frames, max_streams=4096)
    assert r["k"] == 4096
    _check_demux_properties(r, w["sids"], n)
    assert int(r["gathered"].astype(np.uint64).sum()) == int(w["payload"].astype(np.uint64).sum())
    g = r["gathered"].reshape(n, P)
    if sse:
        assert np.all(g[:, :6] == np.frombuffer(b"data: ", np.uint8))
    # the fused call (decode + demux + gather, no descriptors) on the same frames
    r2 = ctx.frame_decode_demux(frames, off, 4096, gathered_cap=n * P, want_desc=False)
    for key in ("uniq_sid", "csr", "gathered_off", "gathered"):
        assert np.array_equal(r2[key], r[key]), key


def test_large_ragged_batch_through_the_32_frame_tiles(ctx):
    """Above ~2.4 M frames the encoder and the gather switch to 32-frame tiles (copy_tile_rounds in
    p2pt.cu).  Ragged payloads of 0..40 bytes with a few long frames and a few undecodable ones, checked
    byte for byte against a numpy construction of encode (protocol.rs:148-154) and of the stable
    per-stream body concatenation (serve.rs:120-124)."""
    rng = np.random.default_rng(0x5EED0006)
    n = 2_600_000
    lens = rng.integers(0, 41, n).astype(np.int64)
    long_ix = rng.choice(n, 2000, replace=False)
    lens[long_ix] = rng.integers(257, 3000, 2000)
    types = np.full(n, 21, np.uint8)
    types[rng.choice(n, 5000, replace=False)] = 22                  # ResEnd frames carry no body
    bad_ix = rng.choice(n, 1000, replace=False)
    types[bad_ix] = 0                                               # not a MessageType (protocol.rs:103-118)
    sids = rng.integers(1, 3001, n).astype(np.uint32)
    pay_off = np.zeros(n + 1, np.uint64); np.cumsum(lens, out=pay_off[1:])
    payload = rng.integers(0, 256, int(pay_off[-1]), dtype=np.uint8)
    frames, off = ctx.frame_encode(types, sids, payload, pay_off)
    o = off[:-1].astype(np.int64)
    assert np.array_equal(off.astype(np.int64), np.concatenate([[0], np.cumsum(lens + 5)]))
    assert np.array_equal(frames[o], types)
    be = sids.astype(">u4").view(np.uint8).reshape(n, 4)
    for b in range(4):
        assert np.array_equal(frames[o + 1 + b], be[:, b])
    fidx = np.repeat(np.arange(n, dtype=np.int64), lens)            # frame of every payload byte
    assert np.array_equal(frames[np.arange(len(payload), dtype=np.int64) + 5 * (fidx + 1)], payload)
    del fidx
    desc = ctx.frame_decode(frames, off)
    bad = np.zeros(n, bool); bad[bad_ix] = True
    assert np.array_equal(desc["status"] != 0, bad) and np.all(desc["status"][bad] == 2)
    r = ctx.frame_demux(desc, frames, max_streams=3000)
    assert r["k"] == 3000
    good = np.flatnonzero(~bad)
    order = good[np.argsort(sids[good], kind="stable")]
    assert np.array_equal(r["perm"][: len(order)].astype(np.int64), order)
    body = np.where(types[order] == 21, lens[order], 0)
    starts = pay_off[:-1].astype(np.int64)[order]
    excl = np.cumsum(body) - body
    src = np.repeat(starts - excl, body) + np.arange(int(body.sum()), dtype=np.int64)
    assert np.array_equal(r["gathered"], payload[src])


@pytest.mark.parametrize("S", [8192, 65536])
def test_c5_full_pipeline_multi_session(ctx, S):
    """configs[4] at 1/8 size and at full size (65 536 sessions x 16 records = 1 048 576 records x 1200 B):
    encode->seal, open->decode->demux->gather keyed by (session, stream id).  The sealed bytes are checked on
    sampled records against Python `cryptography` laid out as SURVEY.md App. A.2 (the layout itself is pinned by
    tests/test_gpu_dtls_openssl.py), the rest by the round trip."""
    from cryptography.hazmat.primitives.ciphers.aead import AESGCM
    from p2p_llm_tunnel_b200 import workload
    R, P = 16, 1200
    keys, ivs = workload.sessions_keys(S, seed_shift=5)
    ctx.sessions_load(keys, ivs, keys, ivs)
    w = workload.pipeline_workload(S, R, P, seed_shift=5)
    n = w["n"]
    rec, rec_off, st = ctx.pipeline_seal(w["session"], w["epoch"], w["seq48"], w["nonce8"], w["types"], w["sids"],
                                         w["payload"], w["pay_off"], w["rec_off"])
    assert not st.any()
    rng = np.random.default_rng(S)
    for i in [0, 1, 15, 16, n // 2, n - 1] + [int(x) for x in rng.integers(0, n, 40)]:
        s = int(w["session"][i])
        frame = bytes([int(w["types"][i])]) + int(w["sids"][i]).to_bytes(4, "big") + \
            w["payload"][int(w["pay_off"][i]):int(w["pay_off"][i + 1])].tobytes()
        hdr = bytes([23, 0xFE, 0xFD]) + int(w["epoch"][i]).to_bytes(2, "big") + int(w["seq48"][i]).to_bytes(6, "big")
        nonce = w["nonce8"][8 * i:8 * i + 8].tobytes()
        aad = hdr[3:11] + hdr[0:3] + len(frame).to_bytes(2, "big")
        exp = hdr + (len(frame) + 24).to_bytes(2, "big") + nonce + AESGCM(keys[s].tobytes()).encrypt(ivs[s].tobytes() + nonce, frame, aad)
        a = int(rec_off[i])
        assert rec[a:a + len(exp)].tobytes() == exp, i
    r = ctx.pipeline_open_demux(w["session"], rec, rec_off, max_streams=S * 4, gathered_cap=n * P)
    assert not r["desc"]["status"].any() and r["k"] == S * 4
    _check_demux_properties(r, w["sids"], n, w["session"])
    assert np.array_equal(r["uniq_session"], np.repeat(np.arange(S, dtype=np.uint32), 4))
    assert np.array_equal(r["gathered"].reshape(n, P), w["payload"].reshape(n, P)[r["perm"]])
