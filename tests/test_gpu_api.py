"""GPU: the host-side mirror of the reference interface (protocol.TunnelMessage, dtls.CryptoGcm) --
these read like the reference's own tests (/root/reference/tunnel/src/protocol.rs:265-460) -- and the
device-pointer entry points driven from torch tensors."""
import json

import numpy as np
import pytest

import oracle_binding as ob

pytestmark = pytest.mark.gpu


def test_roundtrips_like_the_reference(ctx):
    from p2p_llm_tunnel_b200.protocol import (Agree, Hello, MAX_BODY_CHUNK, MessageType, TunnelMessage)
    d = TunnelMessage.decode(TunnelMessage.hello(Hello()).encode(ctx), ctx)
    assert d.msg_type == MessageType.Hello and d.stream_id == 0 and json.loads(d.payload)["proto"] == "httptunnel"
    d = TunnelMessage.decode(TunnelMessage.agree(Agree()).encode(ctx), ctx)
    assert d.msg_type == MessageType.Agree and json.loads(d.payload)["features"] == ["sse"]
    hdr = {"stream_id": 42, "method": "POST", "path": "/v1/chat/completions", "headers": {"content-type": "application/json"}}
    d = TunnelMessage.decode(TunnelMessage.req_headers(hdr).encode(ctx), ctx)
    assert d.msg_type == MessageType.ReqHeaders and d.stream_id == 42 and json.loads(d.payload)["method"] == "POST"
    d = TunnelMessage.decode(TunnelMessage.req_body(7, b"hello world").encode(ctx), ctx)
    assert (d.msg_type, d.stream_id, d.payload) == (MessageType.ReqBody, 7, b"hello world")
    d = TunnelMessage.decode(TunnelMessage.req_end(7).encode(ctx), ctx)
    assert (d.msg_type, d.stream_id, d.payload) == (MessageType.ReqEnd, 7, b"")
    d = TunnelMessage.decode(TunnelMessage.res_headers({"stream_id": 99, "status": 200, "headers": {}}).encode(ctx), ctx)
    assert d.msg_type == MessageType.ResHeaders and d.stream_id == 99
    sse = b'data: {"token": "hello"}\n\n'
    d = TunnelMessage.decode(TunnelMessage.res_body(99, sse).encode(ctx), ctx)
    assert (d.msg_type, d.stream_id, d.payload) == (MessageType.ResBody, 99, sse)
    d = TunnelMessage.decode(TunnelMessage.res_end(99).encode(ctx), ctx)
    assert d.msg_type == MessageType.ResEnd and d.payload == b""
    d = TunnelMessage.decode(TunnelMessage.error(5, "upstream timeout").encode(ctx), ctx)
    assert (d.msg_type, d.stream_id, d.payload) == (MessageType.Error, 5, b"upstream timeout")
    # negative / boundary cases
    for bad, text in ((b"", "too short"), (bytes([1, 0, 0]), "too short"), (bytes([255, 0, 0, 0, 0]), "unknown message type")):
        with pytest.raises(ValueError) as e:
            TunnelMessage.decode(bad, ctx)
        assert text in str(e.value)
    d = TunnelMessage.decode(bytes([12, 0, 0, 0, 0]), ctx)
    assert (d.msg_type, d.stream_id, d.payload) == (MessageType.ReqEnd, 0, b"")
    assert TunnelMessage.decode(TunnelMessage.req_end(0xFFFFFFFF).encode(ctx), ctx).stream_id == 0xFFFFFFFF
    big = bytes([0xAB]) * MAX_BODY_CHUNK
    enc = TunnelMessage.res_body(1, big).encode(ctx)
    assert len(enc) == 5 + MAX_BODY_CHUNK and TunnelMessage.decode(enc, ctx).payload == big
    assert TunnelMessage.decode(TunnelMessage.res_body(1, b"").encode(ctx), ctx).payload == b""


def test_chunker_and_demux_reassemble_a_body(ctx):
    from p2p_llm_tunnel_b200.protocol import MAX_BODY_CHUNK, TunnelMessage, chunk_body, demux_batch, encode_batch
    body_a = bytes((i * 7) & 0xFF for i in range(2 * MAX_BODY_CHUNK + 123))
    body_b = b"short body"
    msgs = chunk_body(1, body_a) + [TunnelMessage.res_body(2, body_b)]
    msgs = [msgs[0], msgs[3], msgs[1], TunnelMessage.ping(), msgs[2], TunnelMessage.res_end(2), TunnelMessage.res_end(1)]
    buf, off = encode_batch(msgs, ctx)
    frames = [buf[int(off[i]):int(off[i + 1])].tobytes() for i in range(len(msgs))]
    r = demux_batch(frames, ctx)
    assert sorted(r) == [0, 1, 2]
    assert r[1][1] == body_a and r[2][1] == body_b and r[0][1] == b""
    assert [int(m.msg_type) for m in r[1][0]] == [21, 21, 21, 22]


def test_crypto_gcm_mirror(ctx):
    from cryptography.hazmat.primitives.ciphers.aead import AESGCM
    from p2p_llm_tunnel_b200.dtls import CryptoGcm, ErrDecrypt, ErrNotEnoughRoomForNonce, record_header
    lk, rk = bytes(range(16)), bytes(range(16, 32))
    a = CryptoGcm(lk, b"\x01\x02\x03\x04", rk, b"\x05\x06\x07\x08", ctx=ctx)
    payload = b"application data over the tunnel" * 10
    raw = record_header(23, 1, 5, len(payload)) + payload
    rec = a.encrypt(None, raw, nonce=bytes(8))
    aad = raw[3:11] + raw[0:3] + len(payload).to_bytes(2, "big")
    assert rec[21:] == AESGCM(lk).encrypt(b"\x01\x02\x03\x04" + bytes(8), payload, aad)
    assert int.from_bytes(rec[11:13], "big") == 8 + len(payload) + 16
    # the peer: its local key is our remote key
    b = CryptoGcm(rk, b"\x05\x06\x07\x08", lk, b"\x01\x02\x03\x04", ctx=ctx)
    out = b.decrypt(rec)
    assert out == rec[:13] + payload
    with pytest.raises(ErrDecrypt):
        b.decrypt(rec[:-1] + bytes([rec[-1] ^ 1]))
    with pytest.raises(ErrNotEnoughRoomForNonce):
        b.decrypt(rec[:20])
    # random nonce path: two seals differ, both open
    r1, r2 = b.encrypt(None, raw), b.encrypt(None, raw)
    assert r1 != r2
    a2 = CryptoGcm(lk, b"\x01\x02\x03\x04", rk, b"\x05\x06\x07\x08", ctx=ctx)
    assert a2.decrypt(r1) == a2.decrypt(r2) == r1[:13] + payload


def test_small_calls_take_the_mapped_buffer_and_match_the_copy_path(ctx):
    """Host-pointer calls whose staging fits `small_call_bytes` (default 64 KiB) run out of a mapped pinned
    buffer - one launch, one sync, no copies; the option set to 0 forces the device-arena path.  Both must
    give the same bytes for encode, decode, seal, open and the fused calls, ragged sizes included."""
    from p2p_llm_tunnel_b200 import workload
    from p2p_llm_tunnel_b200.dtls import record_header
    rng = np.random.default_rng(0x5EED0007)
    keys, ivs = workload.sessions_keys(8)
    ctx.sessions_load(keys, ivs, keys, ivs)
    for n in (1, 2, 17):
        lens = rng.integers(0, 300, n).astype(np.uint64)
        lens[0] = 64
        pay_off = np.zeros(n + 1, np.uint64); np.cumsum(lens, out=pay_off[1:])
        payload = rng.integers(0, 256, int(pay_off[-1]), dtype=np.uint8)
        types = rng.choice(np.array([10, 11, 12, 20, 21, 22, 99], np.uint8), n)
        sids = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
        session = rng.integers(0, 8, n).astype(np.uint32)
        epoch = np.ones(n, np.uint16); seq = np.arange(n, dtype=np.uint64)
        nonce = rng.integers(0, 256, 8 * n, dtype=np.uint8)
        got = {}
        for mode in (65536, 0):
            ctx.set_option("small_call_bytes", mode)
            frames, foff = ctx.frame_encode(types, sids, payload, pay_off)
            desc = ctx.frame_decode(frames, foff)
            rec, roff, st = ctx.pipeline_seal(session, epoch, seq, nonce, types, sids, payload, pay_off)
            out, _, desc2 = ctx.pipeline_open(session, rec, roff)
            raw_off = np.zeros(n + 1, np.uint64); np.cumsum(lens + np.uint64(13), out=raw_off[1:])
            raw = np.concatenate([np.frombuffer(record_header(23, 1, i, int(lens[i])), np.uint8).tolist() +
                                  payload[int(pay_off[i]):int(pay_off[i + 1])].tolist() for i in range(n)]).astype(np.uint8)
            rec3, roff3, st3 = ctx.dtls_seal(session, nonce, raw, raw_off)
            out3, _, olen3, st4 = ctx.dtls_open(session, rec3, roff3)
            od = ctx.pipeline_open_demux(session, rec, roff, max_streams=n, gathered_cap=int(pay_off[-1]) + 8)
            got[mode] = dict(frames=frames, desc=desc, rec=rec, st=st, out=out, desc2=desc2,
                             rec3=rec3, st3=st3, out3=out3, olen3=olen3, st4=st4,
                             od_k=np.array([od["k"]]), od_desc=od["desc"], od_uniq=od["uniq_sid"], od_usess=od["uniq_session"],
                             od_csr=od["csr"], od_perm=od["perm"], od_g=od["gathered"], od_goff=od["gathered_off"])
        ctx.set_option("small_call_bytes", 262144)
        for k in got[0]:
            if k in ("out", "out3"):        # an opened record is 24 bytes shorter than its slot; the rest is undefined
                ro = roff if k == "out" else roff3
                for i in range(n):
                    a, b = int(ro[i]), int(ro[i + 1]) - 24
                    assert np.array_equal(got[0][k][a:b], got[65536][k][a:b]), (k, i)
            else:
                assert np.array_equal(got[0][k], got[65536][k]), k
        assert not got[0]["st"].any() and not got[0]["st3"].any() and not got[0]["st4"].any()
        assert not got[0]["desc2"]["status"].any() and np.array_equal(got[0]["desc2"]["stream_id"], sids)
        for i in range(n):
            a, L = int(got[0]["olen3"][i]), int(lens[i])
            assert a == 13 + L
            o = int(roff3[i])
            assert np.array_equal(got[0]["out3"][o + 13:o + a], payload[int(pay_off[i]):int(pay_off[i + 1])])
    with pytest.raises(Exception):
        ctx.set_option("small_call_bytes", 1 << 30)


def test_option_paths_give_the_same_results(ctx):
    """The non-default code paths behind p2pt_ctx_set_option: the three-launches-per-pass sort (coop_sort 0),
    host-pointer record calls cut into many chunks over the ctx's four streams (chunk_bytes), and spinning
    waits (blocking_sync 0) must all reproduce the default path's bytes."""
    from p2p_llm_tunnel_b200 import workload
    rng = np.random.default_rng(0x5EED0008)
    S, n = 64, 20000
    keys, ivs = workload.sessions_keys(S)
    ctx.sessions_load(keys, ivs, keys, ivs)
    lens = rng.integers(0, 400, n).astype(np.uint64)
    pay_off = np.zeros(n + 1, np.uint64); np.cumsum(lens, out=pay_off[1:])
    payload = rng.integers(0, 256, int(pay_off[-1]), dtype=np.uint8)
    types = rng.choice(np.array([11, 21, 21, 21, 22, 99, 0], np.uint8), n)      # 0: undecodable after open
    sids = rng.integers(1, 40, n).astype(np.uint32)
    session = np.sort(rng.integers(0, S, n)).astype(np.uint32)
    epoch = np.ones(n, np.uint16); seq = np.arange(n, dtype=np.uint64)
    nonce = rng.integers(0, 256, 8 * n, dtype=np.uint8)

    def run():
        rec, roff, st = ctx.pipeline_seal(session, epoch, seq, nonce, types, sids, payload, pay_off)
        out, _, desc = ctx.pipeline_open(session, rec, roff)
        r = ctx.frame_demux(desc, out, session=session)
        return dict(rec=rec, st=st, desc=desc, k=np.array([r["k"]]), uniq=r["uniq_sid"], usess=r["uniq_session"],
                    csr=r["csr"], perm=r["perm"], gathered=r["gathered"], goff=r["gathered_off"])

    base = run()
    assert not base["st"].any() and (base["desc"]["status"] != 0).sum() == (types == 0).sum()
    try:
        for opt, val in (("coop_sort", 0), ("chunk_bytes", 65536), ("blocking_sync", 0)):
            ctx.set_option(opt, val)
            got = run()
            for k2 in base:
                assert np.array_equal(base[k2], got[k2]), (opt, k2)
            ctx.set_option(opt, {"coop_sort": 1, "chunk_bytes": 16 << 20, "blocking_sync": 1}[opt])
    finally:
        ctx.set_option("coop_sort", 1); ctx.set_option("chunk_bytes", 16 << 20); ctx.set_option("blocking_sync", 1)


def test_gather_capacity_short_buffer_and_wide_offsets(ctx):
    """p2pt_frame_demux_batch_device with a gather buffer that is too small: nothing is written past
    gathered_cap (canary), gathered_off[k] still reports the bytes a full gather needs, and the frames that
    fit are in place.  A gathered_cap >= 4 GiB switches the per-frame destinations to 64-bit words: same
    bytes as the 32-bit path."""
    import torch
    rng = np.random.default_rng(0x5EED0009)
    n = 5000
    lens = rng.integers(0, 500, n).astype(np.uint64)
    pay_off = np.zeros(n + 1, np.uint64); np.cumsum(lens, out=pay_off[1:])
    payload = rng.integers(0, 256, int(pay_off[-1]), dtype=np.uint8)
    types = np.full(n, 21, np.uint8)
    sids = rng.integers(1, 30, n).astype(np.uint32)
    frames, foff = ctx.frame_encode(types, sids, payload, pay_off)
    desc = ctx.frame_decode(frames, foff)
    full = ctx.frame_demux(desc, frames, max_streams=64)
    total = int(full["gathered_off"][full["k"]])
    assert total == int(pay_off[-1]) and len(full["gathered"]) == total
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
    d_desc, d_frames = t(desc), t(frames)
    sb = int(ctx.lib.p2pt_frame_demux_scratch_bytes(n))
    scratch = torch.empty(sb, dtype=torch.uint8, device=dev)
    uniq = torch.empty(65, dtype=torch.int32, device=dev); csr = torch.empty(65, dtype=torch.int64, device=dev)
    goff = torch.empty(65, dtype=torch.int64, device=dev); perm = torch.empty(n, dtype=torch.int32, device=dev)
    k = torch.zeros(1, dtype=torch.int32, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    for cap, claim in ((total // 3, total // 3), (total, 1 << 33)):
        gathered = torch.full((total + 4096,), 0xA5, dtype=torch.uint8, device=dev)
        ctx.call("p2pt_frame_demux_batch_device", n, d_desc, None, d_frames, 64, uniq, None, csr, perm, k, gathered,
                 claim, goff, scratch, sb, s)
        torch.cuda.synchronize()
        g = gathered.cpu().numpy(); go = goff.cpu().numpy()
        assert int(k.item()) == full["k"] and int(go[full["k"]]) == total
        assert np.all(g[cap:] == 0xA5)                                   # nothing past the capacity / the data
        # every frame whose destination fits is there; a frame that does not fit is left out whole
        order = np.argsort(sids, kind="stable")
        dst = np.cumsum(lens[order]) - lens[order]
        for j in rng.choice(n, 400, replace=False):
            o, L = int(dst[j]), int(lens[order[j]])
            src = payload[int(pay_off[order[j]]):int(pay_off[order[j] + 1])]
            if o + L <= cap:
                assert np.array_equal(g[o:o + L], src), j
            else:
                assert np.all(g[max(o, 0):o + L][max(cap - o, 0):] == 0xA5)
        if cap == total:
            assert np.array_equal(g[:total], full["gathered"])


def test_device_pointer_entry_points(ctx):
    import torch
    from p2p_llm_tunnel_b200 import workload
    dev = torch.device("cuda:0")
    keys, ivs = workload.sessions_keys(32)
    ctx.sessions_load(keys, ivs, keys, ivs)
    w = workload.pipeline_workload(32, 8, 1200)
    n = w["n"]
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
    d = {k: t(w[k]) for k in ("session", "epoch", "seq48", "nonce8", "types", "sids", "payload", "pay_off", "rec_off")}
    rec = torch.zeros(int(w["rec_off"][-1]), dtype=torch.uint8, device=dev)
    status = torch.zeros(n, dtype=torch.uint8, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    ctx.call("p2pt_pipeline_seal_device", n, d["session"], d["epoch"], d["seq48"], d["nonce8"], d["types"], d["sids"],
             d["payload"], d["pay_off"], rec, d["rec_off"], status, s)
    torch.cuda.synchronize()
    assert int(status.sum()) == 0
    assert np.array_equal(rec.cpu().numpy(), ob.pipeline_seal(w, keys, ivs))
    out = torch.zeros_like(rec)
    desc = torch.zeros(n * 24, dtype=torch.uint8, device=dev)
    ctx.call("p2pt_pipeline_open_device", n, d["session"], rec, d["rec_off"], out, d["rec_off"], desc, s)
    scratch_bytes = int(ctx.lib.p2pt_frame_demux_scratch_bytes(n))
    scratch = torch.zeros(scratch_bytes, dtype=torch.uint8, device=dev)
    uniq = torch.zeros(n + 1, dtype=torch.int32, device=dev); csr = torch.zeros(n + 1, dtype=torch.int64, device=dev)
    perm = torch.zeros(n, dtype=torch.int32, device=dev); k = torch.zeros(1, dtype=torch.int32, device=dev)
    gathered = torch.zeros(n * 1200, dtype=torch.uint8, device=dev); goff = torch.zeros(n + 1, dtype=torch.int64, device=dev)
    usess = torch.zeros(n + 1, dtype=torch.int32, device=dev)
    ctx.call("p2pt_frame_demux_batch_device", n, desc, d["session"], out, n, uniq, usess, csr, perm, k, gathered,
             n * 1200, goff, scratch, scratch_bytes, s)
    torch.cuda.synchronize()
    from p2p_llm_tunnel_b200._lib import DESC_DTYPE
    dn = desc.cpu().numpy().view(DESC_DTYPE)
    exp = ob.frame_demux(dn, out.cpu().numpy(), w["session"])
    kk = int(k.item())
    assert kk == exp["k"] == 32 * 4
    assert np.array_equal(usess.cpu().numpy()[:kk].astype(np.uint32), exp["uniq_session"])
    assert np.array_equal(uniq.cpu().numpy()[:kk].astype(np.uint32), exp["uniq_sid"])
    assert np.array_equal(csr.cpu().numpy()[:kk + 1].astype(np.uint64), exp["csr"])
    assert np.array_equal(perm.cpu().numpy().astype(np.uint32), exp["perm"])
    assert np.array_equal(gathered.cpu().numpy(), exp["gathered"])


def test_device_kernels_never_write_outside_their_items(ctx):
    """Canary check (compute-sanitizer is not available on this pool): output slots are laid out with
    slack at odd offsets inside a buffer pre-filled with 0xCD; after the kernels ran every byte outside
    the slots must still be 0xCD, the bytes of a slot behind its item must be zero (p2pt.h: nothing
    stale may reach the caller), and the items must match the oracle."""
    import torch
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(99)
    s = torch.cuda.current_stream().cuda_stream
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
    keys = rng.integers(0, 256, (5, 16), dtype=np.uint8); ivs = rng.integers(0, 256, (5, 4), dtype=np.uint8)
    ctx.sessions_load(keys, ivs, keys, ivs)
    n = 700
    lens = rng.integers(0, 600, n).astype(np.uint64); lens[:20] = np.arange(20)
    raw_off = np.zeros(n + 1, np.uint64); np.cumsum(lens + np.uint64(13), out=raw_off[1:])
    raw = rng.integers(0, 256, int(raw_off[-1]), dtype=np.uint8)
    for i in range(n):
        o = int(raw_off[i]); raw[o] = 23; raw[o + 1] = 0xFE; raw[o + 2] = 0xFD
    session = rng.integers(0, 5, n).astype(np.uint32); nonce = rng.integers(0, 256, 8 * n, dtype=np.uint8)
    slot = lens + np.uint64(37) + rng.integers(1, 40, n).astype(np.uint64)
    rec_off = np.zeros(n + 1, np.uint64); np.cumsum(slot, out=rec_off[1:]); rec_off += np.uint64(7)
    total = int(rec_off[-1]) + 64

    def mask_of(off, item_len):
        m = np.zeros(total, bool)
        for i in range(n):
            m[int(off[i]):int(off[i]) + int(item_len[i])] = True
        return m

    for g in (4, 8, 16, 32):
        ctx.set_option("gcm_group_width", g)
        rec = torch.full((total,), 0xCD, dtype=torch.uint8, device=dev)
        status = torch.zeros(n, dtype=torch.uint8, device=dev)
        ctx.call("p2pt_dtls_gcm_seal_batch_device", n, t(session), t(nonce), t(raw), t(raw_off), rec, t(rec_off), status, s)
        torch.cuda.synchronize()
        got = rec.cpu().numpy()
        m = mask_of(rec_off, lens + np.uint64(37))
        inside = np.zeros(total, bool); inside[int(rec_off[0]):int(rec_off[-1])] = True
        assert np.all(got[~inside] == 0xCD), f"seal wrote outside its slots (G={g})"
        assert not got[inside & ~m].any(), f"seal left stale bytes behind an item (G={g})"
        exp, _, _ = ob.dtls_seal(session, keys, ivs, nonce, raw, raw_off, rec_off)
        assert np.array_equal(got[:int(rec_off[-1])][m[:int(rec_off[-1])]], exp[m[:int(rec_off[-1])]])
        # open takes the record length from the input offsets, so its input is packed tight; its
        # OUTPUT slots keep the gaps
        tight_off = np.zeros(n + 1, np.uint64); np.cumsum(lens + np.uint64(37), out=tight_off[1:])
        tight = np.concatenate([got[int(rec_off[i]):int(rec_off[i]) + int(lens[i]) + 37] for i in range(n)])
        out = torch.full((total,), 0xCD, dtype=torch.uint8, device=dev)
        olen = torch.zeros(n, dtype=torch.int32, device=dev)
        ctx.call("p2pt_dtls_gcm_open_batch_device", n, t(session), t(tight), t(tight_off), out, t(rec_off), olen, status, s)
        torch.cuda.synchronize()
        assert int(status.sum()) == 0
        go = out.cpu().numpy()
        mo = mask_of(rec_off, lens + np.uint64(13))
        assert np.all(go[~inside] == 0xCD), f"open wrote outside its slots (G={g})"
        assert not go[inside & ~mo].any(), f"open left stale bytes behind an item (G={g})"
        for i in range(0, n, 7):
            a = int(rec_off[i])
            assert np.array_equal(go[a + 13:a + 13 + int(lens[i])], raw[int(raw_off[i]) + 13:int(raw_off[i + 1])])
    ctx.set_option("gcm_group_width", 0)

    # frame encode with gapped, unaligned slots (short and long payloads mixed)
    plens = rng.integers(0, 900, n).astype(np.uint64); plens[::3] = rng.integers(0, 80, len(plens[::3]))
    pay_off = np.zeros(n + 1, np.uint64); np.cumsum(plens, out=pay_off[1:])
    payload = rng.integers(0, 256, int(pay_off[-1]), dtype=np.uint8)
    fslot = plens + np.uint64(5) + rng.integers(1, 30, n).astype(np.uint64)
    f_off = np.zeros(n + 1, np.uint64); np.cumsum(fslot, out=f_off[1:]); f_off += np.uint64(3)
    total = int(f_off[-1]) + 64
    types = np.full(n, 21, np.uint8); sids = rng.integers(1, 9, n).astype(np.uint32)
    frames = torch.full((total,), 0xCD, dtype=torch.uint8, device=dev)
    ctx.call("p2pt_frame_encode_batch_device", n, t(types), t(sids), t(payload), t(pay_off), frames, t(f_off), s)
    torch.cuda.synchronize()
    gf = frames.cpu().numpy()
    mf = mask_of(f_off, plens + np.uint64(5))
    assert np.all(gf[~mf] == 0xCD), "frame_encode wrote outside its frames"
    expf, _ = ob.frame_encode(types, sids, payload, pay_off, f_off)
    assert np.array_equal(gf[:int(f_off[-1])][mf[:int(f_off[-1])]], expf[mf[:int(f_off[-1])]])
    # the small-frame regime of the encoder (average slot <= 128 bytes): dense slots go through shared memory in
    # tiles of 128 frames, a tile with slack between its slots takes the four-lanes-per-frame copy
    for gaps in (False, True):
        n2 = 3000
        pl2 = rng.integers(0, 110, n2).astype(np.uint64); pl2[:64] = np.arange(64); pl2[500:700] = 64
        po2 = np.zeros(n2 + 1, np.uint64); np.cumsum(pl2, out=po2[1:])
        pay2 = rng.integers(0, 256, int(po2[-1]), dtype=np.uint8)
        sl2 = pl2 + np.uint64(5)
        if gaps:
            sl2[1000:1400] += rng.integers(1, 9, 400).astype(np.uint64)      # only some tiles have slack
        fo2 = np.zeros(n2 + 1, np.uint64); np.cumsum(sl2, out=fo2[1:]); fo2 += np.uint64(11)
        tot2 = int(fo2[-1]) + 64
        ty2 = rng.choice(np.array([11, 21, 20, 22], np.uint8), n2); si2 = rng.integers(0, 2**32, n2, dtype=np.uint64).astype(np.uint32)
        fr2 = torch.full((tot2,), 0xCD, dtype=torch.uint8, device=dev)
        ctx.call("p2pt_frame_encode_batch_device", n2, t(ty2), t(si2), t(pay2), t(po2), fr2, t(fo2), s)
        torch.cuda.synchronize()
        g2 = fr2.cpu().numpy()
        m2 = np.zeros(tot2, bool)
        for i in range(n2):
            m2[int(fo2[i]):int(fo2[i]) + int(pl2[i]) + 5] = True
        assert np.all(g2[~m2] == 0xCD), f"small-frame encode wrote outside its frames (gaps={gaps})"
        e2, _ = ob.frame_encode(ty2, si2, pay2, po2, fo2)
        assert np.array_equal(g2[:int(fo2[-1])][m2[:int(fo2[-1])]], e2[m2[:int(fo2[-1])]]), gaps
    # gather into a canary-filled buffer: exactly the body bytes, nothing after them
    from p2p_llm_tunnel_b200._lib import DESC_DTYPE
    tight_off = np.zeros(n + 1, np.uint64); np.cumsum(plens + np.uint64(5), out=tight_off[1:])
    tight, _ = ob.frame_encode(types, sids, payload, pay_off)
    desc = ob.frame_decode(tight, tight_off)
    body = int(plens.sum())
    gathered = torch.full((body + 256,), 0xCD, dtype=torch.uint8, device=dev)
    sb = int(ctx.lib.p2pt_frame_demux_scratch_bytes(n))
    scratch = torch.zeros(sb, dtype=torch.uint8, device=dev)
    uniq = torch.zeros(n + 1, dtype=torch.int32, device=dev); csr = torch.zeros(n + 1, dtype=torch.int64, device=dev)
    goff = torch.zeros(n + 1, dtype=torch.int64, device=dev); perm = torch.zeros(n, dtype=torch.int32, device=dev)
    k = torch.zeros(1, dtype=torch.int32, device=dev)
    ctx.call("p2pt_frame_demux_batch_device", n, t(desc.view(np.uint8)), None, t(tight), n, uniq, None, csr, perm, k, gathered,
             body, goff, scratch, sb, s)
    torch.cuda.synchronize()
    gg = gathered.cpu().numpy()
    e = ob.frame_demux(desc, tight)
    assert np.array_equal(gg[:body], e["gathered"]) and np.all(gg[body:] == 0xCD)


def test_error_codes_and_empty_batches(ctx):
    """whole-call failures are negative return codes with a message; n = 0 is a no-op everywhere"""
    import p2p_llm_tunnel_b200 as p2pt
    z8 = np.zeros(0, np.uint8); z64 = np.zeros(1, np.uint64)
    out, off = ctx.frame_encode(z8, np.zeros(0, np.uint32), z8, z64)
    assert out.size == 0 and list(off) == [0]
    assert ctx.frame_decode(z8, z64).size == 0
    r = ctx.frame_demux(np.zeros(0, p2pt._lib.DESC_DTYPE), z8)
    assert r["k"] == 0 and list(r["csr"]) == [0]
    keys = np.zeros((1, 16), np.uint8); ivs = np.zeros((1, 4), np.uint8)
    ctx.sessions_load(keys, ivs, keys, ivs)
    rec, roff, st = ctx.dtls_seal(np.zeros(0, np.uint32), z8, z8, z64)
    assert rec.size == 0 and st.size == 0
    # non-monotonic offsets
    with pytest.raises(p2pt.P2ptError) as e:
        ctx.frame_decode(np.zeros(16, np.uint8), np.array([0, 10, 5], np.uint64))
    assert e.value.code == -1 and "non-decreasing" in str(e.value)
    # output slot too small
    with pytest.raises(p2pt.P2ptError) as e:
        ctx.dtls_seal([0], np.zeros(8, np.uint8), np.zeros(50, np.uint8), np.array([0, 50], np.uint64), np.array([0, 60], np.uint64))
    assert e.value.code == -3
    # seal before any session is loaded
    fresh = p2pt.Context(0)
    with pytest.raises(p2pt.P2ptError) as e:
        fresh.dtls_seal([0], np.zeros(8, np.uint8), np.zeros(50, np.uint8), np.array([0, 50], np.uint64))
    assert e.value.code == -4 and "sessions" in str(e.value)
    with pytest.raises(p2pt.P2ptError):
        fresh.set_option("no_such_option", 1)
    assert fresh.get_info("sm_count") == 148 and fresh.get_info("nope") == -1
    fresh.close()
    # threads: two host-pointer calls on one ctx from two threads serialise and both succeed
    import threading
    res = []
    def work(seed):
        rng = np.random.default_rng(seed)
        raw = np.zeros(13 + 500, np.uint8); raw[0] = 23; raw[1] = 0xFE; raw[2] = 0xFD; raw[13:] = rng.integers(0, 256, 500)
        rec, _, st = ctx.dtls_seal([0], rng.integers(0, 256, 8, dtype=np.uint8), raw, np.array([0, 513], np.uint64))
        out, _, olen, st2 = ctx.dtls_open([0], rec, np.array([0, 537], np.uint64))
        res.append(bool(st[0] == 0 and st2[0] == 0 and np.array_equal(out[13:513], raw[13:])))
    ts = [threading.Thread(target=work, args=(i,)) for i in range(8)]
    [t.start() for t in ts]; [t.join() for t in ts]
    assert res == [True] * 8
