"""ctypes view of oracle/liboracle.so (the CPU oracle).  Test infrastructure only."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_LIB = os.path.join(ORACLE_DIR, "liboracle.so")

DESC_DTYPE = np.dtype([("pay_off", "<u8"), ("pay_len", "<u4"), ("stream_id", "<u4"),
                       ("msg_type", "u1"), ("status", "u1"), ("pad", "u1", (6,))])

_lib = None


def lib():
    global _lib
    if _lib is None:
        src = os.path.join(ORACLE_DIR, "p2pt_oracle.c")
        if not os.path.exists(ORACLE_LIB) or os.path.getmtime(ORACLE_LIB) < os.path.getmtime(src):
            subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-o", ORACLE_LIB, src])
        _lib = C.CDLL(ORACLE_LIB)
        _lib.orc_frame_demux_batch.restype = C.c_uint32
    return _lib


def _p(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


def _u8(a):
    return np.ascontiguousarray(a, np.uint8)


def frame_encode(types, sids, payload, pay_off, out_off=None):
    types = _u8(types); sids = np.ascontiguousarray(sids, np.uint32)
    payload = _u8(payload); pay_off = np.ascontiguousarray(pay_off, np.uint64)
    n = len(types)
    if out_off is None:
        out_off = np.zeros(n + 1, np.uint64)
        np.cumsum((pay_off[1:] - pay_off[:-1]) + np.uint64(5), out=out_off[1:])
    out_off = np.ascontiguousarray(out_off, np.uint64)
    out = np.zeros(int(out_off[-1]) if n else 0, np.uint8)
    lib().orc_frame_encode_batch(C.c_uint64(n), _p(types), _p(sids), _p(payload), _p(pay_off), _p(out), _p(out_off))
    return out, out_off


def frame_decode(buf, in_off):
    buf = _u8(buf); in_off = np.ascontiguousarray(in_off, np.uint64)
    n = len(in_off) - 1
    desc = np.zeros(n, DESC_DTYPE)
    lib().orc_frame_decode_batch(C.c_uint64(n), _p(buf), _p(in_off), _p(desc))
    return desc


def frame_demux(desc, buf, session=None):
    desc = np.ascontiguousarray(desc, DESC_DTYPE); buf = _u8(buf)
    n = len(desc)
    if session is not None:
        session = np.ascontiguousarray(session, np.uint32)
    usess = np.zeros(n + 1, np.uint32)
    uniq = np.zeros(n + 1, np.uint32); csr = np.zeros(n + 1, np.uint64); perm = np.zeros(max(n, 1), np.uint32)
    body = (desc["status"] == 0) & ((desc["msg_type"] == 11) | (desc["msg_type"] == 21))
    gathered = np.zeros(max(int(desc["pay_len"][body].sum()), 1), np.uint8)
    goff = np.zeros(n + 1, np.uint64)
    k = lib().orc_frame_demux_batch(C.c_uint64(n), _p(desc), _p(session), _p(buf), _p(uniq), _p(usess), _p(csr),
                                    _p(perm), _p(gathered), _p(goff))
    m = int(csr[k])
    return {"k": k, "uniq_sid": uniq[:k].copy(), "uniq_session": usess[:k].copy(), "csr": csr[:k + 1].copy(), "perm": perm[:m].copy(),
            "gathered_off": goff[:k + 1].copy(), "gathered": gathered[:int(goff[k])].copy()}


def dtls_seal(session, keys, ivs, nonce8, raw, raw_off, rec_off=None):
    session = np.ascontiguousarray(session, np.uint32); keys = _u8(keys); ivs = _u8(ivs)
    nonce8 = _u8(nonce8); raw = _u8(raw); raw_off = np.ascontiguousarray(raw_off, np.uint64)
    n = len(session)
    if rec_off is None:
        rec_off = np.zeros(n + 1, np.uint64)
        np.cumsum((raw_off[1:] - raw_off[:-1]) + np.uint64(24), out=rec_off[1:])
    rec_off = np.ascontiguousarray(rec_off, np.uint64)
    rec = np.zeros(int(rec_off[-1]) if n else 0, np.uint8)
    status = np.zeros(n, np.uint8)
    lib().orc_dtls_seal_batch(C.c_uint64(n), _p(session), _p(keys), _p(ivs), _p(nonce8), _p(raw), _p(raw_off),
                              _p(rec), _p(rec_off), _p(status))
    return rec, rec_off, status


def dtls_open(session, keys, ivs, rec, rec_off, out_off=None):
    session = np.ascontiguousarray(session, np.uint32); keys = _u8(keys); ivs = _u8(ivs)
    rec = _u8(rec); rec_off = np.ascontiguousarray(rec_off, np.uint64)
    n = len(session)
    if out_off is None:
        out_off = rec_off
    out_off = np.ascontiguousarray(out_off, np.uint64)
    out = np.zeros(int(out_off[-1]) if n else 0, np.uint8)
    out_len = np.zeros(n, np.uint32); status = np.zeros(n, np.uint8)
    lib().orc_dtls_open_batch(C.c_uint64(n), _p(session), _p(keys), _p(ivs), _p(rec), _p(rec_off),
                              _p(out), _p(out_off), _p(out_len), _p(status))
    return out, out_off, out_len, status


def pipeline_raw(w):
    """frames -> raw records (hdr13 || frame) exactly as serve.rs:269-276 + the DTLS record layer
    would build them: content type 23, version fe fd, epoch, seq48, length."""
    frames, foff = frame_encode(w["types"], w["sids"], w["payload"], w["pay_off"])
    n = w["n"]
    flen = (foff[1:] - foff[:-1]).astype(np.uint64)
    raw_off = np.zeros(n + 1, np.uint64)
    np.cumsum(flen + np.uint64(13), out=raw_off[1:])
    raw = np.zeros(int(raw_off[-1]), np.uint8)
    for i in range(n):
        o = int(raw_off[i]); L = int(flen[i])
        raw[o:o + 13] = np.frombuffer(bytes([23, 0xFE, 0xFD]) + int(w["epoch"][i]).to_bytes(2, "big") +
                                      int(w["seq48"][i]).to_bytes(6, "big") + (L & 0xFFFF).to_bytes(2, "big"), np.uint8)
        raw[o + 13:o + 13 + L] = frames[int(foff[i]):int(foff[i + 1])]
    return raw, raw_off


def pipeline_seal(w, keys, ivs):
    raw, raw_off = pipeline_raw(w)
    rec, _, status = dtls_seal(w["session"], keys, ivs, w["nonce8"], raw, raw_off, w.get("rec_off"))
    assert (status == 0).all()
    return rec


SCTP_CHUNK_DTYPE = np.dtype([("tsn", "<u4"), ("stream_id", "<u2"), ("ssn", "<u2"), ("ppid", "<u4"),
                             ("flags", "u1"), ("status", "u1"), ("pad", "u1", (2,))])


def crc32c(data):
    lib().orc_crc32c.restype = C.c_uint32
    return lib().orc_crc32c(bytes(data), C.c_size_t(len(data)))


def sctp_packetize(msgs, src_port, dst_port, vtag, sid, ssn, ppid, tsn0, epoch=None, seq48_first=None):
    """list of byte strings -> (packet bytes, pkt_off) with the oracle; scalars broadcast per message"""
    L = lib()
    L.orc_sctp_packetize.restype = C.c_uint64
    L.orc_sctp_packets_bytes.restype = C.c_uint64
    L.orc_sctp_fragments.restype = C.c_uint64
    n = len(msgs)
    bc = lambda x: np.broadcast_to(np.asarray(x), (n,))
    src_port, dst_port, vtag, sid, ssn, ppid, tsn0 = map(bc, (src_port, dst_port, vtag, sid, ssn, ppid, tsn0))
    hdr = epoch is not None
    if hdr:
        epoch, seq48_first = bc(epoch), bc(seq48_first)
    chunks, offs = [], [0]
    for i, m in enumerate(msgs):
        k = int(L.orc_sctp_fragments(C.c_uint64(len(m))))
        cap = int(L.orc_sctp_packets_bytes(C.c_uint64(len(m)))) + (13 * k if hdr else 0)
        out = C.create_string_buffer(cap)
        w = L.orc_sctp_packetize(bytes(m), C.c_uint64(len(m)), C.c_uint16(int(src_port[i])), C.c_uint16(int(dst_port[i])),
                                 C.c_uint32(int(vtag[i])), C.c_uint16(int(sid[i])), C.c_uint16(int(ssn[i])),
                                 C.c_uint32(int(ppid[i])), C.c_uint32(int(tsn0[i])), C.c_int(1 if hdr else 0),
                                 C.c_uint16(int(epoch[i]) if hdr else 0), C.c_uint64(int(seq48_first[i]) if hdr else 0), out)
        assert w == cap
        chunks.append(out.raw)
        # per-packet offsets: all but the last fragment are 1228 (+13) bytes
        for f in range(k):
            ln = len(m) - 1200 * f
            fl = min(max(ln, 0), 1200)
            offs.append(offs[-1] + 28 + ((fl + 3) & ~3) + (13 if hdr else 0))
    return np.frombuffer(b"".join(chunks), np.uint8), np.array(offs, np.uint64)


def sctp_unpack(buf, in_off, headroom=0):
    buf = _u8(buf); in_off = np.ascontiguousarray(in_off, np.uint64)
    n = len(in_off) - 1
    info = np.zeros(n, SCTP_CHUNK_DTYPE); po = np.zeros(n, np.uint64); pl = np.zeros(n, np.uint32)
    for i in range(n):
        a, b = int(in_off[i]), int(in_off[i + 1])
        one = np.zeros(1, SCTP_CHUNK_DTYPE); o = C.c_uint64(0); ln = C.c_uint32(0)
        lib().orc_sctp_unpack(C.c_void_p(buf.ctypes.data + a), C.c_uint64(b - a), C.c_uint32(headroom), _p(one), C.byref(o), C.byref(ln))
        info[i] = one[0]; po[i] = a + o.value; pl[i] = ln.value
    return info, po, pl


def dtls_datagram_split(buf, dgram_off):
    buf = _u8(buf); dgram_off = np.ascontiguousarray(dgram_off, np.uint64)
    lib().orc_dtls_datagram_split.restype = C.c_int64
    offs, dg, status = [], [], []
    for i in range(len(dgram_off) - 1):
        a, b = int(dgram_off[i]), int(dgram_off[i + 1])
        tmp = np.zeros(max((b - a) // 13 + 1, 1), np.uint64)
        k = lib().orc_dtls_datagram_split(C.c_void_p(buf.ctypes.data + a), C.c_uint64(b - a), _p(tmp), C.c_uint64(len(tmp)))
        status.append(6 if k < 0 else 0)
        for j in range(max(k, 0)):
            offs.append(a + int(tmp[j])); dg.append(i)
    return np.array(offs + [int(dgram_off[-1])], np.uint64), np.array(dg, np.uint32), np.array(status, np.uint8)


def replay_filter(session, seq48, state, authentic=None):
    session = np.ascontiguousarray(session, np.uint32); seq48 = np.ascontiguousarray(seq48, np.uint64)
    n = len(session)
    auth = np.ones(n, np.uint8) if authentic is None else np.ascontiguousarray(authentic, np.uint8)
    accept = np.zeros(max(n, 1), np.uint8)
    lib().orc_replay_filter(C.c_uint64(n), _p(session), _p(seq48), _p(auth), _p(state["latest"]), _p(state["mask"]),
                            _p(state["started"]), _p(accept))
    return accept[:n]


def seq_assign(session, next_seq):
    """next_seq (u64[S]) advanced in place; returns (seq48[n], ok[n])"""
    session = np.ascontiguousarray(session, np.uint32)
    n = len(session)
    seq48 = np.zeros(max(n, 1), np.uint64); ok = np.zeros(max(n, 1), np.uint8)
    lib().orc_dtls_seq_assign(C.c_uint64(n), _p(session), C.c_uint32(len(next_seq)), _p(next_seq), _p(seq48), _p(ok))
    return seq48[:n], ok[:n]


def sse_scan(body, body_off):
    body = _u8(body); body_off = np.ascontiguousarray(body_off, np.uint64)
    k = len(body_off) - 1
    lib().orc_sse_scan.restype = C.c_uint64
    cap = int(body.size // 2) + 1
    ev_csr = np.zeros(k + 1, np.uint64); ev_end = np.zeros(cap, np.uint64); done = np.zeros(max(k, 1), np.uint8)
    total = lib().orc_sse_scan(C.c_uint64(k), _p(body), _p(body_off), _p(ev_csr), _p(ev_end), C.c_uint64(cap), _p(done))
    return ev_csr, ev_end[:total].copy(), done[:k].copy()
