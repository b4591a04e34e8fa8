"""ctypes binding of libp2ptunnel_b200.so (include/p2pt.h).

This is the stub a maintainer of the reference would write in Rust (`extern "C"` block, see
INTEGRATION.md); here it is in Python because no Rust toolchain exists in this environment.
The library is loaded lazily and loudly: a missing .so raises, nothing falls back to the CPU.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_NAME = "libp2ptunnel_b200.so"

# whole-call return codes / per-item status bytes (include/p2pt.h)
OK, ERR_ARG, ERR_CUDA, ERR_CAPACITY, ERR_STATE = 0, -1, -2, -3, -4
ST_OK, ST_TOO_SHORT, ST_UNKNOWN_TYPE, ST_AUTH_FAIL, ST_NO_ROOM_FOR_NONCE, ST_PASS_THROUGH, ST_BAD_HEADER = range(7)
ST_SCTP_BAD_CHECKSUM, ST_SCTP_MALFORMED, ST_SCTP_INCOMPLETE = 7, 8, 9
SCTP_CHUNK_DTYPE = np.dtype([("tsn", "<u4"), ("stream_id", "<u2"), ("ssn", "<u2"), ("ppid", "<u4"),
                             ("flags", "u1"), ("status", "u1"), ("pad", "u1", (2,))])
assert SCTP_CHUNK_DTYPE.itemsize == 16

DESC_DTYPE = np.dtype([("pay_off", "<u8"), ("pay_len", "<u4"), ("stream_id", "<u4"),
                       ("msg_type", "u1"), ("status", "u1"), ("pad", "u1", (6,))])
assert DESC_DTYPE.itemsize == 24

_vp, _u64, _u32, _i32, _int = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int32, C.c_int

# name -> (restype, argtypes); every symbol include/p2pt.h declares
SIGNATURES = {
    "p2pt_abi_version": (_i32, []),
    "p2pt_ctx_create": (_i32, [_int, C.POINTER(_vp)]),
    "p2pt_ctx_destroy": (_i32, [_vp]),
    "p2pt_last_error": (C.c_char_p, [_vp]),
    "p2pt_host_alloc": (_vp, [C.c_size_t]),
    "p2pt_host_free": (None, [_vp]),
    "p2pt_launch_count": (_u64, [_vp]),
    "p2pt_sessions_load": (_i32, [_vp, _u32, _vp, _vp, _vp, _vp]),
    "p2pt_frame_encode_batch": (_i32, [_vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp]),
    "p2pt_frame_decode_batch": (_i32, [_vp, _u64, _vp, _vp, _vp]),
    "p2pt_frame_demux_batch": (_i32, [_vp, _u64, _vp, _vp, _vp, _u64, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _u64, _vp]),
    "p2pt_frame_decode_demux_batch": (_i32, [_vp, _u64, _vp, _vp, _vp, _vp, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _u64, _vp]),
    "p2pt_frame_decode_demux_scratch_bytes": (_u64, [_u64]),
    "p2pt_frame_decode_demux_batch_device": (_i32, [_vp, _u64, _vp, _vp, _vp, _vp, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _u64, _vp, _vp, _u64, _vp]),
    "p2pt_dtls_gcm_seal_batch": (_i32, [_vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "p2pt_dtls_gcm_open_batch": (_i32, [_vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "p2pt_pipeline_seal": (_i32, [_vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "p2pt_pipeline_open": (_i32, [_vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp]),
    "p2pt_pipeline_open_demux": (_i32, [_vp, _u64, _vp, _vp, _vp, _vp, _vp, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _u64, _vp]),
    "p2pt_sctp_fragments": (_u64, [_u64]),
    "p2pt_sctp_packets_bytes": (_u64, [_u64]),
    "p2pt_sctp_packetize_batch": (_i32, [_vp, _u64] + [_vp] * 11 + [_vp, _u64, _vp, _u64, _vp]),
    "p2pt_sctp_unpack_reassemble_batch": (_i32, [_vp, _u64, _vp, _vp, _u32, _vp, _vp, _vp, _u32, _vp, _u64, _vp, _vp, _vp, _vp]),
    "p2pt_sctp_packetize_batch_device": (_i32, [_vp, _u64] + [_vp] * 11 + [_vp, _vp, _vp, _vp, _vp]),
    "p2pt_sctp_unpack_batch_device": (_i32, [_vp, _u64, _vp, _vp, _u32, _vp, _vp, _vp, _vp]),
    "p2pt_sctp_reassemble_scratch_bytes": (_u64, [_u64]),
    "p2pt_sctp_reassemble_batch_device": (_i32, [_vp, _u64, _vp, _vp, _vp, _vp, _u32, _vp, _u64, _vp, _vp, _vp, _vp, _vp, _u64, _vp]),
    "p2pt_dtls_datagram_split_batch": (_i32, [_vp, _u64, _vp, _vp, _u64, _vp, _vp, _vp, _vp]),
    "p2pt_dtls_replay_filter_batch": (_i32, [_vp, _u64, _vp, _vp, _vp, _u32, _vp, _vp, _vp, _vp]),
    "p2pt_dtls_datagram_split_scratch_bytes": (_u64, [_u64]),
    "p2pt_dtls_datagram_split_batch_device": (_i32, [_vp, _u64, _vp, _vp, _u64, _vp, _vp, _vp, _vp, _vp, _u64, _vp]),
    "p2pt_dtls_replay_filter_scratch_bytes": (_u64, [_u64]),
    "p2pt_dtls_replay_filter_batch_device": (_i32, [_vp, _u64, _vp, _vp, _vp, _u32, _vp, _vp, _vp, _vp, _vp, _u64, _vp]),
    "p2pt_dtls_seq_assign_batch": (_i32, [_vp, _u64, _vp, _u32, _vp, _vp, _vp]),
    "p2pt_dtls_seq_assign_scratch_bytes": (_u64, [_u64]),
    "p2pt_dtls_seq_assign_batch_device": (_i32, [_vp, _u64, _vp, _u32, _vp, _vp, _vp, _vp, _u64, _vp]),
    "p2pt_sse_scan_batch": (_i32, [_vp, _u64, _vp, _vp, _u64, _vp, _vp, _vp, _vp]),
    "p2pt_sse_scan_scratch_bytes": (_u64, [_u64]),
    "p2pt_sse_scan_onepass_scratch_bytes": (_u64, [_u64, _u64]),
    "p2pt_sse_scan_onepass_batch_device": (_i32, [_vp, _u64, _vp, _vp, _u64, _u64, _vp, _vp, _vp, _vp, _vp, _u64, _vp]),
    "p2pt_sse_scan_batch_device": (_i32, [_vp, _u64, _vp, _vp, _u64, _vp, _vp, _vp, _vp, _vp, _u64, _vp]),
    "p2pt_frame_encode_batch_device": (_i32, [_vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "p2pt_frame_decode_batch_device": (_i32, [_vp, _u64, _vp, _vp, _vp, _vp]),
    "p2pt_frame_demux_scratch_bytes": (_u64, [_u64]),
    "p2pt_frame_demux_batch_device": (_i32, [_vp, _u64, _vp, _vp, _vp, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _u64, _vp, _vp, _u64, _vp]),
    "p2pt_frame_demux_batch_device_noalloc": (_i32, [_vp, _u64, _vp, _vp, _vp, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _u64, _vp, _vp]),
    "p2pt_dtls_gcm_seal_batch_device": (_i32, [_vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "p2pt_dtls_gcm_open_batch_device": (_i32, [_vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "p2pt_pipeline_seal_device": (_i32, [_vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "p2pt_pipeline_open_device": (_i32, [_vp, _u64, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "p2pt_ctx_set_option": (_i32, [_vp, C.c_char_p, C.c_int64]),
    "p2pt_ctx_get_info": (C.c_int64, [_vp, C.c_char_p]),
}


class P2ptError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"p2pt error {code}: {msg}")
        self.code = code


def lib_path():
    # P2PT_LIB lets the tuning scripts load an alternative build of the same library
    return os.environ.get("P2PT_LIB") or os.path.join(_HERE, LIB_NAME)


_lib = None


def load_library():
    """dlopen the CUDA library and bind every symbol; raises if it was not built."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise P2ptError(ERR_STATE, f"{path} not found: run `python -c 'import __graft_entry__ as g; g.build()'` "
                                   "(there is no CPU fallback)")
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the .so does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def _ptr(x):
    """numpy array / torch tensor / int / None -> address."""
    if x is None:
        return None
    if isinstance(x, (int, np.integer)):
        return int(x)
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    if hasattr(x, "data_ptr"):
        return x.data_ptr()
    raise TypeError(f"cannot take the address of {type(x)}")


def _as(a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype)
    return a


class Context:
    """One p2pt_ctx (one per GPU).  Host methods take/return numpy arrays; *_device methods take
    raw device pointers (ints or torch tensors) and a CUDA stream handle."""

    def __init__(self, device=0):
        self.lib = load_library()
        h = _vp()
        rc = self.lib.p2pt_ctx_create(int(device), C.byref(h))
        if rc != OK:
            raise P2ptError(rc, (self.lib.p2pt_last_error(None) or b"").decode())
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.lib.p2pt_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != OK:
            raise P2ptError(rc, (self.lib.p2pt_last_error(self.h) or b"").decode())

    def call(self, name, *args):
        """Raw call with pointer conversion; raises P2ptError on a negative return."""
        fn = getattr(self.lib, name)
        conv = [(_ptr(a) if not isinstance(a, (int, np.integer, bytes)) or a is None else a) for a in args]
        self._check(fn(self.h, *conv))

    def launch_count(self):
        return int(self.lib.p2pt_launch_count(self.h))

    def get_info(self, name):
        return int(self.lib.p2pt_ctx_get_info(self.h, name.encode()))

    def set_option(self, name, value):
        self._check(self.lib.p2pt_ctx_set_option(self.h, name.encode(), int(value)))

    # ---- sessions ------------------------------------------------------------------------------
    def sessions_load(self, local_keys, local_ivs, remote_keys=None, remote_ivs=None):
        lk = _as(local_keys, np.uint8).reshape(-1, 16)
        li = _as(local_ivs, np.uint8).reshape(-1, 4)
        rk = lk if remote_keys is None else _as(remote_keys, np.uint8).reshape(-1, 16)
        ri = li if remote_ivs is None else _as(remote_ivs, np.uint8).reshape(-1, 4)
        n = lk.shape[0]
        assert li.shape[0] == n and rk.shape[0] == n and ri.shape[0] == n
        self._check(self.lib.p2pt_sessions_load(self.h, n, lk.ctypes.data, li.ctypes.data, rk.ctypes.data, ri.ctypes.data))
        self.n_sessions = n
        self._session_owner = None

    # ---- frame codec (host buffers) ------------------------------------------------------------------
    def frame_encode(self, types, sids, payload, pay_off, out_off=None):
        types = _as(types, np.uint8); sids = _as(sids, np.uint32)
        payload = _as(payload, np.uint8); pay_off = _as(pay_off, np.uint64)
        n = len(types)
        if out_off is None:
            lens = (pay_off[1:] - pay_off[:-1]) + np.uint64(5)
            out_off = np.zeros(n + 1, np.uint64)
            np.cumsum(lens, out=out_off[1:])
        out_off = _as(out_off, np.uint64)
        out = np.zeros(int(out_off[-1]) if n else 0, np.uint8)
        self._check(self.lib.p2pt_frame_encode_batch(self.h, n, types.ctypes.data, sids.ctypes.data,
                                                     payload.ctypes.data, pay_off.ctypes.data,
                                                     out.ctypes.data, out_off.ctypes.data))
        return out, out_off

    def frame_decode(self, buf, in_off):
        buf = _as(buf, np.uint8); in_off = _as(in_off, np.uint64)
        n = len(in_off) - 1
        desc = np.zeros(n, DESC_DTYPE)
        self._check(self.lib.p2pt_frame_decode_batch(self.h, n, buf.ctypes.data, in_off.ctypes.data, desc.ctypes.data))
        return desc

    def frame_demux(self, desc, buf=None, max_streams=None, gather=True, session=None):
        desc = np.ascontiguousarray(desc, dtype=DESC_DTYPE)
        n = len(desc)
        if session is not None:
            session = _as(session, np.uint32)
        usess = np.zeros((max_streams if max_streams is not None else max(n, 1)) + 1, np.uint32)
        if max_streams is None:
            max_streams = max(n, 1)
        uniq = np.zeros(max_streams + 1, np.uint32)
        csr = np.zeros(max_streams + 1, np.uint64)
        perm = np.zeros(max(n, 1), np.uint32)
        k = _u32(0)
        gathered = goff = None
        gcap = 0
        if gather:
            buf = _as(buf, np.uint8)
            body = (desc["status"] == 0) & ((desc["msg_type"] == 11) | (desc["msg_type"] == 21))
            gcap = int(desc["pay_len"][body].sum())
            gathered = np.zeros(max(gcap, 1), np.uint8)
            goff = np.zeros(max_streams + 1, np.uint64)
        self._check(self.lib.p2pt_frame_demux_batch(
            self.h, n, desc.ctypes.data, _ptr(session), _ptr(buf) if gather else None, (buf.size if gather else 0),
            max_streams, uniq.ctypes.data, usess.ctypes.data if session is not None else None,
            csr.ctypes.data, perm.ctypes.data, C.addressof(k),
            _ptr(gathered), gcap, _ptr(goff)))
        kk = k.value
        m = int(csr[kk])
        res = {"k": kk, "uniq_sid": uniq[:kk].copy(), "csr": csr[:kk + 1].copy(), "perm": perm[:m].copy()}
        if session is not None:
            res["uniq_session"] = usess[:kk].copy()
        if gather:
            res["gathered_off"] = goff[:kk + 1].copy()
            res["gathered"] = gathered[:int(goff[kk])].copy()
        return res

    def frame_decode_demux(self, buf, in_off, max_streams, gathered_cap=None, session=None, want_desc=False, want_perm=True):
        """decode + demux + gather in one call (p2pt_frame_decode_demux_batch)"""
        buf = _as(buf, np.uint8); in_off = _as(in_off, np.uint64)
        n = len(in_off) - 1
        if session is not None:
            session = _as(session, np.uint32)
        uniq = np.zeros(max_streams + 1, np.uint32); usess = np.zeros(max_streams + 1, np.uint32)
        csr = np.zeros(max_streams + 1, np.uint64); goff = np.zeros(max_streams + 1, np.uint64)
        perm = np.zeros(max(n, 1), np.uint32)
        desc = np.zeros(max(n, 1), DESC_DTYPE)
        gather = gathered_cap is not None
        gathered = np.zeros(max(gathered_cap or 0, 1), np.uint8)
        k = _u32(0)
        self._check(self.lib.p2pt_frame_decode_demux_batch(
            self.h, n, buf.ctypes.data, in_off.ctypes.data, _ptr(session), desc.ctypes.data if want_desc else None,
            max_streams, uniq.ctypes.data, usess.ctypes.data if session is not None else None, csr.ctypes.data,
            perm.ctypes.data if want_perm else None, C.addressof(k),
            gathered.ctypes.data if gather else None, gathered_cap or 0, goff.ctypes.data if gather else None))
        kk = k.value
        m = int(csr[kk])
        res = {"k": kk, "uniq_sid": uniq[:kk].copy(), "csr": csr[:kk + 1].copy()}
        if want_perm:
            res["perm"] = perm[:m].copy()
        if want_desc:
            res["desc"] = desc[:n]
        if session is not None:
            res["uniq_session"] = usess[:kk].copy()
        if gather:
            res["gathered_off"] = goff[:kk + 1].copy()
            res["gathered"] = gathered[:int(goff[kk])].copy()
        return res

    # ---- DTLS record layer (host buffers) --------------------------------------------------------------
    def dtls_seal(self, session, nonce8, raw, raw_off, rec_off=None):
        session = _as(session, np.uint32); nonce8 = _as(nonce8, np.uint8)
        raw = _as(raw, np.uint8); raw_off = _as(raw_off, np.uint64)
        n = len(session)
        if rec_off is None:
            lens = (raw_off[1:] - raw_off[:-1]) + np.uint64(24)
            rec_off = np.zeros(n + 1, np.uint64)
            np.cumsum(lens, out=rec_off[1:])
        rec_off = _as(rec_off, np.uint64)
        rec = np.zeros(int(rec_off[-1]) if n else 0, np.uint8)
        status = np.zeros(n, np.uint8)
        self._check(self.lib.p2pt_dtls_gcm_seal_batch(self.h, n, session.ctypes.data, nonce8.ctypes.data,
                                                      raw.ctypes.data, raw_off.ctypes.data,
                                                      rec.ctypes.data, rec_off.ctypes.data, status.ctypes.data))
        return rec, rec_off, status

    def dtls_open(self, session, rec, rec_off, out_off=None):
        session = _as(session, np.uint32); rec = _as(rec, np.uint8); rec_off = _as(rec_off, np.uint64)
        n = len(session)
        if out_off is None:
            out_off = rec_off
        out_off = _as(out_off, np.uint64)
        out = np.zeros(int(out_off[-1]) if n else 0, np.uint8)
        out_len = np.zeros(n, np.uint32)
        status = np.zeros(n, np.uint8)
        self._check(self.lib.p2pt_dtls_gcm_open_batch(self.h, n, session.ctypes.data, rec.ctypes.data,
                                                      rec_off.ctypes.data, out.ctypes.data, out_off.ctypes.data,
                                                      out_len.ctypes.data, status.ctypes.data))
        return out, out_off, out_len, status

    def pipeline_seal(self, session, epoch, seq48, nonce8, types, sids, payload, pay_off, rec_off=None):
        session = _as(session, np.uint32); epoch = _as(epoch, np.uint16); seq48 = _as(seq48, np.uint64)
        nonce8 = _as(nonce8, np.uint8); types = _as(types, np.uint8); sids = _as(sids, np.uint32)
        payload = _as(payload, np.uint8); pay_off = _as(pay_off, np.uint64)
        n = len(session)
        if rec_off is None:
            lens = (pay_off[1:] - pay_off[:-1]) + np.uint64(5 + 13 + 24)
            rec_off = np.zeros(n + 1, np.uint64)
            np.cumsum(lens, out=rec_off[1:])
        rec_off = _as(rec_off, np.uint64)
        rec = np.zeros(int(rec_off[-1]) if n else 0, np.uint8)
        status = np.zeros(n, np.uint8)
        self._check(self.lib.p2pt_pipeline_seal(self.h, n, session.ctypes.data, epoch.ctypes.data, seq48.ctypes.data,
                                                nonce8.ctypes.data, types.ctypes.data, sids.ctypes.data,
                                                payload.ctypes.data, pay_off.ctypes.data,
                                                rec.ctypes.data, rec_off.ctypes.data, status.ctypes.data))
        return rec, rec_off, status

    def pipeline_open(self, session, rec, rec_off, out_off=None):
        session = _as(session, np.uint32); rec = _as(rec, np.uint8); rec_off = _as(rec_off, np.uint64)
        n = len(session)
        if out_off is None:
            out_off = rec_off
        out_off = _as(out_off, np.uint64)
        out = np.zeros(int(out_off[-1]) if n else 0, np.uint8)
        desc = np.zeros(n, DESC_DTYPE)
        self._check(self.lib.p2pt_pipeline_open(self.h, n, session.ctypes.data, rec.ctypes.data, rec_off.ctypes.data,
                                                out.ctypes.data, out_off.ctypes.data, desc.ctypes.data))
        return out, out_off, desc

    def pipeline_open_demux(self, session, rec, rec_off, max_streams, gathered_cap, want_out=False, want_desc=True, want_perm=True):
        """open -> decode -> demux -> gather in one call; returns dict like frame_demux plus desc.  want_desc / want_perm
        = False passes NULL for them (they then come back zero-filled / empty)."""
        session = _as(session, np.uint32); rec = _as(rec, np.uint8); rec_off = _as(rec_off, np.uint64)
        n = len(session)
        desc = np.zeros(n, DESC_DTYPE)
        out = np.zeros(int(rec_off[-1]) if (n and want_out) else 0, np.uint8)
        uniq = np.zeros(max_streams + 1, np.uint32); usess = np.zeros(max_streams + 1, np.uint32)
        csr = np.zeros(max_streams + 1, np.uint64); goff = np.zeros(max_streams + 1, np.uint64)
        perm = np.zeros(max(n, 1), np.uint32); gathered = np.zeros(max(gathered_cap, 1), np.uint8)
        k = _u32(0)
        self._check(self.lib.p2pt_pipeline_open_demux(
            self.h, n, session.ctypes.data, rec.ctypes.data, rec_off.ctypes.data,
            out.ctypes.data if want_out else None, desc.ctypes.data if want_desc else None, max_streams, uniq.ctypes.data,
            usess.ctypes.data, csr.ctypes.data, perm.ctypes.data if want_perm else None, C.addressof(k),
            gathered.ctypes.data, gathered_cap, goff.ctypes.data))
        kk = k.value
        m = int(csr[kk])
        return {"k": kk, "desc": desc, "out": out, "uniq_sid": uniq[:kk].copy(), "uniq_session": usess[:kk].copy(),
                "csr": csr[:kk + 1].copy(), "perm": perm[:m].copy() if want_perm else perm[:0], "gathered_off": goff[:kk + 1].copy(),
                "gathered": gathered[:int(goff[kk])].copy()}

    # ---- SCTP DATA chunks + CRC32c (host buffers) ------------------------------------------------------------
    def sctp_packetize(self, msg, msg_off, src_port, dst_port, vtag, sid, ssn, ppid, tsn0, epoch=None, seq48_first=None):
        """messages -> packets.  Returns (out bytes, pkt_off[n_pkts + 1]); with epoch/seq48_first every
        packet carries a 13-byte DTLS record header in front (ready for dtls_seal)."""
        msg = _as(msg, np.uint8); msg_off = _as(msg_off, np.uint64)
        n = len(msg_off) - 1
        a16 = lambda x: _as(np.broadcast_to(np.asarray(x), (n,)), np.uint16)
        a32 = lambda x: _as(np.broadcast_to(np.asarray(x), (n,)), np.uint32)
        src_port, dst_port, sid, ssn = a16(src_port), a16(dst_port), a16(sid), a16(ssn)
        vtag, ppid, tsn0 = a32(vtag), a32(ppid), a32(tsn0)
        lens = (msg_off[1:] - msg_off[:-1]).astype(np.int64)
        frags = np.maximum(1, (lens + 1199) // 1200)
        hdr = 0 if epoch is None else 13
        last = lens - (frags - 1) * 1200
        cap = int(((frags - 1) * 1228 + 28 + ((last + 3) & ~3) + hdr * frags).sum())
        npk = int(frags.sum())
        out = np.zeros(max(cap, 1), np.uint8)
        pkt_off = np.zeros(npk + 1, np.uint64)
        got = _u64(0)
        ep = None if epoch is None else a16(epoch)
        sq = None if epoch is None else _as(np.broadcast_to(np.asarray(seq48_first), (n,)), np.uint64)
        self._check(self.lib.p2pt_sctp_packetize_batch(
            self.h, n, msg.ctypes.data, msg_off.ctypes.data, src_port.ctypes.data, dst_port.ctypes.data, vtag.ctypes.data,
            sid.ctypes.data, ssn.ctypes.data, ppid.ctypes.data, tsn0.ctypes.data, _ptr(ep), _ptr(sq),
            out.ctypes.data, cap, pkt_off.ctypes.data, npk + 1, C.addressof(got)))
        assert got.value == npk
        return out[:cap], pkt_off

    def sctp_unpack_reassemble(self, buf, in_off, headroom=0, session=None, max_msgs=None, msg_cap=None, tsn_base=None):
        buf = _as(buf, np.uint8); in_off = _as(in_off, np.uint64)
        n = len(in_off) - 1
        if session is not None:
            session = _as(session, np.uint32)
        if tsn_base is not None:
            tsn_base = _as(np.broadcast_to(np.asarray(tsn_base, np.uint32), (n,)), np.uint32)
        max_msgs = max(n, 1) if max_msgs is None else max_msgs
        msg_cap = max(int(buf.size), 1) if msg_cap is None else msg_cap
        info = np.zeros(n, SCTP_CHUNK_DTYPE)
        msg = np.zeros(msg_cap, np.uint8); msg_off = np.zeros(max_msgs + 1, np.uint64)
        first = np.zeros(max_msgs + 1, np.uint32); status = np.zeros(max_msgs + 1, np.uint8)
        nm = _u32(0)
        self._check(self.lib.p2pt_sctp_unpack_reassemble_batch(
            self.h, n, buf.ctypes.data, in_off.ctypes.data, headroom, _ptr(session), _ptr(tsn_base), info.ctypes.data, max_msgs,
            msg.ctypes.data, msg_cap, msg_off.ctypes.data, first.ctypes.data, status.ctypes.data, C.addressof(nm)))
        m = nm.value
        return {"info": info, "n_msgs": m, "msg_off": msg_off[:m + 1].copy(), "msg": msg[:int(msg_off[m])].copy(),
                "msg_first": first[:m].copy(), "msg_status": status[:m].copy()}

    # ---- DTLS record-layer bookkeeping -------------------------------------------------------------------------
    def dtls_datagram_split(self, buf, dgram_off, max_records=None):
        buf = _as(buf, np.uint8); dgram_off = _as(dgram_off, np.uint64)
        n = len(dgram_off) - 1
        max_records = max(int(buf.size // 13) + 1, 1) if max_records is None else max_records
        rec_off = np.zeros(max_records + 1, np.uint64); rec_dgram = np.zeros(max_records + 1, np.uint32)
        status = np.zeros(max(n, 1), np.uint8); nr = _u32(0)
        self._check(self.lib.p2pt_dtls_datagram_split_batch(self.h, n, buf.ctypes.data, dgram_off.ctypes.data, max_records,
                                                            rec_off.ctypes.data, rec_dgram.ctypes.data, status.ctypes.data,
                                                            C.addressof(nr)))
        k = nr.value
        return rec_off[:k + 1].copy(), rec_dgram[:k].copy(), status[:n].copy()

    def dtls_replay_filter(self, session, seq48, state, authentic=None):
        """state = dict(latest=u64[S], mask=u64[S], started=u8[S]) updated in place; returns accept[n]"""
        session = _as(session, np.uint32); seq48 = _as(seq48, np.uint64)
        n = len(session)
        accept = np.zeros(max(n, 1), np.uint8)
        if authentic is not None:
            authentic = _as(authentic, np.uint8)
        S = len(state["latest"])
        self._check(self.lib.p2pt_dtls_replay_filter_batch(self.h, n, session.ctypes.data, seq48.ctypes.data, _ptr(authentic), S,
                                                           state["latest"].ctypes.data, state["mask"].ctypes.data,
                                                           state["started"].ctypes.data, accept.ctypes.data))
        return accept[:n]

    def dtls_seq_assign(self, session, next_seq):
        """next_seq = u64[S], the number each session's next record gets, advanced in place; returns (seq48[n], ok[n])"""
        session = _as(session, np.uint32)
        n = len(session)
        seq48 = np.zeros(max(n, 1), np.uint64); ok = np.zeros(max(n, 1), np.uint8)
        assert next_seq.dtype == np.uint64 and next_seq.flags.c_contiguous
        self._check(self.lib.p2pt_dtls_seq_assign_batch(self.h, n, session.ctypes.data, len(next_seq), next_seq.ctypes.data,
                                                        seq48.ctypes.data, ok.ctypes.data))
        return seq48[:n], ok[:n]

    def sse_scan(self, body, body_off, max_events=None):
        """per-stream bodies -> (ev_csr[k + 1], ev_end[n_events], done[k])"""
        body = _as(body, np.uint8); body_off = _as(body_off, np.uint64)
        k = len(body_off) - 1
        max_events = max(int(body.size // 2) + 1, 1) if max_events is None else max_events
        ev_csr = np.zeros(k + 1, np.uint64); ev_end = np.zeros(max_events + 1, np.uint64)
        done = np.zeros(max(k, 1), np.uint8); total = _u64(0)
        self._check(self.lib.p2pt_sse_scan_batch(self.h, k, body.ctypes.data, body_off.ctypes.data, max_events,
                                                 ev_csr.ctypes.data, ev_end.ctypes.data, done.ctypes.data, C.addressof(total)))
        return ev_csr, ev_end[:total.value].copy(), done[:k].copy()
