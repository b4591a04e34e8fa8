"""Host-side mirror of webrtc-dtls 0.10.0 `CryptoGcm` (crypto/crypto_gcm.rs; un-vendored, pinned at
/root/reference/tunnel/Cargo.lock:2822-2824) over the C ABI.  Same constructor arguments and
encrypt / decrypt meaning; the explicit nonce, random in the reference, can be injected."""
import os

import numpy as np

from . import _lib

RECORD_LAYER_HEADER_SIZE = 13
CRYPTO_GCM_TAG_LENGTH = 16
CRYPTO_GCM_NONCE_LENGTH = 12


class DtlsError(Exception):
    pass


class ErrNotEnoughRoomForNonce(DtlsError):
    pass


class ErrDecrypt(DtlsError):          # aes-gcm decrypt_in_place Err
    pass


class ErrBadHeader(DtlsError):        # RecordLayerHeader::unmarshal failure
    pass


def record_header(content_type, epoch, seq, length, version=(0xFE, 0xFD)):
    """RecordLayerHeader::marshal (RFC 6347 4.1): type ver(2) epoch(2) seq(6) len(2)."""
    return bytes([content_type, version[0], version[1]]) + int(epoch).to_bytes(2, "big") + \
        int(seq).to_bytes(6, "big") + int(length).to_bytes(2, "big")


class CryptoGcm:
    """CryptoGcm::new(local_key, local_write_iv, remote_key, remote_write_iv)."""

    def __init__(self, local_key, local_write_iv, remote_key, remote_write_iv, ctx=None):
        self.ctx = ctx or _lib.Context(0)
        self._keys = (np.frombuffer(bytes(local_key), np.uint8), np.frombuffer(bytes(local_write_iv[:4]), np.uint8),
                      np.frombuffer(bytes(remote_key), np.uint8), np.frombuffer(bytes(remote_write_iv[:4]), np.uint8))
        self._ensure()

    def _ensure(self):
        # a Context holds one session table; instances sharing a Context reload theirs on demand
        if getattr(self.ctx, "_session_owner", None) is not self:
            self.ctx.sessions_load(*self._keys)
            self.ctx._session_owner = self

    def encrypt(self, pkt_rlh, raw, nonce=None):
        """raw = marshalled header (13) + payload; pkt_rlh is accepted for signature parity (the
        header fields are read from raw, which the reference marshals from pkt_rlh)."""
        self._ensure()
        nonce = os.urandom(8) if nonce is None else bytes(nonce)
        rec, _, status = self.ctx.dtls_seal(np.zeros(1, np.uint32), np.frombuffer(nonce, np.uint8),
                                            np.frombuffer(bytes(raw), np.uint8), np.array([0, len(raw)], np.uint64))
        if status[0] != _lib.ST_OK:
            raise ErrBadHeader(f"seal status {status[0]}")
        return rec.tobytes()

    def decrypt(self, r):
        self._ensure()
        r = bytes(r)
        out, _, out_len, status = self.ctx.dtls_open(np.zeros(1, np.uint32), np.frombuffer(r, np.uint8),
                                                     np.array([0, len(r)], np.uint64))
        st = int(status[0])
        if st in (_lib.ST_OK, _lib.ST_PASS_THROUGH):
            return out[:int(out_len[0])].tobytes()
        if st == _lib.ST_NO_ROOM_FOR_NONCE:
            raise ErrNotEnoughRoomForNonce()
        if st == _lib.ST_BAD_HEADER:
            raise ErrBadHeader()
        raise ErrDecrypt()
