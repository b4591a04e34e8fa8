"""Deterministic synthetic workloads: splitmix64 streams, numpy-vectorised."""
import numpy as np

SEED_PAYLOAD, SEED_SID, SEED_KEYS, SEED_NONCE, SEED_CORRUPT = 0x5EED0001, 0x5EED0002, 0x5EED0003, 0x5EED0004, 0x5EED0005
_M = np.uint64


def splitmix64(seed, n, start=0):
    """n outputs of splitmix64 seeded with `seed`, beginning at output index `start`."""
    with np.errstate(over="ignore"):
        i = np.arange(start + 1, start + n + 1, dtype=np.uint64)
        z = _M(seed) + i * _M(0x9E3779B97F4A7C15)
        z = (z ^ (z >> _M(30))) * _M(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> _M(27))) * _M(0x94D049BB133111EB)
        return z ^ (z >> _M(31))


def rand_bytes(seed, nbytes, start_word=0):
    w = splitmix64(seed, (nbytes + 7) // 8, start_word)
    return w.view(np.uint8)[:nbytes].copy()


# SSE chunk template of /root/reference/tmp/mock_llm.py:45-54 (one streamed token)
SSE_TEMPLATE = (b'data: {"id": "chatcmpl-test", "object": "chat.completion.chunk", "choices": '
                b'[{"index": 0, "delta": {"content": "Hello"}, "finish_reason": null}]}\n\n')


def fixed_offsets(n, size):
    return (np.arange(n + 1, dtype=np.uint64) * np.uint64(size))


def frames_workload(n, payload_len, n_streams=4096, sse=False, seed_shift=0, out_stride=None):
    """C2 / C4: n RES_BODY frames, payload_len bytes each, stream ids uniform over 1..=n_streams,
    packed back to back.  Returns dict with the encoder inputs.  out_stride (C2p: 1216 for 1200-byte
    payloads) pads every output slot to that many bytes."""
    types = np.full(n, 21, np.uint8)
    sids = (splitmix64(SEED_SID + seed_shift, n) % _M(n_streams)).astype(np.uint32) + np.uint32(1)
    if sse:
        t = np.frombuffer((SSE_TEMPLATE * (payload_len // len(SSE_TEMPLATE) + 1))[:payload_len], np.uint8)
        payload = np.tile(t, n)
    else:
        payload = rand_bytes(SEED_PAYLOAD + seed_shift, n * payload_len)
    return {"n": n, "types": types, "sids": sids, "payload": payload,
            "pay_off": fixed_offsets(n, payload_len), "frame_off": fixed_offsets(n, out_stride or payload_len + 5)}


def lifecycle_workload(n_streams, bodies_per_stream, payload_len, corrupt_permille=1, seed_shift=0):
    """C2x: every stream lives the way proxy.rs:118-153 sees it - one ResHeaders frame (JSON, the stream id inside),
    `bodies_per_stream` ResBody frames, one ResEnd - and the streams are interleaved at random with each stream's own
    order kept.  `corrupt_permille` of the frames are then damaged (half truncated below five bytes, half given a tag
    that is no MessageType); which ones is in `corrupt`.  Ragged payloads (headers ~40 B, bodies payload_len, ends 0)."""
    per = bodies_per_stream + 2
    n = n_streams * per
    # a random interleaving that keeps per-stream order: sort frames by a random key, then number the frames of
    # each stream in the order they come out
    key = splitmix64(SEED_SID + 0x200 + seed_shift, n)
    sids = (np.arange(n, dtype=np.uint64) % _M(n_streams)).astype(np.uint32)[np.argsort(key, kind="stable")] + np.uint32(1)
    order = np.argsort(sids, kind="stable")
    pos = np.empty(n, np.int64)
    pos[order] = np.tile(np.arange(per, dtype=np.int64), n_streams)           # index of the frame inside its stream
    types = np.where(pos == 0, 20, np.where(pos == per - 1, 22, 21)).astype(np.uint8)
    hdr = [b'{"stream_id":%d,"status":200,"headers":{}}' % s for s in range(1, n_streams + 1)]
    hlen = np.array([len(h) for h in hdr], np.int64)
    lens = np.where(types == 21, payload_len, 0).astype(np.int64)
    is_h = types == 20
    lens[is_h] = hlen[sids[is_h].astype(np.int64) - 1]
    pay_off = np.zeros(n + 1, np.uint64); np.cumsum(lens, out=pay_off[1:])
    payload = rand_bytes(SEED_PAYLOAD + 0x200 + seed_shift, int(pay_off[-1]))
    for i in np.flatnonzero(is_h):
        o = int(pay_off[i]); payload[o:o + int(lens[i])] = np.frombuffer(hdr[int(sids[i]) - 1], np.uint8)
    frame_off = np.zeros(n + 1, np.uint64); np.cumsum(lens + 5, out=frame_off[1:])
    corrupt = (splitmix64(SEED_CORRUPT + 0x200 + seed_shift, n) % _M(1000)) < _M(corrupt_permille)
    return {"n": n, "types": types, "sids": sids, "payload": payload, "pay_off": pay_off, "frame_off": frame_off,
            "pos": pos, "corrupt": corrupt, "per_stream": per}


def sessions_keys(n_sessions, seed_shift=0):
    keys = rand_bytes(SEED_KEYS + seed_shift, 16 * n_sessions).reshape(n_sessions, 16)
    ivs = rand_bytes(SEED_KEYS + seed_shift + 0x100, 4 * n_sessions).reshape(n_sessions, 4)
    return keys, ivs


def records_workload(n, plaintext_len, n_sessions=1, seed_shift=0):
    """C3: n raw records = hdr13 || plaintext_len random bytes; content type 23, version fe fd,
    epoch 1, seq = index of the record within its session; injected explicit nonces."""
    session = (np.arange(n, dtype=np.uint64) % _M(n_sessions)).astype(np.uint32)
    seq = (np.arange(n, dtype=np.uint64) // _M(n_sessions))
    stride = 13 + plaintext_len
    raw = np.empty((n, stride), np.uint8)
    raw[:, 0] = 23; raw[:, 1] = 0xFE; raw[:, 2] = 0xFD; raw[:, 3] = 0; raw[:, 4] = 1
    for b in range(6):
        raw[:, 5 + b] = ((seq >> _M(8 * (5 - b))) & _M(0xFF)).astype(np.uint8)
    raw[:, 11] = (plaintext_len >> 8) & 0xFF
    raw[:, 12] = plaintext_len & 0xFF
    raw[:, 13:] = rand_bytes(SEED_PAYLOAD + seed_shift, n * plaintext_len).reshape(n, plaintext_len)
    nonce8 = rand_bytes(SEED_NONCE + seed_shift, 8 * n)
    return {"n": n, "session": session, "seq": seq, "nonce8": nonce8, "raw": raw.reshape(-1),
            "raw_off": fixed_offsets(n, stride), "rec_off": fixed_offsets(n, stride + 24)}


def pipeline_workload(n_sessions, recs_per_session, payload_len, streams_per_session=4, seed_shift=0,
                      session_base=0, session_stride=1):
    """C5: n_sessions x recs_per_session RES_BODY frames of payload_len bytes; the records of a
    session are consecutive (seq = 0..recs-1), stream ids 1..=streams_per_session inside a session."""
    n = n_sessions * recs_per_session
    local = np.repeat(np.arange(n_sessions, dtype=np.uint32), recs_per_session)
    seq = np.tile(np.arange(recs_per_session, dtype=np.uint64), n_sessions)
    sids = (seq % _M(streams_per_session)).astype(np.uint32) + np.uint32(1)
    return {"n": n, "session": local, "global_session": local * np.uint32(session_stride) + np.uint32(session_base),
            "epoch": np.ones(n, np.uint16), "seq48": seq,
            "nonce8": rand_bytes(SEED_NONCE + seed_shift, 8 * n), "types": np.full(n, 21, np.uint8), "sids": sids,
            "payload": rand_bytes(SEED_PAYLOAD + seed_shift, n * payload_len),
            "pay_off": fixed_offsets(n, payload_len), "rec_off": fixed_offsets(n, payload_len + 5 + 13 + 24)}
