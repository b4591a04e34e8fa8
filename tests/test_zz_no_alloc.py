"""GPU: the allocation-free demux entry point (p2pt_frame_demux_batch_device_noalloc).

Some callers of this library are not allowed to allocate device memory while they run -- a reader
task living inside a fixed device-memory budget because another tenant (an inference server) owns
the rest of the GPU, or a per-connection loop that cannot take cudaMalloc's device-wide
synchronisation.  For them the demux scratch is a static device array inside the library.  This test
calls that entry point through the C ABI several times in a row (the second batch is smaller than the
first, so the array still holds stale data) and requires the same bytes as the caller-scratch entry point
produces AND as the CPU oracle produces, at the sizes where the library changes strategy (up to 1024
frames a single-block kernel; above it the sort-free path when the caller's max_streams is at most 8192,
else the cooperative radix sort; and the stated maximum exactly),
checks that the amount of free device memory is the same before and after the call, and that a
batch beyond the stated maximum is refused.  It is collected last (file name) and needs nothing from
the other tests, so it passes alone in a fresh process too.
"""
import numpy as np
import pytest

import oracle_binding as ob

pytestmark = pytest.mark.gpu

MAX_FRAMES = 262144


def _frames(rng, n, n_streams, n_sessions):
    lens = rng.integers(0, 96, n).astype(np.uint64)
    pay_off = np.zeros(n + 1, np.uint64); np.cumsum(lens, out=pay_off[1:])
    payload = rng.integers(0, 256, int(pay_off[-1]), dtype=np.uint8)
    types = rng.choice(np.array([11, 21, 21, 21, 20, 22, 99, 3], np.uint8), n)
    sids = rng.integers(1, n_streams + 1, n).astype(np.uint32)
    session = rng.integers(0, n_sessions, n).astype(np.uint32)
    return types, sids, payload, pay_off, session


def _run(ctx, torch, entry, n, d_desc, d_sess, d_buf, gcap, scratch=None, ms=None):
    dev = d_desc.device
    ms = ms or n
    o = {"uniq": torch.zeros(ms + 1, dtype=torch.int32, device=dev), "usess": torch.zeros(ms + 1, dtype=torch.int32, device=dev),
         "csr": torch.zeros(ms + 1, dtype=torch.int64, device=dev), "perm": torch.zeros(n, dtype=torch.int32, device=dev),
         "k": torch.zeros(1, dtype=torch.int32, device=dev), "g": torch.zeros(max(gcap, 1), dtype=torch.uint8, device=dev),
         "goff": torch.zeros(ms + 1, dtype=torch.int64, device=dev)}
    s = torch.cuda.current_stream().cuda_stream
    args = [n, d_desc, d_sess, d_buf, ms, o["uniq"], o["usess"], o["csr"], o["perm"], o["k"], o["g"], gcap, o["goff"]]
    if scratch is not None:
        args += [scratch, scratch.numel()]
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    ctx.call(entry, *args, s)
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    if scratch is None:
        assert free_after == free_before, "the allocation-free entry point changed the free device memory"
    k = int(o["k"].item())
    m = int(o["csr"][k].item())
    return {"k": k, "uniq": o["uniq"][:k].cpu().numpy(), "usess": o["usess"][:k].cpu().numpy(),
            "csr": o["csr"][:k + 1].cpu().numpy(), "perm": o["perm"][:m].cpu().numpy(),
            "goff": o["goff"][:k + 1].cpu().numpy(), "g": o["g"][:int(o["goff"][k].item())].cpu().numpy()}


def test_no_alloc_demux_matches_caller_scratch_entry_point(ctx):
    import torch
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(2026)
    # calls in a row on the one shared array: large, smaller (stale data from the first still there), the
    # single-block sizes (n <= 1024), one frame, and exactly the maximum
    for n, streams, sessions, ms in ((40000, 300, 64, None), (7001, 5, 3, None), (1024, 40, 4, None), (700, 9, 2, None),
                                     (1, 1, 1, None), (MAX_FRAMES, 4096, 16, None), (1025, 1000, 1, None),
                                     (MAX_FRAMES, 2000, 4, 8192), (100000, 300, 4, 1200)):
        types, sids, payload, pay_off, session = _frames(rng, n, streams, sessions)
        buf, off = ctx.frame_encode(types, sids, payload, pay_off)
        buf = buf.copy()
        if n > 1:
            buf[off[:-1][rng.random(n) < 0.01].astype(np.int64)] = 250   # a few undecodable frames
        desc = ctx.frame_decode(buf, off)
        d_desc = torch.from_numpy(desc.view(np.uint8).reshape(-1).copy()).to(dev)
        d_sess = torch.from_numpy(session.view(np.int32)).to(dev)
        d_buf = torch.from_numpy(buf).to(dev)
        gcap = int(payload.size)
        sb = int(ctx.lib.p2pt_frame_demux_scratch_bytes(n))
        scratch = torch.zeros(sb, dtype=torch.uint8, device=dev)
        ref = _run(ctx, torch, "p2pt_frame_demux_batch_device", n, d_desc, d_sess, d_buf, gcap, scratch, ms)
        got = _run(ctx, torch, "p2pt_frame_demux_batch_device_noalloc", n, d_desc, d_sess, d_buf, gcap, None, ms)
        assert got["k"] == ref["k"] and got["k"] > 0
        for key in ("uniq", "usess", "csr", "perm", "goff", "g"):
            assert np.array_equal(got[key], ref[key]), (n, key)         # bit-exact, like every other parity test
        exp = ob.frame_demux(desc, buf, session)                        # and against the CPU oracle directly
        assert got["k"] == exp["k"]
        for key, okey in (("uniq", "uniq_sid"), ("usess", "uniq_session"), ("csr", "csr"), ("perm", "perm"),
                          ("goff", "gathered_off"), ("g", "gathered")):
            assert np.array_equal(got[key].astype(np.uint64), np.asarray(exp[okey]).astype(np.uint64)), (n, key)


def test_no_alloc_demux_refuses_batches_beyond_its_maximum(ctx):
    import torch
    import p2p_llm_tunnel_b200 as p2pt
    dev = torch.device("cuda:0")
    one = torch.zeros(64, dtype=torch.uint8, device=dev)
    with pytest.raises(p2pt.P2ptError) as e:
        ctx.call("p2pt_frame_demux_batch_device_noalloc", MAX_FRAMES + 1, one, None, one, 1, one, one, one, one, one,
                 None, 0, None, torch.cuda.current_stream().cuda_stream)
    assert e.value.code == -3 and "P2PT_NOALLOC_MAX_FRAMES" in str(e.value)
