"""GPU: many threads inside host-pointer calls on ONE context at the same time -- the reference encodes and
sends from many tokio tasks over one connection (/root/reference/tunnel/src/main.rs:18, serve.rs:131) -- and
what a call leaves in the bytes of an output slot that the item does not fill."""
import threading
import time

import numpy as np
import pytest

import oracle_binding as ob

pytestmark = pytest.mark.gpu


def test_eight_threads_on_one_context_are_bit_exact_and_overlap(ctx):
    from p2p_llm_tunnel_b200 import workload
    S, R, P, NT = 2048, 16, 1200, 8
    keys, ivs = workload.sessions_keys(S, seed_shift=21)
    ctx.sessions_load(keys, ivs, keys, ivs)
    w = workload.pipeline_workload(S, R, P, seed_shift=21)
    n = w["n"]
    per = n // NT                                    # whole sessions per thread
    rec_len = P + 5 + 13 + 24

    def piece(i):
        a, b = i * per, (i + 1) * per
        return (w["session"][a:b], w["epoch"][a:b], w["seq48"][a:b], w["nonce8"][8 * a:8 * b], w["types"][a:b], w["sids"][a:b],
                w["payload"][a * P:b * P], workload.fixed_offsets(per, P))

    exp = ob.pipeline_seal({**w, "rec_off": workload.fixed_offsets(n, rec_len)}, keys, ivs)
    out = [None] * NT
    errs = []

    def work(i, reps):
        try:
            for _ in range(reps):
                out[i] = ctx.pipeline_seal(*piece(i))
        except Exception as e:       # noqa: BLE001
            errs.append(e)

    def run(nthreads, reps):
        ts = [threading.Thread(target=work, args=(i, reps)) for i in range(nthreads)]
        t0 = time.perf_counter()
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        return time.perf_counter() - t0

    run(NT, 1)
    assert not errs, errs
    for i in range(NT):
        rec, rec_off, st = out[i]
        assert not st.any()
        assert np.array_equal(rec, exp[i * per * rec_len:(i + 1) * per * rec_len]), i
    # and the consumer side, all threads at once on the same context
    got = [None] * NT

    def open_work(i):
        try:
            rec, rec_off, _ = out[i]
            got[i] = ctx.pipeline_open_demux(w["session"][i * per:(i + 1) * per], rec, rec_off, max_streams=per, gathered_cap=per * P)
        except Exception as e:       # noqa: BLE001
            errs.append(e)
    ts = [threading.Thread(target=open_work, args=(i,)) for i in range(NT)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errs, errs
    for i in range(NT):
        r = got[i]
        assert not r["desc"]["status"].any() and r["k"] == (S // NT) * 4
        pay = w["payload"][i * per * P:(i + 1) * per * P].reshape(per, P)
        assert np.array_equal(r["gathered"].reshape(per, P), pay[r["perm"]])
    # overlap: the same total work from one thread and from eight (copies and kernels of different calls overlap
    # only if calls do not serialise on the context)
    reps = 6

    def serial():
        t0 = time.perf_counter()
        for _ in range(reps):
            for i in range(NT):
                ctx.pipeline_seal(*piece(i))
        return time.perf_counter() - t0
    t1 = min(serial() for _ in range(2))
    t8 = min(run(NT, reps) for _ in range(2))
    print(f"host-pointer fused seal, {n} records per batch: one thread {t1 * 1e3 / reps:.1f} ms, {NT} threads {t8 * 1e3 / reps:.1f} ms, x{t1 / t8:.2f}")
    assert not errs, errs
    assert t1 / t8 > 1.25, (t1, t8)


def test_slot_bytes_beyond_the_item_are_zero(ctx):
    """a record layer must not hand one session's bytes to another: whatever an earlier batch left in the staging
    memory, the bytes of an output slot that the item does not fill come back as zero"""
    rng = np.random.default_rng(5)
    keys = rng.integers(0, 256, (4, 16), dtype=np.uint8); ivs = rng.integers(0, 256, (4, 4), dtype=np.uint8)
    ctx.sessions_load(keys, ivs, keys, ivs)
    n, L = 6000, 300
    # first batch: fills the staging arena with non-zero plaintext
    raw_off = (np.arange(n + 1, dtype=np.uint64) * np.uint64(13 + L))
    raw = np.full(int(raw_off[-1]), 0xA5, np.uint8)
    raw.reshape(n, 13 + L)[:, 0] = 23; raw.reshape(n, 13 + L)[:, 1] = 0xFE; raw.reshape(n, 13 + L)[:, 2] = 0xFD
    session = rng.integers(0, 4, n).astype(np.uint32)
    nonce = rng.integers(0, 256, 8 * n, dtype=np.uint8)
    rec, rec_off, st = ctx.dtls_seal(session, nonce, raw, raw_off)
    assert not st.any()
    # open: the natural layout (out slot = record length) leaves 24 bytes per slot unused
    bad = rec.copy(); bad[int(rec_off[7]) + 40] ^= 1                        # one record that fails is cleared whole
    out, _, olen, st = ctx.dtls_open(session, bad, rec_off)
    o2 = out.reshape(n, 37 + L)
    assert st[7] == 3 and not o2[7].any()
    ok = np.ones(n, bool); ok[7] = False
    assert np.all(olen[ok] == 13 + L) and not o2[ok, 13 + L:].any()
    assert np.array_equal(o2[ok, 13:13 + L], raw.reshape(n, 13 + L)[ok, 13:])
    # seal into slots with slack, and with every kind of item that is refused
    slot = 13 + L + 24 + 19
    rec_off2 = np.arange(n + 1, dtype=np.uint64) * np.uint64(slot)
    rec2, _, st = ctx.dtls_seal(session, nonce, raw, raw_off, rec_off2)
    assert not st.any()
    r2 = rec2.reshape(n, slot)
    assert np.array_equal(r2[:, :13 + L + 24].reshape(-1), rec) and not r2[:, 13 + L + 24:].any()
    # the fused reader without desc / perm (a reader that only wants the per-stream bodies): same bodies
    ra = ctx.pipeline_open_demux(session, rec, rec_off, max_streams=n, gathered_cap=int(rec.size))
    rb = ctx.pipeline_open_demux(session, rec, rec_off, max_streams=n, gathered_cap=int(rec.size), want_desc=False, want_perm=False)
    assert rb["k"] == ra["k"] and np.array_equal(rb["gathered"], ra["gathered"]) and np.array_equal(rb["csr"], ra["csr"])
    assert not rb["desc"]["status"].any() and rb["perm"].size == 0
    # the fused reader with `out` requested
    w_off = np.arange(n + 1, dtype=np.uint64) * np.uint64(37 + L)
    r = ctx.pipeline_open_demux(session, rec, w_off, max_streams=n, gathered_cap=int(rec.size), want_out=True)
    assert not r["out"].reshape(n, 37 + L)[:, 13 + L:].any()


def test_eight_threads_demux_on_one_context_share_the_side_streams(ctx):
    """p2pt_frame_decode_demux_batch from eight threads at once on ONE context: every call takes the sort-free path
    (max_streams <= 8192), whose on-device fallback is enqueued on one of the context's eight side streams - handed out
    round-robin, so concurrent calls share them.  Results must equal the oracle's for every thread and repetition, also
    for the batch that overflows the direct path and comes out of the fallback (capacity error with the exact count)."""
    import p2p_llm_tunnel_b200 as p2pt
    from test_gpu_parity import make_frames
    NT, REPS = 8, 6
    rng = np.random.default_rng(0x7EA0)
    work = []
    for i in range(NT):
        n = int(rng.integers(3000, 40000))
        buf, off = make_frames(rng, n, int(rng.integers(1, 3000)), corrupt=0.02)
        d = ob.frame_decode(buf, off)
        work.append((buf, off, ob.frame_demux(d, buf)))
    # one thread's batch has more streams than the direct path holds
    big_buf, big_off = make_frames(rng, 60000, 20000, corrupt=0)
    big_k = ob.frame_demux(ob.frame_decode(big_buf, big_off), big_buf)["k"]
    assert big_k > 8192
    errs = []

    def run(i):
        try:
            buf, off, exp = work[i]
            for rep in range(REPS):
                if i == 3 and rep % 2:
                    try:
                        ctx.frame_decode_demux(big_buf, big_off, 8192, gathered_cap=len(big_buf))
                        raise AssertionError("expected a capacity error")
                    except p2pt.P2ptError as e:
                        assert e.code == -3
                    continue
                r = ctx.frame_decode_demux(buf, off, 8192, gathered_cap=int(exp["gathered_off"][-1]))
                assert r["k"] == exp["k"]
                for key in ("uniq_sid", "csr", "perm", "gathered_off", "gathered"):
                    assert np.array_equal(r[key], exp[key]), (i, rep, key)
        except Exception as e:       # noqa: BLE001
            errs.append((i, e))

    ts = [threading.Thread(target=run, args=(i,)) for i in range(NT)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errs, errs
