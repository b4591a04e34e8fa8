#!/usr/bin/env python
"""bench.py -- framed + DTLS-GCM sealed GB/s and records/s on N B200s (BASELINE.json `metric`).

Workload (BASELINE.json configs[4], "C5"): 65 536 tunnel sessions x 16 RES_BODY frames x 1200 B per
GPU (weak scaling: every rank holds its own 65 536 sessions; global session id = local * N + rank,
no data-path collective -- sessions share no state).  One step = the full pipeline over that batch:
    provider side  frame encode -> DTLS-GCM seal            (serve.rs:269-276 + dc.send)
    consumer side  DTLS-GCM open -> frame decode -> demux by (session, stream_id) -> body gather
`value`  = payload bytes carried through that whole round trip per second, inputs resident in HBM.
`e2e`    = the same step through the host-pointer C ABI with pinned host buffers (H2D/D2H inside).
`--impl reference` times the reference's CPU path (AES-NI/PCLMULQDQ port, all host threads).
"""
import argparse
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SESSIONS, RECS, PAYLOAD = 65536, 16, 1200
METRIC = "framed+DTLS-GCM sealed GB/s (payload bytes through encode->seal->open->decode->demux)"


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


RECORD_KERNEL_SOURCES = ("gcm_kernels.cuh", "gcm_math.cuh", "common.cuh")


def record_kernel_source_hash():
    """hash of the sources of gcm_record_kernel: profiles/record_kernel_pipes.json and roofline_traffic.json are ncu
    measurements of ONE build, tools/make_roofline_profile.py stamps them with this and bench.py ignores a stale file"""
    h = hashlib.sha256()
    for f in RECORD_KERNEL_SOURCES:
        with open(os.path.join(ROOT, "p2p-llm-tunnel_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def bench_config():
    """the `config` object of both arms (the driver compares them: same workload, same batch)"""
    return {"workload": "C5 full frame+seal/open pipeline: 65536 sessions x 16 RES_BODY frames x 1200 B per GPU, "
                        "sessions sharded by connection (global id = local*N + rank), no collective",
            "records_per_step_per_gpu": SESSIONS * RECS, "payload_bytes": PAYLOAD, "streams_per_session": 4,
            "l2_policy": "inputs (1.26 GB payload + 1.31 GB records per step) are larger than the 126 MB L2"}


def make_workload(rank, world, sessions=SESSIONS):
    from p2p_llm_tunnel_b200 import workload
    w = workload.pipeline_workload(sessions, RECS, PAYLOAD, seed_shift=rank, session_base=rank, session_stride=world)
    keys, ivs = workload.sessions_keys(sessions, seed_shift=rank)
    return w, keys, ivs


# --------------------------------------------------------------------------------------- CPU reference arm
def load_cpufast():
    import __graft_entry__ as g
    g.build_oracle()
    lib = C.CDLL(g.CPUFAST_LIB)
    if not lib.cpufast_supported():
        raise RuntimeError("host CPU lacks AES-NI / PCLMULQDQ")
    return lib


def cpu_pipeline(lib, w, keys, ivs, rec, nthreads, reps=1, backend=0, phases=3):
    """returns best (seal_s, open_s) over reps, and stats; phases=5 times the open WITHOUT the per-stream map (AEAD +
    frame decode only).  backend 0 = the RustCrypto-style port
    (oracle/p2pt_cpu_fast.c), 1 = the same record layout around OpenSSL EVP_aes_128_gcm (oracle/p2pt_cpu_openssl.c)"""
    p = lambda a: C.c_void_p(a.ctypes.data)
    times = (C.c_double * 3)(); stats = (C.c_uint64 * 4)()
    best = None
    for _ in range(reps):
        rc = lib.cpufast_pipeline_ex(C.c_uint64(w["n"]), p(w["session"]), p(keys), p(ivs), C.c_uint32(len(keys)),
                                     p(w["epoch"]), p(w["seq48"]), p(w["nonce8"]), p(w["types"]), p(w["sids"]),
                                     p(w["payload"]), p(w["pay_off"]), p(rec), p(w["rec_off"]), None,
                                     C.c_int(nthreads), C.c_int(phases), times, stats, C.c_int(backend))
        if rc != 0:
            raise RuntimeError(f"cpufast_pipeline_ex failed (backend {backend}, rc {rc})")
        t = (times[0], times[1]) if phases == 3 else (times[0], times[2])
        if best is None or sum(t) < sum(best):
            best = t
    return best, [int(x) for x in stats]


def slice_workload(w, sessions):
    """the first `sessions` sessions of a pipeline workload (records of a session are consecutive)"""
    m = sessions * RECS
    sub = {k: (v[:m] if k not in ("nonce8", "payload", "pay_off", "rec_off") else v) for k, v in w.items() if k != "n"}
    sub["nonce8"] = w["nonce8"][: 8 * m]; sub["pay_off"] = w["pay_off"][: m + 1]; sub["rec_off"] = w["rec_off"][: m + 1]
    sub["n"] = m
    return sub


def openssl_baseline(lib, w, keys, ivs, rec, cores, check=None):
    """the record layout around OpenSSL EVP_aes_128_gcm, all cores on the full batch and one thread on
    1/16 of it.  Returns the `openssl_evp` object of cpu_baseline (None if libcrypto was not there at build time)."""
    if not lib.cpuossl_available():
        return None
    payload_bytes = w["n"] * PAYLOAD
    (ts, to), stats = cpu_pipeline(lib, w, keys, ivs, rec, cores, reps=2, backend=1)
    assert stats[0] == 0 and stats[2] == payload_bytes, stats
    if check is not None:
        assert np.array_equal(rec[: 1 << 20], check), "OpenSSL baseline and GPU records differ"
    (_, tn), _ = cpu_pipeline(lib, w, keys, ivs, rec, cores, reps=2, backend=1, phases=5)
    w1 = slice_workload(w, SESSIONS // 16)
    (ts1, to1), stats1 = cpu_pipeline(lib, w1, keys, ivs, rec, 1, reps=1, backend=1)
    (_, tn1), _ = cpu_pipeline(lib, w1, keys, ivs, rec, 1, reps=1, backend=1, phases=5)
    b1 = w1["n"] * PAYLOAD
    return {"value": payload_bytes / (ts + to) / 1e9, "unit": "GB/s", "cores": cores, "kind": "port",
            "seal_gbs": payload_bytes / ts / 1e9, "open_demux_gbs": payload_bytes / to / 1e9,
            "open_only_gbs": payload_bytes / tn / 1e9,
            "one_thread": {"value": b1 / (ts1 + to1) / 1e9, "seal_gbs": b1 / ts1 / 1e9, "open_demux_gbs": b1 / to1 / 1e9,
                           "open_only_gbs": b1 / tn1 / 1e9, "sample": f"{w1['n']} records (1/16 of the batch)"},
            "note": "open_only = AEAD open + frame decode without the per-stream HashMap<_, Vec<u8>> append (serve.rs:120-124), "
                    "which is what limits open_demux on the CPU",
            "sample": f"the full per-GPU batch ({w['n']} records x {PAYLOAD} B), best of 2, the oracle's record layout around "
                      f"OpenSSL EVP_aes_128_gcm (one EVP context per worker, re-keyed per session, IV per record), "
                      f"frame codec and per-stream map as in the port, {cores} host threads sharded by session"}


def run_reference(args):
    rank = env_int("RANK", 0)
    if rank != 0:
        return
    lib = load_cpufast()
    cores = os.cpu_count() or 1
    # the same batch as the GPU arm's step: 65536 sessions x 16 records x 1200 B (1.26 GB payload) per step
    sessions = SESSIONS
    w, keys, ivs = make_workload(0, 1, sessions)
    rec = np.zeros(int(w["rec_off"][-1]), np.uint8)
    for _ in range(max(min(args.warmup, 2), 1)):
        cpu_pipeline(lib, w, keys, ivs, rec, cores)
    t0 = time.perf_counter()
    tot_seal = tot_open = 0.0
    for _ in range(args.steps):
        (ts, to), stats = cpu_pipeline(lib, w, keys, ivs, rec, cores)
        tot_seal += ts; tot_open += to
    wall = time.perf_counter() - t0
    assert stats[0] == 0 and stats[2] == w["n"] * PAYLOAD, stats
    step_s = wall / args.steps
    payload_bytes = w["n"] * PAYLOAD
    gbs = payload_bytes / step_s / 1e9
    try:
        ossl = openssl_baseline(lib, w, keys, ivs, rec, cores)
    except Exception as e:
        ossl = {"value": None, "error": str(e)}
    line = {
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": "GB/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": bench_config(),
        "records_per_s": w["n"] / step_s,
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "port",
                         "sample": f"the full per-GPU batch every step ({sessions} sessions x {RECS} records x {PAYLOAD} B), "
                                   f"AES-NI + PCLMULQDQ C port of the RustCrypto path, {cores} threads sharded by session",
                         "seal_gbs": payload_bytes * args.steps / tot_seal / 1e9,
                         "open_demux_gbs": payload_bytes * args.steps / tot_open / 1e9,
                         "openssl_evp": ossl},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# --------------------------------------------------------------------------------------- GPU arm
def run_gpu(args):
    import torch
    import torch.distributed as dist
    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL prints its version banner on stdout when the communicator is created; keep stdout for
        # the one JSON line by pointing fd 1 at stderr until the first collective has run
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            t_ = torch.zeros(1, device=dev)
            dist.all_reduce(t_)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    import __graft_entry__ as g
    if not os.path.exists(g.LIB):
        g.build_cuda()
    import p2p_llm_tunnel_b200 as p2pt
    from p2p_llm_tunnel_b200._lib import DESC_DTYPE

    ctx = p2pt.Context(local)
    w, keys, ivs = make_workload(rank, world)
    n = w["n"]
    ctx.sessions_load(keys, ivs, keys, ivs)
    payload_bytes = n * PAYLOAD
    rec_bytes = int(w["rec_off"][-1])
    max_streams = SESSIONS * 4

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    from p2p_llm_tunnel_b200 import sharding

    def max_over_ranks(x):
        return sharding.max_over_ranks(x, device=dev)

    # ---------------- device-resident arm ----------------
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).to(dev)
    d = {k: t(w[k]) for k in ("session", "epoch", "seq48", "nonce8", "types", "sids", "payload", "pay_off", "rec_off")}
    rec = torch.empty(rec_bytes, dtype=torch.uint8, device=dev)
    out = torch.empty(rec_bytes, dtype=torch.uint8, device=dev)
    status = torch.empty(n, dtype=torch.uint8, device=dev)
    desc = torch.empty(n * 24, dtype=torch.uint8, device=dev)
    sbytes = int(ctx.lib.p2pt_frame_demux_scratch_bytes(n))
    scratch = torch.empty(sbytes, dtype=torch.uint8, device=dev)
    uniq = torch.empty(max_streams + 1, dtype=torch.int32, device=dev)
    usess = torch.empty(max_streams + 1, dtype=torch.int32, device=dev)
    csr = torch.empty(max_streams + 1, dtype=torch.int64, device=dev)
    goff = torch.empty(max_streams + 1, dtype=torch.int64, device=dev)
    perm = torch.empty(n, dtype=torch.int32, device=dev)
    kdev = torch.zeros(1, dtype=torch.int32, device=dev)
    gathered = torch.empty(payload_bytes, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()
    s = stream.cuda_stream

    def seal():
        ctx.call("p2pt_pipeline_seal_device", n, d["session"], d["epoch"], d["seq48"], d["nonce8"], d["types"], d["sids"],
                 d["payload"], d["pay_off"], rec, d["rec_off"], status, s)

    def open_():
        ctx.call("p2pt_pipeline_open_device", n, d["session"], rec, d["rec_off"], out, d["rec_off"], desc, s)

    def demux():
        ctx.call("p2pt_frame_demux_batch_device", n, desc, d["session"], out, max_streams, uniq, usess, csr, perm, kdev,
                 gathered, payload_bytes, goff, scratch, sbytes, s)

    def step():
        seal(); open_(); demux()

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    # correctness gate on what was just computed: round trip must reproduce the payload per stream
    dn = desc.cpu().numpy().view(DESC_DTYPE)
    assert int(status.sum()) == 0 and int(dn["status"].sum()) == 0, "pipeline produced failing records"
    assert int(kdev.item()) == max_streams, int(kdev.item())
    pm = perm[: 4096].cpu().numpy().astype(np.int64)
    got = gathered[: 4096 * PAYLOAD].cpu().numpy().reshape(4096, PAYLOAD)
    assert np.array_equal(got, w["payload"].reshape(n, PAYLOAD)[pm]), "round trip does not reproduce the payload"

    ev = lambda: torch.cuda.Event(enable_timing=True)
    marks = [[ev() for _ in range(4)] for _ in range(args.steps)]
    launches0 = ctx.launch_count()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    e0, e1 = ev(), ev()
    e0.record(stream)
    for k in range(args.steps):
        marks[k][0].record(stream); seal()
        marks[k][1].record(stream); open_()
        marks[k][2].record(stream); demux()
        marks[k][3].record(stream)
    e1.record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = ctx.launch_count() - launches0
    dev_ms = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    seal_ms = float(np.mean([m[0].elapsed_time(m[1]) for m in marks]))
    open_ms = float(np.mean([m[1].elapsed_time(m[2]) for m in marks]))
    demux_ms = float(np.mean([m[2].elapsed_time(m[3]) for m in marks]))
    value = world * payload_bytes / (dev_ms * 1e-3) / 1e9

    # ---------------- the other BASELINE.json configs, device-resident (C2, C3, C4, SCTP, bookkeeping) ----------------
    configs = None
    if world == 1 and not args.no_configs:
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import bench_configs
            sampler2 = ClockSampler(local)
            sampler2.start()
            configs = bench_configs.run_all(ctx, dev, with_cpu=False)
            configs["clocks"] = sampler2.stop()
            configs["launches"] = ctx.launch_count() - launches0 - launches
        except Exception as e:
            configs = {"error": str(e)}
        ctx.sessions_load(keys, ivs, keys, ivs)      # the C3 configs loaded their own sessions

    def run_e2e():
        # ---------------- end-to-end arm: host-pointer C ABI, pinned host buffers ----------------
        # What a deployment does with one GPU: the batch is cut into sub-batches of whole sessions
        # (sessions share nothing, so each sub-batch is an independent call), the provider side seals
        # (serve.rs:269-276 + dc.send) while the consumer side opens + demuxes (rtc.rs:88 +
        # proxy.rs:109-170).  Two seal workers and two open workers, each a host thread with its own
        # call lane of the one p2pt_ctx, keep PCIe busy in both directions; every byte goes host -> device -> host inside the
        # timed region, through the host-pointer C ABI only.
        import queue
        NB = env_int("P2PT_BENCH_NB", 8)          # sub-batches per step (8192 sessions = 131072 records each)
        sub_n = n // NB
        sub_streams = max_streams // NB
        NW = env_int("P2PT_BENCH_NW", 3)          # worker threads per direction (all on the one context)
        # ONE context per GPU: its call lanes let all six workers be inside host-pointer calls at the same time
        ctxs = [ctx] * (2 * NW)
        # with several calls in flight the overlap comes from the other workers; one large copy per
        # direction per call is what PCIe moves fastest (measured: 48 GB/s vs 41 GB/s in 16 MiB pieces)
        ctx.set_option("chunk_bytes", env_int("P2PT_BENCH_CHUNK_MB", 256) << 20)

        def pinned(a):
            return torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).pin_memory()
        h = {k: pinned(w[k]) for k in ("session", "epoch", "seq48", "nonce8", "types", "sids", "payload", "pay_off", "rec_off")}
        esz = {"session": 4, "epoch": 2, "seq48": 8, "nonce8": 8, "types": 1, "sids": 4, "pay_off": 8, "rec_off": 8}
        hp = {k: h[k].data_ptr() for k in h}
        h_rec = [torch.empty(rec_bytes, dtype=torch.uint8).pin_memory() for _ in range(2)]
        h_status = torch.empty(n, dtype=torch.uint8).pin_memory()
        h_desc = torch.empty(n * 24, dtype=torch.uint8).pin_memory()
        h_uniq = torch.empty(NB * (sub_streams + 1) * 4, dtype=torch.uint8).pin_memory()
        h_usess = torch.empty(NB * (sub_streams + 1) * 4, dtype=torch.uint8).pin_memory()
        h_csr = torch.empty(NB * (sub_streams + 1) * 8, dtype=torch.uint8).pin_memory()
        h_goff = torch.empty(NB * (sub_streams + 1) * 8, dtype=torch.uint8).pin_memory()
        h_perm = torch.empty(n * 4, dtype=torch.uint8).pin_memory()
        h_gath = torch.empty(payload_bytes, dtype=torch.uint8).pin_memory()
        kks = (C.c_uint32 * NB)()
        errs = []

        def at(name, r0):
            return hp[name] + r0 * esz[name]

        def do_seal(c_, buf, b):
            r0 = b * sub_n
            c_.call("p2pt_pipeline_seal", sub_n, at("session", r0), at("epoch", r0), at("seq48", r0), at("nonce8", r0),
                    at("types", r0), at("sids", r0), hp["payload"], at("pay_off", r0), h_rec[buf].data_ptr(),
                    at("rec_off", r0), h_status.data_ptr() + r0)

        def do_open(c_, buf, b):
            r0 = b * sub_n
            so = b * (sub_streams + 1)
            c_.call("p2pt_pipeline_open_demux", sub_n, at("session", r0), h_rec[buf].data_ptr(), at("rec_off", r0), None,
                    h_desc.data_ptr() + 24 * r0, sub_streams, h_uniq.data_ptr() + 4 * so, h_usess.data_ptr() + 4 * so,
                    h_csr.data_ptr() + 8 * so, h_perm.data_ptr() + 4 * r0, C.addressof(kks) + 4 * b,
                    h_gath.data_ptr() + r0 * PAYLOAD, sub_n * PAYLOAD, h_goff.data_ptr() + 8 * so)

        def e2e_run(steps):
            """steps x NB sub-batches, each sealed then opened+demuxed; 2 seal + 2 open worker threads."""
            jobs = steps * NB
            sealed = [threading.Event() for _ in range(jobs)]
            opened = [threading.Event() for _ in range(jobs)]

            def seal_worker(i):
                try:
                    for q in range(i, jobs, NW):
                        if q >= 2 * NB:
                            opened[q - 2 * NB].wait()          # the record buffer of two steps ago is free again
                        do_seal(ctxs[i], (q // NB) & 1, q % NB)
                        sealed[q].set()
                except Exception as e:
                    errs.append(e)
                    for ev_ in sealed + opened:
                        ev_.set()

            def open_worker(i):
                try:
                    for q in range(i, jobs, NW):
                        sealed[q].wait()
                        do_open(ctxs[NW + i], (q // NB) & 1, q % NB)
                        opened[q].set()
                except Exception as e:
                    errs.append(e)
                    for ev_ in sealed + opened:
                        ev_.set()
            ts = [threading.Thread(target=seal_worker, args=(i,)) for i in range(NW)] + \
                 [threading.Thread(target=open_worker, args=(i,)) for i in range(NW)]
            for t_ in ts:
                t_.start()
            for t_ in ts:
                t_.join()
            if errs:
                raise errs[0]

        e2e_steps = max(1, min(args.steps, 6))
        e2e_run(2)
        assert all(int(kks[b]) == sub_streams for b in range(NB)), list(kks)
        assert int(h_status.sum()) == 0
        assert np.array_equal(h_gath[: 4096 * PAYLOAD].numpy().reshape(4096, PAYLOAD), got), "e2e result differs"
        tail = gathered[-4096 * PAYLOAD:].cpu().numpy()
        assert np.array_equal(h_gath[-4096 * PAYLOAD:].numpy(), tail), "e2e result differs (tail)"
        launches_e0 = ctx.launch_count()
        # three timed runs of e2e_steps steps each, the fastest counts: a run is half a second of six host threads
        # on a shared VM, and one descheduled worker costs a whole step (seen: 15.5 GB/s between two runs of 19.4)
        e2e_s = None
        for _ in range(3):
            barrier()
            t0 = time.perf_counter()
            e2e_run(e2e_steps)
            torch.cuda.synchronize()
            run_s = max_over_ranks(time.perf_counter() - t0) / e2e_steps
            e2e_s = run_s if e2e_s is None else min(e2e_s, run_s)
        barrier()
        launches_e2e = (ctx.launch_count() - launches_e0) // (3 * e2e_steps)
        # one side only, as a deployed box runs it (a provider seals, a consumer opens; the round trip above drags
        # every payload byte over PCIe four times): all six workers on the same call, same sub-batches
        def one_side(fn, steps):
            jobs = steps * NB

            def worker(i):
                try:
                    for q in range(i, jobs, 2 * NW):
                        fn(ctx, (q // NB) & 1, q % NB)
                except Exception as e:
                    errs.append(e)
            ts = [threading.Thread(target=worker, args=(i,)) for i in range(2 * NW)]
            barrier()
            t0_ = time.perf_counter()
            for t_ in ts:
                t_.start()
            for t_ in ts:
                t_.join()
            torch.cuda.synchronize()
            dt = max_over_ranks(time.perf_counter() - t0_) / steps
            if errs:
                raise errs[0]
            return dt
        one_side(do_seal, 1)
        prov_s = one_side(do_seal, e2e_steps)
        one_side(do_open, 1)
        cons_s = one_side(do_open, e2e_steps)
        sides = {"provider_only": {"value": world * payload_bytes / prov_s / 1e9, "unit": "GB/s", "ms_per_step": prov_s * 1e3,
                                   "h2d_bytes_per_step": int(payload_bytes + n * (4 + 2 + 8 + 8 + 1 + 4 + 8 + 8)),
                                   "d2h_bytes_per_step": int(rec_bytes + n),
                                   "call": "p2pt_pipeline_seal (serve.rs:269-276 + dc.send): payload up, sealed records down"},
                 "consumer_only": {"value": world * payload_bytes / cons_s / 1e9, "unit": "GB/s", "ms_per_step": cons_s * 1e3,
                                   "h2d_bytes_per_step": int(rec_bytes + n * (4 + 8)),
                                   "d2h_bytes_per_step": int(payload_bytes + n * 28 + NB * (sub_streams + 1) * 24),
                                   "call": "p2pt_pipeline_open_demux (rtc.rs:88 + proxy.rs:109-170): records up, per-stream bodies down"}}
        # latency of ONE sub-batch through both calls with nothing else running, for reference
        t0 = time.perf_counter(); do_seal(ctxs[0], 0, 0); t_seal = time.perf_counter() - t0
        t0 = time.perf_counter(); do_open(ctxs[NW], 0, 0); t_open = time.perf_counter() - t0
        meta = n * (4 + 2 + 8 + 8 + 1 + 4 + 8 + 8)
        h2d = payload_bytes + meta + rec_bytes + n * (4 + 8)
        d2h = rec_bytes + n + payload_bytes + n * 24 + n * 4 + NB * (sub_streams + 1) * (4 + 4 + 8 + 8)
        e2e_val = world * payload_bytes / e2e_s / 1e9

        # single small frame, batch of 1, host buffers: launch + copies dominate; the CPU wins this one
        from p2p_llm_tunnel_b200.protocol import TunnelMessage
        lat = {}
        try:
            sse = b'data: {"token": "hello"}\n\n' + b" " * 37          # 64-byte SSE chunk
            m1 = TunnelMessage.res_body(99, sse)
            for _ in range(5):
                f1 = m1.encode(ctxs[-1]); TunnelMessage.decode(f1, ctxs[-1])
            t0 = time.perf_counter()
            for _ in range(50):
                f1 = m1.encode(ctxs[-1])
            lat["frame_encode_batch1_us"] = (time.perf_counter() - t0) / 50 * 1e6
            t0 = time.perf_counter()
            for _ in range(50):
                TunnelMessage.decode(f1, ctxs[-1])
            lat["frame_decode_batch1_us"] = (time.perf_counter() - t0) / 50 * 1e6
            one = {k: w[k][:1].copy() for k in ("session", "epoch", "seq48", "types", "sids")}
            one_pay = np.frombuffer(sse, np.uint8); one_off = np.array([0, 64], np.uint64)
            for _ in range(3):
                ctxs[-1].pipeline_seal(one["session"], one["epoch"], one["seq48"], w["nonce8"][:8], one["types"], one["sids"], one_pay, one_off)
            t0 = time.perf_counter()
            for _ in range(50):
                r1 = ctxs[-1].pipeline_seal(one["session"], one["epoch"], one["seq48"], w["nonce8"][:8], one["types"], one["sids"], one_pay, one_off)
            lat["pipeline_seal_batch1_us"] = (time.perf_counter() - t0) / 50 * 1e6
            ctxs[-1].pipeline_open_demux(one["session"], r1[0], r1[1], max_streams=1, gathered_cap=64)
            t0 = time.perf_counter()
            for _ in range(50):
                ctxs[-1].pipeline_open_demux(one["session"], r1[0], r1[1], max_streams=1, gathered_cap=64)
            lat["pipeline_open_demux_batch1_us"] = (time.perf_counter() - t0) / 50 * 1e6
            lat["note"] = "wall clock per call through the Python binding, one 64-byte RES_BODY frame, host buffers"
        except Exception as e:
            lat = {"error": str(e)}

        # PCIe as this box delivers it: pinned 1 GiB copies, each direction alone and both at once
        pcie = None
        try:
            nb = 1 << 30
            hp = torch.empty(nb, dtype=torch.uint8).pin_memory(); hq = torch.empty(nb, dtype=torch.uint8).pin_memory()
            dp = torch.empty(nb, dtype=torch.uint8, device=dev); dq = torch.empty(nb, dtype=torch.uint8, device=dev)
            s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

            def timed(fn):
                torch.cuda.synchronize(); t = time.perf_counter(); fn(); torch.cuda.synchronize(); return time.perf_counter() - t

            def up():
                with torch.cuda.stream(s1):
                    dp.copy_(hp, non_blocking=True)

            def down():
                with torch.cuda.stream(s2):
                    hq.copy_(dq, non_blocking=True)
            up(); down(); torch.cuda.synchronize()
            t_up = min(timed(up) for _ in range(3)); t_dn = min(timed(down) for _ in range(3))
            t_bi = min(timed(lambda: (up(), down())) for _ in range(3))
            pcie = {"h2d_gbs": nb / t_up / 1e9, "d2h_gbs": nb / t_dn / 1e9, "bidir_each_gbs": nb / t_bi / 1e9,
                    "theoretical_gbs_per_dir": 63.0}
        except Exception as e:
            pcie = {"error": str(e)}
        if world > 1:
            # what one GPU gets while ALL ranks copy both ways at once: four back-to-back 1 GiB pairs after a
            # barrier, slowest rank counts (the ceiling of the end-to-end arm at N GPUs)
            ok = "error" not in pcie
            barrier()
            t_all = 0.0
            if ok:
                torch.cuda.synchronize(); t_ = time.perf_counter()
                for _ in range(4):
                    up(); down()
                torch.cuda.synchronize(); t_all = time.perf_counter() - t_
            t_all = max_over_ranks(t_all)
            if ok and t_all > 0:
                pcie["bidir_each_gbs_all_ranks_at_once"] = 4 * nb / t_all / 1e9


        return e2e_val, e2e_s, e2e_steps, launches_e2e, t_seal, t_open, h2d, d2h, pcie, lat, sub_n, NB, sides

    if args.no_e2e:
        e2e_val = e2e_s = t_seal = t_open = None
        e2e_steps = launches_e2e = h2d = d2h = sub_n = NB = 0
        pcie = lat = sides = None
    else:
        e2e_val, e2e_s, e2e_steps, launches_e2e, t_seal, t_open, h2d, d2h, pcie, lat, sub_n, NB, sides = run_e2e()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---------------- roofline of the dominant kernel ----------------
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    # algorithmic bytes per record, fused encode->seal: read P + 43, write P + 43
    seal_bytes = 2 * PAYLOAD + 86
    stages = {"seal_ms": seal_ms, "open_ms": open_ms, "demux_gather_ms": demux_ms,
              "seal_gbs_payload": payload_bytes / seal_ms / 1e6, "open_gbs_payload": payload_bytes / open_ms / 1e6,
              "demux_gather_gbs_hbm": n * (2 * PAYLOAD + 41) / demux_ms / 1e6,
              "demux_gather_frac_of_hbm": n * (2 * PAYLOAD + 41) / demux_ms / 1e6 / peak}
    dom = "gcm_record_kernel<4,kPipeSeal>" if seal_ms >= open_ms else "gcm_record_kernel<4,kPipeOpen>"
    dom_ms = max(seal_ms, open_ms)
    dom_bytes = seal_bytes if seal_ms >= open_ms else (2 * PAYLOAD + 91)
    achieved = n * dom_bytes / dom_ms / 1e6
    hbm = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
           "peak_source": peak_src, "algorithmic_bytes_per_record": dom_bytes}
    # AES-GCM without AES / carry-less-multiply instructions is bound by the integer ALU pipe (LOP3 / PRMT / SHF /
    # IADD3), not by HBM.  The instruction counts per 16-byte block come from ncu (tools/make_roofline_profile.py
    # writes them, stamped with the hash of the kernel's sources); a file that describes another build is ignored
    # and the line falls back to the HBM figure.
    roofline = dict(hbm, kernel=dom, traffic=None)
    src_hash = record_kernel_source_hash()
    pipes = os.path.join(ROOT, "profiles", "record_kernel_pipes.json")
    try:
        pj = json.load(open(pipes))
        if pj.get("source_hash") != src_hash:
            roofline["alu_pipe"] = {"stale": True, "note": f"{os.path.basename(pipes)} describes build {pj.get('source_hash')}, "
                                                           f"this build is {src_hash}; re-run tools/make_roofline_profile.py"}
        else:
            which = "seal" if seal_ms >= open_ms else "open"
            k = pj[which]
            sm_hz = (clocks.get("sm_mhz") or 1965.0) * 1e6
            alu_winst = n * k["lane_slots_per_record"] / 32.0 * k["alu_pipe_inst_per_lane_slot"]
            alu_peak = ctx.get_info("sm_count") * pj["alu_warp_inst_per_clk_per_sm"] * sm_hz
            a_ = alu_winst / (dom_ms * 1e-3)
            roofline = {"bound": "alu", "kernel": dom, "achieved": a_ / 1e9, "peak": alu_peak / 1e9, "unit": "G warp-inst/s",
                        "frac": a_ / alu_peak, "traffic": k.get("dram_bytes_per_launch"),
                        "pipe": "integer ALU (LOP3 / PRMT / SHF / IADD3): %.0f of %.0f instructions per 16-byte block and lane"
                                % (k["alu_pipe_inst_per_lane_slot"], k["total_inst_per_lane_slot"]),
                        "peak_source": pj["alu_peak_source"], "sm_mhz_used": sm_hz / 1e6,
                        "ncu": k.get("ncu"), "source": pj.get("source"), "source_hash": src_hash,
                        "hbm": hbm,
                        "note": "bound by the integer ALU pipe, not HBM (DESIGN.md 4); `hbm` is the secondary figure: algorithmic "
                                "bytes per record x records / kernel time against the measured copy peak"}
    except FileNotFoundError:
        pass
    except Exception as e:
        roofline["alu_pipe"] = {"error": str(e)}

    cpu_baseline = None
    if world == 1 and not args.no_cpu:
        try:
            lib = load_cpufast()
            cores = os.cpu_count() or 1
            hw = {k: w[k] for k in w}
            crec = np.zeros(rec_bytes, np.uint8)
            (ts, to), stats = cpu_pipeline(lib, hw, keys, ivs, crec, cores, reps=2)
            assert stats[0] == 0 and stats[2] == payload_bytes
            assert np.array_equal(crec[: 1 << 20], rec[: 1 << 20].cpu().numpy()), "CPU baseline and GPU records differ"
            cpu_baseline = {"value": payload_bytes / (ts + to) / 1e9, "unit": "GB/s", "cores": cores, "kind": "port",
                            "sample": f"the full per-GPU batch ({n} records x {PAYLOAD} B), best of 2, AES-NI + PCLMULQDQ "
                                      f"C port of the RustCrypto path on {cores} host threads sharded by session",
                            "seal_gbs": payload_bytes / ts / 1e9, "open_demux_gbs": payload_bytes / to / 1e9}
            try:
                cpu_baseline["openssl_evp"] = openssl_baseline(lib, hw, keys, ivs, crec, cores, check=rec[: 1 << 20].cpu().numpy())
            except Exception as e:
                cpu_baseline["openssl_evp"] = {"value": None, "error": str(e)}
        except Exception as e:  # the CPU arm must never take the GPU numbers down with it
            cpu_baseline = {"value": None, "unit": "GB/s", "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {e}"}

    line = {
        "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": dev_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": bench_config(),
        "records_per_s": world * n / (dev_ms * 1e-3),
        "stages": stages,
        "configs": configs,
        "roofline": roofline,
        "cpu_baseline": cpu_baseline,
        "e2e": None if args.no_e2e else {"value": e2e_val, "unit": "GB/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_s * 1e3, "steps": e2e_steps, "launches_per_step": int(launches_e2e),
                "one_subbatch_alone_ms": {"pipeline_seal": t_seal * 1e3, "pipeline_open_demux": t_open * 1e3,
                                          "records": sub_n},
                "pcie": pcie,
                "provider_only": sides["provider_only"], "consumer_only": sides["consumer_only"],
                # against what PCIe delivered per direction with both directions busy (on every rank at once for N > 1)
                "frac_of_pcie": (max(h2d, d2h) / e2e_s / 1e9) / pcie.get("bidir_each_gbs_all_ranks_at_once", pcie["bidir_each_gbs"])
                if pcie and "bidir_each_gbs" in pcie else None,
                "note": f"each step = the same {n}-record batch as {NB} sub-batches of whole sessions through "
                        "p2pt_pipeline_seal then p2pt_pipeline_open_demux (host pointers, pinned memory); 3 seal + 3 open "
                        "worker threads on ONE p2pt_ctx (its call lanes); fill and drain of the worker pipeline are inside the timing; "
                        f"each step moves {h2d / 1e9:.2f} GB up and {d2h / 1e9:.2f} GB down over PCIe; "
                        "fastest of three timed runs of `steps` steps (max over ranks each)"},
        "gpu_launches": int(launches),
        "single_frame_latency": lat,
        "clocks": clocks,
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end (host buffer) leg; for profiling runs")
    ap.add_argument("--no-configs", action="store_true", help="skip the device-resident C2 / C3 / C4 / SCTP stage timings")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
