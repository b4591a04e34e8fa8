"""CPU: host-side logic that needs no GPU -- workload generator determinism, the math header the
kernels use (compiled for the host) against the oracle, message-type mirror, negotiation."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_splitmix64_known_values():
    from p2p_llm_tunnel_b200 import workload
    # reference values of splitmix64 with seed 0 (public test vector of the generator)
    v = workload.splitmix64(0, 3)
    assert [int(x) for x in v] == [0xE220A8397B1DCDAF, 0x6E789E6AA1B965F4, 0x06C45D188009454F]
    a = workload.splitmix64(0x5EED0001, 10)
    b = workload.splitmix64(0x5EED0001, 5, start=5)
    assert np.array_equal(a[5:], b)


def test_workloads_are_deterministic_and_shaped():
    from p2p_llm_tunnel_b200 import workload
    w1 = workload.frames_workload(1000, 1200)
    w2 = workload.frames_workload(1000, 1200)
    assert np.array_equal(w1["payload"], w2["payload"]) and np.array_equal(w1["sids"], w2["sids"])
    assert w1["sids"].min() >= 1 and w1["sids"].max() <= 4096
    assert w1["frame_off"][-1] == 1000 * 1205
    r = workload.records_workload(100, 1200, n_sessions=7)
    assert r["raw"].size == 100 * 1213 and r["raw"][0] == 23 and r["raw"][1] == 0xFE and r["raw"][2] == 0xFD
    assert int.from_bytes(r["raw"][1213 * 15 + 5: 1213 * 15 + 11].tobytes(), "big") == 15 // 7
    s = workload.frames_workload(10, 64, sse=True)
    assert bytes(s["payload"][:6]) == b"data: "
    p = workload.pipeline_workload(8, 16, 1200)
    assert p["n"] == 128 and p["sids"].max() == 4 and p["seq48"][17] == 1


def test_message_type_mirror():
    from p2p_llm_tunnel_b200.protocol import MessageType, Agree, Hello
    assert MessageType.from_u8(21) is MessageType.ResBody and MessageType.from_u8(255) is None
    assert sorted(int(m) for m in MessageType) == [1, 2, 3, 4, 10, 11, 12, 20, 21, 22, 99]
    # negotiation (protocol.rs:464-549)
    a = Agree.from_hello(Hello(min_version=1, max_version=5, features=["sse", "zstd"]))
    assert a.version == 1 and a.features == ["sse"]
    with pytest.raises(ValueError):
        Agree.from_hello(Hello(min_version=2, max_version=3))
    with pytest.raises(ValueError):
        Agree.from_hello(Hello(proto="other"))


def test_device_gf128_math_on_host_matches_oracle(tmp_path):
    """gcm_math.cuh is __host__ __device__: compile it for the host and compare the integer-multiply
    carry-less GF(2^128) product with the oracle's SP 800-38D bitwise multiply."""
    src = tmp_path / "t.cpp"
    src.write_text(r'''
#include "gcm_math.cuh"
#include "p2pt_oracle.h"
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
using namespace p2pt;
static U128 from_bytes(const uint8_t* b){ uint32_t w[4]; memcpy(w,b,16); return block_to_u128(w[0],w[1],w[2],w[3]); }
static void to_bytes(const U128& x, uint8_t* b){ uint32_t w[4]={bswap32(x.w[3]),bswap32(x.w[2]),bswap32(x.w[1]),bswap32(x.w[0])}; memcpy(b,w,16);}
int main(){ srand(7); int bad=0;
 for(int it=0;it<5000;it++){ uint8_t x[16],y[16],z[16],z2[16];
  for(int i=0;i<16;i++){x[i]=rand();y[i]=rand();}
  if(it==0){memset(x,0xff,16);memset(y,0xff,16);} if(it==1){memset(x,0,16);x[0]=0x80;} if(it==2){memset(y,0,16);y[15]=1;}
  if(it==3){memset(x,0,16);} if(it==4){memset(x,0xff,16);memset(y,0,16);y[0]=0x80;}
  orc_gf128_mul(x,y,z); GhashKey k; ghash_key_prepare(from_bytes(y),k); to_bytes(gf128_mul(from_bytes(x),k),z2);
  if(memcmp(z,z2,16)) bad++;
  // the form the record kernel uses: multiplier times alpha^-1 once, no shift in the product
  GhashKey kp; ghash_key_prepare(gf128_premul(from_bytes(y)),kp); to_bytes(gf128_mul_pre(from_bytes(x),kp),z2);
  if(memcmp(z,z2,16)) bad++; }
 printf("bad=%d\n",bad); return bad!=0; }
''')
    exe = tmp_path / "t"
    subprocess.check_call(["g++", "-O2", "-I", os.path.join(ROOT, "p2p-llm-tunnel_b200", "csrc"), "-I",
                           os.path.join(ROOT, "oracle"), "-x", "c++", str(src), os.path.join(ROOT, "oracle", "p2pt_oracle.c"),
                           "-o", str(exe)])
    out = subprocess.check_output([str(exe)]).decode()
    assert "bad=0" in out


def test_cpp_reader_routes_headers_like_a_json_parser(tmp_path):
    """The C++ reader mirror routes REQ_HEADERS / RES_HEADERS frames by the `stream_id` INSIDE their JSON payload
    (proxy.rs:129, serve.rs:113-118 parse the payload with serde_json), with a scanner that interprets nothing else of the
    JSON.  Host-only logic: compiled for the CPU and compared with Python's json module on well-formed and broken
    payloads."""
    import json
    cases = [
        '{"stream_id":42,"status":200,"headers":{}}',
        '{ "stream_id" : 7 , "method":"POST"}',
        '{"headers":{"stream_id":9},"stream_id":3}',                       # the nested one is not it
        '{"note":"\\"stream_id\\":99","stream_id":5}',                     # inside a string, with escaped quotes
        '{"a":[{"stream_id":1}],"stream_id":4294967295}',                  # u32::MAX
        '{"stream_id":4294967296}',                                        # does not fit u32: serde fails
        '{"stream_id":-1}', '{"stream_id":"12"}', '{"stream_id":null}', '{"status":200}',
        '{"stream_id":12', '{"stream_id":12}}', '', 'not json', '{"stream_id":1,"s":"unterminated}',
        '{"k":"\\\\","stream_id":8}',                                      # a string that ends in a backslash pair
        '{"stream_id":\n\t17}',
    ]
    want = []
    for c in cases:
        try:
            v = json.loads(c)
            sid = v.get("stream_id") if isinstance(v, dict) else None
            ok = isinstance(sid, int) and not isinstance(sid, bool) and 0 <= sid <= 0xFFFFFFFF
            want.append(sid if ok else -1)
        except ValueError:
            want.append(-1)
    src = tmp_path / "j.cpp"
    src.write_text(r'''
#include "p2pt_host.hpp"
#include <cstdio>
#include <iostream>
#include <string>
int main() {
    std::string line;
    while (std::getline(std::cin, line, '\x1e')) {
        const uint8_t* p = (const uint8_t*)line.data();
        auto sid = p2pt_host::reader::json_looks_complete(p, line.size()) ? p2pt_host::reader::json_u32_field(p, line.size(), "stream_id")
                                                                          : std::nullopt;
        std::printf("%lld\n", sid ? (long long)*sid : -1LL);
    }
    return 0;
}
''')
    import p2p_llm_tunnel_b200 as p2pt
    libdir = os.path.dirname(p2pt.lib_path())
    exe = tmp_path / "j"
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "p2p-llm-tunnel_b200", "host"),
                           str(src), "-o", str(exe), "-L", libdir, "-lp2ptunnel_b200", "-Wl,-rpath," + libdir])
    out = subprocess.run([str(exe)], input="\x1e".join(cases).encode(), capture_output=True, check=True).stdout.decode().split()
    got = [int(x) for x in out]
    # an empty last case yields no record from getline: compare what came out, then the count
    assert got == want[:len(got)] and len(got) >= len(cases) - 1, list(zip(cases, want, got))


def test_committed_bench_lines_carry_the_contract():
    """profiles/r02_bench_final.json and r02_reference_final.json (the lines `bench.py` printed on the B200 box) have every
    key the measurement contract names, with consistent values."""
    import json
    line = json.load(open(os.path.join(ROOT, "profiles", "r02_bench_final.json")))
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
                "dtype", "data", "config", "roofline", "cpu_baseline", "e2e", "gpu_launches", "clocks", "configs"):
        assert key in line, key
    assert line["unit"] == "GB/s" and line["higher_is_better"] is True and line["scaling"] == "weak" and line["dtype"] == "u8"
    assert line["vs_baseline"] is None and line["data"] == "synthetic" and "workload" in line["config"] and line["warmup"] >= 3
    assert line["gpu_launches"] > 0 and line["n_gpus"] == 1
    rf = line["roofline"]
    for key in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert key in rf, key
    assert abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-9 and 0.5 < rf["frac"] < 1.0
    assert rf["hbm"]["bound"] == "hbm" and abs(rf["hbm"]["frac"] - rf["hbm"]["achieved"] / rf["hbm"]["peak"]) < 1e-9
    cb = line["cpu_baseline"]
    for key in ("value", "unit", "cores", "kind", "sample"):
        assert key in cb, key
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and "openssl_evp" in cb
    e2e = line["e2e"]
    for key in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"):
        assert key in e2e, key
    assert e2e["h2d_bytes_per_step"] > 2e9 and e2e["d2h_bytes_per_step"] > 2e9 and 0 < e2e["value"] < line["value"]
    ck = line["clocks"]
    assert ck["sm_mhz"] > 0.9 * ck["sm_max_mhz"] and not set(ck["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    # value = payload bytes of the batch / step time
    n = line["config"]["records_per_step_per_gpu"] * line["config"]["payload_bytes"]
    assert abs(line["value"] - n / line["ms_per_step"] / 1e6) < 0.01 * line["value"]
    ref = json.load(open(os.path.join(ROOT, "profiles", "r02_reference_final.json")))
    assert ref["impl"] == "reference" and ref["metric"] == line["metric"] and ref["unit"] == line["unit"] and ref["config"] == line["config"]
    assert ref["e2e"]["h2d_bytes_per_step"] == 0 and ref["e2e"]["d2h_bytes_per_step"] == 0 and ref["e2e"]["value"] == ref["value"]
    assert ref["cpu_baseline"]["value"] == ref["value"] and ref["cpu_baseline"]["kind"] in ("port", "reference")
