"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol include/p2pt.h declares.
No compute call is made (there is no GPU here and the library has no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "p2pt.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(p2pt_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_expected_surface():
    syms = declared_symbols()
    for must in ("p2pt_ctx_create", "p2pt_frame_encode_batch", "p2pt_frame_decode_batch", "p2pt_frame_demux_batch",
                 "p2pt_dtls_gcm_seal_batch", "p2pt_dtls_gcm_open_batch", "p2pt_pipeline_seal", "p2pt_pipeline_open",
                 "p2pt_sessions_load", "p2pt_last_error"):
        assert must in syms


def test_library_exports_every_declared_symbol(built):
    import p2p_llm_tunnel_b200 as p2pt
    lib = C.CDLL(p2pt.lib_path())
    for s in declared_symbols():
        assert hasattr(lib, s), f"{s} declared in include/p2pt.h but not exported"
    lib.p2pt_abi_version.restype = C.c_int32
    assert lib.p2pt_abi_version() == 2


def test_python_binding_covers_every_symbol(built):
    from p2p_llm_tunnel_b200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_symbols()
    _lib.load_library()


def test_no_cpu_fallback_without_a_device(built):
    """ctx_create must fail loudly (never fall back) when no CUDA device is present."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the failure path is exercised on the CPU box")
    import p2p_llm_tunnel_b200 as p2pt
    with pytest.raises(p2pt.P2ptError) as e:
        p2pt.Context(0)
    assert "no CUDA device" in str(e.value) or "CUDA" in str(e.value)


def test_product_does_not_reference_the_oracle():
    pkg = os.path.join(ROOT, "p2p-llm-tunnel_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in text and "p2pt_oracle" not in text and "oracle_binding" not in text, f


def test_header_is_plain_c_and_links_from_c(built, tmp_path):
    """include/p2pt.h through a C99 compiler with -pedantic -Werror (what cgo / bindgen / a Rust `extern "C"` block see), and a
    C program linked against the library calling the entry points that need no device."""
    import subprocess
    import p2p_llm_tunnel_b200 as p2pt
    src = tmp_path / "c_abi.c"
    src.write_text('#include "p2pt.h"\n#include <stdio.h>\n'
                   'int main(void) {\n'
                   '    p2pt_desc d; p2pt_sctp_chunk c; (void)d; (void)c;\n'
                   '    if (sizeof(p2pt_desc) != 24 || sizeof(p2pt_sctp_chunk) != 16) return 2;\n'
                   '    printf("%d %llu %llu %llu\\n", (int)p2pt_abi_version(), (unsigned long long)p2pt_sctp_fragments(65413),\n'
                   '           (unsigned long long)p2pt_sctp_packets_bytes(1200), (unsigned long long)p2pt_frame_demux_scratch_bytes(1000));\n'
                   '    return 0;\n}\n')
    libdir = os.path.dirname(p2pt.lib_path())
    exe = tmp_path / "c_abi"
    subprocess.check_call(["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"), str(src),
                           "-o", str(exe), "-L", libdir, "-lp2ptunnel_b200", "-Wl,-rpath," + libdir])
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()
    assert int(out[0]) >= 2 and out[1] == "55" and out[2] == "1228" and int(out[3]) > 0
