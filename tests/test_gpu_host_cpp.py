"""The C++ host mirror (p2p-llm-tunnel_b200/host/p2pt_host.hpp): the reference's codec unit tests
(/root/reference/tunnel/src/protocol.rs:265-460) restated in C++, and the two batched reader loops against their
per-frame forms (proxy.rs:109-170, serve.rs:82-156), run against it on the GPU; on a box
without a GPU the same binary must fail loudly instead of computing anything on the CPU."""
import subprocess

import pytest


@pytest.mark.gpu
def test_reference_codec_tests_in_cpp(built):
    r = subprocess.run([built.HOST_TEST], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    print(r.stdout)
    assert r.returncode == 0, r.stdout
    assert "26 tests, 0 failed checks" in r.stdout
    for name in ("test_roundtrip_res_body", "test_decode_unknown_type", "test_large_payload", "test_crypto_gcm_known_record",
                 "test_proxy_reader_matches_the_per_frame_loop", "test_serve_reassembler_matches_the_per_frame_loop",
                 "test_replay_detector_matches_the_sliding_window", "test_sequence_numbers_count_per_session",
                 "test_unpack_datagrams_and_drop_the_truncated_one", "test_sse_scan_finds_the_events_of_every_stream"):
        assert f"ok   {name}" in r.stdout


def test_cpp_host_mirror_has_no_cpu_fallback(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = subprocess.run([built.HOST_TEST], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert r.returncode == 2 and "no CUDA device" in r.stdout
