#!/bin/bash
# SCTP tests and stage times for alternative builds of the library (build/var/<name>.so), then the in-tree build
for v in "$@"; do
  export P2PT_LIB=$PWD/build/var/$v.so; echo "== $v"
  timeout 300 python -m pytest tests/test_gpu_sctp.py -q -m gpu 2>&1 | tail -1
  ONLY=SCTP python tools/bench_configs.py 2>/dev/null | grep -A3 "unpack_crc\|\"packetize" | grep "\"ms"
done
unset P2PT_LIB; echo "== in-tree"
ONLY=SCTP python tools/bench_configs.py 2>/dev/null | grep -A3 "unpack_crc\|\"packetize" | grep "\"ms"
