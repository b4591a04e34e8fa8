#!/bin/bash
# usage: tools/lib_variants.sh "<pytest selection>" <ONLY filter of bench_configs> <variant>...   (variants = build/var/<name>.so;
# the in-tree build is timed last)
sel="$1"; only="$2"; shift 2
for v in "$@"; do
  export P2PT_LIB=$PWD/build/var/$v.so; echo "== $v"
  timeout 600 python -m pytest $sel -q -m gpu -x 2>&1 | tail -3
  ONLY=$only python tools/bench_configs.py 2>/dev/null | grep -B1 -A1 "\"ms\"" | grep "ms\|{" | paste - - | sed 's/ \+/ /g'
done
unset P2PT_LIB; echo "== in-tree"
ONLY=$only python tools/bench_configs.py 2>/dev/null | grep -B1 -A1 "\"ms\"" | grep "ms\|{" | paste - - | sed 's/ \+/ /g'
