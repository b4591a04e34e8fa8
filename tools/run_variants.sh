#!/bin/bash
# Times every alternative build of the library under build/var/ (made with extra -D flags):
# P2PT_LIB points the Python binding at the variant.  Usage: tools/run_variants.sh <pattern> <command...>
pat=$1; shift
for f in build/var/${pat}.so; do
  echo "== $f"
  P2PT_LIB=$f "$@" 2>&1 | tail -${TAILN:-3}
done
