#!/bin/bash
# ncu launch list (kernel times) of the direct demux kernels for every build/var/<pattern>.so on C4 (and C2 with ARGS="1000000 1200")
pat=$1
for f in build/var/${pat}.so; do
  echo "== $f"
  P2PT_LIB=$f ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file /tmp/l.csv python tools/prof_codec.py ${ARGS:-10000000 64} 1 > /dev/null 2>&1
  python tools/ncu_launch_table.py /tmp/l.csv | grep -E "direct_rank|direct_gather"
done
