#!/bin/bash
# One GPU-box pass that produces the round's committed evidence under gpurun_out/ (copied to profiles/ afterwards):
# GPU test suite, the bench line and the reference arm, the ncu launch list of the bench step, per-config stage times
# with the CPU port beside them, and ncu captures of the direct demux kernels on C4.
tag=${1:-r02}
o=gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu 2>&1 | tail -3 > $o/${tag}_gputest.txt; cat $o/${tag}_gputest.txt
python bench.py > $o/${tag}_bench_final.json 2> $o/bench.err; tail -c 600 $o/${tag}_bench_final.json; echo
python bench.py --impl reference --steps 3 --warmup 1 > $o/${tag}_reference_final.json 2>> $o/bench.err
python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-configs > /dev/null 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_bench_launches.csv \
      python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-configs > /dev/null 2>&1
python tools/launch_summary.py $o/${tag}_bench_launches.csv > $o/${tag}_bench_launches_summary.txt 2>&1; tail -12 $o/${tag}_bench_launches_summary.txt
python tools/bench_configs.py > $o/${tag}_configs_final.json 2> $o/configs.err; tail -2 $o/configs.err
python tools/prof_codec.py 10000000 64 2 > /dev/null 2>&1 && \
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv \
      --log-file $o/${tag}_c4_launches.csv python tools/prof_codec.py 10000000 64 1 > /dev/null 2>&1
python tools/ncu_launch_table.py $o/${tag}_c4_launches.csv > $o/${tag}_demux_direct_launches.txt; tail -14 $o/${tag}_demux_direct_launches.txt
ncu --set full --import-source on --clock-control none -k regex:demux_direct_rank_kernel --launch-skip 1 -c 1 -o $o/${tag}_rank -f python tools/prof_codec.py 10000000 64 1 > /dev/null 2>&1
ncu --set full --import-source on --clock-control none -k regex:demux_direct_gather_kernel --launch-skip 1 -c 1 -o $o/${tag}_gather -f python tools/prof_codec.py 10000000 64 1 > /dev/null 2>&1
ls -la $o/*.ncu-rep | tail -3
