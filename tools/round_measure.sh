#!/bin/bash
# One GPU box: the round's bench lines, the ncu launch list and one full capture of every kernel.  usage: tools/round_measure.sh r02x
# (numbers printed under ncu are never bench values: the bench lines come from the plain runs first; one ncu pass per box call)
R=${1:-rXX}; O=gpurun_out; mkdir -p $O
timeout 600 python bench.py --steps 5 --warmup 3 > $O/${R}_bench_default_1gpu.json 2> $O/${R}_bench.err; echo "bench rc $?"      # configs[1] + [2], [3], [0] + e2e_fd
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/${R}_bench_reference_1gpu.json 2>> $O/${R}_bench.err; echo "ref rc $?"
timeout 200 python tools/level_sweep.py 1024 text > $O/${R}_level_sweep.txt 2>&1
if [ "$2" = "launches" ]; then
  timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${R}_launches_bench4g_L6.csv \
      python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extra-configs > $O/${R}_ncu_launches.log 2>&1; echo "ncu launches rc $?"
  python tools/launch_summary.py $O/${R}_launches_bench4g_L6.csv > $O/${R}_launches_summary.txt
elif [ "$2" = "full" ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_ -s 5 -c 5 -f -o $O/${R}_all_L6_text512m \
      python tools/run_one.py 512 6 > $O/${R}_ncu_full.log 2>&1; echo "ncu full rc $?"    # then here: python tools/ncu_summary.py $O/${R}_all_L6_text512m.ncu-rep
fi
# cycle counters of k_lz77 need the LZPROF=1 build (make -C gzippier_b200/csrc LZPROF=1; never the shipped library): tools/lzprof.py
tail -n 1 $O/${R}_bench_default_1gpu.json | head -c 600; echo; tail -3 $O/${R}_bench.err
