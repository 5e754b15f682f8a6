#!/bin/bash
# One GPU box: the round's bench lines, the ncu launch list and one full capture of k_lz77.  usage: tools/round_measure.sh r01f
# (numbers printed under ncu are never bench values: the bench lines come from the plain runs first)
R=${1:-rXX}; O=gpurun_out; mkdir -p $O
timeout 400 python bench.py --steps 5 --warmup 3 > $O/${R}_bench_L6.json 2> $O/${R}_bench.err; echo "bench rc $?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/${R}_bench_ref.json 2>> $O/${R}_bench.err; echo "ref rc $?"
timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --level 6 --corpus zipf > $O/${R}_bench_L6_zipf.json 2>> $O/${R}_bench.err
timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --level 9 --corpus mixed > $O/${R}_bench_L9_mixed.json 2>> $O/${R}_bench.err
timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --level 1 --corpus random > $O/${R}_bench_L1_random.json 2>> $O/${R}_bench.err
timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --level 1 --corpus text > $O/${R}_bench_L1_text.json 2>> $O/${R}_bench.err
timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --level 9 --corpus text > $O/${R}_bench_L9_text.json 2>> $O/${R}_bench.err
timeout 200 python tools/level_sweep.py 1024 text > $O/${R}_level_sweep.txt 2>&1
timeout 100 python tools/lzprof.py 256 text > $O/${R}_lzprof.txt 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${R}_launches_bench1g_L6.csv python bench.py --mib 1024 --steps 2 --warmup 1 --no-cpu-baseline > $O/${R}_ncu_launches.log 2>&1; echo "ncu launches rc $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_lz77 -s 1 -c 1 -f -o $O/${R}_lz77_L6_text512m python tools/run_one.py 512 6 > $O/${R}_ncu_full.log 2>&1; echo "ncu full rc $?"
cat $O/${R}_bench_L6.json; tail -3 $O/${R}_bench.err
