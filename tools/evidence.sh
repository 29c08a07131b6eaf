#!/bin/bash
# Round evidence on one GPU (run under gpurun): tests, plain bench runs, then the ncu launch
# list and one --set full capture of the same commands.  usage: tools/evidence.sh <tag>
tag=${1:-r01x}
out=gpurun_out
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; tail -2 $out/${tag}_pytest.log
timeout 900 python bench.py > $out/${tag}_bench_soup1m.json 2> $out/${tag}_bench_soup1m.err; echo "bench soup1m rc=$?"
timeout 900 python bench.py --workload flake10m > $out/${tag}_bench_flake10m.json 2> $out/${tag}_bench_flake10m.err; echo "bench flake10m rc=$?"
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err; echo "bench reference rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $out/${tag}_launches_soup1m.csv \
  python bench.py --steps 3 --warmup 3 --no-cpu > $out/${tag}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'trace_kernel_dyn' -c 4 -f -o /tmp/${tag}_trace \
  python bench.py --steps 2 --warmup 1 --no-cpu --staged > $out/${tag}_ncu_trace.log 2>&1; echo "ncu trace rc=$?"
ncu -i /tmp/${tag}_trace.ncu-rep --page raw --csv > $out/${tag}_ncu_trace_raw.csv 2>/dev/null
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'bounds_kernel|morton|chunk_|scatter_kernel|build_walk|build_top' -c 26 -f -o /tmp/${tag}_build \
  python tools/prof_build.py flake10m 30 2 > $out/${tag}_ncu_build.log 2>&1; echo "ncu build rc=$?"
ncu -i /tmp/${tag}_build.ncu-rep --page raw --csv > $out/${tag}_ncu_build_raw.csv 2>/dev/null
# configs 4 and 5 of BASELINE.json (100M-triangle builds with 63- and 30-bit keys; 1e9 random rays,
# plain and with RT_TRACE_REORDER)
timeout 900 python tools/run_configs.py 4 > $out/${tag}_config4_terrain100m.jsonl 2> $out/${tag}_config4.err; echo "config 4 rc=$?"
timeout 900 python tools/run_configs.py 5 1e9 > $out/${tag}_config5_random1b_n1.jsonl 2> $out/${tag}_config5.err; echo "config 5 rc=$?"
timeout 900 python tools/run_configs.py 5 1e9 reorder >> $out/${tag}_config5_random1b_n1.jsonl 2>> $out/${tag}_config5.err; echo "config 5 reorder rc=$?"
ls -la $out | tail -20
