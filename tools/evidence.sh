#!/bin/bash
# Round-2 evidence on one GPU (run under gpurun): plain bench runs, sort vs CUB, then the ncu
# launch list of the bench command and one --set full capture of the trace launches.
# usage: tools/evidence.sh <tag>
tag=${1:-r02x}
out=gpurun_out
mkdir -p $out
timeout 600 python bench.py > $out/${tag}_bench_flake10m.json 2> $out/${tag}_bench_flake10m.err; echo "bench flake10m rc=$?"
timeout 300 python bench.py --workload soup1m --no-configs > $out/${tag}_bench_soup1m.json 2> $out/${tag}_bench_soup1m.err; echo "bench soup1m rc=$?"
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err; echo "bench reference rc=$?"
timeout 300 python tools/bench_sort.py > $out/${tag}_sort_vs_cub.jsonl 2> $out/${tag}_sort.err; echo "sort rc=$?"
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu --no-configs --staged > $out/${tag}_plain_short.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $out/${tag}_launches_flake10m.csv \
  python bench.py --steps 3 --warmup 3 --no-cpu --no-configs --staged > $out/${tag}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
timeout 300 python tools/prof_bench_trace.py > $out/${tag}_prof_trace_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:trace_kernel -c 4 -f -o $out/${tag}_trace \
  python tools/prof_bench_trace.py > $out/${tag}_ncu_trace.log 2>&1; echo "ncu trace rc=$?"
ncu -i $out/${tag}_trace.ncu-rep --page raw --csv > $out/${tag}_trace_raw.csv 2>/dev/null
cat $out/${tag}_prof_trace_plain.log
ls -la $out | tail -20
