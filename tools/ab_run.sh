#!/bin/bash
# usage: tools/ab_run.sh <workload> <lib>...   -> quick_bench of every library (VOLRT_LIB), twice, interleaved
wl=$1; shift
for rep in 1 2; do
  for lib in "$@"; do
    echo -n "$lib: "; VOLRT_LIB=$lib tools/quick_bench.sh $wl
  done
done
