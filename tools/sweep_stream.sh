#!/bin/bash
# Rebuilds bed_ingest.o with different compile-time parameters of the streaming scanner and times the bench step.
# usage (on a GPU box): tools/sweep_stream.sh "-DSALSA_STREAM_CHUNK=262144" "-DSALSA_STREAM_RING=16384 -DSALSA_STREAM_WARPS=11" ...
cd "$(dirname "$0")/.."
for flags in "$@"; do
    make -C salsa_b200/csrc -s clean >/dev/null 2>&1
    make -C salsa_b200/csrc -s -j8 NVFLAGS_EXTRA="$flags" >/dev/null 2>&1 || { echo "{\"flags\": \"$flags\", \"error\": \"build failed\"}"; continue; }
    r=$(SALSA_BENCH_ABLATION=$ABL python bench.py --steps 5 --warmup 3 --no-extras 2>/dev/null | tail -1)
    echo "{\"flags\": \"$flags\", \"result\": $r}"
done
make -C salsa_b200/csrc -s clean >/dev/null 2>&1
make -C salsa_b200/csrc -s -j8 >/dev/null 2>&1
