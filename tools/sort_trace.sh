#!/bin/bash
# per-entry-point GPU time of one DirectSort: bash tools/sort_trace.sh <N> <out>
FHEB200_TRACE=1 timeout 1500 python tools/sort_latency.py $1 > $2.json 2> $2.trace
cat $2.json; cat $2.trace
