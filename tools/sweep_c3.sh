#!/bin/bash
# Developer tool: one-pass edit + filter + compaction kernel, tile size A/B on the GPU box (C3 batch of 2^22 reads).
cd "$(dirname "$0")/.."
run() { echo -n "$* : "; env "$@" python tools/bench_configs.py --configs C3 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('kernel_ms %.3f submit_ms %.3f'%(d['kernel_ms'],d['submit_ms']))"; }
for cfg in "$@"; do run $cfg; done
