#!/bin/bash
# Developer tool: C5 (1-20 kb reads) per-batch timing of library variants (HPGQ_LIB=...), run on the GPU box.
cd "$(dirname "$0")/.."
run() { echo -n "$* : "; env "$@" python tools/bench_configs.py --configs C5 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('kernel_ms %.3f submit_ms %.3f gbs %.0f'%(d['kernel_ms'],d['submit_ms'],d['gbs_total']))"; }
for cfg in "$@"; do run $cfg; done
