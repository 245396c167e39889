#!/bin/bash
# Sweeps the tiled kernel's ring depth / tile size on the GPU box (developer tool, not part of the bench contract).
cd "$(dirname "$0")/.."
run() {
  echo -n "$* : "
  env "$@" python bench.py --steps 3 --warmup 2 --reads 34000000 --no-e2e --no-cpu 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('value %.3e reads/s  kernel %.0f GB/s frac %.3f  kernel_ms %.4f' % (d['value'], r['achieved'], r['frac'], r['kernel_ms_mean']))
        break
else:
    print('FAILED')
"
}
for cfg in "$@"; do run $cfg; done
