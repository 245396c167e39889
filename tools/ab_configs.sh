#!/bin/bash
# Developer tool: bench.py's configs legs (device-resident, named sizes) under different environments, on the GPU box.
#   tools/ab_configs.sh C3 "A=0" "HPGQ_FUSE=1"
cd "$(dirname "$0")/.."
cfg="$1"; shift
for e in "$@"; do
  env $e python bench.py --no-e2e --no-cpu --verify-reads 0 --reads 20000000 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); v = d['configs']['$cfg']
print('$e', '$cfg %.2f ms frac %.3f' % (v['ms'], v['frac']), v.get('checksum_job', ''))"
done
