for w in 4 2 3 6 8; do
  echo "== waves $w"
  HPGQ_POS_WAVES=$w python tools/bench_configs.py --configs C5 --steps 8 --warmup 3 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['kernel_ms'], d['submit_ms'])"
done
python -m pytest tests/test_gpu_parity.py -q -m gpu -k "long_reads or group_kernel" 2>&1 | tail -1
