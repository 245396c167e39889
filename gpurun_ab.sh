for rep in 1 2; do
echo "== split tail"; python tools/probe_fixed.py 150 | tail -2; python tools/probe_fixed.py 100 | tail -1
echo "== head"; HPGQ_LIB=/root/repo/hpg-fastq_b200/csrc/build_head/libhpgq_cuda.so python tools/probe_fixed.py 150 | tail -2; HPGQ_LIB=/root/repo/hpg-fastq_b200/csrc/build_head/libhpgq_cuda.so python tools/probe_fixed.py 100 | tail -1
done
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "fixed_length or flush or uniform or bench_shaped or foreign or full_options or lowercase" 2>&1 | tail -2
