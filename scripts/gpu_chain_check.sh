set -x
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for w in c2c_3d_512_fp32 r2c_1d_1048576x256_fp32 c2r_1d_1048576x256_fp32 c2c_1d_4096x65536_fp32; do
  timeout 300 python bench.py --workload $w --steps 20 --warmup 3 --no-e2e --no-cpu 2>&1 | tail -1 | cut -c1-400
  FFTW_B200_NO_CHAIN=1 timeout 300 python bench.py --workload $w --steps 20 --warmup 3 --no-e2e --no-cpu 2>&1 | tail -1 | cut -c1-400
done
