set -x
for lag in 1 2 3 4 6 8 14; do
  FFTW_B200_CHAIN_LAG=$lag timeout 300 python bench.py --workload c2c_3d_512_fp32 --steps 20 --warmup 3 --no-e2e --no-cpu 2>&1 | tail -1 | sed -E "s/.*\"ms_per_step\": ([0-9.]+).*/lag $lag 512^3 \1 ms/"
done
for lag in 1 2 4; do
  FFTW_B200_CHAIN_LAG=$lag timeout 300 python bench.py --workload r2c_1d_1048576x256_fp32 --steps 20 --warmup 3 --no-e2e --no-cpu 2>&1 | tail -1 | sed -E "s/.*\"ms_per_step\": ([0-9.]+).*/lag $lag r2c \1 ms/"
done
FFTW_B200_CHAIN_LAG=4 timeout 600 ncu --set full --clock-control none --import-source on -k regex:b2_chain_kernel -s 3 -c 1 -o gpurun_out/r1_chain512_lag4 -f python bench.py --workload c2c_3d_512_fp32 --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_chain.log 2>&1
tail -3 gpurun_out/ncu_chain.log
