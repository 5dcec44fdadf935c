# bench a set of workloads, device-resident only; prints "workload ms"
for w in c2c_1d_4096x65536_fp32 c2c_3d_512_fp32 r2c_1d_1048576x256_fp32 c2r_1d_1048576x256_fp32 c2c_1d_1024_fp64; do
  for nc in 1 0; do
    if [ $nc = 1 ]; then export FFTW_B200_NO_CHAIN=1; else unset FFTW_B200_NO_CHAIN; fi
    timeout 300 python bench.py --workload $w --steps 20 --warmup 3 --no-e2e --no-cpu 2>&1 | tail -1 | sed -E "s/.*\"ms_per_step\": ([0-9.]+).*\"workload\": \"([a-z0-9_]+)\".*\"frac\": ([0-9.]+).*/\2 nochain=$nc \1 ms frac \3/" | cut -c1-200
  done
done
