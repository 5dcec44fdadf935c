echo "=== default (selective prefetch)"; timeout 300 python scripts/gpu_axis_probe.py fp32 2>&1 | grep -v "^   "
echo "=== E=32 for 512/1024"; B2FFT_LIB=variants/e32/libfftw3.so timeout 300 python scripts/gpu_axis_probe.py fp32 2>&1 | grep -v "^   "
echo "=== E=32 for 512/1024, prefetch everywhere"; FFTW_B200_PF=148 B2FFT_LIB=variants/e32/libfftw3.so timeout 300 python scripts/gpu_axis_probe.py fp32 2>&1 | grep -v "^   "
