# 8 GPUs: the default bench (driver path) and slab-only variants of the round-2 schedule
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544"
timeout 600 $TR bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err; echo "bench rc=$?"
slab() { timeout 300 $TR bench.py --gpus 8 --workload c2c_3d_2048_slab_fp32 --steps 8 --warmup 3 --no-slab-checks 2>/dev/null | grep "^{" | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', round(d['ms_per_step'],3), 'ms', {k: round(v,2) for k,v in d['config']['stage_ms'].items()}, 'min', round(d['slab']['min_step_ms'],3))"; }
slab default
B2FFT_SLAB_V2_CHUNKS=4 slab chunks4
B2FFT_SLAB_V2_CHUNKS=16 slab chunks16
B2FFT_SLAB_V2_LAG=1 slab lag1
B2FFT_SLAB_V2_LAG=3 slab lag3
B2FFT_SLAB_NO_V2=1 slab round1_schedule
