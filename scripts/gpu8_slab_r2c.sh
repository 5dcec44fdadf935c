# 8 GPUs: flags vs NCCL barrier, split vs one compute stream, chunk / group counts
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544"
slab() { timeout 200 $TR bench.py --gpus 8 --workload c2c_3d_2048_slab_fp32 --steps 8 --warmup 3 --no-slab-checks 2>/dev/null | grep "^{" | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1', round(d['ms_per_step'],3), 'ms', {k: round(v,2) for k,v in d['config']['stage_ms'].items()}, 'min', round(d['slab']['min_step_ms'],3))"; }
slab "flags_split_g2_ch8(default)"
B2FFT_SLAB_V2_CHUNKS=4 slab flags_split_g2_ch4
B2FFT_SLAB_V2_GROUPS=1 slab flags_split_g1_ch8
B2FFT_SLAB_V2_CHUNKS=16 slab flags_split_g2_ch16
B2FFT_SLAB_V2_NCCL_BARRIER=1 slab nccl_split_g2_ch8
B2FFT_SLAB_V2_ONE_STREAM=1 slab flags_onestream_g2_ch8
