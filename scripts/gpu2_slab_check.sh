# 2-GPU box: slab tests over NCCL (fused and DMA exchange), then 1024^3 / 2048^3 timings of the three
# exchange variants.  Usage: bash scripts/gpu2_slab_check.sh [nproc]
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi -L | head -8
timeout 900 python -m pytest tests/test_gpu_slab.py -m gpu -x -q 2>&1 | tail -3
for wl in c2c_3d_1024_slab_fp32 c2c_3d_2048_slab_fp32; do
  for v in fused dma dmaN; do
    unset B2FFT_SLAB_NO_FUSED B2FFT_SLAB_PEER_STREAMS
    [ $v = dma ] && export B2FFT_SLAB_NO_FUSED=1
    [ $v = dmaN ] && export B2FFT_SLAB_NO_FUSED=1 B2FFT_SLAB_PEER_STREAMS=1
    echo "== $wl P=$N exchange=$v"
    timeout 600 $TR --nproc-per-node $N --master-port $((29800+RANDOM%100)) bench.py --gpus $N --workload $wl --steps 8 --warmup 3 --no-e2e --no-cpu 2>&1 \
      | grep -E '^\{|rror' | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: print(l[:300]); continue
    print(d['config']['workload'], 'P=%d' % d['n_gpus'], '%.3f ms' % d['ms_per_step'], '%.0f GF' % d['value'], d['config']['stage_ms'], [x for x in d['plan'] if 'exchange' in x][:1])
"
  done
done
