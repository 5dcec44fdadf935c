# 2 / 4-GPU box: slab tests, then cfg5 with the hybrid exchange against the plain DMA exchange
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
[ -z "$SKIP_TESTS" ] && timeout 600 python -m pytest tests/test_gpu_slab.py -m gpu -x -q 2>&1 | tail -2
for v in hybrid dma; do
  unset B2FFT_SLAB_NO_HYBRID
  [ $v = dma ] && export B2FFT_SLAB_NO_HYBRID=1
  echo "== 2048^3 P=$N exchange=$v"
  timeout 300 $TR --nproc-per-node $N --master-port $((29600+RANDOM%300)) bench.py --gpus $N --workload c2c_3d_2048_slab_fp32 --steps 6 --warmup 3 --no-e2e --no-cpu 2>&1 \
    | grep -E '^\{|rror' | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: print(l[:300]); continue
    c = d['config']
    print(c['workload'], 'P=%d' % d['n_gpus'], '%.3f ms' % d['ms_per_step'], '%.0f GF' % d['value'], c.get('stage_ms'), [x.strip()[:70] for x in d['plan'] if 'exchange' in x][:1])
"
done
