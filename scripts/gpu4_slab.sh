# 4-GPU box: cfg5 (2048^3) and 1024^3 slab with the default exchange choice
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
N=${1:-4}
for wl in c2c_3d_2048_slab_fp32 c2c_3d_1024_slab_fp32; do
  timeout 300 $TR --nproc-per-node $N --master-port $((29600+RANDOM%300)) bench.py --gpus $N --workload $wl --steps 6 --warmup 3 --no-e2e --no-cpu 2>&1 \
    | grep -E '^\{|rror' | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: print(l[:300]); continue
    c = d['config']
    print(c['workload'], 'P=%d' % d['n_gpus'], '%.3f ms' % d['ms_per_step'], '%.0f GF' % d['value'], c.get('stage_ms'), [x.strip()[:80] for x in d['plan'] if 'exchange' in x][:1])
"
done
