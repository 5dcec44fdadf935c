# 8-GPU box: the driver's scaling run for the default workload at N=8, then cfg5 (2048^3 slab) and 1024^3
# with the exchange variants (default choice / forced fused / DMA on one stream / DMA with per-peer streams).
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
N=${1:-8}
nvidia-smi -L | wc -l
fmt='
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: print(l[:300]); continue
    c = d["config"]
    print(c["workload"], "P=%d" % d["n_gpus"], "%.3f ms" % d["ms_per_step"], "%.0f GF" % d["value"], c.get("stage_ms"), [x.strip()[:90] for x in d["plan"] if "exchange" in x][:1])
'
timeout 600 $TR --nproc-per-node $N --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --no-e2e 2>&1 | grep -E '^\{' | python -c "$fmt"
for wl in c2c_3d_2048_slab_fp32 c2c_3d_1024_slab_fp32; do
  for v in default fused dma dmaN; do
    unset B2FFT_SLAB_NO_FUSED B2FFT_SLAB_PEER_STREAMS B2FFT_SLAB_FUSED
    [ $v = fused ] && export B2FFT_SLAB_FUSED=1
    [ $v = dma ] && export B2FFT_SLAB_NO_FUSED=1
    [ $v = dmaN ] && export B2FFT_SLAB_NO_FUSED=1 B2FFT_SLAB_PEER_STREAMS=1
    echo "== $wl P=$N exchange=$v"
    timeout 600 $TR --nproc-per-node $N --master-port $((29600+RANDOM%300)) bench.py --gpus $N --workload $wl --steps 8 --warmup 3 --no-e2e --no-cpu 2>&1 \
      | grep -E '^\{|rror' | python -c "$fmt"
  done
done
