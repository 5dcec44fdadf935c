# 8-GPU box: cfg5 (2048^3 slab, DMA exchange) against the number of copy streams and of plane chunks
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
N=${1:-8}
fmt='
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: print(l[:300]); continue
    c = d["config"]
    print(c["workload"], "P=%d" % d["n_gpus"], "%.3f ms" % d["ms_per_step"], "%.0f GF" % d["value"], c.get("stage_ms"))
'
for v in "1 8" "2 8" "3 8" "1 16" "2 4"; do
  set -- $v
  export B2FFT_SLAB_PEER_STREAMS=$1 B2FFT_SLAB_CHUNKS=$2
  echo "== 2048^3 P=$N copy streams=$1 chunks=$2"
  timeout 300 $TR --nproc-per-node $N --master-port $((29600+RANDOM%300)) bench.py --gpus $N --workload c2c_3d_2048_slab_fp32 --steps 6 --warmup 3 --no-e2e --no-cpu 2>&1 \
    | grep -E '^\{|rror' | python -c "$fmt"
done
