# 8-GPU box: default bench under torchrun at N=2,4,8 (the driver's scaling run), the slab test, and the
# cfg5 slab workloads at P=8,4,2.
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
nvidia-smi -L | head -8
for n in 2 4 8; do
  timeout 600 $TR --nproc-per-node $n --master-port $((29500+n)) bench.py --gpus $n --steps 10 --warmup 3 2>&1 | grep -E '^\{' | cut -c1-700
done
timeout 900 python -m pytest tests/test_gpu_slab.py -m gpu -x -q 2>&1 | tail -3
for n in 8 4 2; do
  timeout 900 $TR --nproc-per-node $n --master-port $((29600+n)) bench.py --gpus $n --workload c2c_3d_2048_slab_fp32 --steps 5 --warmup 2 --no-e2e --no-cpu 2>&1 | grep -E '^\{|rror' | cut -c1-900
done
timeout 600 $TR --nproc-per-node 8 --master-port 29700 bench.py --gpus 8 --workload c2c_3d_1024_slab_fp32 --steps 10 --warmup 3 --no-e2e --no-cpu 2>&1 | grep -E '^\{|rror' | cut -c1-900
