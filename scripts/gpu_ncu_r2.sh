# ncu captures for profiles/traffic.json: per-kernel DRAM bytes and durations of the cfg 3 / cfg 4 kernels
B="--steps 1 --warmup 3 --no-e2e --no-cpu --no-slab --no-configs"
for w in c2c_3d_512_fp32 r2c_1d_1048576x256_fp32 c2r_1d_1048576x256_fp32; do
  case $w in c2c_3d_512_fp32) k=3;; r2c_*) k=2;; *) k=3;; esac
  python bench.py --workload $w $B > gpurun_out/ncu_plain_$w.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:b2_ -s $((3*k)) -c $k -o gpurun_out/r2_prof_$w -f python bench.py --workload $w $B > gpurun_out/ncu_$w.log 2>&1
  echo "$w ncu rc=$?"
  # text summaries travel back; the raw report (25 MB each) does not fit the 64 MiB return limit
  ncu -i gpurun_out/r2_prof_$w.ncu-rep --page details > gpurun_out/r2_ncu_full_$w.txt 2>&1
  ncu -i gpurun_out/r2_prof_$w.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv, sys
rows = list(csv.reader(sys.stdin))
hdr = rows[0]
keep = [i for i, h in enumerate(hdr) if h in ('Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread', 'smsp__inst_executed.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'lts__t_sector_hit_rate.pct', 'launch__grid_size', 'launch__block_size', 'smsp__issue_active.avg.pct_of_peak_sustained_active')]
w = csv.writer(sys.stdout)
for r in rows:
    w.writerow([r[i] for i in keep])
" > gpurun_out/r2_ncu_raw_$w.csv
  rm -f gpurun_out/r2_prof_$w.ncu-rep
done
# launch list of the default step (shares, not absolutes)
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-slab --no-configs > gpurun_out/ncu_plain_default.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/r2_launches_default.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-slab --no-configs > gpurun_out/ncu_default.log 2>&1
echo "launch list rc=$?"
