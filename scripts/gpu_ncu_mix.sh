# ncu --set full of three mixed-radix kernels (second launch of each); text + csv summaries travel back, the report does not
python scripts/gpu_mix_one.py > gpurun_out/ncu_plain_mix.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_plain_mix.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:b2_mix -c 6 -o gpurun_out/r2_prof_mix -f python scripts/gpu_mix_one.py > gpurun_out/ncu_mix.log 2>&1
echo "ncu rc=$?"
ncu -i gpurun_out/r2_prof_mix.ncu-rep --page details > gpurun_out/r2_ncu_full_mix.txt 2>&1
ncu -i gpurun_out/r2_prof_mix.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv, sys
rows = list(csv.reader(sys.stdin))
hdr = rows[0]
keep = [i for i, h in enumerate(hdr) if h in ('Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread', 'smsp__inst_executed.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'lts__t_sector_hit_rate.pct', 'launch__grid_size', 'launch__block_size', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_warps', 'sm__maximum_warps_per_active_cycle_pct')]
w = csv.writer(sys.stdout)
for r in rows:
    w.writerow([r[i][:90] for i in keep])
" > gpurun_out/r2_ncu_raw_mix.csv
rm -f gpurun_out/r2_prof_mix.ncu-rep
cat gpurun_out/r2_ncu_raw_mix.csv | cut -c1-400
