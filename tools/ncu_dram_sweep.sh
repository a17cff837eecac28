set -x
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,sm__issue_active.avg.pct_of_peak_sustained_elapsed,lts__throughput.avg.pct_of_peak_sustained_elapsed
run() { name=$1; shift; timeout 200 ncu --metrics $M --clock-control none -s 6 -c 1 --csv --log-file gpurun_out/sweep_$name.csv python tools/kbench.py "$@" --iters 1 > /dev/null 2>&1; }
run epix_cm2 --workload cm --cm-mode 2
run epix_cm7 --workload cm --cm-mode 7
run jf_cm3 --workload cm --det jungfrau4M --cm-mode 3
run jf_cm4 --workload cm --det jungfrau4M --cm-mode 4
run jf_calib --workload jungfrau
python - <<PY
import csv,glob
for f in sorted(glob.glob("gpurun_out/sweep_*.csv")):
    rows=[r for r in csv.reader(open(f)) if len(r)>10]
    h=rows[0]; ik=h.index("Kernel Name"); im=h.index("Metric Name"); iv=h.index("Metric Value")
    print(f, rows[1][ik][:60])
    for r in rows[1:]: print("   ", r[im], r[iv])
PY
