set -x
python tools/kbench.py --workload cm --cm-mode 4 --iters 1 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:cm_banks_epix -s 3 -c 1 -o gpurun_out/r02_cm_groups python tools/kbench.py --workload cm --cm-mode 4 --iters 1 > gpurun_out/ncu1.log 2>&1
python tools/kbench.py --workload image --cm-mode 4 --iters 1 >> gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:cm_banks_epix -s 3 -c 1 -o gpurun_out/r02_cm_groups_img python tools/kbench.py --workload image --cm-mode 4 --iters 1 > gpurun_out/ncu2.log 2>&1
python tools/kbench.py --workload cm --det jungfrau4M --cm-mode 4 --iters 1 >> gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:cm_banks_kernel -s 3 -c 1 -o gpurun_out/r02_jf_cm_banks python tools/kbench.py --workload cm --det jungfrau4M --cm-mode 4 --iters 1 > gpurun_out/ncu3.log 2>&1
python tools/kbench.py --workload image_nocm --iters 1 >> gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none -k regex:image_scatter -s 3 -c 1 -o gpurun_out/r02_image_scatter python tools/kbench.py --workload image_nocm --iters 1 > gpurun_out/ncu4.log 2>&1
python tools/kbench.py --workload jungfrau --iters 1 >> gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none -k regex:jf_calib_map -s 3 -c 1 -o gpurun_out/r02_jf_map python tools/kbench.py --workload jungfrau --iters 1 > gpurun_out/ncu5.log 2>&1
# launch lists of the bench command, workload by workload (cold-cache, serialised: shares, not absolutes)
python bench.py --steps 3 --warmup 3 --also "" --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 0 -c 200 --csv --log-file gpurun_out/r02_launches_calib.csv python bench.py --steps 3 --warmup 3 --also "" --no-cpu-baseline > gpurun_out/ncu6.log 2>&1
python bench.py --workload image --steps 3 --warmup 3 --also "" --no-cpu-baseline >> gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 0 -c 200 --csv --log-file gpurun_out/r02_launches_image.csv python bench.py --workload image --steps 3 --warmup 3 --also "" --no-cpu-baseline > gpurun_out/ncu7.log 2>&1
ls -la gpurun_out/*.ncu-rep gpurun_out/*.csv
