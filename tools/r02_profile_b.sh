# round 2, second pass: the kernels changed after tools/r02_profile.sh ran (image epilogue of the two-group kernel; Jungfrau row/column kernel)
set -x
timeout 120 python tools/kbench.py --workload image --cm-mode 4 --iters 1 > gpurun_out/plain_b.log 2>&1 &&
timeout 300 ncu --set full --clock-control none --import-source on -k regex:cm_banks_epix -s 3 -c 1 -f -o gpurun_out/r02_cm_groups_img python tools/kbench.py --workload image --cm-mode 4 --iters 1 > gpurun_out/ncu_b1.log 2>&1
timeout 120 python tools/kbench.py --workload cm --det jungfrau4M --cm-mode 3 --iters 1 >> gpurun_out/plain_b.log 2>&1 &&
timeout 300 ncu --set full --clock-control none --import-source on -k regex:cm_calib_kernel -s 3 -c 1 -f -o gpurun_out/r02_jf_words3 python tools/kbench.py --workload cm --det jungfrau4M --cm-mode 3 --iters 1 > gpurun_out/ncu_b2.log 2>&1
timeout 120 python bench.py --workload image --steps 3 --warmup 3 --also "" --no-cpu-baseline >> gpurun_out/plain_b.log 2>&1 &&
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -s 0 -c 200 --csv --log-file gpurun_out/r02_launches_image.csv python bench.py --workload image --steps 3 --warmup 3 --also "" --no-cpu-baseline > gpurun_out/ncu_b3.log 2>&1
ls -la gpurun_out/*.ncu-rep gpurun_out/*.csv | tail -5
