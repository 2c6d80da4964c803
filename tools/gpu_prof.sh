cd /root/repo
timeout 300 python tools/prof_level0.py --frames 129 --iters 100 --levels 4 > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gn_level_kernel' -s 3 -c 1 -f -o gpurun_out/prof_gn python tools/prof_level0.py --frames 129 --iters 100 --levels 4 > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/prof_plain.log | cut -c1-300
