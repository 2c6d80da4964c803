cd /root/repo
VSLAMB200_BOUND=1 timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:gn_level_kernel<.bool.1>' -s 2 -c 1 -f -o gpurun_out/prof_bound python tools/phase_stats.py --frames 257 > gpurun_out/ncu_bound.log 2>&1; echo ncu=$?; tail -3 gpurun_out/ncu_bound.log
