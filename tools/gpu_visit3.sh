cd /root/repo
VSLAMB200_INC=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gn_inc_kernel' -s 2 -c 1 -f -o gpurun_out/prof_inc python tools/phase_stats.py --frames 129 > gpurun_out/ncu_inc.log 2>&1; echo ncu=$?; tail -3 gpurun_out/ncu_inc.log
