# ncu --set full of the level-0 launch of the moment kernel on BASELINE configs[4] (1 frame against 8 key frames, 1280x720,
# 16 alignments per call): tools/config_stats.py keyframes, first align call, 4th gn_mom_kernel launch = level 0
cd /root/repo
mkdir -p gpurun_out
timeout 600 python tools/config_stats.py keyframes > gpurun_out/r02_config4_plain.log 2>&1 || exit 1
tail -4 gpurun_out/r02_config4_plain.log
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:gn_mom_kernel -s 3 -c 1 -f -o gpurun_out/r02_config4_mom \
  python tools/config_stats.py keyframes > gpurun_out/r02_config4_ncu.log 2>&1; echo ncu=$?
