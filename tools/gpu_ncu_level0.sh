# ncu --set full of the level-0 launch of the level kernel on 257 frames (256 jobs), all iterations
cd /root/repo
mkdir -p gpurun_out
K=${KERNEL:-moments}
NAME=${NAME:-gn_mom_kernel}
VSLAMB200_KERNEL=$K VSLAMB200_PARTS=1 timeout 600 python tools/phase_stats.py --frames ${FRAMES:-257} --timing 0 > gpurun_out/ncu_plain.log 2>&1 || exit 1
VSLAMB200_KERNEL=$K VSLAMB200_PARTS=1 timeout 1500 ncu --set full --clock-control none --import-source on -k regex:$NAME -s 7 -c 1 -f -o gpurun_out/${OUT:-prof_mom} \
  python tools/phase_stats.py --frames ${FRAMES:-257} --timing 0 > gpurun_out/ncu_full.log 2>&1; echo ncu=$?
tail -3 gpurun_out/ncu_full.log
