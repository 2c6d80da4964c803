# A/B of library builds in tools/ab on one box: level times and level-0 phase clocks (999 jobs by default)
cd /root/repo
for v in ${VARIANTS}; do echo "== $v"; VSLAMB200_LIB=$PWD/tools/ab/lib_$v.so VSLAMB200_KERNEL=${KERNEL:-moments} timeout 300 python tools/phase_stats.py --frames ${FRAMES:-1000} 2>&1 | head -2; done
echo "== working tree"; VSLAMB200_KERNEL=${KERNEL:-moments} timeout 300 python tools/phase_stats.py --frames ${FRAMES:-1000} 2>&1 | head -2
