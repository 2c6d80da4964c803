cd /root/repo
mkdir -p gpurun_out
for c in ${CLUSTERS:-1 2 4}; do echo "== cluster $c"; VSLAMB200_KERNEL=${KERNEL:-moments} VSLAMB200_CLUSTER=$c timeout 600 python tools/phase_stats.py --frames ${FRAMES:-1000} 2>&1 | head -${LINES:-3}; done 2>&1 | tee gpurun_out/sweep.txt
