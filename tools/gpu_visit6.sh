cd /root/repo
VSLAMB200_BOUND=1 timeout 300 python tools/phase_stats.py --frames 257 > gpurun_out/phase_b.log 2>&1; head -5 gpurun_out/phase_b.log
