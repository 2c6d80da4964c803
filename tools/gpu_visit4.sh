cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q -k "bounded" > gpurun_out/pytest_b.log 2>&1; echo pytest=$?; tail -8 gpurun_out/pytest_b.log
VSLAMB200_BOUND=0 timeout 300 python tools/phase_stats.py --frames 257 > gpurun_out/phase_d.log 2>&1; head -3 gpurun_out/phase_d.log
VSLAMB200_BOUND=1 timeout 300 python tools/phase_stats.py --frames 257 > gpurun_out/phase_b.log 2>&1; head -4 gpurun_out/phase_b.log
VSLAMB200_BOUND=1 VSLAMB200_INC_CHECK=1 timeout 300 python tools/phase_stats.py --frames 65 > gpurun_out/phase_bc.log 2>&1; head -6 gpurun_out/phase_bc.log
