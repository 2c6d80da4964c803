cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q -k "incremental" > gpurun_out/pytest_inc.log 2>&1; echo pytest=$?; tail -8 gpurun_out/pytest_inc.log
VSLAMB200_INC=1 timeout 300 python tools/phase_stats.py --frames 257 > gpurun_out/phase_inc.log 2>&1; cat gpurun_out/phase_inc.log
