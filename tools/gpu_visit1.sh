cd /root/repo
timeout 600 python -m pytest tests -m gpu -x -q -k "bilinear or levenberg" > gpurun_out/pytest_ext.log 2>&1; echo pytest=$?; tail -3 gpurun_out/pytest_ext.log
python tools/phase_stats.py --frames 257 > gpurun_out/phase_dense.log 2>&1; cat gpurun_out/phase_dense.log
VSLAMB200_INC=1 python tools/phase_stats.py --frames 257 > gpurun_out/phase_inc.log 2>&1; cat gpurun_out/phase_inc.log
