cd /root/repo
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_all.log 2>&1; echo pytest=$?; tail -3 gpurun_out/pytest_all.log
FRAMES=700 bash tools/ab.sh 2>&1 | grep "level 0\|==\|ms_gn"
