cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q -k "pyramid or gradient or preprocess or frame or ingest or kitti" > gpurun_out/pytest_pre.log 2>&1; echo pytest=$?; tail -3 gpurun_out/pytest_pre.log
python tools/pre_stats.py 2>&1 | tail -2
