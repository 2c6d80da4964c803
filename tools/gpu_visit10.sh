cd /root/repo
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_all.log 2>&1; echo pytest=$?; tail -3 gpurun_out/pytest_all.log
python bench.py --no-cpu --frames 200 --steps 2 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value',d['value'],'latency',d['latency'])"
