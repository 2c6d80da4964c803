# one GPU visit: tests, bench, full ncu capture of the level kernel
cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo pytest=$?; tail -5 gpurun_out/pytest.log
timeout 600 python bench.py --frames 400 --steps 2 --warmup 3 --cpu-pairs 2 > gpurun_out/bench400.log 2>&1; echo bench=$?; tail -1 gpurun_out/bench400.log
if [ "$1" = "ncu" ]; then
timeout 300 python tools/prof_level0.py --frames 257 --iters 4 > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gn_level_kernel' -s 1 -c 1 -f -o gpurun_out/prof_gn python tools/prof_level0.py --frames 257 --iters 4 > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/prof_plain.log
fi
