set -e
cd /root/repo
python bench.py --frames 24 --steps 1 --warmup 1 --no-cpu > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'residual_kernel|accumulate_kernel' -s 60 -c 4 -o gpurun_out/prof_r1a python bench.py --frames 24 --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/plain.log
