# (the full capture runs with VSLAMB200_PARTS=1: one launch per level for the whole batch, the launch bench.py's roofline times)
# round profile: default bench (both arms), ncu launch list of the same command, one full capture of the level-0 kernel
cd /root/repo
R=${1:-r01}
( time python bench.py ) > gpurun_out/${R}_bench.log 2>&1; echo bench=$?; tail -4 gpurun_out/${R}_bench.log | cut -c1-400
( time python bench.py --impl reference ) > gpurun_out/${R}_bench_ref.log 2>&1; echo ref=$?; tail -4 gpurun_out/${R}_bench_ref.log | cut -c1-300
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${R}_launches.csv \
  python bench.py --no-cpu > gpurun_out/${R}_ncu_list.log 2>&1; echo list=$?
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'gn_level_kernel' -s 3 -c 1 -f -o gpurun_out/${R}_gn_level0 \
  env VSLAMB200_PARTS=1 python bench.py --no-cpu --steps 1 > gpurun_out/${R}_ncu_full.log 2>&1; echo full=$?
ls -la gpurun_out | tail -8
