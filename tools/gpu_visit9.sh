cd /root/repo
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/pre_launches.csv python tools/pre_stats.py > gpurun_out/pre_ncu.log 2>&1; echo ncu=$?
tail -1 gpurun_out/pre_ncu.log
