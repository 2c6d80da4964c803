# round 2, first visit: parity of the moment kernel, phase clocks of both level kernels, FFMA2 issue rate
cd /root/repo
mkdir -p gpurun_out
tools/bin/mb_ffma2 > gpurun_out/mb_ffma2.txt 2>&1
timeout 1500 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "evaluate or per_iteration or moment_kernel or batch_of_pairs or real_tum" > gpurun_out/t1.log 2>&1; echo tests=$? >> gpurun_out/t1.log
for k in moments dense; do VSLAMB200_KERNEL=$k timeout 600 python tools/phase_stats.py --frames 257 > gpurun_out/phase_$k.txt 2>&1; done
tail -5 gpurun_out/t1.log; cat gpurun_out/mb_ffma2.txt gpurun_out/phase_moments.txt gpurun_out/phase_dense.txt
