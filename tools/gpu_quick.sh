# quick visit: a few parity tests of the moment kernel + phase clocks (257 frames)
cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "${TESTS:-evaluate or moment_kernel or batch_of_pairs}" > gpurun_out/t1.log 2>&1; echo tests=$? >> gpurun_out/t1.log
tail -4 gpurun_out/t1.log
for k in ${KERNELS:-moments}; do VSLAMB200_KERNEL=$k timeout 600 python tools/phase_stats.py --frames ${FRAMES:-257} 2>&1 | tee gpurun_out/phase_$k.txt; done
