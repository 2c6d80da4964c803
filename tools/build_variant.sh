# tools/build_variant.sh NAME "-DFLAG ..." : builds tools/ab/lib_NAME.so from the working tree with extra nvcc flags
set -e
cd /root/repo/vslam_b200/csrc
mkdir -p /tmp/t/var_$1 /root/repo/tools/ab
for f in preprocess lists level_misc gn_level gn_mom ext api; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $2 -c $f.cu -o /tmp/t/var_$1/$f.o &
done
wait
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o /root/repo/tools/ab/lib_$1.so /tmp/t/var_$1/*.o
