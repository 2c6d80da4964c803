# A/B comparison of library builds on ONE box: tools/ab/lib_*.so (git-ignored, built here), each run twice, interleaved
cd /root/repo
for rep in 1 2; do
  for v in tools/ab/lib_*.so; do
    echo "== $v rep $rep"
    VSLAMB200_LIB=$PWD/$v VSLAMB200_INC=${INC:-0} python tools/phase_stats.py --frames ${FRAMES:-257} 2>&1 | head -3
  done
done
