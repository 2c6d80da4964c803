cd /root/repo
for c in 8 4 2; do
echo "== cluster $c"
VSLAMB200_CLUSTER=$c timeout 300 python tools/phase_stats.py --frames 700 2>&1 | head -3
done
echo "== default"
timeout 300 python tools/phase_stats.py --frames 700 2>&1 | head -3
