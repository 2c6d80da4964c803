# level-kernel policy on small batches (what one rank of an 8-GPU strong-scaling run sees): cluster size x stream parts
cd /root/repo
mkdir -p gpurun_out
for P in ${PAIRS:-512}; do
for parts in 0 1; do
for cl in 0 2 4 8 16; do
  VSLAMB200_CLUSTER=$cl VSLAMB200_PARTS=$parts timeout 200 python bench.py --no-cpu --pairs $P --steps 5 --warmup 2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('pairs $P parts $parts cluster $cl: ms_per_step %.2f levels %s lists %.2f' % (d['ms_per_step'], [round(x,2) for x in d['roofline']['all_levels']['ms']], d['roofline']['ms_lists']))"
done; done; done
