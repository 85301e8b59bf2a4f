#!/bin/bash
# kernel-own durations and DRAM bytes of the costmap kernels (ncu launch list).   bash tools/r02g_probe.sh
o=gpurun_out
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'warp_|propagate' -c 800 --csv --log-file $o/r02g_costmap_launches.csv \
    python bench.py --steps 1 --warmup 0 --no-cpu --no-e2e --no-extra --batch 4096 > $o/ncu_r02g.log 2>&1
python - <<'PY'
import csv, collections
rows=[r for r in csv.reader(l for l in open('gpurun_out/r02g_costmap_launches.csv') if l.startswith('"'))]
h=rows[0]; ki,mi,vi,gi=h.index("Kernel Name"),h.index("Metric Name"),h.index("Metric Value"),h.index("Grid Size")
agg=collections.OrderedDict()
for r in rows[1:]:
    k=(r[ki].split('(')[0], r[gi]); a=agg.setdefault(k,collections.defaultdict(float)); a[r[mi]]+=float(r[vi].replace(',','')); a['n:'+r[mi]]+=1
for k,a in agg.items():
    n=a['n:gpu__time_duration.sum']
    print(k, 'n=%d'%n, 'avg_us=%.1f'%(a['gpu__time_duration.sum']/n/1e3), 'dram_rd_MB=%.1f wr_MB=%.1f per launch'%(a['dram__bytes_read.sum']/n/1e6, a['dram__bytes_write.sum']/n/1e6))
PY
