GFORB_POISON_PYR=1 timeout -s KILL 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout -s KILL 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -k regex:pyr -c 8 --csv --log-file gpurun_out/pyr6.csv python bench.py --steps 2 --warmup 1 > gpurun_out/pyr6.log 2>&1
python - <<'P'
import csv
rows=[r for r in csv.reader(open('gpurun_out/pyr6.csv')) if len(r)>10]
h=rows[0]; ik=h.index('Kernel Name'); im=h.index('Metric Name'); iv=h.index('Metric Value'); ii=h.index('ID')
out={}
for r in rows[1:]:
    out.setdefault((r[ii], r[ik][:30]), {})[r[im].split('__')[1][:12]]=r[iv]
for k,v in out.items(): print(k, v)
P
