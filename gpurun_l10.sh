ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file /tmp/l.csv python bench.py --steps 2 --warmup 3 --no-cpu --variant 1 > /dev/null 2>&1
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(open('/tmp/l.csv')) if len(r)>5]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value'); mi=hdr.index('Metric Name')
agg={}
for r in rows[1:]:
    try: v=float(r[vi].replace(',',''))
    except: continue
    a=agg.setdefault((r[ki][:36],r[mi]),[0,0.0]); a[0]+=1; a[1]+=v
for (n,m),a in sorted(agg.items(), key=lambda t:(t[0][0],t[0][1])):
    if a[0]>=4 and ('k_' in n): print("%-38s %-24s n=%3d avg=%12.1f"%(n,m,a[0],a[1]/a[0]))
PY
