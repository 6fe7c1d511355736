"""Stall-sample summary of an ncu report with source counters: totals by stall reason, the top instructions, samples by
instruction-index bin.  usage: python tools/ncu_stalls.py report.ncu-rep [top_n]"""
import csv, sys, subprocess
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address'][0]
H = rows[hi]; data = [r for r in rows[hi + 1:] if len(r) >= len(H)]
ix = {n: i for i, n in enumerate(H)}
stalls = [n for n in H if n.startswith('stall_') and 'Not Issued' not in n]
tot = sum(int(r[ix['# Samples']]) for r in data)
agg = {n[6:]: sum(int(r[ix[n]]) for r in data) for n in stalls}
nw = max(int(r[ix['Instructions Executed']]) for r in data)
print('total samples', tot, 'static instrs', len(data), 'dynamic warp instrs per warp', sum(int(r[ix['Instructions Executed']]) for r in data) / nw)
print({k: "%d (%.1f%%)" % (v, 100.0 * v / tot) for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
per = sorted([(int(r[ix['# Samples']]), k, r[ix['Source']].strip()[:80]) for k, r in enumerate(data)], reverse=True)
for s, k, src in per[:top]:
    r = data[k]; st = {n[6:]: int(r[ix[n]]) for n in stalls if int(r[ix[n]]) > tot * 0.0003}
    print(s, k, src, st)
b = {}
for s, k, src in per:
    b[k // 200] = b.get(k // 200, 0) + s
print('samples per 200-instruction bin:', [(k * 200, v) for k, v in sorted(b.items())])
