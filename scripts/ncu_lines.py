"""Per-source-line stall summary of one kernel from an ncu report (needs -lineinfo and --import-source on).
usage: python scripts/ncu_lines.py report.ncu-rep [top]"""
import collections, csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
cur = None; hdr = None; agg = {}; src = {}
KEYS = ['stall_barrier', 'stall_long_sb', 'stall_lg', 'stall_short_sb', 'stall_mio', 'stall_wait', 'stall_math',
        'stall_branch_resolving', 'stall_membar']
for r in csv.reader(out.splitlines()):
    if not r: continue
    if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r[0] == 'Line No': hdr = r; ix = {k: i for i, k in enumerate(hdr)}; continue
    if hdr and len(r) >= len(hdr) - 1 and r[0].isdigit():
        key = (cur, int(r[0])); src[key] = r[1]
        c = agg.setdefault(key, collections.Counter())
        def g(k):
            try: return int(r[ix[k]] or 0)
            except (ValueError, IndexError): return 0
        c['n'] += g('# Samples'); c['ie'] += g('Instructions Executed')
        for k in KEYS: c[k[6:]] += g(k)
T = collections.Counter()
for c in agg.values(): T.update(c)
print('total', dict(T))
for key, c in sorted(agg.items(), key=lambda kv: -kv[1]['n'])[:top]:
    print(f"{key[0][:18]:18s} {key[1]:4d} {c['n']:6d} {c['ie']:9d}  {src[key].strip()[:72]:72s}",
          {k: v for k, v in c.items() if k not in ('n', 'ie') and v})
