"""Per-source-line instruction / stall-sample shares from an .ncu-rep (needs -lineinfo)."""
import csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
cur = None; per = {}; hdr = None
for row in csv.reader(out.splitlines()):
    if len(row) == 2 and row[0] == 'File Path': cur = row[1].split('/')[-1]; continue
    if len(row) > 5 and row[0] == 'Line No': hdr = row; continue
    if hdr and len(row) > 8 and row[0] != '':
        try:
            ln = int(row[0]); ie = int(row[hdr.index('Instructions Executed')])
            te = int(row[hdr.index('Thread Instructions Executed')]); sm = int(row[hdr.index('# Samples')])
        except Exception:
            continue
        per[(cur, ln)] = (ie, te, sm, row[1].strip()[:95])
tot = sum(v[0] for v in per.values()); tots = sum(v[2] for v in per.values())
print('total warp inst', tot, 'thread inst', sum(v[1] for v in per.values()), 'samples', tots)
for k, v in sorted(per.items(), key=lambda kv: -kv[1][int(sys.argv[3]) if len(sys.argv) > 3 else 2])[:top]:
    print(f"{k[0]:18s}:{k[1]:4d} inst={v[0]/tot*100:5.1f}% thr/inst={v[1]/max(v[0],1):5.1f} samp={v[2]/tots*100:5.1f}% | {v[3]}")
