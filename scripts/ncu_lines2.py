"""Per-source-line hot spots sorted by executed instructions or samples (file-total rows dropped)."""
import csv, subprocess, sys, collections
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
key = sys.argv[4] if len(sys.argv) > 4 else "ins"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr=None; agg=collections.OrderedDict(); fname=""
for r in rows:
    if r and r[0]=="File Path": fname=r[1].split("/")[-1]; continue
    if r and r[0]=="Line No": hdr=r; iS,iI=hdr.index("# Samples"),hdr.index("Instructions Executed"); continue
    if hdr is None or len(r)<=iI or r[0]=='': continue
    try: smp,ins=int(r[iS] or 0),int(r[iI] or 0)
    except ValueError: continue
    k=(fname,r[0],r[1].strip()); a=agg.setdefault(k,[0,0]); a[0]+=smp; a[1]+=ins
tot=sum(a[0] for a in agg.values()) or 1; toti=sum(a[1] for a in agg.values()) or 1
print(f"samples={tot} inst={toti}")
for (f,ln,src),(smp,ins) in sorted(agg.items(), key=lambda kv:-kv[1][1 if key=="ins" else 0])[:top]:
    print(f"{f[:16]:>16}:{ln:>4} {100*smp/tot:5.1f}% smp {100*ins/toti:5.1f}% ins | {src[:95]}")
