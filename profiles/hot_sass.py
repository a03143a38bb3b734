"""Top stall-sample SASS lines of an ncu report (source page): python profiles/hot_sass.py rep [N]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
si, ai, src = hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Address"), hdr.index("Source")
body = [r for r in rows[2:] if len(r) > si and r[si].isdigit()]
tot = sum(int(r[si]) for r in body)
print(f"# {rows[0][1][:100]}  total samples {tot}")
idx = {id(r): i for i, r in enumerate(body)}
for r in sorted(body, key=lambda r: -int(r[si]))[:n]:
    print(f"{int(r[si]):6d} {100*int(r[si])/tot:5.1f}%  #{idx[id(r)]:5d}  {r[src].strip()[:100]}")
