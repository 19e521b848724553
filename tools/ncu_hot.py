"""usage: python tools/ncu_hot.py report.ncu-rep [N] -- the N SASS instructions with the most stall samples, plus totals
per opcode class (reads `ncu --page source --csv`)."""
import csv, subprocess, sys
rep = sys.argv[1]
N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(out))
h = rows[1]
ia, isrc, isamp, iex = h.index("Address"), h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
ilong = h.index("stall_long_sb")
body = [r for r in rows[2:] if len(r) == len(h)]
tot = sum(int(r[isamp] or 0) for r in body)
totex = sum(int(r[iex] or 0) for r in body)
print("total samples", tot, "warp instructions", totex, "static", len(body))
for i, r in enumerate(body):
    r.append(i)
top = sorted(body, key=lambda r: -int(r[isamp] or 0))[:N]
for r in sorted(top, key=lambda r: r[-1]):
    print(f"{r[-1]:5d} {int(r[isamp] or 0):7d} {100.0*int(r[isamp] or 0)/tot:5.1f}% long_sb={r[ilong]:>6s} exec={int(r[iex] or 0):>10d}  {r[isrc][:90]}")
# cumulative samples by region of 50 instructions
print("-- samples / executed per block of 50 static instructions")
for b in range(0, len(body), 50):
    s = sum(int(r[isamp] or 0) for r in body[b:b+50]); e = sum(int(r[iex] or 0) for r in body[b:b+50])
    print(f"{b:5d}: samples {100.0*s/tot:5.1f}%  executed {100.0*e/totex:5.1f}%")
