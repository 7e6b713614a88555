"""Per-instruction view of an .ncu-rep source page: SASS, executed count, stall samples, shared-memory wavefronts.
Usage: python tools/ncu_source.py REP [min_samples]"""
import csv, io, subprocess, sys

rep = sys.argv[1]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
lines = txt.splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith('"Address"'))
end = next((i for i in range(start + 1, len(lines)) if lines[i].startswith('"Kernel Name"')), len(lines))  # first launch only
rows = list(csv.reader(io.StringIO("\n".join(lines[start:end]))))
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
tot_s = sum(int(r[ix["# Samples"]] or 0) for r in rows[1:] if len(r) == len(hdr))
tot_i = sum(int(r[ix["Instructions Executed"]] or 0) for r in rows[1:] if len(r) == len(hdr))
print(f"total samples {tot_s} total warp-instr {tot_i}")
for r in rows[1:]:
    if len(r) != len(hdr):
        continue
    s = int(r[ix["# Samples"]] or 0)
    n = int(r[ix["Instructions Executed"]] or 0)
    wf = r[ix["L1 Wavefronts Shared"]]
    wfi = r[ix["L1 Wavefronts Shared Ideal"]]
    tg = r[ix["L1 Tag Requests Global"]]
    print(f"{r[ix['Address']][-5:]} {n:>9} {float(r[ix['Avg. Threads Executed']] or 0):5.1f} {s:>6} {100.0*s/max(tot_s,1):5.2f}% wf={wf:>9}/{wfi:>9} tag={tg:>8}  {r[ix['Source']][:90]}")
