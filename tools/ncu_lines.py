"""tools/ncu_lines.py -- per-source-line hot spots from an .ncu-rep (needs -lineinfo and --import-source on):
    python tools/ncu_lines.py <report.ncu-rep> [launch-index] [top-N] [inst]      (inst: sort by instruction share)
prints share of warp instructions, average active threads and stall samples per CUDA source line."""
import csv, subprocess, sys, io
rep = sys.argv[1]; top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
cur = ""; seen = set(); hdr = None; data = []; kernel_idx = -1; want = int(sys.argv[2]) if len(sys.argv) > 2 else 0
for r in rows:
    if not r: continue
    if r[0] == "File Path":
        if kernel_idx < 0 or r[1] in seen: kernel_idx += 1; seen = set()
        seen.add(r[1]); cur = r[1].split("/")[-1]
    elif r[0] == "Line No": hdr = r
    elif hdr and r[0].isdigit() and kernel_idx == want and len(r) > 12:
        ie, te, sm = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
        try: data.append((int(r[ie]), int(r[te]), int(r[sm]), cur, int(r[0]), r[1].strip()[:100]))
        except ValueError: pass
tot = sum(d[0] for d in data); ts = sum(d[2] for d in data)
print(f"total warp inst {tot}, thread inst/warp inst {sum(d[1] for d in data)/max(tot,1):.2f}, samples {ts}")
by = 0 if (len(sys.argv) > 4 and sys.argv[4] == "inst") else 2
for d in sorted(data, key=lambda x: -x[by])[:top]:
    print(f"inst {d[0]/tot:6.3f} thr {d[1]/max(d[0],1):5.1f} samp {d[2]/max(ts,1):6.3f} {d[3]}:{d[4]:<4d} {d[5]}")
