"""Per-source-line summary (stall samples, warp instructions) of one kernel launch in an .ncu-rep.
usage: python tools/ncu_lines.py <rep> <launch-skip> [top]"""
import csv, subprocess, sys
rep, skip = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", skip,
                      "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
fpath, hdr, recs = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fpath = r[1].split("/")[-1]
    elif r[0] == "Function Name":
        print("#", r[1][:100])
    elif r[0] == "Line No":
        hdr = r
    elif hdr and len(r) >= 8 and r[2] == "-":        # aggregated source line
        try:
            recs.append((fpath, int(r[0]), r[1].strip()[:90], int(r[4] or 0), int(r[7] or 0)))
        except ValueError:
            pass
ts = sum(x[3] for x in recs) or 1
ti = sum(x[4] for x in recs) or 1
print(f"# total samples {ts}, warp instructions {ti}")
for f, ln, src, smp, ins in sorted(recs, key=lambda x: -x[3])[:top]:
    print(f"{100*smp/ts:5.1f}% smp {100*ins/ti:5.1f}% ins  {f}:{ln}  {src}")
