"""Top stall-sampled SASS lines of one kernel in an ncu report (needs --import-source on / -lineinfo).
Usage: python tools/ncu_hot.py report.ncu-rep [kernel-regex] [launch-skip] [top-n]"""
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    pat = sys.argv[2] if len(sys.argv) > 2 else "."
    skip = sys.argv[3] if len(sys.argv) > 3 else "0"
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{pat}",
                          "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    print(rows[0][1][:120])
    head, body = rows[1], rows[2:]
    col = {h: i for i, h in enumerate(head)}
    stalls = [h for h in head if h.startswith("stall_") and "Not Issued" not in h]
    data = []
    for n, r in enumerate(body):
        try:
            data.append((int(r[col["# Samples"]]), int(r[col["Instructions Executed"]]), n, r))
        except (ValueError, IndexError):
            pass
    tot = sum(d[0] for d in data)
    print("samples", tot, "warp-instructions", sum(d[1] for d in data))
    agg = {s: sum(int(d[3][col[s]] or 0) for d in data) for s in stalls}
    print("stall totals:", {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
    for smp, ex, n, r in sorted(data, reverse=True)[:top]:
        why = ",".join(f"{s[6:]}={r[col[s]]}" for s in stalls if r[col[s]] not in ("0", ""))
        print(f"{smp:6d} {100 * smp / max(tot, 1):5.1f}%  ex={ex:9d} line={n:5d}  {r[col['Source']][:70]:70s} {why}")


if __name__ == "__main__":
    main()
