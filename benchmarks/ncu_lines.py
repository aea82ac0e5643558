"""Summarise an ncu source page (``--page source --csv --print-source cuda,sass``) by CUDA source line."""
import csv
import sys


def main(path, top=40):
    rows = list(csv.reader(open(path)))
    hdr = None
    cur_file = ""
    items = []
    for r in rows:
        if len(r) >= 2 and r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
            continue
        if len(r) > 4 and r[0] == "Line No":
            hdr = r
            i_s = hdr.index("# Samples")
            i_i = hdr.index("Instructions Executed")
            stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") or h.startswith("Stall")]
            continue
        if hdr is None or len(r) <= i_i or not r[0]:
            continue
        try:
            items.append((int(r[i_s]), int(r[i_i]), cur_file, int(r[0]), r[1].strip()[:100], r))
        except ValueError:
            continue
    tot = sum(i[0] for i in items)
    print("total samples", tot)
    items.sort(key=lambda x: -x[0])
    names = [h for h in hdr]
    for s, ins, f, line, src, r in items[:top]:
        # top-2 stall reasons for the line
        reasons = []
        for i, h in enumerate(names):
            if h.startswith("stall_"):
                try:
                    v = int(r[i])
                except ValueError:
                    continue
                if v:
                    reasons.append((v, h[6:]))
        reasons.sort(reverse=True)
        rs = ",".join(f"{n}:{v}" for v, n in reasons[:3])
        print(f"{100 * s / max(tot, 1):5.1f}% inst={ins:9d} {f}:{line:4d} {src}   [{rs}]")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
