#!/usr/bin/env python
"""Per source line share of executed warp instructions and stall samples of one kernel, in line order, from the CSV of
`ncu -i rep --page source --csv --print-source cuda,sass`.   python tools/ncu_lines.py src.csv k_select [min_pct]"""
import csv
import sys


def main():
    path, want = sys.argv[1], sys.argv[2]
    thr = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
    text = open(sys.argv[4]).read().split("\n") if len(sys.argv) > 4 else None      # the source as it was compiled
    rows = list(csv.reader(open(path)))
    i = 0
    while i < len(rows):
        r = rows[i]
        if r and r[0] == "Function Name" and want in r[1] and i + 1 < len(rows) and rows[i + 1] and rows[i + 1][0] == "Line No":
            hdr = rows[i + 1]
            col = {}
            for j, n in enumerate(hdr):
                col.setdefault(n, j)
            if "Instructions Executed" not in col:
                i += 1
                continue
            body = []
            i += 2
            while i < len(rows) and rows[i] and len(rows[i]) >= len(hdr) and rows[i][0] != "Line No":
                r = rows[i]
                if len(r) > len(hdr):         # a source line with quotes or commas of its own: the metric columns are the last ones
                    extra = len(r) - len(hdr)
                    r = [r[0], ",".join(r[1:2 + extra])] + r[2 + extra:]
                if r[0].isdigit():            # source lines; the rows between them are their SASS instructions
                    body.append(r)
                i += 1
            f = lambda r, c: float(r[col[c]]) if r[col[c]] not in ("", "-") else 0.0
            ex, ss, th = "Instructions Executed", "Warp Stall Sampling (All Samples)", "Thread Instructions Executed"
            tot, tots = sum(f(r, ex) for r in body) or 1, sum(f(r, ss) for r in body) or 1
            print(f"== {want}: {tot:.0f} warp instructions, {tots:.0f} stall samples; columns: inst% stall% avg-threads line source")
            for r in body:
                if 100 * f(r, ex) / tot >= thr or 100 * f(r, ss) / tots >= thr:
                    at = f(r, th) / f(r, ex) if f(r, ex) else 0
                    print(f"{100 * f(r, ex) / tot:6.2f} {100 * f(r, ss) / tots:6.2f} {at:5.1f}  L{r[0]:>5s} {(text[int(r[0]) - 1] if text and int(r[0]) <= len(text) else r[1]).strip()[:130]}")
            return
        i += 1


if __name__ == "__main__":
    main()
