#!/usr/bin/env python
"""Print the instructions / source lines with the most warp-stall samples from `ncu --page source --csv`."""
import csv
import sys


def main(path, frac=0.008):
    rows = list(csv.reader(open(path)))
    hdr = None
    block = []
    blocks = []
    for r in rows:
        if "# Samples" in r and "Source" in r:
            if hdr is not None:
                blocks.append((hdr, block))
            hdr, block = r, []
        elif hdr is not None and len(r) == len(hdr):
            block.append(r)
    if hdr is not None:
        blocks.append((hdr, block))
    for hdr, data in blocks:
        ix = {k: i for i, k in enumerate(hdr)}

        def num(r, k):
            try:
                return int(float(r[ix[k]] or 0))
            except (ValueError, KeyError):
                return 0
        tot = sum(num(r, "# Samples") for r in data) or 1
        print(f"--- {len(data)} rows, {tot} samples")
        for r in data:
            s = num(r, "# Samples")
            if s > tot * frac:
                print(f"{s / tot * 100:5.1f}%  long_sb={num(r, 'stall_long_sb'):6d} lg={num(r, 'stall_lg'):5d} "
                      f"drain={num(r, 'stall_drain'):5d} wait={num(r, 'stall_wait'):5d} short={num(r, 'stall_short_sb'):5d} "
                      f"mio={num(r, 'stall_mio'):5d}  {r[ix['Source']][:110]}")


if __name__ == "__main__":
    main(sys.argv[1], float(sys.argv[2]) if len(sys.argv) > 2 else 0.008)
