#!/usr/bin/env python
"""Hot-loop breakdown from `ncu --page source --csv --print-source sass`: exec levels, opcode mix, stall reasons."""
import collections
import csv
import sys


def main(path, show=0):
    rows = list(csv.reader(open(path)))
    hdr = rows[1]
    blk = []
    for r in rows[2:]:
        if len(r) < 10 or r[0] == 'Kernel Name':
            break
        blk.append(r)
    ie, isrc, isamp = hdr.index('Instructions Executed'), hdr.index('Source'), hdr.index('# Samples')
    tot = sum(float(r[ie]) for r in blk)
    mx = max(float(r[ie]) for r in blk)
    print("instructions", len(blk), "executed", "%.4e" % tot, "max level", "%.4e" % mx)
    lv = collections.Counter()
    for r in blk:
        lv[round(float(r[ie]) / mx, 2)] += 1
    print("exec levels (fraction of max: #instr):", sorted(lv.items(), reverse=True)[:12])
    ops, opsn = collections.Counter(), collections.Counter()
    for r in blk:
        t = r[isrc].split()
        op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
        ops[op] += float(r[ie]); opsn[op] += 1
    for k, v in ops.most_common(22):
        print("  %-10s exec %.3e  %5.1f%%" % (k, v, 100 * v / tot))
    # static opcode histogram of the loop bodies: instructions grouped by how often they execute (fraction of the maximum)
    byl = collections.defaultdict(collections.Counter)
    for r in blk:
        t = r[isrc].split()
        op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
        byl[round(float(r[ie]) / mx, 2)][op] += 1
    for level, cnt in sorted(byl.items(), reverse=True):
        n = sum(cnt.values())
        if n < 20 or level < 0.05:
            continue
        keys = ('DFMA', 'DMUL', 'DADD', 'LDS', 'SHFL', 'MUFU', 'F2F', 'FSEL', 'IMAD', 'BRA', 'UBLKCP', 'SYNCS', 'REDG')
        print("  body executed %.2f x max: %3d instructions: " % (level, n) + ", ".join("%s %d" % (k, cnt[k]) for k in keys if cnt[k])
              + ", FP64 %d" % (cnt['DFMA'] + cnt['DMUL'] + cnt['DADD']))
    st = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
    acc = collections.Counter()
    for r in blk:
        for h in st:
            v = r[hdr.index(h)]
            if v:
                acc[h] += float(v)
    s = sum(acc.values())
    print("stalls:", ", ".join("%s %.1f%%" % (k.replace('stall_', ''), 100 * v / s) for k, v in acc.most_common(8)))
    if show:
        ts = sum(float(r[isamp] or 0) for r in blk)
        for r in blk:
            if float(r[ie]) > 0.25 * mx:
                print("%5.2f%% %4.2f  %s" % (100 * float(r[isamp] or 0) / ts, float(r[ie]) / mx, r[isrc][:90]))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0)
