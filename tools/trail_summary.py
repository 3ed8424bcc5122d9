import sys, collections
for fn in sys.argv[1:]:
    rows = {}
    for line in open(fn):
        parts = line.split()
        cta = int(parts[2].rstrip(':'))
        rows[cta] = [int(x, 16) for x in parts[3:19]]
    print("==", fn, "ctas", len(rows))
    for cl in sorted(set(c // 8 for c in rows)):
        ws = [rows.get(cl * 8 + r) for r in range(8)]
        s = []
        for r, w in enumerate(ws):
            if w is None: s.append("--"); continue
            s.append("s%d.c%x/%02x b%d p[%s]%s" % (w[0] >> 8, w[0] & 255, w[3], w[1],
                     ",".join("%d:%d" % ((x >> 28), x & 0xffff) for x in w[8:12]), ",".join("%x" % x for x in w[12:16])))
        print("cluster %2d" % cl)
        for r, t in enumerate(s): print("   r%d %s" % (r, t))
