"""Hottest SASS instructions of each kernel in an `ncu --page source --csv --print-source sass` export."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name = rows[i][1]; hdr = rows[i + 1]; j = i + 2; body = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            if len(rows[j]) == len(hdr): body.append(rows[j])
            j += 1
        ci = {h: k for k, h in enumerate(hdr)}
        tot = sum(int(r[ci["# Samples"]] or 0) for r in body)
        print("====", name[:90], "samples", tot, "instructions", len(body))
        stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
        agg = {s: sum(int(r[ci[s]] or 0) for r in body) for s in stalls}
        print("  ", {k: round(v / max(tot, 1), 3) for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
        order = sorted(range(len(body)), key=lambda k: -int(body[k][ci["# Samples"]] or 0))[:top]
        for k in sorted(order):
            r = body[k]
            s = int(r[ci["# Samples"]] or 0)
            why = sorted(((int(r[ci[x]] or 0), x[6:]) for x in stalls), reverse=True)[:2]
            print("  %5d %5.2f%%  #%4d %-70s %s" % (s, 100.0 * s / max(tot, 1), k, r[ci["Source"]][:70], " ".join("%s=%d" % (b, a) for a, b in why if a)))
        i = j
    else:
        i += 1
