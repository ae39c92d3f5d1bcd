"""Hot spots of an `ncu --page source --csv` dump (SASS view): instruction mix by opcode and the most
sampled instructions with their dominant stall reason.  usage: ncu_sass_hot.py dump.csv [top]"""
import csv
import sys
from collections import Counter


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    rows = list(csv.reader(open(path)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    ix = {h: i for i, h in enumerate(hdr)}
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
    tot_s = sum(int(r[ix["# Samples"]]) for r in body)
    tot_i = sum(int(r[ix["Instructions Executed"]]) for r in body)
    print("# %d SASS instructions, %d warp-instructions executed, %d samples" % (len(body), tot_i, tot_s))
    by_op_i, by_op_s = Counter(), Counter()
    for r in body:
        op = r[ix["Source"]].split()
        op = op[1] if op and op[0].startswith("@") else (op[0] if op else "?")
        op = op.split(".")[0]
        by_op_i[op] += int(r[ix["Instructions Executed"]])
        by_op_s[op] += int(r[ix["# Samples"]])
    print("opcode,executed_share,sample_share")
    for op, n in by_op_i.most_common(22):
        print("%s,%.3f,%.3f" % (op, n / tot_i, by_op_s[op] / max(tot_s, 1)))
    tot_st = Counter()
    for r in body:
        for s in stalls:
            tot_st[s] += int(r[ix[s]])
    print("stall totals:", ", ".join("%s %.2f" % (k[6:], v / max(tot_s, 1)) for k, v in tot_st.most_common(8)))
    print("idx,samples_share,executed,sass,top_stall")
    order = sorted(range(len(body)), key=lambda i: -int(body[i][ix["# Samples"]]))[:top]
    for i in sorted(order):
        r = body[i]
        st = max(stalls, key=lambda s: int(r[ix[s]]))
        print("%d,%.3f,%s,%s,%s" % (i, int(r[ix["# Samples"]]) / max(tot_s, 1), r[ix["Instructions Executed"]],
                                    r[ix["Source"]].strip()[:90], st[6:]))


if __name__ == "__main__":
    main()
