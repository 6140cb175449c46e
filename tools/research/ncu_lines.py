"""Developer aid: per-source-line instruction and stall-sample shares from `ncu -i X.ncu-rep --page source --csv
--print-source cuda,sass` (each SASS address counted once, under the first line that lists it)."""
import collections
import csv
import sys


def load(path):
    rows = list(csv.reader(open(path)))
    cur = line = None
    agg = collections.defaultdict(lambda: [0, 0, 0, 0, 0])
    src = {}
    seen = set()

    def I(x):
        try:
            return int(x)
        except ValueError:
            return 0
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name":
            continue
        if r[0] == "Line No":
            h = r
            ii, it, isamp, ib, il = (h.index(k) for k in ("Instructions Executed", "Thread Instructions Executed",
                                                          "# Samples", "stall_barrier", "stall_long_sb"))
            continue
        if r[0] != "":
            line = (cur, I(r[0]))
            src[line] = r[1]
        if len(r) > ii and r[2].startswith("0x") and r[2] not in seen:
            seen.add(r[2])
            a = agg[line]
            a[0] += I(r[ii]); a[1] += I(r[it]); a[2] += I(r[isamp]); a[3] += I(r[ib]); a[4] += I(r[il])
    return agg, src


if __name__ == "__main__":
    agg, src = load(sys.argv[1])
    ranges = [tuple(x.split(":")) for x in sys.argv[2:]]
    tot = sum(v[0] for v in agg.values())
    ts = sum(v[2] for v in agg.values())
    print("total inst", tot, "samples", ts)
    for name, a, b in ranges:
        s = [v for k, v in agg.items() if k[0] == "kernels.cu" and int(a) <= k[1] <= int(b)]
        print("%-16s inst %5.1f%% samples %5.1f%% barrier %5.1f%% long_sb %5.1f%%" % (
            name, sum(x[0] for x in s) / tot * 100, sum(x[2] for x in s) / ts * 100, sum(x[3] for x in s) / ts * 100,
            sum(x[4] for x in s) / ts * 100))
    print("top lines by samples")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][2])[:45]:
        print("%-12s:%5d inst %5.2f%% lanes %4.1f samp %5.2f%% bar %4.2f%% lsb %4.2f%%  %s" % (
            k[0][:12], k[1], v[0] / tot * 100, v[1] / max(v[0], 1), v[2] / ts * 100, v[3] / ts * 100, v[4] / ts * 100,
            src.get(k, "")[:80].strip()))
