"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel into a small JSON
for profiles/.   python scripts/launch_summary.py gpurun_out/launches.csv profiles/launches_<name>_rNN.json "<command>" """
import collections, csv, json, re, sys


def main(src, dst, command):
    rows = [r for r in csv.reader(open(src)) if len(r) > 10]
    h = rows[0]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        v = float(r[vi].replace(",", ""))
        v = {"ns": v / 1e3, "nsecond": v / 1e3, "us": v, "usecond": v, "ms": v * 1e3, "msecond": v * 1e3}[r[ui]]
        k = re.sub(r"\(.*", "", r[ki]).replace("<unnamed>::", "").replace("void ", "")
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    ks = [{"kernel": k, "launches": v[0], "us": round(v[1], 1), "share": round(v[1] / tot, 4)}
          for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])]
    conv = sum(k["us"] for k in ks if k["kernel"].startswith(("conv_halo", "conv_tc")))
    json.dump({"command": command,
               "note": "ncu --metrics gpu__time_duration.sum --clock-control none; per-launch times are cold-cache "
                       "and serialised: compare shares, not absolutes",
               "launches": len(rows) - 1, "total_us": round(tot, 1), "tcgen05_conv_share": round(conv / tot, 4),
               "kernels": ks[:40]}, open(dst, "w"), indent=1)
    for k in ks[:16]:
        print(f"{k['us']:10.1f} us {100 * k['share']:5.1f}% n={k['launches']:5d} {k['kernel'][:80]}")
    print("total", tot, "conv share", conv / tot)


if __name__ == "__main__":
    main(*sys.argv[1:4])
