"""Turn the ncu CSV logs of a bench.py run into the summaries kept under profiles/.

  python profiles/summarise_ncu.py launches <launches.csv> <out.txt> "<header comment>"
  python profiles/summarise_ncu.py traffic  <traffic.csv>  <out.txt> <out.json> "<header comment>"

`launches`: kernel, launch count, total ns, share of the captured time (gpu__time_duration.sum).
`traffic`:  per launch duration + dram__bytes_read.sum / dram__bytes_write.sum, and the per-family averages
            bench.py reads for `roofline.traffic`.
"""
import csv
import io
import json
import sys
from collections import OrderedDict, defaultdict

FAMILIES = OrderedDict([
    ("conv_nhwc_wgrad_kernel", "conv_nhwc_wgrad (TMA-fed)"),
    ("conv_nhwc_kernel", "conv_nhwc (TMA-fed fwd / dgrad)"),
    ("conv_wgrad_igemm_kernel", "conv_wgrad_igemm"),
    ("conv_igemm_kernel", "conv_igemm (fwd / dgrad)"),
    ("skinny_gemm_kernel", "skinny_gemm (FC 262144<->256)"),
])


def rows(path):
    text = open(path, errors="replace").read().splitlines()
    start = next(i for i, l in enumerate(text) if l.startswith('"ID"'))
    return list(csv.DictReader(io.StringIO("\n".join(text[start:]))))


def to_float(v):
    return float(v.replace(",", ""))


def scale(unit, kind):
    u = unit.strip().lower()
    if kind == "time":
        return {"ns": 1.0, "nsecond": 1.0, "us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6, "s": 1e9,
                "second": 1e9}[u]
    return {"byte": 1.0, "bytes": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}[u]


def short(name):
    name = name.split("(")[0].replace("void ", "")
    return name[:70]


def launches(path, out, header):
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in rows(path):
        if r["Metric Name"] != "gpu__time_duration.sum":
            continue
        k = short(r["Kernel Name"])
        val = to_float(r["Metric Value"])
        if val != val:      # ncu writes nan for a launch it could not measure (the process was ending)
            continue
        tot[k] += val * scale(r["Metric Unit"], "time")
        cnt[k] += 1
    total = sum(tot.values())
    with open(out, "w") as f:
        f.write("# " + header + "\n")
        f.write("# kernel, launches, total ns, share of captured time (total %.3f ms over %d launches)\n"
                % (total / 1e6, sum(cnt.values())))
        for k in sorted(tot, key=tot.get, reverse=True):
            f.write("%-72s %6d %14d   %.3f\n" % (k, cnt[k], tot[k], tot[k] / total))


def traffic(path, out, out_json, header):
    per = OrderedDict()
    for r in rows(path):
        e = per.setdefault(r["ID"], {"kernel": short(r["Kernel Name"]), "grid": r.get("Grid Size", "")})
        m = r["Metric Name"]
        if m == "gpu__time_duration.sum":
            e["ns"] = to_float(r["Metric Value"]) * scale(r["Metric Unit"], "time")
        elif m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            e[m] = to_float(r["Metric Value"]) * scale(r["Metric Unit"], "bytes")
    fam = OrderedDict()
    with open(out, "w") as f:
        f.write("# " + header + "\n# id kernel grid duration_us dram_read_MB dram_write_MB\n")
        for i, e in per.items():
            rd, wr = e.get("dram__bytes_read.sum", 0.0), e.get("dram__bytes_write.sum", 0.0)
            f.write("%3s %-30s %-16s %8.1f %8.1f %8.1f\n" % (i, e["kernel"][:30], e["grid"], e["ns"] / 1e3,
                                                           rd / 1e6, wr / 1e6))
            for key, label in FAMILIES.items():
                if key in e["kernel"]:
                    if key == "skinny_gemm_kernel" and wr > 4e8:    # the same kernel with the Adam epilogue: it writes p, m, v
                        label = "skinny_gemm wgrad + Adam (FC 262144<->256)"
                    a = fam.setdefault(label, {"launches": 0, "rd": 0.0, "wr": 0.0, "ns": 0.0})
                    a["launches"] += 1
                    a["rd"] += rd
                    a["wr"] += wr
                    a["ns"] += e["ns"]
                    break
    js = OrderedDict()
    for label, a in fam.items():
        n = a["launches"]
        js[label] = {"launches": n, "dram_bytes_per_launch": (a["rd"] + a["wr"]) / n,
                     "read_bytes_per_launch": a["rd"] / n, "write_bytes_per_launch": a["wr"] / n,
                     "avg_duration_us_under_ncu": a["ns"] / n / 1e3}
    js["_source"] = "%s (ncu dram__bytes_read.sum + dram__bytes_write.sum; %s)" % (out, header)
    json.dump(js, open(out_json, "w"), indent=1)


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3], sys.argv[4])
    else:
        traffic(sys.argv[2], sys.argv[3], sys.argv[4], sys.argv[5])
