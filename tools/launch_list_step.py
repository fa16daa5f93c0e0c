"""Sum of one time step out of an ncu launch list (--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum
--csv): the launches between the last two k_tridiag launches.  Prints per-kernel time / DRAM bytes / rate and the totals
that go into profiles/traffic.json.   usage: python tools/launch_list_step.py profiles/<list>.csv"""
import collections
import csv
import re
import sys


def load(path):
    rows = csv.DictReader([l for l in open(path) if l.startswith('"')])
    K, order = {}, []
    scale = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0, "byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for x in rows:
        i = int(x["ID"])
        if i not in K:
            K[i] = {"name": x["Kernel Name"]}
            order.append(i)
        K[i][x["Metric Name"]] = float(x["Metric Value"].replace(",", "")) * scale.get(x["Metric Unit"], 1)
    return [K[i] for i in order]


def last_step(ks):
    tri = [n for n, k in enumerate(ks) if "tridiag" in k["name"]]
    return ks[tri[-2] + 1:tri[-1] + 1]


if __name__ == "__main__":
    step = last_step(load(sys.argv[1]))
    tot_t = sum(k["gpu__time_duration.sum"] for k in step)
    tot_b = 0.0
    for k in step:
        t = k["gpu__time_duration.sum"]
        b = k.get("dram__bytes_read.sum", 0) + k.get("dram__bytes_write.sum", 0)
        tot_b += b
        name = re.sub(r"\(.*", "", k["name"]).replace("void ", "")
        print("%-48s %9.1f us %7.3f GB %5.2f TB/s %5.1f%%" % (name[:48], t * 1e6, b / 1e9, b / t / 1e12 if t else 0, 100 * t / tot_t))
    print("step: %d launches, %.3f ms serialised, %.3f GB DRAM (%d bytes)" % (len(step), tot_t * 1e3, tot_b / 1e9, int(tot_b)))
