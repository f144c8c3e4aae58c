"""Writes profiles/r02_ncu_metrics.json, the counters bench.py's `roofline` block quotes with their source: for each
capture given, the kernel's DRAM bytes, L2 bytes, issue-slot utilisation, active lanes per instruction and duration.

  python scripts/ncu_metrics.py k_wavefront=gpurun_out/prof.ncu-rep:"1080p x 8 spp" [k_wavefront.traffic=gpurun_out/traffic_1024.csv:"1080p x 1024 spp"] ...

A `.ncu-rep` is read through `ncu -i --page raw --csv`; a `.csv` is the --metrics log of scripts/gpu_traffic.sh (its DRAM
bytes and duration override the rep's: that pass profiles the very launch the bench times)."""
import csv, io, json, os, subprocess, sys
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
OUT = Path(os.environ.get("NCU_METRICS_OUT") or REPO / "profiles" / "r02_ncu_metrics.json")  # the GPU box writes into gpurun_out/


def from_rep(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, r = rows[0], rows[1], rows[2]
    idx = {h: i for i, h in enumerate(hdr)}

    def val(name, scale_unit=True):
        if name not in idx:
            return None
        v = float(r[idx[name]].replace(",", ""))
        u = units[idx[name]]
        if scale_unit:
            v *= {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0,
                  "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9, "second": 1.0}.get(u, 1.0)
        return v

    out = {"kernel": r[idx["Kernel Name"]].strip(), "duration_s": val("gpu__time_duration.sum"),
           "dram_bytes": (val("dram__bytes_read.sum") or 0) + (val("dram__bytes_write.sum") or 0),
           "lts_bytes": val("lts__t_bytes.sum"), "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active"),
           "lanes_per_inst": val("smsp__thread_inst_executed_per_inst_executed.ratio"), "inst_executed": val("smsp__inst_executed.sum"),
           "warps_active_pct": val("sm__warps_active.avg.pct_of_peak_sustained_active"), "registers": val("launch__registers_per_thread"),
           "l1_hit_pct": val("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": val("lts__t_sector_hit_rate.pct"),
           "dram_throughput_pct": val("dram__throughput.avg.pct_of_peak_sustained_elapsed")}
    if out["lts_bytes"] is None:  # --set full on this ncu carries sectors / the L2->crossbar rate, not lts__t_bytes
        if val("lts__t_sectors.sum") is not None:
            out["lts_bytes"], out["lts_bytes_from"] = 32.0 * val("lts__t_sectors.sum"), "lts__t_sectors.sum x 32 B"
        elif val("derived__lts__lts2xbar_bytes.sum.per_second", scale_unit=False) is not None and out["duration_s"]:
            u = units[idx["derived__lts__lts2xbar_bytes.sum.per_second"]]
            rate = val("derived__lts__lts2xbar_bytes.sum.per_second", scale_unit=False) * {"Gbyte/s": 1e9, "Tbyte/s": 1e12, "Mbyte/s": 1e6, "Kbyte/s": 1e3}.get(u, 1.0)
            out["lts_bytes"], out["lts_bytes_from"] = rate * out["duration_s"], "derived__lts__lts2xbar_bytes.sum.per_second x duration (L2 -> SM only)"
    clk = val("smsp__cycles_elapsed.avg.per_second", scale_unit=False)
    if clk is not None:
        u = units[idx["smsp__cycles_elapsed.avg.per_second"]]
        out["sm_mhz"] = clk * {"Ghz": 1e3, "Mhz": 1.0, "Khz": 1e-3, "hz": 1e-6, "GHz": 1e3, "MHz": 1.0}.get(u, 1e-6)
    return out


def from_csv(path):
    rows = [r for r in csv.reader(open(path)) if r and not r[0].startswith("==")]
    hdr = rows[0]
    idx = {h: i for i, h in enumerate(hdr)}
    out = {}
    for r in rows[1:]:
        name, unit, v = r[idx["Metric Name"]], r[idx["Metric Unit"]], float(r[idx["Metric Value"]].replace(",", ""))
        v *= {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "usecond": 1e-6, "msecond": 1e-3, "nsecond": 1e-9}.get(unit, 1.0)
        out[name] = out.get(name, 0.0) + v
    return {"dram_bytes": out.get("dram__bytes_read.sum", 0.0) + out.get("dram__bytes_write.sum", 0.0),
            "duration_s": out.get("gpu__time_duration.sum")}


def main():
    try:
        result = json.loads(OUT.read_text())
    except Exception:
        result = {}
    for arg in sys.argv[1:]:
        key, rest = arg.split("=", 1)
        path, _, note = rest.partition(":")
        kernel, _, sub = key.partition(".")
        entry = result.setdefault(kernel, {})
        data = from_csv(path) if path.endswith(".csv") else from_rep(path)
        if sub:
            entry[sub] = dict(data, capture=note, file=path)
            if sub == "traffic":
                entry["dram_bytes_full_launch"] = data["dram_bytes"]
        else:
            entry.update(data)
            entry["capture"] = note
            entry["file"] = path
    OUT.write_text(json.dumps(result, indent=1) + "\n")
    print(json.dumps(result, indent=1))


if __name__ == "__main__":
    main()
