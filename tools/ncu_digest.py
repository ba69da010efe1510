#!/usr/bin/env python
"""Turns `ncu -i report.ncu-rep --page raw --csv` output into a small JSON digest per kernel launch (what profiles/ keeps):
duration, DRAM bytes read/written and the rate against the measured HBM peak, global-load sectors per request (sector
efficiency of the record / interval loads: 1 = fully coalesced 32-byte use, 32 = one sector per lane), L1/L2 hit rates,
issue-slot utilisation, active threads per instruction, registers.

    python tools/ncu_digest.py raw.csv out.json [peak_GBps]
"""
import csv
import json
import re
import sys


def main(src, dst, peak=6536.4):
    rows = list(csv.reader(open(src)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}

    def val(r, name, scale=None):
        i = col.get(name)
        if i is None or i >= len(r) or r[i] in ("", "n/a"):
            return None
        v = float(r[i].replace(",", ""))
        u = units[i]
        if scale == "bytes":
            v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(u, 1)
        if scale == "ms":
            v *= {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1, "msecond": 1, "s": 1e3, "second": 1e3}.get(u, 1)
        return v
    out = []
    for r in rows[2:]:
        if len(r) < len(hdr) // 2:
            continue
        name = re.sub(r"\(.*", "", r[col["Kernel Name"]]).replace("void ", "").replace("irf::", "")
        ms = val(r, "gpu__time_duration.sum", "ms")
        rd, wr = val(r, "dram__bytes_read.sum", "bytes"), val(r, "dram__bytes_write.sum", "bytes")
        sec, req = val(r, "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum"), val(r, "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum")
        d = {"kernel": name, "grid": r[col["Grid Size"]] if "Grid Size" in col else None, "block": r[col["Block Size"]] if "Block Size" in col else None,
             "duration_ms": None if ms is None else round(ms, 4), "dram_read_GB": None if rd is None else round(rd / 1e9, 3), "dram_write_GB": None if wr is None else round(wr / 1e9, 3)}
        if ms and rd is not None and wr is not None:
            d["dram_GBps"] = round((rd + wr) / 1e9 / (ms / 1e3), 1)
            d["dram_frac_of_measured_peak"] = round(d["dram_GBps"] / peak, 4)
        d["global_load_sectors_per_request"] = None if not req else round(sec / req, 2)
        d["l1_hit_pct"] = val(r, "l1tex__t_sector_hit_rate.pct")
        d["l2_hit_pct"] = val(r, "lts__t_sector_hit_rate.pct")
        d["issue_active_pct"] = val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active")
        d["threads_per_instruction"] = val(r, "smsp__thread_inst_executed_per_inst_executed.ratio")
        d["warp_instructions"] = val(r, "smsp__inst_executed.sum")
        d["warps_active_pct_of_peak"] = val(r, "sm__warps_active.avg.pct_of_peak_sustained_active")
        d["registers"] = val(r, "launch__registers_per_thread")
        out.append(d)
    json.dump({"source": src, "hbm_peak_GBps": peak, "launches": out}, open(dst, "w"), indent=1)
    for d in out:
        print(d["kernel"][:30].ljust(30), d["duration_ms"], "ms", d.get("dram_GBps"), "GB/s", d.get("dram_frac_of_measured_peak"), "sect/req", d["global_load_sectors_per_request"], "thr/inst", d["threads_per_instruction"])


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], float(sys.argv[3]) if len(sys.argv) > 3 else 6536.4)
