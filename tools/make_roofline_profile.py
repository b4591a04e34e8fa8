"""Turn one ncu capture of the two record kernels into the files bench.py reads for its roofline object.

  on the GPU box (after `python tools/prof_target.py 65536 1` has exited 0 without ncu):
      ncu --set full --clock-control none --import-source on -k regex:gcm_record -c 2 \
          -o gpurun_out/rNN_record python tools/prof_target.py 65536 1
  here:
      python tools/make_roofline_profile.py gpurun_out/rNN_record.ncu-rep rNN

Writes profiles/record_kernel_pipes.json (instruction counts per 16-byte block and lane from the source page,
pipe utilisation and DRAM bytes from the raw page, for the fused seal <4,2> and open <4,3> kernels), a readable
summary profiles/<tag>_ncu_record.txt and the opcode histograms profiles/<tag>_{seal,open}_opcode_hist.txt.
The JSON carries `source_hash` = bench.record_kernel_source_hash(): bench.py ignores the file when the kernel
sources have changed since the capture, so a kernel edit cannot leave a stale roofline in the bench line."""
import collections
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

ALU = {"LOP3", "PRMT", "SHF", "IADD3", "ISETP", "VIADD", "SEL", "LEA", "IADD", "VIMNMX", "PLOP3", "P2R", "R2P", "BMSK",
       "SGXT", "FLO", "POPC", "VIADDMNMX", "IABS", "MOV", "CS2R", "VABSDIFF", "LOP", "BREV"}
RECORDS, SLOTS = 1048576, 80          # the capture's launch: 65536 sessions x 16 records x 1200 B, 78 slots -> 80 lane slots (G = 4)
WARP_SLOTS = RECORDS * SLOTS / 32.0

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
           "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
           "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
           "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
           "sass__inst_executed_local_loads", "sass__inst_executed_local_stores", "sm__cycles_elapsed.max"]


def ncu(rep, *args):
    return subprocess.run(["ncu", "-i", rep] + list(args), capture_output=True, text=True, check=True).stdout


def raw_page(rep):
    rows = list(csv.reader(io.StringIO(ncu(rep, "--page", "raw", "--csv"))))
    hdr, units = rows[0], rows[1]
    out = []
    for r in rows[2:]:
        d = {h: (v, u) for h, v, u in zip(hdr, r, units)}
        out.append(d)
    return out


def to_bytes(v, unit):
    x = float(v.replace(",", ""))
    return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit, 1)


def main():
    rep, tag = sys.argv[1], sys.argv[2]
    raw = raw_page(rep)
    out = {"source_hash": bench.record_kernel_source_hash(),
           "source": f"profiles/{tag}_ncu_record.txt + profiles/{tag}_{{seal,open}}_opcode_hist.txt, written by tools/make_roofline_profile.py from one "
                     "`ncu --set full --clock-control none --import-source on` capture of tools/prof_target.py 65536 1 (the bench's launch size)",
           "alu_warp_inst_per_clk_per_sm": 2.0,
           "alu_peak_source": "2 warp instructions per clock per SM on the integer ALU pipe (LOP3 / PRMT / SHF / IADD3), measured on this pool's "
                              "B200 with tools/alu_peak.cu (profiles/r01_pipe_peaks_measured.json) x SM clock sampled during the bench x 148 SMs"}
    text = [f"ncu --set full --clock-control none --import-source on -k regex:gcm_record -c 2, python tools/prof_target.py 65536 1",
            f"(the bench's own launch size: {RECORDS} records x 1200 B payload per launch; source hash {out['source_hash']})"]
    for d in raw:
        name = d["Kernel Name"][0]
        if "gcm_record_kernel" not in name:
            continue
        which = "seal" if ", 2>" in name else "open"
        text.append("-----")
        text.append(f"  Kernel Name = {name}")
        m = {}
        for k in METRICS:
            if k in d:
                v, u = d[k]
                m[k] = (v, u)
                text.append(f"  {k} [{u}] = {v}")
        f = lambda k: float(m[k][0].replace(",", ""))
        dram = to_bytes(*m["dram__bytes_read.sum"]) + to_bytes(*m["dram__bytes_write.sum"])
        out[which] = {"kernel": name.split("(")[0].strip(), "lane_slots_per_record": SLOTS,
                      "total_inst_per_lane_slot": f("smsp__inst_executed.sum") / WARP_SLOTS,
                      "dram_bytes_per_launch": dram, "dram_bytes_per_record": dram / RECORDS,
                      "ncu": {"alu_pipe_busy": f("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active") / 100,
                              "issue_slots_busy": f("smsp__issue_active.avg.pct_of_peak_sustained_active") / 100,
                              "fma_pipe_busy": f("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active") / 100,
                              "lsu_pipe_busy": f("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active") / 100,
                              "dram_pct_of_peak": f("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed") / 100,
                              "warps_active_pct": f("sm__warps_active.avg.pct_of_peak_sustained_active") / 100,
                              "local_loads_per_launch": f("sass__inst_executed_local_loads"),
                              "registers_per_thread": f("launch__registers_per_thread"),
                              "duration_ms_under_ncu": f("gpu__time_duration.sum") * {"us": 1e-3, "ms": 1, "ns": 1e-6, "s": 1e3}.get(m["gpu__time_duration.sum"][1], 1)}}
    # instruction mix from the source page, one kernel at a time (launch order in the capture: seal, then open)
    for which in ("seal", "open"):
        if which not in out:
            continue
        k = out[which]
        src = ncu(rep, "--page", "source", "--csv", "--kernel-name", "regex:gcm_record", "--launch-skip", "0" if which == "seal" else "1",
                  "--launch-count", "1")
        assert ("(int)2>" if which == "seal" else "(int)3>") in src.split("\n", 1)[0], src.split("\n", 1)[0]
        rows = list(csv.reader(io.StringIO(src)))
        hi = next(i for i, r in enumerate(rows) if "Source" in r and "Instructions Executed" in r)
        hdr = rows[hi]
        si, ei, ss = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
        stall_cols = [(j, c) for j, c in enumerate(hdr) if c.startswith("stall_") and "Not Issued" not in c]
        agg, stalls, tot, stot = collections.Counter(), collections.Counter(), 0, 0
        for r in rows[hi + 1:]:
            try:
                n = int(r[ei]); smp = int(r[ss])
            except (ValueError, IndexError):
                continue
            op = r[si].strip().split()
            if not op:
                continue
            o = op[1] if op[0].startswith("@") else op[0]
            o = ".".join(o.split(".")[:2]) if o.startswith(("LD", "ST", "IMAD", "SHFL")) else o.split(".")[0]
            agg[o] += n; tot += n; stot += smp
            for j, c in stall_cols:
                try:
                    stalls[c] += int(r[j])
                except ValueError:
                    pass
        # the source page's counts come from its own replay passes and can be a whole multiple of the launch's
        # smsp__inst_executed (2x in the r02 capture): the MIX is taken from it, the total from the raw page
        scale = k["total_inst_per_lane_slot"] * WARP_SLOTS / tot
        agg = collections.Counter({o: v * scale for o, v in agg.items()})
        tot_src, tot = tot, tot * scale
        alu = sum(v for o, v in agg.items() if o in ALU)
        fma = sum(v for o, v in agg.items() if o.startswith("IMAD") or o in ("IMUL", "FFMA", "FMUL"))
        lds = sum(v for o, v in agg.items() if o.startswith("LDS"))
        k["alu_pipe_inst_per_lane_slot"] = alu / WARP_SLOTS
        k["fma_pipe_inst_per_lane_slot"] = fma / WARP_SLOTS
        k["lds_inst_per_lane_slot"] = lds / WARP_SLOTS
        k["source_page_count_scale"] = scale
        lines = [f"total warp instructions {tot:.0f} (smsp__inst_executed.sum; the source page counted {tot_src}, mix scaled by {scale:.3f}), "
                 f"per warp-slot {tot / WARP_SLOTS:.1f}; samples {stot}",
                 f"ALU pipe {alu / WARP_SLOTS:.1f}, FMA pipe {fma / WARP_SLOTS:.1f}, LDS {lds / WARP_SLOTS:.1f} per warp-slot"]
        for o, v in agg.most_common(30):
            lines.append(f"{o:14s} {v:14.0f} {v / tot * 100:5.1f}%  per-slot {v / WARP_SLOTS:7.1f}")
        lines.append("stall samples: " + ", ".join(f"{c[6:]}={v / max(stot, 1) * 100:.1f}%" for c, v in stalls.most_common(9)))
        with open(os.path.join(ROOT, "profiles", f"{tag}_{which}_opcode_hist.txt"), "w") as fh:
            fh.write("\n".join(lines) + "\n")
    with open(os.path.join(ROOT, "profiles", f"{tag}_ncu_record.txt"), "w") as fh:
        fh.write("\n".join(text) + "\n")
    with open(os.path.join(ROOT, "profiles", "record_kernel_pipes.json"), "w") as fh:
        json.dump(out, fh, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
