"""profiles/r2_counters.json from an ncu counter pass of one 10,000-stop pass (tools/pass_shard.py --once):

  ncu --metrics smsp__inst_executed.sum,smsp__inst_executed_pipe_alu.sum,...,dram__bytes_read.sum,dram__bytes_write.sum,\\
      gpu__time_duration.sum --clock-control none --csv --log-file counters.csv python tools/pass_shard.py --world 1 --once
  python tools/ncu_counters.py counters.csv 10000 > profiles/r2_counters.json

bench.py divides these per-launch instruction counts by its own live durations (roofline.achieved).  The LAST launch of
each kernel in the file is taken (the timed pass; earlier ones belong to the set-up)."""
import csv
import json
import sys


def main():
    path, stops = sys.argv[1], int(sys.argv[2])
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    ki, mi, vi, ii = (hdr.index(k) for k in ("Kernel Name", "Metric Name", "Metric Value", "ID"))
    launches = {}
    for r in rows[1:]:
        launches.setdefault(int(r[ii]), {"kernel": r[ki]})[r[mi]] = float(r[vi].replace(",", ""))
    last = {}
    for _, d in sorted(launches.items()):
        name = d["kernel"]
        for key in ("eval_inter_fused_kernel", "eval_inter_kernel", "eval_intra_warp_kernel", "eval_intra_kernel",
                    "matrix_kernel<0>", "state_routes_kernel", "state_route_tables_kernel", "position_unit_vectors_kernel"):
            if key in name:
                last[key] = d

    def pick(d):
        return {"warp_inst": d["smsp__inst_executed.sum"], "alu_warp_inst": d["smsp__inst_executed_pipe_alu.sum"],
                "fma_warp_inst": d["smsp__inst_executed_pipe_fma.sum"], "fp64_warp_inst": d["smsp__inst_executed_pipe_fp64.sum"],
                "lsu_warp_inst": d["smsp__inst_executed_pipe_lsu.sum"], "xu_warp_inst": d["smsp__inst_executed_pipe_xu.sum"],
                "dram_bytes": d["dram__bytes_read.sum"] + d["dram__bytes_write.sum"],
                "ncu_duration_us": d["gpu__time_duration.sum"] / 1e3, "kernel": d["kernel"][:90]}

    inter = pick(last.get("eval_inter_fused_kernel") or last["eval_inter_kernel"])
    intra = pick(last.get("eval_intra_warp_kernel") or last["eval_intra_kernel"])
    ev = {k: inter[k] + intra[k] for k in inter if k not in ("kernel", "ncu_duration_us")}
    ev["parts"] = {"tile_kernel": inter, "same_route_kernel": intra}
    out = {f"pass_{stops}": {"eval_moves": ev,
                             "state": {k: pick(last[k]) for k in ("position_unit_vectors_kernel", "state_routes_kernel",
                                                                  "state_route_tables_kernel") if k in last}},
           "source": path, "note": "per launch, one GPU, whole instance (a 1/N shard executes 1/N of the tile kernel's "
                                   "instructions); counts do not depend on clocks"}
    if "matrix_kernel<0>" in last:
        out[f"pass_{stops}"]["matrix_kernel"] = pick(last["matrix_kernel<0>"])
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
