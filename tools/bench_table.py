"""Markdown table of the bench lines under profiles/ with a given tag:  python tools/bench_table.py r2n [r2r ...]
(later tags override earlier ones for the same workload name)."""
import glob
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    rows = {}
    for tag in sys.argv[1:]:
        for path in sorted(glob.glob(os.path.join(ROOT, "profiles", tag + "_bench_*.json"))):
            name = os.path.basename(path)[len(tag) + 7:-5]
            try:
                d = json.load(open(path))
            except Exception:
                continue
            if d.get("impl") == "reference":
                continue
            rows[name] = (tag, d)
    print("| line (profiles/<tag>_bench_<name>.json) | PCG it/s | e2e it/s | dominant kernel | ms / launch | algorithmic GB/s (of measured / of 8 TB/s) | iteration frac | parity (cost / unknowns) | CPU port it/s (cores) | clocks |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    for name, (tag, d) in rows.items():
        r, p, c = d.get("roofline") or {}, d.get("parity") or {}, d.get("cpu_baseline") or {}
        it = (r.get("iteration") or {}).get("frac")
        par = "—" if not p else "%s (%.1e / %.1e)" % ("ok" if p.get("ok") else "**FAIL**", p.get("rel_err_cost", 0), p.get("max_rel_err_unknowns", 0))
        cpu = "—" if not c else "%.3g (%d)" % (c["value"], c["cores"])
        print("| %s `%s` | %.1f | %.1f | %s | %s | %s | %s | %s | %s | %s |" % (
            name, tag, d["value"], d["e2e"]["value"], r.get("kernel", "—"),
            "%.4f" % r["ms_per_launch"] if r.get("ms_per_launch") else "—",
            "%.0f (%.0f %% / %.0f %%)" % (r["achieved"], 100 * r["frac"], 100 * r["frac_of_nominal_8TBs"]) if r.get("achieved") else "—",
            "%.2f" % it if it else "—", par, cpu,
            "%s MHz%s" % (int(d["clocks"]["sm_mhz"]) if d["clocks"].get("sm_mhz") else "?", " " + ",".join(d["clocks"]["reasons"]) if d["clocks"].get("reasons") else "")))


if __name__ == "__main__":
    main()
