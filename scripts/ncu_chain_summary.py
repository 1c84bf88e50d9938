"""profiles/<tag>_chain_summary.md from the reports scripts/ncu_chain.sh leaves in gpurun_out/."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ncu_summary as ns

tag = sys.argv[1] if len(sys.argv) > 1 else "r1s"
out = [f"# ncu summary {tag}: matching -> residual chain", "",
       "`ncu --set full --clock-control none --import-source on`, one launch per kernel, over "
       "`python scripts/bench_trajectory.py --poses 500` (2000 poses; the nn_match launch is one "
       "60 000-pair slice of the growing-window schedule, the other three the resident chain of "
       "the widest round).", ""]
for name in ("nn", "res"):
    rep = os.path.join(ns.ROOT, "gpurun_out", f"{tag}_chain_{name}.ncu-rep")
    if os.path.exists(rep):
        table, _ = ns.full_table(rep)
        out += [table, ""]
open(os.path.join(ns.PROF, f"{tag}_chain_summary.md"), "w").write("\n".join(out))
print("\n".join(out))
