"""one-line digest of a bench.py JSON line: python tools/show_bench.py file.json"""
import json
import sys

j = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
c = j["config"]
print("value", round(j["value"]), "ms", round(j["ms_per_step"], 1), "vcycles", c.get("vcycles_per_solve"), "bottom_iters",
      c.get("bottom_iters_per_solve"), "final_rel_res", c.get("final_rel_residual"),
      {k: round(v, 1) for k, v in c.get("phase_ms_per_solve", {}).items()}, "roofline_frac", round(j["roofline"]["frac"], 3),
      "hbm_frac", round(c.get("vcycle_hbm_frac_per_gpu", 0), 3), "launches", j.get("gpu_launches"),
      "e2e", round(j["e2e"]["value"]) if j.get("e2e") else None)
