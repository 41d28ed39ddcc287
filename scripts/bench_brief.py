"""Print the fields of a bench.py JSON line (stdin) that the exploration scripts look at."""
import json, sys
l = json.loads(sys.stdin.read())
e = l["e2e"]
print(sys.argv[1] if len(sys.argv) > 1 else "", "value %.1fM" % (l["value"] / 1e6), "ms/step %.4f" % l["ms_per_step"],
      "issue %.4f" % l.get("host_issue_ms_per_step", -1), "e2e %.1fM" % (e["value"] / 1e6), [round(x / 1e6, 1) for x in (e.get("passes") or [])],
      "single %.1fM" % (e.get("single_caller_value", 0) / 1e6), "samples", l["clocks"]["samples"],
      "build %.2fM vec/s" % (l["build"]["vec_per_s"] / 1e6))
