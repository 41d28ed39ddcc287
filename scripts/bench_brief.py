"""Print the fields of a bench.py JSON line (stdin) that the exploration scripts look at."""
import json, sys
txt = sys.stdin.read().strip()
if not txt:
    print(sys.argv[1] if len(sys.argv) > 1 else "", "NO JSON LINE")
    sys.exit(0)
l = json.loads(txt.splitlines()[-1])
tag = sys.argv[1] if len(sys.argv) > 1 else ""


def brief(name, b):
    e = b.get("e2e", {})
    r = b.get("roofline", {})
    p = b.get("parity") or {}
    cb = b.get("cpu_baseline") or {}
    print(f"{tag} {name}: value {b.get('value', 0) / 1e6:.2f}M ms/step {b.get('ms_per_step', 0):.4f} e2e {e.get('value', 0) / 1e6:.2f}M"
          f" (1 caller {e.get('pinned_1_caller', 0) / 1e6:.2f}M, pageable {e.get('pageable_1_caller', 0) / 1e6:.2f}M)"
          f" frac {r.get('frac', 0):.2f} solo {r.get('algorithmic_frac_solo') or 0:.2f} dram {r.get('dram_frac') or 0:.3f}"
          f" E/q {r.get('dist_evals_per_query', 0):.0f} recall {b.get('recall_at_10')}"
          f" parity {[p.get(k) for k in ('rows_equal', 'dists_bit_equal', 'counts_equal')]} cpu {(cb.get('value') or 0) / 1e6:.3f}M"
          f" (1 thread {(cb.get('single_thread_value') or 0) / 1e3:.1f}k)")


brief("headline", l)
if "build" in l:
    print(tag, "build", {k: (round(v, 1) if isinstance(v, float) else v) for k, v in l["build"].items() if k != "semantics"})
for name, b in (l.get("sides") or {}).items():
    if isinstance(b, dict):
        brief(name, b)
        if "extended" in b and isinstance(b["extended"], dict):
            brief(name + ".extended", b["extended"])
            print(tag, name, "recall_vs_ef", b["extended"].get("recall_vs_ef"))
        if isinstance(b.get("extended_build"), dict):
            eb = b["extended_build"]
            brief(name + ".extended_build", eb)
            print(tag, name, "extended_build recall_vs_ef", eb.get("recall_vs_ef"), "build vec/s %.0f" % eb["build"]["vec_per_s"])
        if "build" in b:
            print(tag, name, "build vec/s %.0f kernel %.0f" % (b["build"]["vec_per_s"], b["build"]["kernel_vec_per_s"]))
fl = (l.get("parity") or {}).get("flavours")
if fl:
    print(tag, "flavours", {k: (v["queries_with_id_difference"], v["tie_class"], v["non_tie"]) for k, v in fl["search"].items()},
          {k: v["lists_different"] for k, v in fl["build"].items()}, "%.1fs" % fl["seconds"])
print(tag, "clocks", l.get("clocks"))
