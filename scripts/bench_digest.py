"""Reads one bench.py JSON line on stdin and prints the numbers that matter, one per line."""
import json
import sys

d = json.loads(sys.stdin.read().strip().splitlines()[-1])
c = d["config"]
print("headline %s: %.3e %s, %.3f us/step, roofline frac %.3f (streams %s, pipeline %s)" % (
    c.get("workload", "?")[:40], d["value"], d["unit"], d["ms_per_step"] * 1e3, (d.get("roofline") or {}).get("frac", float("nan")),
    c.get("streams"), c.get("pipeline")))
for k in sorted(k for k in c if k == "serial_single_stream" or (k.startswith("overlapped_") and k.endswith("_streams"))):
    if k in c:
        print("  %s: %.3f us/step, frac %.3f" % (k, c[k]["ms_per_step"] * 1e3, c[k]["frac"]))
if d.get("e2e"):
    print("  e2e %.3e %s (h2d %d B/step)" % (d["e2e"]["value"], d["e2e"]["unit"], d["e2e"]["h2d_bytes_per_step"]))
if d.get("cpu_baseline"):
    b = d["cpu_baseline"]
    print("  cpu %.3e (%d cores, %s); 1 thread %.3e; C1 %s" % (b["value"], b["cores"], b.get("cpu_model"), b.get("one_thread", {}).get("value", float("nan")),
                                                            b.get("config1_T20_B32", {}).get("value")))
for name, e in (c.get("other_workloads") or {}).items():
    if "error" in e:
        print("  other %s: ERROR %s" % (name, e["error"]))
        continue
    print("  other %s: %.3e %s, %.4f ms/step%s%s" % (name, e["value"], e["unit"], e["ms_per_step"],
          ", frac %.3f" % e["roofline"]["frac"] if "roofline" in e else "",
          ", 3 streams: %.4f ms frac %.3f" % (e["overlapped_3_streams"]["ms_per_step"], e["overlapped_3_streams"]["frac"]) if "overlapped_3_streams" in e else ""))
    for k in ("breakdown_ms", "grad_exchange", "launches_per_step", "k8_check"):
        if k in e:
            print("     %s: %s" % (k, e[k]))
print("  clocks %s, gpu_launches %s, timing %s" % (d.get("clocks"), d.get("gpu_launches"), {k: v for k, v in c.get("timing", {}).items() if k != "method"}))
