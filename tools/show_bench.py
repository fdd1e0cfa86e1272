"""Print the interesting keys of a bench.py JSON line."""
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
r = d.get("roofline", {})
print("value %.0f  e2e %.0f (%.3f of link ceiling %.0f)  ms/step %.3f  e2e ms %.3f  clocks %s" % (
    d["value"], d["e2e"]["value"], d["e2e"].get("frac_of_link_ceiling", 0), d["e2e"].get("link_ceiling_scans_s", 0),
    d["ms_per_step"], d["e2e"]["ms_per_step"], (d.get("clocks") or {}).get("sm_mhz")))
if r:
    print("roofline %.1f TF/s frac %.3f  conv layers ms %s  segnet %.3f  glue %.3f" % (
        r["achieved"], r["frac"], [round(x, 3) for x in r["conv_layers_ms"]], r["segnet_ms"], r["segnet_minus_convs_ms"]))
for k in ("latency_ms_p50_batch1", "latency_ms_p50_batch1_graph", "latency_ms_p50_batch1_host_to_host", "latency_ms_p50_batch1_graph_variable_n"):
    if k in d:
        print(k, round(d[k], 4))
if "stage_ms" in d:
    print("stage_ms", {k: round(v, 4) for k, v in d["stage_ms"].items() if isinstance(v, float)})
if "hbm_stages" in d:
    print("hbm frac", {k: round(v["frac_of_measured_hbm"], 3) for k, v in d["hbm_stages"].items() if isinstance(v, dict)})
for k in ("precision_split", "cpu_baseline", "parity"):
    if k in d:
        print(k, json.dumps(d[k])[:400])
if "torch_cuda_baseline" in d:
    print("torch_cuda_baseline", json.dumps(d["torch_cuda_baseline"])[:900])
