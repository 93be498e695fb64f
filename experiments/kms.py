import json, sys
d = json.loads(open(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/bench_new.log").read().strip().splitlines()[-1])
print({k: round(v["ms"], 3) for k, v in d["kernel_ms"].items()})
print("value", round(d["value"]), "ms", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"]), "roof", round(d["roofline"]["achieved"]), d["clocks"])
