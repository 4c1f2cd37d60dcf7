#!/usr/bin/env python
"""Store the first-losses trajectory of bench.py runs as the golden record (benchmarks/golden_losses.json).

    scripts/update_golden.py <bench json> [<bench json> ...]

Only for runs whose other checks passed; the key is (world, batch, fc1 precision)."""
import json, os, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
path = os.path.join(root, "benchmarks", "golden_losses.json")
try:
    golden = json.load(open(path))
except Exception:
    golden = {}
for f in sys.argv[1:]:
    line = [l for l in open(f).read().splitlines() if l.startswith("{")][-1]
    j = json.loads(line)
    g = j["checks"]["golden_losses"]
    if j["checks"]["param_update_norm"] > 0 and j["checks"]["replica_drift"] == 0.0:
        golden[g["key"]] = g["first_losses"]
        print("stored", g["key"], g["first_losses"][:3], "...")
json.dump(golden, open(path, "w"), indent=1, sort_keys=True)
