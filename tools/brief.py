"""Reads bench.py JSON lines on stdin and prints the key numbers in one compact line each (for parameter sweeps)."""
import json
import sys
for line in sys.stdin:
    line = line.strip()
    if not line.startswith("{"):
        continue
    d = json.loads(line)
    k = d.get("kernel_ms_per_step", {})
    nb = d.get("nblist", {})
    print(f"{sys.argv[1] if len(sys.argv) > 1 else ''} ms/step={d['ms_per_step']:.4f} ns/day={d['value']:.1f} e2e_ms={d['e2e']['ms_per_step']:.3f} | " +
          " ".join(f"{n}={v*1e3:.0f}us" for n, v in k.items() if v > 0) +
          f" | tiles={nb.get('tiles')} sparse={nb.get('sparse_tiles')} units={nb.get('units')} rebuilds={nb.get('rebuilds')}/{nb.get('steps')} eff={nb.get('tile_efficiency', 0):.3f}"
          f" roof={d['roofline']['frac']:.3f}")
