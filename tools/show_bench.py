"""Pretty-prints the last JSON line of a bench log."""
import json
import sys

line = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
for k, v in line.items():
    if k in ("kernels", "pipeline") and isinstance(v, list):
        print(k)
        for r in v:
            print("   ", r)
    else:
        print(k, v)
