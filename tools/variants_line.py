"""stdin: bench.py output; argv[1]: library path -> one line 'lib value {stage: ms}'."""
import json
import sys

d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(sys.argv[1], round(d["value"]), {k: round(v, 3) for k, v in d["stages_ms_per_step"].items()})
