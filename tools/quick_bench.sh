#!/bin/bash
# quick A/B helper: prints it/s and per-stage ms of bench.py for the library currently in place
python bench.py --steps 10 --warmup 3 --no-cpu-baseline "$@" > /tmp/qb.json 2>/tmp/qb.err || { tail -5 /tmp/qb.err; exit 1; }
python - <<'PY'
import json
d = json.loads(open("/tmp/qb.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"]["value"], {k: round(v["ms"], 4) for k, v in d["roofline"]["stages"].items()})
PY
