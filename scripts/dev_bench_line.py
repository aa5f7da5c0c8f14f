"""Developer helper: reads bench.py's JSON line on stdin and prints value, e2e, ms/step and the stage times."""
import json, sys
a = json.loads(sys.stdin.read())
print(sys.argv[1] if len(sys.argv) > 1 else "", a["value"], a["e2e"]["value"], a["ms_per_step"], a["config"]["stage_ms_last_step"])
