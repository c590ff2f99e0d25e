"""Print one line per bench JSON file given on the command line."""
import json
import sys

for f in sys.argv[1:]:
    try:
        line = [x for x in open(f) if x.startswith("{")][-1]
        r = json.loads(line)
        print(f.split("/")[-1], round(r["value"]), "img/s", round(r["ms_per_step"] * 1e3, 1), "us", "e2e",
              round(r.get("e2e", {}).get("value", 0)), "launches/step", round(r["gpu_launches"] / r["steps"], 2))
    except Exception as e:  # noqa: BLE001
        print(f, "FAILED", e)
        print(open(f).read()[-1500:])
