"""Development helper: formats the output of tools/variants.sh as one line per library variant."""
import sys, json
name = "?"
for l in sys.stdin:
    if l.startswith("=="):
        name = l.split()[1].split("/")[-1] + " " + l.split()[2]
    elif l.startswith("{"):
        d = json.loads(l)
        print("%-44s total %.2f ms  float %.1f GCUPS  all %.1f" % (name, d["best_ms"], d["float_gcups"], d["gcups"]))
    else:
        print(name, l.strip()[:200])
