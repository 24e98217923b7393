"""Print the one-line essentials of bench JSON files given on the command line."""
import json, sys
for f in sys.argv[1:]:
    try:
        b = json.loads(open(f).read().strip().splitlines()[-1])
        st = b["stage_ms"]
        print(f, b["config"]["workload"], "ms/step %.1f" % b["ms_per_step"], " ".join("%s=%.1f" % (k, v) for k, v in st.items() if v > 0.5),
              "frac %.3f" % b["roofline"]["frac"], "E %.15e" % b["energy"], "t3", b["counts"]["t3"])
    except Exception as e:
        print(f, "unreadable:", e)
