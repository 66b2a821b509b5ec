"""Development helper: print the demodulator numbers of a bench.py JSON line (file given as argv[1])."""
import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
m = d["demodulator"]
print("demod ms %.4f  GB/s %.0f  | fixed6 %.2f  early %.2f  pusch %.2f" % (m["ms_per_step"], m["roofline"]["achieved"], d["value"], d["early_stop"]["value"], d["pusch_slot"]["value"]))
