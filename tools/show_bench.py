"""Print the interesting numbers of a bench.py JSON line (argv[1])."""
import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
tr = d.pop("train", None)
if "value" in d and "config" in d and "us_per_decoder_step" in d["config"]:
    print("AR  frames/s %.0f  ms/step %.2f  us/frame %.2f  e2e %.0f" % (d["value"], d["ms_per_step"], d["config"]["us_per_decoder_step"], d["e2e"]["value"]))
    print("   ", d["config"]["phase_us_per_frame"])
    print("    roofline frac %.3f" % d["roofline"]["frac"])
if tr:
    print("TRAIN frames/s %.0f  ms/step %.2f" % (tr["value"], tr["ms_per_step"]), tr["phase_ms"])
    print("   ", tr.get("decoder_bwd_phase_us_per_frame"))
