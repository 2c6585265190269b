#!/usr/bin/env python
"""Development helper: the headline numbers of a bench.py JSON line.  usage: python tools/show_bench.py file.json"""
import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("value %.1f fps (%.4f ms)  e2e %.1f fps  serial %.4f ms" % (d["value"], d["ms_per_step"], d["e2e"]["value"],
                                                                   d["config"].get("serial_ms_per_picture", 0)))
print("stage_ms", d["roofline"]["stage_ms"])
print("clocks", d["clocks"], "identical", d.get("e2e_frames_identical"), "cpu", d["cpu_baseline"]["value"])
