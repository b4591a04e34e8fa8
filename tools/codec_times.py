"""One line of device-resident codec stage times for a config of tools/bench_configs.py (ONLY=C2 / C4)."""
import json, os, subprocess, sys
here = os.path.dirname(os.path.abspath(__file__))
d = json.loads(subprocess.run([sys.executable, os.path.join(here, "bench_configs.py")], capture_output=True, text=True, check=True).stdout)
for k, c in d.items():
    if isinstance(c, dict) and "encode" in c:
        print(k, {s: round(c[s]["ms"], 3) for s in ("encode", "decode", "demux_no_gather", "demux_gather", "decode_demux_gather")},
              "enc frac", round(c["encode"]["frac_of_hbm_peak"], 3), "d+d+g frac", round(c["decode_demux_gather"]["frac_of_hbm_peak"], 3))
