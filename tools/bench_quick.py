"""seal / open / demux stage times of the bench's device-resident step (C5), nothing else: for comparing builds
(P2PT_LIB=build/var/x.so python tools/bench_quick.py)"""
import json, subprocess, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "5", "--warmup", "3", "--no-configs", "--no-e2e"],
                     capture_output=True, text=True)
line = [l for l in out.stdout.splitlines() if l.startswith("{")]
if not line:
    print(out.stderr[-400:]); sys.exit(1)
d = json.loads(line[-1]); s = d.get("stages", {})
print(f"value {d['value']:.1f} GB/s  step {d['ms_per_step']:.3f} ms  seal {s.get('seal_ms', 0):.3f}  open {s.get('open_ms', 0):.3f}  demux+gather {s.get('demux_gather_ms', 0):.3f}")
