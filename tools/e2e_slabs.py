"""End-to-end assemblies/s of thcm_b200_assemble_host for several slab counts of its upload/compute/download pipeline
(IEMIC_SLABS; measured on one B200 at 1 degree: 2 -> 280, 4 -> 353, 6 -> 376, 8 -> 385, 16 -> 368 assemblies/s)."""
import os, sys, json, subprocess
for sl in (2, 4, 6, 8, 12):
    env = dict(os.environ, IEMIC_SLABS=str(sl))
    r = subprocess.run([sys.executable, "bench.py", "--no-cpu-baseline", "--skip-gmres"], capture_output=True, text=True, env=env)
    d = json.loads(r.stdout.strip().splitlines()[-1])
    print(sl, round(d["e2e"]["value"], 1), flush=True)
