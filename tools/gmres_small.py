"""Small-problem GMRES: the one-launch cycle (gmres_cycle.cu) against the kernel-per-operation path, history and time.

python tools/gmres_small.py            runs itself twice (IEMIC_GMRES_CYCLE=1 / 0) and compares the histories
python tools/gmres_small.py --child    one leg: prints a JSON line per case
"""
import json
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))

CASES = [("2dmoc", 500, 500, 1, "cgs2"), ("2dmoc", 500, 500, 1, "cgs2f"), ("2dmoc", 100, 300, 0, "cgs2f"), ("natl8", 30, 119, 1, "cgs2"), ("natl8", 30, 119, 0, "cgs2f"),
         ("rand_periodic", 30, 119, 1, "cgs2f"), ("global4", 30, 119, 1, "cgs2f")]


def child():
    import numpy as np
    import torch
    import cases
    import iemic_b200 as ib
    for name, m, maxit, prec, ortho in CASES:
        cfg, mask, pars = cases.case(name)
        G = ib.Ocean(cfg, mask)
        for k, v in pars.items():
            G.setPar(k, v)
        x = cases.state(cfg, mask, seed=6)
        dx = torch.from_numpy(x).cuda()
        G.computeJacobian(dx)
        n = G.nrows
        ones = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
        ones[:n] = 1.0
        b = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
        G.applyMatrix(ones, b)
        d_binv = None
        if prec:
            d_binv = torch.zeros(n * 6, dtype=torch.float64, device="cuda")
            G.block_jacobi(d_binv)
        best, hist, it, rc = 1e30, None, 0, 0
        for rep in range(4):
            d_x = torch.zeros(G.nvec, dtype=torch.float64, device="cuda")
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            rc, hist, it = G.solve(b, d_x, m=m, maxit=maxit, tol=1e-30, ortho=ib.CGS2F if ortho == "cgs2f" else ib.CGS2, prec=ib.PREC_BLOCKJACOBI if prec else ib.PREC_NONE,
                                   d_blockinv=d_binv, check_every=8)
            torch.cuda.synchronize()
            if rep:
                best = min(best, time.perf_counter() - t0)
        err = float((d_x[:n] - 1.0).abs().max())
        print(json.dumps({"case": name + "/" + ortho, "m": m, "prec": prec, "n": int(n), "rc": rc, "iters": it, "seconds": best, "it_per_s": it / best,
                          "x_err": err, "hist": [float(h) for h in hist[:it + 1]]}), flush=True)


def main():
    out = {}
    for flag in ("1", "0"):
        env = dict(os.environ, IEMIC_GMRES_CYCLE=flag)
        env.setdefault("IEMIC_GMRES_CYCLE_ROWS", "400000") # so that the 4-degree grid shows what the cooperative kernel does there
        p = subprocess.run([sys.executable, os.path.abspath(__file__), "--child"], env=env, capture_output=True, text=True, timeout=170)
        if p.returncode != 0:
            print("leg", flag, "failed:", p.stderr[-2000:])
            sys.exit(1)
        out[flag] = [json.loads(l) for l in p.stdout.splitlines() if l.startswith("{")]
    bad = 0
    for a, b in zip(out["1"], out["0"]):
        import numpy as np
        ha, hb = np.array(a["hist"]), np.array(b["hist"])
        k = min(ha.size, hb.size)
        # same Krylov space, different summation order: compare where the residual is well above the rounding floor
        sel = hb[:k] > 1e-9
        rel = float(np.max(np.abs(ha[:k][sel] - hb[:k][sel]) / hb[:k][sel])) if sel.any() else 0.0
        ok = a["iters"] == b["iters"] and a["rc"] == b["rc"] and rel < 1e-5
        bad += not ok
        print(f'{a["case"]:20s} m={a["m"]:3d} prec={a["prec"]} n={a["n"]:7d} iters {a["iters"]}/{b["iters"]} rc {a["rc"]}/{b["rc"]} '
              f'one-launch {a["it_per_s"]:9.0f} it/s  per-kernel {b["it_per_s"]:9.0f} it/s  hist rel diff {rel:.2e} x_err {a["x_err"]:.1e}/{b["x_err"]:.1e} '
              f'{"ok" if ok else "MISMATCH"}')
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    child() if "--child" in sys.argv else main()
