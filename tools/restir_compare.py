"""Per-frame ReSTIR DI comparison of the CUDA passes with the oracle (matched RNG streams): mismatch rates of the initial-candidate
and output reservoirs, and the image error. Usage: python tools/restir_compare.py [W H frames boxes lights]"""
import importlib
import sys

import numpy as np

sys.path.insert(0, ".")
pt = importlib.import_module("hiprt-path-tracer_b200")
wl = importlib.import_module("hiprt-path-tracer_b200.workloads")
from oracle import ref  # noqa: E402  (test tooling: the oracle is the checker)

RES = np.dtype([("M", "<i4"), ("weight_sum", "<f4"), ("UCW", "<f4"), ("tri", "<i4"), ("point", "<f4", 3), ("target", "<f4"), ("flags", "u1"), ("pad", "u1", 3)])


def res(raw):
    return np.ascontiguousarray(raw).view(RES).reshape(raw.shape[0], raw.shape[1])


def main():
    overrides = {k: float(v) if "." in v else int(v) for k, v in (arg.split("=") for arg in sys.argv[1:] if "=" in arg)}
    a = [int(v) for v in sys.argv[1:] if "=" not in v]
    w, h, frames, boxes, lights = (a + [128, 96, 4, 150, 3000][len(a):])[:5]
    cfg = wl.many_lights(w, h, boxes=boxes, lights=lights)
    cfg["settings"] = dict(cfg["settings"], **{"restir_di_settings." + k: v for k, v in overrides.items()})
    print("overrides", overrides)
    ctx = pt.Context(0)
    r = wl.setup_oracle(ref, pt.structs, cfg)
    g = wl.setup_gpu(pt, ctx, cfg, seed_schedule="cpu")
    prev_bad = None
    for f in range(frames):
        r.render(1)
        g.render_frame()
        g.synchronize()
        ia, ib = res(r.buffer("restir_initial_candidates")), res(g.download("restir_initial_candidates"))
        oa, ob = res(r.buffer("restir_output_reservoirs")), res(g.download("restir_output_reservoirs"))
        ibad = (ia["tri"] != ib["tri"]) | (ia["M"] != ib["M"])
        obad = (oa["tri"] != ob["tri"]) | (oa["M"] != ob["M"])
        line = f"frame {f}: initial mismatch {ibad.mean():.5f}  output mismatch {obad.mean():.5f} (M only {(oa['M'] != ob['M']).mean():.5f})"
        if prev_bad is not None:
            # how many of this frame's mismatches sit within 32 px of a mismatch of the previous frame (cascade) ?
            ys, xs = np.nonzero(prev_bad | ibad)
            near = np.zeros_like(obad)
            for y, x in zip(ys, xs):
                near[max(0, y - 32):y + 33, max(0, x - 32):x + 33] = True
            line += f"  near-previous {(obad & near).sum()}/{obad.sum()}"
        ok = ~obad & np.isfinite(oa["UCW"])
        rel = np.abs(oa["UCW"][ok] - ob["UCW"][ok]) / np.maximum(np.abs(oa["UCW"][ok]), 1e-12)
        line += f"  UCW rel err max {rel.max():.2e} mean {rel.mean():.2e}"
        pa, pb = r.pixels(), g.download("pixels")
        line += f"  image rel {np.abs(pa - pb).mean() / pa.mean():.2e} mean ratio {pb.mean() / pa.mean():.6f}"
        print(line)
        if f == 1 and not overrides:
            ys, xs = np.nonzero(obad)
            for y, x in list(zip(ys, xs))[:24]:
                print(f"  ({x},{y}) oracle M {oa['M'][y,x]} tri {oa['tri'][y,x]} UCW {oa['UCW'][y,x]:.4g} ws {oa['weight_sum'][y,x]:.4g} | gpu M {ob['M'][y,x]} tri {ob['tri'][y,x]} UCW {ob['UCW'][y,x]:.4g} ws {ob['weight_sum'][y,x]:.4g} | initial M {ia['M'][y,x]} prev-out M {prev_out['M'][y,x]} prim {prim[y,x]}")
            ys, xs = np.nonzero(~obad & (np.abs(oa["UCW"] - ob["UCW"]) > 1e-2 * np.abs(oa["UCW"])))
            for y, x in list(zip(ys, xs))[:12]:
                print(f"  UCW ({x},{y}) oracle M {oa['M'][y,x]} tri {oa['tri'][y,x]} UCW {oa['UCW'][y,x]:.4g} ws {oa['weight_sum'][y,x]:.4g} tf {oa['target'][y,x]:.4g} | gpu UCW {ob['UCW'][y,x]:.4g} ws {ob['weight_sum'][y,x]:.4g} tf {ob['target'][y,x]:.4g}")
        prev_out = oa.copy()
        prim = r.buffer("first_hit_prim_index").reshape(oa.shape)
        prev_bad = obad
    g.close()


if __name__ == "__main__":
    main()
