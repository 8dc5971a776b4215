import sys, os, numpy as np
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
from rtu_common import rtu, tracer_b
R = rtu()
z = np.load("tests/golden/rtu_multi3_d3_96.npz"); g = {k: z[k] for k in z.files}
karg = g["karg"].view(R.KARG_B_DT).copy()
port = R.RtuPort(); img = R.Image(g, karg)
rays = port.primary_rays(img, 1)
W = 96
sel = [29 * W + x for x in range(22, 30)] + [30 * W + x for x in range(22, 30)]
h0 = port.trace_rays(img, rays[sel])
ex = g["triEx"].view(np.float32).reshape(-1, 16)
blas = g["blas"].view(R.BLAS_B_DT)
print("blas reflectivity", blas["reflectivity"], "max_depth", karg["max_depth"])
def bounce(r, h):
    out = []
    for ray, hit in zip(r, h):
        o, d = ray[:3], ray[3:]
        I = o + d * hit["dist"]
        e = ex[hit["triIdx"]]
        N = e[3:6] * hit["bx"] + e[6:9] * hit["by"] + e[0:3] * hit["bz"]
        N = N / np.linalg.norm(N)
        Rv = d - 2 * N * np.dot(N, d); Rv = Rv / np.linalg.norm(Rv)
        out.append(np.concatenate([I + Rv * np.float32(0.001), Rv]).astype(np.float32))
    return np.array(out, dtype=np.float32)
tr = tracer_b(g, karg, 1)
r1 = bounce(rays[sel], h0)
for strict in (1, 0):
    tr.set_order(strict)
    hg = tr.trace_rays(r1); hp = port.trace_rays(img, r1)
    print("strict", strict, "secondary identical:", hg.tobytes() == hp.tobytes())
    for a, b, p in zip(hg, hp, h0):
        if a.tobytes() != b.tobytes(): print("   primary blas", p["blasIdx"], "gpu", a, "port", b)
tr.close()
