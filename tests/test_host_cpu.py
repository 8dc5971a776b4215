"""CPU tests (no GPU) of the host side: the C-ABI library loads and exports every declared symbol, the header and the
ctypes binding agree, the band partition covers every row exactly once, and the world_size-2 framebuffer gather (gloo)
reassembles a frame rendered in bands by the oracle."""
import ctypes
import os
import re
import socket
import subprocess
import sys

import numpy as np
import pytest

from rt_common import ROOT, pyoracle, scenes, vx

multigpu = __import__("importlib").import_module("vortex-raytracing_b200.multigpu")


def test_library_exports_every_declared_symbol():
    if not os.path.exists(vx.LIB_PATH):
        subprocess.run(["make", "-C", os.path.join(ROOT, "vortex-raytracing_b200", "csrc")], check=True, stdout=subprocess.DEVNULL)
    header = open(os.path.join(ROOT, "include", "vortex_b200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(vx_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 30
    assert declared == set(vx.EXPORTS), declared ^ set(vx.EXPORTS)
    L = ctypes.CDLL(vx.LIB_PATH)          # loads without a GPU (no compute call is made)
    for name in declared:
        assert hasattr(L, name), name
    L.vx_b200_last_error.restype = ctypes.c_char_p
    assert L.vx_b200_last_error() is not None


def test_no_cpu_fallback_without_gpu():
    """on a box without a CUDA device vx_dev_open must fail loudly (non-zero + message), never compute on the CPU"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    L = vx.lib()
    h = ctypes.c_void_p()
    assert L.vx_dev_open(ctypes.byref(h)) != 0
    assert b"no CUDA device" in L.vx_b200_last_error() or b"failed" in L.vx_b200_last_error()
    with pytest.raises(vx.VxError):
        vx.Device()


def test_product_never_imports_oracle():
    """the product path must not route through the oracle (or any CPU fallback)"""
    pkg_dir = os.path.join(ROOT, "vortex-raytracing_b200")
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert "pyoracle" not in src and "liboracle_port" not in src and "libref_raycast" not in src, f


@pytest.mark.parametrize("H,world,band", [(1080, 1, 8), (1080, 2, 8), (1080, 3, 8), (4320, 8, 8), (37, 4, 8), (5, 8, 4), (64, 2, 16)])
def test_band_partition_covers_every_row_once(H, world, band):
    seen = np.zeros(H, dtype=np.int32)
    for r in range(world):
        rows = multigpu.owned_rows(r, world, H, band)
        seen[rows] += 1
        assert all((y // band) % world == r for y in rows)
    assert (seen == 1).all()
    assert sum(multigpu.rays_per_rank(10, H, world, band)) == 10 * H


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


_WORKER = r"""
import os, sys, importlib
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "tests"))
from rt_common import pyoracle, scenes
mg = importlib.import_module("vortex-raytracing_b200.multigpu")
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
po = pyoracle()
orc = po.PortRaycast()
W, H, band = {W}, {H}, {band}
sc = orc.scene_create([scenes.heightfield(12)])
karg = sc.setup(W, H)
full, _ = sc.render()
fb = np.zeros((H, W), dtype=np.uint32)
rows = mg.owned_rows(rank, world, H, band)
# each rank renders only its own bands (row ranges), as a GPU rank would after vx_b200_set_partition
for y in rows:
    part, _ = sc.render(rows=(int(y), int(y) + 1))
    fb[y] = part[y]
t = torch.from_numpy(fb.view(np.int32))
mg.gather_framebuffer(t, rank, world, band)
if rank == 0:
    assert np.array_equal(t.numpy().view(np.uint32), full), "gathered frame differs from the single-rank frame"
else:
    other = np.setdiff1d(np.arange(H), rows)
    assert not t.numpy()[other].any()
dist.barrier()
dist.destroy_process_group()
print("OK", rank)
"""


@pytest.mark.parametrize("W,H,band,world", [(40, 64, 8, 2), (33, 37, 8, 2), (16, 20, 4, 3)])
def test_framebuffer_gather_gloo(tmp_path, W, H, band, world):
    po = pyoracle()
    if not po.have_port():
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "port"], check=True, stdout=subprocess.DEVNULL)
    script = tmp_path / "worker.py"
    script.write_text(_WORKER.format(root=ROOT, W=W, H=H, band=band))
    port = _free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    for p in procs:
        out, _ = p.communicate(timeout=240)
        assert p.returncode == 0, out
        assert "OK" in out


host = __import__("importlib").import_module("vortex-raytracing_b200.host")

HOST_CASES = [
    ("cornell", lambda: [scenes.cornell_box()]),
    ("heightfield64", lambda: [scenes.heightfield(64)]),
    ("multi4", lambda: [scenes.heightfield(16), scenes.icosphere((0, 0.5, 0), 0.4, 2, reflectivity=0.5), scenes.cornell_box(0.3),
                        scenes.icosphere((0.5, 0.5, 0.3), 0.2, 1)]),
    ("degenerate", lambda: [{"tris": np.zeros((3, 9), np.float32), "triex": np.zeros((3, 15), np.float32),
                             "tex": np.full((1, 1), 0x808080, np.uint32), "reflectivity": 0.0}, scenes.cornell_box()]),
]


def test_host_library_exports():
    header = open(os.path.join(ROOT, "include", "vortex_b200_host.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(vxh_[a-z0-9_]+)\s*\(", header))
    assert declared == set(host.EXPORTS) | set(host.RTX_EXPORTS) | set(host.ASSET_EXPORTS)
    L = ctypes.CDLL(host.LIB_PATH)
    for n in declared:
        assert hasattr(L, n), n


@pytest.mark.parametrize("name,mk", HOST_CASES, ids=[c[0] for c in HOST_CASES])
@pytest.mark.parametrize("threads", [1, 4])
def test_host_builders_emit_the_reference_bytes(name, mk, threads):
    """the product's host builders (task-parallel BVH2, TLAS, Tracer::setup maths) vs the oracle: identical upload bytes"""
    po = pyoracle()
    orc = po.best()
    meshes = mk()
    a = host.HostScene(meshes, threads=threads)
    b = orc.scene_create(meshes)
    ka, kb = a.setup(96, 54, spp=2, max_depth=3), b.setup(96, 54, spp=2, max_depth=3)
    ba, bb = a.buffers(), b.buffers()
    for n in po.BUFFER_NAMES:
        assert ba[n].tobytes() == bb[n].tobytes(), n
    assert ba["tlas_root"] == bb["tlas_root"] and ba["bvh_nodes_used"] == bb["bvh_nodes_used"]
    for f in ka.dtype.names:
        if not f.endswith("_addr"):
            assert ka[f].tobytes() == kb[f].tobytes(), f
        else:
            assert int(ka[f][0]) == 0


def test_host_builder_large_parallel_equals_serial():
    m = [scenes.heightfield(160)]
    a, b = host.HostScene(m, threads=1), host.HostScene(m, threads=8)
    a.setup(8, 8), b.setup(8, 8)
    ba, bb = a.buffers(), b.buffers()
    for n in ("bvh", "triIdx", "tlas", "blas"):
        assert ba[n].tobytes() == bb[n].tobytes(), n


def test_parity_helpers_generate_the_reference_primary_rays():
    """vortex-raytracing_b200/parity.py: the numpy GenerateRay used by bench.py's parity block equals the oracle's own primary rays
    (render.h:190-208) bit for bit on a small frame, and the band sampler hits every GPU of a round-robin deal"""
    import importlib
    par = importlib.import_module("vortex-raytracing_b200.parity")
    po = importlib.import_module("oracle.pyoracle")
    orc = po.best()
    sc = orc.scene_create([scenes.heightfield(8)])
    karg = sc.setup(96, 56)
    _, rays = sc.trace_primary(want_rays=True, threads=2)
    mine = par.primary_rays(karg, range(56))
    assert mine.shape == rays.shape
    assert np.array_equal(mine[:, :3], rays[:, :3])
    assert np.abs(mine[:, 3:] - rays[:, 3:]).max() <= 1.2e-7          # same formula in float32; the last bit may differ
    bench = importlib.import_module("bench")
    for gpus in (1, 2, 4, 8):
        bands = bench.sample_bands(4320, gpus)
        assert len(bands) == 8 and len({b % gpus for b in bands}) == min(gpus, 8)
    a = np.zeros(4, dtype=po.HIT_DT)
    a["dist"] = [1.0, 2.0, 1e30, 3.0]
    a["triIdx"] = [1, 2, 0, 3]
    b = a.copy()
    b["triIdx"][0] = 9                                                  # same t, other triangle: a tie
    b["dist"][3], b["triIdx"][3] = 4.0, 7                               # other t: not a tie
    hp = par.hit_parity(b, a)
    assert (hp["mismatch"], hp["ties"], hp["non_tie"]) == (2, 1, 1)
