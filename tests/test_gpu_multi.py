"""Multi-GPU tests (need >= 2 GPUs on the box; skipped otherwise).  Multi-GPU is a property of the backend: ONE
vx_device_h spans N GPUs (vx_b200_select_gpus / VORTEX_B200_GPUS) — uploads are sliced over the GPUs' PCIe links and replicated
over NVLink, every GPU renders its bands, the framebuffer is read band-wise.  Everything such a device returns must equal what a
one-GPU device returns, which the other test files compare with the CPU oracle; the frame is compared with the oracle here too."""
import hashlib
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from rt_common import ROOT, assert_hit_gates, best_oracle, channel_diff, hit_parity, pyoracle, scenes, tracer_for, vx

pytestmark = pytest.mark.gpu
po = pyoracle()


def n_gpus():
    import torch
    return torch.cuda.device_count()


needs2 = pytest.mark.skipif("n_gpus() < 2", reason="needs 2 GPUs")


def scene_and_args(oracle, w=960, h=540):
    sc = oracle.scene_create([scenes.heightfield(224), scenes.icosphere((0, 0.5, 0), 0.3, 2, reflectivity=0.4)])
    karg = sc.setup(w, h, max_depth=2)
    return sc, karg


@needs2
@pytest.mark.parametrize("env", [{}, {"VORTEX_B200_FB": "peer"}, {"VORTEX_B200_COLLECTIVE": "nccl"}, {"VORTEX_B200_SLICE_MIN": "4096"}],
                         ids=["default", "fb_peer", "nccl_allgather", "slice_everything"])
def test_group_device_equals_single_gpu_and_oracle(env):
    oracle = best_oracle()
    sc, karg = scene_and_args(oracle)
    w, h = 960, 540
    one = tracer_for(sc, karg)
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        two = tracer_for(sc, karg, gpus=2)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    try:
        assert two.dev.caps(0xA) == 2 and two.dev.caps(0x3) == 2 * one.dev.caps(0x3)      # clusters = GPUs, cores = all SMs
        for mode in (1, 0):
            for strict in (1, 0):
                one.set_mode(mode, strict)
                two.set_mode(mode, strict)
                a, b = one.run().copy(), two.run().copy()
                assert np.array_equal(a, b), (mode, strict)
                st = two.dev.stats()
                assert st["gpus"] == 2 and st["group_ms"] > 0 and st["rays"] == one.dev.stats()["rays"]
                if strict:
                    ref, _ = sc.render(mode=mode, threads=oracle.hw_threads())
                    assert np.array_equal(b, ref), mode
        # a read that does not cover the whole image gathers the bands on GPU 0 first
        two.set_mode(1, 0)
        one.set_mode(1, 0)
        a = one.run().copy()
        two.dev.start(two.krnl, two.args)
        two.dev.ready_wait()
        part = np.zeros((47, w), dtype=np.uint32)
        two.out.copy_from_dev(part, src_offset=3 * w * 4)
        assert np.array_equal(part, a[3:50])
        whole = np.zeros((h, w), dtype=np.uint32)
        two.out.copy_from_dev(whole)
        assert np.array_equal(whole, a)
        # hit records: rays sharded by index over the GPUs, read back piece-wise or whole
        rays = scenes.random_rays(70_001, (-1.5, -1.2, -1.5), (1.5, 1.2, 1.5), seed=9)
        want = one.trace_rays(rays)
        assert two.trace_rays(rays).tobytes() == want.tobytes()
        assert two.dev.stats()["rays"] == len(rays)
        assert two.trace_rays_host(rays).tobytes() == want.tobytes()
        # a partial read of the hit records gathers the pieces on GPU 0 first (peer copies, or ncclSend/ncclRecv)
        rb, hb = two.dev.upload_bytes(rays), two.dev.mem_alloc(len(rays) * 24)
        two.dev.trace_rays(two.args, rb, len(rays), hb)
        two.dev.ready_wait()
        tail = np.zeros(len(rays) - 20_000, dtype=want.dtype)
        hb.copy_from_dev(tail, src_offset=20_000 * 24)
        assert tail.tobytes() == want[20_000:].tobytes()
        rb.free()
        hb.free()
        assert hit_parity(two.trace_primary(), one.trace_primary())["mismatch"] == 0
        assert_hit_gates(hit_parity(want, sc.trace_rays(rays, threads=oracle.hw_threads())), False)
        # a partial upload into a replicated buffer reaches every GPU
        blas = one.host_scene["blas"]
        for t in (one, two):
            t.bufs["blas"].copy_to_dev(np.ascontiguousarray(blas[160:320]), dst_offset=160)
        assert np.array_equal(two.run(), a)
    finally:
        one.close()
        two.close()


@needs2
def test_group_readback_and_gpu_build():
    import torch
    oracle = best_oracle()
    sc, karg = scene_and_args(oracle, 640, 360)
    two = tracer_for(sc, karg, gpus=2)
    one = tracer_for(sc, karg)
    try:
        two.set_mode(1, 0)
        one.set_mode(1, 0)
        want = one.run().copy()
        host = torch.zeros((360, 640), dtype=torch.int32).pin_memory()
        two._upload_args()
        two.dev.set_readback(two.out, host.data_ptr(), 640 * 360 * 4)
        two.dev.start(two.krnl, two.args)
        two.dev.ready_wait()
        assert np.array_equal(host.numpy().view(np.uint32), want)          # the image is in host memory when vx_ready_wait returns
        two.out.copy_from_dev((host.data_ptr(), 640 * 360 * 4))             # ... and the reference's own download call still works
        assert np.array_equal(host.numpy().view(np.uint32), want)
        other = np.zeros((360, 640), dtype=np.uint32)
        two.out.copy_from_dev(other)                                        # any other destination is an ordinary band-wise read
        assert np.array_equal(other, want)
        two.dev.set_readback(two.out, None, 0)
        with pytest.raises(vx.VxError):
            two.dev.set_readback(two.out, other.ctypes.data, other.nbytes)  # pageable memory is refused
        # GPU BVH build on a group: GPU 0 builds, the tree is replicated; the frame equals the one-GPU device's with the same call
        ntri = one.host_scene["tri"].nbytes // 36
        n0 = 2 * 224 * 224
        for t in (one, two):
            t.dev.build_bvh(t.bufs["tri"], n0, t.bufs["bvh"], t.bufs["triIdx"])
        a, b = one.run().copy(), two.run().copy()
        assert np.array_equal(a, b) and ntri > n0
        assert channel_diff(a, want) <= 1
    finally:
        one.close()
        two.close()


REF_APP = os.path.join(ROOT, "oracle", "_ref", "raycast")


@needs2
@pytest.mark.skipif(not (os.path.exists(REF_APP) and os.path.exists(os.path.join(ROOT, "oracle", "_ref", "assets", "teapot.obj"))),
                    reason="the compiled reference app (oracle/_ref/raycast + assets) did not travel with the snapshot")
def test_unmodified_reference_app_on_all_gpus(tmp_path):
    """the reference's own raycast binary, VORTEX_DRIVER=b200 and VORTEX_B200_GPUS=<all GPUs of the box>: same PPM as its CPU path"""
    lib_dir = os.path.dirname(vx.LIB_PATH)
    kernel = tmp_path / "kernel.vxbin"
    kernel.write_bytes(vx.DUMMY_KERNEL_IMAGE)
    for gpus in sorted({2, n_gpus()}):
        out = tmp_path / f"out{gpus}.ppm"
        env = dict(os.environ, VORTEX_DRIVER="b200", VORTEX_B200_ORDER="1", VORTEX_B200_GPUS=str(gpus), VORTEX_B200_SLICE_MIN="4096",
                   LD_LIBRARY_PATH=lib_dir + ":" + os.path.dirname(REF_APP) + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
        r = subprocess.run([REF_APP, "-k", str(kernel), "-o", str(out), "-n3", "-d4", "-w64", "-h64"], cwd=os.path.dirname(REF_APP), env=env,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
        assert r.returncode == 0, r.stdout
        assert hashlib.md5(out.read_bytes()).hexdigest() == "d8ecb7285f1920568d8f32abe0ad600c", (gpus, r.stdout[-400:])


@needs2
def test_bench_under_torchrun_two_gpus():
    """the driver's launch shape: torchrun with one rank per GPU; rank 0 drives ONE vx device over both GPUs"""
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "bench.py"), "--gpus", "2", "--steps", "3", "--warmup", "1",
           "--workload", "cfg1_1080p_primary+shadow_100k", "--no-cpu", "--extra", "0"]
    r = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["n_gpus"] == 2 and line["config"]["torchrun_world"] == 2
    assert line["parity"]["green"] and line["parity"]["non_tie"] == 0 and line["parity"]["max_channel_diff"] <= 1
    assert line["parity"]["e2e_frame_equals_timed_frame"] and line["e2e"]["with_readback"]["same_frame"]
    assert line["gpu_launches"] >= 2 * 3
