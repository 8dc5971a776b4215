"""Host-side driver of the RT-unit ("raytracing", layout B) path over the vx_* C ABI — the call sequence of the
reference's `Tracer` in tests/regression/raytracing/tracer.cpp: init (:91-193: open device, upload the kernel and the
miss / closest-hit / any-hit shader images, allocate the ten scene buffers + framebuffer + SBT), setup (:195-258: copy
the buffers, fill the SBT, write the RT unit's DCRs 0x006..0x009 with the 32-bit TLAS/BLAS/qBVH/triangle addresses),
run (:260-287: upload kernel_arg_t, vx_start, vx_ready_wait, download the framebuffer).

Scene buffers are the reference's layout-B byte arrays (tri, triEx, triIdx, mat, tex, bvh, qbvh, blas, tlas[quantized]).
"""
import numpy as np

from . import vx

KARG_B_DT = np.dtype([   # tests/regression/raytracing/common.h:164-195 — 216 bytes
    ("dst_width", "<u4"), ("dst_height", "<u4"), ("dst_addr", "<u8"),
    ("tri_addr", "<u8"), ("triEx_addr", "<u8"), ("triIdx_addr", "<u8"), ("mat_addr", "<u8"), ("tex_addr", "<u8"),
    ("bvh_addr", "<u8"), ("qBvh_addr", "<u8"), ("blas_addr", "<u8"), ("tlas_addr", "<u8"), ("tlas_root", "<u4"),
    ("camera_pos", "<f4", 3), ("camera_forward", "<f4", 3), ("camera_right", "<f4", 3), ("camera_up", "<f4", 3),
    ("viewplane", "<f4", 2), ("samples_per_pixel", "<u4"), ("max_depth", "<u4"),
    ("light_pos", "<f4", 3), ("light_color", "<f4", 3), ("ambient_color", "<f4", 3), ("background_color", "<f4", 3),
    ("_pad", "<u4"), ("sbt_addr", "<u8"),
])
assert KARG_B_DT.itemsize == 216
HIT_DT = np.dtype([("dist", "<f4"), ("bx", "<f4"), ("by", "<f4"), ("bz", "<f4"), ("blasIdx", "<u4"), ("triIdx", "<u4")])

SCENE_BUFFERS = ("tri", "triEx", "triIdx", "tlas", "blas", "bvh", "qbvh", "mat", "tex")   # allocation order of tracer.cpp:156-186
_ADDR_FIELD = {"tri": "tri_addr", "triEx": "triEx_addr", "triIdx": "triIdx_addr", "tlas": "tlas_addr", "blas": "blas_addr",
               "bvh": "bvh_addr", "qbvh": "qBvh_addr", "mat": "mat_addr", "tex": "tex_addr"}
CAMERA_FIXED, CAMERA_PINHOLE = 0, 1   # kernel.cpp:28-39 (what the shipped kernel runs) / kernel.cpp:13-26


class RtxTracer:
    def __init__(self, width, height, spp=1, max_depth=1, cuda_device=None, gpus=None):
        self.width, self.height = int(width), int(height)
        self.karg = np.zeros(1, dtype=KARG_B_DT)
        self.karg["dst_width"], self.karg["dst_height"] = self.width, self.height
        self.karg["samples_per_pixel"], self.karg["max_depth"] = spp, max_depth
        self.dev = vx.Device(cuda_device, gpus=gpus)
        self.bufs, self.shaders = {}, []
        self.krnl = self.out = self.args = self.sbt = None

    # Tracer::init (tracer.cpp:144-190)
    def init(self, scene_buffers):
        self.krnl = self.dev.upload_kernel_bytes(vx.DUMMY_KERNEL_IMAGE)
        # miss.vxbin / closest.vxbin / anyhit.vxbin: RISC-V images, accepted and ignored by this backend
        self.shaders = [self.dev.upload_kernel_bytes(vx.dummy_image(v)) for v in (0x80100000, 0x80200000, 0x80300000)]
        for name in SCENE_BUFFERS:
            host = scene_buffers[name]
            b = self.dev.mem_alloc(max(int(host.nbytes), 4), vx.VX_MEM_READ)
            self.bufs[name] = b
            self.karg[_ADDR_FIELD[name]] = b.address
        self.out = self.dev.mem_alloc(self.width * self.height * 4, vx.VX_MEM_WRITE)
        self.karg["dst_addr"] = self.out.address
        self.sbt = self.dev.mem_alloc(8 * 4, vx.VX_MEM_READ)
        self.karg["sbt_addr"] = self.sbt.address
        self.host_scene = scene_buffers
        return self

    # Tracer::setup (tracer.cpp:195-258)
    def setup(self, karg_fields, camera=CAMERA_PINHOLE):
        for k in ("tlas_root", "camera_pos", "camera_forward", "camera_right", "camera_up", "viewplane", "light_pos",
                  "light_color", "ambient_color", "background_color"):
            self.karg[k] = karg_fields[k]
        for name in SCENE_BUFFERS:
            host = self.host_scene[name]
            if host.nbytes:
                self.bufs[name].copy_to_dev(np.ascontiguousarray(host))
        sbt = np.array([self.shaders[0].address, self.shaders[1].address, 0, self.shaders[2].address], dtype=np.uint64)
        self.sbt.copy_to_dev(sbt)
        mask = 0xFFFFFFFF   # (uint32_t)(kernel_arg_.x_addr), tracer.cpp:252-256
        self.dev.dcr_write(vx.VX_DCR_BASE_RTX_TLAS_PTR, int(self.karg["tlas_addr"][0]) & mask)
        self.dev.dcr_write(vx.VX_DCR_BASE_RTX_BLAS_PTR, int(self.karg["blas_addr"][0]) & mask)
        self.dev.dcr_write(vx.VX_DCR_BASE_RTX_BVH_PTR, int(self.karg["qBvh_addr"][0]) & mask)
        self.dev.dcr_write(vx.VX_DCR_BASE_RTX_TRI_PTR, int(self.karg["tri_addr"][0]) & mask)
        self.dev.dcr_write(vx.VX_DCR_B200_RT_CAMERA, int(camera))
        return self

    def _upload_args(self):
        if self.args is not None:
            self.args.free()
        self.args = self.dev.upload_bytes(self.karg)

    # Tracer::run (tracer.cpp:270-283)
    def run(self, host_out=None):
        self._upload_args()
        self.dev.start(self.krnl, self.args)
        self.dev.ready_wait(vx.VX_MAX_TIMEOUT)
        if host_out is None:
            host_out = np.empty((self.height, self.width), dtype=np.uint32)
        self.out.copy_from_dev(host_out)
        return host_out

    def launch(self):
        self.dev.start(self.krnl, self.args)
        self.dev.ready_wait(vx.VX_MAX_TIMEOUT)

    def trace_primary(self):
        self._upload_args()
        n = self.width * self.height
        hits = self.dev.mem_alloc(n * 24)
        try:
            self.dev.trace_primary(self.args, hits)
            self.dev.ready_wait()
            out = np.empty(n, dtype=HIT_DT)
            hits.copy_from_dev(out)
        finally:
            hits.free()
        return out

    def trace_rays(self, rays):
        self._upload_args()
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        n = rays.shape[0]
        rb = self.dev.upload_bytes(rays)
        hits = self.dev.mem_alloc(n * 24)
        try:
            self.dev.trace_rays(self.args, rb, n, hits)
            self.dev.ready_wait()
            out = np.empty(n, dtype=HIT_DT)
            hits.copy_from_dev(out)
        finally:
            hits.free()
            rb.free()
        return out

    def trace_rays_host(self, rays, out=None):
        """the same hit records with rays and records left in host memory (vx_b200_trace_rays_host: chunked, pipelined)"""
        self._upload_args()
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        n = rays.shape[0]
        if out is None:
            out = np.empty(n, dtype=HIT_DT)
        if n:
            self.dev.trace_rays_host(self.args, rays.ctypes.data, n, out.ctypes.data)
        return out

    def set_order(self, strict_order=False):
        self.dev.dcr_write(vx.VX_DCR_B200_RT_ORDER, int(bool(strict_order)))

    def set_mode(self, mode=0):
        """VX_DCR_B200_RT_MODE: bit0 shadow rays, bit1 diffuse bounces (the builder-defined extension of BASELINE configs 1-3)"""
        self.dev.dcr_write(vx.VX_DCR_B200_RT_MODE, int(mode))

    def set_program(self, program=None):
        """None = the fused kernel; n = shader program n written against the traceRay/getWork/getAttr/commit shim
        (csrc/vx_raytrace_b200.cuh; 0 = the three shipped shaders, 1 = the cut-out any-hit demo)"""
        self.dev.dcr_write(vx.VX_DCR_B200_RT_KERNEL, 0 if program is None else 2 + int(program))

    def close(self):
        for b in list(self.bufs.values()) + self.shaders + [self.out, self.args, self.krnl, self.sbt]:
            if b is not None:
                b.free()
        self.bufs, self.shaders, self.out, self.args, self.krnl, self.sbt = {}, [], None, None, None, None
        self.dev.close()
