"""Host-side driver of the raycast path over the vx_* C ABI — the call sequence of the reference's
`Tracer` (tests/regression/raycast/tracer.cpp): init (:89-150: open device, upload kernel image, allocate the
seven scene buffers + framebuffer and record their addresses in kernel_arg_t), setup (:152-219: copy the scene
buffers to the device), run (:221-247: upload kernel_arg_t, vx_start, vx_ready_wait, download the framebuffer).

The scene buffers are the reference's layout-A byte arrays (tri, triEx, triIdx, tlas, blas, bvh, tex), from the
host builder (vortex-raytracing_b200/host) or any Vortex host application.
"""
import numpy as np

from . import vx

# numpy view of the 192-byte kernel_arg_t (tests/regression/raycast/common.h:120-147)
KARG_DT = np.dtype([
    ("dst_width", "<u4"), ("dst_height", "<u4"), ("dst_addr", "<u8"),
    ("tri_addr", "<u8"), ("triEx_addr", "<u8"), ("triIdx_addr", "<u8"), ("tex_addr", "<u8"),
    ("bvh_addr", "<u8"), ("blas_addr", "<u8"), ("tlas_addr", "<u8"), ("tlas_root", "<u4"),
    ("camera_pos", "<f4", 3), ("camera_forward", "<f4", 3), ("camera_right", "<f4", 3), ("camera_up", "<f4", 3),
    ("viewplane", "<f4", 2), ("samples_per_pixel", "<u4"), ("max_depth", "<u4"),
    ("light_pos", "<f4", 3), ("light_color", "<f4", 3), ("ambient_color", "<f4", 3), ("background_color", "<f4", 3),
    ("_pad", "<u4"),
])
assert KARG_DT.itemsize == 192
HIT_DT = np.dtype([("dist", "<f4"), ("bx", "<f4"), ("by", "<f4"), ("bz", "<f4"), ("blasIdx", "<u4"), ("triIdx", "<u4")])
assert HIT_DT.itemsize == 24

SCENE_BUFFERS = ("tri", "triEx", "triIdx", "tlas", "blas", "bvh", "tex")
_ADDR_FIELD = {"tri": "tri_addr", "triEx": "triEx_addr", "triIdx": "triIdx_addr", "tlas": "tlas_addr", "blas": "blas_addr",
               "bvh": "bvh_addr", "tex": "tex_addr"}


class Tracer:
    def __init__(self, width, height, spp=1, max_depth=1, cuda_device=None, gpus=None):
        self.width, self.height = int(width), int(height)
        self.karg = np.zeros(1, dtype=KARG_DT)
        self.karg["dst_width"], self.karg["dst_height"] = self.width, self.height
        self.karg["samples_per_pixel"], self.karg["max_depth"] = spp, max_depth
        self.dev = vx.Device(cuda_device, gpus=gpus)
        self.bufs = {}
        self.krnl = None
        self.out = None
        self.args = None

    # Tracer::init (tracer.cpp:111-146)
    def init(self, scene_buffers):
        self.krnl = self.dev.upload_kernel_bytes(vx.DUMMY_KERNEL_IMAGE)
        for name in SCENE_BUFFERS:
            host = scene_buffers[name]
            b = self.dev.mem_alloc(max(int(host.nbytes), 4), vx.VX_MEM_READ)
            self.bufs[name] = b
            self.karg[_ADDR_FIELD[name]] = b.address
        self.out = self.dev.mem_alloc(self.width * self.height * 4, vx.VX_MEM_WRITE)
        self.karg["dst_addr"] = self.out.address
        self.host_scene = scene_buffers
        return self

    # Tracer::setup (tracer.cpp:152-219): camera/light fields come from the host library's setup math
    def setup(self, karg_fields):
        for k in ("tlas_root", "camera_pos", "camera_forward", "camera_right", "camera_up", "viewplane", "light_pos",
                  "light_color", "ambient_color", "background_color"):
            self.karg[k] = karg_fields[k]
        for name in SCENE_BUFFERS:
            host = self.host_scene[name]
            if host.nbytes:
                self.bufs[name].copy_to_dev(np.ascontiguousarray(host))
        return self

    def _upload_args(self):
        if self.args is not None:
            self.args.free()
        self.args = self.dev.upload_bytes(self.karg)

    # Tracer::run (tracer.cpp:231-240); `host_out` may be a numpy array or (pinned address, nbytes)
    def run(self, host_out=None):
        self._upload_args()
        self.dev.start(self.krnl, self.args)
        self.dev.ready_wait(vx.VX_MAX_TIMEOUT)
        if host_out is None:
            host_out = np.empty((self.height, self.width), dtype=np.uint32)
        self.out.copy_from_dev(host_out)
        return host_out

    def launch(self):
        """vx_start + vx_ready_wait only (inputs and framebuffer stay resident)."""
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

    def set_mode(self, mode=0, strict_order=False, kernel=None):
        self.dev.dcr_write(vx.VX_DCR_B200_RT_MODE, int(mode))
        self.dev.dcr_write(vx.VX_DCR_B200_RT_ORDER, int(bool(strict_order)))
        if kernel is not None:   # 0 = default (one pixel per lane over the 4-wide records), 1 = round 1's BVH2-walk kernel
            self.dev.dcr_write(vx.VX_DCR_B200_RT_KERNEL, int(kernel))

    def close(self):
        for b in list(self.bufs.values()) + [self.out, self.args, self.krnl]:
            if b is not None:
                b.free()
        self.bufs, self.out, self.args, self.krnl = {}, None, None, None
        self.dev.close()
