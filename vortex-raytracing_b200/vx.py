"""ctypes binding of libvortex-b200.so — the vx_* runtime C ABI (include/vortex_b200.h).

Function names, argument order and error behaviour are the reference's (runtime/include/vortex.h:80-145):
every call returns 0 on success; this wrapper raises VxError on a non-zero return.  There is no CPU
fallback: loading fails loudly when the CUDA library has not been built, and vx_dev_open() fails when no
B200 is visible.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libvortex-b200.so")

VX_MEM_READ, VX_MEM_WRITE, VX_MEM_READ_WRITE = 1, 2, 3
VX_MAX_TIMEOUT = 24 * 60 * 60 * 1000
VX_DCR_BASE_RTX_TLAS_PTR, VX_DCR_BASE_RTX_BLAS_PTR, VX_DCR_BASE_RTX_BVH_PTR, VX_DCR_BASE_RTX_TRI_PTR = 6, 7, 8, 9
VX_DCR_B200_RT_MODE, VX_DCR_B200_RT_LAYOUT, VX_DCR_B200_RT_ORDER, VX_DCR_B200_RT_CAMERA, VX_DCR_B200_RT_KERNEL = 0x0F0, 0x0F1, 0x0F2, 0x0F3, 0x0F4
MODE_SHADOW, MODE_DIFFUSE = 1, 2

# every symbol include/vortex_b200.h declares (tests check that the library exports all of them)
EXPORTS = (
    "vx_dev_init vx_dev_open vx_dev_close vx_dev_caps vx_mem_alloc vx_mem_reserve vx_mem_free vx_mem_access "
    "vx_mem_address vx_mem_info vx_copy_to_dev vx_copy_from_dev vx_start vx_ready_wait vx_dcr_read vx_dcr_write "
    "vx_mpm_query vx_upload_kernel_bytes vx_upload_kernel_file vx_upload_bytes vx_upload_file vx_dump_perf "
    "vx_b200_select_device vx_b200_select_gpus vx_b200_set_readback vx_b200_flush_l2 vx_b200_set_partition vx_b200_build_bvh vx_b200_build_bvh_static vx_b200_build_qbvh vx_b200_export_buffer vx_b200_set_remote_framebuffer vx_b200_device_ptr vx_b200_mark_written vx_b200_probe_gather vx_b200_stream vx_b200_last_kernel_ms "
    "vx_b200_enable_stats vx_b200_get_stats vx_b200_trace_primary vx_b200_trace_rays vx_b200_trace_rays_host vx_b200_last_error"
).split()


class VxError(RuntimeError):
    pass


class Stats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("box_tests", C.c_uint64), ("tri_tests", C.c_uint64), ("blas_entered", C.c_uint64),
                ("shaded_hits", C.c_uint64), ("kernel_launches", C.c_uint64), ("prep_launches", C.c_uint64),
                ("kernel_ms", C.c_float), ("prep_ms", C.c_float),
                ("lane_steps", C.c_uint64), ("warp_steps", C.c_uint64), ("warps", C.c_uint64),
                ("gpus", C.c_uint32), ("group_ms", C.c_float), ("kernel_ms_min", C.c_float), ("kernel_ms_max", C.c_float),
                ("latency_dist", (C.c_uint64 * 5) * 3), ("steps_hist", C.c_uint64 * 16)]


_lib = None


def lib():
    """Load the CUDA backend.  Raises if it was not built (run __graft_entry__.build() or make -C csrc)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise VxError(f"{LIB_PATH} is missing: build it with `make -C vortex-raytracing_b200/csrc` "
                          "(the b200 backend has no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        H, B, u64, u32, i32, vp = C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_int, C.c_void_p
        sig = {
            "vx_dev_open": [C.POINTER(H)], "vx_dev_close": [H], "vx_dev_caps": [H, u32, C.POINTER(u64)],
            "vx_mem_alloc": [H, u64, i32, C.POINTER(B)], "vx_mem_reserve": [H, u64, u64, i32, C.POINTER(B)],
            "vx_mem_free": [B], "vx_mem_access": [B, u64, u64, i32], "vx_mem_address": [B, C.POINTER(u64)],
            "vx_mem_info": [H, C.POINTER(u64), C.POINTER(u64)], "vx_copy_to_dev": [B, vp, u64, u64],
            "vx_copy_from_dev": [vp, B, u64, u64], "vx_start": [H, B, B], "vx_ready_wait": [H, u64],
            "vx_dcr_read": [H, u32, C.POINTER(u32)], "vx_dcr_write": [H, u32, u32], "vx_mpm_query": [H, u32, u32, C.POINTER(u64)],
            "vx_upload_kernel_bytes": [H, vp, u64, C.POINTER(B)], "vx_upload_kernel_file": [H, C.c_char_p, C.POINTER(B)],
            "vx_upload_bytes": [H, vp, u64, C.POINTER(B)], "vx_upload_file": [H, C.c_char_p, C.POINTER(B)],
            "vx_b200_select_device": [i32], "vx_b200_select_gpus": [i32], "vx_b200_set_readback": [H, B, vp, u64], "vx_b200_flush_l2": [H],
            "vx_b200_set_partition": [H, u32, u32, u32], "vx_b200_device_ptr": [B, C.POINTER(u64)], "vx_b200_mark_written": [B], "vx_b200_probe_gather": [H, u64, u32, C.POINTER(C.c_double)],
            "vx_b200_stream": [H, C.POINTER(u64)], "vx_b200_last_kernel_ms": [H, C.POINTER(C.c_float)],
            "vx_b200_enable_stats": [H, i32], "vx_b200_get_stats": [H, C.POINTER(Stats)],
            "vx_b200_trace_primary": [H, B, B], "vx_b200_trace_rays": [H, B, B, u64, B],
            "vx_b200_trace_rays_host": [H, B, C.c_void_p, u64, C.c_void_p],
            "vx_b200_build_bvh": [H, B, u64, u64, B, u64, B, u64, C.POINTER(u64)],
            "vx_b200_build_bvh_static": [H, B, u64, u64, B, u64, B, u64, C.POINTER(u64)],
            "vx_b200_build_qbvh": [H, B, u64, u64, B, B, u64, C.POINTER(u64), C.POINTER(C.c_float)],
            "vx_b200_export_buffer": [B, vp, C.POINTER(u64)], "vx_b200_set_remote_framebuffer": [H, vp, u64, u64],
        }
        for name, args in sig.items():
            f = getattr(L, name)
            f.argtypes = args
            f.restype = C.c_int
        L.vx_b200_last_error.restype = C.c_char_p
        _lib = L
    return _lib


def _check(rc, what):
    if rc != 0:
        msg = lib().vx_b200_last_error().decode(errors="replace")
        raise VxError(f"{what} returned {rc}" + (f": {msg}" if msg else ""))


def _ptr(a):
    if isinstance(a, np.ndarray):
        if not a.flags["C_CONTIGUOUS"]:
            raise ValueError("host array must be C-contiguous")
        return a.ctypes.data, a.nbytes, a
    if isinstance(a, (bytes, bytearray)):
        buf = (C.c_char * len(a)).from_buffer_copy(a)
        return C.addressof(buf), len(a), buf
    if isinstance(a, tuple):   # (address, nbytes): e.g. a pinned torch tensor's data_ptr()
        return a[0], a[1], None
    raise TypeError(type(a))


class Buffer:
    def __init__(self, dev, handle, size):
        self.dev, self.h, self.size = dev, handle, size

    @property
    def address(self):
        v = C.c_uint64()
        _check(lib().vx_mem_address(self.h, C.byref(v)), "vx_mem_address")
        return v.value

    def mark_written(self):
        """the device memory was written without copy_to_dev (NCCL receive): derived GPU records are rebuilt at the next start"""
        _check(lib().vx_b200_mark_written(self.h), "vx_b200_mark_written")

    @property
    def device_ptr(self):
        v = C.c_uint64()
        _check(lib().vx_b200_device_ptr(self.h, C.byref(v)), "vx_b200_device_ptr")
        return v.value

    def ipc_export(self):
        """(64-byte CUDA IPC handle of the device arena, offset of this buffer in it) for vx_b200_set_remote_framebuffer"""
        hb = C.create_string_buffer(64)
        off = C.c_uint64()
        _check(lib().vx_b200_export_buffer(self.h, hb, C.byref(off)), "vx_b200_export_buffer")
        return hb.raw, off.value

    def copy_to_dev(self, host, dst_offset=0):
        p, n, _keep = _ptr(host)
        _check(lib().vx_copy_to_dev(self.h, p, dst_offset, n), "vx_copy_to_dev")

    def copy_from_dev(self, host, src_offset=0):
        p, n, _keep = _ptr(host)
        _check(lib().vx_copy_from_dev(p, self.h, src_offset, n), "vx_copy_from_dev")

    def free(self):
        if self.h:
            _check(lib().vx_mem_free(self.h), "vx_mem_free")
            self.h = None


class Device:
    """vx_device_h with the reference's call names as methods."""

    def __init__(self, cuda_device=None, gpus=None):
        """gpus: how many GPUs of the box this ONE device handle spans (vx_b200_select_gpus; default env VORTEX_B200_GPUS, else 1)"""
        L = lib()
        if cuda_device is not None:
            L.vx_b200_select_device(int(cuda_device))
        L.vx_b200_select_gpus(int(gpus) if gpus else 0)
        h = C.c_void_p()
        try:
            _check(L.vx_dev_open(C.byref(h)), "vx_dev_open")
        finally:
            L.vx_b200_select_gpus(0)
        self.h = h

    def close(self):
        if self.h:
            _check(lib().vx_dev_close(self.h), "vx_dev_close")
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def caps(self, caps_id):
        v = C.c_uint64()
        _check(lib().vx_dev_caps(self.h, caps_id, C.byref(v)), "vx_dev_caps")
        return v.value

    def mem_alloc(self, size, flags=VX_MEM_READ_WRITE):
        b = C.c_void_p()
        _check(lib().vx_mem_alloc(self.h, size, flags, C.byref(b)), "vx_mem_alloc")
        return Buffer(self, b, size)

    def mem_reserve(self, address, size, flags=0):
        b = C.c_void_p()
        _check(lib().vx_mem_reserve(self.h, address, size, flags, C.byref(b)), "vx_mem_reserve")
        return Buffer(self, b, size)

    def mem_info(self):
        f, u = C.c_uint64(), C.c_uint64()
        _check(lib().vx_mem_info(self.h, C.byref(f), C.byref(u)), "vx_mem_info")
        return f.value, u.value

    def upload_bytes(self, host):
        p, n, _keep = _ptr(host)
        b = C.c_void_p()
        _check(lib().vx_upload_bytes(self.h, p, n, C.byref(b)), "vx_upload_bytes")
        return Buffer(self, b, n)

    def upload_kernel_bytes(self, content):
        p, n, _keep = _ptr(content)
        b = C.c_void_p()
        _check(lib().vx_upload_kernel_bytes(self.h, p, n, C.byref(b)), "vx_upload_kernel_bytes")
        return Buffer(self, b, n)

    def start(self, kernel, arguments):
        _check(lib().vx_start(self.h, kernel.h, arguments.h), "vx_start")

    def ready_wait(self, timeout=VX_MAX_TIMEOUT):
        _check(lib().vx_ready_wait(self.h, timeout), "vx_ready_wait")

    def dcr_write(self, addr, value):
        _check(lib().vx_dcr_write(self.h, addr, value), "vx_dcr_write")

    def dcr_read(self, addr):
        v = C.c_uint32()
        _check(lib().vx_dcr_read(self.h, addr, C.byref(v)), "vx_dcr_read")
        return v.value

    def mpm_query(self, addr, core_id=0):
        v = C.c_uint64()
        _check(lib().vx_mpm_query(self.h, addr, core_id, C.byref(v)), "vx_mpm_query")
        return v.value

    # ---- b200 extensions ----
    def build_bvh(self, tri, tri_count, bvh, tri_idx, tri_first=0, node_first=0, slot_first=0):
        """GPU build of the layout-A BVH2 of triangles [tri_first, tri_first+tri_count) of buffer `tri` into `bvh` / `tri_idx`;
        returns the number of nodes written (2*tri_count-1)."""
        used = C.c_uint64()
        _check(lib().vx_b200_build_bvh(self.h, tri.h, tri_first, tri_count, bvh.h, node_first, tri_idx.h, slot_first, C.byref(used)),
               "vx_b200_build_bvh")
        return used.value

    def build_bvh_static(self, tri, tri_count, bvh, tri_idx, tri_first=0, node_first=0, slot_first=0):
        """build_bvh without any allocation (scratch in a static device array; at most 65 536 triangles, larger counts raise)."""
        used = C.c_uint64()
        _check(lib().vx_b200_build_bvh_static(self.h, tri.h, tri_first, tri_count, bvh.h, node_first, tri_idx.h, slot_first, C.byref(used)),
               "vx_b200_build_bvh_static")
        return used.value

    def build_qbvh(self, tri, tri_count, tri_ex, qbvh, tri_first=0, node_first=0):
        """GPU build of the layout-B quantized BVH4 of triangles [tri_first, tri_first+tri_count); `tri` / `tri_ex` are permuted
        in place into leaf order.  Returns (nodes written, root box as 6 floats)."""
        used = C.c_uint64()
        box = (C.c_float * 6)()
        _check(lib().vx_b200_build_qbvh(self.h, tri.h, tri_first, tri_count, tri_ex.h if tri_ex is not None else None, qbvh.h,
                                        node_first, C.byref(used), box), "vx_b200_build_qbvh")
        return used.value, np.array(list(box), dtype=np.float32)

    def set_remote_framebuffer(self, handle, offset=0, nbytes=0):
        """handle: the 64 bytes of Buffer.ipc_export() from the image owner's process (None detaches)"""
        hb = C.create_string_buffer(bytes(handle), 64) if handle is not None else None
        _check(lib().vx_b200_set_remote_framebuffer(self.h, hb, C.c_uint64(offset), C.c_uint64(nbytes)), "vx_b200_set_remote_framebuffer")

    def set_partition(self, rank, world, band_rows=8):
        _check(lib().vx_b200_set_partition(self.h, rank, world, band_rows), "vx_b200_set_partition")

    def stream(self):
        v = C.c_uint64()
        _check(lib().vx_b200_stream(self.h, C.byref(v)), "vx_b200_stream")
        return v.value

    def last_kernel_ms(self):
        v = C.c_float()
        _check(lib().vx_b200_last_kernel_ms(self.h, C.byref(v)), "vx_b200_last_kernel_ms")
        return v.value

    def probe_gather(self, buffer_bytes=32 << 20, record_bytes=32):
        """random-record gather bandwidth (GB/s) over a scratch buffer: the L2 -> SM roofline denominator (vx_b200_probe_gather)"""
        v = C.c_double()
        _check(lib().vx_b200_probe_gather(self.h, int(buffer_bytes), int(record_bytes), C.byref(v)), "vx_b200_probe_gather")
        return v.value

    def enable_stats(self, on=True):
        _check(lib().vx_b200_enable_stats(self.h, int(on)), "vx_b200_enable_stats")

    def stats(self):
        s = Stats()
        _check(lib().vx_b200_get_stats(self.h, C.byref(s)), "vx_b200_get_stats")
        out = {n: getattr(s, n) for n, _ in Stats._fields_}
        out["latency_dist"] = [[int(v) for v in row] for row in s.latency_dist]
        out["steps_hist"] = [int(v) for v in s.steps_hist]
        return out

    def set_readback(self, framebuffer, host_ptr, nbytes):
        """every later vx_start of `framebuffer` ends with the copy of the image into pinned host memory at `host_ptr`
        (each GPU copies its own bands); None cancels"""
        _check(lib().vx_b200_set_readback(self.h, framebuffer.h, C.c_void_p(host_ptr) if host_ptr else None, C.c_uint64(nbytes)),
               "vx_b200_set_readback")

    def flush_l2(self):
        _check(lib().vx_b200_flush_l2(self.h), "vx_b200_flush_l2")

    def trace_primary(self, arguments, hits):
        _check(lib().vx_b200_trace_primary(self.h, arguments.h, hits.h), "vx_b200_trace_primary")

    def trace_rays(self, arguments, rays, count, hits):
        _check(lib().vx_b200_trace_rays(self.h, arguments.h, rays.h, count, hits.h), "vx_b200_trace_rays")

    def trace_rays_host(self, arguments, rays_ptr, count, hits_ptr):
        """rays / hit records in host memory (raw addresses), chunked and pipelined by the backend; synchronous"""
        _check(lib().vx_b200_trace_rays_host(self.h, arguments.h, C.c_void_p(rays_ptr), count, C.c_void_p(hits_ptr)), "vx_b200_trace_rays_host")


# a minimal stand-in for kernel.vxbin: {min_vma, max_vma} header + 1 byte (runtime/stub/utils.cpp:25-61).
# The b200 backend reserves the range like any driver and never executes the image.
DUMMY_KERNEL_IMAGE = np.array([0x80000000, 0x80001000], dtype=np.uint64).tobytes() + b"\x00" * 16


def dummy_image(vma):
    """stand-in for a shader image linked at `vma` (the raytracing test links miss/closest/anyhit at 0x8010/20/30_0000,
    tests/regression/raytracing/Makefile:56-59)"""
    return np.array([vma, vma + 0x1000], dtype=np.uint64).tobytes() + b"\x00" * 16
