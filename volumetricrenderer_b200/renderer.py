"""Host-side mirror of the reference's renderer surface for the hot path, over the libvolrt C ABI.

Reference call site (source/Etna/TestMain.cpp)            ->  here
    renderer.Init(); AddRenderPass("BasePass", ...) :121-166  ->  Renderer(width, height, format)
    vkc::Texture3D texture(pixelData, extent)       :128      ->  Renderer.upload_volume(texels)
    the start-up bake                               :43-92    ->  Renderer.generate_volume()
    object/globalUniformBuffer.Update(&osd/&gsd)    :248-249  ->  Renderer.set_params(ubo0, ubo1, ...)
    EnqueueRenderPass + BeginFrame/EndFrame         :194-255  ->  Renderer.render()
    renderer.Shutdown()                             :260      ->  Renderer.close()

Every method is one C-ABI call into libvolrt.so (CUDA, sm_100a).  There is no CPU path: errors
(including "no CUDA device") raise VolrtError carrying vr_last_error().
"""
import ctypes as C

import numpy as np

from . import capi

_OUT_DTYPE = {capi.OUT_RGBA8_UNORM: np.uint8, capi.OUT_RGBA8_SRGB: np.uint8, capi.OUT_RGBA16F: np.float16,
              capi.OUT_RGBA32F: np.float32}
_VOL_DTYPE = {capi.VOL_RGBA8_UNORM: (np.uint8, 4), capi.VOL_R16F: (np.float16, 1), capi.VOL_R8_UNORM: (np.uint8, 1)}


class VolrtError(RuntimeError):
    def __init__(self, status, message):
        name = capi.STATUS_NAMES[status] if 0 <= status < len(capi.STATUS_NAMES) else str(status)
        super().__init__("%s: %s" % (name, message))
        self.status = status


class PinnedArray:
    """numpy view over page-locked host memory from vr_host_alloc (frees on close / GC)."""

    def __init__(self, shape, dtype):
        self._lib = capi.load()
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self.ptr = self._lib.vr_host_alloc(nbytes)
        if not self.ptr:
            raise VolrtError(capi.VR_ERR_OUT_OF_MEMORY, "vr_host_alloc(%d) failed" % nbytes)
        buf = (C.c_char * nbytes).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dtype).reshape(shape)

    def close(self):
        if self.ptr:
            self.array = None
            self._lib.vr_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Renderer:
    def __init__(self, width=1280, height=720, fmt=capi.OUT_RGBA8_UNORM, device=0, max_views=1, devices=None):
        """devices: CUDA ordinals of several GPUs behind this ONE context (vr_create_info.device_count): the frame lives on
        devices[0], vr_render cuts it into row bands (or a view batch into view shards) over all of them."""
        self._lib = capi.load()
        self._ctx = C.c_void_p()
        self.width, self.height, self.format, self.max_views = width, height, fmt, max_views
        info = capi.vr_create_info(C.sizeof(capi.vr_create_info), device, width, height, fmt, max_views, 0)
        if devices is not None and len(devices) > 1:
            info.device_count = len(devices)
            for i, d in enumerate(devices):
                info.devices[i] = d
            info.device = devices[0]
        st = self._lib.vr_create(C.byref(info), C.byref(self._ctx))
        if st != capi.VR_OK:
            msg = self._lib.vr_last_error(None)
            self._ctx = C.c_void_p()
            raise VolrtError(st, (msg or b"").decode())
        self.view_count = 1
        self._keep = []

    # ---- plumbing
    def _check(self, st):
        if st != capi.VR_OK:
            raise VolrtError(st, (self._lib.vr_last_error(self._ctx) or b"").decode())

    def close(self):
        if self._ctx:
            self._lib.vr_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- volume
    def upload_volume(self, texels, fmt=capi.VOL_RGBA8_UNORM, layout=capi.LAYOUT_AUTO,
                      address_mode=capi.ADDRESS_MIRRORED_REPEAT):
        """texels: array shaped (depth, height, width[, 4]), tightly packed, x fastest."""
        dt, ch = _VOL_DTYPE[fmt]
        t = np.ascontiguousarray(texels)
        if t.dtype == np.uint16 and dt == np.float16:
            t = t.view(np.float16)
        if t.dtype != dt:
            raise TypeError("volume dtype %s does not match format (want %s)" % (t.dtype, np.dtype(dt)))
        if t.ndim != (4 if ch == 4 else 3) or (ch == 4 and t.shape[3] != 4):
            raise ValueError("volume must be shaped (d,h,w%s)" % (",4" if ch == 4 else ""))
        d, h, w = t.shape[:3]
        desc = capi.vr_volume_desc(C.sizeof(capi.vr_volume_desc), w, h, d, fmt, address_mode, layout)
        self._check(self._lib.vr_upload_volume(self._ctx, C.byref(desc), t.ctypes.data))
        self._vol_shape, self._vol_fmt = t.shape, fmt

    def default_noise_desc(self):
        d = capi.vr_noise_desc()
        self._check(self._lib.vr_noise_desc_default(C.byref(d)))
        return d

    def generate_volume(self, desc=None, fbm_params=None, **kw):
        """Bake on the device (TestMain.cpp:43-92).  kw overrides fields of the default recipe."""
        d = desc if desc is not None else self.default_noise_desc()
        for k, v in kw.items():
            setattr(d, k, v)
        self._check(self._lib.vr_generate_volume(self._ctx, C.byref(d), C.byref(fbm_params) if fbm_params is not None else None))
        n = d.size
        self._vol_fmt = d.format
        self._vol_shape = (n, n, n, 4) if d.format == capi.VOL_RGBA8_UNORM else (n, n, n)

    def download_volume(self, out=None):
        dt, _ = _VOL_DTYPE[self._vol_fmt]
        if out is None:
            out = np.empty(self._vol_shape, dt)
        self._check(self._lib.vr_download_volume(self._ctx, out.ctypes.data, out.nbytes))
        return out

    # ---- volume files (volfile.py has the host-only half)
    def upload_volume_file(self, path, layout=capi.LAYOUT_AUTO, flags=capi.FILE_VERIFY):
        from . import volfile
        self._check(self._lib.vr_upload_volume_file(self._ctx, path.encode(), layout, flags))
        i = volfile.info(path)
        self._vol_fmt = i["format"]
        self._vol_shape = (i["depth"], i["height"], i["width"]) + ((4,) if i["format"] == capi.VOL_RGBA8_UNORM else ())

    def save_volume_file(self, path, with_pyramid=False):
        self._check(self._lib.vr_save_volume_file(self._ctx, path.encode(), capi.FILE_WITH_PYRAMID if with_pyramid else 0))

    def volume_pyramid(self):
        """min-max pyramid of the resident volume (device-computed) -> list of (bz,by,bx,channels,2) float32 arrays"""
        from . import volfile
        d, h, w = self._vol_shape[:3]
        n = self._lib.vr_pyramid_floats(w, h, d, self._vol_fmt, None)
        flat = np.empty(n, np.float32)
        self._check(self._lib.vr_volume_pyramid(self._ctx, flat.ctypes.data_as(C.POINTER(C.c_float)), flat.nbytes))
        return volfile.split_pyramid(flat, w, h, d, self._vol_fmt)

    # ---- FastNoise2 node trees (noisetree.py builds / compiles them)
    def noise_tree_grid(self, tree, shape, frequency, seed, start=(0, 0, 0), want_grid=True):
        """tree: encoded string, noisetree.N or noisetree.CompiledTree; shape = (nz, ny, nx) -> (grid float32, (min, max))"""
        from . import noisetree
        own = None
        if isinstance(tree, noisetree.N):
            tree = tree.encode()
        if not isinstance(tree, noisetree.CompiledTree):
            tree = own = noisetree.CompiledTree(tree)
        nz, ny, nx = shape
        out = np.empty(shape, np.float32) if want_grid else None
        mm = (C.c_float * 2)()
        try:
            self._check(self._lib.vr_noise_tree_grid(self._ctx, tree.handle, start[0], start[1], start[2], nx, ny, nz, frequency, seed,
                                                     out.ctypes.data_as(C.POINTER(C.c_float)) if want_grid else None, mm))
        finally:
            if own is not None:
                own.close()
        return out, (mm[0], mm[1])

    # ---- params
    def default_params(self):
        p = capi.vr_params()
        self._check(self._lib.vr_params_default(C.byref(p)))
        return p

    def set_params(self, ubo0=None, ubo1=None, view_ubos=None, params=None, **kw):
        """ubo0 / ubo1: the byte images the host writes (arrays or bytes; 192 B and 140 or 144 B).
        view_ubos: uint8[count, 336] for a multi-view batch.  kw sets any other vr_params field."""
        p = params if params is not None else self.default_params()
        keep = []
        if ubo0 is not None:
            b0 = np.frombuffer(np.ascontiguousarray(ubo0).tobytes(), np.uint8).copy()
            p.ubo0, p.ubo0_size = b0.ctypes.data, b0.size
            keep.append(b0)
        if ubo1 is not None:
            b1 = np.frombuffer(np.ascontiguousarray(ubo1).tobytes(), np.uint8).copy()
            p.ubo1, p.ubo1_size = b1.ctypes.data, b1.size
            keep.append(b1)
        if view_ubos is not None:
            bv = np.ascontiguousarray(view_ubos, np.uint8)
            if bv.ndim != 2 or bv.shape[1] != 336:
                raise ValueError("view_ubos must be uint8[count, 336]")
            p.view_ubos, p.view_count = bv.ctypes.data, bv.shape[0]
            keep.append(bv)
        tree = kw.get("density_tree")
        if tree is not None and hasattr(tree, "handle"):        # a noisetree.CompiledTree: pass its handle
            kw = dict(kw, density_tree=tree.handle)
            kw.setdefault("density_mode", capi.DENSITY_TREE)
        for k, v in kw.items():
            cur = getattr(p, k)
            if hasattr(cur, "__len__"):
                vals = [float(x) for x in np.asarray(v, np.float32).ravel()]
                if len(vals) != len(cur):
                    raise ValueError("%s needs %d values" % (k, len(cur)))
                for i, x in enumerate(vals):
                    cur[i] = x
            else:
                setattr(p, k, v)
        self._check(self._lib.vr_set_params(self._ctx, C.byref(p)))
        self.view_count = max(1, p.view_count)
        return p

    # ---- render
    def new_frame_buffer(self, views=None, pinned=True):
        """Host buffer shaped (views, H, W, 4) in the context's output format."""
        shape = (views or self.view_count, self.height, self.width, 4)
        if pinned:
            pa = PinnedArray(shape, _OUT_DTYPE[self.format])
            self._keep.append(pa)
            return pa.array
        return np.empty(shape, _OUT_DTYPE[self.format])

    def render(self, out=None, rows=None, views=None, flags=0, to_host=True):
        """Render; returns the frame(s) as (views, rows, W, 4) host array (or None if to_host=False)."""
        r0, r1 = rows if rows is not None else (0, self.height)
        v0, v1 = views if views is not None else (0, self.view_count)
        t = capi.vr_target()
        t.struct_size = C.sizeof(capi.vr_target)
        t.row_begin, t.row_end, t.view_begin, t.view_end, t.flags = r0, r1, v0, v1, flags
        if r0 == 0 and r1 == 0:
            raise ValueError("empty row band")
        if to_host:
            if out is None:
                out = np.empty((self.view_count, self.height, self.width, 4), _OUT_DTYPE[self.format])
            if out.shape[-3:] != (self.height, self.width, 4) or not out.flags.c_contiguous:
                raise ValueError("out must be a contiguous (views, H, W, 4) array")
            t.memory_kind, t.dst = capi.MEM_HOST, out.ctypes.data
        self._check(self._lib.vr_render(self._ctx, C.byref(t)))
        if not to_host:
            return None
        full = out.reshape(-1, self.height, self.width, 4)
        return full[v0:v1, r0:r1]

    def render_device(self, dst_ptr, rows=None, row_pitch=0, flags=capi.RENDER_DIRECT):
        """Render a row band straight into device memory (own or peer-mapped) — the NVLink gather."""
        r0, r1 = rows if rows is not None else (0, self.height)
        t = capi.vr_target()
        t.struct_size = C.sizeof(capi.vr_target)
        t.memory_kind, t.dst, t.row_pitch = capi.MEM_DEVICE, dst_ptr, row_pitch
        t.row_begin, t.row_end, t.flags = r0, r1, flags
        self._check(self._lib.vr_render(self._ctx, C.byref(t)))

    def wait(self):
        self._check(self._lib.vr_wait(self._ctx))

    def flush_l2(self):
        self._check(self._lib.vr_flush_l2(self._ctx))

    def measure_peaks(self):
        out = (C.c_float * 4)()
        self._check(self._lib.vr_measure_peaks(self._ctx, out))
        return {"fp32_tflops": out[0], "tex_gfetch": out[1], "sms": int(out[2]), "sm_max_mhz": out[3]}

    def measure_tex_stream(self):
        """GB/s of the texture path streaming the resident one-channel volume out of DRAM once (vr_measure_tex_stream)"""
        g = C.c_float()
        self._check(self._lib.vr_measure_tex_stream(self._ctx, C.byref(g)))
        return g.value

    def stats(self):
        s = capi.vr_stats()
        s.struct_size = C.sizeof(capi.vr_stats)
        self._check(self._lib.vr_get_stats(self._ctx, C.byref(s)))
        return {k: getattr(s, k) for k, _ in capi.vr_stats._fields_ if k != "struct_size"}

    def frame_info(self):
        ptr, pitch, stride = C.c_void_p(), C.c_size_t(), C.c_size_t()
        self._check(self._lib.vr_frame_info(self._ctx, C.byref(ptr), C.byref(pitch), C.byref(stride)))
        return ptr.value, pitch.value, stride.value

    def row_bands(self, parts):
        b = (C.c_uint32 * (parts + 1))()
        self._check(self._lib.vr_row_bands(self._ctx, parts, b))
        return list(b)

    def row_cost(self):
        cost = np.zeros(self.height, np.float32)
        self._check(self._lib.vr_row_cost(self._ctx, cost.ctypes.data_as(C.POINTER(C.c_float))))
        return cost

    def frame_export(self):
        h = (C.c_char * 64)()
        self._check(self._lib.vr_frame_export(self._ctx, C.cast(h, C.c_void_p)))
        return bytes(h)

    def frame_import(self, handle):
        buf = (C.c_char * 64).from_buffer_copy(handle)
        mapped = C.c_void_p()
        self._check(self._lib.vr_frame_import(self._ctx, C.cast(buf, C.c_void_p), C.byref(mapped)))
        return mapped.value

    # ---- device-side completion of the multi-GPU gather (include/volrt.h: vr_gather_*)
    def gather_acquire(self, frame0, frames, slot, releases):
        self._check(self._lib.vr_gather_acquire(self._ctx, frame0, frames, slot, releases))

    def gather_signal(self, frame0, frames, slot):
        self._check(self._lib.vr_gather_signal(self._ctx, frame0, frames, slot))

    def gather_collect(self, slot, arrivals, out):
        """GPU 0: when `arrivals` bands have landed in frame slot `slot`, copy it to out (H, W, 4; page-locked) and release
        the slot; returns at once, vr_wait completes it"""
        if out.shape != (self.height, self.width, 4) or not out.flags.c_contiguous:
            raise ValueError("out must be a contiguous (H, W, 4) array")
        self._check(self._lib.vr_gather_collect(self._ctx, slot, arrivals, out.ctypes.data, 0))

    def frame_unmap(self, mapped):
        self._check(self._lib.vr_frame_unmap(self._ctx, mapped))
