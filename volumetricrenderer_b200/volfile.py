"""Volume files (include/volrt.h, "volume files"): header + x-fastest texels (+ min-max pyramid), SHA-256 of the texels.

Host-side mirror of the library's file entry points; nothing here needs a device.  The reference has no volume
file at all (its volume lives in memory, TestMain.cpp:51-92; its only loader is stb_image for 2-D pictures,
VulkanTexture.cpp:14-55), so the format is this build's.
"""
import ctypes as C

import numpy as np

from . import capi

_DTYPE = {capi.VOL_RGBA8_UNORM: (np.uint8, 4), capi.VOL_R16F: (np.float16, 1), capi.VOL_R8_UNORM: (np.uint8, 1)}


class VolumeFileError(IOError):
    pass


def sha256(data):
    """SHA-256 by the library's own implementation (the one that signs volume files) -> 32 bytes"""
    b = np.frombuffer(bytes(data) if not isinstance(data, np.ndarray) else np.ascontiguousarray(data).view(np.uint8).reshape(-1), np.uint8)
    out = (C.c_uint8 * 32)()
    st = capi.load().vr_sha256(b.ctypes.data if b.size else None, b.size, out)
    if st != capi.VR_OK:
        raise ValueError("vr_sha256: status %d" % st)
    return bytes(out)


def pyramid_shapes(width, height, depth, fmt):
    """[(bz, by, bx, channels, 2)] per level, level 0 first"""
    ch = 4 if fmt == capi.VOL_RGBA8_UNORM else 1
    b = [(depth + 7) // 8, (height + 7) // 8, (width + 7) // 8]
    out = []
    while True:
        out.append((b[0], b[1], b[2], ch, 2))
        if b == [1, 1, 1]:
            return out
        b = [(x + 1) // 2 for x in b]


def split_pyramid(flat, width, height, depth, fmt):
    """flat float32 array (all levels back to back) -> list of per-level arrays shaped (bz, by, bx, channels, 2)"""
    out, o = [], 0
    for shp in pyramid_shapes(width, height, depth, fmt):
        n = int(np.prod(shp))
        out.append(flat[o:o + n].reshape(shp))
        o += n
    assert o == flat.size
    return out


def write(path, texels, fmt, address_mode=capi.ADDRESS_MIRRORED_REPEAT, pyramid=None):
    """texels shaped (d, h, w[, 4]); pyramid: flat float32 array of every level, or None"""
    dt, ch = _DTYPE[fmt]
    t = np.ascontiguousarray(texels)
    if t.dtype == np.uint16 and dt == np.float16:
        t = t.view(np.float16)
    if t.dtype != dt or t.ndim != (4 if ch == 4 else 3):
        raise TypeError("texels do not match the format")
    d, h, w = t.shape[:3]
    desc = capi.vr_volume_desc(C.sizeof(capi.vr_volume_desc), w, h, d, fmt, address_mode, capi.LAYOUT_AUTO)
    lv = C.c_uint32(0)
    pp = None
    if pyramid is not None:
        pyramid = np.ascontiguousarray(pyramid, np.float32).reshape(-1)
        if pyramid.size != capi.load().vr_pyramid_floats(w, h, d, fmt, C.byref(lv)):
            raise ValueError("pyramid has the wrong size")
        pp = pyramid.ctypes.data_as(C.POINTER(C.c_float))
    st = capi.load().vr_volume_file_write(path.encode(), C.byref(desc), t.ctypes.data, pp, lv.value)
    if st != capi.VR_OK:
        raise VolumeFileError("vr_volume_file_write(%s): %s" % (path, capi.STATUS_NAMES[st]))


def info(path):
    i = capi.vr_volume_file_info()
    i.struct_size = C.sizeof(i)
    st = capi.load().vr_volume_file_info_read(path.encode(), C.byref(i))
    if st != capi.VR_OK:
        raise VolumeFileError("vr_volume_file_info_read(%s): %s" % (path, capi.STATUS_NAMES[st]))
    return {"width": i.width, "height": i.height, "depth": i.depth, "format": i.format, "address_mode": i.address_mode,
            "pyramid_levels": i.pyramid_levels, "texel_bytes": i.texel_bytes, "pyramid_bytes": i.pyramid_bytes,
            "sha256": bytes(i.sha256)}


def read(path, verify=True, want_pyramid=False):
    """-> (texels shaped (d,h,w[,4]), info dict[, flat pyramid])"""
    i = info(path)
    dt, ch = _DTYPE[i["format"]]
    shape = (i["depth"], i["height"], i["width"]) + ((4,) if ch == 4 else ())
    t = np.empty(shape, dt)
    pyr = np.empty(i["pyramid_bytes"] // 4, np.float32) if want_pyramid else None
    st = capi.load().vr_volume_file_read(path.encode(), t.ctypes.data, t.nbytes,
                                         pyr.ctypes.data_as(C.POINTER(C.c_float)) if want_pyramid else None,
                                         pyr.nbytes if want_pyramid else 0, capi.FILE_VERIFY if verify else 0)
    if st != capi.VR_OK:
        raise VolumeFileError("vr_volume_file_read(%s): %s" % (path, capi.STATUS_NAMES[st]))
    return (t, i, pyr) if want_pyramid else (t, i)
